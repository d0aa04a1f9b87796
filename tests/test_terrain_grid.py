"""Sub-terrain GRID mode of the terrain manager (examples/locomotion/terrain_manager.py:104-133,179-214,308-335).
The height field is Genesis' `gs.morphs.Terrain(n_subterrains, subterrain_types)` (genesis/utils/terrain.py parse_terrain over the
Isaac Gym tile generators), restated in go2_b200/terrain.py and pinned here by sha256 digests of the REFERENCE functions' own output
(oracle/ref_terrain_tiles.py runs them - the numba-compiled spline included - in the build container and prints the digests).
Replayed-draw parity of the env in this mode: tests/test_golden_dr.py, fixture walk_grid_c1."""
import hashlib
import json
import os

import numpy as np
import pytest
import torch

from util import ROOT, make_env

BACKENDS = ["hostemu", pytest.param("cuda", marks=pytest.mark.gpu)]
ALL8 = [["flat_terrain", "random_uniform_terrain", "sloped_terrain", "pyramid_sloped_terrain"],
        ["discrete_obstacles_terrain", "wave_terrain", "stairs_terrain", "pyramid_stairs_terrain"]]
# (types, tile size, horizontal scale, vertical scale, randomize, parameters) after np.random.seed(123) -> digest of the reference's field
REF_DIGESTS = [
    ([["flat_terrain", "random_uniform_terrain"], ["random_uniform_terrain", "flat_terrain"]], (8.0, 8.0), 0.25, 0.005, True,
     {"random_uniform_terrain": {"min_height": -0.08, "max_height": 0.08}}, (65, 65), "d340162471566daf8706aaca162074253966620b66effaafeee65737ab19ed0b"),
    (ALL8, (12.0, 12.0), 0.25, 0.005, False, None, (97, 193), "04c9eca2f14b60337ec1ec77223e9a1c3a99cc3f9001cbf2c5ece30df7ab1eab"),
    ([["pyramid_stairs_terrain", "discrete_obstacles_terrain"], ["random_uniform_terrain", "wave_terrain"]], (6.0, 9.0), 0.1, 0.005, True,
     {"random_uniform_terrain": {"step": 0.02, "downsampled_scale": 0.3}}, (121, 181), "78e2697becc4b702724174776f51587e4760f1b07186cfe3e18a3b091ff291f1"),
    ([["fractal_terrain", "stepping_stones_terrain"], ["stepping_stones_terrain", "fractal_terrain"]], (6.0, 8.0), 0.25, 0.005, True,
     {"fractal_terrain": {"levels": 5, "scale": 0.4}, "stepping_stones_terrain": {"stone_size": 0.75, "platform_size": 1.0}}, (49, 65),
     "39563c94648b1014424e6aa6cf9683720f467e54d41389dcd4bdb2f1fc01fd0f"),
    ([["stepping_stones_terrain", "fractal_terrain"]], (8.0, 5.0), 0.25, 0.005, True, None, (33, 41), "0dc8328f7c77657d57d11510016dc3d3e38843be330a6e39f3953b798bd78ef6"),
]


def test_generated_fields_match_the_reference_generators():
    from go2_b200.terrain import build_subterrain_grid, not_a_knot_spline

    for types, size, hs, vs, rnd, params, shape, sha in REF_DIGESTS:
        np.random.seed(123)
        hf = build_subterrain_grid(types, size, hs, vs, rnd, params)
        assert hf.shape == shape and hf.dtype == np.float32 and np.isfinite(hf).all()
        assert hashlib.sha256(np.ascontiguousarray(hf).tobytes()).hexdigest() == sha
    # randomize=False leaves numpy's global generator untouched and always gives the same tiles
    np.random.seed(5)
    st = np.random.get_state()[1].copy()
    a = build_subterrain_grid(ALL8, (12.0, 12.0), 0.25, 0.005, False, None)
    assert np.array_equal(np.random.get_state()[1], st) and np.array_equal(a, build_subterrain_grid(ALL8, (12.0, 12.0), 0.25, 0.005, False, None))
    # flat tiles are zero inside; a border shared with a bumpy neighbour keeps the larger of the two values
    hf = build_subterrain_grid([["flat_terrain", "wave_terrain"]], (4.0, 4.0), 0.25, 0.005, False, None)
    assert (hf[:, :16] == 0).all() and hf[:, 16].min() >= 0 and hf[:, 17:].min() < 0
    # the spline interpolates its knots and reproduces a cubic exactly (not-a-knot ends)
    x = np.linspace(0.0, 3.0, 9)
    y = np.stack([x**3 - 2 * x, np.sin(x)], 1)
    assert np.allclose(not_a_knot_spline(x, y, x), y, atol=1e-5)
    xv = np.linspace(0.0, 3.0, 50)
    assert np.allclose(not_a_knot_spline(x, y, xv)[:, 0], xv**3 - 2 * xv, atol=2e-4)
    with pytest.raises(NotImplementedError):
        build_subterrain_grid([["lunar_terrain"]], (4.0, 4.0), 0.25, 0.005, False, None)


def _grid_cfgs():
    cfgs = json.load(open(os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200/go2_b200/cfgs/walk.json")))
    cfgs[0]["terrain"] = {"enabled": True, "subterrain_types": [["flat_terrain", "random_uniform_terrain", "wave_terrain"], ["pyramid_sloped_terrain", "flat_terrain", "stairs_terrain"]],
                          "subterrain_size": (6.0, 8.0), "horizontal_scale": 0.25, "vertical_scale": 0.005, "randomize": True, "spawn_height_offset": 0.05,
                          "boundary_margin": 1.0, "np_seed": 17,
                          "subterrain_parameters": {"random_uniform_terrain": {"min_height": -0.05, "max_height": 0.05}, "stairs_terrain": {"step_width": 1.0, "step_height": 0.02}}}
    cfgs[0]["base_init_pos"] = [0.0, 0.0, 0.45]
    return cfgs


@pytest.mark.parametrize("backend", BACKENDS)
def test_grid_spawn_blind_heights_and_boundary(backend, hostemu_lib):
    cfgs = _grid_cfgs()
    B = 192
    env = make_env("walk_terrain", cfgs, B, backend, hostemu_lib, seed=4)
    T = env.terrain
    assert T.enabled and not T.use_custom_heightmap and (T.n_rows, T.n_cols) == (2, 3)
    assert T.max_wander == pytest.approx(min(6.0, 8.0) / 2 - 1.0) and T._flat_row_indices == []
    assert T._height_field is None and T._height_field_np.shape == (2 * 24 + 1, 3 * 32 + 1)
    assert tuple(T._terrain_origin) == (-6.0, -12.0, 0.0)
    env.reset()
    sp = T._env_spawn_pos.clone()
    rows, cols = T._env_tile_row, T._env_tile_col
    # every tile is used; the spawn point is the tile centre + a jitter inside (half size - margin - 0.5); z ignores the ground
    assert sorted(set(zip(rows.tolist(), cols.tolist()))) == [(r, c) for r in range(2) for c in range(3)]
    ctr = T._tile_centers.to(sp.device)[rows, cols]
    assert float((sp[:, 0] - ctr[:, 0]).abs().max()) <= 3.0 - 1.5 + 1e-5 and float((sp[:, 1] - ctr[:, 1]).abs().max()) <= 4.0 - 1.5 + 1e-5
    assert float((sp[:, 0] - ctr[:, 0]).std()) > 0.5 and float((sp[:, 1] - ctr[:, 1]).std()) > 0.8
    assert torch.allclose(sp[:, 2], torch.full_like(sp[:, 2], 0.45 + 0.05)) and torch.equal(env.core.buf["qpos"][:, :3], sp)
    assert T.get_terrain_type_string(0) == cfgs[0]["terrain"]["subterrain_types"][int(rows[0])][int(cols[0])]
    zero = torch.zeros(B, env.num_actions, device=env.device)
    for _ in range(30):
        env.step(zero)
    # the manager is blind (reference: no height field to read from a generated terrain) - but the robots do stand on the field
    assert bool((env._ground_height == 0).all()) and bool((T.get_height_at_robot(env.base_pos[:, :2]) == 0).all())
    hs, vs = T.horizontal_scale, T.vertical_scale
    ix = ((env.base_pos[:, 0].cpu().numpy() + 6.0) / hs).round().astype(int)
    iy = ((env.base_pos[:, 1].cpu().numpy() + 12.0) / hs).round().astype(int)
    under = torch.as_tensor(T._height_field_np[ix, iy] * vs, device=env.device)
    on_stairs = (rows == 1) & (cols == 2) & (env.episode_length_buf >= 30)
    assert int(on_stairs.sum()) > 5 and float(under[on_stairs].min()) >= 0.02 - 1e-6
    clr = env.base_pos[:, 2] - under
    assert float(clr[on_stairs].min()) > 0.15 and float(clr[on_stairs].max()) < 0.6  # on top of the steps (2 .. 12 cm; spawn heights ignore the ground, so tall tiles would swallow the robot)
    # boundary: an env whose spawn point is moved > max_wander away is reset in the next tick and gets a new tile / spawn point
    T._env_spawn_pos[7, 0] = env.base_pos[7, 0] + T.max_wander + 0.3
    old = T._env_spawn_pos[7].clone()
    _, _, done, _ = env.step(zero)
    assert bool(done[7]) and not torch.equal(T._env_spawn_pos[7], old)


def test_respawn_on_a_tile(hostemu_lib):
    """go2_eval_walk.py:417-439 respawn_on_tile on a tile grid: centre of the chosen (clamped) tile, z from the (blind) height query."""
    from go2_b200 import eval as g2eval

    env = make_env("walk_terrain", _grid_cfgs(), 6, "hostemu", hostemu_lib, seed=4)
    env.reset()
    T = env.terrain
    g2eval.respawn(env, [1, 4], tile=(1, 7))  # column clamped to the last one
    q = env.core.buf["qpos"]
    want = T._tile_centers[1, 2]
    for e in (1, 4):
        assert float(q[e, 0]) == float(want[0]) and float(q[e, 1]) == float(want[1]) and abs(float(q[e, 2]) - (0.45 + 0.05)) < 1e-6
        assert torch.equal(T._env_spawn_pos[e], q[e, :3]) and int(T._env_tile_row[e]) == 1 and int(T._env_tile_col[e]) == 2
        assert T.get_terrain_type_string(e) == "stairs_terrain" and int(env.episode_length_buf[e]) == 0
    keep = T._env_spawn_pos[0].clone()
    g2eval.respawn(env, [0])  # tile=None: back to the centre of the env's own tile
    r, c = int(T._env_tile_row[0]), int(T._env_tile_col[0])
    assert torch.allclose(env.core.buf["qpos"][0, :2], T._tile_centers[r, c, :2]) and (T._tile_of(0, 6.0, 2)[0] == r)
    assert not torch.equal(keep, T._env_spawn_pos[0]) or torch.allclose(keep[:2], T._tile_centers[r, c, :2])
