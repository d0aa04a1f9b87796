"""Walk env on a custom height map (examples/locomotion/go2_env_Omni_kplearn_varied_terrain.py + terrain_manager.py, heightmap mode;
SURVEY 8(f)-4) with the kernel's own draws: spawn sampling, ground-relative base height, boundary reset.  Replayed-draw parity against
the unmodified reference classes: tests/test_golden_dr.py, fixture walk_terrain_c1."""
import json
import os
import sys

import numpy as np
import pytest
import torch

from util import ROOT, make_env

BACKENDS = ["hostemu", pytest.param("cuda", marks=pytest.mark.gpu)]


def _cfgs():
    cfgs = json.load(open(os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200/go2_b200/cfgs/walk.json")))
    rng = np.random.default_rng(3)
    hf = np.zeros((160, 120), dtype=np.float32)  # 8 m x 6 m at 5 cm: left half flat, right half a ramp with bumps
    hf[80:] = (np.linspace(0.0, 0.3, 80)[:, None] + 0.02 * rng.standard_normal((80, 120))).astype(np.float32)
    cfgs[0]["terrain"] = {"enabled": True, "height_field": hf.tolist(), "horizontal_scale": 0.05, "vertical_scale": 1.0,
                          "spawn_height_offset": 0.05, "boundary_margin": 1.0}
    cfgs[0]["base_init_pos"] = [0.0, 0.0, 0.45]
    return cfgs, hf


@pytest.mark.parametrize("backend", BACKENDS)
def test_spawn_ground_height_and_boundary_reset(backend, hostemu_lib):
    cfgs, hf = _cfgs()
    B = 64
    env = make_env("walk_terrain", cfgs, B, backend, hostemu_lib, seed=8)
    T = env.terrain
    assert T.enabled and T.use_custom_heightmap and T.max_wander == pytest.approx(min(8.0, 6.0) / 2 - 1.0)
    env.reset()
    sp = T._env_spawn_pos.clone()
    q = env.core.buf["qpos"]
    # spawn: uniform in the safe area (half size - margin - 0.5), on top of the interpolated ground
    assert float(sp[:, 0].abs().max()) <= 4.0 - 1.5 + 1e-6 and float(sp[:, 1].abs().max()) <= 3.0 - 1.5 + 1e-6
    assert float(sp[:, 0].std()) > 0.8 and float(sp[:, 1].std()) > 0.4  # not all at the centre
    ground = T._query_height_batch(sp[:, 0], sp[:, 1])
    assert torch.allclose(sp[:, 2], ground + 0.45 + 0.05, atol=1e-6) and torch.equal(q[:, :3], sp)
    assert float(ground.max()) > 0.05 and int((ground == 0.0).sum()) > 5  # both halves of the map are used
    # a few ticks: the ground height under the base follows the host-side bilinear lookup; robots stand on the map, not on z = 0
    zero = torch.zeros(B, env.num_actions, device=env.device)
    for _ in range(25):
        env.step(zero)
    gh = T.get_height_at_robot(env.base_pos[:, :2])
    assert torch.allclose(env._ground_height, gh, atol=1e-6)
    up = ground > 0.1
    assert int(up.sum()) > 3 and float((env.base_pos[up, 2] - gh[up]).min()) > 0.2  # standing ON the ramp
    # boundary: an env whose spawn point is moved > max_wander away is reset in the next tick and gets a new spawn point
    T._env_spawn_pos[5, 0] = env.base_pos[5, 0] + T.max_wander + 0.3
    old = T._env_spawn_pos[5].clone()
    _, _, done, _ = env.step(zero)
    assert bool(done[5]) and not torch.equal(T._env_spawn_pos[5], old)
    assert bool(T.check_boundary(env.base_pos[:, :2]).sum() == 0)


def test_dropin_modules_of_the_terrain_variant(hostemu_lib):
    sys.path.insert(0, os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200/dropin"))
    try:
        import go2_env_Omni_kplearn_varied_terrain as mod
        import go2_env_terrain as mod2
        import terrain_manager as tm
    finally:
        sys.path.pop(0)
    from go2_b200.envs import Go2EnvTerrain, Go2EnvWalkTerrain, TerrainManager

    assert mod.Go2Env is Go2EnvWalkTerrain and mod2.Go2Env is Go2EnvTerrain and tm.TerrainManager is TerrainManager
    g = TerrainManager({"enabled": True, "subterrain_types": [["flat_terrain"]]}, 4)  # grid mode: tests/test_terrain_grid.py
    assert not g.use_custom_heightmap and (g.n_rows, g.n_cols) == (1, 1) and g._flat_row_indices == [0] and g._height_field is None
    with pytest.raises(NotImplementedError):
        TerrainManager({"enabled": True, "subterrain_types": [["lunar_terrain"]]}, 4)
    assert not TerrainManager({}, 4).enabled


def test_respawn_on_the_height_map(hostemu_lib):
    from go2_b200 import eval as g2eval

    cfgs, hf = _cfgs()
    env = make_env("walk_terrain", cfgs, 4, "hostemu", hostemu_lib, seed=8)
    env.reset()
    g2eval.respawn(env, [2])
    want_z = float(env.terrain._query_height_batch(torch.zeros(1), torch.zeros(1))) + 0.45 + 0.05
    q = env.core.buf["qpos"][2]
    assert float(q[0]) == 0.0 and float(q[1]) == 0.0 and abs(float(q[2]) - want_z) < 1e-6
    assert torch.equal(env.terrain._env_spawn_pos[2], q[:3])
