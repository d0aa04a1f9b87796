"""Stair terrain-row assignment (go2_env_stair.py:809-854).  The reference splits the n envs reset in one tick into int(0.4 n) frontier,
int(0.3 n) near-frontier and the remaining easy rows and shuffles them, so what a reset env gets depends on HOW MANY envs reset in
that tick: alone (n = 1) or in a pair (n = 2) it always lands on an easy row.  The kernel draws per env with exactly the
probabilities int(0.4 n) / n, int(0.3 n) / n, rest (csrc/go2_env.cuh: stair_pick_row); n is only known after the step kernel, so the
post-pass redraws the rows of the tick's reset envs and redoes what the spawn point feeds (post_stair_rows)."""
import json
import os

import pytest
import torch

from util import ROOT, make_env

BACKENDS = ["hostemu", pytest.param("cuda", marks=pytest.mark.gpu)]


def _env(hostemu_lib, B, seed, backend="hostemu"):
    cfgs = json.load(open(os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200/go2_b200/cfgs/stair.json")))
    env = make_env("stair", cfgs, B, backend, hostemu_lib, seed=seed)
    env.curriculum.level = 0.7  # frontier row = int(0.7 * 12) = 8: near = {6, 7}, easy = 0..5
    env.reset()
    return env


def _time_out(env, idx):
    env.episode_length_buf[torch.as_tensor(idx, device=env.device)] = env.max_episode_length  # these envs time out in the next tick


@pytest.mark.parametrize("backend", BACKENDS)
def test_row_split_follows_the_reset_count_of_the_tick(backend, hostemu_lib):
    B = 48
    env = _env(hostemu_lib, B, seed=4, backend=backend)
    rows0 = env._env_terrain_row.clone()
    # reset(): n = B, the large-n split (about 40 / 30 / 30)
    assert int((rows0 == 8).sum()) in range(10, 30) and int(((rows0 == 6) | (rows0 == 7)).sum()) in range(5, 25) and int(rows0.max()) == 8
    zero = torch.zeros(B, env.num_actions, device=env.device)
    easy_only = 0
    for trial in range(12):  # n = 1 and n = 2: int(0.4 n) = int(0.3 n) = 0 -> always an easy row
        idx = [(5 * trial) % B] if trial % 2 == 0 else [(7 * trial) % B, (7 * trial + 3) % B]
        _time_out(env, idx)
        _, _, done, _ = env.step(zero)
        assert sorted(torch.nonzero(done).flatten().tolist()) == sorted(idx)
        r = env._env_terrain_row[torch.as_tensor(idx, device=env.device)]
        assert int(r.max()) <= 5, (trial, r.tolist())
        easy_only += len(idx)
        # what the spawn point feeds was redone for the final row: qpos, base_pos, last_x, privileged-obs tail
        for e in idx:
            cx, cy, cz = env._terrain_row_centers[int(env._env_terrain_row[e])]
            q = env.core.buf["qpos"][e]
            assert abs(float(q[0]) - cx) < 1e-6 and abs(float(q[1]) - cy) < 1e-6
            assert torch.equal(env.base_pos[e], q[:3]) and float(env._last_base_pos_x[e]) == float(q[0])
            pr = env.privileged_obs_buf[e]
            k = env.num_obs + 55
            assert abs(float(pr[k]) - float(env._env_terrain_row[e]) / 12.0) < 1e-6
            scan = pr[k + 1:k + 1 + 77].reshape(11, 7)
            assert torch.isfinite(scan).all() and float(scan.max()) < 0.0 and float(scan.min()) > -1.5
    assert easy_only == 18
    # n = 10: int(4.0) = 4 / int(3.0) = 3 / 3 -> every class occurs over a few ticks
    seen = set()
    for trial in range(8):
        idx = [(trial + 4 * j) % B for j in range(10)]
        _time_out(env, idx)
        env.step(zero)
        seen |= set(env._env_terrain_row[torch.as_tensor(idx, device=env.device)].tolist())
    assert 8 in seen and (6 in seen or 7 in seen) and min(seen) <= 5


def test_locked_rows_and_replay_are_untouched(hostemu_lib):
    env = _env(hostemu_lib, 8, seed=9)
    env._lock_terrain_rows = True
    env._env_terrain_row[:] = 3
    _time_out(env, [2])
    env.step(torch.zeros(8, env.num_actions))
    assert env._env_terrain_row.tolist() == [3] * 8
