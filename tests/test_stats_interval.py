"""`stats_sync_interval`: k = 1 reads the device-side episode statistics every tick (the reference's timing), k > 1 reads the
accumulated sums asynchronously every k ticks.  The sums must be the same numbers, only later: with the curriculum's effect on the
physics switched off both envs run the same trajectory, and everything `_after_resets` ever receives adds up to the same totals."""
import json
import os

import numpy as np
import torch

from go2_b200 import abi
from util import ROOT, make_env


def _run(hostemu_lib, k, ticks=40):
    cfgs = json.load(open(os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200/go2_b200/cfgs/walk.json")))
    cfgs[0]["curriculum"]["enabled"] = False  # the level must not feed back into DR ranges: identical physics for every k
    cfgs[0]["episode_length_s"] = 0.2  # 10 ticks: time-outs every episode, plus falls from the large actions
    env = make_env("walk", cfgs, 16, "hostemu", hostemu_lib, seed=3, stats_sync_interval=k)
    seen, calls = np.zeros(env.core.L.stats_len, dtype=np.float64), []
    orig = env._after_resets

    def spy(stats):
        seen[:] += np.asarray(stats, dtype=np.float64)
        calls.append(env._step_count)
        return orig(stats)

    env._after_resets = spy
    env.reset()
    gen = torch.Generator().manual_seed(0)
    for _ in range(ticks):
        env.step(2.0 * torch.randn(16, env.num_actions, generator=gen))
    env.flush_stats()
    return env, seen, calls


def test_interval_reads_deliver_the_same_sums_later(hostemu_lib):
    e1, s1, c1 = _run(hostemu_lib, 1)
    e4, s4, c4 = _run(hostemu_lib, 4)
    assert torch.equal(e1.core.buf["qpos"], e4.core.buf["qpos"]) and torch.equal(e1.obs_buf, e4.obs_buf)  # same trajectory
    assert s1[abi.STATS_RESETS] >= 16 * 3 and s1[abi.STATS_TIMEOUTS] > 0
    assert s1[abi.STATS_RESETS] == s4[abi.STATS_RESETS] and s1[abi.STATS_TIMEOUTS] == s4[abi.STATS_TIMEOUTS]  # integer counts: exact
    np.testing.assert_allclose(s4, s1, rtol=1e-5, atol=1e-5)  # float sums: same terms, other summation order
    # k = 4 consumed them in fewer, later calls (every 4th tick, one interval behind, rest at flush)
    mid = c4[1:-2]
    assert len(c4) < len(c1) and all(b - a == 4 for a, b in zip(mid, mid[1:]))


def test_error_bits_reach_the_host_through_the_statistics(hostemu_lib):
    """Genesis polls the solver's errno every 10 substeps and raises (rigid_solver.py:1189-1213); here the statistics row counts the
    env-ticks that raised an error bit and the errno buffer is polled when that count is non-zero."""
    import pytest

    cfgs = json.load(open(os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200/go2_b200/cfgs/walk.json")))
    env = make_env("walk", cfgs, 8, "hostemu", hostemu_lib, seed=3)
    env.reset()
    env.step(torch.zeros(8, env.num_actions))
    assert "errno_env_ticks" not in env.extras
    stats = np.zeros(env.core.L.stats_len, dtype=np.float32)
    stats[abi.STATS_ERR] = 2
    env.core.buf["errno"][3] = 8  # NaN acceleration: the env was force-reset, reported but not raised
    env._consume_stats(stats)
    assert env.extras["errno_env_ticks"] == 2
    env.core.buf["errno"][5] = 2  # contact overflow: raises the reference's message
    with pytest.raises(RuntimeError, match="max number of contact pairs"):
        env._consume_stats(stats)
