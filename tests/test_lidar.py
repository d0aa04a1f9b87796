"""Lidar variant of the stair env (examples/locomotion/go2_env_stair_lidar.py) with the kernel's OWN random draws
(the replayed-draw parity against the unmodified reference env is tests/test_golden_dr.py, fixture stair_lidar_c1): the noise model's
event rates follow the configured probabilities x the DR level, the clean critic scan equals a direct heightfield lookup."""
import json
import os

import numpy as np
import pytest
import torch

from util import ROOT, make_env

BACKENDS = ["hostemu", pytest.param("cuda", marks=pytest.mark.gpu)]


@pytest.mark.parametrize("backend", BACKENDS)
def test_lidar_noise_rates_and_critic_scan(backend, hostemu_lib):
    cfgs = json.load(open(os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200/go2_b200/cfgs/stair_lidar.json")))
    cfgs[0]["curriculum"]["level_init"] = 1.0  # DR level 1: dropout 0.15, blackout 0.05, latency 0.10 per env-tick
    cfgs[0]["episode_length_s"] = 100.0
    B, T = 256, 30
    env = make_env("stair_lidar", cfgs, B, backend, hostemu_lib, seed=6)
    assert (env.num_obs, env.num_privileged_obs) == (64, 165) and env._lidar.num_actor_scan == 15 and env._lidar.num_critic_scan == 45
    assert env._lidar_noise_level == pytest.approx(1.0)
    env.reset()
    zero = torch.zeros(B, env.num_actions, device=env.device)
    n_cells = n_zero = n_black = n_lat = n_ticks = 0
    prev = env._actor_scan_buf.clone()
    for t in range(T):
        obs, _, done, extras = env.step(zero)
        scan = obs[:, 49:]
        assert torch.equal(scan, env._actor_scan_buf)
        live = ~done
        black = (scan == 0).all(1) & live
        lat = (scan == prev).all(1) & live & ~black & (t > 0)
        rest = live & ~black & ~lat
        n_black += int(black.sum()); n_lat += int(lat.sum()); n_ticks += int(live.sum())
        n_zero += int((scan[rest].abs() < 0.06).sum()); n_cells += int(rest.sum()) * 15  # dropped cell = 0 + the 1 cm vertical offset; the ground is >= 0.25 m below
        assert bool((scan[done] == 0).all()) and float(scan.abs().max()) <= 1.0 + 1e-6  # reset envs: zeroed; clip 1.0
        # clean critic scan = heightfield lookups at the 9 x 5 grid rotated by the yaw (pose AFTER the tick, envs not reset)
        crit = extras["observations"]["critic"][:, 64 + 56:]
        q, p = env.base_quat, env.base_pos
        yaw = torch.atan2(2 * (q[:, 0] * q[:, 3] + q[:, 1] * q[:, 2]), 1 - 2 * (q[:, 2] ** 2 + q[:, 3] ** 2))
        lx = torch.as_tensor(env._lidar.critic_local_x, device=env.device)[None]
        ly = torch.as_tensor(env._lidar.critic_local_y, device=env.device)[None]
        wx = p[:, 0:1] + torch.cos(yaw)[:, None] * lx - torch.sin(yaw)[:, None] * ly
        wy = p[:, 1:2] + torch.sin(yaw)[:, None] * lx + torch.cos(yaw)[:, None] * ly
        want = (env._get_terrain_height(wx, wy) - p[:, 2:3]).clamp(-1.0, 1.0)
        ok = (crit - want).abs() < 1e-5
        assert float(ok[live].float().mean()) > 0.995  # (a sample point within rounding of a cell edge may pick the neighbour cell)
        assert bool((crit[done] == 0).all())
        prev = scan.clone()
    assert 0.12 < n_zero / n_cells < 0.18, n_zero / n_cells          # per-cell dropout 0.15
    assert 0.03 < n_black / n_ticks < 0.075, n_black / n_ticks       # blackout 0.05
    assert 0.06 < n_lat / n_ticks < 0.13, n_lat / n_ticks            # latency 0.10 (of the envs not blacked out)
