"""Stair-info variant of the stair env (examples/locomotion/go2_env_stair_lidar_2.py) with the kernel's OWN random draws
(the replayed-draw parity against the unmodified reference env is tests/test_golden_dr.py, fixture stair_info_c1): the clean copy in
the privileged obs equals an edge detection done here with torch on the same heightfield, the noise model's event rates follow the
configured probabilities x the DR level, the variable-depth terrain equals the reference builder's output, body_contact counts links."""
import hashlib
import json
import os

import numpy as np
import pytest
import torch

from util import ROOT, make_env

BACKENDS = ["hostemu", pytest.param("cuda", marks=pytest.mark.gpu)]
# sha256 of the int16 heightfield the REFERENCE's build_stair_terrain (go2_env_stair_lidar_2.py:83) returns for the shipped
# go2_stair_train_lidar_2 terrain cfg after np.random.seed(0) / (7) - generated in the build container, where /root/reference exists
HF_SHA = {0: "e1bbb449008ddeeb2f08b73b68e744c0c2a2291d53ea2e4b682512f41240448e", 7: "7f9b46b1be1e32911e0a623d8165000548fda3d7629629605205ac6001416bbf"}


def _cfgs():
    return json.load(open(os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200/go2_b200/cfgs/stair_info.json")))


def test_variable_depth_terrain_matches_the_reference_builder():
    from go2_b200.terrain import build_stair_terrain_variable_depth

    tc = _cfgs()[0]["terrain"]
    for seed, sha in HF_SHA.items():
        np.random.seed(seed)
        hf, info = build_stair_terrain_variable_depth(tc)
        assert hf.dtype == np.int16 and hashlib.sha256(np.ascontiguousarray(hf).tobytes()).hexdigest() == sha
        assert sorted(info["step_depths_m"]) == pytest.approx(np.linspace(0.22, 0.30, 13).tolist())
    # explicit depths (extension used by the fixture): same layout as the shuffle that produced them, no RNG touched
    st = np.random.get_state()[1].copy()
    hf2, info2 = build_stair_terrain_variable_depth(dict(tc, step_depths_m=info["step_depths_m"]))
    assert np.array_equal(hf2, hf) and np.array_equal(np.random.get_state()[1], st)
    # every row keeps one length: a short ascent is padded at platform height, a short descent at ground level
    width = int(round(tc["row_width_m"] / tc["horizontal_scale"]))
    for r in range(13):
        col = hf[:, r * width]
        assert col.max() == 6 * round(info["step_heights_m"][r] / tc["vertical_scale"]) and col[-1] == 0 and col[0] == 0


@pytest.mark.parametrize("backend", BACKENDS)
def test_stair_info_clean_copy_noise_rates_and_body_contact(backend, hostemu_lib):
    cfgs = _cfgs()
    cfgs[0]["curriculum"]["level_init"] = 1.0   # DR level 1: flip 0.03, blackout 0.05, latency 0.10 per env-tick
    cfgs[0]["episode_length_s"] = 100.0
    cfgs[0]["terrain"]["flat_before_m"] = 0.5   # spawn 0.25 m before the first riser: the detector (range 0.27 m) sees an edge
    cfgs[0]["body_contact_names"] = ["FR_thigh", "FL_thigh", "RR_calf", "RL_calf"]
    B, T = 256, 30
    np.random.seed(5)
    env = make_env("stair_info", cfgs, B, backend, hostemu_lib, seed=6, export_reward_terms=True)
    S = env._stair_cfg
    assert (env.num_obs, env.num_privileged_obs) == (53, 113) and env._stair_noise_level == pytest.approx(1.0)
    env.reset()
    dev = env.device
    zero = torch.zeros(B, env.num_actions, device=dev)
    hts, dps = torch.tensor(env._terrain_step_heights, device=dev).float(), torch.tensor(env._terrain_step_depths, device=dev).float()
    sx = S.edge_sample_x.to(dev)
    mid = (sx[:-1] + sx[1:]) / 2
    n_live = n_black = n_lat = n_flip = n_dir = n_edge = n_cand = 0
    prev = env._stair_info_buf.clone()
    i_bc = list(env.reward_scales).index("body_contact")
    for t in range(T):
        rows_before = env._env_terrain_row.clone()
        obs, _, done, extras = env.step(zero)
        info, live = obs[:, 49:], ~done
        assert torch.equal(info, env._stair_info_buf) and bool((info[done] == 0).all())
        # ---- clean copy = edge detection on the pose after the tick (envs not reset)
        clean = extras["observations"]["critic"][:, 53 + 56:]
        assert torch.equal(clean, env._stair_info_clean) and bool((clean[done] == 0).all())
        q, p = env.base_quat, env.base_pos
        yaw = torch.atan2(2 * (q[:, 0] * q[:, 3] + q[:, 1] * q[:, 2]), 1 - 2 * (q[:, 2] ** 2 + q[:, 3] ** 2))
        h = env._get_terrain_height(p[:, 0:1] + torch.cos(yaw)[:, None] * sx[None], p[:, 1:2] + torch.sin(yaw)[:, None] * sx[None])
        dh = h[:, 1:] - h[:, :-1]
        sig = dh.abs() > S.edge_threshold
        has, first = sig.any(1), sig.int().argmax(1)
        want = torch.stack([torch.where(has, mid[first], torch.full_like(mid[first], S.edge_max_range)) * S.edge_dist_scale,
                            hts[rows_before] * S.height_scale, dps[rows_before] * S.depth_scale,
                            torch.where(has, torch.sign(dh[torch.arange(B, device=dev), first]), torch.zeros(B, device=dev)) * S.direction_scale], 1)
        ok = ((clean - want).abs() < 1e-5).all(1)
        assert float(ok[live].float().mean()) > 0.99  # (a sample within rounding of a cell edge may pick the neighbour cell)
        # ---- noisy copy: riser height / tread depth within 5 sigma of the row's values, edge distance inside [0, max range]
        assert bool(((info[:, 1] / S.height_scale - hts[rows_before]).abs()[live] < 5 * S.height_noise_std).all())
        assert bool(((info[:, 2] / S.depth_scale - dps[rows_before]).abs()[live] < 5 * S.depth_noise_std).all())
        assert float(info[:, 0].min()) >= 0.0 and float(info[:, 0].max()) <= S.edge_max_range * S.edge_dist_scale + 1e-6
        black = live & (info[:, 0] == np.float32(S.edge_max_range) * np.float32(S.edge_dist_scale)) & (info[:, 3] == 0) & (clean[:, 3] != 0)
        # latency replays the previous edge distance: recognisable where that one was strictly inside (0, max range) (a noisy
        # distance clamped to an end of the range repeats by itself)
        maxs = np.float32(S.edge_max_range) * np.float32(S.edge_dist_scale)
        cand = live & (prev[:, 0] > 0) & (prev[:, 0] < maxs) & (t > 0)
        lat = cand & (info[:, 0] == prev[:, 0]) & (info[:, 3] == prev[:, 3])
        rest = live & ~black & ~lat & (clean[:, 3] != 0)
        n_dir += int((live & (clean[:, 3] != 0)).sum()); n_black += int(black.sum())
        n_live += int(live.sum()); n_lat += int(lat.sum()); n_cand += int(cand.sum())
        n_flip += int((info[rest, 3] == -clean[rest, 3]).sum()); n_edge += int(rest.sum())
        # ---- body_contact = number of listed links with |net contact force| > threshold, times scale * dt
        cf = env.robot.get_links_net_contact_force()
        cnt = sum((cf[:, l].norm(dim=1) > env._body_contact_threshold).float() for l in env._body_contact_link_indices)
        got = env.reward_terms[:, i_bc]
        assert torch.allclose(got[live], (cnt * env.reward_scales["body_contact"])[live], atol=1e-6)
        prev = info.clone()
    assert n_dir > 0.3 * n_live                                  # the detector does see the riser in this layout
    assert 0.03 < n_black / n_dir < 0.075, n_black / n_dir       # blackout 0.05 (observable only where the clean direction is non-zero)
    assert 0.06 < n_lat / n_cand < 0.14, n_lat / n_cand          # latency 0.10 (x 0.95 not blacked out)
    assert 0.015 < n_flip / n_edge < 0.05, n_flip / n_edge       # direction flip 0.03
