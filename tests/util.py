"""Shared helpers for the parity tests: build an env on a backend, teacher-force it from a golden fixture."""
import json
import os

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests/golden")


def load_golden(name):
    data = dict(np.load(os.path.join(GOLDEN, name + ".npz")))
    with open(os.path.join(GOLDEN, name + "_cfg.json")) as f:
        cfgs = json.load(f)
    return data, cfgs


def make_env(kind, cfgs, B, backend, hostemu_lib=None, **kw):
    from go2_b200.envs import Go2EnvBaseFlavour, Go2EnvStair, Go2EnvWalk

    cls = {"walk": Go2EnvWalk, "stair": Go2EnvStair}.get(kind, Go2EnvBaseFlavour)
    cfgs = json.loads(json.dumps(cfgs))
    if backend == "hostemu":
        return cls(B, *cfgs, device="cpu", lib_path=hostemu_lib, **kw)
    return cls(B, *cfgs, device="cuda:0", **kw)


def _t(env, a, dtype=torch.float32):
    return torch.as_tensor(np.ascontiguousarray(a), device=env.device).to(dtype)


def force_state(env, g, t, kind):
    """Overwrite the env's persistent state with the fixture's pre-step state at step t."""
    b = env.core.buf
    B = env.num_envs
    b["qpos"].copy_(_t(env, g["qpos"][t]))
    b["qvel"].copy_(_t(env, g["qvel"][t]))
    b["qacc_ws"].copy_(_t(env, g["qacc_ws"][t]))
    b["warm"].copy_(_t(env, g["warm"][t][:, None], torch.int32))
    cache = g["cache"][t]  # [B, npairs, 3]
    ncache = np.zeros((B, 1), dtype=np.int32)
    cpair = np.zeros(tuple(b["cache_pair"].shape), dtype=np.int32)
    cn = np.zeros(tuple(b["cache_n"].shape), dtype=np.float32)
    for e in range(B):
        nz = np.nonzero(np.abs(cache[e]).sum(1) > 0)[0]
        ncache[e, 0] = len(nz)
        for k, p in enumerate(nz):
            cpair[e, k] = p
            cn[e, 3 * k:3 * k + 3] = cache[e, p]
    b["ncache"].copy_(_t(env, ncache, torch.int32))
    b["cache_pair"].copy_(_t(env, cpair, torch.int32))
    b["cache_n"].copy_(_t(env, cn))
    for name, key in (("last_actions", "last_actions"), ("last_dof_vel", "last_dof_vel"), ("dof_pos", "dof_pos"), ("dof_vel", "dof_vel"),
                      ("commands", "commands"), ("base_pos", "base_pos"), ("base_quat", "base_quat"), ("base_lin_vel", "base_lin_vel"),
                      ("base_ang_vel", "base_ang_vel"), ("proj_g", "proj_g")):
        b[name].copy_(_t(env, g[key][t]))
    b["ep_len"].copy_(_t(env, g["ep_len"][t][:, None], torch.int32))
    if g["ep_sums"].shape[1] > 0:
        b["ep_sums"].copy_(_t(env, g["ep_sums"][t].T))
    if kind == "stair":
        b["terrain_row"].copy_(_t(env, g["terrain_row"][t][:, None], torch.int32))
        b["last_x"].copy_(_t(env, g["last_x"][t][:, None]))
        env._lock_terrain_rows = True
    if kind in ("walk", "stair"):
        b["hist"].copy_(_t(env, g["hist"][t].reshape(B, -1)))
        env._action_history_write_idx = int(g["hist_idx"][t])
        b["applied"].copy_(_t(env, g["applied"][t]))
        b["delay"].copy_(_t(env, g["delay"][t][:, None], torch.int32))
        b["air_time"].copy_(_t(env, g["air_time"][t]))
        b["foot_contact"].copy_(_t(env, g["foot_contact"][t], torch.uint8))


def close(a, b, atol, rtol):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return np.abs(a - b) <= atol + rtol * np.abs(b)


def np_(x):
    return x.detach().cpu().numpy()
