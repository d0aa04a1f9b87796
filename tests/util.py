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
    t = cfgs[0].get("terrain") if isinstance(cfgs[0], dict) else None
    if isinstance(t, dict) and isinstance(t.get("height_field"), str) and t["height_field"].startswith("npz:"):
        t["height_field"] = data[t["height_field"][4:]].tolist()  # a custom height map lives in the npz (too bulky as JSON text)
    return data, cfgs


def make_env(kind, cfgs, B, backend, hostemu_lib=None, **kw):
    from go2_b200.envs import Go2EnvBaseFlavour, Go2EnvStair, Go2EnvStairInfo, Go2EnvStairLidar, Go2EnvTerrain, Go2EnvWalk, Go2EnvWalkTerrain

    cls = {"walk": Go2EnvWalk, "stair": Go2EnvStair, "stair_lidar": Go2EnvStairLidar, "stair_info": Go2EnvStairInfo, "walk_terrain": Go2EnvWalkTerrain, "terrain_env": Go2EnvTerrain}.get(kind, Go2EnvBaseFlavour)
    cfgs = json.loads(json.dumps(cfgs))
    t = cfgs[0].get("terrain") if isinstance(cfgs[0], dict) else None
    if isinstance(t, dict) and "np_seed" in t:  # a terrain generated with numpy's global generator (grid mode of the terrain manager)
        np.random.seed(int(t["np_seed"]))
    if backend == "hostemu":
        from hostemu_core import HostEmuCore

        return cls(B, *cfgs, device="cpu", lib_path=hostemu_lib, core_cls=HostEmuCore, **kw)
    return cls(B, *cfgs, device="cuda:0", fast_math=(backend == "cuda_fast"), **kw)


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
    if kind in ("stair", "stair_lidar", "stair_info"):
        b["terrain_row"].copy_(_t(env, g["terrain_row"][t][:, None], torch.int32))
        b["last_x"].copy_(_t(env, g["last_x"][t][:, None]))
        env._lock_terrain_rows = True
    if kind in ("walk", "stair", "stair_lidar", "stair_info", "walk_terrain", "terrain_env"):
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


# ---------------------------------------------------------------------------------------------------
# DR-on fixtures (oracle/gen_golden_dr.py): teacher forcing + replay of the reference's random draws
# ---------------------------------------------------------------------------------------------------
class Derived:
    """Env-level buffers the DR fixtures leave out because they follow exactly from the stored arrays (the generator
    asserts bit-equality with the reference env's own buffers at every tick): last_actions, the action-delay ring and the
    episode sums.  Advanced tick by tick with the FIXTURE's actions / resets / reward terms."""

    def __init__(self, g, kind, clip):
        T, B, A = g["actions"].shape
        self.kind, self.clip = kind, float(clip)
        self.last_actions = np.zeros((B, A), np.float32)
        self.ep_sums = np.zeros((B, g["o_rew_terms"].shape[2]), np.float32)
        self.hist = None

    def start(self, hist_slots):
        if self.kind != "base":
            self.hist = np.zeros((self.last_actions.shape[0], hist_slots, self.last_actions.shape[1]), np.float32)

    def advance(self, g, t):
        a = np.clip(g["actions"][t], -self.clip, self.clip).astype(np.float32)
        done = g["o_reset"][t].astype(bool)
        terms = g["o_rew_terms"][t]
        self.last_actions = a.copy()
        if self.kind != "base":
            self.hist[:, int(g["hist_idx"][t])] = a
            self.hist[done] = 0.0
            self.ep_sums = (self.ep_sums + terms).astype(np.float32)
            self.ep_sums[done] = 0.0
        else:
            self.ep_sums[done] = 0.0
            self.ep_sums = (self.ep_sums + terms).astype(np.float32)


def pre(g, key, t):
    """Pre-state array of tick t; t == T reads the state after the last tick."""
    return g[key][t] if t < g[key].shape[0] else g["final_" + key]


def force_state_dr(env, g, t, kind, der):
    """Overwrite the env's persistent state (device buffers, global DR in force, host counters) with the fixture's state
    before tick t, and queue the fixture's random draws / global-DR candidates of tick t for replay."""
    from go2_b200 import abi

    b = env.core.buf
    B = env.num_envs
    md = list(env.motors_dof_idx)
    qpos, qvel = g["qpos"][t], g["qvel"][t]
    b["qpos"].copy_(_t(env, qpos))
    b["qvel"].copy_(_t(env, qvel))
    b["qacc_ws"].copy_(_t(env, g["qacc_ws"][t]))
    b["warm"].copy_(_t(env, g["warm"][t][:, None], torch.int32))
    ids = g["cache_pair"][t].astype(np.int32)
    b["ncache"].copy_(_t(env, (ids >= 0).sum(1)[:, None], torch.int32))
    b["cache_pair"].copy_(_t(env, np.maximum(ids, 0), torch.int32))
    b["cache_n"].copy_(_t(env, g["cache_n"][t].reshape(B, -1)))
    b["last_actions"].copy_(_t(env, der.last_actions))
    b["last_dof_vel"].copy_(_t(env, qvel[:, md]))
    b["dof_vel"].copy_(_t(env, qvel[:, md]))
    b["dof_pos"].copy_(_t(env, qpos[:, [d + 1 for d in md]]))
    b["commands"].copy_(_t(env, g["commands"][t]))
    b["ep_len"].copy_(_t(env, g["ep_len"][t][:, None], torch.int32))
    if der.ep_sums.shape[1] > 0:
        b["ep_sums"].copy_(_t(env, der.ep_sums))
    if kind in ("walk", "stair", "stair_lidar", "stair_info", "walk_terrain", "terrain_env"):
        b["hist"].copy_(_t(env, der.hist.reshape(B, -1)))
        env._action_history_write_idx = int(g["hist_idx"][t])
        b["delay"].copy_(_t(env, g["delay"][t][:, None], torch.int32))
        b["air_time"].copy_(_t(env, g["air_time"][t]))
        b["foot_contact"].copy_(_t(env, g["foot_contact"][t], torch.uint8))
        for name in ("kp_factors", "kd_factors", "motor_strength", "grav_off"):
            b[name].copy_(_t(env, g[name][t]))
        if "push_force" in g:
            b["push_force"].copy_(_t(env, g["push_force"][t]))
            b["push_remain"].copy_(_t(env, g["push_remain"][t][:, None], torch.int32))
            env._push_counter = int(g["push_counter"][t])
        if float(g["level"][t]) != float(env.curriculum.level):
            env.curriculum.level = float(g["level"][t])
            env._apply_curriculum_level(force=True)
        # global physics DR in force + the values the reference's reset_idx draws in this tick (used if an env resets)
        gl = np.zeros(abi.GLOBALS_LEN, np.float32)
        gl[abi.G_FRIC_GROUND] = gl[abi.G_FRIC_ROBOT] = g["g_friction"][t]
        gl[abi.G_MASS_SHIFT] = g["g_mass"][t]
        gl[abi.G_COM_SHIFT:abi.G_COM_SHIFT + 3] = g["g_com"][t]
        for i, l in enumerate(env._hip_link_indices):
            gl[abi.G_MASS_SHIFT + l] = g["g_leg"][t][i]
        gl[abi.G_DR_COUNTER] = g["g_counter"][t]
        # the other parity slot = the values of the previous tick (what the previous closing FK ran on); an env reset in the
        # previous tick had its FK refreshed with the new values
        gp = gl.copy()
        if t > 0:
            gp[abi.G_MASS_SHIFT] = g["g_mass"][t - 1]
            gp[abi.G_COM_SHIFT:abi.G_COM_SHIFT + 3] = g["g_com"][t - 1]
            for i, l in enumerate(env._hip_link_indices):
                gp[abi.G_MASS_SHIFT + l] = g["g_leg"][t - 1][i]
        par = env.core.tick_parity()
        b["globals"].copy_(_t(env, np.stack([gl, gp] if par == 0 else [gp, gl])))
        b["fk_fresh"].copy_(_t(env, (g["o_reset"][t - 1] if t > 0 else np.ones(B))[:, None], torch.uint8))
        env._cand_override = {"friction": float(pre(g, "g_friction", t + 1)), "mass": float(pre(g, "g_mass", t + 1)),
                              "com": [float(x) for x in pre(g, "g_com", t + 1)], "leg": [float(x) for x in pre(g, "g_leg", t + 1)]}
    if kind in ("stair", "stair_lidar", "stair_info"):
        b["terrain_row"].copy_(_t(env, g["terrain_row"][t][:, None], torch.int32))
        b["last_x"].copy_(_t(env, g["last_x"][t][:, None]))
    if kind in ("walk_terrain", "terrain_env"):
        b["spawn_pos"].copy_(_t(env, g["spawn_pos"][t]))
    if kind == "stair_lidar":
        b["actor_scan"].copy_(_t(env, g["actor_scan"][t]))
        env._lidar_noise_level = float(g["lidar_noise_level"][t])
    if kind == "stair_info":
        b["actor_scan"].copy_(_t(env, g["actor_scan"][t]))
        env._stair_noise_level = float(g["lidar_noise_level"][t])
    env.replay_draws(g["replay"][t])
