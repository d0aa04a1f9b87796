"""Headless evaluation loop.

Replaces the viewer / keyboard loop of `go2_eval_walk.py`, `go2_eval_stairs.py`, `go2_eval_base.py` (no graphics
stack on this box): a scripted command schedule instead of key presses, the same per-step policy call
(`obs -> actions`, under `torch.no_grad()`), metrics written as JSON instead of being drawn.

    python -m go2_b200.eval --task walk --envs 1024 --steps 500 --ckpt logs/go2-walk/model_188000.pt --out eval.json
"""
from __future__ import annotations

import argparse
import json
import os

import torch

from . import checkpoint
from .envs import Go2EnvBaseFlavour, Go2EnvStair, Go2EnvWalk

_ENVS = {"walk": Go2EnvWalk, "stair": Go2EnvStair, "crouch": Go2EnvBaseFlavour, "jump": Go2EnvBaseFlavour}


def evaluate(env, policy, steps, command_schedule=None, terrain_row=None):
    """Roll `policy` for `steps` ticks.  `command_schedule`: list of (first_step, [vx, vy, wz]) applied to every env
    (like the eval scripts' keyboard overrides, which also disable command resampling by rewriting `env.commands`
    every step); `terrain_row`: stair only - lock every env on that difficulty row (go2_eval_stairs.py)."""
    if terrain_row is not None and hasattr(env, "_env_terrain_row"):
        env._env_terrain_row[:] = int(terrain_row)
        env._lock_terrain_rows = True
    obs, _ = env.reset()
    B, dev = env.num_envs, env.device
    sched = sorted(command_schedule or [], key=lambda x: x[0])
    cur_cmd = None
    n_done = torch.zeros((), device=dev)
    n_timeout = torch.zeros((), device=dev)
    rew_sum = torch.zeros(B, device=dev)
    err_lin = torch.zeros((), device=dev)
    err_ang = torch.zeros((), device=dev)
    x0 = env.base_pos[:, 0].clone()
    with torch.no_grad():
        for t in range(steps):
            while sched and sched[0][0] <= t:
                cur_cmd = torch.tensor(sched.pop(0)[1], device=dev, dtype=torch.float32)
            if cur_cmd is not None:
                env.commands[:] = cur_cmd
            obs, rew, done, extras = env.step(policy(obs))
            rew_sum += rew
            n_done += done.sum()
            n_timeout += extras["time_outs"].sum()
            err_lin += torch.linalg.norm(env.commands[:, :2] - env.base_lin_vel[:, :2], dim=1).mean()
            err_ang += (env.commands[:, 2] - env.base_ang_vel[:, 2]).abs().mean()
    out = {
        "envs": B, "steps": steps, "episodes_ended": int(n_done), "time_outs": int(n_timeout), "falls": int(n_done - n_timeout),
        "mean_reward_per_step": float(rew_sum.mean() / steps), "mean_lin_vel_error": float(err_lin / steps),
        "mean_ang_vel_error": float(err_ang / steps), "mean_base_height": float(env.base_pos[:, 2].mean()),
        "mean_forward_distance": float((env.base_pos[:, 0] - x0).mean()),
    }
    if hasattr(env, "_env_terrain_row"):
        out["mean_terrain_row"] = float(env._env_terrain_row.float().mean())
    return out


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--task", default="walk", choices=sorted(_ENVS))
    ap.add_argument("--envs", type=int, default=1024)
    ap.add_argument("--steps", type=int, default=500)
    ap.add_argument("--cfgs", default=None, help="cfgs.pkl of a training run (default: the task's shipped cfg table)")
    ap.add_argument("--ckpt", default=None, help="rsl-rl model_<iter>.pt (default: zero actions = stand at the default pose)")
    ap.add_argument("--cmd", type=float, nargs=3, default=None, help="fixed command vx vy wz")
    ap.add_argument("--terrain-row", type=int, default=None)
    ap.add_argument("--device", default="cuda:0")
    ap.add_argument("--out", default=None)
    a = ap.parse_args(argv)
    if a.cfgs:
        cfgs = checkpoint.load_cfgs(a.cfgs)[:4]
    else:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "cfgs", a.task + ".json")) as f:
            cfgs = json.load(f)
    env = _ENVS[a.task](a.envs, *cfgs, device=a.device)
    if a.ckpt:
        policy = checkpoint.load_actor(a.ckpt, device=env.device)
    else:
        zeros = torch.zeros((a.envs, env.num_actions), device=env.device)
        policy = lambda obs: zeros  # noqa: E731
    res = evaluate(env, policy, a.steps, [(0, a.cmd)] if a.cmd else None, a.terrain_row)
    txt = json.dumps(res, indent=1)
    print(txt)
    if a.out:
        with open(a.out, "w") as f:
            f.write(txt)
    return res


if __name__ == "__main__":
    main()
