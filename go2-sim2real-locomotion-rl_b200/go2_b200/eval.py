"""Headless evaluation loop.

Replaces the viewer / keyboard loop of `go2_eval_walk.py`, `go2_eval_stairs.py`, `go2_eval_base.py` (no graphics
stack on this box): a scripted command schedule instead of key presses, the same per-step policy call
(`obs -> actions`, under `torch.no_grad()`), metrics written as JSON instead of being drawn.

    python -m go2_b200.eval --task walk --envs 1024 --steps 500 --ckpt logs/go2-walk/model_188000.pt --out eval.json
"""
from __future__ import annotations

import argparse
import json
import math
import os

import torch

from . import checkpoint
from .envs import Go2EnvBaseFlavour, Go2EnvStair, Go2EnvStairInfo, Go2EnvStairLidar, Go2EnvWalk

_ENVS = {"walk": Go2EnvWalk, "stair": Go2EnvStair, "stair_lidar": Go2EnvStairLidar, "stair_info": Go2EnvStairInfo, "crouch": Go2EnvBaseFlavour, "jump": Go2EnvBaseFlavour}

# training-time randomisation the eval scripts strip from the cfg before building the env (go2_eval_walk.py:519-545)
_EVAL_STRIP = (
    "friction_range", "kp_range", "kd_range", "kp_factor_range", "kd_factor_range", "obs_noise", "obs_noise_level", "action_noise_std",
    "push_interval_s", "push_force_range", "push_duration_s", "mass_shift_range", "com_shift_range", "leg_mass_shift_range",
    "dynamic_payload_range", "dynamic_payload_interval_s", "dynamic_payload_prob", "gravity_offset_range", "motor_strength_range",
    "init_pos_z_range", "init_euler_range", "min_delay_steps", "max_delay_steps",
)
GO2_APPROX_MASS = 15.0  # go2_eval_walk.py:89: the mass the gravity-perturbation force is scaled by
_OBS_NOISE_STD = {"ang_vel": 0.2, "gravity": 0.05, "commands": 0.0, "dof_pos": 0.01, "dof_vel": 1.5, "actions": 0.0}  # :61-68


def eval_cfgs(cfgs, terminate=False, kp=None, kd=None):
    """The cfg edits of the eval scripts (go2_eval_walk.py:506-591): training DR and the curriculum off, no privileged
    obs, no rewards, a 100 s episode and (unless `terminate`) the fall terminations disabled."""
    import copy

    env_cfg, obs_cfg, reward_cfg, command_cfg = (copy.deepcopy(c) for c in cfgs[:4])
    for k in _EVAL_STRIP:
        env_cfg.pop(k, None)
    if "curriculum" in env_cfg:
        env_cfg["curriculum"]["enabled"] = False
    if kp is not None and not env_cfg.get("pls_enable", False):
        env_cfg["kp"], env_cfg["kd"] = float(kp), float(kd if kd is not None else env_cfg["kd"])
    obs_cfg["num_privileged_obs"] = None
    env_cfg["episode_length_s"] = 100.0
    if not terminate:
        for k in ("roll", "pitch"):
            env_cfg[f"termination_if_{k}_greater_than"] = 1e9
        for k in ("z_vel", "y_vel"):
            if f"termination_if_{k}_greater_than" in env_cfg:
                env_cfg[f"termination_if_{k}_greater_than"] = 1e9
    reward_cfg["reward_scales"] = {}
    return env_cfg, obs_cfg, reward_cfg, command_cfg


def respawn(env, envs_idx, terrain_row=None, tile=None):
    """Put the given envs back on their spawn point with the default pose and cleared action buffers: `respawn_at_start` of
    go2_eval_stairs.py:314-361 (spawn = centre of the env's terrain row + base height) and the flat-ground branch of
    `respawn_on_tile` of go2_eval_walk.py:399-488 (flat ground, tile grid: centre of tile `tile` = (row, col), custom height map: map
    centre), for any set of envs.  `terrain_row` moves them to that stair row first."""
    ii = torch.as_tensor(envs_idx, device=env.device).long().reshape(-1)
    n = int(ii.numel())
    if n == 0:
        return
    spawn = env.base_init_pos.expand(n, 3).clone()
    if getattr(env, "_use_terrain", False):
        if terrain_row is not None:
            env._env_terrain_row[ii] = int(terrain_row)
        centers = torch.tensor(env._terrain_row_centers, device=env.device, dtype=torch.float32)
        spawn = centers[env._env_terrain_row[ii].long()].clone()
        spawn[:, 2] += env.base_init_pos[2]
    T = getattr(env, "terrain", None)
    if T is not None and T.enabled:
        xy = torch.zeros(n, 2, device=env.device)  # custom height map (go2_eval_walk.py:441-464): map centre, on top of the interpolated ground
        if getattr(T, "_tile_centers", None) is not None:  # tile grid (:417-439): centre of the (clamped) tile; tile=None keeps each env's tile
            if tile is not None:
                r = torch.full((n,), min(int(tile[0]), T.n_rows - 1), device=env.device, dtype=torch.long)
                c = torch.full((n,), min(int(tile[1]), T.n_cols - 1), device=env.device, dtype=torch.long)
            else:
                r, c = T._env_tile_row[ii], T._env_tile_col[ii]
            xy = T._tile_centers.to(env.device)[r, c, :2].clone()
        spawn = torch.cat([xy, (T._query_height_batch(xy[:, 0], xy[:, 1]) + float(env.base_init_pos[2]) + T.spawn_height_offset)[:, None]], 1)
        T._env_spawn_pos[ii] = spawn
    env.dof_pos[ii] = env.default_dof_pos
    env.dof_vel[ii] = 0.0
    env.robot.set_dofs_position(env.default_dof_pos.expand(n, 12), env.motors_dof_idx, zero_velocity=True, envs_idx=ii)
    env.base_pos[ii] = spawn
    env.base_quat[ii] = env.base_init_quat
    env.robot.set_pos(spawn, zero_velocity=False, envs_idx=ii)
    env.robot.set_quat(env.base_init_quat.expand(n, 4), zero_velocity=False, envs_idx=ii)
    env.robot.zero_all_dofs_velocity(ii)
    for name in ("last_actions", "last_dof_vel", "_applied_actions", "base_lin_vel", "base_ang_vel"):
        b = getattr(env, name, None)
        if b is not None:
            b[ii] = 0.0
    if getattr(env, "_action_history", None) is not None:
        env._action_history[ii] = 0.0
    env.episode_length_buf[ii] = 0
    if hasattr(env, "_last_base_pos_x"):
        env._last_base_pos_x[ii] = spawn[:, 0]


class EvalScenario:
    """The eval-time disturbance knobs of go2_eval_walk.py:31-112, batched over all envs (the script runs one env).

    ground / robot friction (:553-554), motor strength (:556-559), constant payload + gravity-perturbation force and
    timed random pushes on the base link (:603-632, :712-741), a payload that changes at a given step (:697-707), gravity
    observation offset + observation noise on the policy input only (:746-755), action noise (:765-770), an action delay
    line (:775-779); plus respawn of fallen envs (what the operator does by key press, go2_eval_walk.py:399-488)."""

    def __init__(self, ground_friction=None, robot_friction=None, motor_strength_scale=None, payload_mass=0.0, dynamic_payload=None,
                 push=None, gravity_perturb=None, gravity_obs_offset=None, obs_noise_level=0.0, obs_noise_std=None, action_noise_std=0.0,
                 action_delay_steps=0, respawn_on_fall=False, fall_height=0.15, force_link=1, seed=0):
        self.ground_friction, self.robot_friction, self.motor_strength_scale = ground_friction, robot_friction, motor_strength_scale
        self.payload_mass = float(payload_mass)
        self.dynamic_payload = dynamic_payload  # (step, mass)
        self.push = push  # dict(force_range=(lo, hi), duration_s=(lo, hi), interval_s=, z_scale=)
        self.gravity_perturb, self.gravity_obs_offset = gravity_perturb, gravity_obs_offset
        self.obs_noise_level, self.obs_noise_std = float(obs_noise_level), dict(obs_noise_std or _OBS_NOISE_STD)
        self.action_noise_std, self.action_delay_steps = float(action_noise_std), int(action_delay_steps)
        self.respawn_on_fall, self.fall_height, self.seed = bool(respawn_on_fall), float(fall_height), int(seed)
        self.force_link = int(force_link)  # index into env.robot.links; the eval scripts use links[1] (go2_eval_walk.py:606), 0 = the base
        self.respawns = 0

    def setup(self, env):
        B, dev = env.num_envs, env.device
        self.gen = torch.Generator(device="cpu").manual_seed(self.seed)
        if self.ground_friction is not None:
            env.ground.set_friction(self.ground_friction)
        if self.robot_friction is not None:
            env.robot.set_friction(self.robot_friction)
        if self.motor_strength_scale is not None and hasattr(env, "_motor_strength"):
            env._motor_strength[:] = self.motor_strength_scale
        self.base_link = env.robot.links[self.force_link].idx
        self.const_force = torch.zeros(B, 3, device=dev)
        self.const_force[:, 2] -= self.payload_mass * 9.81
        if self.gravity_perturb is not None:
            self.const_force[:, 0] += self.gravity_perturb[0] * GO2_APPROX_MASS
            self.const_force[:, 1] += self.gravity_perturb[1] * GO2_APPROX_MASS
        self.push_remaining = torch.zeros(B, dtype=torch.long, device=dev)
        self.push_force = torch.zeros(B, 3, device=dev)
        self.noise_vec = None
        self.delay_line = [torch.zeros(B, env.num_actions, device=dev) for _ in range(self.action_delay_steps + 1)]
        self.respawns = 0

    def _rand(self, lo, hi, n, dev):
        return (lo + (hi - lo) * torch.rand(n, generator=self.gen)).to(dev)

    def _forces(self, env, t):
        B, dev = env.num_envs, env.device
        if self.dynamic_payload is not None and t == int(self.dynamic_payload[0]):
            self.const_force[:, 2] += (self.payload_mass - float(self.dynamic_payload[1])) * 9.81
            self.payload_mass = float(self.dynamic_payload[1])
            env.robot.set_mass_shift([self.payload_mass], [0])
        total = self.const_force.clone()
        if self.push is not None:
            dt = env.dt
            every = max(1, int(self.push["interval_s"] / dt))
            d0, d1 = (max(1, int(s / dt)) for s in self.push["duration_s"])
            if t % every == 0:
                idle = self.push_remaining == 0
                mag = self._rand(*self.push["force_range"], B, dev)
                th = self._rand(0.0, 2.0 * math.pi, B, dev)
                f = torch.stack([torch.cos(th) * mag, torch.sin(th) * mag, float(self.push.get("z_scale", 0.0)) * mag], dim=1)
                dur = torch.randint(d0, d1 + 1, (B,), generator=self.gen).to(dev)
                self.push_force = torch.where(idle[:, None], f, self.push_force)
                self.push_remaining = torch.where(idle, dur, self.push_remaining)
            active = self.push_remaining > 0
            total += self.push_force * active[:, None]
            self.push_remaining -= active.long()
        if bool((total != 0).any()):
            env.scene.sim.rigid_solver.apply_links_external_force(force=total, links_idx=[self.base_link])

    def before_step(self, env, t, obs, policy):
        """Everything the eval loop does between two env.step calls; returns the actions to apply."""
        self._forces(env, t)
        x = obs.clone()
        if self.gravity_obs_offset is not None:
            x[:, 3:6] += torch.as_tensor(self.gravity_obs_offset, device=x.device, dtype=x.dtype)
        if self.obs_noise_level > 0:
            if self.noise_vec is None:
                sizes = [("ang_vel", 3), ("gravity", 3), ("commands", 3), ("dof_pos", 12), ("dof_vel", 12), ("actions", x.shape[1] - 33)]
                self.noise_vec = torch.cat([torch.full((n,), self.obs_noise_std.get(k, 0.0) * self.obs_noise_level) for k, n in sizes])[: x.shape[1]].to(x.device)
            x = x + torch.randn(x.shape, generator=self.gen).to(x.device) * self.noise_vec
        a = policy(x)
        if self.action_noise_std > 0:
            npos = getattr(env, "num_pos_actions", env.num_actions)
            a = a.clone()
            a[:, :npos] += torch.randn((a.shape[0], npos), generator=self.gen).to(a.device) * self.action_noise_std
        if self.action_delay_steps > 0:
            self.delay_line.append(a.clone())
            self.delay_line.pop(0)
            a = self.delay_line[0]
        return a

    @staticmethod
    def _ground_z(env):
        if getattr(env, "_use_terrain", False):
            return env._get_terrain_height(env.base_pos[:, 0], env.base_pos[:, 1])
        return 0.0

    def after_step(self, env):
        """Put fallen envs back on their spawn pose (the R key of the eval scripts); returns how many."""
        if not self.respawn_on_fall:
            return 0
        fallen = torch.nonzero(env.base_pos[:, 2] - self._ground_z(env) < self.fall_height).flatten()
        n = int(fallen.numel())
        if n:
            respawn(env, fallen)
            for line in self.delay_line:
                line[fallen] = 0.0
            self.respawns += n
        return n


def evaluate(env, policy, steps, command_schedule=None, terrain_row=None, scenario=None):
    """Roll `policy` for `steps` ticks.  `scenario`: an EvalScenario (disturbances of the eval scripts).  `command_schedule`: list of (first_step, [vx, vy, wz]) applied to every env
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
    if scenario is not None:
        scenario.setup(env)
    with torch.no_grad():
        for t in range(steps):
            while sched and sched[0][0] <= t:
                cur_cmd = torch.tensor(sched.pop(0)[1], device=dev, dtype=torch.float32)
            if cur_cmd is not None:
                env.commands[:] = cur_cmd
            act = scenario.before_step(env, t + 1, obs, policy) if scenario is not None else policy(obs)  # the scripts count steps from 1
            obs, rew, done, extras = env.step(act)
            if scenario is not None:
                scenario.after_step(env)
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
    if scenario is not None:
        out["respawns"] = int(scenario.respawns)
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
    ap.add_argument("--eval-cfg", action="store_true", help="apply the eval scripts' cfg edits (DR / curriculum / terminations off)")
    ap.add_argument("--ground-friction", type=float, default=None)
    ap.add_argument("--robot-friction", type=float, default=None)
    ap.add_argument("--payload", type=float, default=0.0, help="constant payload mass [kg] as a downward force on the base")
    ap.add_argument("--push", type=float, nargs=3, default=None, metavar=("FORCE", "INTERVAL_S", "DURATION_S"))
    ap.add_argument("--obs-noise", type=float, default=0.0, help="observation noise level (go2_eval_walk.py OBS_NOISE_LEVEL)")
    ap.add_argument("--action-noise", type=float, default=0.0)
    ap.add_argument("--action-delay", type=int, default=0)
    ap.add_argument("--respawn", action="store_true", help="respawn fallen envs")
    ap.add_argument("--device", default="cuda:0")
    ap.add_argument("--out", default=None)
    a = ap.parse_args(argv)
    if a.cfgs:
        cfgs = checkpoint.load_cfgs(a.cfgs)[:4]
    else:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "cfgs", a.task + ".json")) as f:
            cfgs = json.load(f)
    if a.eval_cfg:
        cfgs = eval_cfgs(cfgs)
    env = _ENVS[a.task](a.envs, *cfgs[:4], device=a.device)
    scenario = None
    if any(v not in (None, 0, 0.0, False) for v in (a.ground_friction, a.robot_friction, a.payload, a.push, a.obs_noise, a.action_noise, a.action_delay, a.respawn)):
        push = dict(force_range=(a.push[0], a.push[0]), interval_s=a.push[1], duration_s=(a.push[2], a.push[2])) if a.push else None
        scenario = EvalScenario(a.ground_friction, a.robot_friction, None, a.payload, None, push, obs_noise_level=a.obs_noise,
                                action_noise_std=a.action_noise, action_delay_steps=a.action_delay, respawn_on_fall=a.respawn)
    if a.ckpt:
        policy = checkpoint.load_actor(a.ckpt, device=env.device)
    else:
        zeros = torch.zeros((a.envs, env.num_actions), device=env.device)
        policy = lambda obs: zeros  # noqa: E731
    res = evaluate(env, policy, a.steps, [(0, a.cmd)] if a.cmd else None, a.terrain_row, scenario)
    txt = json.dumps(res, indent=1)
    print(txt)
    if a.out:
        with open(a.out, "w") as f:
            f.write(txt)
    return res


if __name__ == "__main__":
    main()
