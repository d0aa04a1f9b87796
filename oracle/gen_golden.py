"""TEST INFRASTRUCTURE ONLY - generates tests/golden/*.npz by running the UNMODIFIED reference env classes
(/root/reference/examples/locomotion/final/go2_env_{walk,base}.py) on top of the oracle-backed `genesis`
shim (oracle/genesis_shim).  Runs only in the build container (needs /root/reference); the fixtures travel.

For every step the fixture holds the env + solver state BEFORE the step (so that another implementation
can be teacher-forced from it), the action, and the outputs AFTER the step.

    python oracle/gen_golden.py            # writes tests/golden/walk_det.npz, crouch.npz, jump.npz + cfg json
"""
import ast
import contextlib
import io
import json
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("GO2_REFERENCE_ROOT", "/root/reference")
FINAL = os.path.join(REF, "examples/locomotion/final")
sys.path[:0] = [os.path.join(ROOT, "oracle/genesis_shim"), os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200"), FINAL, ROOT]
OUT = os.path.join(ROOT, "tests/golden")

DR_KEYS = ["friction_range", "kp_factor_range", "kd_factor_range", "obs_noise", "obs_noise_level", "action_noise_std",
           "push_interval_s", "push_force_range", "push_duration_s", "init_pos_z_range", "init_euler_range", "mass_shift_range",
           "com_shift_range", "leg_mass_shift_range", "gravity_offset_range", "motor_strength_range"]


def load_funcs(path, names):
    """exec only the named top-level functions of a reference script (its imports need rsl-rl)."""
    tree = ast.parse(open(path).read())
    body = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in names]
    ns = {}
    exec(compile(ast.Module(body=body, type_ignores=[]), path, "exec"), ns)
    return types.SimpleNamespace(**{k: ns[k] for k in names})


def walk_det_cfgs():
    """go2_train_walk.get_cfgs() with every random draw removed or collapsed to a point."""
    env_cfg, obs_cfg, reward_cfg, command_cfg = load_funcs(os.path.join(FINAL, "go2_train_walk.py"), ["get_cfgs"]).get_cfgs()
    for k in DR_KEYS:
        env_cfg.pop(k, None)
    command_cfg.update(lin_vel_x_range=[0.5, 0.5], lin_vel_y_range=[0.1, 0.1], ang_vel_range=[0.2, 0.2])
    return env_cfg, obs_cfg, reward_cfg, command_cfg


def walk_short_cfgs():
    """Same, with short episodes and tight tilt limits so that time-outs and falls (resets) occur often."""
    env_cfg, obs_cfg, reward_cfg, command_cfg = walk_det_cfgs()
    env_cfg.update(episode_length_s=1.0, termination_if_roll_greater_than=8, termination_if_pitch_greater_than=8, resampling_time_s=0.4)
    return env_cfg, obs_cfg, reward_cfg, command_cfg


def stair_det_cfgs():
    """go2_train_stair.get_cfgs() with the random draws removed (terrain, curriculum level and rewards kept)."""
    env_cfg, obs_cfg, reward_cfg, command_cfg = load_funcs(os.path.join(FINAL, "go2_train_stair.py"), ["get_cfgs"]).get_cfgs()
    for k in DR_KEYS:
        env_cfg.pop(k, None)
    env_cfg.update(episode_length_s=1.6, resampling_time_s=0.6)
    command_cfg.update(lin_vel_x_range=[0.5, 0.5], lin_vel_y_range=[0.0, 0.0], ang_vel_range=[0.0, 0.0])
    return env_cfg, obs_cfg, reward_cfg, command_cfg


def stair_place(env):
    """One env per difficulty row (rows locked), robots dropped at staggered x over the first flight of stairs so
    that feet, calves and the trunk meet step edges of every height."""
    B = env.num_envs
    env._env_terrain_row[:] = torch.arange(B) % env._num_terrain_rows
    env._lock_terrain_rows = True
    env.reset()
    pos = env.base_pos.clone()
    pos[:, 0] = 1.6 + 0.37 * torch.arange(B, dtype=torch.float32)
    pos[:, 2] = env._get_terrain_height(pos[:, 0], pos[:, 1]) + 0.40
    idx = torch.arange(B)
    env.base_pos[:] = pos
    env.robot.set_pos(pos, zero_velocity=False, envs_idx=idx)
    env._last_base_pos_x[:] = pos[:, 0]


def base_cfgs(script):
    env_cfg, obs_cfg, reward_cfg, command_cfg = load_funcs(os.path.join(FINAL, script), ["get_cfgs"]).get_cfgs()
    return env_cfg, obs_cfg, reward_cfg, command_cfg


def snapshot(env, kind):
    r = env.scene._rigid
    qws, warm, cache = r.get_warm_state()
    s = dict(qpos=r.get_qpos(), qvel=r.get_qvel(), qacc_ws=qws, warm=warm, cache=cache,
             last_actions=env.last_actions.numpy().copy(), last_dof_vel=env.last_dof_vel.numpy().copy(),
             dof_pos=env.dof_pos.numpy().copy(), dof_vel=env.dof_vel.numpy().copy(), commands=env.commands.numpy().copy(),
             ep_len=env.episode_length_buf.numpy().copy(), base_pos=env.base_pos.numpy().copy(), base_quat=env.base_quat.numpy().copy(),
             base_lin_vel=env.base_lin_vel.numpy().copy(), base_ang_vel=env.base_ang_vel.numpy().copy(),
             proj_g=env.projected_gravity.numpy().copy(),
             ep_sums=np.stack([v.numpy().copy() for v in env.episode_sums.values()]) if env.episode_sums else np.zeros((0, env.num_envs)))
    if kind in ("walk", "stair"):
        s.update(hist=env._action_history.numpy().copy(), hist_idx=np.array(env._action_history_write_idx),
                 applied=env._applied_actions.numpy().copy(), delay=env._delay_steps.numpy().copy(),
                 air_time=env._feet_air_time.numpy().copy(), foot_contact=env._foot_contact.numpy().copy())
    if kind == "stair":
        s.update(terrain_row=env._env_terrain_row.numpy().astype(np.int32), last_x=env._last_base_pos_x.numpy().copy())
    return s


def outputs(env, obs, rew, done, extras, kind):
    r = env.scene._rigid
    o = dict(o_obs=obs.numpy().copy(), o_rew=rew.numpy().copy(), o_reset=done.numpy().astype(np.uint8),
             o_timeout=extras["time_outs"].numpy().copy(), o_qpos=r.get_qpos(), o_qvel=r.get_qvel(),
             o_base_pos=env.base_pos.numpy().copy(), o_commands=env.commands.numpy().copy(), o_ep_len=env.episode_length_buf.numpy().copy(),
             o_ncon=r.get_solver_stats()[:, 0].copy())
    if kind in ("walk", "stair"):
        o.update(o_priv=extras["observations"]["critic"].numpy().copy(), o_foot_contact=env._foot_contact.numpy().copy(),
                 o_air_time=env._feet_air_time.numpy().copy())
    return o


def run(env_cls, cfgs, kind, B, T, act_std, seed, act_dim, after_reset=None):
    cfgs_json = json.loads(json.dumps(cfgs))  # pristine copy (the ctor multiplies reward scales by dt in place)
    torch.manual_seed(1)
    with contextlib.redirect_stdout(io.StringIO()):
        env = env_cls(B, *cfgs)
    env.reset()
    if after_reset is not None:
        after_reset(env)
    g = torch.Generator().manual_seed(seed)
    rec = {}
    for t in range(T):
        a = act_std * torch.randn(B, act_dim, generator=g)
        pre = snapshot(env, kind)
        obs, rew, done, extras = env.step(a.clone())
        post = outputs(env, obs, rew, done, extras, kind)
        for k, v in {**pre, **post, "actions": a.numpy()}.items():
            rec.setdefault(k, []).append(np.asarray(v))
    out = {k: np.stack(v).astype(np.float32 if np.asarray(v[0]).dtype.kind == "f" else np.asarray(v[0]).dtype) for k, v in rec.items()}
    return out, cfgs_json


def main():
    os.makedirs(OUT, exist_ok=True)
    from go2_env_walk import Go2Env as WalkEnv
    from go2_env_base import Go2Env as BaseEnv
    from go2_env_stair import Go2Env as StairEnv

    jobs = [
        ("walk_det", WalkEnv, walk_det_cfgs(), "walk", 12, 200, 0.5, 1234, 16),
        ("walk_short", WalkEnv, walk_short_cfgs(), "walk", 8, 160, 0.8, 99, 16),
        ("crouch", BaseEnv, base_cfgs("go2_train_crouch.py"), "base", 8, 160, 0.6, 1234, 12),
        ("jump", BaseEnv, base_cfgs("go2_train_jump.py"), "base", 8, 160, 0.6, 4321, 12),
        ("stair_det", StairEnv, stair_det_cfgs(), "stair", 13, 120, 0.5, 77, 16, stair_place),
    ]
    only = sys.argv[1:]
    for name, cls, cfgs, kind, B, T, std, seed, ad, *hook in jobs:
        if only and name not in only:
            continue
        out, cj = run(cls, cfgs, kind, B, T, std, seed, ad, hook[0] if hook else None)
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
        with open(os.path.join(OUT, name + "_cfg.json"), "w") as f:
            json.dump(cj, f)
        print(name, {k: v.shape for k, v in list(out.items())[:4]}, "resets", int(out["o_reset"].sum()),
              "size MB", os.path.getsize(os.path.join(OUT, name + ".npz")) / 1e6)


if __name__ == "__main__":
    main()
