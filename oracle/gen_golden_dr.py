"""TEST INFRASTRUCTURE ONLY - golden fixtures with domain randomisation ON and every random draw recorded.

Runs the UNMODIFIED reference env classes (/root/reference/examples/locomotion/final/go2_env_{walk,stair,base}.py)
with their shipped configurations on the oracle-backed `genesis` shim and records, per control tick,
  * the state BEFORE the tick (solver state, DR buffers, global DR values, host counters) - to teacher-force from,
  * every random number the reference drew IN the tick (observation / action noise, pushes, command resampling,
    per-episode DR, spawn pose, terrain row, global friction / mass / COM draws) - captured by wrapping the RNG
    primitives the env calls (module-level gs_rand_float / gs_rand_int, torch.randn_like); the env code itself is
    untouched,
  * the outputs AFTER the tick, including each reward term.
An implementation that replays the recorded draws must reproduce the outputs (tests/test_golden_dr.py).

    python oracle/gen_golden_dr.py [name ...]     # walk_c1 walk_hard stair_c1 base_allrew
"""
import contextlib
import io
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from gen_golden import FINAL, OUT, REF, ROOT, load_funcs  # noqa: E402  (also sets sys.path for the shim + reference)

sys.path.insert(0, os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200"))
from go2_b200 import abi  # noqa: E402  (replay row layout only)

MAXC = 24  # compact contact-normal cache entries stored per env (pairs in contact at the previous detection)


class DrawLog:
    """Wraps gs_rand_float / gs_rand_int of an env module and torch.randn_like; logs (caller function, result)."""

    def __init__(self, module):
        self.m = module
        self.calls = []
        self._of, self._oi, self._orn = module.gs_rand_float, getattr(module, "gs_rand_int", None), torch.randn_like
        self._orl, self._or, self._ornn, self._ori = torch.rand_like, torch.rand, torch.randn, torch.randint
        module.gs_rand_float = self._wrap(self._of, "f")
        if hasattr(module, "gs_rand_int"):
            module.gs_rand_int = self._wrap(self._oi, "i")
        torch.randn_like = self._wrap_randn
        # the lidar variant's noise model draws with torch.rand_like / torch.rand / torch.randn (go2_env_stair_lidar.py:973-997);
        # only calls made from its _update_actor_lidar_scan are logged (gs_rand_float itself calls torch.rand)
        torch.rand_like = self._wrap_t(self._orl, "ul")
        torch.rand = self._wrap_t(self._or, "u")
        torch.randn = self._wrap_t(self._ornn, "nn")
        torch.randint = self._wrap_t(self._ori, "ri")  # terrain_manager._sample_spawn_grid: tile row / column

    def _wrap(self, fn, kind):
        def w(*a, **k):
            r = fn(*a, **k)
            self.calls.append((sys._getframe(1).f_code.co_name, kind, r.clone()))
            return r

        return w

    def _wrap_t(self, fn, kind):
        def w(*a, **k):
            r = fn(*a, **k)
            caller = sys._getframe(1).f_code.co_name
            if caller in ("_update_actor_lidar_scan", "_update_stair_info", "_sample_spawn_heightmap", "_sample_spawn_grid"):
                self.calls.append((caller, kind, r.clone()))
            return r

        return w

    def _wrap_randn(self, x, *a, **k):
        r = self._orn(x, *a, **k)
        self.calls.append((sys._getframe(1).f_code.co_name, "n", r.clone()))
        return r

    def restore(self):
        self.m.gs_rand_float, torch.randn_like = self._of, self._orn
        torch.rand_like, torch.rand, torch.randn, torch.randint = self._orl, self._or, self._ornn, self._ori
        if hasattr(self.m, "gs_rand_int"):
            self.m.gs_rand_int = self._oi

    def take(self):
        c, self.calls = self.calls, []
        return c


def compact_cache(cache):
    """[B, npairs, 3] dense contact-normal cache -> ([B, MAXC] pair ids (-1 = none), [B, MAXC, 3])."""
    B = cache.shape[0]
    ids = -np.ones((B, MAXC), dtype=np.int16)
    nrm = np.zeros((B, MAXC, 3), dtype=np.float32)
    for e in range(B):
        nz = np.nonzero(np.abs(cache[e]).sum(1) > 0)[0]
        assert len(nz) <= MAXC
        ids[e, :len(nz)] = nz
        nrm[e, :len(nz)] = cache[e, nz]
    return ids, nrm


def pre_state(env, kind):
    r = env.scene._rigid
    qws, warm, cache = r.get_warm_state()
    ids, nrm = compact_cache(cache)
    # NOT stored, because tests/test_golden_dr.py rebuilds them exactly from the stored arrays (checked here at generation
    # time by check_derived): last_actions / action history / applied actions (from `actions` + resets), last_dof_vel and
    # dof_pos / dof_vel (from qpos / qvel), episode sums (running float32 sums of o_rew_terms)
    s = dict(qpos=r.get_qpos(), qvel=r.get_qvel(), qacc_ws=qws, warm=warm, cache_pair=ids, cache_n=nrm,
             commands=env.commands.numpy().copy(), ep_len=env.episode_length_buf.numpy().copy())
    if kind in ("walk", "stair", "stair_lidar", "stair_info", "walk_terrain", "terrain_env"):
        s.update(hist_idx=np.array(env._action_history_write_idx), delay=env._delay_steps.numpy().copy(),
                 air_time=env._feet_air_time.numpy().copy(), foot_contact=env._foot_contact.numpy().copy(),
                 kp_factors=env._kp_factors.numpy().copy(), kd_factors=env._kd_factors.numpy().copy(), motor_strength=env._motor_strength.numpy().copy(),
                 grav_off=env._gravity_offset.numpy().copy(), level=np.array(float(env.curriculum.level)),
                 g_friction=np.array(float(env._current_friction[0])), g_mass=np.array(float(env._current_mass_shift[0])),
                 g_com=env._current_com_shift[0].numpy().copy(), g_leg=env._current_leg_mass_shifts[0].numpy().copy(),
                 g_counter=np.array(int(env._global_dr_reset_counter)))
        if env.push_force_range_hard is not None:
            s.update(push_force=env._push_stored_force.numpy().copy(), push_remain=env._push_remaining.numpy().copy(), push_counter=np.array(env._push_counter))
    if kind in ("stair", "stair_lidar", "stair_info"):
        s.update(terrain_row=env._env_terrain_row.numpy().astype(np.int32), last_x=env._last_base_pos_x.numpy().copy())
    if kind in ("walk_terrain", "terrain_env"):
        s.update(spawn_pos=env.terrain._env_spawn_pos.numpy().copy())
    if kind == "stair_lidar":
        s.update(actor_scan=env._actor_scan_buf.numpy().copy(), lidar_noise_level=np.array(float(env._lidar_noise_level)))
    if kind == "stair_info":
        s.update(actor_scan=env._stair_info_buf.numpy().copy(), lidar_noise_level=np.array(float(env._stair_noise_level)))
    return s


def draws_of_tick(env, kind, calls, done, ep_len_pre, A):
    """Sort the tick's logged RNG calls into one replay row per env (layout abi.RP_*)."""
    B = env.num_envs
    row = np.zeros((B, abi.REPLAY_LEN), dtype=np.float32)
    flags = dict(push_resampled=0)
    ridx = np.nonzero(done)[0]
    res_steps = int(env.env_cfg["resampling_time_s"] / env.dt)
    cidx = np.nonzero((ep_len_pre + 1) % res_steps == 0)[0]  # envs resampled by step() (go2_env_walk.py:1046-1051)
    by = {}
    for fn, kind_, r in calls:
        by.setdefault(fn, []).append((kind_, r.numpy()))
    if kind == "base":  # go2_env_base.py:118-121,146-151,238: three draws per call, the step-level call even when nobody resamples
        rc = by.pop("_resample_commands", [])
        sets = [rc[i:i + 3] for i in range(0, len(rc), 3)]
        targets = [(cidx, abi.RP_CMD)] + ([(ridx, abi.RP_CMD_RESET)] if len(ridx) else [])
        assert len(sets) == len(targets) and not by
        for (idx, col), st in zip(targets, sets):
            for j in range(3):
                row[idx, col + j] = st[j][1]
        return row, flags
    for _, r in by.pop("step", []):  # action noise (go2_env_walk.py:1001-1003)
        assert r.shape == (B, 12)
        row[:, abi.RP_ACT_NOISE:abi.RP_ACT_NOISE + 12] = r
    lid = by.pop("_update_actor_lidar_scan", [])
    if lid:  # order of go2_env_stair_lidar.py:973-997: cell normals, cell uniforms, offset normal, latency uniform, blackout uniform
        assert [k for k, _ in lid] == ["n", "ul", "nn", "u", "u"], [k for k, _ in lid]
        n_sc = lid[0][1].shape[1]
        row[:, abi.RP_LIDAR_N:abi.RP_LIDAR_N + n_sc] = lid[0][1]
        row[:, abi.RP_LIDAR_U:abi.RP_LIDAR_U + n_sc] = lid[1][1]
        row[:, abi.RP_LIDAR_V] = lid[2][1][:, 0]
        row[:, abi.RP_LIDAR_V + 1] = lid[3][1]
        row[:, abi.RP_LIDAR_V + 2] = lid[4][1]
    sti = by.pop("_update_stair_info", [])
    if sti:  # order of go2_env_stair_lidar_2.py:939-967: edge / height / depth normals, flip / latency / blackout uniforms
        assert [k for k, _ in sti] == ["n", "n", "n", "u", "u", "u"], [k for k, _ in sti]
        for j in range(3):
            row[:, abi.RP_LIDAR_N + j] = sti[j][1]
            row[:, abi.RP_LIDAR_U + j] = sti[3 + j][1]
    for _, r in by.pop("_add_obs_noise", []):
        assert r.shape == (B, env.num_obs)
        # entries with zero noise std never reach the obs; dropping their draws keeps the fixture small
        row[:, abi.RP_OBS_NOISE:abi.RP_OBS_NOISE + env.num_obs] = r * (env.obs_noise_vec.numpy() != 0)
    p = by.pop("_apply_push", [])
    if p:  # fx, fy, duration for ALL envs (go2_env_walk.py:878-885)
        assert [k for k, _ in p] == ["f", "f", "i"]
        for j in range(3):
            row[:, abi.RP_PUSH + j] = p[j][1]
        flags["push_resampled"] = 1
    rc = by.pop("_resample_commands", [])
    if env._compound_commands:
        sets = [rc[i:i + 3] for i in range(0, len(rc), 3)]
        targets = []
        if len(cidx):
            targets.append((cidx, abi.RP_CMD))
        if len(ridx):
            targets.append((ridx, abi.RP_CMD_RESET))
        assert len(sets) == len(targets), (len(sets), len(targets))
        for (idx, col), st in zip(targets, sets):
            for j in range(3):
                assert st[j][1].shape == (len(idx),)
                row[idx, col + j] = st[j][1]
        # the reference zeroes the standing envs' commands after the draw; the replay row carries the drawn values
    else:
        raise NotImplementedError("non-compound command replay")
    if len(ridx):
        n = len(ridx)
        kk = by.pop("_randomize_kp_kd", [])
        if env.pls_enable:
            want = int("kp_factor_range" in env.env_cfg) + int("kd_factor_range" in env.env_cfg)
            assert len(kk) == want
            i = 0
            if "kp_factor_range" in env.env_cfg:
                row[ridx, abi.RP_KPF:abi.RP_KPF + 12] = kk[i][1]; i += 1
            if "kd_factor_range" in env.env_cfg:
                row[ridx, abi.RP_KDF:abi.RP_KDF + 12] = kk[i][1]
        else:
            raise NotImplementedError("fixed-gain manual PD replay")
        for _, r in by.pop("_randomize_gravity_offset", []):
            row[ridx, abi.RP_GO:abi.RP_GO + 3] = r
        for _, r in by.pop("_randomize_motor_strength", []):
            row[ridx, abi.RP_MS:abi.RP_MS + 12] = r
        for _, r in by.pop("_randomize_delay", []):
            row[ridx, abi.RP_DELAY] = r
        rr = by.pop("reset_idx", [])  # init z, roll, pitch in call order (go2_env_walk.py:1190-1203)
        i = 0
        if "init_pos_z_range" in env.env_cfg and not (kind in ("walk_terrain", "terrain_env") and env.terrain.enabled):  # (not drawn on a terrain)
            row[ridx, abi.RP_INIT_Z] = rr[i][1]; i += 1
        if "init_euler_range" in env.env_cfg:
            row[ridx, abi.RP_ROLL] = rr[i][1]; row[ridx, abi.RP_PITCH] = rr[i + 1][1]; i += 2
        assert i == len(rr)
        sg_ = by.pop("_sample_spawn_grid", [])
        if sg_:  # terrain_manager.py:311-328: tile rows, tile columns, x jitter uniforms, y jitter uniforms
            assert [k for k, _ in sg_] == ["ri", "ri", "u", "u"], [k for k, _ in sg_]
            row[ridx, abi.RP_TILE] = sg_[0][1]
            row[ridx, abi.RP_TILE + 1] = sg_[1][1]
            row[ridx, abi.RP_SPAWN] = sg_[2][1]
            row[ridx, abi.RP_SPAWN + 1] = sg_[3][1]
        sp_ = by.pop("_sample_spawn_heightmap", [])
        if sp_:  # terrain_manager.py:293-296: x uniforms, then y uniforms
            assert [k for k, _ in sp_] == ["u", "u"]
            row[ridx, abi.RP_SPAWN] = sp_[0][1]
            row[ridx, abi.RP_SPAWN + 1] = sp_[1][1]
        if kind in ("stair", "stair_lidar", "stair_info"):
            by.pop("_assign_terrain_rows", None)
            row[ridx, abi.RP_ROW] = env._env_terrain_row.numpy()[ridx]
        for k in ("_randomize_friction", "_randomize_mass", "_randomize_leg_mass"):
            by.pop(k, None)  # global draws: recorded through the values in force after the tick (g_* of the next pre-state)
        _ = n
    assert not by, f"unclassified RNG calls: {list(by)}"
    return row, flags


def outputs(env, obs, rew, done, extras, kind, terms):
    r = env.scene._rigid
    # the state after the tick (qpos, qvel, commands, ep_len, DR buffers ...) is the NEXT tick's pre-state (`final_*` after the last)
    o = dict(o_obs=obs.numpy().copy(), o_rew=rew.numpy().copy(), o_reset=done.numpy().astype(np.uint8), o_timeout=extras["time_outs"].numpy().copy(),
             o_ncon=r.get_solver_stats()[:, 0].copy(), o_rew_terms=terms)
    if kind in ("walk", "stair", "stair_lidar", "stair_info", "walk_terrain", "terrain_env"):
        pr = extras["observations"]["critic"].numpy()
        assert np.array_equal(pr[:, :env.num_obs], o["o_obs"])
        o.update(o_priv_tail=pr[:, env.num_obs:].copy(), o_foot_contact=env._foot_contact.numpy().copy())
    return o


class Derived:
    """The env-level buffers the fixture leaves out, rebuilt from the stored arrays exactly like the test does; asserts
    bit-equality with the reference env's own buffers every tick."""

    def __init__(self, env, kind):
        B, A = env.num_envs, env.num_actions
        self.kind = kind
        self.last_actions = np.zeros((B, A), np.float32)
        self.ep_sums = np.zeros((B, len(env.reward_scales)), np.float32)
        if kind != "base":
            self.hist = np.zeros((B, env._delay_buf_size, A), np.float32)

    def check(self, env):
        assert np.array_equal(self.last_actions, env.last_actions.numpy())
        if env.episode_sums:
            assert np.array_equal(self.ep_sums, np.stack([v.numpy() for v in env.episode_sums.values()], 1))
        if self.kind != "base":
            assert np.array_equal(self.hist, env._action_history.numpy())
        r = env.scene._rigid
        md = list(env.motors_dof_idx)
        assert np.array_equal(env.last_dof_vel.numpy(), r.get_qvel()[:, md].astype(np.float32))
        assert np.array_equal(env.dof_vel.numpy(), r.get_qvel()[:, md].astype(np.float32))
        assert np.array_equal(env.dof_pos.numpy(), r.get_qpos()[:, [d + 1 for d in md]].astype(np.float32))

    def advance(self, env, actions, hist_idx, done, terms):
        a = np.clip(actions, -env.env_cfg["clip_actions"], env.env_cfg["clip_actions"]).astype(np.float32)
        self.last_actions = a.copy()
        if self.kind != "base":
            self.hist[:, hist_idx] = a
            self.hist[done] = 0.0
            # (last_actions is overwritten with this tick's actions AFTER reset_idx: go2_env_walk.py:1100-1101)
            self.ep_sums = (self.ep_sums + terms).astype(np.float32)
            self.ep_sums[done] = 0.0
        else:  # base env: reset first, then the rewards of the post-reset state (go2_env_base.py:163-170)
            self.ep_sums[done] = 0.0
            self.ep_sums = (self.ep_sums + terms).astype(np.float32)


def instrument_rewards(env):
    """Record every scaled reward term of a tick: the step loop calls `reward_func() * scale` per name in dict order
    (go2_env_walk.py:1076-1080); wrapping the bound functions in the env's OWN dict leaves the class code untouched."""
    rec = {}
    for name, fn in list(env.reward_functions.items()):
        def make(name, fn):
            def w():
                v = fn()
                rec[name] = (v * env.reward_scales[name]).numpy().copy()
                return v
            return w
        env.reward_functions[name] = make(name, fn)
    return rec


def run(env_cls, module, cfgs, kind, B, T, act_std, seed, act_dim, after_build=None, random_ep_len=True):
    cfgs_json = json.loads(json.dumps(cfgs, default=lambda o: "ndarray"))  # pristine copy (the ctor multiplies reward scales by dt in place)
    torch.manual_seed(1)
    with contextlib.redirect_stdout(io.StringIO()):
        env = env_cls(B, *cfgs)
    if hasattr(env, "_terrain_step_depths"):  # stair-info variant: the rows' tread depths come out of np.random.shuffle; kept with the cfg
        cfgs_json[0]["terrain"]["step_depths_m"] = [float(d) for d in env._terrain_step_depths]
    log = DrawLog(module)
    try:
        env.reset()
        if after_build is not None:
            after_build(env)
        if random_ep_len:  # OnPolicyRunner.learn(init_at_random_ep_len=True): go2_train_walk.py:476
            env.episode_length_buf = torch.randint_like(env.episode_length_buf, high=int(env.max_episode_length))
        rec_terms = instrument_rewards(env)
        names = list(env.reward_scales.keys())
        g = torch.Generator().manual_seed(seed)
        rec = {}
        log.take()
        env.last_actions[:] = 0.0  # (already zero after reset; the base env keeps them across reset() otherwise)
        der = Derived(env, kind)
        der.ep_sums = np.stack([v.numpy().copy() for v in env.episode_sums.values()], 1) if env.episode_sums else der.ep_sums
        for t in range(T):
            a = act_std * torch.randn(B, act_dim, generator=g)
            der.check(env)
            pre = pre_state(env, kind)
            hist_idx = int(getattr(env, "_action_history_write_idx", 0))
            obs, rew, done, extras = env.step(a.clone())
            terms = np.stack([rec_terms[n] for n in names], 1) if names else np.zeros((B, 0), np.float32)
            der.advance(env, a.numpy(), hist_idx, done.numpy().astype(bool), terms)
            row, flags = draws_of_tick(env, kind, log.take(), done.numpy(), pre["ep_len"], act_dim)
            post = outputs(env, obs, rew, done, extras, kind, terms)
            for k, v in {**pre, **post, "actions": a.numpy(), "replay": row, "push_resampled": np.array(flags["push_resampled"])}.items():
                rec.setdefault(k, []).append(np.asarray(v))
        gjk = env.scene._rigid.get_gjk_stats()
        print(f"  GJK-preferred detections (narrowphase.py:724-736) sphere pairs {gjk[0]}, other pairs {gjk[1]} first + {gjk[2]} perturbed, of {gjk[3]} MPR collisions in {B * T} env-ticks")
        final = pre_state(env, kind)  # state after the last tick (DR buffers / globals the last tick's resets drew)
    finally:
        log.restore()
    out = {}
    for k, v in rec.items():
        v0 = np.asarray(v[0])
        keep64 = k == "level"  # host-side python float: the push interval / noise levels are derived from it in double precision
        out[k] = np.stack(v).astype(np.float32 if (v0.dtype.kind == "f" and not keep64) else v0.dtype)
    for k, v in final.items():
        out["final_" + k] = np.asarray(v).astype(np.float32 if (np.asarray(v).dtype.kind == "f" and k != "level") else np.asarray(v).dtype)
    return out, cfgs_json, names


def walk_cfgs():
    return load_funcs(os.path.join(FINAL, "go2_train_walk.py"), ["get_cfgs"]).get_cfgs()


def walk_hard_cfgs():
    """Same cfg at curriculum level 1.0 (full noise / push / DR ranges), short episodes and tight tilt limits so that
    falls, time-outs, command resampling and friction redraws all occur within the horizon."""
    env_cfg, obs_cfg, reward_cfg, command_cfg = walk_cfgs()
    env_cfg["curriculum"].update(level_init=1.0, global_dr_update_interval=6)
    reward_cfg["reward_scales"].update(energy=-2e-4, torque_load=-1e-4)  # zero-scaled in the shipped cfg: exercised here
    env_cfg.update(episode_length_s=2.0, resampling_time_s=0.7, push_interval_s=0.8,
                   termination_if_roll_greater_than=20, termination_if_pitch_greater_than=20)
    return env_cfg, obs_cfg, reward_cfg, command_cfg


def stair_cfgs():
    env_cfg, obs_cfg, reward_cfg, command_cfg = load_funcs(os.path.join(FINAL, "go2_train_stair.py"), ["get_cfgs"]).get_cfgs()
    env_cfg.update(episode_length_s=3.0, resampling_time_s=1.0)  # horizon of the fixture: resets / resampling inside 400 ticks
    return env_cfg, obs_cfg, reward_cfg, command_cfg


def stair_lidar_cfgs():
    """examples/locomotion/go2_stair_train_lidar.get_cfgs() (actor obs 49 + 15 lidar heights, critic scan 9 x 5) at a curriculum level
    where the two-phase DR schedule gives a non-zero lidar noise level, with the fixture horizon of stair_c1."""
    env_cfg, obs_cfg, reward_cfg, command_cfg = load_funcs(os.path.join(REF, "examples/locomotion/go2_stair_train_lidar.py"), ["get_cfgs"]).get_cfgs()
    env_cfg.update(episode_length_s=3.0, resampling_time_s=1.0)
    env_cfg["curriculum"]["level_init"] = 0.95  # past the DR gate: every branch of the lidar noise model fires
    return env_cfg, obs_cfg, reward_cfg, command_cfg


def stair_info_cfgs():
    """examples/locomotion/go2_stair_train_lidar_2.get_cfgs() (actor obs 49 + 4 stair-info values, rows with individual tread depths,
    body_contact reward with a low threshold so that it fires) past the DR gate, with the fixture horizon of stair_c1."""
    env_cfg, obs_cfg, reward_cfg, command_cfg = load_funcs(os.path.join(REF, "examples/locomotion/go2_stair_train_lidar_2.py"), ["get_cfgs"]).get_cfgs()
    env_cfg.update(episode_length_s=3.0, resampling_time_s=1.0)
    env_cfg["curriculum"]["level_init"] = 0.95
    env_cfg["terrain"]["flat_before_m"] = 0.5   # spawn 0.25 m before the first riser: the edge detector (range 0.27 m) sees it from the start
    env_cfg["body_contact_names"] = ["FR_thigh", "FL_thigh", "RR_calf"]  # a foot link as well, so that the count is non-zero in the fixture
    np.random.seed(21)  # the env ctor shuffles the rows' tread depths with numpy's global generator
    return env_cfg, obs_cfg, reward_cfg, command_cfg


def walk_terrain_cfgs():
    """examples/locomotion/go2_train_Omni_kplearn_varied_terrain.get_cfgs() (heightmap mode: left half flat, right half smoothed random
    bumps drawn with numpy's global RNG, seeded here), short episodes, and a small map so that the boundary reset occurs."""
    import ast

    path = os.path.join(REF, "examples/locomotion/go2_train_Omni_kplearn_varied_terrain.py")
    tree = ast.parse(open(path).read())
    keep = [n for n in tree.body if isinstance(n, (ast.FunctionDef, ast.Assign))]  # module constants (TERRAIN_MODE, HEIGHTMAP_*) + functions
    ns = {"np": np}
    exec(compile(ast.Module(body=keep, type_ignores=[]), path, "exec"), ns)
    ns["HEIGHTMAP_SIZE_X"], ns["HEIGHTMAP_SIZE_Y"] = 6.0, 5.0  # 120 x 100 cells: max_wander = 1.5 m
    np.random.seed(11)
    with contextlib.redirect_stdout(io.StringIO()):
        env_cfg, obs_cfg, reward_cfg, command_cfg = ns["get_cfgs"]()
    env_cfg.update(episode_length_s=4.0, resampling_time_s=1.0)
    env_cfg["curriculum"]["level_init"] = 1.0  # full command ranges: the robots cover ground
    return env_cfg, obs_cfg, reward_cfg, command_cfg


def walk_grid_cfgs():
    """The same train script with TERRAIN_MODE = "grid" (make_terrain_cfg_grid: 2 x 2 tiles of 8 m, two flat and two random_uniform
    ones generated by Genesis' terrain parser with numpy's global RNG, seeded here), short episodes."""
    import ast

    path = os.path.join(REF, "examples/locomotion/go2_train_Omni_kplearn_varied_terrain.py")
    tree = ast.parse(open(path).read())
    keep = [n for n in tree.body if isinstance(n, (ast.FunctionDef, ast.Assign))]
    ns = {"np": np}
    exec(compile(ast.Module(body=keep, type_ignores=[]), path, "exec"), ns)
    ns["TERRAIN_MODE"] = "grid"
    with contextlib.redirect_stdout(io.StringIO()):
        env_cfg, obs_cfg, reward_cfg, command_cfg = ns["get_cfgs"]()
    assert "subterrain_types" in env_cfg["terrain"] and env_cfg["terrain"].get("height_field") is None
    env_cfg["terrain"]["subterrain_size"] = list(env_cfg["terrain"]["subterrain_size"])
    env_cfg["terrain"]["boundary_margin"] = 3.0  # max_wander = 1 m: the boundary reset occurs inside the fixture
    env_cfg.update(episode_length_s=4.0, resampling_time_s=1.0)
    env_cfg["curriculum"]["level_init"] = 1.0
    env_cfg["terrain"]["np_seed"] = 29  # (fixture key) numpy's global generator is seeded with this right before the env is built
    np.random.seed(29)
    return env_cfg, obs_cfg, reward_cfg, command_cfg


def terrain_env_cfgs():
    """go2_env_terrain.py has no train script in the tree: the height-map cfg above with what the class adds switched on - one more
    privileged-obs entry (ground height), the stumble reward (low threshold so that it fires) and the foot rewards it relies on."""
    env_cfg, obs_cfg, reward_cfg, command_cfg = walk_terrain_cfgs()
    obs_cfg["num_privileged_obs"] = obs_cfg["num_privileged_obs"] + 1
    reward_cfg["reward_scales"]["stumble"] = -0.05
    reward_cfg["stumble_threshold"] = 1.0
    return env_cfg, obs_cfg, reward_cfg, command_cfg


def base_allrew_cfgs():
    """go2_train_crouch.get_cfgs() with EVERY reward term of go2_env_base.py switched on (the shipped scripts use 6-10 of the 20;
    logs/go2-crouch/cfgs.pkl names crouch_2 / y_stability / crouch_speed but lacks the termination keys the final env class
    reads, so it cannot be run as is), non-zero command ranges and short episodes."""
    env_cfg, obs_cfg, reward_cfg, command_cfg = load_funcs(os.path.join(FINAL, "go2_train_crouch.py"), ["get_cfgs"]).get_cfgs()
    _, _, jr, _ = load_funcs(os.path.join(FINAL, "go2_train_jump.py"), ["get_cfgs"]).get_cfgs()
    for k, v in jr.items():
        if k != "reward_scales":
            reward_cfg.setdefault(k, v)
    reward_cfg.setdefault("tracking_sigma", 0.25)
    reward_cfg.setdefault("base_height_target", 0.3)
    reward_cfg["reward_scales"] = {
        "tracking_lin_vel": 1.0, "tracking_ang_vel": 0.5, "lin_vel_z": -1.0, "base_height": -5.0, "action_rate": -0.05, "similar_to_default": -0.1,
        "jump_impulse": 6.0, "jump_apex": 20.0, "xy_stability": 12.0, "orientation": 3.0, "no_shake": 1.0, "crouch": 6.0, "crouch_2": 4.0,
        "ground_penalty": 10.0, "crouch_target": 50.0, "no_fall": 2.0, "y_stability": 3.0, "torque_load": 0.5, "crouch_progress": 50.0,
        "crouch_speed": 1.5}
    env_cfg.update(episode_length_s=1.5, resampling_time_s=0.5, termination_if_roll_greater_than=35, termination_if_pitch_greater_than=35,
                   termination_if_z_vel_greater_than=100.0, termination_if_y_vel_greater_than=100.0)
    command_cfg.update(lin_vel_x_range=[-0.5, 0.5], lin_vel_y_range=[-0.2, 0.2], ang_vel_range=[-0.4, 0.4])
    return env_cfg, obs_cfg, reward_cfg, command_cfg


def main():
    os.makedirs(OUT, exist_ok=True)
    import go2_env_base
    import go2_env_stair
    import go2_env_walk

    sys.path.insert(0, os.path.join(REF, "examples/locomotion"))
    import go2_env_stair_lidar
    import go2_env_Omni_kplearn_varied_terrain as go2_env_wt
    import terrain_manager  # noqa: F401  (the env imports it inside its ctor)

    import go2_env_terrain
    import go2_env_stair_lidar_2

    jobs = {
        "stair_info_c1": lambda: (go2_env_stair_lidar_2.Go2Env, go2_env_stair_lidar_2, stair_info_cfgs(), "stair_info", 16, 300, 0.5, 2121, 16),
        "terrain_env_c1": lambda: (go2_env_terrain.Go2Env, go2_env_terrain, terrain_env_cfgs(), "terrain_env", 16, 300, 0.6, 616, 16),
        "walk_grid_c1": lambda: (go2_env_wt.Go2Env, go2_env_wt, walk_grid_cfgs(), "walk_terrain", 16, 300, 0.5, 818, 16),
        "walk_terrain_c1": lambda: (go2_env_wt.Go2Env, go2_env_wt, walk_terrain_cfgs(), "walk_terrain", 16, 300, 0.5, 515, 16),
        "stair_lidar_c1": lambda: (go2_env_stair_lidar.Go2Env, go2_env_stair_lidar, stair_lidar_cfgs(), "stair_lidar", 16, 300, 0.5, 909, 16),
        "walk_c1": lambda: (go2_env_walk.Go2Env, go2_env_walk, walk_cfgs(), "walk", 32, 1000, 0.5, 1234, 16),
        "walk_hard": lambda: (go2_env_walk.Go2Env, go2_env_walk, walk_hard_cfgs(), "walk", 16, 300, 0.6, 4242, 16),
        "stair_c1": lambda: (go2_env_stair.Go2Env, go2_env_stair, stair_cfgs(), "stair", 16, 400, 0.5, 777, 16),
        "base_allrew": lambda: (go2_env_base.Go2Env, go2_env_base, base_allrew_cfgs(), "base", 8, 240, 0.6, 31, 12),
    }
    only = sys.argv[1:] or list(jobs)
    for name in only:
        cls, mod, cfgs, kind, B, T, std, seed, ad = jobs[name]()
        hf = None
        if isinstance(cfgs[0].get("terrain", {}).get("height_field", None), np.ndarray):
            hf = cfgs[0]["terrain"]["height_field"]  # kept as an array for the reference env, stored in the npz for the tests
        out, cj, names = run(cls, mod, cfgs, kind, B, T, std, seed, ad)
        if hf is not None:
            out["cfg_height_field"] = hf
            cj[0]["terrain"]["height_field"] = "npz:cfg_height_field"
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
        with open(os.path.join(OUT, name + "_cfg.json"), "w") as f:
            json.dump(cj, f)
        print(name, "B,T", B, T, "resets", int(out["o_reset"].sum()), "timeouts", int(out["o_timeout"].sum()),
              "pushes", int(out["push_resampled"].sum()), "mean contacts", float(out["o_ncon"].mean()),
              "size MB", round(os.path.getsize(os.path.join(OUT, name + ".npz")) / 1e6, 2), "terms", names)


if __name__ == "__main__":
    main()
