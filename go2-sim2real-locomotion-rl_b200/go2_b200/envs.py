"""Drop-in `Go2Env` classes backed by the fused CUDA step (walk / base flavours).

Same constructor, attributes and return values as the reference env classes
(examples/locomotion/final/go2_env_walk.py:154-171,985-1109,1145-1245 and go2_env_base.py), so rsl-rl 2.2.4's
OnPolicyRunner and the repo's train scripts can use them unchanged.  What the reference does per step with
~80 torch kernels + ~20 quadrants kernels happens here in ONE kernel launch; the Python side only keeps the
scalar state machines the reference also keeps on the host (curriculum, global domain randomisation, push
schedule, delay-ring write index) and turns them into the per-step parameter block.
"""
from __future__ import annotations

import ctypes as C
import math

import numpy as np
import torch

from . import abi
from .core import Go2Core
from .model import load_model, motors_dof_idx

DT = 0.02


def _clamp01(x):
    return float(max(0.0, min(1.0, x)))


def _lerp(a, b, t):
    return a + (b - a) * _clamp01(t)


def _lerp_range(a, b, t):
    t = 0.0 if t < 0.0 else (1.0 if t > 1.0 else float(t))  # == _clamp01, inlined: this runs ~15 times per tick on the host
    return [a[0] + (b[0] - a[0]) * t, a[1] + (b[1] - a[1]) * t]


class CurriculumManager:
    """Metric-gated difficulty level (behaviour of go2_env_walk.py:42-151 CurriculumManager; same constructor).
    `rng` (optional numpy Generator) drives sample_level(); the reference draws from python's `random` module."""

    def __init__(self, cfg: dict = None, rng=None):
        self.cfg = cfg or {}
        c = self.cfg
        g = c.get
        self.enabled = bool(g("enabled", False))
        self.level = float(g("level_init", 0.0))
        self.level_min, self.level_max = float(g("level_min", 0.0)), float(g("level_max", 1.0))
        self.alpha = float(g("ema_alpha", 0.05))
        self.ready = (float(g("ready_timeout_rate", 0.7)), float(g("ready_tracking", 0.6)), float(g("ready_fall_rate", 0.30)))
        self.ready_need, self.hard_rate, self.hard_need = int(g("ready_streak", 3)), float(g("hard_fall_rate", 0.55)), int(g("hard_streak", 2))
        self.step_up, self.step_down, self.cooldown_updates = float(g("step_up", 0.02)), float(g("step_down", 0.01)), int(g("cooldown_updates", 1))
        self.mix = (float(g("mix_prob_current", 0.80)), float(g("mix_level_low", 0.0)), float(g("mix_level_high", 0.6)))
        self.ema = {"timeout_rate": None, "tracking": None, "fall_rate": None}
        self._ready_streak = self._hard_streak = self._cooldown = 0
        self._rng = rng if rng is not None else np.random.default_rng()

    def sample_level(self):
        if not self.enabled:
            return 1.0
        if self._rng.random() < self.mix[0]:
            return _clamp01(self.level)
        hi = min(self.level, self.mix[2])
        lo = min(self.mix[1], hi)
        return _clamp01(lo + (hi - lo) * self._rng.random())

    def update(self, timeout_rate, tracking, fall_rate):
        if not self.enabled:
            return False
        for k, x in (("timeout_rate", timeout_rate), ("tracking", tracking), ("fall_rate", fall_rate)):
            old = self.ema[k]
            self.ema[k] = float(x) if old is None else float((1.0 - self.alpha) * old + self.alpha * float(x))
        if self._cooldown > 0:
            self._cooldown -= 1
        ready = self.ema["timeout_rate"] >= self.ready[0] and self.ema["tracking"] >= self.ready[1] and self.ema["fall_rate"] <= self.ready[2]
        hard = self.ema["fall_rate"] >= self.hard_rate
        self._ready_streak = self._ready_streak + 1 if ready else 0
        self._hard_streak = self._hard_streak + 1 if hard else 0
        old = self.level
        if self._hard_streak >= self.hard_need:
            self.level = max(self.level_min, self.level - self.step_down)
            self._hard_streak = self._ready_streak = 0
            self._cooldown = self.cooldown_updates
        elif self._ready_streak >= self.ready_need and self._cooldown == 0:
            self.level = min(self.level_max, self.level + self.step_up)
            self._ready_streak = 0
            self._cooldown = self.cooldown_updates
        self.level = _clamp01(self.level)
        return self.level != old

    def state_dict(self):
        return {"enabled": self.enabled, "level": float(self.level), "timeout_rate_ema": self.ema["timeout_rate"],
                "tracking_ema": self.ema["tracking"], "fall_rate_ema": self.ema["fall_rate"]}

    # attribute names of the reference class (go2_env_walk.py:58-90), read by logging code
    timeout_rate_ema = property(lambda self: self.ema["timeout_rate"])
    tracking_ema = property(lambda self: self.ema["tracking"])
    fall_rate_ema = property(lambda self: self.ema["fall_rate"])


Curriculum = CurriculumManager


def _inv_quat(q):
    return [q[0], -q[1], -q[2], -q[3]]


def _default_device():
    try:
        import genesis as gs  # the shim (or a real install) decides the device, like the reference does

        return gs.device
    except Exception:
        return torch.device("cuda:0")


_DEFAULT_CORE = [Go2Core]  # the test-suite swaps in its host emulation of the kernel source here (tests/hostemu_core.py)


# default arithmetic of new envs: False = IEEE division / sqrt (libgo2_b200.so), True = approximate (libgo2_b200_fast.so, ~9 % faster).
# The drop-in `genesis.init()` sets it like the reference does for its own solver: fast_math = not debug (genesis/__init__.py:289).
DEFAULT_FAST_MATH = [False]


class _Go2EnvBase:
    """State shared by the walk / stair / base flavours: buffers, views, rsl-rl facing attributes."""

    KIND = abi.ENV_WALK

    def __init__(self, num_envs, env_cfg, obs_cfg, reward_cfg, command_cfg, show_viewer=False, *, device=None, seed=1,
                 env_offset=0, total_envs=None, lib_path=None, stats_sync_interval=1, sync_ranks=False, core_cls=None,
                 replay=False, export_reward_terms=False, solver_dense=False, fast_math=None):
        """Keyword-only extras (none is needed by the train / eval scripts):
        stats_sync_interval  ticks between reads of the device-side episode statistics.  1 = read (and wait) every tick, the
                             reference's timing for extras["episode"] and the curriculum; k > 1 = an asynchronous double-buffered
                             read every k ticks (no tick waits for the host; the curriculum sees the numbers k ticks late).
        sync_ranks           multi-process runs (torch.distributed initialised).  False (default): every rank is an independent env
                             shard with its own curriculum / global-DR state machine - what N copies of the reference script
                             would be; no collective on the step path.  True: the per-tick statistics row is all-reduced between
                             the step kernel and the post-pass, so all ranks run ONE state machine, as a single env of
                             world x num_envs would (shard-invariant results) - at the price of a per-tick lockstep that also
                             aligns the ranks' host-link bursts (8 GPUs, host buffers: 152 M vs 194 M env-steps/s).
        fast_math            True = the library build with approximate division / sqrt (the reference's own fast_math arithmetic class,
                             ~9 % faster), False = IEEE (bit-faithful parity against the CPU oracle); None = DEFAULT_FAST_MATH.
        replay               allocate the draw-replay buffer (fixture replay, see replay_draws()).
        export_reward_terms  keep every scaled reward term of the tick in `reward_terms` [B, n_rew].
        solver_dense         diagnostic: every env takes the dense 18x18 Newton path."""
        if show_viewer:
            raise RuntimeError("show_viewer=True: no graphics stack on this box (viewer unavailable)")
        self.num_envs = int(num_envs)
        self.num_obs = obs_cfg["num_obs"]
        self.num_actions = env_cfg["num_actions"]
        self.num_commands = command_cfg["num_commands"]
        self.device = torch.device(device) if device is not None else torch.device(_default_device())
        self.dt = DT
        self.max_episode_length = math.ceil(env_cfg["episode_length_s"] / self.dt)
        self.env_cfg, self.obs_cfg, self.reward_cfg, self.command_cfg = env_cfg, obs_cfg, reward_cfg, command_cfg
        self.obs_scales = obs_cfg["obs_scales"]
        self.reward_scales = reward_cfg["reward_scales"]
        self._model = load_model()
        self.motors_dof_idx = motors_dof_idx(self._model, env_cfg["joint_names"])
        self._seed = int(seed)
        self._np_rng = np.random.default_rng(self._seed)
        self._env_offset = int(env_offset)
        self._total_envs = int(total_envs) if total_envs is not None else self.num_envs
        self.fast_math = bool(DEFAULT_FAST_MATH[0] if fast_math is None else fast_math)
        self._lib_path = lib_path or (abi.CUDA_LIB_FAST_PATH if self.fast_math and core_cls is None and _DEFAULT_CORE[0] is Go2Core else None)
        self._core_cls = core_cls or _DEFAULT_CORE[0]
        self.stats_sync_interval = int(stats_sync_interval)
        self._sync_ranks = bool(sync_ranks)
        self._opt_replay, self._opt_rew_terms, self._opt_dense = bool(replay), bool(export_reward_terms), bool(solver_dense)
        self._step_count = 0
        self._ticks_since_stats = 0
        self._stats_slot = 0
        self._stats_inflight = [False, False]
        self._replay_next = False
        self._ext_on = False
        self._kernel_event_hook = None
        self.extras = {"observations": {}}

    # ---- cfg -> EnvConfig fields common to all flavours
    def _fill_common(self, c: abi.EnvConfig):
        e = self.env_cfg
        c.kind = self.KIND
        c.num_envs, c.env_offset = self.num_envs, self._env_offset
        c.num_actions, c.num_obs = self.num_actions, self.num_obs
        c.max_episode_length = self.max_episode_length
        c.resample_steps = int(e["resampling_time_s"] / self.dt)
        names = list(self.reward_scales.keys())
        c.n_rew = len(names)
        for k, name in enumerate(names):
            if name not in abi.REWARD_IDS:
                raise AttributeError(f"Reward function '_reward_{name}' not found in Go2Env.")
            self.reward_scales[name] *= self.dt  # the reference mutates the cfg dict in place (go2_env_walk.py:255-265)
            c.rew_id[k] = abi.REWARD_IDS[name]
            c.rew_scale[k] = float(self.reward_scales[name])
        self._reward_names = names
        for j, d in enumerate(self.motors_dof_idx):
            c.motors_dof[j] = d
            c.default_dof_pos[j] = float(e["default_joint_angles"][e["joint_names"][j]])
        c.action_scale, c.clip_actions = float(e["action_scale"]), float(e["clip_actions"])
        c.kp, c.kd = float(e["kp"]), float(e["kd"])
        big = 1e30
        c.term_roll = float(e.get("termination_if_roll_greater_than", big))
        c.term_pitch = float(e.get("termination_if_pitch_greater_than", big))
        c.term_zvel = float(e.get("termination_if_z_vel_greater_than", big))
        c.term_yvel = float(e.get("termination_if_y_vel_greater_than", big))
        for j in range(3):
            c.base_init_pos[j] = float(e["base_init_pos"][j])
        q = [float(v) for v in e["base_init_quat"]]
        for j in range(4):
            c.base_init_quat[j] = q[j]
            c.inv_base_init_quat[j] = _inv_quat(q)[j]
        s = self.obs_scales
        c.obs_scale_lin_vel, c.obs_scale_ang_vel = float(s["lin_vel"]), float(s["ang_vel"])
        c.obs_scale_dof_pos, c.obs_scale_dof_vel = float(s["dof_pos"]), float(s["dof_vel"])
        r = self.reward_cfg
        c.tracking_sigma = float(r.get("tracking_sigma", 0.25))
        c.base_height_target = float(r.get("base_height_target", 0.3))
        c.feet_height_target = float(r.get("feet_height_target", 0.075))
        c.feet_air_time_target = float(r.get("feet_air_time_target", 0.1))
        c.rp[0] = float(r.get("jump_apex_height", 0.55))
        c.rp[1] = float(r.get("jump_apex_sigma", 0.05))
        c.seed_lo = self._seed & 0xFFFFFFFF
        c.seed_hi = (self._seed >> 32) & 0xFFFFFFFF
        c.replay_rows = self.num_envs if self._opt_replay else 0
        c.export_rew_terms = int(self._opt_rew_terms)
        c.solver_dense = int(self._opt_dense)

    def _bind(self, c: abi.EnvConfig):
        self.core = self._core_cls(c, device=self.device, model=self._model, lib_path=self._lib_path)
        b = self.core.buf
        B = self.num_envs
        self._obs_view = b["obs"]   # the kernel's persistent output row; obs_buf is re-bound to a fresh copy every step
        self.obs_buf = b["obs"].clone()
        self.reward_terms = b["rew_terms"] if self._opt_rew_terms else None
        self.rew_buf = b["rew"].view(-1)
        self.reset_buf = b["reset"].view(-1).view(torch.bool)
        self._episode_length_buf = b["ep_len"].view(-1)
        self.commands = b["commands"]
        self.last_actions = b["last_actions"]
        self.dof_pos, self.dof_vel, self.last_dof_vel = b["dof_pos"], b["dof_vel"], b["last_dof_vel"]
        self.base_pos, self.base_quat = b["base_pos"], b["base_quat"]
        self.base_lin_vel, self.base_ang_vel = b["base_lin_vel"], b["base_ang_vel"]
        self.projected_gravity, self.base_euler = b["proj_g"], b["base_euler"]
        self.actions = torch.zeros((B, self.num_actions), device=self.device, dtype=torch.float32)
        self.default_dof_pos = torch.tensor([c.default_dof_pos[j] for j in range(12)], device=self.device, dtype=torch.float32)
        self.base_init_pos = torch.tensor(list(c.base_init_pos), device=self.device)
        self.base_init_quat = torch.tensor(list(c.base_init_quat), device=self.device)
        self.inv_base_init_quat = torch.tensor(list(c.inv_base_init_quat), device=self.device)
        self.commands_scale = torch.tensor([c.obs_scale_lin_vel, c.obs_scale_lin_vel, c.obs_scale_ang_vel], device=self.device)
        self.global_gravity = torch.tensor([0.0, 0.0, -1.0], device=self.device).repeat(B, 1)
        self.episode_sums = {name: b["ep_sums"][:, k] for k, name in enumerate(self._reward_names)}
        self.extras["time_outs"] = b["timeout"].view(-1)
        self._ecfg = c

    # rsl-rl rebinds this attribute (`env.episode_length_buf = torch.randint_like(...)`): keep the device buffer in sync
    @property
    def episode_length_buf(self):
        return self._episode_length_buf

    @episode_length_buf.setter
    def episode_length_buf(self, value):
        self._episode_length_buf.copy_(torch.as_tensor(value, device=self.device).to(torch.int32))

    def get_observations(self):
        self.extras["observations"]["critic"] = self._critic_obs()
        return self.obs_buf, self.extras

    def _critic_obs(self):
        return self.obs_buf

    def close(self):
        self.core.close()

    def step_host(self, h_actions, h_obs, h_rew, h_done, h_priv=None, h_timeouts=None):
        """Same tick through the C ABI with HOST buffers (numpy or pinned torch CPU tensors): the action H2D copy
        and the obs / rew / done D2H copies happen inside the call, which returns after the stream is idle."""
        sp = self._make_params(physics=True)
        if self._multi_rank():
            # the statistics all-reduce sits between the step and the post-pass launch
            import torch.distributed as dist

            b = self.core.buf
            sp.defer_post = 1
            staged = self.core.step_host_begin(h_actions, sp, h_obs, h_rew, h_done)
            dist.all_reduce(b["stats"][self.core.tick_parity()])
            self.core.post(sp)
            pairs = [(h_timeouts, b["timeout"].view(-1)), (h_priv, b["priv"])]
            if staged:
                pairs = [(h_rew, b["rew"].view(-1)), (h_done, b["reset"].view(-1)), (h_obs, b["obs"])] + pairs
            for dst, src in pairs:
                if dst is not None:
                    torch.as_tensor(dst).copy_(src, non_blocking=True)
            torch.cuda.current_stream(self.device).synchronize()
            self._after_launch()
        else:
            self.core.step_host(h_actions, sp, h_obs, h_rew, h_done, h_priv, h_timeouts)
            self._after_launch()
        self.actions = self.last_actions
        self._stats_tick()
        return h_obs, h_rew, h_done, self.extras

    # ---- one tick on the device: step kernel (+ statistics all-reduce over the env shards) + post-pass kernel
    def _multi_rank(self):
        import torch.distributed as dist

        return self._sync_ranks and dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1

    def _launch(self, actions, sp):
        multi, hook = self._multi_rank(), self._kernel_event_hook
        if actions is None:
            sp.defer_post = int(multi)
            self.core.reset_all(sp)
        else:
            sp.defer_post = int(multi or hook is not None)
            if hook is not None:  # measurement aid: events bracketing the step kernel alone
                e0, e1 = hook()
                e0.record()
            self.core.step(actions, sp)
            if hook is not None:
                e1.record()
        if sp.defer_post:
            if multi:
                import torch.distributed as dist

                # this tick's statistics row, summed over the env shards on the stream (NCCL) - no host synchronisation
                dist.all_reduce(self.core.buf["stats"][self.core.tick_parity()])
            self.core.post(sp)
        self._after_launch()

    def _after_launch(self):
        if self._ext_on:  # clear_external_force after the tick's substeps (simulator.py:283)
            ex = self.core.buf["ext_force"]
            ex.zero_()
            ex[:, 0] = -1.0
            self._ext_on = False
        self._replay_next = False

    def _stats_tick(self):
        """Host side of reset_idx that needs the episode statistics (curriculum aggregates, extras['episode'])."""
        k = self.stats_sync_interval
        if k <= 0:
            return
        self._ticks_since_stats += 1
        if self._ticks_since_stats < k:
            return
        self._ticks_since_stats = 0
        if k == 1:
            self._consume_stats(self.core.read_stats())
            return
        # asynchronous, double-buffered: enqueue this interval's read, consume the previous interval's (long complete)
        cur, prev = self._stats_slot, self._stats_slot ^ 1
        self.core.read_stats_async(cur)
        self._stats_inflight[cur] = True
        if self._stats_inflight[prev]:
            self._consume_stats(self.core.stats_fetch(prev, block=True))
            self._stats_inflight[prev] = False
        self._stats_slot = prev

    def flush_stats(self):
        """Consume every outstanding statistics read (blocking) - e.g. before reading extras at the end of a rollout."""
        for slot in (self._stats_slot, self._stats_slot ^ 1):
            if self._stats_inflight[slot]:
                self._consume_stats(self.core.stats_fetch(slot, block=True))
                self._stats_inflight[slot] = False
        self._consume_stats(self.core.read_stats())
        self._ticks_since_stats = 0

    def _consume_stats(self, stats):
        """One read of the accumulated statistics: the host side of reset_idx, and - like Genesis' check_errno, which polls the
        solver's errno every 10 substeps (rigid_solver.py:1189-1213) - the error check: the statistics carry the number of env-ticks
        that raised an error bit, so the per-env errno buffer is only read (a sync) when there is something to report."""
        self._after_resets(stats)
        if float(stats[abi.STATS_ERR]) > 0:
            self.extras["errno_env_ticks"] = int(round(float(stats[abi.STATS_ERR])))
            self.poll_errno()  # raises on contact / pair overflow; NaN envs were force-reset (bits 4 / 8 are returned, not raised)

    def replay_draws(self, rows):
        """Fixture replay: the NEXT step reads every random draw from `rows` ([B, REPLAY_LEN], layout abi.RP_*) instead of Philox."""
        if not self._opt_replay:
            raise RuntimeError("construct the env with replay=True")
        rows = torch.as_tensor(rows, device=self.device, dtype=torch.float32)
        self.core.buf["replay"][:, :rows.shape[1]].copy_(rows)  # (fixtures recorded before a layout extension are narrower)
        self._replay_next = True

    # ---- global physics DR in force (device-resident, both double-buffer slots)
    def _set_global(self, col, value):
        # only the slot the next tick reads: the other one keeps the values the last closing FK ran on (go2_env.cuh, t_env_step)
        self.core.buf["globals"][self.core.tick_parity(), col] = float(value)

    def global_dr_values(self):
        """Friction / mass / COM shifts currently in force (synchronises: a small D2H read)."""
        g = self.core.buf["globals"][self.core.tick_parity()].tolist()
        return {"friction_ground": g[abi.G_FRIC_GROUND], "friction_robot": g[abi.G_FRIC_ROBOT],
                "mass_shift": g[abi.G_MASS_SHIFT:abi.G_MASS_SHIFT + abi.NLINK], "com_shift": g[abi.G_COM_SHIFT:abi.G_COM_SHIFT + 3],
                "reset_counter": g[abi.G_DR_COUNTER]}

    def poll_errno(self):
        """Raise like Genesis' check_errno (rigid_solver.py:1189-1213) if any env hit an error bit."""
        bits = 0
        for v in torch.unique(self.core.buf["errno"]).tolist():
            bits |= int(v)
        if bits & 32:
            raise RuntimeError("go2_step_host: the action upload did not reach the device within 2 s (copy stream stalled)")
        if bits & 2:
            raise RuntimeError("Exceeding max number of contact pairs. Please increase 'max_collision_pairs'.")
        if bits & 1:
            raise RuntimeError("Exceeding max number of colliding geom pairs per env.")
        return bits  # 4 / 8: non-finite force / acceleration (env was force-reset); 16: a non-sphere pair met the reference's GJK-fallback condition


class Go2EnvWalk(_Go2EnvBase):
    """Walk env (16-dim PLS action, delay ring, manual PD, pushes, curriculum, DR, 49/104-d obs)."""

    KIND = abi.ENV_WALK

    def __init__(self, num_envs, env_cfg, obs_cfg, reward_cfg, command_cfg, show_viewer=False, **kw):
        super().__init__(num_envs, env_cfg, obs_cfg, reward_cfg, command_cfg, show_viewer, **kw)
        e = env_cfg
        self.num_privileged_obs = obs_cfg.get("num_privileged_obs", None)
        self.num_pos_actions = e.get("num_pos_actions", 12)
        self.simulate_action_latency = e.get("simulate_action_latency", True)
        self.pls_enable = bool(e.get("pls_enable", False))
        self.num_stiffness_actions = 4 if self.pls_enable else 0
        c = abi.EnvConfig()
        self._fill_common(c)
        c.num_pos_actions = self.num_pos_actions
        c.num_priv = int(self.num_privileged_obs or 0)
        if self.pls_enable:
            self.pls_kp_range = e["pls_kp_range"]
            self.pls_kp_default = float(e["pls_kp_default"])
            self.pls_kp_action_scale = float(e["pls_kp_action_scale"])
            c.pls_kp_default, c.pls_kp_scale = self.pls_kp_default, self.pls_kp_action_scale
            c.pls_kp_lo, c.pls_kp_hi = float(self.pls_kp_range[0]), float(self.pls_kp_range[1])
        tl = e.get("torque_limits", None) or [23.7] * self.num_pos_actions
        for j in range(12):
            c.torque_limit[j] = float(tl[j])
        self._use_manual_pd = self.pls_enable or ("kp_factor_range" in e)
        c.ctrl = abi.CTRL_MANUAL_PLS if self.pls_enable else (abi.CTRL_MANUAL_FIXED if self._use_manual_pd else abi.CTRL_GENESIS_PD)
        # foot contact tracking (go2_env_walk.py:270-276,531-597)
        link_names = [l["name"] for l in self._model["links"]]
        self._foot_reward_names = ["feet_air_time", "foot_slip", "foot_clearance", "feet_stance"]
        self._has_foot_contact_rewards = any(n in self.reward_scales and self.reward_scales[n] != 0.0 for n in self._foot_reward_names)
        self._foot_link_indices = []
        if self._has_foot_contact_rewards:
            foot_names = e.get("foot_names", [])
            if not foot_names:
                raise ValueError("env_cfg['foot_names'] is required for feet_air_time/foot_slip/foot_clearance rewards.")
            for fn in foot_names:
                if fn not in link_names:
                    raise RuntimeError(f"Foot link '{fn}' NOT FOUND. Available: {sorted(link_names)}")
                self._foot_link_indices.append(link_names.index(fn))
            self._n_feet = len(foot_names)
            self._has_link_vel = self._has_link_pos = True
            self._foot_contact_threshold = float(e.get("foot_contact_threshold", 1.0))
            c.has_foot_tracking = 1
            c.foot_contact_threshold = self._foot_contact_threshold
            for i in range(4):
                c.foot_link[i] = self._foot_link_indices[i]
        self._has_leg_mass_dr = "leg_mass_shift_range" in e
        self._hip_link_indices = [link_names.index(n) for n in ("FR_hip", "FL_hip", "RR_hip", "RL_hip")] if self._has_leg_mass_dr else []
        # global physics DR (one value for every env): applied on the device by the post-pass kernel
        c.has_friction_dr, c.has_mass_dr, c.has_com_dr = int("friction_range" in e), int("mass_shift_range" in e), int("com_shift_range" in e)
        c.has_leg_mass_dr = int(self._has_leg_mass_dr)
        c.global_dr_interval = int((e.get("curriculum", {}) or {}).get("global_dr_update_interval", 200))
        for i, l in enumerate(self._hip_link_indices):
            c.hip_link[i] = l
        # DR maxima
        self.obs_noise_components = e.get("obs_noise", None)
        self.obs_noise_level_max = float(e.get("obs_noise_level", 0.0)) if self.obs_noise_components else 0.0
        self.action_noise_std_max = float(e.get("action_noise_std", 0.0)) if ("action_noise_std" in e) else 0.0
        self.push_interval_s_hard = float(e.get("push_interval_s", 5.0))
        self.push_force_range_hard = e.get("push_force_range", None)
        self._push_counter = 0
        if self.push_force_range_hard is not None:
            self._base_link_idx = 2  # reference quirk: robot.links[1].idx == global link 2 == FL_hip (go2_env_walk.py:339-342)
            c.has_push, c.push_link = 1, 1
            dur = e.get("push_duration_s", [0.05, 0.15])
            self._push_duration_steps = [max(1, int(dur[0] / self.dt)), max(1, int(dur[1] / self.dt))]
            c.push_dur_lo, c.push_dur_hi = self._push_duration_steps
        # commands
        self._cmd_curriculum = bool(command_cfg.get("cmd_curriculum", False))
        self._cmd_start_frac = float(command_cfg.get("cmd_curriculum_start_frac", 0.1))
        self._rel_standing_envs = float(command_cfg.get("rel_standing_envs", 0.0))
        self._compound_commands = bool(command_cfg.get("compound_commands", True))
        self._cmd_full_ranges = {k: list(command_cfg[k + "_range"]) for k in ("lin_vel_x", "lin_vel_y", "ang_vel")}
        self._cmd_cur_ranges = {k: list(v) for k, v in self._cmd_full_ranges.items()}
        c.compound_commands = int(self._compound_commands)
        c.n_standing = int(self._rel_standing_envs * self._total_envs)  # by GLOBAL env index
        # delay ring
        self._min_delay_steps = int(e.get("min_delay_steps", 0))
        self._max_delay_steps = int(e.get("max_delay_steps", 2))
        self._delay_buf_size = self._max_delay_steps + 1
        self._action_history_write_idx = 0
        c.hist_size, c.min_delay, c.max_delay = self._delay_buf_size, self._min_delay_steps, self._max_delay_steps
        # per-episode DR switches
        c.has_kp_factor, c.has_kd_factor = int("kp_factor_range" in e), int("kd_factor_range" in e)
        c.has_kp_range = int((not self.pls_enable) and "kp_range" in e)
        c.has_motor_strength, c.has_gravity_offset = int("motor_strength_range" in e), int("gravity_offset_range" in e)
        c.has_delay = 1
        if "init_pos_z_range" in e:
            c.has_init_z, c.init_z_lo, c.init_z_hi = 1, float(e["init_pos_z_range"][0]), float(e["init_pos_z_range"][1])
        if "init_euler_range" in e:
            c.has_init_euler = 1
            c.init_euler_lo, c.init_euler_hi = math.radians(e["init_euler_range"][0]), math.radians(e["init_euler_range"][1])
        self._fill_flavour(c)
        self._bind(c)
        self._after_bind()
        b = self.core.buf
        self.privileged_obs_buf = b["priv"] if self.num_privileged_obs else self.obs_buf
        self._applied_actions = b["applied"]
        self._action_history = b["hist"].view(self.num_envs, self._delay_buf_size, self.num_actions)
        self._delay_steps = b["delay"].view(-1)
        self._kp_factors, self._kd_factors, self._motor_strength = b["kp_factors"], b["kd_factors"], b["motor_strength"]
        self._effective_kp, self._effective_kd = b["eff_kp"], b["eff_kd"]
        self._gravity_offset, self._current_push_force = b["grav_off"], b["cur_push"]
        self._push_stored_force, self._push_remaining = b["push_force"], b["push_remain"].view(-1)
        self._target_dof_pos = b["target"]
        if self._has_foot_contact_rewards:
            self._foot_contact = b["foot_contact"].view(torch.bool)
            self._last_foot_contact = b["last_foot_contact"].view(torch.bool)
            self._feet_air_time = b["air_time"]
        # curriculum + DR ranges (go2_env_walk.py:457-514)
        self.curr_cfg = e.get("curriculum", {}) or {}
        cc = self.curr_cfg
        self.curriculum = CurriculumManager(cc, self._np_rng)
        self._curr_ep_total, self._curr_timeout_total, self._curr_tracking_sum, self._curr_tracking_n = 0, 0.0, 0.0, 0
        self.curr_update_every_episodes = int(cc.get("update_every_episodes", 2048))
        kp_nom, kd_nom = float(e.get("kp", 60.0)), float(e.get("kd", 2.0))
        self.friction_easy = cc.get("friction_easy", [0.6, 0.9]); self.friction_hard = e.get("friction_range", self.friction_easy)
        self.kp_easy = cc.get("kp_easy", [0.9 * kp_nom, 1.1 * kp_nom]); self.kp_hard = e.get("kp_range", self.kp_easy)
        self.kd_easy = cc.get("kd_easy", [0.75 * kd_nom, 1.25 * kd_nom]); self.kd_hard = e.get("kd_range", self.kd_easy)
        self.kp_factor_easy = cc.get("kp_factor_easy", [0.95, 1.05]); self.kp_factor_hard = e.get("kp_factor_range", self.kp_factor_easy)
        self.kd_factor_easy = cc.get("kd_factor_easy", [0.85, 1.15]); self.kd_factor_hard = e.get("kd_factor_range", self.kd_factor_easy)
        self.mass_easy = cc.get("mass_shift_easy", [-0.2, 0.5]); self.mass_hard = e.get("mass_shift_range", self.mass_easy)
        self.com_easy = cc.get("com_shift_easy", [-0.005, 0.005]); self.com_hard = e.get("com_shift_range", self.com_easy)
        self.gravity_offset_easy = cc.get("gravity_offset_easy", [-0.2, 0.2]); self.gravity_offset_hard = e.get("gravity_offset_range", self.gravity_offset_easy)
        self.motor_strength_easy = cc.get("motor_strength_easy", [0.97, 1.03]); self.motor_strength_hard = e.get("motor_strength_range", self.motor_strength_easy)
        self.leg_mass_easy = cc.get("leg_mass_shift_easy", [-0.1, 0.1]); self.leg_mass_hard = e.get("leg_mass_shift_range", self.leg_mass_easy)
        self.delay_easy_max = int(cc.get("delay_easy_max_steps", 1))
        self.push_start = float(cc.get("push_start", 0.30))
        self.push_interval_s_easy = float(cc.get("push_interval_easy_s", 6.0))
        self._global_dr_update_interval = int(cc.get("global_dr_update_interval", 200))
        self._cand_override = None  # fixture replay: {"friction": f, "mass": m, "com": [3], "leg": [4]} for the next tick
        self._kp_genesis, self._kd_genesis = (0.0, 0.0) if self._use_manual_pd else (kp_nom, kd_nom)
        self._obs_noise_level_cur = self._action_noise_std_cur = 0.0
        self._push_enable_cur = False
        self._push_interval = int(self.push_interval_s_easy / self.dt) if self.push_force_range_hard is not None else 1
        self._push_force_range_cur = [0.0, 0.0]
        self._delay_max_cur = self.delay_easy_max
        self.obs_noise_vec = None
        self._apply_curriculum_level(force=True)
        self.robot = _RobotFacade(self)
        self.ground = _GroundFacade(self)
        self.scene = _SceneFacade(self)

    def _fill_flavour(self, c):
        """Hook for flavours built on the walk env (stair): extra EnvConfig fields before the handle is created."""

    def _after_bind(self):
        """Hook: the native handle exists (upload terrain etc.)."""

    def _dr_level(self, lvl):
        """DR difficulty as a function of the curriculum level (walk: identical)."""
        return lvl

    # ---- curriculum-driven runtime values (go2_env_walk.py:611-686)
    def _rebuild_obs_noise_vec(self, level):
        if self.obs_noise_components is None or self.obs_noise_level_max <= 0.0:
            self.obs_noise_vec = None
            return
        nl = _lerp(0.0, self.obs_noise_level_max, level)
        nc, s = self.obs_noise_components, self.obs_scales
        nv = np.zeros(self.num_obs, dtype=np.float32)
        nv[0:3] = nc.get("ang_vel", 0.0) * s["ang_vel"] * nl
        nv[3:6] = nc.get("gravity", 0.0) * nl
        nv[9:21] = nc.get("dof_pos", 0.0) * s["dof_pos"] * nl
        nv[21:33] = nc.get("dof_vel", 0.0) * s["dof_vel"] * nl
        self.obs_noise_vec = nv
        self._obs_noise_c = (C.c_float * abi.NOBS_MAX)(*[float(v) for v in nv[:abi.NOBS_MAX]])  # copied into StepParams in one go

    def _apply_curriculum_level(self, force=False):
        if not self.curriculum.enabled and not force:
            return
        lvl = float(self.curriculum.level) if self.curriculum.enabled else 1.0
        cmd_lvl = lvl
        lvl = self._dr_level(lvl)  # stair: two-phase DR schedule; the command curriculum keeps the terrain level
        self._obs_noise_level_cur = _lerp(0.0, self.obs_noise_level_max, lvl)
        self._action_noise_std_cur = _lerp(0.0, self.action_noise_std_max, lvl)
        self._rebuild_obs_noise_vec(lvl)
        if self.push_force_range_hard is None:
            self._push_enable_cur, self._push_force_range_cur, self._push_interval = False, [0.0, 0.0], 10**9
        elif lvl < self.push_start:
            self._push_enable_cur, self._push_force_range_cur = False, [0.0, 0.0]
            self._push_interval = int(self.push_interval_s_easy / self.dt)
        else:
            s = _clamp01((lvl - self.push_start) / max(1e-6, (1.0 - self.push_start)))
            lo, hi = float(self.push_force_range_hard[0]), float(self.push_force_range_hard[1])
            self._push_force_range_cur = [lo * s, hi * s]
            self._push_interval = max(1, int(_lerp(self.push_interval_s_easy, self.push_interval_s_hard, s) / self.dt))
            self._push_enable_cur = True
        self._delay_max_cur = int(round(_lerp(float(self.delay_easy_max), float(self._max_delay_steps), lvl)))
        if self._cmd_curriculum:
            frac = _lerp(self._cmd_start_frac, 1.0, cmd_lvl)
            for k, fr in self._cmd_full_ranges.items():
                ctr, half = (fr[0] + fr[1]) / 2.0, (fr[1] - fr[0]) / 2.0
                self._cmd_cur_ranges[k] = [ctr - half * frac, ctr + half * frac]
        else:
            self._cmd_cur_ranges = {k: list(v) for k, v in self._cmd_full_ranges.items()}
        cur = self.extras.setdefault("curriculum", {})
        cur.update(self.curriculum.state_dict())
        cur.update({"obs_noise_level_cur": float(self._obs_noise_level_cur), "action_noise_std_cur": float(self._action_noise_std_cur),
                    "push_enable": bool(self._push_enable_cur), "push_force_range_cur": list(self._push_force_range_cur),
                    "push_interval_steps": int(self._push_interval), "delay_max_cur": int(self._delay_max_cur),
                    "cmd_ranges": {k: list(v) for k, v in self._cmd_cur_ranges.items()}})
        self._curriculum_extras(cur, cmd_lvl, lvl)

    def _curriculum_extras(self, cur, lvl, dr_lvl):
        """Hook: flavour-specific entries of extras['curriculum']."""

    def _sample_t(self):
        """DR interpolation parameter of this tick's resets (go2_env_walk.py:1162)."""
        return self.curriculum.sample_level()

    # ---- per-step parameter block
    def _make_params(self, physics=True):
        sp = abi.StepParams()
        sp.step_count = self._step_count & 0xFFFFFFFF
        sp.do_physics = 1 if physics else 0
        sp.hist_write_idx = self._action_history_write_idx
        if physics:
            self._action_history_write_idx = (self._action_history_write_idx + 1) % self._delay_buf_size
        sp.push_enable = int(self.push_force_range_hard is not None and self._push_enable_cur)
        if physics and sp.push_enable:
            sp.push_resample = int(self._push_counter % self._push_interval == 0)
            self._push_counter += 1
        sp.push_lo, sp.push_hi = float(self._push_force_range_cur[0]), float(self._push_force_range_cur[1])
        sp.action_noise_std = float(self._action_noise_std_cur)
        if self.obs_noise_vec is not None:
            sp.obs_noise_on = 1
            sp.obs_noise = self._obs_noise_c  # the proprioceptive entries (a lidar scan appended to the obs carries its own noise model)
        for i, k in enumerate(("lin_vel_x", "lin_vel_y", "ang_vel")):
            sp.cmd_lo[i], sp.cmd_hi[i] = float(self._cmd_cur_ranges[k][0]), float(self._cmd_cur_ranges[k][1])
        t = self._t_sample = self._sample_t()
        sp.kpf_lo, sp.kpf_hi = _lerp_range(self.kp_factor_easy, self.kp_factor_hard, t)
        sp.kdf_lo, sp.kdf_hi = _lerp_range(self.kd_factor_easy, self.kd_factor_hard, t)
        sp.kp_lo, sp.kp_hi = _lerp_range(self.kp_easy, self.kp_hard, t)
        sp.kd_lo, sp.kd_hi = _lerp_range(self.kd_easy, self.kd_hard, t)
        sp.ms_lo, sp.ms_hi = _lerp_range(self.motor_strength_easy, self.motor_strength_hard, t)
        sp.go_lo, sp.go_hi = _lerp_range(self.gravity_offset_easy, self.gravity_offset_hard, t)
        sp.delay_lo = self._min_delay_steps
        sp.delay_hi = max(self._min_delay_steps, min(self._delay_max_cur, self._max_delay_steps))
        sp.kp_genesis, sp.kd_genesis = float(self._kp_genesis), float(self._kd_genesis)
        # candidates of the global physics DR: what reset_idx would draw if an env resets in this tick
        # (go2_env_walk.py:737-756,803-848); the post-pass kernel decides on the device
        e, rng, ov = self.env_cfg, self._np_rng, self._cand_override or {}

        def draw(easy, hard):
            lo, hi = _lerp_range(easy, hard, t)
            return float((hi - lo) * rng.random() + lo)

        if "friction_range" in e:
            sp.cand_friction = float(ov["friction"]) if "friction" in ov else draw(self.friction_easy, self.friction_hard)
        if "mass_shift_range" in e:
            sp.cand_mass = float(ov["mass"]) if "mass" in ov else draw(self.mass_easy, self.mass_hard)
        if "com_shift_range" in e:
            for j in range(3):
                sp.cand_com[j] = float(ov["com"][j]) if "com" in ov else draw(self.com_easy, self.com_hard)
        if self._has_leg_mass_dr:
            for j in range(4):
                sp.cand_leg[j] = float(ov["leg"][j]) if "leg" in ov else draw(self.leg_mass_easy, self.leg_mass_hard)
        self._cand_override = None
        sp.replay_on = int(self._replay_next)
        sp.ext_on = int(self._ext_on)
        self._flavour_params(sp)
        self._step_count += 1
        return sp

    def _flavour_params(self, sp):
        """Hook: flavour-specific StepParams fields."""

    # ---- host side of reset_idx: curriculum aggregates, global DR, episode log (go2_env_walk.py:688-756,803-848,1228-1235)
    def _after_resets(self, stats):
        n = int(round(float(stats[abi.STATS_RESETS])))
        if n <= 0:
            return
        if self.curriculum.enabled:
            self._curr_ep_total += n
            self._curr_timeout_total += float(stats[abi.STATS_TIMEOUTS])
            self._curr_tracking_sum += float(stats[abi.STATS_TRACKING])
            self._curr_tracking_n += n
            if self._curr_ep_total >= self.curr_update_every_episodes:
                tr = float(self._curr_timeout_total / max(1, self._curr_ep_total))
                if self.curriculum.update(tr, float(self._curr_tracking_sum / max(1, self._curr_tracking_n)), float(1.0 - tr)):
                    self._apply_curriculum_level()
                self._curr_ep_total, self._curr_timeout_total, self._curr_tracking_sum, self._curr_tracking_n = 0, 0.0, 0.0, 0
        self.extras["episode"] = {"rew_" + name: float(stats[abi.STATS_REW0 + k]) / n for k, name in enumerate(self._reward_names)}

    def _critic_obs(self):
        return self.privileged_obs_buf if self.num_privileged_obs else self.obs_buf

    def get_privileged_observations(self):
        return self.privileged_obs_buf if self.num_privileged_obs is not None else None

    def step(self, actions):
        sp = self._make_params(physics=True)
        self._launch(actions, sp)
        self.actions = self.last_actions  # after the tick last_actions holds the clipped policy output
        # a fresh tensor every step, like the reference's `self.obs_buf = torch.cat(...)` (go2_env_walk.py:1084): rsl-rl keeps
        # the reference to obs_t across env.step() until it stores the transition
        self.obs_buf = self._obs_view.clone()
        self._stats_tick()
        self.extras["observations"]["critic"] = self._critic_obs()
        return self.obs_buf, self.rew_buf, self.reset_buf, self.extras

    def reset(self):
        sp = self._make_params(physics=False)
        self._launch(None, sp)
        self.flush_stats()
        return self.obs_buf, None


class Go2EnvStair(Go2EnvWalk):
    """Stair env (go2_env_stair.py): the walk env on a stair heightfield - difficulty rows with frontier-heavy
    assignment, terrain-aware spawn, terrain-relative base-height / foot-clearance rewards, lin_vel_z dead zone,
    orientation_roll_only, forward_progress, two-phase DR schedule, 182-d privileged obs (terrain row + 11x7 scan)."""

    KIND = abi.ENV_STAIR

    @staticmethod
    def _build_terrain(terrain_cfg):
        from .terrain import build_stair_terrain

        return build_stair_terrain(terrain_cfg)

    def _fill_flavour(self, c):
        e = self.env_cfg
        self._terrain_cfg = e.get("terrain", None)
        self._use_terrain = self._terrain_cfg is not None and self._terrain_cfg.get("enabled", False)
        self._lock_terrain_rows = False
        dr = e.get("dr_schedule", None)
        self._dr_enabled = dr is not None
        if dr is not None:
            self._dr_phase1_level = float(dr.get("phase1_level", 0.15))
            self._dr_terrain_gate = float(dr.get("terrain_gate", 0.85))
        c.lin_vel_z_deadzone = float(self.reward_cfg.get("lin_vel_z_deadzone", 0.0))
        if self._use_terrain:
            self._hf, self._terrain_info = self._build_terrain(self._terrain_cfg)
            ti = self._terrain_info
            self._num_terrain_rows = ti["num_difficulty_rows"]
            if self._num_terrain_rows > abi.MAX_TERRAIN_ROWS:
                raise ValueError(f"at most {abi.MAX_TERRAIN_ROWS} terrain difficulty rows are supported")
            self._terrain_row_centers, self._terrain_step_heights = ti["row_centers"], ti["step_heights_m"]
            self._terrain_origin, self._h_scale, self._v_scale = ti["terrain_origin"], ti["horizontal_scale"], ti["vertical_scale"]
            hs = self._terrain_cfg.get("height_scan", {})
            self._scan_nx, self._scan_ny = int(hs.get("num_x", 11)), int(hs.get("num_y", 7))
            self._scan_n = self._scan_nx * self._scan_ny
            xr, yr = hs.get("x_range", [-0.5, 0.5]), hs.get("y_range", [-0.3, 0.3])
            c.has_terrain, c.n_terrain_rows = 1, self._num_terrain_rows
            for r, ctr in enumerate(self._terrain_row_centers):
                for j in range(3):
                    c.row_center[r][j] = float(ctr[j])
            c.scan_nx, c.scan_ny = self._scan_nx, self._scan_ny
            c.scan_x0, c.scan_x1, c.scan_y0, c.scan_y1 = float(xr[0]), float(xr[1]), float(yr[0]), float(yr[1])
        else:
            self._num_terrain_rows, self._terrain_row_centers, self._scan_n = 1, [(0.0, 0.0, 0.0)], 0
            c.n_terrain_rows = 1

    def _after_bind(self):
        b = self.core.buf
        self._env_terrain_row = b["terrain_row"].view(-1)
        self._last_base_pos_x = b["last_x"].view(-1)
        if self._use_terrain:
            self._hf_heights = torch.tensor(self._hf.astype(np.float32) * np.float32(self._v_scale), device=self.device)
            self.core.set_terrain(self._hf.astype(np.float32) * np.float32(self._v_scale), self._h_scale, self._terrain_origin)

    # go2_env_stair.py:972-988
    def _get_dr_level(self, terrain_level):
        if not self._dr_enabled:
            return terrain_level
        if terrain_level < self._dr_terrain_gate:
            return self._dr_phase1_level
        prog = _clamp01((terrain_level - self._dr_terrain_gate) / max(1e-6, 1.0 - self._dr_terrain_gate))
        return _lerp(self._dr_phase1_level, 1.0, prog)

    _dr_level = _get_dr_level

    def _sample_t(self):  # go2_env_stair.py:1506-1511: DR level of the resets follows the two-phase schedule
        return self._get_dr_level(self.curriculum.level if self.curriculum.enabled else 1.0)

    def _curriculum_extras(self, cur, lvl, dr_lvl):
        cur["terrain_max_row"] = int(lvl * (self._num_terrain_rows - 1)) if self._use_terrain else 0
        cur["dr_level"] = float(dr_lvl)

    def _flavour_params(self, sp):
        lvl = float(self.curriculum.level) if self.curriculum.enabled else 1.0
        mr = int(lvl * (self._num_terrain_rows - 1))
        sp.terrain_max_row = max(0, min(mr, self._num_terrain_rows - 1))
        sp.lock_terrain_rows = int(self._lock_terrain_rows)
        # reset(): every env of every rank resets, the count the row split depends on is known up front; in a regular tick the
        # post-pass kernel redraws the rows of the reset envs once the tick's (all-rank) reset count is known
        sp.n_resets_hint = 0 if sp.do_physics else self._total_envs

    def _after_resets(self, stats):
        super()._after_resets(stats)
        if self._use_terrain and int(round(float(stats[abi.STATS_RESETS]))) > 0 and "episode" in self.extras:
            self.extras["episode"]["terrain_mean_row"] = float(self._env_terrain_row.float().mean())  # over all envs (not only the reset ones)

    # go2_env_stair.py:758-770
    def _get_terrain_height(self, x, y):
        if not self._use_terrain:
            return torch.zeros_like(x)
        ox, oy, _ = self._terrain_origin
        col = ((x - ox) / self._h_scale).long().clamp(0, self._hf_heights.shape[0] - 1)
        row = ((y - oy) / self._h_scale).long().clamp(0, self._hf_heights.shape[1] - 1)
        return self._hf_heights[col, row]


class LidarScanConfig:
    """Body-frame sample points and noise parameters of the simulated lidar (go2_env_stair_lidar.py:298-383, same cfg keys and
    attribute names): `actor_scan` = explicit `x_points` / `y_points` or `num_x` / `num_y` / `x_range` / `y_range` (default 5 x 3,
    forward-biased), `critic_scan` = dense surround grid (default 9 x 5), `noise`, `height_clip`, `height_scale`."""

    def __init__(self, cfg: dict, device=None):
        self.device = device
        self.enabled = bool(cfg.get("enabled", True))
        self.num_actor_scan = self.num_critic_scan = 0
        if not self.enabled:
            return
        a = cfg.get("actor_scan", {})
        if "x_points" in a:
            ax, ay = np.asarray(a["x_points"], np.float32), np.asarray(a["y_points"], np.float32)
        else:
            xr, yr = a.get("x_range", [0.0, 0.8]), a.get("y_range", [-0.15, 0.15])
            ax = torch.linspace(float(xr[0]), float(xr[1]), int(a.get("num_x", 5))).numpy()
            ay = torch.linspace(float(yr[0]), float(yr[1]), int(a.get("num_y", 3))).numpy()
        self._actor_nx, self._actor_ny = len(ax), len(ay)
        gx, gy = np.meshgrid(ax, ay, indexing="ij")
        self.actor_local_x, self.actor_local_y = gx.reshape(-1).astype(np.float32), gy.reshape(-1).astype(np.float32)
        self.num_actor_scan = self._actor_nx * self._actor_ny
        c = cfg.get("critic_scan", {})
        self._critic_nx, self._critic_ny = int(c.get("num_x", 9)), int(c.get("num_y", 5))
        xr, yr = c.get("x_range", [-0.4, 0.8]), c.get("y_range", [-0.3, 0.3])
        cx = torch.linspace(float(xr[0]), float(xr[1]), self._critic_nx).numpy()  # torch.linspace like the reference (f32 rounding)
        cy = torch.linspace(float(yr[0]), float(yr[1]), self._critic_ny).numpy()
        gx, gy = np.meshgrid(cx, cy, indexing="ij")
        self.critic_local_x, self.critic_local_y = gx.reshape(-1).astype(np.float32), gy.reshape(-1).astype(np.float32)
        self.num_critic_scan = self._critic_nx * self._critic_ny
        n = cfg.get("noise", {})
        self.height_noise_std = float(n.get("height_noise_std", 0.02))
        self.dropout_prob = float(n.get("dropout_prob", 0.15))
        self.full_blackout_prob = float(n.get("full_blackout_prob", 0.05))
        self.latency_prob = float(n.get("latency_prob", 0.1))
        self.vertical_offset_std = float(n.get("vertical_offset_std", 0.01))
        self.height_clip = float(cfg.get("height_clip", 1.0))
        self.height_scale = float(cfg.get("height_scale", 1.0))


class Go2EnvStairLidar(Go2EnvStair):
    """Stair env with the simulated lidar of examples/locomotion/go2_env_stair_lidar.py: a sparse forward-biased height
    grid with a sim-to-real noise model appended to the ACTOR obs (49 + 15 = 64), a clean dense grid at the tail of the critic's
    privileged obs instead of the 11 x 7 height scan.  The "lidar" is a heightfield lookup in the reference too (no ray casting):
    `_compute_lidar_scan` :895-937, `_update_actor_lidar_scan` :939-1002, `_update_critic_lidar_scan` :1004-1015."""

    def _fill_flavour(self, c):
        super()._fill_flavour(c)
        self._lidar = L = LidarScanConfig(self.env_cfg.get("lidar", {}), self.device)
        self._lidar_noise_level = 0.0
        if not L.enabled:
            return
        if L.num_actor_scan > abi.MAX_ACTOR_SCAN or L.num_critic_scan > abi.MAX_CRITIC_SCAN:
            raise ValueError(f"lidar grids of at most {abi.MAX_ACTOR_SCAN} actor / {abi.MAX_CRITIC_SCAN} critic points are supported")
        if self.num_obs != 33 + self.num_actions + L.num_actor_scan:
            raise ValueError(f"obs_cfg['num_obs'] = {self.num_obs} != 33 + {self.num_actions} actions + {L.num_actor_scan} lidar points")
        c.lidar_on, c.n_actor_scan, c.n_critic_scan = 1, L.num_actor_scan, L.num_critic_scan
        for i in range(L.num_actor_scan):
            c.actor_lx[i], c.actor_ly[i] = float(L.actor_local_x[i]), float(L.actor_local_y[i])
        for i in range(L.num_critic_scan):
            c.critic_lx[i], c.critic_ly[i] = float(L.critic_local_x[i]), float(L.critic_local_y[i])
        c.lidar_noise_std, c.lidar_dropout, c.lidar_blackout = L.height_noise_std, L.dropout_prob, L.full_blackout_prob
        c.lidar_latency, c.lidar_voffset_std, c.lidar_clip, c.lidar_scale = L.latency_prob, L.vertical_offset_std, L.height_clip, L.height_scale

    def _after_bind(self):
        super()._after_bind()
        self._actor_scan_buf = self.core.buf["actor_scan"]  # [B, num_actor_scan]; the previous scan until the next tick overwrites it

    @property
    def _critic_scan_buf(self):
        k = self.num_obs + 56
        return self.privileged_obs_buf[:, k:k + self._lidar.num_critic_scan]

    def _curriculum_extras(self, cur, lvl, dr_lvl):
        super()._curriculum_extras(cur, lvl, dr_lvl)
        self._lidar_noise_level = float(dr_lvl)  # go2_env_stair_lidar.py:1220
        cur["lidar_noise_level"] = float(dr_lvl)

    def _flavour_params(self, sp):
        super()._flavour_params(sp)
        sp.lidar_noise_level = float(self._lidar_noise_level)


class StairInfoConfig:
    """Parameters of the 4-value stair info (go2_env_stair_lidar_2.py:331-391, same cfg keys and attribute names): `edge_detection`
    (`max_range`, `threshold`, `num_samples`), the four observation scales and the `noise` model."""

    NUM_STAIR_INFO = 4

    def __init__(self, cfg: dict, device=None):
        self.device = device
        self.enabled = bool(cfg.get("enabled", True))
        if not self.enabled:
            return
        ed = cfg.get("edge_detection", {})
        self.edge_max_range = float(ed.get("max_range", 0.27))
        self.edge_threshold = float(ed.get("threshold", 0.02))
        self.edge_num_samples = int(ed.get("num_samples", 14))
        self.edge_dist_scale = float(cfg.get("edge_dist_scale", 1.0 / 0.27))
        self.height_scale = float(cfg.get("height_scale", 1.0 / 0.20))
        self.depth_scale = float(cfg.get("depth_scale", 1.0 / 0.30))
        self.direction_scale = float(cfg.get("direction_scale", 1.0))
        n = cfg.get("noise", {})
        self.edge_noise_std = float(n.get("edge_noise_std", 0.04))
        self.height_noise_std = float(n.get("height_noise_std", 0.01))
        self.depth_noise_std = float(n.get("depth_noise_std", 0.01))
        self.direction_flip_prob = float(n.get("direction_flip_prob", 0.03))
        self.full_blackout_prob = float(n.get("full_blackout_prob", 0.05))
        self.latency_prob = float(n.get("latency_prob", 0.10))
        self.edge_sample_x = torch.linspace(0.0, self.edge_max_range, self.edge_num_samples)  # f32 rounding of the reference


class Go2EnvStairInfo(Go2EnvStair):
    """Stair env with the 4-value stair info of examples/locomotion/go2_env_stair_lidar_2.py instead of height grids:
    [distance to the next height change on the forward axis, riser height and tread depth of the env's terrain row, direction] with
    a sim-to-real noise model appended to the ACTOR obs (49 + 4 = 53), the clean copy at the tail of the privileged obs (113); rows
    with individual tread depths (`build_stair_terrain` :83-203, numpy-global shuffle) and the `body_contact` reward (:1865-1880)."""

    @staticmethod
    def _build_terrain(terrain_cfg):
        from .terrain import build_stair_terrain_variable_depth

        return build_stair_terrain_variable_depth(terrain_cfg)

    def _fill_flavour(self, c):
        super()._fill_flavour(c)
        self._stair_cfg = S = StairInfoConfig(self.env_cfg.get("stair_info", {}), self.device)
        self._stair_noise_level = 0.0
        e = self.env_cfg
        self._has_body_contact_penalty = float(self.reward_cfg.get("reward_scales", {}).get("body_contact", 0.0)) != 0.0
        self._body_contact_link_indices = []
        if self._has_body_contact_penalty:  # _init_body_contact_tracking :1098-1137: unknown names are skipped
            names = [l["name"] for l in self._model["links"]]
            self._body_contact_link_indices = [names.index(n) for n in e.get("body_contact_names", ["FR_thigh", "FL_thigh"]) if n in names]
            self._body_contact_threshold = float(e.get("body_contact_threshold", 1.0))
            if len(self._body_contact_link_indices) > 4:
                raise ValueError("at most 4 body_contact_names are supported")
            if not self._body_contact_link_indices:
                self._has_body_contact_penalty = False
            c.n_body_contact = len(self._body_contact_link_indices)
            for i, l in enumerate(self._body_contact_link_indices):
                c.body_contact_link[i] = l
            c.body_contact_threshold = self._body_contact_threshold if self._has_body_contact_penalty else 0.0
        if not S.enabled:
            return
        if S.edge_num_samples > abi.MAX_EDGE_SAMPLES:
            raise ValueError(f"at most {abi.MAX_EDGE_SAMPLES} edge-detection samples are supported")
        if self.num_obs != 33 + self.num_actions + S.NUM_STAIR_INFO:
            raise ValueError(f"obs_cfg['num_obs'] = {self.num_obs} != 33 + {self.num_actions} actions + 4 stair-info values")
        c.lidar_on, c.n_actor_scan, c.n_critic_scan = 2, S.NUM_STAIR_INFO, S.NUM_STAIR_INFO
        c.si_n_samples = S.edge_num_samples
        sx = S.edge_sample_x
        mid = (sx[:-1] + sx[1:]) / 2
        for i in range(S.edge_num_samples):
            c.si_edge_x[i] = float(sx[i])
        for i in range(S.edge_num_samples - 1):
            c.si_edge_mid[i] = float(mid[i])
        c.si_max_range, c.si_threshold = S.edge_max_range, S.edge_threshold
        for i, v in enumerate((S.edge_dist_scale, S.height_scale, S.depth_scale, S.direction_scale)):
            c.si_scale[i] = v
        c.si_edge_noise_std, c.si_height_noise_std, c.si_depth_noise_std = S.edge_noise_std, S.height_noise_std, S.depth_noise_std
        c.si_flip_prob, c.si_blackout_prob, c.si_latency_prob = S.direction_flip_prob, S.full_blackout_prob, S.latency_prob
        if self._use_terrain:
            self._terrain_step_depths = self._terrain_info["step_depths_m"]
            hs, ds = self._terrain_step_heights, self._terrain_step_depths
        else:
            hs, ds = [0.0], [0.25]
        for r in range(len(hs)):
            c.si_row_height[r], c.si_row_depth[r] = float(np.float32(hs[r])), float(np.float32(ds[r]))

    def _after_bind(self):
        super()._after_bind()
        self._stair_info_buf = self.core.buf["actor_scan"]  # [B, 4]; the previous tick's info until the next tick overwrites it

    @property
    def _stair_info_clean(self):
        k = self.num_obs + 56
        return self.privileged_obs_buf[:, k:k + 4]

    def _curriculum_extras(self, cur, lvl, dr_lvl):
        super()._curriculum_extras(cur, lvl, dr_lvl)
        self._stair_noise_level = float(dr_lvl)  # go2_env_stair_lidar_2.py:1188
        cur["stair_noise_level"] = float(dr_lvl)

    def _flavour_params(self, sp):
        super()._flavour_params(sp)
        sp.lidar_noise_level = float(self._stair_noise_level)


class TerrainManager:
    """Host-side view of the terrain state of `Go2EnvWalkTerrain`, with the attribute / method names of the reference's
    examples/locomotion/terrain_manager.py that the env and eval scripts touch.  Both modes: a custom `height_field` (heightmap mode,
    :63-98) and the sub-terrain GRID (`subterrain_types`, :104-133,179-214) - there the field is generated here by the restatement of
    Genesis' terrain parser (`go2_b200.terrain.build_subterrain_grid`), and, like the reference's manager, which cannot read the
    field of a Genesis-generated terrain (`post_build` :216-262 finds no attribute), every height query returns 0."""

    def __init__(self, terrain_cfg: dict, num_envs: int, device=None):
        self.cfg = terrain_cfg or {}
        self.enabled = bool(self.cfg.get("enabled", False))
        self.device, self.num_envs = device, num_envs
        if not self.enabled:
            return
        self.use_custom_heightmap = self.cfg.get("height_field", None) is not None
        self.horizontal_scale = float(self.cfg.get("horizontal_scale", 0.25))
        self.vertical_scale = float(self.cfg.get("vertical_scale", 0.005))
        self.spawn_height_offset = float(self.cfg.get("spawn_height_offset", 0.05))
        self.boundary_margin = float(self.cfg.get("boundary_margin", 1.0))
        self._env_spawn_pos = None  # bound to the device buffer by the env
        if not self.use_custom_heightmap:
            from .terrain import build_subterrain_grid

            self.subterrain_types = [list(r) for r in self.cfg["subterrain_types"]]
            self.n_rows, self.n_cols = len(self.subterrain_types), len(self.subterrain_types[0])
            self.subterrain_size = tuple(self.cfg.get("subterrain_size", (8.0, 8.0)))
            self.randomize = bool(self.cfg.get("randomize", True))
            self.max_wander = min(self.subterrain_size) / 2.0 - self.boundary_margin
            self._flat_row_indices = [r for r, row in enumerate(self.subterrain_types) if all(t == "flat_terrain" for t in row)]
            self._terrain_size_x, self._terrain_size_y = self.n_rows * self.subterrain_size[0], self.n_cols * self.subterrain_size[1]
            self._terrain_origin = np.array([-self._terrain_size_x / 2.0, -self._terrain_size_y / 2.0, 0.0])
            self._tile_centers = torch.zeros((self.n_rows, self.n_cols, 3), dtype=torch.float32)
            for r in range(self.n_rows):
                for c in range(self.n_cols):
                    self._tile_centers[r, c, 0] = self._terrain_origin[0] + (r + 0.5) * self.subterrain_size[0]
                    self._tile_centers[r, c, 1] = self._terrain_origin[1] + (c + 0.5) * self.subterrain_size[1]
            # collision field (gs.morphs.Terrain(n_subterrains=..., subterrain_types=...) -> genesis/utils/terrain.py parse_terrain)
            self._height_field_np = build_subterrain_grid(self.subterrain_types, self.subterrain_size, self.horizontal_scale, self.vertical_scale,
                                                          self.randomize, self.cfg.get("subterrain_parameters", None))
            self._height_field = None  # what the manager itself sees
            return
        self._height_field_np = np.array(self.cfg["height_field"], dtype=np.float32)  # terrain_manager.py:63-76
        rows, cols = self._height_field_np.shape
        self._terrain_size_x, self._terrain_size_y = rows * self.horizontal_scale, cols * self.horizontal_scale
        self.max_wander = min(self._terrain_size_x, self._terrain_size_y) / 2.0 - self.boundary_margin
        self._terrain_origin = np.array([-self._terrain_size_x / 2.0, -self._terrain_size_y / 2.0, 0.0])
        self._height_field = self._height_field_np
        self.subterrain_types, self.n_rows, self.n_cols, self._tile_centers = None, 1, 1, None

    def _tile_of(self, k, size, n):
        o = float(self._terrain_origin[k])
        return ((self._env_spawn_pos[:, k] - o) / size).floor().long().clamp(0, n - 1)

    @property
    def _env_tile_row(self):  # tile the env spawned on (heightmap mode: 0), recovered from the spawn point
        if self.use_custom_heightmap:
            return torch.zeros(self.num_envs, dtype=torch.long, device=self._env_spawn_pos.device)
        return self._tile_of(0, self.subterrain_size[0], self.n_rows)

    @property
    def _env_tile_col(self):
        if self.use_custom_heightmap:
            return torch.zeros(self.num_envs, dtype=torch.long, device=self._env_spawn_pos.device)
        return self._tile_of(1, self.subterrain_size[1], self.n_cols)

    def get_env_tile_rows(self):
        return self._env_tile_row

    def _query_height_batch(self, world_x, world_y):
        """Bilinear ground height (terrain_manager.py:354-392), float64 numpy arithmetic like the reference."""
        hf = self._height_field
        if hf is None:
            return torch.zeros(len(world_x), device=world_x.device, dtype=torch.float32)
        R, Cc = hf.shape
        lx = np.clip((np.asarray(world_x.detach().cpu()) - self._terrain_origin[0]) / self.horizontal_scale, 0, R - 2).astype(np.float64)
        ly = np.clip((np.asarray(world_y.detach().cpu()) - self._terrain_origin[1]) / self.horizontal_scale, 0, Cc - 2).astype(np.float64)
        ix, iy = lx.astype(np.int32), ly.astype(np.int32)
        fx, fy = lx - ix, ly - iy
        ix1, iy1 = np.minimum(ix + 1, R - 1), np.minimum(iy + 1, Cc - 1)
        h = hf[ix, iy] * (1 - fx) * (1 - fy) + hf[ix1, iy] * fx * (1 - fy) + hf[ix, iy1] * (1 - fx) * fy + hf[ix1, iy1] * fx * fy
        return torch.tensor(h * self.vertical_scale, device=world_x.device, dtype=torch.float32)

    def get_height_at_robot(self, base_pos_xy):
        return self._query_height_batch(base_pos_xy[:, 0], base_pos_xy[:, 1])

    def check_boundary(self, base_pos_xy):
        return torch.norm(base_pos_xy - self._env_spawn_pos[:, :2], dim=1) > self.max_wander

    def get_terrain_type_string(self, env_idx: int) -> str:
        if self.use_custom_heightmap:
            return "custom_heightmap"
        return self.subterrain_types[int(self._env_tile_row[env_idx])][int(self._env_tile_col[env_idx])]


class Go2EnvWalkTerrain(Go2EnvWalk):
    """Walk env on a custom heightmap: examples/locomotion/go2_env_Omni_kplearn_varied_terrain.py + terrain_manager.py in heightmap mode
    (SURVEY 8(f)-4).  The heightfield replaces the ground plane (same terrain narrow phase as the stair env); resets spawn the robot at a
    random point of the map on top of the interpolated ground, an env is also reset when it wanders further than `max_wander` from
    its spawn point, and the base-height reward is measured above the ground under the base."""

    def _fill_flavour(self, c):
        super()._fill_flavour(c)
        self.terrain = T = TerrainManager(self.env_cfg.get("terrain", {}), self.num_envs, self.device)
        if not T.enabled:
            return
        c.tm_on = 1
        c.tm_max_wander, c.tm_spawn_offset = float(T.max_wander), float(T.spawn_height_offset)
        if T.use_custom_heightmap:
            c.tm_safe_half_x = float(T._terrain_size_x / 2.0 - T.boundary_margin - 0.5)
            c.tm_safe_half_y = float(T._terrain_size_y / 2.0 - T.boundary_margin - 0.5)
        else:  # grid mode: random tile, centre + jitter (terrain_manager.py:308-335); height queries see nothing
            c.tm_grid_rows, c.tm_grid_cols, c.tm_blind = T.n_rows, T.n_cols, 1
            c.tm_tile_x, c.tm_tile_y = float(T.subterrain_size[0]), float(T.subterrain_size[1])
            c.tm_safe_half_x = float(T.subterrain_size[0] / 2.0 - T.boundary_margin - 0.5)
            c.tm_safe_half_y = float(T.subterrain_size[1] / 2.0 - T.boundary_margin - 0.5)
        c.tm_hs, c.tm_ox, c.tm_oy = float(T.horizontal_scale), float(T._terrain_origin[0]), float(T._terrain_origin[1])

    def _after_bind(self):
        super()._after_bind()
        T = self.terrain
        if T.enabled:
            self.core.set_terrain(T._height_field_np * np.float32(T.vertical_scale), T.horizontal_scale, T._terrain_origin)
            T._env_spawn_pos = self.core.buf["spawn_pos"]
            self._ground_height = self.core.buf["ground_h"].view(-1)
        else:
            self._ground_height = torch.zeros(self.num_envs, device=self.device)


class Go2EnvTerrain(Go2EnvWalkTerrain):
    """examples/locomotion/go2_env_terrain.py: the height-map walk env plus the ground height under the base as one more privileged-obs
    entry (:1139-1141), foot clearance measured above that ground height (:1355-1363), the `stumble` reward (contact force on
    non-foot links above `stumble_threshold`, :1380-1394) and `terrain_row_mean` in the episode log."""

    def _fill_flavour(self, c):
        super()._fill_flavour(c)
        c.tm_variant = 1
        c.stumble_threshold = float(self.reward_cfg.get("stumble_threshold", 5.0))

    def _after_resets(self, stats):
        super()._after_resets(stats)
        if self.terrain.enabled and int(round(float(stats[abi.STATS_RESETS]))) > 0 and "episode" in self.extras:
            T = self.terrain  # go2_env_terrain.py:1242-1244: mean tile row of the envs reset in this tick (heightmap mode: row 0)
            m = self.reset_buf.bool()
            self.extras["episode"]["terrain_row_mean"] = 0.0 if (T.use_custom_heightmap or not bool(m.any())) else float(T._env_tile_row[m].float().mean())


class Go2EnvBaseFlavour(_Go2EnvBase):
    """Genesis stock Go2 env used by the crouch / jump scripts (go2_env_base.py): 12-dim action, 1-step
    latency through last_actions, solver-side PD, reset BEFORE rewards, 45-d obs, no privileged obs."""

    KIND = abi.ENV_BASE

    def __init__(self, num_envs, env_cfg, obs_cfg, reward_cfg, command_cfg, show_viewer=False, **kw):
        super().__init__(num_envs, env_cfg, obs_cfg, reward_cfg, command_cfg, show_viewer, **kw)
        self.num_privileged_obs = None
        self.simulate_action_latency = True
        c = abi.EnvConfig()
        self._fill_common(c)
        c.num_pos_actions, c.num_priv, c.ctrl = 12, 0, abi.CTRL_GENESIS_PD
        c.simulate_action_latency = 1
        c.hist_size = 1
        c.compound_commands = 1
        for j in range(12):
            c.torque_limit[j] = 1e30
        self._bind(c)
        self._kp_genesis, self._kd_genesis = float(env_cfg["kp"]), float(env_cfg["kd"])
        self.robot = _RobotFacade(self)
        self.ground = _GroundFacade(self)
        self.scene = _SceneFacade(self)

    def _make_params(self, physics=True):
        sp = abi.StepParams()
        sp.step_count = self._step_count & 0xFFFFFFFF
        sp.do_physics = 1 if physics else 0
        cc = self.command_cfg
        for i, k in enumerate(("lin_vel_x_range", "lin_vel_y_range", "ang_vel_range")):
            sp.cmd_lo[i], sp.cmd_hi[i] = float(cc[k][0]), float(cc[k][1])
        sp.kp_genesis, sp.kd_genesis = float(self._kp_genesis), float(self._kd_genesis)
        sp.delay_lo = sp.delay_hi = 0
        sp.replay_on = int(self._replay_next)
        sp.ext_on = int(self._ext_on)
        self._step_count += 1
        return sp

    def _after_resets(self, stats):
        n = int(round(float(stats[abi.STATS_RESETS])))
        if n <= 0:
            return
        ep_s = float(self.env_cfg["episode_length_s"])  # go2_env_base.py:231-236: mean(sum) / episode_length_s
        self.extras["episode"] = {"rew_" + name: float(stats[abi.STATS_REW0 + k]) / n / ep_s for k, name in enumerate(self._reward_names)}

    def get_privileged_observations(self):
        return None

    def step(self, actions):
        sp = self._make_params(physics=True)
        self._launch(actions, sp)
        self.actions = self.last_actions
        self.obs_buf = self._obs_view.clone()  # fresh tensor per step (go2_env_base.py:175 re-binds obs_buf with torch.cat)
        self._stats_tick()
        self.extras["observations"]["critic"] = self.obs_buf
        return self.obs_buf, self.rew_buf, self.reset_buf, self.extras

    def reset(self):
        sp = self._make_params(physics=False)
        self._launch(None, sp)
        self.flush_stats()
        return self.obs_buf, None


# ---------------------------------------------------------------------------------------------------
# Facades for the scene-level attribute surface the eval scripts poke: env.robot.*,
# env.ground.set_friction, env.scene.sim.rigid_solver.apply_links_external_force.  They edit the same
# device buffers / host scalars the step kernel reads.
# ---------------------------------------------------------------------------------------------------
class _Link:
    def __init__(self, name, idx):
        self.name, self.idx = name, idx


class _Joint:
    def __init__(self, name, dof_start):
        self.name, self.dof_start = name, dof_start


class _RobotFacade:
    def __init__(self, env):
        self._env = env
        self.links = [_Link(l["name"], i + 1) for i, l in enumerate(env._model["links"])]  # .idx is global (ground = 0)
        self._joints = {l["joint_name"]: _Joint(l["joint_name"], l["dof_start"]) for l in env._model["links"]}

    def get_joint(self, name):
        return self._joints[name]

    def _idx(self, envs_idx):
        e = self._env
        return torch.arange(e.num_envs, device=e.device) if envs_idx is None else torch.as_tensor(envs_idx, device=e.device).long()

    def _refreshed(self, ii):
        self._env.core.buf["fk_fresh"][ii] = 1  # Genesis refreshes the env's FK in every state setter (rigid_solver.py:1922-1937)

    def zero_all_dofs_velocity(self, envs_idx=None):
        ii = self._idx(envs_idx)
        self._env.core.buf["qvel"][ii] = 0.0
        self._refreshed(ii)

    def set_dofs_position(self, position, dofs_idx_local=None, zero_velocity=True, envs_idx=None):
        ii = self._idx(envs_idx)
        dofs = list(range(18)) if dofs_idx_local is None else list(dofs_idx_local)
        pos = torch.as_tensor(position, device=self._env.device, dtype=torch.float32).reshape(len(ii), len(dofs))
        q = self._env.core.buf["qpos"]
        for k, d in enumerate(dofs):
            q[ii, d + 1] = pos[:, k]
        self._refreshed(ii)
        if zero_velocity:
            self.zero_all_dofs_velocity(envs_idx)

    def set_pos(self, pos, zero_velocity=True, envs_idx=None):
        ii = self._idx(envs_idx)
        p = torch.as_tensor(pos, device=self._env.device, dtype=torch.float32).reshape(len(ii), 3)
        self._env.core.buf["qpos"][ii, 0:3] = p
        self._refreshed(ii)
        if zero_velocity:
            self.zero_all_dofs_velocity(envs_idx)

    def set_quat(self, quat, zero_velocity=True, envs_idx=None):
        ii = self._idx(envs_idx)
        q = torch.as_tensor(quat, device=self._env.device, dtype=torch.float32).reshape(len(ii), 4)
        self._env.core.buf["qpos"][ii, 3:7] = q
        self._refreshed(ii)
        if zero_velocity:
            self.zero_all_dofs_velocity(envs_idx)

    def get_pos(self):
        return self._env.core.buf["qpos"][:, 0:3].contiguous()

    def get_quat(self):
        q = self._env.core.buf["qpos"][:, 3:7]
        return q / q.norm(dim=1, keepdim=True)

    def get_dofs_position(self, dofs=None):
        q = self._env.core.buf["qpos"]
        return torch.stack([q[:, d + 1] for d in (dofs if dofs is not None else range(6, 18))], dim=1)

    def get_dofs_velocity(self, dofs=None):
        v = self._env.core.buf["qvel"]
        return torch.stack([v[:, d] for d in (dofs if dofs is not None else range(18))], dim=1)

    def get_dofs_control_force(self, dofs=None):
        return self._env.core.buf["torque"]

    def _links(self):
        self._env.core.export_links()

    def get_links_pos(self):
        self._links()
        return self._env.core.buf["link_pos"].view(self._env.num_envs, 13, 3)

    def get_links_vel(self):
        self._links()
        return self._env.core.buf["link_vel"].view(self._env.num_envs, 13, 3)

    def get_vel(self):
        return self.get_links_vel()[:, 0]

    def get_links_net_contact_force(self):
        return self._env.core.buf["contact_force"].view(self._env.num_envs, 13, 3)

    # global (all envs) physics parameters, Genesis semantics (rigid_geom.py:327-336, accessor.py:337-361)
    def set_friction(self, mu):
        self._env._set_global(abi.G_FRIC_ROBOT, mu)

    def set_mass_shift(self, shifts, links_idx_local):
        for s, l in zip(shifts, links_idx_local):
            self._env._set_global(abi.G_MASS_SHIFT + int(l), float(s))

    def set_COM_shift(self, shifts, links_idx_local):
        for s, l in zip(shifts, links_idx_local):
            if int(l) != 0:
                raise NotImplementedError("COM shift is supported for the base link only (all the envs use)")
            v = np.asarray(torch.as_tensor(s).cpu(), dtype=np.float32).reshape(3)
            for j in range(3):
                self._env._set_global(abi.G_COM_SHIFT + j, float(v[j]))

    def set_dofs_kp(self, kp, dofs=None):
        self._env._kp_genesis = float(kp[0])

    def set_dofs_kv(self, kv, dofs=None):
        self._env._kd_genesis = float(kv[0])


class _GroundFacade:
    def __init__(self, env):
        import types

        self._env = env
        # go2_eval_stairs.py:592-601 looks the ground plane up among scene.entities by its URDF file name
        self.morph = types.SimpleNamespace(file="urdf/plane/plane.urdf") if not getattr(env, "_use_terrain", False) else types.SimpleNamespace()

    def set_friction(self, mu):
        self._env._set_global(abi.G_FRIC_GROUND, mu)


class _SceneFacade:
    """env.scene: only `sim.rigid_solver.apply_links_external_force` (both eval scripts) is meaningful on the fused step."""

    def __init__(self, env):
        import types

        self.sim = types.SimpleNamespace(rigid_solver=self)
        self._env = env

    @property
    def entities(self):
        return [self._env.ground, self._env.robot]

    def apply_links_external_force(self, force, links_idx, envs_idx=None):
        """World-frame force at the link origin for the NEXT tick's two substeps, then cleared (abd/misc.py:655-715,
        simulator.py:283).  `links_idx`: ONE global link index (robot.links[i].idx); force [n_envs, 3] or [n_envs, 1, 3]."""
        e = self._env
        links = [int(l) for l in (links_idx.tolist() if hasattr(links_idx, "tolist") else list(links_idx))]
        if len(links) != 1:
            raise NotImplementedError("one link per apply_links_external_force call (what the envs and eval scripts use)")
        local = links[0] - 1
        if not 0 <= local < abi.NLINK:
            raise ValueError(f"link index {links[0]} is not a robot link")
        ii = torch.arange(e.num_envs, device=e.device) if envs_idx is None else torch.as_tensor(envs_idx, device=e.device).long().reshape(-1)
        f = torch.as_tensor(force, device=e.device, dtype=torch.float32).reshape(len(ii), 3)
        ex = e.core.buf["ext_force"]
        if e._ext_on:
            prev = ex[ii, 0]
            if bool(((prev >= 0) & (prev != float(local))).any()):
                raise NotImplementedError("external forces on two different links of one env in the same tick")
            f = f + ex[ii, 1:4] * (prev >= 0).unsqueeze(1).float()  # forces accumulate until the step clears them
        ex[ii, 0] = float(local)
        ex[ii, 1:4] = f
        e._ext_on = True

    def step(self):
        raise NotImplementedError("scene.step() is fused into Go2Env.step()")
