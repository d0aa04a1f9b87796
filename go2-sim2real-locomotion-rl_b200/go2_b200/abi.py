"""ctypes binding of the C ABI in include/go2_b200.h (struct mirrors of csrc/go2_types.cuh).

`Go2Lib` is deliberately dumb: it loads one shared library by path and exposes the entry points.  The
package always loads `libgo2_b200.so` (CUDA, sm_100a) and raises if it is missing - there is no CPU
fallback on the product path.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

NLINK, NV, NQ, NHINGE = 13, 18, 19, 12
MAX_REW, NOBS_MAX, MAX_CPAIRS = 32, 49, 24
MAX_TERRAIN_ROWS = 16
MAX_ACTOR_SCAN, MAX_CRITIC_SCAN = 32, 96
MAX_EDGE_SAMPLES = 32

_HERE = os.path.dirname(os.path.abspath(__file__))
CUDA_LIB_PATH = os.path.join(_HERE, "libgo2_b200.so")
# the same sources compiled with approximate division / square root (nvcc -prec-div=false -prec-sqrt=false -ftz=true): the arithmetic class
# of the reference itself, whose solver is always JIT-compiled with fast_math on (genesis/__init__.py:289 `fast_math=not debug`).
# libgo2_b200.so keeps IEEE division / sqrt and is what the bit-faithful parity against the CPU oracle is stated for.
CUDA_LIB_FAST_PATH = os.path.join(_HERE, "libgo2_b200_fast.so")

ENV_WALK, ENV_STAIR, ENV_BASE = 0, 1, 2
CTRL_MANUAL_PLS, CTRL_MANUAL_FIXED, CTRL_GENESIS_PD = 0, 1, 2

REWARD_IDS = {
    name: i
    for i, name in enumerate(
        [
            "tracking_lin_vel", "tracking_ang_vel", "lin_vel_z", "base_height", "action_rate", "similar_to_default",
            "orientation_penalty", "dof_acc", "dof_vel", "ang_vel_xy", "feet_air_time", "foot_slip", "foot_clearance",
            "joint_tracking", "energy", "torque_load", "stand_still", "stand_still_vel", "feet_stance",
            "jump_impulse", "jump_apex", "xy_stability", "orientation", "no_shake", "crouch", "crouch_2", "ground_penalty",
            "crouch_target", "no_fall", "y_stability", "crouch_progress", "crouch_speed",
            "orientation_roll_only", "forward_progress",
            "stumble",
            "body_contact",
        ]
    )
}

BUF_NAMES = [
    "qpos", "qvel", "qacc_ws", "warm", "ncache", "cache_pair", "cache_n",
    "actions_in", "obs", "priv", "rew", "reset", "timeout", "ep_len", "commands", "last_actions", "applied", "hist",
    "last_dof_vel", "dof_pos", "dof_vel", "target", "base_pos", "base_quat", "base_lin_vel", "base_ang_vel", "proj_g",
    "base_euler", "kp_factors", "kd_factors", "motor_strength", "eff_kp", "eff_kd", "grav_off", "push_force", "cur_push",
    "push_remain", "delay", "foot_contact", "last_foot_contact", "air_time", "ep_sums", "errno", "link_pos", "link_vel",
    "contact_force", "torque", "stats", "ncontact", "nnewton", "terrain_row", "last_x",
    "globals", "rew_terms", "replay", "fk_fresh", "ext_force", "actor_scan", "spawn_pos", "ground_h",
]
BUF = {n: i for i, n in enumerate(BUF_NAMES)}
STATS_RESETS, STATS_TIMEOUTS, STATS_TRACKING, STATS_ERR, STATS_REW0 = 0, 1, 2, 3, 4
# B_GLOBALS slot layout / B_REPLAY row layout (csrc/go2_types.cuh)
G_FRIC_GROUND, G_FRIC_ROBOT, G_MASS_SHIFT, G_COM_SHIFT, G_DR_COUNTER, GLOBALS_LEN = 0, 1, 2, 15, 18, 20
RP_ACT_NOISE, RP_OBS_NOISE, RP_PUSH, RP_CMD, RP_KPF, RP_KDF, RP_MS, RP_GO, RP_DELAY = 0, 12, 61, 64, 67, 79, 91, 103, 106
RP_INIT_Z, RP_ROLL, RP_PITCH, RP_CMD_RESET, RP_ROW, RP_KP, RP_KD = 107, 108, 109, 110, 113, 114, 115
RP_LIDAR_N, RP_LIDAR_U, RP_LIDAR_V = 116, 116 + MAX_ACTOR_SCAN, 116 + 2 * MAX_ACTOR_SCAN  # lidar variant: cell normals, cell uniforms, (offset, latency, blackout)
RP_SPAWN = RP_LIDAR_V + 3  # heightmap walk variant: spawn x, y uniforms
RP_TILE = RP_SPAWN + 2  # grid mode of the terrain manager: tile row, column
REPLAY_LEN = RP_TILE + 2

_i, _f, _u = C.c_int32, C.c_float, C.c_uint32


class EnvConfig(C.Structure):
    _fields_ = [
        ("kind", _i), ("ctrl", _i), ("num_envs", _i), ("env_offset", _i),
        ("num_actions", _i), ("num_pos_actions", _i), ("num_obs", _i), ("num_priv", _i),
        ("max_episode_length", _i), ("resample_steps", _i), ("hist_size", _i), ("n_rew", _i),
        ("rew_id", _i * MAX_REW), ("rew_scale", _f * MAX_REW), ("motors_dof", _i * NHINGE),
        ("default_dof_pos", _f * NHINGE), ("torque_limit", _f * NHINGE),
        ("action_scale", _f), ("clip_actions", _f), ("kp", _f), ("kd", _f),
        ("pls_kp_default", _f), ("pls_kp_scale", _f), ("pls_kp_lo", _f), ("pls_kp_hi", _f),
        ("term_roll", _f), ("term_pitch", _f), ("term_zvel", _f), ("term_yvel", _f),
        ("base_init_pos", _f * 3), ("base_init_quat", _f * 4), ("inv_base_init_quat", _f * 4),
        ("obs_scale_lin_vel", _f), ("obs_scale_ang_vel", _f), ("obs_scale_dof_pos", _f), ("obs_scale_dof_vel", _f),
        ("tracking_sigma", _f), ("base_height_target", _f), ("feet_height_target", _f), ("feet_air_time_target", _f),
        ("foot_contact_threshold", _f), ("foot_link", _i * 4), ("has_foot_tracking", _i), ("push_link", _i),
        ("push_dur_lo", _i), ("push_dur_hi", _i),
        ("has_push", _i), ("has_init_z", _i), ("has_init_euler", _i), ("has_kp_factor", _i), ("has_kd_factor", _i),
        ("has_motor_strength", _i), ("has_gravity_offset", _i), ("has_delay", _i), ("compound_commands", _i), ("n_standing", _i),
        ("init_z_lo", _f), ("init_z_hi", _f), ("init_euler_lo", _f), ("init_euler_hi", _f),
        ("min_delay", _i), ("max_delay", _i), ("has_kp_range", _i), ("simulate_action_latency", _i),
        ("rp", _f * 16), ("seed_lo", _u), ("seed_hi", _u),
        ("has_terrain", _i), ("n_terrain_rows", _i), ("row_center", (_f * 3) * MAX_TERRAIN_ROWS),
        ("scan_nx", _i), ("scan_ny", _i), ("scan_x0", _f), ("scan_x1", _f), ("scan_y0", _f), ("scan_y1", _f),
        ("lin_vel_z_deadzone", _f),
        ("has_friction_dr", _i), ("has_mass_dr", _i), ("has_com_dr", _i), ("has_leg_mass_dr", _i), ("global_dr_interval", _i),
        ("hip_link", _i * 4), ("solver_dense", _i), ("export_rew_terms", _i),  ("replay_rows", _i),
        ("lidar_on", _i), ("n_actor_scan", _i), ("n_critic_scan", _i),
        ("actor_lx", _f * MAX_ACTOR_SCAN), ("actor_ly", _f * MAX_ACTOR_SCAN), ("critic_lx", _f * MAX_CRITIC_SCAN), ("critic_ly", _f * MAX_CRITIC_SCAN),
        ("lidar_noise_std", _f), ("lidar_dropout", _f), ("lidar_blackout", _f), ("lidar_latency", _f), ("lidar_voffset_std", _f),
        ("lidar_clip", _f), ("lidar_scale", _f),
        ("si_n_samples", _i), ("si_edge_x", _f * MAX_EDGE_SAMPLES), ("si_edge_mid", _f * MAX_EDGE_SAMPLES),
        ("si_max_range", _f), ("si_threshold", _f), ("si_scale", _f * 4),
        ("si_edge_noise_std", _f), ("si_height_noise_std", _f), ("si_depth_noise_std", _f), ("si_flip_prob", _f), ("si_blackout_prob", _f),
        ("si_latency_prob", _f), ("si_row_height", _f * MAX_TERRAIN_ROWS), ("si_row_depth", _f * MAX_TERRAIN_ROWS),
        ("n_body_contact", _i), ("body_contact_link", _i * 4), ("body_contact_threshold", _f),
        ("tm_on", _i), ("tm_variant", _i), ("stumble_threshold", _f), ("tm_max_wander", _f), ("tm_safe_half_x", _f), ("tm_safe_half_y", _f), ("tm_spawn_offset", _f),
        ("tm_grid_rows", _i), ("tm_grid_cols", _i), ("tm_blind", _i),
        ("tm_hs", C.c_double), ("tm_ox", C.c_double), ("tm_oy", C.c_double), ("tm_tile_x", C.c_double), ("tm_tile_y", C.c_double),
    ]


class StepParams(C.Structure):
    _fields_ = [
        ("step_count", _u), ("hist_write_idx", _i), ("push_resample", _i), ("push_enable", _i),
        ("push_lo", _f), ("push_hi", _f), ("action_noise_std", _f), ("obs_noise", _f * NOBS_MAX), ("obs_noise_on", _i),
        ("cmd_lo", _f * 3), ("cmd_hi", _f * 3),
        ("kpf_lo", _f), ("kpf_hi", _f), ("kdf_lo", _f), ("kdf_hi", _f), ("ms_lo", _f), ("ms_hi", _f), ("go_lo", _f), ("go_hi", _f),
        ("kp_lo", _f), ("kp_hi", _f), ("kd_lo", _f), ("kd_hi", _f),
        ("delay_lo", _i), ("delay_hi", _i),
        ("kp_genesis", _f), ("kd_genesis", _f),
        ("do_physics", _i), ("force_reset_all", _i), ("terrain_max_row", _i), ("lock_terrain_rows", _i), ("lidar_noise_level", _f), ("n_resets_hint", _i),
        ("cand_friction", _f), ("cand_mass", _f), ("cand_com", _f * 3), ("cand_leg", _f * 4),
        ("parity", _i), ("defer_post", _i), ("replay_on", _i), ("ext_on", _i),
        ("host_obs", C.c_void_p), ("host_rew", C.c_void_p), ("host_done", C.c_void_p),
        ("act_flag", C.c_void_p), ("act_seq", C.c_uint),
    ]


_EXPORTS = [
    "go2_abi_sizes", "go2_create", "go2_destroy", "go2_last_error", "go2_buffer_info", "go2_step", "go2_reset_all",
    "go2_step_host", "go2_read_stats", "go2_export_links", "go2_launch_count", "go2_set_terrain",
    "go2_post", "go2_tick_parity", "go2_step_host_begin", "go2_read_stats_async", "go2_stats_fetch", "go2_launch_shape",
]
# include/go2_policy.h (CUDA library only; the host-emulation test library has no tensor cores to emulate)
_POLICY_EXPORTS = ["go2_mlp_create", "go2_mlp_destroy", "go2_mlp_forward", "go2_policy_act", "go2_policy_act_critic", "go2_gae", "go2_mlp_forward_noalloc"]
_DTYPES = {0: np.float32, 1: np.int32, 2: np.uint8}


class Go2Lib:
    """Thin loader; raises if the shared library or any declared symbol is missing."""

    def __init__(self, path=CUDA_LIB_PATH):
        if not os.path.exists(path):
            raise RuntimeError(
                f"go2_b200: native library not found at {path}. Build it with __graft_entry__.build() "
                "(nvcc -gencode arch=compute_100a,code=sm_100a). There is no CPU fallback."
            )
        self.path = path
        self.lib = C.CDLL(path)
        for name in _EXPORTS:
            if not hasattr(self.lib, name):
                raise RuntimeError(f"go2_b200: {path} does not export {name}")
        self.lib.go2_last_error.restype = C.c_char_p
        self.lib.go2_launch_count.restype = C.c_longlong
        sizes = (C.c_int * 4)()
        self.lib.go2_abi_sizes(sizes)
        if sizes[0] != C.sizeof(EnvConfig) or sizes[1] != C.sizeof(StepParams) or sizes[2] != len(BUF_NAMES):
            raise RuntimeError(f"go2_b200: ABI mismatch {list(sizes)} vs {C.sizeof(EnvConfig)}, {C.sizeof(StepParams)}, {len(BUF_NAMES)}")
        self.stats_len = sizes[3]

    def create(self, model_iv, model_fv, cfg: EnvConfig, device: int):
        h = C.c_void_p()
        iv = np.ascontiguousarray(model_iv, dtype=np.int32)
        fv = np.ascontiguousarray(model_fv, dtype=np.float64)
        rc = self.lib.go2_create(iv.ctypes.data_as(C.c_void_p), len(iv), fv.ctypes.data_as(C.c_void_p), len(fv),
                                 C.byref(cfg), C.sizeof(cfg), int(device), C.byref(h))
        if rc != 0:
            msg = self.lib.go2_last_error(h).decode() if h else ""
            raise RuntimeError(f"go2_create failed rc={rc} {msg}")
        return h

    def check(self, h, rc, what):
        if rc != 0:
            raise RuntimeError(f"{what} failed rc={rc}: {self.lib.go2_last_error(h).decode()}")

    def buffer_info(self, h, buf_id):
        ptr, dt, r, c = C.c_void_p(), C.c_int(), C.c_int(), C.c_int()
        self.check(h, self.lib.go2_buffer_info(h, int(buf_id), C.byref(ptr), C.byref(dt), C.byref(r), C.byref(c)), "go2_buffer_info")
        return ptr.value, _DTYPES[dt.value], r.value, c.value
