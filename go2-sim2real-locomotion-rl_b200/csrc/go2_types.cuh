// go2_b200 - shared POD types for the fused Go2 env-step kernel (device + host).
//
// Replaces, for the Go2 scenes only, the struct-of-fields zoo of the reference
// (genesis/utils/array_class.py: LinksInfo/DofsInfo/GeomsInfo/ColliderInfo ...) by one constant table
// (Go2Model, lives in __constant__ memory) and one per-step parameter block (StepParams, passed by value).
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define G2_HD __host__ __device__ __forceinline__
#define G2_HDN __host__ __device__ __noinline__
#else
#define G2_HD inline
#define G2_HDN
#endif

namespace go2 {

constexpr int NLINK = 13;   // robot links (robot-local index; ground = -1)
constexpr int NV = 18;      // dofs: 6 free + 12 hinges
constexpr int NQ = 19;
constexpr int NHINGE = 12;
constexpr int NGEOM = 28;   // geom 0 = ground box, 1..27 robot
constexpr int MAX_PAIRS_STATIC = 352;
constexpr int MAX_LINK_PAIRS = 96, MAX_PAIR_RUNS = 160;  // 3 x 32-bit overlap masks
constexpr int MAX_CPAIRS = 24;   // colliding geom pairs kept per env (contact-normal cache entries)
constexpr int MAX_CONTACTS = 32; // contacts per env kept for the solve (reference cap: 150; overflow -> errno bit)
constexpr int MAX_ROWS = 4 * MAX_CONTACTS + NHINGE;
constexpr int NTRI = NV * (NV + 1) / 2;  // packed lower triangle
constexpr int MAX_REW = 32;      // reward terms
constexpr int NACT_MAX = 16;
constexpr int HIST_MAX = 4;      // action-delay ring slots (max_delay_steps + 1 <= 4)
constexpr int NOBS_MAX = 49;     // proprioceptive obs entries (the lidar variant appends up to MAX_ACTOR_SCAN heights after them)
constexpr int MAX_ACTOR_SCAN = 32, MAX_CRITIC_SCAN = 96;  // simulated-lidar height grids (go2_env_stair_lidar.py LidarScanConfig)
constexpr int SCAN_MAX = 77;
constexpr int MAX_EDGE_SAMPLES = 32;  // forward samples of the stair-info edge detector (go2_env_stair_lidar_2.py: 14)
constexpr int MAX_TERRAIN_ROWS = 16;  // difficulty rows of the stair terrain (go2_env_stair.py:61)
constexpr int TERRAIN_MIP_SHIFT = 2;      // 4 x 4 cells per block of the max-height level
constexpr int TERRAIN_PAIR_CONTACTS = 5;  // n_contacts_per_pair (collider.py)

constexpr float EPSF = 1.1920929e-07f;  // gs.EPS for precision="32" (genesis/__init__.py:229)
constexpr float CCD_EPS = 1e-9f;        // collider/mpr.py:16-20
constexpr float CCD_TOL = 1e-6f;
constexpr int CCD_ITERS = 50;
constexpr float MC_PERTURB = 1e-2f;     // collider/collider.py:78-80
constexpr float MC_TOL = 1e-2f;
constexpr int SOLVER_ITERS = 50;        // options/solvers.py:474-519
constexpr float SOLVER_TOL = 1e-6f;
constexpr int LS_ITERS = 50;
constexpr float LS_TOL = 1e-2f;

enum GeomType { GEOM_SPHERE = 1, GEOM_CYLINDER = 3, GEOM_BOX = 5, GEOM_TERRAIN = 7 };
enum ErrBits { ERR_CPAIR_OVERFLOW = 1, ERR_CONTACT_OVERFLOW = 2, ERR_NAN_FORCE = 4, ERR_NAN_ACC = 8, ERR_GJK_WANTED = 16, ERR_ACTION_WAIT = 32 /* the action upload of go2_step_host did not arrive within 2 s */ };
enum EnvKind { ENV_WALK = 0, ENV_STAIR = 1, ENV_BASE = 2 };
enum CtrlKind { CTRL_MANUAL_PLS = 0, CTRL_MANUAL_FIXED = 1, CTRL_GENESIS_PD = 2 };

struct Go2Model {
    int parent[NLINK];
    float l_pos[NLINK][3], l_quat[NLINK][4];
    float inertial_pos[NLINK][3], inertial_quat[NLINK][4], inertial_i[NLINK][9], mass[NLINK];
    float axis[NLINK][3];
    float limit_lo[NV], limit_hi[NV], frange[NV], armature[NV];
    float invweight_link[NLINK], invweight_dof[NV];
    float meaninertia, dt, gravity[3], sol_params[7];
    int g_link[NGEOM], g_type[NGEOM];
    float g_pos[NGEOM][3], g_quat[NGEOM][4], g_size[NGEOM][3], g_half[NGEOM][3];
    int npairs;
    uint8_t pair_a[MAX_PAIRS_STATIC], pair_b[MAX_PAIRS_STATIC];
    // two-level broad phase (built by unpack_model from the pair table): the distinct (link, link) pairs behind the geom pairs
    // (box index 0 = ground, l + 1 = robot link l) and the runs of consecutive geom pairs that share one of them
    int n_lp, n_run;
    uint8_t lp_a[MAX_LINK_PAIRS], lp_b[MAX_LINK_PAIRS], run_lp[MAX_PAIR_RUNS];
    uint16_t run_end[MAX_PAIR_RUNS];
    // Heightfield terrain replacing the ground box (stair env; go2_set_terrain).  hf is [terrain_r][terrain_c] metres,
    // shared by all envs; cell (r, c) sits at terrain_org + (r, c) * terrain_sh  (collider.py:374-394).
    int terrain_on, terrain_r, terrain_c;
    float terrain_sh, terrain_org[3], terrain_zmax, terrain_zbot;  // zbot = min(hf) - 1 (prism bottoms)
    const float* terrain_hf;
    // max-height pyramid level of the heightfield: mip[(r >> TERRAIN_MIP_SHIFT)][(c >> TERRAIN_MIP_SHIFT)] = max of the block's cells
    const float* terrain_mip;
    int terrain_mip_r, terrain_mip_c;
};

// Reward terms (union of the walk / stair / base env zoos).
enum RewardId {
    R_TRACKING_LIN_VEL = 0, R_TRACKING_ANG_VEL, R_LIN_VEL_Z, R_BASE_HEIGHT, R_ACTION_RATE, R_SIMILAR_TO_DEFAULT,
    R_ORIENTATION_PENALTY, R_DOF_ACC, R_DOF_VEL, R_ANG_VEL_XY, R_FEET_AIR_TIME, R_FOOT_SLIP, R_FOOT_CLEARANCE,
    R_JOINT_TRACKING, R_ENERGY, R_TORQUE_LOAD, R_STAND_STILL, R_STAND_STILL_VEL, R_FEET_STANCE,
    // base env (crouch / jump)
    R_JUMP_IMPULSE, R_JUMP_APEX, R_XY_STABILITY, R_ORIENTATION, R_NO_SHAKE, R_CROUCH, R_CROUCH_2, R_GROUND_PENALTY,
    R_CROUCH_TARGET, R_NO_FALL, R_Y_STABILITY, R_CROUCH_PROGRESS, R_CROUCH_SPEED,
    // stair
    R_ORIENTATION_ROLL_ONLY, R_FORWARD_PROGRESS,
    // go2_env_terrain.py
    R_STUMBLE,
    // go2_env_stair_lidar_2.py
    R_BODY_CONTACT,
    R_COUNT
};

// Static env configuration, compiled once from the four cfg dicts (host: go2_b200/config.py).
struct EnvConfig {
    int kind;             // EnvKind
    int ctrl;             // CtrlKind
    int num_envs, env_offset;  // env_offset = global id of local env 0 (multi-GPU sharding)
    int num_actions, num_pos_actions, num_obs, num_priv;
    int max_episode_length, resample_steps;
    int hist_size;        // delay ring slots
    int n_rew;            // number of active reward terms, evaluated in this order
    int rew_id[MAX_REW];
    float rew_scale[MAX_REW];  // already multiplied by dt, like the reference ctor does
    int motors_dof[NHINGE];    // dof index of action channel j
    float default_dof_pos[NHINGE], torque_limit[NHINGE];
    float action_scale, clip_actions;
    float kp, kd;         // Genesis PD gains (CTRL_GENESIS_PD) / nominal
    float pls_kp_default, pls_kp_scale, pls_kp_lo, pls_kp_hi;
    float term_roll, term_pitch, term_zvel, term_yvel;  // degrees / m/s; <0 => check disabled
    float base_init_pos[3], base_init_quat[4], inv_base_init_quat[4];
    float obs_scale_lin_vel, obs_scale_ang_vel, obs_scale_dof_pos, obs_scale_dof_vel;
    float tracking_sigma, base_height_target, feet_height_target, feet_air_time_target;
    float foot_contact_threshold;
    int foot_link[4];     // robot-local link index of each foot (cfg order FR, FL, RR, RL)
    int has_foot_tracking;
    int push_link;        // robot-local link the push force is applied to (reference quirk: FL_hip)
    int push_dur_lo, push_dur_hi;
    int has_push, has_init_z, has_init_euler, has_kp_factor, has_kd_factor, has_motor_strength, has_gravity_offset,
        has_delay, compound_commands, n_standing;
    float init_z_lo, init_z_hi, init_euler_lo, init_euler_hi;  // radians
    int min_delay, max_delay;
    int has_kp_range;
    int simulate_action_latency;  // base env: 1-step latency through last_actions
    // base-env reward parameters
    float rp[16];
    uint32_t seed_lo, seed_hi;
    // stair env (go2_env_stair.py): difficulty rows, height scan of the privileged obs, reward variants
    int has_terrain, n_terrain_rows;
    float row_center[MAX_TERRAIN_ROWS][3];
    int scan_nx, scan_ny;
    float scan_x0, scan_x1, scan_y0, scan_y1;
    float lin_vel_z_deadzone;
    // global (one value for every env, Genesis semantics) physics DR applied by the post-pass kernel whenever at least one
    // env reset in the tick (go2_env_walk.py:737-756, 803-848): which draws exist, the friction throttle, the hip links
    int has_friction_dr, has_mass_dr, has_com_dr, has_leg_mass_dr, global_dr_interval;
    int hip_link[4];          // robot-local link of each per-leg mass shift, cfg order FR, FL, RR, RL
    int solver_dense;         // 1 = every env uses the dense 18x18 Newton path (diagnostic; default 0 = arrow unless legs touch)
    int export_rew_terms;     // 1 = B_REW_TERMS receives every scaled reward term of the tick
    int replay_rows;          // rows of B_REPLAY (0 = no replay buffer; num_envs = draws can be replayed, see StepParams.replay_on)
    // simulated lidar of examples/locomotion/go2_env_stair_lidar.py: body-frame sample points of the actor scan (appended to the obs,
    // with the sim-to-real noise model) and of the clean critic scan (tail of the privileged obs instead of the height scan)
    int lidar_on, n_actor_scan, n_critic_scan;
    float actor_lx[MAX_ACTOR_SCAN], actor_ly[MAX_ACTOR_SCAN], critic_lx[MAX_CRITIC_SCAN], critic_ly[MAX_CRITIC_SCAN];
    float lidar_noise_std, lidar_dropout, lidar_blackout, lidar_latency, lidar_voffset_std, lidar_clip, lidar_scale;
    // lidar_on == 2: the 4-value stair info of examples/locomotion/go2_env_stair_lidar_2.py instead of the height grids
    // (n_actor_scan = n_critic_scan = 4): [distance to the next height change along the forward axis, riser height of the env's
    // terrain row, tread depth of the row, direction +1 up / -1 down / 0 none] x si_scale, noisy copy in the obs, clean copy at the
    // tail of the privileged obs.  si_edge_x = forward sample points, si_edge_mid = midpoints between consecutive samples.
    int si_n_samples;
    float si_edge_x[MAX_EDGE_SAMPLES], si_edge_mid[MAX_EDGE_SAMPLES];
    float si_max_range, si_threshold, si_scale[4];
    float si_edge_noise_std, si_height_noise_std, si_depth_noise_std, si_flip_prob, si_blackout_prob, si_latency_prob;
    float si_row_height[MAX_TERRAIN_ROWS], si_row_depth[MAX_TERRAIN_ROWS];
    // body_contact reward of the same file (:1865-1880): number of the listed links whose net contact force exceeds the threshold
    int n_body_contact, body_contact_link[4];
    float body_contact_threshold;
    // walk env on a custom heightmap (examples/locomotion/go2_env_Omni_kplearn_varied_terrain.py + terrain_manager.py, heightmap mode):
    // random spawn points, reset when the base wanders too far from its spawn point, base-height reward relative to the bilinearly
    // interpolated ground height (computed in double like the reference's numpy code: spacing / origin kept in double)
    int tm_on;
    int tm_variant;       // 1 = go2_env_terrain.py: ground height in the privileged obs, foot clearance above the ground under the base
    float stumble_threshold;
    float tm_max_wander, tm_safe_half_x, tm_safe_half_y, tm_spawn_offset;
    // sub-terrain GRID mode of terrain_manager.py (tm_grid_rows > 0): a reset picks a random tile and spawns at its centre + a uniform
    // jitter of +-tm_safe_half_x / y; tm_blind = 1: height queries return 0 (the reference's manager cannot read the height field of
    // a Genesis-generated terrain, terrain_manager.py:236-262), collision still runs on the field
    int tm_grid_rows, tm_grid_cols, tm_blind;
    double tm_hs, tm_ox, tm_oy;
    double tm_tile_x, tm_tile_y;  // tile size in metres (grid mode)
};

// Per-step parameters: everything the reference keeps as host-side python scalars (curriculum-driven
// values, global DR values, ring-buffer write index, push schedule).  Passed by value at each launch.
struct StepParams {
    uint32_t step_count;       // RNG counter component
    int hist_write_idx;        // ring slot written this step
    int push_resample;         // 1 if push_counter % push_interval == 0 at this step
    int push_enable;
    float push_lo, push_hi;
    float action_noise_std;
    float obs_noise[NOBS_MAX]; // per-entry std (obs_noise_vec); all zero => no noise draws
    int obs_noise_on;
    float cmd_lo[3], cmd_hi[3];
    // reset-time DR ranges, already lerped with this step's t_sample
    float kpf_lo, kpf_hi, kdf_lo, kdf_hi, ms_lo, ms_hi, go_lo, go_hi;
    float kp_lo, kp_hi, kd_lo, kd_hi;  // fixed-gain manual PD (non-PLS) base gain ranges
    int delay_lo, delay_hi;
    float kp_genesis, kd_genesis;  // dofs kp/kv for CTRL_GENESIS_PD
    int do_physics;            // 0 => reset-only launch
    int force_reset_all;
    int terrain_max_row;       // stair: int(curriculum_level * (rows - 1))
    int lock_terrain_rows;     // stair eval: keep each env's row across resets
    float lidar_noise_level;   // lidar variant: the DR level scaling the scan noise model (_lidar_noise_level)
    int n_resets_hint;         // stair: number of envs (all ranks) this launch resets when the host knows it (reset of all envs), else 0
    // candidates of the global physics DR (already lerped with this tick's t_sample); the post-pass kernel puts them in
    // force if an env reset in this tick - B_GLOBALS holds the values in force
    float cand_friction, cand_mass, cand_com[3], cand_leg[4];
    int parity;                // double-buffer index of B_GLOBALS / B_STATS for this tick (filled by the library)
    int defer_post;            // 1 = go2_step launches the step kernel only; the caller reduces B_STATS over ranks, then go2_post
    int replay_on;             // 1 = every random draw of the tick is read from B_REPLAY instead of Philox (fixture replay)
    int ext_on;                // 1 = B_EXT_FORCE holds external link forces for this tick (apply_links_external_force)
    // go2_step_host with pinned (device-mappable) result buffers: the step kernel streams obs / rew / done of finished warps
    // straight into host memory (filled by the library; the host leaves them 0)
    float* host_obs;
    float* host_rew;
    uint8_t* host_done;
    // go2_step_host: the action rows travel on a copy stream WHILE the kernel's opening phase (state load, FK, velocities) runs; the
    // copy stream writes act_seq to *act_flag after the rows, thread 0 of every block waits for it before env_pre (filled by the
    // library; null = the actions are already in place)
    const unsigned int* act_flag;
    unsigned int act_seq;
};

// B_GLOBALS: [2][GLOBALS_LEN] global physics DR values in force (slot = StepParams.parity)
constexpr int G_FRIC_GROUND = 0, G_FRIC_ROBOT = 1, G_MASS_SHIFT = 2, G_COM_SHIFT = G_MASS_SHIFT + NLINK, G_DR_COUNTER = G_COM_SHIFT + 3,
              GLOBALS_LEN = 20;

// B_REPLAY row layout (one row per env): the values the reference drew in the replayed tick
constexpr int RP_ACT_NOISE = 0, RP_OBS_NOISE = 12, RP_PUSH = 61 /* fx fy duration */, RP_CMD = 64, RP_KPF = 67, RP_KDF = 79, RP_MS = 91,
              RP_GO = 103, RP_DELAY = 106, RP_INIT_Z = 107, RP_ROLL = 108, RP_PITCH = 109, RP_CMD_RESET = 110, RP_ROW = 113, RP_KP = 114,
              RP_KD = 115, RP_LIDAR_N = 116 /* per-cell normals */, RP_LIDAR_U = 116 + MAX_ACTOR_SCAN /* per-cell uniforms */,
              RP_LIDAR_V = 116 + 2 * MAX_ACTOR_SCAN /* offset normal, latency u, blackout u */, RP_SPAWN = RP_LIDAR_V + 3 /* spawn x, y uniforms */, RP_TILE = RP_SPAWN + 2 /* grid mode: tile row, column */,
              REPLAY_LEN = RP_TILE + 2;

}  // namespace go2
