// go2_b200 - device buffer table.  The library owns one arena; the Python side wraps each entry as a
// non-owning torch view (see go2_b200/core.py).
// every buffer is env-major ([B][K]): the env-level tensors the reference exposes as torch attributes get the
// reference's shapes and the physics state can be teacher-forced from the host.
#pragma once
#include "go2_types.cuh"

namespace go2 {

enum BufId {
    // ---- physics state, env-major
    B_QPOS = 0,      // f32 [B][19]
    B_QVEL,          // f32 [B][18]
    B_QACC_WS,       // f32 [B][18]
    B_WARM,          // i32 [B]
    B_NCACHE,        // i32 [B]
    B_CACHE_PAIR,    // i32 [B][MAX_CPAIRS]
    B_CACHE_N,       // f32 [B][MAX_CPAIRS*3]
    // ---- env tensors, batch first
    B_ACTIONS_IN,    // f32 [B][A]   (staging copy of the policy output for host-buffer calls)
    B_OBS,           // f32 [B][num_obs]
    B_PRIV,          // f32 [B][num_priv]
    B_REW,           // f32 [B]
    B_RESET,         // u8  [B]
    B_TIMEOUT,       // f32 [B]
    B_EP_LEN,        // i32 [B]
    B_COMMANDS,      // f32 [B][3]
    B_LAST_ACTIONS,  // f32 [B][A]
    B_APPLIED,       // f32 [B][A]
    B_HIST,          // f32 [B][H][A]
    B_LAST_DOF_VEL,  // f32 [B][12]
    B_DOF_POS,       // f32 [B][12]
    B_DOF_VEL,       // f32 [B][12]
    B_TARGET,        // f32 [B][12]
    B_BASE_POS,      // f32 [B][3]
    B_BASE_QUAT,     // f32 [B][4]
    B_BASE_LIN_VEL,  // f32 [B][3]
    B_BASE_ANG_VEL,  // f32 [B][3]
    B_PROJ_G,        // f32 [B][3]
    B_BASE_EULER,    // f32 [B][3]
    B_KP_FACTORS,    // f32 [B][12]
    B_KD_FACTORS,    // f32 [B][12]
    B_MOTOR_STRENGTH,// f32 [B][12]
    B_EFF_KP,        // f32 [B][12]
    B_EFF_KD,        // f32 [B][12]
    B_GRAV_OFF,      // f32 [B][3]
    B_PUSH_FORCE,    // f32 [B][3]  stored push force
    B_CUR_PUSH,      // f32 [B][3]  force applied this step
    B_PUSH_REMAIN,   // i32 [B]
    B_DELAY,         // i32 [B]
    B_FOOT_CONTACT,  // u8  [B][4]
    B_LAST_FOOT_CONTACT, // u8 [B][4]
    B_AIR_TIME,      // f32 [B][4]
    B_EP_SUMS,       // f32 [B][n_rew]
    B_ERRNO,         // i32 [B]
    B_LINK_POS,      // f32 [B][13][3]  robot link origins (get_links_pos)
    B_LINK_VEL,      // f32 [B][13][3]
    B_CONTACT_FORCE, // f32 [B][13][3]  net contact force per link (get_links_net_contact_force)
    B_TORQUE,        // f32 [B][12]     control force in action order (get_dofs_control_force)
    B_STATS,         // f32 [3][STATS_LEN] episode statistics: rows 0/1 = this tick (by StepParams.parity), row 2 = accumulator
    B_NCONTACT,      // i32 [B]  diagnostics: contacts in the last substep
    B_NNEWTON,       // i32 [B]  diagnostics: Newton iterations in the last substep
    B_TERRAIN_ROW,   // i32 [B]  stair: difficulty row of the env
    B_LAST_X,        // f32 [B]  stair: base x at the previous tick (forward_progress)
    B_GLOBALS,       // f32 [2][GLOBALS_LEN] global physics DR in force (friction, mass / COM shifts, reset counter), double-buffered
    B_REW_TERMS,     // f32 [B][n_rew]  scaled reward terms of the tick (cfg.export_rew_terms)
    B_REPLAY,        // f32 [B][REPLAY_LEN] replayed random draws (cfg.replay_rows, StepParams.replay_on)
    B_FK_FRESH,      // u8  [B]  1 = the env's kinematics were refreshed (reset / state setter) after the last global mass / COM change
    B_EXT_FORCE,     // f32 [B][4]  external force of the tick: robot-local link (as float), fx, fy, fz (world frame, at the link origin)
    B_ACTOR_SCAN,    // f32 [B][n_actor_scan]  lidar variant: the actor scan in the obs (kept: a stale scan is replayed on "latency")
    B_SPAWN_POS,     // f32 [B][3]  heightmap walk variant: spawn point of the env (boundary reset, terrain_manager._env_spawn_pos)
    B_GROUND_H,      // f32 [B]     heightmap walk variant: ground height under the base at the last tick (_ground_height)
    B_COUNT
};

constexpr int STATS_RESETS = 0, STATS_TIMEOUTS = 1, STATS_TRACKING = 2, STATS_ERR = 3, STATS_REW0 = 4;
constexpr int STATS_LEN = STATS_REW0 + MAX_REW;

struct Buffers {
    void* p[B_COUNT];
    int B;
};

}  // namespace go2
