// go2_b200 - buffer-table shapes and initial contents (host-side, shared by the CUDA library and the
// CPU emulation harness under tests/host_emul).
#pragma once
#include "go2_buffers.cuh"
#include <stddef.h>
#include <vector>

namespace go2 {

inline void compute_layout(const EnvConfig& c, int* dtype, int* rows, int* cols) {
    int B = c.num_envs, A = c.num_actions;
    auto set = [&](int id, int dt, int r, int cc) { dtype[id] = dt; rows[id] = r; cols[id] = cc; };
    set(B_QPOS, 0, B, NQ); set(B_QVEL, 0, B, NV); set(B_QACC_WS, 0, B, NV); set(B_WARM, 1, B, 1); set(B_NCACHE, 1, B, 1);
    set(B_CACHE_PAIR, 1, B, MAX_CPAIRS); set(B_CACHE_N, 0, B, MAX_CPAIRS * 3);
    set(B_ACTIONS_IN, 0, B, A); set(B_OBS, 0, B, c.num_obs); set(B_PRIV, 0, B, c.num_priv > 0 ? c.num_priv : 1);
    set(B_REW, 0, B, 1); set(B_RESET, 2, B, 1); set(B_TIMEOUT, 0, B, 1); set(B_EP_LEN, 1, B, 1); set(B_COMMANDS, 0, B, 3);
    set(B_LAST_ACTIONS, 0, B, A); set(B_APPLIED, 0, B, A); set(B_HIST, 0, B, c.hist_size * A);
    set(B_LAST_DOF_VEL, 0, B, NHINGE); set(B_DOF_POS, 0, B, NHINGE); set(B_DOF_VEL, 0, B, NHINGE); set(B_TARGET, 0, B, NHINGE);
    set(B_BASE_POS, 0, B, 3); set(B_BASE_QUAT, 0, B, 4); set(B_BASE_LIN_VEL, 0, B, 3); set(B_BASE_ANG_VEL, 0, B, 3);
    set(B_PROJ_G, 0, B, 3); set(B_BASE_EULER, 0, B, 3);
    set(B_KP_FACTORS, 0, B, NHINGE); set(B_KD_FACTORS, 0, B, NHINGE); set(B_MOTOR_STRENGTH, 0, B, NHINGE);
    set(B_EFF_KP, 0, B, NHINGE); set(B_EFF_KD, 0, B, NHINGE); set(B_GRAV_OFF, 0, B, 3); set(B_PUSH_FORCE, 0, B, 3);
    set(B_CUR_PUSH, 0, B, 3); set(B_PUSH_REMAIN, 1, B, 1); set(B_DELAY, 1, B, 1);
    set(B_FOOT_CONTACT, 2, B, 4); set(B_LAST_FOOT_CONTACT, 2, B, 4); set(B_AIR_TIME, 0, B, 4);
    set(B_EP_SUMS, 0, B, c.n_rew > 0 ? c.n_rew : 1); set(B_ERRNO, 1, B, 1);
    set(B_LINK_POS, 0, B, NLINK * 3); set(B_LINK_VEL, 0, B, NLINK * 3); set(B_CONTACT_FORCE, 0, B, NLINK * 3);
    set(B_TORQUE, 0, B, NHINGE); set(B_STATS, 0, 3, STATS_LEN); set(B_NCONTACT, 1, B, 1); set(B_NNEWTON, 1, B, 1);
    set(B_TERRAIN_ROW, 1, B, 1); set(B_LAST_X, 0, B, 1);
    set(B_GLOBALS, 0, 2, GLOBALS_LEN); set(B_REW_TERMS, 0, c.export_rew_terms ? B : 1, c.n_rew > 0 ? c.n_rew : 1);
    set(B_REPLAY, 0, c.replay_rows > 0 ? c.replay_rows : 1, REPLAY_LEN); set(B_FK_FRESH, 2, B, 1); set(B_EXT_FORCE, 0, B, 4);
    set(B_SPAWN_POS, 0, c.tm_on ? B : 1, 3); set(B_GROUND_H, 0, c.tm_on ? B : 1, 1);
    set(B_ACTOR_SCAN, 0, c.lidar_on && c.n_actor_scan > 0 ? B : 1, c.n_actor_scan > 0 ? c.n_actor_scan : 1);
}

// max-height level of the heightfield (terrain_contacts' conservative cull): one value per 4 x 4 block of cells
inline void build_terrain_mip(const float* hf, int rows, int cols, std::vector<float>& mip, int& mr, int& mc) {
    const int bs = 1 << TERRAIN_MIP_SHIFT;
    mr = (rows + bs - 1) / bs;
    mc = (cols + bs - 1) / bs;
    mip.assign((size_t)mr * mc, -1e30f);
    for (int r = 0; r < rows; ++r)
        for (int c = 0; c < cols; ++c) {
            float& t = mip[(size_t)(r >> TERRAIN_MIP_SHIFT) * mc + (c >> TERRAIN_MIP_SHIFT)];
            float v = hf[(size_t)r * cols + c];
            t = v > t ? v : t;
        }
}

inline size_t compute_offsets(const int* dtype, const int* rows, const int* cols, size_t* offs) {
    size_t off = 0;
    for (int i = 0; i < B_COUNT; ++i) {
        size_t bytes = (size_t)rows[i] * cols[i] * (dtype[i] == 2 ? 1 : 4);
        offs[i] = off;
        off += (bytes + 255) / 256 * 256;
    }
    return off;
}

// initial state = scene.build(): qpos0 (base at base_init pose, hinges 0), DR factors 1, delay 1
inline void init_arena(const EnvConfig& c, const size_t* offs, char* a) {
    int B = c.num_envs;
    float* q = (float*)(a + offs[B_QPOS]);
    for (int b = 0; b < B; ++b) {
        for (int j = 0; j < 3; ++j) q[b * NQ + j] = c.base_init_pos[j];
        for (int j = 0; j < 4; ++j) q[b * NQ + 3 + j] = c.base_init_quat[j];
    }
    const int ones[3] = {B_KP_FACTORS, B_KD_FACTORS, B_MOTOR_STRENGTH};
    for (int t = 0; t < 3; ++t) {
        float* f = (float*)(a + offs[ones[t]]);
        for (int i = 0; i < B * NHINGE; ++i) f[i] = 1.0f;
    }
    float* ekp = (float*)(a + offs[B_EFF_KP]);
    float* ekd = (float*)(a + offs[B_EFF_KD]);
    for (int i = 0; i < B * NHINGE; ++i) { ekp[i] = c.kp; ekd[i] = c.kd; }
    int* dl = (int*)(a + offs[B_DELAY]);
    for (int b = 0; b < B; ++b) dl[b] = 1;
    unsigned char* rb = (unsigned char*)(a + offs[B_RESET]);
    for (int b = 0; b < B; ++b) rb[b] = 1;
    unsigned char* ff = (unsigned char*)(a + offs[B_FK_FRESH]);
    for (int b = 0; b < B; ++b) ff[b] = 1;
    float* gl = (float*)(a + offs[B_GLOBALS]);  // friction 1 on both entities (URDF default), no shifts
    for (int s = 0; s < 2; ++s) gl[s * GLOBALS_LEN + G_FRIC_GROUND] = gl[s * GLOBALS_LEN + G_FRIC_ROBOT] = 1.0f;
    float* ex = (float*)(a + offs[B_EXT_FORCE]);
    for (int b = 0; b < B; ++b) ex[b * 4] = -1.0f;
}

}  // namespace go2
