// TEST INFRASTRUCTURE ONLY - compiles the kernel source (csrc/go2_env.cuh, the exact per-thread code the
// CUDA kernel runs) with the host compiler and runs it env-by-env on host memory.  It exists because the
// build container has no GPU: it lets the CPU test-suite exercise the kernel *source* against the oracle.
// It is never loaded by the go2_b200 package (which hard-fails without the CUDA library); only
// tests/ loads it, by explicit path.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>
#include "../../include/go2_b200.h"
#include "../../go2-sim2real-locomotion-rl_b200/csrc/go2_env.cuh"
#include "../../go2-sim2real-locomotion-rl_b200/csrc/go2_model_host.hpp"
#include "../../go2-sim2real-locomotion-rl_b200/csrc/go2_layout.hpp"

using namespace go2;

struct go2_handle {
    Go2Model model;
    EnvConfig cfg;
    Buffers bf;
    int dtype[B_COUNT], rows[B_COUNT], cols[B_COUNT];
    std::vector<char> arena;
    std::vector<float> hf, mip;
    long long launches;
    int parity, post_pending;
    std::string err;
};

extern "C" {

void go2_abi_sizes(int* out4) {
    out4[0] = (int)sizeof(EnvConfig);
    out4[1] = (int)sizeof(StepParams);
    out4[2] = (int)B_COUNT;
    out4[3] = (int)STATS_LEN;
}

int go2_create(const int32_t* iv, int n_iv, const double* fv, int n_fv, const void* cfg, int cfg_bytes, int device, go2_handle** out) {
    (void)device;
    *out = nullptr;
    if (cfg_bytes != (int)sizeof(EnvConfig)) return -2;
    go2_handle* h = new go2_handle();
    h->launches = 0;
    h->parity = 0;
    h->post_pending = 0;
    memcpy(&h->cfg, cfg, sizeof(EnvConfig));
    if (!unpack_model(h->model, iv, n_iv, fv, n_fv)) { delete h; return -3; }
    compute_layout(h->cfg, h->dtype, h->rows, h->cols);
    size_t offs[B_COUNT];
    size_t bytes = compute_offsets(h->dtype, h->rows, h->cols, offs);
    h->arena.assign(bytes, 0);
    for (int i = 0; i < B_COUNT; ++i) h->bf.p[i] = h->arena.data() + offs[i];
    h->bf.B = h->cfg.num_envs;
    init_arena(h->cfg, offs, h->arena.data());
    *out = h;
    return 0;
}
void go2_destroy(go2_handle* h) { delete h; }
const char* go2_last_error(go2_handle* h) { return h ? h->err.c_str() : "null handle"; }
int go2_buffer_info(go2_handle* h, int id, void** ptr, int* dtype, int* rows, int* cols) {
    if (!h || id < 0 || id >= B_COUNT) return -1;
    *ptr = h->bf.p[id]; *dtype = h->dtype[id]; *rows = h->rows[id]; *cols = h->cols[id];
    return 0;
}
static void post(go2_handle* h, const StepParams& sp) {
    const EnvConfig& cfg = h->cfg;
    float* stats = (float*)h->bf.p[B_STATS];
    const float* tick = stats + sp.parity * STATS_LEN;
    const float* cur = (const float*)h->bf.p[B_GLOBALS] + sp.parity * GLOBALS_LEN;
    float* nxt = (float*)h->bf.p[B_GLOBALS] + (sp.parity ^ 1) * GLOBALS_LEN;
    float g[GLOBALS_LEN];
    bool changed = post_globals(cfg, sp, tick[STATS_RESETS], cur, g);
    if (changed && sp.do_physics && cfg.num_priv > 0)
        for (int b = 0; b < cfg.num_envs; ++b) post_patch_priv(cfg, g, (float*)h->bf.p[B_PRIV] + (size_t)b * cfg.num_priv);
    if (post_stair_rows_due(cfg, sp, tick[STATS_RESETS])) {
        const int n = (int)(tick[STATS_RESETS] + 0.5f);
        for (int b = 0; b < cfg.num_envs; ++b) post_stair_rows(h->model, cfg, sp, h->bf, b, n);
    }
    for (int i = 0; i < GLOBALS_LEN; ++i) nxt[i] = g[i];
    for (int i = 0; i < STATS_LEN; ++i) {
        stats[2 * STATS_LEN + i] += tick[i];
        stats[(sp.parity ^ 1) * STATS_LEN + i] = 0.0f;
    }
    h->launches += 1;
    h->parity ^= 1;
    h->post_pending = 0;
}
static int run(go2_handle* h, const float* actions, StepParams sp) {
    if (h->post_pending) return -3;
    sp.parity = h->parity;
    float* st = (float*)h->bf.p[B_STATS] + sp.parity * STATS_LEN;
    Phys* p = new Phys();
    EnvLocal* e = new EnvLocal();
    EnvStepOut out;
    for (int b = 0; b < h->cfg.num_envs; ++b) {
        t_env_step(h->model, h->cfg, sp, h->bf, actions, b, true, *p, *e, out);
        if (out.reset) {
            st[STATS_RESETS] += 1.0f;
            if (out.timeout) st[STATS_TIMEOUTS] += 1.0f;
            st[STATS_TRACKING] += out.tracking_per_sec;
            for (int k = 0; k < h->cfg.n_rew; ++k) st[STATS_REW0 + k] += out.ep_stat[k];
        }
        if (p->err & ~ERR_GJK_WANTED) st[STATS_ERR] += 1.0f;
    }
    delete p;
    delete e;
    h->launches += 1;
    if (sp.defer_post) { h->post_pending = 1; return 0; }
    post(h, sp);
    return 0;
}
int go2_step(go2_handle* h, const float* actions, const void* params, int params_bytes, void* stream) {
    (void)stream;
    if (params_bytes != (int)sizeof(StepParams)) return -2;
    StepParams sp;
    memcpy(&sp, params, sizeof(sp));
    return run(h, actions, sp);
}
int go2_post(go2_handle* h, const void* params, int params_bytes, void* stream) {
    (void)stream;
    if (params_bytes != (int)sizeof(StepParams) || !h->post_pending) return -2;
    StepParams sp;
    memcpy(&sp, params, sizeof(sp));
    sp.parity = h->parity;
    post(h, sp);
    return 0;
}
int go2_tick_parity(go2_handle* h) { return h ? h->parity : -1; }
int go2_reset_all(go2_handle* h, const void* params, int params_bytes, void* stream) {
    (void)stream;
    if (params_bytes != (int)sizeof(StepParams)) return -2;
    StepParams sp;
    memcpy(&sp, params, sizeof(sp));
    sp.do_physics = 0;
    sp.force_reset_all = 1;
    return run(h, (const float*)h->bf.p[B_ACTIONS_IN], sp);
}
int go2_step_host(go2_handle* h, const float* h_actions, const void* params, int params_bytes, float* h_obs, float* h_rew,
                  uint8_t* h_done, float* h_priv, float* h_timeouts, void* stream) {
    const EnvConfig& c = h->cfg;
    size_t B = c.num_envs;
    memcpy(h->bf.p[B_ACTIONS_IN], h_actions, B * c.num_actions * 4);
    int r = go2_step(h, (const float*)h->bf.p[B_ACTIONS_IN], params, params_bytes, stream);
    if (r) return r;
    memcpy(h_obs, h->bf.p[B_OBS], B * c.num_obs * 4);
    memcpy(h_rew, h->bf.p[B_REW], B * 4);
    memcpy(h_done, h->bf.p[B_RESET], B);
    if (h_priv && c.num_priv > 0) memcpy(h_priv, h->bf.p[B_PRIV], B * c.num_priv * 4);
    if (h_timeouts) memcpy(h_timeouts, h->bf.p[B_TIMEOUT], B * 4);
    return 0;
}
static float g_stat_slots[2][STATS_LEN];
int go2_read_stats_async(go2_handle* h, int slot, void* stream) {
    (void)stream;
    float* acc = (float*)h->bf.p[B_STATS] + 2 * STATS_LEN;
    memcpy(g_stat_slots[slot & 1], acc, sizeof(float) * STATS_LEN);
    memset(acc, 0, sizeof(float) * STATS_LEN);
    return 0;
}
int go2_stats_fetch(go2_handle* h, int slot, float* host_out, int n, int block) {
    (void)h; (void)block;
    if (n > STATS_LEN) n = STATS_LEN;
    memcpy(host_out, g_stat_slots[slot & 1], sizeof(float) * n);
    return 0;
}
int go2_step_host_begin(go2_handle* h, const float* h_actions, const void* params, int params_bytes, float* h_obs, float* h_rew,
                        uint8_t* h_done, void* stream, int* staged) {
    (void)h_obs; (void)h_rew; (void)h_done;
    *staged = 1;
    memcpy(h->bf.p[B_ACTIONS_IN], h_actions, (size_t)h->cfg.num_envs * h->cfg.num_actions * 4);
    return go2_step(h, (const float*)h->bf.p[B_ACTIONS_IN], params, params_bytes, stream);
}
int go2_read_stats(go2_handle* h, float* host_out, int n, void* stream) {
    go2_read_stats_async(h, 0, stream);
    return go2_stats_fetch(h, 0, host_out, n, 1);
}
int go2_export_links(go2_handle* h, void* stream) {
    (void)stream;
    Phys* p = new Phys();
    int B = h->bf.B;
    for (int b = 0; b < B; ++b) {
        for (int i = 0; i < NQ; ++i) p->qpos[i] = ((float*)h->bf.p[B_QPOS])[b * NQ + i];
        for (int i = 0; i < NV; ++i) p->qvel[i] = ((float*)h->bf.p[B_QVEL])[b * NV + i];
        kinematics(h->model, (const float*)h->bf.p[B_GLOBALS] + h->parity * GLOBALS_LEN, *p);
        velocities(h->model, *p);
        float* lp = (float*)h->bf.p[B_LINK_POS] + (size_t)b * NLINK * 3;
        float* lv = (float*)h->bf.p[B_LINK_VEL] + (size_t)b * NLINK * 3;
        for (int l = 0; l < NLINK; ++l) {
            v3 v = link_vel_world(*p, l);
            lp[l * 3] = p->lpos[l].x; lp[l * 3 + 1] = p->lpos[l].y; lp[l * 3 + 2] = p->lpos[l].z;
            lv[l * 3] = v.x; lv[l * 3 + 1] = v.y; lv[l * 3 + 2] = v.z;
        }
    }
    delete p;
    h->launches += 1;
    return 0;
}
void go2_launch_shape(go2_handle* h, int* out3) { out3[0] = h->cfg.num_envs; out3[1] = 1; out3[2] = 0; }
int go2_set_terrain(go2_handle* h, const float* hf_m, int rows, int cols, float h_scale, const float* origin3) {
    if (!h || !hf_m || rows < 2 || cols < 2) return -1;
    h->hf.assign(hf_m, hf_m + (size_t)rows * cols);
    Go2Model& m = h->model;
    float zmax = hf_m[0], zmin = hf_m[0];
    for (size_t i = 1; i < h->hf.size(); ++i) { zmax = hf_m[i] > zmax ? hf_m[i] : zmax; zmin = hf_m[i] < zmin ? hf_m[i] : zmin; }
    m.terrain_on = 1; m.terrain_r = rows; m.terrain_c = cols; m.terrain_sh = h_scale;
    for (int j = 0; j < 3; ++j) m.terrain_org[j] = origin3[j];
    m.terrain_zmax = zmax; m.terrain_zbot = zmin - 1.0f;
    m.terrain_hf = h->hf.data();
    build_terrain_mip(h->hf.data(), rows, cols, h->mip, m.terrain_mip_r, m.terrain_mip_c);
    m.terrain_mip = h->mip.data();
    m.g_type[0] = GEOM_TERRAIN;
    return 0;
}
long long go2_launch_count(go2_handle* h) { return h ? h->launches : 0; }
// debug aid of the test-suite: contacts the collision phase finds for env b's current qpos (first substep of a tick);
// out rows: pos[3], normal[3], penetration, mu, link a, link b
int go2_debug_contacts(go2_handle* h, int b, float* out, int max_rows) {
    Phys* p = new Phys();
    load_phys(h->bf, b, *p);
    const float* gl = (const float*)h->bf.p[B_GLOBALS] + h->parity * GLOBALS_LEN;
    kinematics(h->model, gl, *p);
    velocities(h->model, *p);
    detect_collisions(h->model, *p, gl[G_FRIC_GROUND], gl[G_FRIC_ROBOT]);
    int n = p->n_contacts < max_rows ? p->n_contacts : max_rows;
    for (int i = 0; i < n; ++i) {
        const Contact& c = p->con[i];
        float* o = out + 10 * i;
        o[0] = c.pos.x; o[1] = c.pos.y; o[2] = c.pos.z; o[3] = c.normal.x; o[4] = c.normal.y; o[5] = c.normal.z;
        o[6] = c.pen; o[7] = c.mu; o[8] = (float)c.la; o[9] = (float)c.lb;
    }
    delete p;
    return n;
}
}
