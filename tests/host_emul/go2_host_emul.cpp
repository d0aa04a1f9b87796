// TEST INFRASTRUCTURE ONLY - compiles the kernel source (csrc/go2_env.cuh, the exact per-warp code the
// CUDA kernel runs, lane loops executed serially) with the host compiler and runs it env-by-env on host memory.  It exists because the
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
    std::vector<float> hf;
    long long launches;
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
static int run(go2_handle* h, const float* actions, const StepParams& sp) {
    float* st = (float*)h->bf.p[B_STATS];
    memset(st, 0, sizeof(float) * STATS_LEN);
    WarpS* w = new WarpS();
    Phys* p = new Phys();
    const char* lay = getenv("GO2_B200_LAYOUT");
    bool warp = lay && !strcmp(lay, "warp");
    for (int b = 0; b < h->cfg.num_envs; ++b) {
        if (warp) w_env_step(h->model, h->cfg, sp, h->bf, actions, b, *w);
        else { t_env_step(h->model, h->cfg, sp, h->bf, actions, b, true, *p, w->e, w->out); w->s.err = p->err; }
        const EnvStepOut& out = w->out;
        if (out.reset) {
            st[STATS_RESETS] += 1.0f;
            if (out.timeout) st[STATS_TIMEOUTS] += 1.0f;
            st[STATS_TRACKING] += out.tracking_per_sec;
            for (int k = 0; k < h->cfg.n_rew; ++k) st[STATS_REW0 + k] += out.ep_stat[k];
        }
        if (w->s.err) st[STATS_ERR] += 1.0f;
    }
    delete w;
    delete p;
    h->launches += 1;
    return 0;
}
int go2_step(go2_handle* h, const float* actions, const void* params, int params_bytes, void* stream) {
    (void)stream;
    if (params_bytes != (int)sizeof(StepParams)) return -2;
    StepParams sp;
    memcpy(&sp, params, sizeof(sp));
    return run(h, actions, sp);
}
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
int go2_read_stats(go2_handle* h, float* host_out, int n, void* stream) {
    (void)stream;
    if (n > STATS_LEN) n = STATS_LEN;
    memcpy(host_out, h->bf.p[B_STATS], sizeof(float) * n);
    return 0;
}
int go2_export_links(go2_handle* h, const void* params, int params_bytes, void* stream) {
    (void)stream;
    if (params_bytes != (int)sizeof(StepParams)) return -2;
    StepParams sp;
    memcpy(&sp, params, sizeof(sp));
    WarpS* w = new WarpS();
    int B = h->bf.B;
    for (int b = 0; b < B; ++b) {
        EnvS& es = w->s;
        for (int i = 0; i < NQ; ++i) es.qpos[i] = ((float*)h->bf.p[B_QPOS])[b * NQ + i];
        for (int i = 0; i < NV; ++i) es.qvel[i] = ((float*)h->bf.p[B_QVEL])[b * NV + i];
        w_kinematics(h->model, sp.mass_shift, sp.com_shift, es);
        w_velocities(h->model, es);
        float* lp = (float*)h->bf.p[B_LINK_POS] + (size_t)b * NLINK * 3;
        float* lv = (float*)h->bf.p[B_LINK_VEL] + (size_t)b * NLINK * 3;
        for (int l = 0; l < NLINK; ++l) {
            v3 v = link_vel_world(es, l);
            lp[l * 3] = es.lpos[l].x; lp[l * 3 + 1] = es.lpos[l].y; lp[l * 3 + 2] = es.lpos[l].z;
            lv[l * 3] = v.x; lv[l * 3 + 1] = v.y; lv[l * 3 + 2] = v.z;
        }
    }
    delete w;
    h->launches += 1;
    return 0;
}
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
    m.g_type[0] = GEOM_TERRAIN;
    return 0;
}
long long go2_launch_count(go2_handle* h) { return h ? h->launches : 0; }
#ifdef GO2_LS_STATS
void go2_ls_stats(long* out) {  // [calls, A[64], B[64], end[4]]
    out[0] = g_ls_calls;
    for (int i = 0; i < 64; ++i) { out[1 + i] = g_lsA[i]; out[65 + i] = g_lsB[i]; }
    for (int i = 0; i < 4; ++i) out[129 + i] = g_lsend[i];
    for (int i = 0; i < 64; ++i) out[133 + i] = g_nfl[i];
    for (int i = 0; i < 3; ++i) out[197 + i] = g_refac[i];
    for (int i = 0; i < 4; ++i) out[200 + i] = g_pairkind[i];
}
#endif
}
