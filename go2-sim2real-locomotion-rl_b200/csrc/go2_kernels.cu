// go2_b200 - CUDA kernels + C ABI (include/go2_b200.h) of the fused Go2 env step, sm_100a.
//
// One THREAD per environment (go2_env.cuh: t_env_step).  A thread loads its env's persistent state, runs the whole
// control tick (torque law, 2 rigid substeps incl. collision + Newton solve, rewards, reset, observations) in
// registers / thread-local memory and writes the state back once: one launch per tick, no global-memory round trips
// between phases.  A second, tiny launch (go2_post_kernel) resolves the one grid-wide dependency of a tick: "did any
// env reset" -> global physics DR, privileged-obs patch, statistics accumulation (see go2_env.cuh: post_globals).
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <cstdlib>
#include <string>
#include <vector>
#include "../../include/go2_b200.h"
#include "go2_env.cuh"
#include "go2_model_host.hpp"
#include "go2_layout.hpp"

using namespace go2;

constexpr int N_STAT_SLOTS = 2;  // pinned host slots for asynchronous statistics reads

struct go2_handle {
    int device;
    Go2Model* d_model;
    Go2Model model;      // host copy (patched by go2_set_terrain)
    float* d_hf;         // terrain heightfield (metres), shared by all envs
    float* d_mip;        // its 4 x 4-block max level
    EnvConfig* d_cfg;
    EnvConfig cfg;
    Buffers bf;
    int dtype[B_COUNT], rows[B_COUNT], cols[B_COUNT];
    char* arena;
    size_t arena_bytes;
    float* h_stats;      // pinned, [N_STAT_SLOTS][STATS_LEN]
    cudaEvent_t stat_ev[N_STAT_SLOTS];
    long long launches;
    int tpb, nblocks, small;  // launch shape of the step kernel; small = the 255-register instantiation
    int parity;          // double-buffer index the NEXT tick uses (B_GLOBALS / B_STATS)
    int post_pending;    // a deferred post-pass is outstanding (StepParams.defer_post)
    // go2_step_host: action rows + a sequence word go over their own (non-blocking) stream while the step kernel already runs
    cudaStream_t copy_stream;
    cudaEvent_t copy_gate;     // work queued on the caller's stream before this tick (the previous kernel may still read the rows)
    unsigned int* h_seq;       // pinned source of the sequence word
    unsigned int* d_seq;       // device word the kernel polls (StepParams.act_flag)
    unsigned int seq;
    std::string err;
};

#ifndef GO2_TPB_SMALL
#define GO2_TPB_SMALL 256
#endif
#ifndef GO2_TPB
#define GO2_TPB 256  // max envs (threads) per block; the launch picks blockDim <= GO2_TPB
#endif
#ifndef GO2_MIN_BLOCKS
#define GO2_MIN_BLOCKS 2  // resident blocks per SM the register budget is sized for: 2 x 224 threads x 128 registers at 65,536 envs
                          // (measured +1 % over one 448-thread block per SM, -DGO2_TPB=512 -DGO2_MIN_BLOCKS=1; profiles/r02_experiments.md)
#endif

__device__ __forceinline__ void accumulate_stats(const EnvConfig* gcfg, const StepParams& sp, const Buffers& bf, const EnvStepOut& out, int err) {
    float* st = (float*)bf.p[B_STATS] + sp.parity * STATS_LEN;
    if (out.reset) {
        atomicAdd(&st[STATS_RESETS], 1.0f);
        if (out.timeout) atomicAdd(&st[STATS_TIMEOUTS], 1.0f);
        atomicAdd(&st[STATS_TRACKING], out.tracking_per_sec);
        for (int k = 0; k < gcfg->n_rew; ++k) atomicAdd(&st[STATS_REW0 + k], out.ep_stat[k]);
    }
    if (err & ~ERR_GJK_WANTED) atomicAdd(&st[STATS_ERR], 1.0f);  // envs that raised an error bit this tick (bit 16 is a diagnostic, not an error)
}

// one thread per env
__device__ __forceinline__ void step_body(const Go2Model* __restrict__ gm, const EnvConfig* __restrict__ gcfg, const StepParams& sp,
                                          const Buffers& bf, const float* __restrict__ actions) {
    int b = blockIdx.x * blockDim.x + threadIdx.x;
    bool valid = b < bf.B;
    Phys p;
    EnvLocal e;
    EnvStepOut out;
    t_env_step(*gm, *gcfg, sp, bf, actions, b, valid, p, e, out);
    if (valid) accumulate_stats(gcfg, sp, bf, out, p.err);
    if (sp.host_obs) {
        // Host-buffer call with pinned result buffers: the 32 envs of a warp own one contiguous block of the env-major obs /
        // rew / done arrays, so a finished warp re-reads its block (L2) and streams it to the mapped host buffers with
        // full 128-byte stores - the PCIe writes of early warps overlap the late warps' compute, no D2H copy follows.
        __syncwarp();
        const int lane = threadIdx.x & 31, b0 = b - lane, nv = min(32, bf.B - b0);
        if (nv > 0) {
            const int no = gcfg->num_obs;
            const float* src = (const float*)bf.p[B_OBS] + (size_t)b0 * no;
            float* dst = sp.host_obs + (size_t)b0 * no;
            for (int w = lane; w < nv * no; w += 32) dst[w] = __ldcg(src + w);
            if (lane < nv) {
                sp.host_rew[b0 + lane] = __ldcg((const float*)bf.p[B_REW] + b0 + lane);
                sp.host_done[b0 + lane] = __ldcg((const uint8_t*)bf.p[B_RESET] + b0 + lane);
            }
        }
    }
}
// large batches: up to GO2_TPB threads per block, 128 registers (two 224-thread blocks per SM at 65,536 envs)
__global__ void __launch_bounds__(GO2_TPB, GO2_MIN_BLOCKS) go2_step_kernel(const Go2Model* __restrict__ gm, const EnvConfig* __restrict__ gcfg,
                                                           StepParams sp, Buffers bf, const float* __restrict__ actions) {
    step_body(gm, gcfg, sp, bf, actions);
}
// batches of at most GO2_TPB_SMALL envs per SM: same code with up to 255 registers (fewer thread-local round trips on the
// dependent chain; measured +7 % at 4,096, +16 % at 16,384 and 32,768 envs, +10 % on the stair cfg at 8,192)
__global__ void __launch_bounds__(GO2_TPB_SMALL) go2_step_kernel_small(const Go2Model* __restrict__ gm, const EnvConfig* __restrict__ gcfg,
                                                                       StepParams sp, Buffers bf, const float* __restrict__ actions) {
    step_body(gm, gcfg, sp, bf, actions);
}

// Post-pass of a tick (go2_env.cuh: post_globals).  Every block recomputes the (deterministic) decision from this tick's
// statistics row and the globals in force, and patches its envs' privileged obs; block 0 also publishes the next
// tick's globals slot, folds the statistics into the accumulator row and clears the next tick's statistics row.
__global__ void __launch_bounds__(256) go2_post_kernel(const Go2Model* __restrict__ gm, const EnvConfig* __restrict__ gcfg, StepParams sp,
                                                       Buffers bf) {
    const EnvConfig& cfg = *gcfg;
    float* stats = (float*)bf.p[B_STATS];
    const float* tick = stats + sp.parity * STATS_LEN;
    const float* cur = (const float*)bf.p[B_GLOBALS] + sp.parity * GLOBALS_LEN;
    float* nxt = (float*)bf.p[B_GLOBALS] + (sp.parity ^ 1) * GLOBALS_LEN;
    float g[GLOBALS_LEN];
    bool changed = post_globals(cfg, sp, tick[STATS_RESETS], cur, g);
    if (changed && sp.do_physics && cfg.num_priv > 0) {
        for (int b = blockIdx.x * blockDim.x + threadIdx.x; b < bf.B; b += gridDim.x * blockDim.x)
            post_patch_priv(cfg, g, (float*)bf.p[B_PRIV] + (size_t)b * cfg.num_priv);
    }
    if (post_stair_rows_due(cfg, sp, tick[STATS_RESETS])) {  // stair: rows of the envs reset in this tick, with the true reset count
        const int n = (int)(tick[STATS_RESETS] + 0.5f);
        for (int b = blockIdx.x * blockDim.x + threadIdx.x; b < bf.B; b += gridDim.x * blockDim.x) post_stair_rows(*gm, cfg, sp, bf, b, n);
    }
    if (blockIdx.x == 0) {
        if (threadIdx.x < GLOBALS_LEN) nxt[threadIdx.x] = g[threadIdx.x];
        if (threadIdx.x < STATS_LEN) {
            stats[2 * STATS_LEN + threadIdx.x] += tick[threadIdx.x];
            stats[(sp.parity ^ 1) * STATS_LEN + threadIdx.x] = 0.0f;
        }
    }
}

// get_links_pos / get_links_vel on demand: forward kinematics + link velocities of every env (one thread per env)
__global__ void __launch_bounds__(128) go2_links_kernel(const Go2Model* __restrict__ gm, int parity, Buffers bf) {
    int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= bf.B) return;
    Phys p;
    for (int i = 0; i < NQ; ++i) p.qpos[i] = ((const float*)bf.p[B_QPOS])[b * NQ + i];
    for (int i = 0; i < NV; ++i) p.qvel[i] = ((const float*)bf.p[B_QVEL])[b * NV + i];
    kinematics(*gm, (const float*)bf.p[B_GLOBALS] + parity * GLOBALS_LEN, p);
    velocities(*gm, p);
    for (int l = 0; l < NLINK; ++l) {
        float* lp = (float*)bf.p[B_LINK_POS] + ((size_t)b * NLINK + l) * 3;
        float* lv = (float*)bf.p[B_LINK_VEL] + ((size_t)b * NLINK + l) * 3;
        v3 v = link_vel_world(p, l);
        lp[0] = p.lpos[l].x; lp[1] = p.lpos[l].y; lp[2] = p.lpos[l].z;
        lv[0] = v.x; lv[1] = v.y; lv[2] = v.z;
    }
}

static bool ck(go2_handle* h, cudaError_t e, const char* what) {
    if (e == cudaSuccess) return true;
    h->err = std::string(what) + ": " + cudaGetErrorString(e);
    return false;
}

extern "C" {

void go2_abi_sizes(int* out4) {
    out4[0] = (int)sizeof(EnvConfig);
    out4[1] = (int)sizeof(StepParams);
    out4[2] = (int)B_COUNT;
    out4[3] = (int)STATS_LEN;
}

int go2_create(const int32_t* model_iv, int n_iv, const double* model_fv, int n_fv, const void* cfg, int cfg_bytes, int device,
               go2_handle** out) {
    *out = nullptr;
    if (cfg_bytes != (int)sizeof(EnvConfig)) return -2;
    go2_handle* h = new go2_handle();
    h->device = device;
    h->launches = 0;
    h->parity = 0;
    h->post_pending = 0;
    memcpy(&h->cfg, cfg, sizeof(EnvConfig));
    {
        // k blocks per SM over all waves, as large as allowed: the warps of a block run phase-locked (barriers in
        // t_env_step / solve_constraints) and share instruction fetches, so big blocks win; the block size is
        // the smallest multiple of 32 that covers the envs with k * n_SM blocks.
        int n_sm = 148;
        cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, device);
        int B = h->cfg.num_envs, tpb = GO2_TPB;
        for (int k = 1; k <= 64; ++k) {
            int t = ((B + n_sm * k - 1) / (n_sm * k) + 31) / 32 * 32;
            if (t <= GO2_TPB) { tpb = t; break; }
        }
        h->tpb = tpb;
        h->small = (B <= GO2_TPB_SMALL * n_sm && tpb <= GO2_TPB_SMALL) ? 1 : 0;  // one wave of <= 256-thread blocks
        h->nblocks = (B + tpb - 1) / tpb;
    }
    const EnvConfig& c = h->cfg;
    h->d_hf = nullptr;
    h->d_mip = nullptr;
    Go2Model& m = h->model;
    if (!unpack_model(m, model_iv, n_iv, model_fv, n_fv)) { delete h; return -3; }
    if (c.num_envs <= 0 || c.num_actions > NACT_MAX || c.num_obs - (c.lidar_on ? c.n_actor_scan : 0) > NOBS_MAX || c.n_actor_scan > MAX_ACTOR_SCAN || c.n_critic_scan > MAX_CRITIC_SCAN || c.hist_size > HIST_MAX || c.n_rew > MAX_REW) { delete h; return -4; }
    if (c.replay_rows != 0 && c.replay_rows != c.num_envs) { delete h; return -4; }
    if (cudaSetDevice(device) != cudaSuccess) { delete h; return -5; }
    compute_layout(c, h->dtype, h->rows, h->cols);
    size_t offs[B_COUNT];
    h->arena_bytes = compute_offsets(h->dtype, h->rows, h->cols, offs);
    if (cudaMalloc(&h->arena, h->arena_bytes) != cudaSuccess) { delete h; return -6; }
    for (int i = 0; i < B_COUNT; ++i) h->bf.p[i] = h->arena + offs[i];
    h->bf.B = c.num_envs;
    cudaMalloc(&h->d_model, sizeof(Go2Model));
    cudaMalloc(&h->d_cfg, sizeof(EnvConfig));
    cudaMemcpy(h->d_model, &m, sizeof(Go2Model), cudaMemcpyHostToDevice);
    cudaMemcpy(h->d_cfg, &h->cfg, sizeof(EnvConfig), cudaMemcpyHostToDevice);
    cudaMallocHost(&h->h_stats, sizeof(float) * STATS_LEN * N_STAT_SLOTS);
    for (int i = 0; i < N_STAT_SLOTS; ++i) cudaEventCreateWithFlags(&h->stat_ev[i], cudaEventDisableTiming);
    h->copy_stream = nullptr; h->copy_gate = nullptr; h->h_seq = nullptr; h->d_seq = nullptr; h->seq = 0;
    {   // overlapped action upload of go2_step_host (step_host_begin); GO2_B200_SERIAL_UPLOAD=1 keeps the copy on the caller's stream
        const char* off = getenv("GO2_B200_SERIAL_UPLOAD");
        if (!(off && off[0] == '1')) {
            bool ok = cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking) == cudaSuccess &&
                      cudaEventCreateWithFlags(&h->copy_gate, cudaEventDisableTiming) == cudaSuccess &&
                      cudaMallocHost(&h->h_seq, sizeof(unsigned int)) == cudaSuccess && cudaMalloc(&h->d_seq, sizeof(unsigned int)) == cudaSuccess;
            if (ok) { *h->h_seq = 0; cudaMemset(h->d_seq, 0, sizeof(unsigned int)); }
            else { cudaGetLastError(); if (h->copy_stream) cudaStreamDestroy(h->copy_stream); h->copy_stream = nullptr; }
        }
    }
    cudaFuncSetAttribute(go2_step_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 0);  // all of L1 for thread-local data
    cudaFuncSetAttribute(go2_step_kernel_small, cudaFuncAttributePreferredSharedMemoryCarveout, 0);
    std::string init(h->arena_bytes, 0);
    init_arena(c, offs, &init[0]);
    cudaMemcpy(h->arena, init.data(), h->arena_bytes, cudaMemcpyHostToDevice);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) h->err = cudaGetErrorString(e);
    *out = h;
    return e == cudaSuccess ? 0 : -7;
}

void go2_destroy(go2_handle* h) {
    if (!h) return;
    cudaSetDevice(h->device);
    cudaFree(h->arena);
    cudaFree(h->d_model);
    if (h->d_hf) cudaFree(h->d_hf);
    if (h->d_mip) cudaFree(h->d_mip);
    cudaFree(h->d_cfg);
    cudaFreeHost(h->h_stats);
    for (int i = 0; i < N_STAT_SLOTS; ++i) cudaEventDestroy(h->stat_ev[i]);
    if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
    if (h->copy_gate) cudaEventDestroy(h->copy_gate);
    if (h->h_seq) cudaFreeHost(h->h_seq);
    if (h->d_seq) cudaFree(h->d_seq);
    delete h;
}

const char* go2_last_error(go2_handle* h) { return h ? h->err.c_str() : "null handle"; }

int go2_buffer_info(go2_handle* h, int id, void** ptr, int* dtype, int* rows, int* cols) {
    if (!h || id < 0 || id >= B_COUNT) return -1;
    *ptr = h->bf.p[id];
    *dtype = h->dtype[id];
    *rows = h->rows[id];
    *cols = h->cols[id];
    return 0;
}

static int launch_post(go2_handle* h, const StepParams& sp, cudaStream_t s) {
    const EnvConfig& c = h->cfg;
    // the privileged-obs patch is the only per-env work; one block is enough for the flavours without it
    int blocks = ((c.num_priv > 0 && c.kind != ENV_BASE) || c.kind == ENV_STAIR) ? (c.num_envs + 255) / 256 : 1;
    if (blocks > 1184) blocks = 1184;
    go2_post_kernel<<<blocks, 256, 0, s>>>(h->d_model, h->d_cfg, sp, h->bf);
    h->launches += 1;
    h->parity ^= 1;
    h->post_pending = 0;
    return ck(h, cudaGetLastError(), "go2_post_kernel") ? 0 : -1;
}

static int launch_step(go2_handle* h, const float* d_actions, const void* params, int params_bytes, cudaStream_t s, bool host_call = false) {
    if (params_bytes != (int)sizeof(StepParams)) { h->err = "StepParams size mismatch"; return -2; }
    if (h->post_pending) { h->err = "go2_post must follow a go2_step with defer_post before the next tick"; return -3; }
    StepParams sp;
    memcpy(&sp, params, sizeof(sp));
    sp.parity = h->parity;
    if (!host_call) sp.host_obs = nullptr, sp.host_rew = nullptr, sp.host_done = nullptr, sp.act_flag = nullptr;  // only go2_step_host* sets them
    if (sp.replay_on && h->cfg.replay_rows != h->cfg.num_envs) { h->err = "replay_on needs EnvConfig.replay_rows == num_envs"; return -4; }
    if (cudaSetDevice(h->device) != cudaSuccess) { h->err = "cudaSetDevice"; return -1; }
    if (h->small)
        go2_step_kernel_small<<<h->nblocks, h->tpb, 0, s>>>(h->d_model, h->d_cfg, sp, h->bf, d_actions);
    else
        go2_step_kernel<<<h->nblocks, h->tpb, 0, s>>>(h->d_model, h->d_cfg, sp, h->bf, d_actions);
    h->launches += 1;
    if (!ck(h, cudaGetLastError(), "go2_step_kernel")) return -1;
    if (sp.defer_post) {
        h->post_pending = 1;
        return 0;
    }
    return launch_post(h, sp, s);
}

int go2_step(go2_handle* h, const float* d_actions, const void* params, int params_bytes, void* stream) {
    return launch_step(h, d_actions, params, params_bytes, (cudaStream_t)stream);
}

int go2_post(go2_handle* h, const void* params, int params_bytes, void* stream) {
    if (params_bytes != (int)sizeof(StepParams)) { h->err = "StepParams size mismatch"; return -2; }
    if (!h->post_pending) { h->err = "go2_post without a pending deferred post-pass"; return -3; }
    StepParams sp;
    memcpy(&sp, params, sizeof(sp));
    sp.parity = h->parity;
    if (cudaSetDevice(h->device) != cudaSuccess) { h->err = "cudaSetDevice"; return -1; }
    return launch_post(h, sp, (cudaStream_t)stream);
}

int go2_tick_parity(go2_handle* h) { return h ? h->parity : -1; }

int go2_reset_all(go2_handle* h, const void* params, int params_bytes, void* stream) {
    if (params_bytes != (int)sizeof(StepParams)) { h->err = "StepParams size mismatch"; return -2; }
    StepParams sp;
    memcpy(&sp, params, sizeof(sp));
    sp.do_physics = 0;
    sp.force_reset_all = 1;
    return launch_step(h, (const float*)h->bf.p[B_ACTIONS_IN], &sp, sizeof(sp), (cudaStream_t)stream);
}

// the device address of a host buffer the kernels may write directly (pinned memory under unified addressing), else null
static void* mapped(const void* hp) {
    if (!hp) return nullptr;
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, hp) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return at.type == cudaMemoryTypeHost ? at.devicePointer : nullptr;
}

// H2D actions + step kernel (+ post-pass unless StepParams.defer_post).  Result buffers that are pinned are written by the
// step kernel itself; *staged = 1 tells the caller that obs / rew / done still need device-to-host copies.
static int step_host_begin(go2_handle* h, const float* h_actions, const void* params, int params_bytes, float* h_obs, float* h_rew,
                           uint8_t* h_done, cudaStream_t s, int* staged) {
    const EnvConfig& c = h->cfg;
    size_t B = c.num_envs;
    if (params_bytes != (int)sizeof(StepParams)) { h->err = "StepParams size mismatch"; return -2; }
    if (cudaSetDevice(h->device) != cudaSuccess) { h->err = "cudaSetDevice"; return -1; }
    StepParams sp;
    memcpy(&sp, params, sizeof(sp));
    void *mo = mapped(h_obs), *mr = mapped(h_rew), *md = mapped(h_done);
    const bool direct = mo && mr && md;
    sp.host_obs = direct ? (float*)mo : nullptr;
    sp.host_rew = direct ? (float*)mr : nullptr;
    sp.host_done = direct ? (uint8_t*)md : nullptr;
    *staged = direct ? 0 : 1;
    if (h->copy_stream && sp.do_physics) {
        // The opening phase of the kernel (state load, FK, velocities: ~0.1 ms at 65,536 envs) needs no actions, so the 4 MB upload
        // runs beside it instead of before it: rows, then the sequence word, on the copy stream (ordered); the kernel waits for the
        // word before env_pre.  Both copies are queued BEFORE the launch, so they can never sit behind the kernel.
        h->seq += 1;
        *h->h_seq = h->seq;
        if (!ck(h, cudaEventRecord(h->copy_gate, s), "copy gate")) return -1;
        if (!ck(h, cudaStreamWaitEvent(h->copy_stream, h->copy_gate, 0), "copy gate wait")) return -1;
        if (!ck(h, cudaMemcpyAsync(h->bf.p[B_ACTIONS_IN], h_actions, B * c.num_actions * 4, cudaMemcpyHostToDevice, h->copy_stream), "H2D actions")) return -1;
        if (!ck(h, cudaMemcpyAsync(h->d_seq, h->h_seq, sizeof(unsigned int), cudaMemcpyHostToDevice, h->copy_stream), "H2D action seq")) return -1;
        sp.act_flag = h->d_seq;
        sp.act_seq = h->seq;
    } else {
        if (!ck(h, cudaMemcpyAsync(h->bf.p[B_ACTIONS_IN], h_actions, B * c.num_actions * 4, cudaMemcpyHostToDevice, s), "H2D actions")) return -1;
    }
    return launch_step(h, (const float*)h->bf.p[B_ACTIONS_IN], &sp, sizeof(sp), s, true);
}

int go2_step_host_begin(go2_handle* h, const float* h_actions, const void* params, int params_bytes, float* h_obs, float* h_rew,
                        uint8_t* h_done, void* stream, int* staged) {
    return step_host_begin(h, h_actions, params, params_bytes, h_obs, h_rew, h_done, (cudaStream_t)stream, staged);
}

int go2_step_host(go2_handle* h, const float* h_actions, const void* params, int params_bytes, float* h_obs, float* h_rew,
                  uint8_t* h_done, float* h_priv, float* h_timeouts, void* stream) {
    cudaStream_t s = (cudaStream_t)stream;
    const EnvConfig& c = h->cfg;
    size_t B = c.num_envs;
    int staged = 1;
    int r = step_host_begin(h, h_actions, params, params_bytes, h_obs, h_rew, h_done, s, &staged);
    if (r) return r;
    if (h->post_pending) { h->err = "go2_step_host does not support defer_post (use go2_step_host_begin + go2_post)"; return -3; }
    if (staged) {
        // the small results first: a consumer polling rew / done is not held back by the bulk rows
        cudaMemcpyAsync(h_rew, h->bf.p[B_REW], B * 4, cudaMemcpyDeviceToHost, s);
        cudaMemcpyAsync(h_done, h->bf.p[B_RESET], B, cudaMemcpyDeviceToHost, s);
        cudaMemcpyAsync(h_obs, h->bf.p[B_OBS], B * c.num_obs * 4, cudaMemcpyDeviceToHost, s);
    }
    if (h_timeouts) cudaMemcpyAsync(h_timeouts, h->bf.p[B_TIMEOUT], B * 4, cudaMemcpyDeviceToHost, s);
    if (h_priv && c.num_priv > 0) cudaMemcpyAsync(h_priv, h->bf.p[B_PRIV], B * c.num_priv * 4, cudaMemcpyDeviceToHost, s);
    return ck(h, cudaStreamSynchronize(s), "step_host sync") ? 0 : -1;
}

// Enqueue: accumulator row -> pinned slot, then clear the accumulator (stream-ordered after the copy).
int go2_read_stats_async(go2_handle* h, int slot, void* stream) {
    cudaStream_t s = (cudaStream_t)stream;
    if (slot < 0 || slot >= N_STAT_SLOTS) { h->err = "bad stats slot"; return -2; }
    if (cudaSetDevice(h->device) != cudaSuccess) { h->err = "cudaSetDevice"; return -1; }
    float* acc = (float*)h->bf.p[B_STATS] + 2 * STATS_LEN;
    if (!ck(h, cudaMemcpyAsync(h->h_stats + slot * STATS_LEN, acc, sizeof(float) * STATS_LEN, cudaMemcpyDeviceToHost, s), "D2H stats")) return -1;
    if (!ck(h, cudaMemsetAsync(acc, 0, sizeof(float) * STATS_LEN, s), "clear stats")) return -1;
    return ck(h, cudaEventRecord(h->stat_ev[slot], s), "stats event") ? 0 : -1;
}

// Wait for the slot's copy and hand it out.  `block` = 0: returns 1 without waiting if the copy has not finished.
int go2_stats_fetch(go2_handle* h, int slot, float* host_out, int n, int block) {
    if (slot < 0 || slot >= N_STAT_SLOTS) { h->err = "bad stats slot"; return -2; }
    if (!block) {
        cudaError_t q = cudaEventQuery(h->stat_ev[slot]);
        if (q == cudaErrorNotReady) return 1;
        if (!ck(h, q, "stats query")) return -1;
    } else if (!ck(h, cudaEventSynchronize(h->stat_ev[slot]), "stats wait")) return -1;
    if (n > STATS_LEN) n = STATS_LEN;
    memcpy(host_out, h->h_stats + slot * STATS_LEN, sizeof(float) * n);
    return 0;
}

int go2_read_stats(go2_handle* h, float* host_out, int n, void* stream) {
    int r = go2_read_stats_async(h, 0, stream);
    if (r) return r;
    return go2_stats_fetch(h, 0, host_out, n, 1);
}

int go2_export_links(go2_handle* h, void* stream) {
    int B = h->cfg.num_envs;
    if (cudaSetDevice(h->device) != cudaSuccess) { h->err = "cudaSetDevice"; return -1; }
    go2_links_kernel<<<(B + 127) / 128, 128, 0, (cudaStream_t)stream>>>(h->d_model, h->parity, h->bf);
    h->launches += 1;
    return ck(h, cudaGetLastError(), "go2_links_kernel") ? 0 : -1;
}

int go2_set_terrain(go2_handle* h, const float* hf_m, int rows, int cols, float h_scale, const float* origin3) {
    if (!h || !hf_m || rows < 2 || cols < 2 || h_scale <= 0.0f) return -1;
    cudaSetDevice(h->device);
    if (h->d_hf) cudaFree(h->d_hf);
    size_t n = (size_t)rows * cols;
    if (!ck(h, cudaMalloc(&h->d_hf, n * sizeof(float)), "terrain alloc")) return -1;
    if (!ck(h, cudaMemcpy(h->d_hf, hf_m, n * sizeof(float), cudaMemcpyHostToDevice), "terrain upload")) return -1;
    Go2Model& m = h->model;
    float zmax = hf_m[0], zmin = hf_m[0];
    for (size_t i = 1; i < n; ++i) { zmax = hf_m[i] > zmax ? hf_m[i] : zmax; zmin = hf_m[i] < zmin ? hf_m[i] : zmin; }
    m.terrain_on = 1; m.terrain_r = rows; m.terrain_c = cols; m.terrain_sh = h_scale;
    for (int j = 0; j < 3; ++j) m.terrain_org[j] = origin3[j];
    m.terrain_zmax = zmax; m.terrain_zbot = zmin - 1.0f;  // collider.py:386-389 terrain_xyz_maxmin
    m.terrain_hf = h->d_hf;
    {
        std::vector<float> mip;
        build_terrain_mip(hf_m, rows, cols, mip, m.terrain_mip_r, m.terrain_mip_c);
        if (h->d_mip) cudaFree(h->d_mip);
        if (!ck(h, cudaMalloc(&h->d_mip, mip.size() * sizeof(float)), "terrain mip alloc")) return -1;
        if (!ck(h, cudaMemcpy(h->d_mip, mip.data(), mip.size() * sizeof(float), cudaMemcpyHostToDevice), "terrain mip upload")) return -1;
        m.terrain_mip = h->d_mip;
    }
    m.g_type[0] = GEOM_TERRAIN;
    return ck(h, cudaMemcpy(h->d_model, &m, sizeof(Go2Model), cudaMemcpyHostToDevice), "model upload") ? 0 : -1;
}

#ifdef G2_PHASE_CLOCKS
// profiling build only: read (and clear) the per-phase clock table of go2_env.cuh
void go2_phase_clocks(unsigned long long* out8) {
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out8, go2::g2_phase_clk, sizeof(unsigned long long) * 8);
    unsigned long long z[8] = {0};
    cudaMemcpyToSymbol(go2::g2_phase_clk, z, sizeof(z));
}
#endif

// launch shape of the step kernel: [blocks, threads per block, 255-register instantiation?]
void go2_launch_shape(go2_handle* h, int* out3) {
    out3[0] = h->nblocks; out3[1] = h->tpb; out3[2] = h->small;
}

long long go2_launch_count(go2_handle* h) { return h ? h->launches : 0; }

}  // extern "C"
