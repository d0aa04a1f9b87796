// go2_b200 - fused actor / critic MLP of the rollout step on the 5th-generation tensor cores.
//
// Replaces, for one control tick, rsl-rl 2.2.4's `PPO.act`: ActorCritic.act (actor MLP 49-512-256-128-16 with ELU,
// Normal(mean, std).sample(), log_prob) and ActorCritic.evaluate (critic MLP 104-512-256-128-1), i.e. what
// OnPolicyRunner.learn does between two env.step calls (rsl_rl/algorithms/ppo.py act(), modules/actor_critic.py).
// The reference runs 8 cuBLAS fp32 GEMMs + ~20 elementwise launches per tick; here ONE kernel launch per network:
//
//   * one CTA owns a tile of 128 envs (= the 128 TMEM lanes) and walks the 4 layers without leaving the SM;
//   * weights are pre-packed on the host into the exact shared-memory image of each (layer, N-chunk, 32-wide K stage)
//     and streamed with 1-D TMA bulk copies (cp.async.bulk + mbarrier complete_tx) into a 4-deep ring;
//   * each stage is 2 x tcgen05.mma (kind::f16, bf16 operands, fp32 accumulate, M = 128, N <= 256, K = 16) issued by one
//     thread, accumulators in TMEM (up to 512 columns), completion signalled with tcgen05.commit -> mbarrier;
//   * the epilogue (tcgen05.ld 32x32b, bias, ELU, bf16 pack) writes the next layer's A operand straight back into shared
//     memory in the K-major core-matrix layout the next tcgen05.mma reads - activations never touch HBM;
//   * the last epilogue adds the Gaussian exploration noise (counter-based Philox keyed by seed / env / step) and writes
//     actions, mean, sigma, log-prob (actor) or the value (critic) to the rollout-storage rows of this tick.
//
// Operand layout (no swizzle, K-major): element (row r, k) of a tile with R rows lives at byte ((k/8)*R + r)*16 + (k%8)*2,
// i.e. 8x16-byte core matrices, SBO = 128 B between 8-row groups, LBO = R*16 B between the two K halves of one MMA.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

#include <stdarg.h>
#include <stdio.h>

#include <mutex>
#include <string>
#include <vector>

#include "go2_math.cuh"

#define GO2_NOALLOC_MAX_DEVICES 64

namespace go2pol {

constexpr int TILE_M = 128;        // envs per CTA = TMEM lanes
constexpr int MAX_LAYERS = 4;
constexpr int NCOMPUTE = 256;      // 8 warps: two per TMEM lane quarter (they split the columns of an epilogue)
constexpr int NTHREADS = NCOMPUTE + 32;  // + one TMA producer warp (keeps the weight ring full, never joins the layer barriers)
constexpr int KPAD = 64;           // layer widths are padded to multiples of this
constexpr int KSTAGE = 32;         // K elements per weight stage (2 MMAs of K = 16)
constexpr int NSTAGE = 4;          // depth of the TMA weight ring
constexpr int NCHUNK = 256;        // max N of one tcgen05.mma
constexpr int STAGE_BYTES = (KSTAGE / 8) * NCHUNK * 16;  // 16 KB
constexpr int ACT_BYTES = (512 / 8) * TILE_M * 16;        // widest activation tile: 128 x 512 bf16 = 128 KB
constexpr int BIAS_FLOATS = 512 + 512 + 512 + 16;
constexpr int SMEM_BYTES = ACT_BYTES + NSTAGE * STAGE_BYTES + 256 + BIAS_FLOATS * 4;

struct MlpDesc {
    int n_layers;
    int K[MAX_LAYERS];      // padded to a multiple of 64
    int N[MAX_LAYERS];      // padded to a multiple of 16 (<= 512)
    int n_in, n_out;        // real input / output widths
    const __nv_bfloat16* w[MAX_LAYERS];  // packed [n_chunk][K/8][Nc][8]
    const float* b[MAX_LAYERS];          // [N] (zero padded)
};

struct ActOut {            // all optional (nullptr = skip); row-major [B][n_out] / [B]
    float* out;            // network output (action mean / value)
    float* actions;        // mean + sigma * z
    float* sigma;          // broadcast std
    float* logp;           // sum log N(a; mean, sigma)
    const float* std;      // [n_out]
    uint32_t seed_lo, seed_hi, step, env_offset;
};

// ---- PTX wrappers ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes),
                 "r"(bar) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// K-major, no swizzle: start address, LBO (between the K halves), SBO (between 8-row groups), version 1 (sm_100)
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    return (uint64_t)((addr >> 4) & 0x3FFFu) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16) | ((uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32) |
           (1ull << 46);
}
// kind::f16 instruction descriptor: D = f32, A = B = bf16, both K-major, M = 128
__device__ __forceinline__ uint32_t instr_desc(int n) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(TILE_M >> 4) << 24);
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* v) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, "
        "%20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]),
          "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]),
          "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]),
          "=r"(v[31])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* v) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]),
          "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}
// exp(x) - 1 through the fast exponential: its absolute error (~1e-7) is far below the bf16 rounding of the result
__device__ __forceinline__ float elu(float x) { return x > 0.0f ? x : __expf(x) - 1.0f; }


// ---- the fused MLP kernel: 8 compute warps + 1 producer warp, one 128-env tile per CTA -----------------------------------------------
// blockIdx.y selects the network (0 / 1): actor and critic of one tick go out in ONE launch (twice the tiles per wave)
struct MlpJob {
    MlpDesc d;
    const float* in;
    ActOut o;
};
__global__ void __launch_bounds__(NTHREADS, 1) go2_mlp_kernel(MlpJob j0, MlpJob j1, int B) {
    const MlpDesc& d = blockIdx.y ? j1.d : j0.d;
    const float* __restrict__ in = blockIdx.y ? j1.in : j0.in;
    const ActOut& o = blockIdx.y ? j1.o : j0.o;
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* act = smem;
    uint8_t* wbuf = smem + ACT_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + ACT_BYTES + NSTAGE * STAGE_BYTES);  // full[4], empty[4], layer
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * NSTAGE + 2);
    float* sbias = reinterpret_cast<float*>(smem + ACT_BYTES + NSTAGE * STAGE_BYTES + 256);
    const uint32_t full0 = smem_u32(&bars[0]), empty0 = smem_u32(&bars[NSTAGE]), layer_bar = smem_u32(&bars[2 * NSTAGE]);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int trow_i = (warp & 3) * 32 + lane;   // row of the tile = TMEM lane this thread reads
    const int half = warp >> 2;                   // which half of the columns of an epilogue this warp takes
    const int row = blockIdx.x * TILE_M + trow_i;

    if (tid == 0) {
        for (int i = 0; i < NSTAGE; ++i) { mbar_init(full0 + 8 * i, 1); mbar_init(empty0 + 8 * i, 1); }
        mbar_init(layer_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(tmem_slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // biases of all layers -> shared memory (epilogues read them as broadcasts)
    int boff[MAX_LAYERS];
    {
        int acc = 0;
        for (int l = 0; l < d.n_layers; ++l) {
            boff[l] = acc;
            for (int i = tid; i < d.N[l]; i += NTHREADS) sbias[acc + i] = d.b[l][i];
            acc += d.N[l];
        }
    }
    // input tile -> bf16 A operand (K-major core matrices); rows past B and columns past n_in are zero
    {
        const int kchunks = d.K[0] >> 3;
        const int grow = blockIdx.x * TILE_M + (tid & (TILE_M - 1));
        for (int kc = tid >> 7; kc < kchunks && tid < NCOMPUTE; kc += NCOMPUTE / TILE_M) {
            float x[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                int k = kc * 8 + e;
                x[e] = (grow < B && k < d.n_in) ? in[(size_t)grow * d.n_in + k] : 0.0f;
            }
            uint4 v = make_uint4(pack_bf16(x[0], x[1]), pack_bf16(x[2], x[3]), pack_bf16(x[4], x[5]), pack_bf16(x[6], x[7]));
            *reinterpret_cast<uint4*>(act + ((size_t)kc * TILE_M + (tid & (TILE_M - 1))) * 16) = v;
        }
    }
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    // stage table walk (warp 0 only): stage = (layer, n-chunk, 32-wide k stage), weights contiguous per stage
    int total = 0;
    for (int l = 0; l < d.n_layers; ++l) total += ((d.N[l] + NCHUNK - 1) / NCHUNK) * (d.K[l] / KSTAGE);
    auto issue_tma = [&](int s) {
        const __nv_bfloat16* src = nullptr;
        uint32_t bytes = 0;
        int r = s;
        for (int l = 0; l < d.n_layers; ++l) {
            int nch = (d.N[l] + NCHUNK - 1) / NCHUNK, ks = d.K[l] / KSTAGE, nc_w = d.N[l] < NCHUNK ? d.N[l] : NCHUNK;
            if (r < nch * ks) {
                int nc = r / ks, k = r % ks;
                src = d.w[l] + ((size_t)nc * (d.K[l] >> 3) + (size_t)k * (KSTAGE >> 3)) * nc_w * 8;
                bytes = (KSTAGE >> 3) * nc_w * 16;
                break;
            }
            r -= nch * ks;
        }
        uint32_t bar = full0 + 8 * (s % NSTAGE);
        mbar_expect_tx(bar, bytes);
        tma_bulk_g2s(smem_u32(wbuf + (size_t)(s % NSTAGE) * STAGE_BYTES), src, bytes, bar);
    };
    if (warp == NCOMPUTE / 32) {
        // producer warp: stage s goes into ring slot s % NSTAGE as soon as the MMAs of stage s - NSTAGE have released it
        if (lane == 0) {
            for (int i = 0; i < NSTAGE && i < total; ++i) issue_tma(i);
            for (int i = NSTAGE; i < total; ++i) {
                mbar_wait(empty0 + 8 * (i % NSTAGE), ((i / NSTAGE) - 1) & 1);
                issue_tma(i);
            }
        }
        __syncwarp();
    }

    int s = 0;
    for (int l = 0; l < d.n_layers && warp < NCOMPUTE / 32; ++l) {
        const int Kl = d.K[l], Nl = d.N[l];
        const int nch = (Nl + NCHUNK - 1) / NCHUNK, kst = Kl / KSTAGE, nc_w = Nl < NCHUNK ? Nl : NCHUNK;
        if (warp == 0) {
            for (int nc = 0; nc < nch; ++nc) {
                for (int ks = 0; ks < kst; ++ks, ++s) {
                    if (lane == 0) {
                        const int b = s % NSTAGE;
                        mbar_wait(full0 + 8 * b, (s / NSTAGE) & 1);
                        tc_fence_after();
                        const uint32_t a_base = smem_u32(act) + (uint32_t)(ks * (KSTAGE >> 3)) * TILE_M * 16;
                        const uint32_t b_base = smem_u32(wbuf + (size_t)b * STAGE_BYTES);
                        const uint32_t idesc = instr_desc(nc_w);
#pragma unroll
                        for (int j = 0; j < KSTAGE / 16; ++j) {
                            uint64_t ad = smem_desc(a_base + (uint32_t)j * 2 * TILE_M * 16, TILE_M * 16, 128);
                            uint64_t bd = smem_desc(b_base + (uint32_t)j * 2 * nc_w * 16, nc_w * 16, 128);
                            tc_mma_bf16(tmem_base + (uint32_t)nc * NCHUNK, ad, bd, idesc, (ks > 0 || j > 0) ? 1u : 0u);
                        }
                        tc_commit(empty0 + 8 * b);
                        if (nc == nch - 1 && ks == kst - 1) tc_commit(layer_bar);
                    }
                    __syncwarp();
                }
            }
        }
        mbar_wait(layer_bar, l & 1);
        tc_fence_after();
        const uint32_t trow = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
        const float* bias = sbias + boff[l];
        if (l + 1 < d.n_layers) {
            for (int c0 = 32 * half; c0 < Nl; c0 += 64) {
                uint32_t v[32];
                tmem_ld32(trow + (uint32_t)c0, v);
                float x[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) x[i] = elu(__uint_as_float(v[i]) + bias[c0 + i]);
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    uint4 w4 = make_uint4(pack_bf16(x[8 * q], x[8 * q + 1]), pack_bf16(x[8 * q + 2], x[8 * q + 3]),
                                          pack_bf16(x[8 * q + 4], x[8 * q + 5]), pack_bf16(x[8 * q + 6], x[8 * q + 7]));
                    *reinterpret_cast<uint4*>(act + ((size_t)((c0 >> 3) + q) * TILE_M + trow_i) * 16) = w4;
                }
            }
            fence_async_smem();
        } else if (half == 0) {
            uint32_t v[16];
            tmem_ld16(trow, v);
            if (row < B) {
                const int no = d.n_out;
                float mean[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) mean[i] = __uint_as_float(v[i]) + bias[i];
                if (o.out)
                    for (int i = 0; i < no; ++i) o.out[(size_t)row * no + i] = mean[i];
                if (o.actions) {
                    float lp = 0.0f;
                    for (int i4 = 0; i4 < no; i4 += 4) {
                        go2::u4 r = go2::philox(o.env_offset + (uint32_t)row, o.step, 0x504F4Cu, (uint32_t)(i4 >> 2), o.seed_lo, o.seed_hi);
                        float z[4];
                        go2::boxmuller(r.a, r.b, z[0], z[1]);
                        go2::boxmuller(r.c, r.d, z[2], z[3]);
                        for (int j = 0; j < 4 && i4 + j < no; ++j) {
                            float sd = o.std[i4 + j];
                            o.actions[(size_t)row * no + i4 + j] = mean[i4 + j] + sd * z[j];
                            if (o.sigma) o.sigma[(size_t)row * no + i4 + j] = sd;
                            lp += -0.5f * z[j] * z[j] - logf(sd) - 0.9189385332f;
                        }
                    }
                    if (o.logp) o.logp[row] = lp;
                }
            }
        }
        tc_fence_before();
        asm volatile("bar.sync 1, %0;" ::"n"(NCOMPUTE) : "memory");  // compute warps only
        tc_fence_after();
    }
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
}

// Generalised advantage estimation over a [T][B] rollout (rsl_rl/storage/rollout_storage.py compute_returns), one thread
// per env walking backwards in time; coalesced over envs.  Advantages are NOT normalised here (a reduction over T*B).
__global__ void go2_gae_kernel(const float* __restrict__ rew, const uint8_t* __restrict__ done, const float* __restrict__ val,
                               const float* __restrict__ last_val, int T, int B, float gamma, float lam, float* __restrict__ ret,
                               float* __restrict__ adv) {
    int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    float a = 0.0f, next_v = last_val[b];
    for (int t = T - 1; t >= 0; --t) {
        size_t i = (size_t)t * B + b;
        float nt = 1.0f - (float)(done[i] != 0);
        float delta = rew[i] + nt * gamma * next_v - val[i];
        a = delta + nt * gamma * lam * a;
        ret[i] = a + val[i];
        adv[i] = a;
        next_v = val[i];
    }
}


// ---- no-allocation variant of the MLP forward (go2_mlp_forward_noalloc) --------------------------------------------------------------
// go2_mlp_create packs the weights on the host into cudaMalloc'ed buffers.  Callers that may not allocate (CUDA graph capture, tight
// serving loops) get the same kernel fed from ONE file-scope device array: a small kernel packs the f32 torch-layout parameters (device
// memory of the caller) into it, zero padding included, and go2_mlp_kernel streams the packed image from there.  Every GPU that loads
// the library holds the array until exit; all calls share it (calls that overlap on the GPU race).
constexpr size_t NOALLOC_SCRATCH_BYTES = 1u << 20;  // = GO2_MLP_NOALLOC_SCRATCH_BYTES of include/go2_policy.h
// 1 MiB: the reference's actor (49-512-256-128-12) needs 401,152 B, its critic (104-...-1) 466,688 B
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_SCRATCH_BYTES];

struct PackLayer {
    const float* W;         // [nr][kr] row-major (torch.nn.Linear.weight), device
    const float* bias;      // [nr], device
    __nv_bfloat16* w_out;   // packed [n_chunk][K/8][ncw][8] (the layout go2_mlp_create produces)
    float* b_out;           // [N] zero padded
    int kr, nr, K, N, ncw;
};
struct PackJob {
    PackLayer l[MAX_LAYERS];
};
// blockIdx.y = layer; every element of the padded image is written (stale contents of an earlier call never survive)
__global__ void __launch_bounds__(256) go2_mlp_pack_kernel(PackJob j) {
    const PackLayer& L = j.l[blockIdx.y];
    const size_t total = (size_t)L.K * L.N, stride = (size_t)gridDim.x * blockDim.x;
    const int k8 = L.K / 8;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += stride) {
        const int e = (int)(idx & 7);
        size_t r = idx >> 3;
        const int n = (int)(r % L.ncw);
        r /= L.ncw;
        const int kc = (int)(r % k8), nc = (int)(r / k8);
        const int gn = nc * L.ncw + n, gk = kc * 8 + e;
        const float v = (gn < L.nr && gk < L.kr) ? L.W[(size_t)gn * L.kr + gk] : 0.0f;
        L.w_out[idx] = __float2bfloat16_rn(v);
    }
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < (size_t)L.N; idx += stride)
        L.b_out[idx] = (int)idx < L.nr ? L.bias[idx] : 0.0f;
}

}  // namespace go2pol

// ---- host side + C ABI (declared in include/go2_b200.h) -------------------------------------------------------
using namespace go2pol;

struct go2_mlp {
    MlpDesc d;
    int device;
    std::vector<void*> allocs;
    std::string err;
};

static uint16_t f2bf(float f) {  // round to nearest even
    uint32_t u;
    memcpy(&u, &f, 4);
    if ((u & 0x7fffffffu) > 0x7f800000u) return 0x7fc0;
    u += 0x7fffu + ((u >> 16) & 1u);
    return (uint16_t)(u >> 16);
}
static int pad_to(int x, int m) { return (x + m - 1) / m * m; }

extern "C" {

int go2_mlp_create(const int32_t* dims, int n_layers, const float* params, int device, go2_mlp** out) {
    if (!dims || !params || !out || n_layers < 1 || n_layers > MAX_LAYERS) return 1;
    if (dims[n_layers] > 16 || dims[0] > 512) return 2;
    for (int l = 1; l < n_layers; ++l)
        if (dims[l] > 512) return 2;
    if (cudaSetDevice(device) != cudaSuccess) return 3;
    go2_mlp* h = new go2_mlp();
    h->device = device;
    MlpDesc& d = h->d;
    memset(&d, 0, sizeof(d));
    d.n_layers = n_layers;
    d.n_in = dims[0];
    d.n_out = dims[n_layers];
    const float* p = params;
    for (int l = 0; l < n_layers; ++l) {
        const int kr = dims[l], nr = dims[l + 1];
        // hidden widths are padded with zero units (ELU(0) = 0 feeds zero weights of the next layer): to a multiple of KPAD, and
        // to exactly 2 * NCHUNK when wider than one N-chunk - the kernel's chunk loop runs whole NCHUNK-wide chunks
        int N = (l + 1 < n_layers) ? pad_to(nr, KPAD) : 16;
        if (N > NCHUNK) N = 2 * NCHUNK;
        const int K = l == 0 ? pad_to(kr, KPAD) : d.N[l - 1];
        d.K[l] = K;
        d.N[l] = N;
        const float* W = p;          // [nr][kr] row-major (torch.nn.Linear.weight)
        const float* bsrc = p + (size_t)nr * kr;
        p += (size_t)nr * kr + nr;
        const int nch = (N + NCHUNK - 1) / NCHUNK, ncw = N < NCHUNK ? N : NCHUNK;
        std::vector<uint16_t> pk((size_t)K * N, 0);
        for (int nc = 0; nc < nch; ++nc)
            for (int kc = 0; kc < K / 8; ++kc)
                for (int n = 0; n < ncw; ++n)
                    for (int e = 0; e < 8; ++e) {
                        int gn = nc * ncw + n, gk = kc * 8 + e;
                        float v = (gn < nr && gk < kr) ? W[(size_t)gn * kr + gk] : 0.0f;
                        pk[(((size_t)nc * (K / 8) + kc) * ncw + n) * 8 + e] = f2bf(v);
                    }
        std::vector<float> bb(N, 0.0f);
        for (int n = 0; n < nr; ++n) bb[n] = bsrc[n];
        void *dw = nullptr, *db = nullptr;
        if (cudaMalloc(&dw, pk.size() * 2) != cudaSuccess || cudaMalloc(&db, bb.size() * 4) != cudaSuccess) { delete h; return 4; }
        h->allocs.push_back(dw);
        h->allocs.push_back(db);
        cudaMemcpy(dw, pk.data(), pk.size() * 2, cudaMemcpyHostToDevice);
        cudaMemcpy(db, bb.data(), bb.size() * 4, cudaMemcpyHostToDevice);
        d.w[l] = (const __nv_bfloat16*)dw;
        d.b[l] = (const float*)db;
    }
    for (int l = 0; l + 1 < n_layers; ++l)
        if (d.K[l + 1] != d.N[l]) { delete h; return 5; }
    if (cudaFuncSetAttribute(go2_mlp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES) != cudaSuccess) { delete h; return 6; }
    *out = h;
    return 0;
}

void go2_mlp_destroy(go2_mlp* h) {
    if (!h) return;
    cudaSetDevice(h->device);
    for (void* p : h->allocs) cudaFree(p);
    delete h;
}

static int launch_mlp(go2_mlp* h, const float* d_in, int B, const ActOut& o, void* stream) {
    if (!h || !d_in || B <= 0) return 1;
    int grid = (B + TILE_M - 1) / TILE_M;
    MlpJob j;
    j.d = h->d;
    j.in = d_in;
    j.o = o;
    go2_mlp_kernel<<<dim3(grid, 1), NTHREADS, SMEM_BYTES, (cudaStream_t)stream>>>(j, j, B);
    return cudaGetLastError() == cudaSuccess ? 0 : 7;
}

int go2_mlp_forward(go2_mlp* h, const float* d_in, int B, float* d_out, void* stream) {
    ActOut o;
    memset(&o, 0, sizeof(o));
    o.out = d_out;
    return launch_mlp(h, d_in, B, o, stream);
}

int go2_policy_act(go2_mlp* actor, const float* d_obs, int B, const float* d_std, uint64_t seed, uint32_t step, uint32_t env_offset,
                   float* d_actions, float* d_mu, float* d_sigma, float* d_logp, void* stream) {
    if (!d_std || !d_actions) return 1;
    ActOut o;
    memset(&o, 0, sizeof(o));
    o.out = d_mu;
    o.actions = d_actions;
    o.sigma = d_sigma;
    o.logp = d_logp;
    o.std = d_std;
    o.seed_lo = (uint32_t)seed;
    o.seed_hi = (uint32_t)(seed >> 32);
    o.step = step;
    o.env_offset = env_offset;
    return launch_mlp(actor, d_obs, B, o, stream);
}

int go2_policy_act_critic(go2_mlp* actor, go2_mlp* critic, const float* d_obs, const float* d_critic_obs, int B, const float* d_std,
                          uint64_t seed, uint32_t step, uint32_t env_offset, float* d_actions, float* d_mu, float* d_sigma, float* d_logp,
                          float* d_values, void* stream) {
    if (!actor || !critic || !d_obs || !d_critic_obs || !d_std || !d_actions || !d_values || B <= 0) return 1;
    MlpJob ja, jc;
    memset(&ja.o, 0, sizeof(ActOut));
    memset(&jc.o, 0, sizeof(ActOut));
    ja.d = actor->d; ja.in = d_obs;
    ja.o.out = d_mu; ja.o.actions = d_actions; ja.o.sigma = d_sigma; ja.o.logp = d_logp; ja.o.std = d_std;
    ja.o.seed_lo = (uint32_t)seed; ja.o.seed_hi = (uint32_t)(seed >> 32); ja.o.step = step; ja.o.env_offset = env_offset;
    jc.d = critic->d; jc.in = d_critic_obs; jc.o.out = d_values;
    go2_mlp_kernel<<<dim3((B + TILE_M - 1) / TILE_M, 2), NTHREADS, SMEM_BYTES, (cudaStream_t)stream>>>(ja, jc, B);
    return cudaGetLastError() == cudaSuccess ? 0 : 7;
}

int go2_gae(const float* d_rewards, const uint8_t* d_dones, const float* d_values, const float* d_last_values, int T, int B, float gamma,
            float lam, float* d_returns, float* d_advantages, void* stream) {
    if (T <= 0 || B <= 0) return 1;
    go2_gae_kernel<<<(B + 255) / 256, 256, 0, (cudaStream_t)stream>>>(d_rewards, d_dones, d_values, d_last_values, T, B, gamma, lam, d_returns,
                                                                      d_advantages);
    return cudaGetLastError() == cudaSuccess ? 0 : 7;
}

}  // extern "C"

// ---- go2_mlp_forward_noalloc: no cudaMalloc / cudaFree / allocating library call anywhere on this path ----------------------------------
static std::mutex g_na_mu;
static unsigned char* g_na_ptr[GO2_NOALLOC_MAX_DEVICES] = {nullptr};  // DEVICE address of g_noalloc_scratch, per device

static int na_fail(char* err, int err_len, int rc, const char* fmt, ...) {
    if (err && err_len > 0) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(err, (size_t)err_len, fmt, ap);
        va_end(ap);
    }
    return rc;
}

// In host code `g_noalloc_scratch` is only a stand-in symbol: its device address is looked up once per device, with that device current.
static int na_device_scratch(int device, unsigned char** out, char* err, int err_len) {
    if (device < 0 || device >= GO2_NOALLOC_MAX_DEVICES) return na_fail(err, err_len, 3, "device %d outside [0, %d)", device, GO2_NOALLOC_MAX_DEVICES);
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return na_fail(err, err_len, 3, "cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
    std::lock_guard<std::mutex> lk(g_na_mu);
    if (!g_na_ptr[device]) {
        unsigned char* ptr = nullptr;
        e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch);
        if (e != cudaSuccess || !ptr) return na_fail(err, err_len, 9, "cudaGetSymbolAddress(scratch) on device %d: %s", device, cudaGetErrorString(e));
        e = cudaFuncSetAttribute(go2_mlp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
        if (e != cudaSuccess) return na_fail(err, err_len, 6, "cudaFuncSetAttribute(go2_mlp_kernel): %s", cudaGetErrorString(e));
        g_na_ptr[device] = ptr;
    }
    *out = g_na_ptr[device];
    return 0;
}

extern "C" {

int go2_mlp_forward_noalloc(const int32_t* dims, int n_layers, const float* d_params, int device, const float* d_in, int B, float* d_out,
                            void* stream, char* err, int err_len) {
    if (err && err_len > 0) err[0] = 0;
    if (!dims || !d_params || !d_in || !d_out || B <= 0) return na_fail(err, err_len, 1, "null pointer or B <= 0");
    if (n_layers < 1 || n_layers > MAX_LAYERS) return na_fail(err, err_len, 1, "n_layers %d outside [1, %d]", n_layers, MAX_LAYERS);
    if (dims[n_layers] > 16 || dims[n_layers] < 1 || dims[0] > 512 || dims[0] < 1)
        return na_fail(err, err_len, 2, "input width %d (max 512) / output width %d (max 16) not supported", dims[0], dims[n_layers]);
    for (int l = 1; l < n_layers; ++l)
        if (dims[l] > 512 || dims[l] < 1) return na_fail(err, err_len, 2, "hidden width %d of layer %d not in [1, 512]", dims[l], l);
    // the padding rules of go2_mlp_create, and the scratch carve-up (256-byte aligned pieces)
    MlpJob j;
    memset(&j, 0, sizeof(j));
    PackJob pj;
    memset(&pj, 0, sizeof(pj));
    size_t off[MAX_LAYERS][2], need = 0, max_elems = 0;
    const float* p = d_params;
    j.d.n_layers = n_layers;
    j.d.n_in = dims[0];
    j.d.n_out = dims[n_layers];
    for (int l = 0; l < n_layers; ++l) {
        const int kr = dims[l], nr = dims[l + 1];
        int N = (l + 1 < n_layers) ? pad_to(nr, KPAD) : 16;
        if (N > NCHUNK) N = 2 * NCHUNK;
        const int K = l == 0 ? pad_to(kr, KPAD) : j.d.N[l - 1];
        j.d.K[l] = K;
        j.d.N[l] = N;
        off[l][0] = need;
        need += ((size_t)K * N * 2 + 255) / 256 * 256;
        off[l][1] = need;
        need += ((size_t)N * 4 + 255) / 256 * 256;
        PackLayer& L = pj.l[l];
        L.W = p;
        L.bias = p + (size_t)nr * kr;
        p += (size_t)nr * kr + nr;
        L.kr = kr; L.nr = nr; L.K = K; L.N = N; L.ncw = N < NCHUNK ? N : NCHUNK;
        if ((size_t)K * N > max_elems) max_elems = (size_t)K * N;
    }
    if (need > NOALLOC_SCRATCH_BYTES)
        return na_fail(err, err_len, 8, "network needs %zu bytes of packed weights, the static scratch array holds %zu (use go2_mlp_create + go2_mlp_forward)",
                       need, (size_t)NOALLOC_SCRATCH_BYTES);
    unsigned char* scratch = nullptr;
    if (int rc = na_device_scratch(device, &scratch, err, err_len)) return rc;
    for (int l = 0; l < n_layers; ++l) {
        pj.l[l].w_out = (__nv_bfloat16*)(scratch + off[l][0]);
        pj.l[l].b_out = (float*)(scratch + off[l][1]);
        j.d.w[l] = pj.l[l].w_out;
        j.d.b[l] = pj.l[l].b_out;
    }
    j.in = d_in;
    j.o.out = d_out;
    cudaStream_t st = (cudaStream_t)stream;
    int pack_blocks = (int)((max_elems + 256 * 8 - 1) / (256 * 8));
    go2_mlp_pack_kernel<<<dim3(pack_blocks, n_layers), 256, 0, st>>>(pj);
    go2_mlp_kernel<<<dim3((B + TILE_M - 1) / TILE_M, 1), NTHREADS, SMEM_BYTES, st>>>(j, j, B);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return na_fail(err, err_len, 7, "kernel launch: %s", cudaGetErrorString(e));
    return 0;
}

}  // extern "C"
