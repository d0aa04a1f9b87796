/* go2_b200 - C ABI of the B200-native batched Go2 environment step.
 *
 * This is the drop-in boundary for the hot path `Go2Env.step(actions)` of
 * saifahmadgit/go2-sim2real-locomotion-rl.  The reference has no FFI for this path: its env classes
 * call the Genesis scene API in-process (examples/locomotion/final/go2_env_walk.py:985-1109 ->
 * genesis/engine/scene.py:979 Scene.step -> genesis/engine/solvers/rigid/rigid_solver.py:1116 substep).
 * Each entry point below names the reference interface it replaces.  Plain pointers and sizes only; no
 * exceptions cross the boundary (int return codes, 0 = ok, message via go2_last_error).
 *
 * Memory: the library owns every device buffer; go2_buffer_info() hands out raw device pointers that the
 * host side wraps as non-owning views (lifetime = handle).  Threading: one host thread per handle; all
 * work is enqueued on the stream passed by the caller; distinct handles are independent.
 */
#ifndef GO2_B200_H
#define GO2_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct go2_handle go2_handle;

/* sizeof(EnvConfig), sizeof(StepParams), B_COUNT, STATS_LEN - lets a binding verify its struct mirrors. */
void go2_abi_sizes(int* out4);

/* Replaces gs.Scene(...) + add_entity(plane, go2) + scene.build(n_envs)  (go2_env_walk.py:211-244,
 * rigid_solver.py:337 RigidSolver.build).  `model_iv/model_fv` is the flat model table produced by
 * go2_b200.model.flatten_model; `cfg` points to an EnvConfig (csrc/go2_types.cuh). */
int go2_create(const int32_t* model_iv, int n_iv, const double* model_fv, int n_fv, const void* cfg, int cfg_bytes,
               int device, go2_handle** out);
void go2_destroy(go2_handle* h);
const char* go2_last_error(go2_handle* h);

/* Buffer table (csrc/go2_buffers.cuh BufId).  dtype: 0 = f32, 1 = i32, 2 = u8.  rows/cols describe the
 * 2-D shape in memory order. Replaces the qd_to_torch zero-copy views of genesis/utils/misc.py. */
int go2_buffer_info(go2_handle* h, int buf_id, void** dev_ptr, int* dtype, int* rows, int* cols);

/* One control tick for every env: replaces Go2Env.step (go2_env_walk.py:985-1109 /
 * go2_env_base.py:123-192) including scene.step() (2 substeps).  `d_actions`: device f32 [B][A].
 * `params` points to a StepParams.  Asynchronous on `stream` (a cudaStream_t, 0 = default stream).
 * Two launches: the fused step kernel and the post-pass kernel (global physics DR of reset_idx,
 * go2_env_walk.py:737-756,803-848, + statistics accumulation).  With StepParams.defer_post = 1 only the
 * step kernel is launched, so that a multi-rank caller can all-reduce this tick's statistics row
 * (B_STATS row go2_tick_parity()) before it calls go2_post with the same StepParams. */
int go2_step(go2_handle* h, const float* d_actions, const void* params, int params_bytes, void* stream);
int go2_post(go2_handle* h, const void* params, int params_bytes, void* stream);
/* Double-buffer index (B_STATS / B_GLOBALS row) the next step launch - or a pending deferred post-pass - uses. */
int go2_tick_parity(go2_handle* h);

/* Replaces Go2Env.reset() -> reset_idx(all envs) (go2_env_walk.py:1242-1245). */
int go2_reset_all(go2_handle* h, const void* params, int params_bytes, void* stream);

/* Same tick with HOST buffers: copies actions H2D, steps, delivers obs / rew / done (and, if non-NULL, privileged obs and
 * time_outs) to the host buffers and synchronises the stream before returning.  Pinned obs / rew / done buffers are written
 * directly by the step kernel (mapped host memory, overlapped with the kernel); pageable ones are staged through D2H copies. */
int go2_step_host(go2_handle* h, const float* h_actions, const void* params, int params_bytes, float* h_obs, float* h_rew,
                  uint8_t* h_done, float* h_priv, float* h_timeouts, void* stream);
/* First half of go2_step_host for callers that insert work between the step and the post-pass (StepParams.defer_post: the
 * multi-rank statistics all-reduce): H2D of the actions + step kernel.  If h_obs / h_rew / h_done are pinned (device-mappable)
 * the step kernel writes them itself as its warps finish - no device-to-host copy - and *staged = 0; otherwise *staged = 1 and
 * the caller copies B_OBS / B_REW / B_RESET.  The results are complete once `stream` has been synchronised. */
int go2_step_host_begin(go2_handle* h, const float* h_actions, const void* params, int params_bytes, float* h_obs, float* h_rew,
                        uint8_t* h_done, void* stream, int* staged);

/* Episode statistics accumulated on the device since the previous read (n_resets, n_timeouts, sum
 * tracking/sec, number of env-ticks with an errno bit, then the per-reward-term sums): replaces the per-key
 * `.mean().item()` host syncs of reset_idx (go2_env_walk.py:1228-1235) and the curriculum aggregates
 * (:688-729).  go2_read_stats copies the accumulator to the host, clears it and synchronises.  The async
 * pair does the same without stalling the tick: go2_read_stats_async enqueues copy + clear into pinned slot
 * 0 or 1; go2_stats_fetch returns 0 and the numbers once that copy has completed (block = 1 waits,
 * block = 0 returns 1 if it has not). */
int go2_read_stats(go2_handle* h, float* host_out, int n, void* stream);
int go2_read_stats_async(go2_handle* h, int slot, void* stream);
int go2_stats_fetch(go2_handle* h, int slot, float* host_out, int n, int block);

/* Recompute link poses / velocities of every env into B_LINK_POS / B_LINK_VEL
 * (replaces RigidEntity.get_links_pos / get_links_vel, rigid_entity.py:2241-2406). */
int go2_export_links(go2_handle* h, void* stream);

/* Launch shape of the step kernel: out3 = {blocks, threads per block, 1 if the 255-register instantiation}. */
void go2_launch_shape(go2_handle* h, int* out3);

/* Replace the ground box by a heightfield terrain shared by all envs: replaces
 * scene.add_entity(gs.morphs.Terrain(height_field, horizontal_scale, vertical_scale, pos))
 * (go2_env_stair.py:423-433; collider.py:374-394 _init_terrain_state).  `hf_m`: host f32 [rows][cols] heights in
 * metres (height_field * vertical_scale), cell (r, c) at origin + (r, c) * h_scale.  Call before the first step. */
int go2_set_terrain(go2_handle* h, const float* hf_m, int rows, int cols, float h_scale, const float* origin3);

/* Number of kernels this library has launched on the handle since creation. */
long long go2_launch_count(go2_handle* h);

#ifdef __cplusplus
}
#endif
#endif
