/* go2_b200 - C ABI of the fused rollout-side networks: what rsl-rl 2.2.4 runs between two
 * env.step calls while collecting a rollout.  Same conventions as go2_b200.h: plain pointers and sizes, int return
 * codes (0 = ok), all work enqueued on the caller's stream, device pointers are f32 row-major.
 *
 * The kernels are tcgen05 / TMEM / TMA-bulk code for sm_100a (csrc/go2_policy.cu): bf16 operands, fp32 accumulation.
 */
#ifndef GO2_POLICY_H
#define GO2_POLICY_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct go2_mlp go2_mlp;

/* An ELU MLP with `n_layers` Linear layers (<= 4), dims[0..n_layers] = input, hidden..., output widths (hidden <= 512,
 * output <= 16).  `params` = for each layer the torch.nn.Linear weight [out][in] row-major followed by its bias [out]
 * (host f32).  Replaces rsl_rl/modules/actor_critic.py ActorCritic.actor / .critic (nn.Sequential of Linear + ELU). */
int go2_mlp_create(const int32_t* dims, int n_layers, const float* params, int device, go2_mlp** out);
void go2_mlp_destroy(go2_mlp* h);

/* d_out[B][dims[n_layers]] = mlp(d_in[B][dims[0]]).  Replaces ActorCritic.act_inference / ActorCritic.evaluate. */
int go2_mlp_forward(go2_mlp* h, const float* d_in, int B, float* d_out, void* stream);

/* Replaces ActorCritic.act + get_actions_log_prob (Normal(mean, std).sample(), log_prob summed over actions) as used
 * by PPO.act (rsl_rl/algorithms/ppo.py): d_mu = actor(d_obs), d_actions = d_mu + d_std * z with z ~ N(0, 1) from a
 * counter-based generator keyed by (seed, env_offset + row, step), d_sigma = broadcast d_std, d_logp[B].
 * d_mu / d_sigma / d_logp may be NULL. */
int go2_policy_act(go2_mlp* actor, const float* d_obs, int B, const float* d_std, uint64_t seed, uint32_t step, uint32_t env_offset,
                   float* d_actions, float* d_mu, float* d_sigma, float* d_logp, void* stream);

/* go2_policy_act for the actor and go2_mlp_forward for the critic (d_values[B] = critic(d_critic_obs)) in ONE launch:
 * what PPO.act does per tick (ppo.py: actions = actor_critic.act(obs); values = actor_critic.evaluate(critic_obs)). */
int go2_policy_act_critic(go2_mlp* actor, go2_mlp* critic, const float* d_obs, const float* d_critic_obs, int B, const float* d_std,
                          uint64_t seed, uint32_t step, uint32_t env_offset, float* d_actions, float* d_mu, float* d_sigma, float* d_logp,
                          float* d_values, void* stream);

/* Replaces RolloutStorage.compute_returns (rsl_rl/storage/rollout_storage.py) without the advantage normalisation:
 * rewards / values / returns / advantages are [T][B] f32, dones [T][B] u8, last_values [B]. */
int go2_gae(const float* d_rewards, const uint8_t* d_dones, const float* d_values, const float* d_last_values, int T, int B, float gamma,
            float lam, float* d_returns, float* d_advantages, void* stream);

/* go2_mlp_forward without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * d_out[B][dims[n_layers]] = mlp(d_in[B][dims[0]]) like go2_mlp_create + go2_mlp_forward, but `d_params` is DEVICE f32 memory of the
 * caller in the layout go2_mlp_create takes on the host (per layer: Linear weight [out][in] row-major, then bias [out]); it is packed
 * on the device into a file-scope `static __device__` array of GO2_MLP_NOALLOC_SCRATCH_BYTES = 1 MiB that every GPU loading the
 * library holds until exit, and the same tcgen05 kernel runs from there.  The call performs no cudaMalloc / cudaFree / allocating
 * library routine; the first call on a device looks up the array's device address (cudaGetSymbolAddress) and sets the kernel's
 * shared-memory attribute - make it once outside a graph capture.  A network whose packed image exceeds the array is rejected with
 * rc 8 and a message in `err` (NUL-terminated, may be NULL).  All calls on a device share the array: two calls that overlap on the
 * GPU (different streams, no ordering) race; calls on one stream are safe, each call rewrites every byte it reads.
 * rc: 0 ok, 1 bad argument, 2 unsupported widths, 3 device, 6 / 9 set-up, 7 launch, 8 scratch too small. */
#define GO2_MLP_NOALLOC_SCRATCH_BYTES (1u << 20)
int go2_mlp_forward_noalloc(const int32_t* dims, int n_layers, const float* d_params, int device, const float* d_in, int B, float* d_out,
                            void* stream, char* err, int err_len);

#ifdef __cplusplus
}
#endif
#endif
