/* ppo_kernels.h - C-ABI of the device-resident PPO v2 update (BASELINE config 4).
 *
 * The reference's agent is Python/torch (`/root/reference/ppo_v2.py`), driven once per control step by
 * `parallelized_base_v2.py:116-166`. Everything that is not a plain GEMM (those go to cuBLAS from the host
 * layer, `diffusion-piano_b200/ppo/agent.py`) is one of the kernels below; one minibatch update is captured
 * as a CUDA graph and replayed, with its cursor, RNG counter, Adam step count and learning rates living in
 * device memory so that a replay needs no host value.
 *
 *   ppo_relu_ln_fwd / _bwd   <- the `Linear -> ReLU -> LayerNorm [-> Dropout(0.1)]` blocks of Actor / Critic
 *                               (ppo_v2.py:54-63, 79-92): bias add, ReLU, LayerNorm, dropout in one pass over
 *                               the GEMM output; backward gives dY and the bias / gamma / beta gradients.
 *   ppo_policy_head          <- Tanh + Normal(mean, exp(clamp(log_std))) log_prob / entropy + clipped surrogate
 *                               (ppo_v2.py:64,69-74,262-271) and its gradient.
 *   ppo_value_head           <- `nn.MSELoss()(value_pred, b_returns)` (ppo_v2.py:274-275) and its gradient.
 *   ppo_sample_actions       <- `PPOAgent.select_actions` (ppo_v2.py:202-209): dist.sample(), log_prob.sum(1).
 *   ppo_prepare_batch        <- `RunningMeanStd.__call__` (ppo_v2.py:113-131) + returns + the index-wise GAE
 *                               recursion + advantage normalisation (ppo_v2.py:212-249).
 *   ppo_gather_minibatch     <- `DataLoader(TensorDataset(...), batch_size, shuffle=True)` (ppo_v2.py:254-260).
 *   ppo_grad_sumsq + ppo_clip_adam <- `clip_grad_norm_` + `Adam.step` (ppo_v2.py:277-287, 157-168).
 *   ppo_plateau_step         <- `ReduceLROnPlateau.step` (ppo_v2.py:171-184, 305-308).
 *   ppo_advance              <- the loop counters of ppo_v2.py:252-260.
 *
 * Conventions as in piano_step.h: device pointers, caller-owned, contiguous; `stream` is a cudaStream_t;
 * nothing synchronises the host; return 0 = ok, negative = error (ppo_last_error()). All kernels are
 * deterministic (column sums are reduced in a fixed order, no floating-point atomics).
 */
#ifndef PPO_KERNELS_H_
#define PPO_KERNELS_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Device-resident control block of one agent (all fields written by kernels only). */
typedef struct ppo_ctrl {
  int32_t cursor;        /* first sample of the current minibatch within the epoch's permutation */
  int32_t epoch;         /* epochs finished since ppo_ctrl was last reset (selects the permutation) */
  uint64_t rng_counter;  /* bumped once per minibatch and once per select_actions */
  uint64_t seed;
  /* per optimiser [0] = actor, [1] = critic */
  float lr[2];
  int32_t adam_t[2];
  float norm2[2];        /* squared gradient norm of the current minibatch */
  /* ReduceLROnPlateau state */
  float best[2];
  int32_t num_bad[2];
  /* statistics of the last minibatch / batch for logging and the schedulers */
  float actor_loss, entropy, mean_ratio, critic_loss, mean_value, mean_reward_n;
  int32_t pad_;
  /* RunningMeanStd (float64 as in the reference) */
  double rms_mean, rms_var, rms_count;
} ppo_ctrl;

const char* ppo_last_error(void);
const char* ppo_version(void);
int ppo_ctrl_size(void); /* sizeof(ppo_ctrl), for the binding's self-check */

/* ctrl <- zeros, seed, learning rates, plateau bests (+inf for the critic's "min", -inf for the actor's "max"),
 * RunningMeanStd (0, 1, 1e-4). */
int ppo_ctrl_init(ppo_ctrl* ctrl_dev, uint64_t seed, float lr_actor, float lr_critic, void* stream);

/* fp32 GEMMs for small M (a minibatch: 128 rows), row-major contiguous operands, fixed summation order. They
 * replace the cuBLAS calls behind `nn.Linear` forward / backward (ppo_v2.py:54-63,79-92 through autograd) where
 * cuBLAS picks a 32-CTA SIMT kernel; large M (whole-batch inference) stays on cuBLAS.
 *   ppo_gemm_nt: C[M,N] = A[M,K] . B[N,K]^T    (forward:  activations x weight^T)
 *   ppo_gemm_nn: C[M,K] = A[M,N] . B[N,K]      (backward: dA = dY x weight)
 *   ppo_gemm_tn: C[N,K] = A[M,N]^T . B[M,K]    (backward: dW = dY^T x activations) */
int ppo_gemm_nt(const float* A, const float* B, float* C, int M, int N, int K, void* stream);
int ppo_gemm_nn(const float* A, const float* B, float* C, int M, int N, int K, void* stream);
int ppo_gemm_tn(const float* A, const float* B, float* C, int M, int N, int K, void* stream);



/* out = dropout(LayerNorm(relu(y + bias))); rows x H, H a multiple of 32 and <= 256. h_save (post-ReLU) and
 * stats [rows,2] = (mean, rstd) are what the backward needs (NULL at inference; h_save may alias y).
 * dropout_p > 0: keep-mask from mask_in [rows,H] bytes if given, else Philox(seed, rng_counter, rng_stream, i);
 * the mask used is written to mask_out when non-NULL. */
int ppo_relu_ln_fwd(const float* y, const float* bias, const float* gamma, const float* beta, float* out,
                    float* h_save, float* stats, int rows, int H, float dropout_p, const uint8_t* mask_in,
                    uint8_t* mask_out, const ppo_ctrl* ctrl_dev, uint32_t rng_stream, void* stream);

/* Backward of the block above given dout. ppo_relu_ln_bwd_dy writes dy [rows,H] (what the next dA GEMM waits
 * for); ppo_relu_ln_colsums OVERWRITES dbias / dgamma / dbeta [H] with the column sums (it reads dy, so it runs
 * after _dy - on another stream if the caller wants it off its dependent chain); ppo_relu_ln_bwd does both in
 * order on one stream. Deterministic (fixed summation order). */
int ppo_relu_ln_bwd_dy(const float* dout, const float* h, const float* stats, const float* gamma,
                       const uint8_t* mask, float dropout_p, float* dy, int rows, int H, void* stream);
/* ppo_relu_ln_bwd_dy fused with the dA GEMM of the Linear below the block: dy [rows,H] as above, and
 * dA [rows,K] = dy . W with W [H,K] the weight of this block's own Linear (the gradient flowing to the block
 * below). One dependent launch instead of two. */
int ppo_ln_bwd_gemm_nn(const float* dout, const float* h, const float* stats, const float* gamma,
                       const uint8_t* mask, float dropout_p, float* dy, const float* W, float* dA, int rows, int H,
                       int K, void* stream);
int ppo_relu_ln_colsums(const float* dout, const float* h, const float* stats, const uint8_t* mask,
                        float dropout_p, const float* dy, float* dbias, float* dgamma, float* dbeta, int rows,
                        int H, void* stream);
int ppo_relu_ln_bwd(const float* dout, const float* h, const float* stats, const float* gamma,
                    const uint8_t* mask, float dropout_p, float* dy, float* dbias, float* dgamma, float* dbeta,
                    int rows, int H, void* stream);

/* y4 [B,A] is the last Linear's output WITHOUT bias. Writes dz [B,A] (gradient of the actor loss with respect
 * to y4), db4 [A], dlog_std [A], optionally logp_out [B]; ctrl->actor_loss / entropy / mean_ratio.
 * scratch: >= 130 * ceil(B/8) floats; ticket: one zero-initialised uint32 (left at zero). */
int ppo_policy_head(const float* y4, const float* b4, const float* log_std, const float* actions,
                    const float* old_logp, const float* adv, float clip_eps, float entropy_coef, float* dz,
                    float* db4, float* dlog_std, float* logp_out, ppo_ctrl* ctrl_dev, float* scratch,
                    uint32_t* ticket, int B, int A, void* stream);


/* y4 [B] without bias. dv [B] = 2 (v - ret) / B, db4 [1]; ctrl->critic_loss / mean_value. */
int ppo_value_head(const float* y4, const float* b4, const float* returns, float* dv, float* db4,
                   ppo_ctrl* ctrl_dev, int B, void* stream);

/* The same with the critic's last Linear (w4 [H3]) and its dA inside: v = a3 . w4 + b4, dA3 [B,H3] = dv w4. */
int ppo_value_head_fused(const float* a3, const float* w4, const float* b4, const float* returns, float* dv,
                         float* dA3, float* db4, ppo_ctrl* ctrl_dev, int B, int H3, void* stream);


/* values[i] = y4[i] + b4 (critic inference on the whole batch). */
int ppo_add_bias(const float* y4, const float* b4, float* out, int n, void* stream);

/* actions = tanh(y4 + b4) + exp(clamp(log_std)) * N(0,1); logp = sum log_prob. noise_in [N,A] replaces the
 * Philox draws when given (tests). Bumps nothing: the caller runs ppo_advance(..., 0, 0) afterwards. */
int ppo_sample_actions(const float* y4, const float* b4, const float* log_std, const float* noise_in,
                       float* actions, float* logp, const ppo_ctrl* ctrl_dev, int N, int A, void* stream);

/* One block. rewards [N] f32 raw; dones [N] f32 (0/1); values, next_values [N]. Updates the RunningMeanStd in
 * ctrl, writes rewards_n, returns, adv_raw (may be NULL), adv (normalised), ctrl->mean_reward_n. */
int ppo_prepare_batch(const float* rewards, const float* dones, const float* values, const float* next_values,
                      float gamma, float lam, float* rewards_n, float* returns, float* adv_raw, float* adv,
                      ppo_ctrl* ctrl_dev, int N, void* stream);

/* Rows perm[cursor .. cursor+B) of the batch into contiguous minibatch buffers. perm [perm_epochs, N] is an
 * explicit table (row epoch % perm_epochs; tests); perm NULL: the permutation is a keyed Feistel bijection of
 * [0,N) chosen by (seed, epoch, perm_salt). */
int ppo_gather_minibatch(const float* states, const float* actions, const float* old_logp, const float* adv,
                         const float* returns, const int32_t* perm, int perm_epochs, const ppo_ctrl* ctrl_dev,
                         uint32_t perm_salt,
                         int N, int S, int A, int B, float* b_states, float* b_actions, float* b_logp,
                         float* b_adv, float* b_returns, void* stream);
/* The permutation above, materialised (tests): out[i] = perm_epoch(i). */
int ppo_permutation(const ppo_ctrl* ctrl_dev, uint32_t perm_salt, int N, int32_t* out, void* stream);

/* ctrl->norm2[which] = sum(grad^2) (fixed-order two-level reduction); bumps ctrl->adam_t[which].
 * grad 16-byte aligned; scratch >= 592 floats, ticket as above. */
int ppo_grad_sumsq(const float* grad, int n, float* scratch, uint32_t* ticket, ppo_ctrl* ctrl_dev, int which,
                   void* stream);
/* ppo_grad_sumsq for callers that cannot allocate device memory (CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc). NO COUNTERPART IN THE REFERENCE.
 * Same kernel, same launch shape, same result (bit for bit) as ppo_grad_sumsq; the scratch (592 block
 * partial sums + the ticket word) is a file-scope device array inside the library - 2560 bytes on every GPU that
 * loads it, until the process exits - so the call allocates nothing (no cudaMalloc / cudaFree, no allocating
 * library routine). Limit: n <= PPO_GRAD_SUMSQ_STATIC_MAX_N, larger n is rejected (-1, ppo_last_error() names the
 * limit). The array is shared by every call on that GPU and keeps the last call's contents (the call clears the
 * ticket on `stream` first): calls that overlap on the GPU - different streams, or graphs replayed concurrently -
 * RACE; serialise them. The first call on a device resolves the array's address (cudaGetSymbolAddress): make it
 * before any stream capture. */
#define PPO_GRAD_SUMSQ_STATIC_MAX_N (1 << 22) /* floats: 16 MiB of gradient, 10x both PPO networks together */
int ppo_grad_sumsq_static(const float* grad, int n, ppo_ctrl* ctrl_dev, int which, void* stream);

/* n a multiple of 4, buffers 16-byte aligned.
 * g = grad * grad_scale * min(1, max_norm / (sqrt(norm2) * grad_scale + 1e-6)); Adam(b1, b2, eps) with
 * lr = ctrl->lr[which], t = ctrl->adam_t[which]. grad_scale = 1/world_size after a summing all-reduce. */
int ppo_clip_adam(float* theta, const float* grad, float* m, float* v, int n, const ppo_ctrl* ctrl_dev, int which,
                  float max_norm, float grad_scale, float beta1, float beta2, float eps, void* stream);

/* cursor += B; past N: cursor = 0, epoch += 1. rng_counter += 1. (B = 0: only the RNG counter.) */
int ppo_advance(ppo_ctrl* ctrl_dev, int N, int B, void* stream);
/* cursor = 0 (start of an update); epoch keeps counting so that every epoch has its own permutation. */
int ppo_begin_update(ppo_ctrl* ctrl_dev, void* stream);

/* ReduceLROnPlateau(factor, patience, threshold 1e-4 rel, eps 1e-8): actor (which 0) mode max on
 * ctrl->mean_reward_n, critic (which 1) mode min on ctrl->critic_loss. */
int ppo_plateau_step(ppo_ctrl* ctrl_dev, float factor, int patience, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PPO_KERNELS_H_ */
