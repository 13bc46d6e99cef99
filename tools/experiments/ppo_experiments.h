/* ppo_experiments.h - entry points of tools/experiments/ppo_fusion_experiments.cu (measured slower than the default
 * path; not part of the product library). Declarations as they stood in include/ppo_kernels.h in round 1. */
#ifndef PPO_EXPERIMENTS_H_
#define PPO_EXPERIMENTS_H_
#include "../../include/ppo_kernels.h"
#ifdef __cplusplus
extern "C" {
#endif

/* One block forward in one launch: out [rows,H] = dropout(LayerNorm(relu(x [rows,K] . W [H,K]^T + bias))), with
 * h_save / stats / masks as ppo_relu_ln_fwd. The H / 32 CTAs that share 4 rows form a thread-block cluster and
 * exchange their row partial sums through distributed shared memory. H a multiple of 32, <= 256. */
int ppo_linear_relu_ln_fwd(const float* x, const float* W, const float* bias, const float* gamma, const float* beta,
                           float* out, float* h_save, float* stats, int rows, int H, int K, float dropout_p,
                           const uint8_t* mask_in, uint8_t* mask_out, const ppo_ctrl* ctrl_dev, uint32_t rng_stream,
                           void* stream);

/* ppo_gemm_nt with the previous block's epilogue as its prologue (one dependent launch instead of two):
 * A = dropout(LayerNorm(relu(y_prev + bias))) [M,K] is computed on the fly (K = that block's width, a multiple of 32,
 * <= 256), C [M,N] = A . W[N,K]^T; a_out / h_out [M,K] and stats [M,2] receive what ppo_relu_ln_fwd would have
 * written (h_out must not alias y_prev here). dropout_p > 0 reads keep-masks from `mask` [M,K], made by
 * ppo_dropout_mask with the Philox stream relu_ln_fwd would use (rng_stream + block index). */
int ppo_ln_gemm_nt(const float* y_prev, const float* bias, const float* gamma, const float* beta,
                   const uint8_t* mask, float dropout_p, float* a_out, float* h_out, float* stats, const float* W,
                   float* C, int M, int N, int K, void* stream);

int ppo_dropout_mask(uint8_t* mask, int rows, int H, float dropout_p, const ppo_ctrl* ctrl_dev, uint32_t rng_stream,
                     void* stream);

/* The same with the last Linear and its dA inside (one dependent launch instead of three): a3 [B,H3] is the third
 * block's output, W4 [A,H3] the last Linear's weight; also writes dA3 [B,H3] = dz . W4. H3 <= 256. */
int ppo_policy_head_fused(const float* a3, const float* W4, const float* b4, const float* log_std,
                          const float* actions, const float* old_logp, const float* adv, float clip_eps,
                          float entropy_coef, float* dz, float* dA3, float* db4, float* dlog_std, float* logp_out,
                          ppo_ctrl* ctrl_dev, float* scratch, uint32_t* ticket, int B, int A, int H3, void* stream);

/* measurement probe: v[i] += 1, launched plainly (mode 0), with programmatic dependent launch (1), or with it
 * and an early trigger (2) - the per-node cost of a dependent chain (tools/microbench_chain.py). */
int ppo_pdl_probe(float* v, int n, int mode, void* stream);

/* Gradient all-reduce over NVLink peer memory fused with ppo_grad_sumsq (several GPUs, one process each):
 * grads_dev [world] = every rank's gradient buffer as mapped into THIS process (symmetric memory), flags_dev
 * [world] = every rank's zero-initialised flag array of ppo_allreduce_flag_words() uint32 (symmetric memory too),
 * out [n] = sum over ranks (rank order, bit-identical on all ranks), ctrl->norm2[which] = |out|^2,
 * ctrl->adam_t[which] += 1. epoch_dev: one uint32 initialised to 1; error_dev: one uint32, set to 1 if a peer did
 * not arrive within ~2 s. Every rank must launch it the same number of times (stands for the NCCL all-reduce the
 * reference's sharded form needs between `backward()` and `clip_grad_norm_`, ppo_v2.py:277-287). `which | 256`
 * skips the closing barrier - a TIMING experiment only (tools/probe_peer_allreduce.py), unsafe otherwise.
 * Status: exact against NCCL on 2 x B200 and graph-replayable, but 27.9 us per call against 19.2 us for NCCL's
 * all-reduce of the same 0.91 MB (+ ~3 us for ppo_grad_sumsq): each cross-GPU flag barrier costs ~9 us here, so
 * the agent keeps NCCL (profiles/r01_ppo_summary.md). world <= 8. */
int ppo_allreduce_sumsq(const float* const* grads_dev, uint32_t* const* flags_dev, float* out, int n, int rank,
                        int world, uint32_t* epoch_dev, uint32_t* error_dev, float* scratch, uint32_t* ticket,
                        ppo_ctrl* ctrl_dev, int which, void* stream);

int ppo_allreduce_flag_words(void);

#ifdef __cplusplus
}
#endif
#endif
