// ppo_fusion_experiments.cu - kernels that were built, tested and MEASURED SLOWER than the default path of the PPO
// update (profiles/r01_ppo_summary.md), kept out of the product library libppo_kernels.so since round 2:
//   ppo_linear_relu_ln_fwd  Linear + bias + ReLU + LayerNorm + dropout as one thread-block-cluster kernel (row
//                           statistics through distributed shared memory): 21.4 ms per update against 21.1 ms
//   ppo_ln_gemm_nt (+ ppo_dropout_mask)  a block's epilogue as the next GEMM's prologue: 26.6 ms against 21.1 ms
//   ppo_policy_head_fused   the last Linear and its dA inside the policy head: 29.9 ms against 25.3 ms
//   ppo_allreduce_sumsq     gradient all-reduce over NVLink peer memory fused with the squared norm: exact and graph
//                           replayable, 27.9 us against NCCL's 19.2 + 3 us on 2 GPUs
//   ppo_pdl_probe           dependent-launch microbenchmark
// This translation unit includes the product source so that the helpers are shared; tools/experiments/build.sh
// builds libppo_experiments.so (every product entry point + these), which the probes under tools/ load.
#include "../../diffusion-piano_b200/csrc/ppo_kernels.cu"

namespace {

__global__ void pdl_probe_kernel(float* v, int n, int early) {
  if (early) cudaTriggerProgrammaticLaunchCompletion();
  cudaGridDependencySynchronize();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) v[i] += 1.f;
}


// One block forward in ONE launch: Linear + bias + ReLU + LayerNorm + dropout. The tiling is gemm_nt_kernel's (a CTA
// = 4 rows x 32 columns, one warp per 4 x 8 tile); the N / 32 CTAs that share 4 rows form a thread-block CLUSTER and
// exchange their per-row partial sums through distributed shared memory (two cluster barriers: mean, then the
// centred second moment - the same two-pass variance as relu_ln_fwd_kernel), so that the 4 x N outputs are
// normalised without leaving the chip and without a second kernel. N a multiple of 32, <= 256 (cluster <= 8).
__global__ void __launch_bounds__(128) gemm_nt_ln_cluster_kernel(
    const float* __restrict__ A, const float* __restrict__ Bm, const float* __restrict__ bias,
    const float* __restrict__ gamma, const float* __restrict__ beta, float* __restrict__ out,
    float* __restrict__ h_save, float* __restrict__ stats, int M, int N, int K, float p,
    const uint8_t* __restrict__ mask_in, uint8_t* __restrict__ mask_out, const ppo_ctrl* ctrl, uint32_t rng_stream) {
  PPO_PDL_SYNC();  // programmatic dependent launch: see launch_k
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  __shared__ float wpart[4][4];  // [warp][row]
  __shared__ float cpart[2][4];  // this CTA's partial per row: pass 0 sum, pass 1 centred squares
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int m0 = blockIdx.y * 4, n0 = blockIdx.x * 32 + warp * 8;
  const float* ap[4];
  const float* bp[8];
#pragma unroll
  for (int i = 0; i < 4; ++i) ap[i] = A + (size_t)min(m0 + i, M - 1) * K;
#pragma unroll
  for (int j = 0; j < 8; ++j) bp[j] = Bm + (size_t)min(n0 + j, N - 1) * K;
  float acc[32];
#pragma unroll
  for (int t = 0; t < 32; ++t) acc[t] = 0.f;
#pragma unroll 2
  for (int k0 = 0; k0 < K; k0 += 32) {
    const int k = k0 + lane;
    const bool ok = k < K;
    float xv[4], wv[8];
#pragma unroll
    for (int i = 0; i < 4; ++i) xv[i] = ok ? ap[i][k] : 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) wv[j] = ok ? bp[j][k] : 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[i * 8 + j] = fmaf(xv[i], wv[j], acc[i * 8 + j]);
  }
#pragma unroll
  for (int off = 16, n = 16; off >= 1; off >>= 1, n >>= 1) {
    const bool up = (lane & off) != 0;
#pragma unroll
    for (int t = 0; t < 16; ++t)
      if (t < n) {
        const float send = up ? acc[t] : acc[t + n], keep = up ? acc[t + n] : acc[t];
        acc[t] = keep + __shfl_xor_sync(0xffffffffu, send, off);
      }
  }
  // lane l holds the output (row l >> 3, column n0 + (l & 7))
  const int r = lane >> 3, col = n0 + (lane & 7), row = m0 + r;
  const float h = fmaxf(acc[0] + bias[col], 0.f);
  const unsigned nranks = cluster.num_blocks();
  float mu = 0.f, rstd = 0.f;
#pragma unroll
  for (int pass = 0; pass < 2; ++pass) {
    float v = pass == 0 ? h : (h - mu) * (h - mu);
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    v += __shfl_xor_sync(0xffffffffu, v, 4);  // sum over the warp's 8 columns of row r
    if ((lane & 7) == 0) wpart[warp][r] = v;
    __syncthreads();
    if (threadIdx.x < 4) cpart[pass][threadIdx.x] = wpart[0][threadIdx.x] + wpart[1][threadIdx.x] +
                                                    wpart[2][threadIdx.x] + wpart[3][threadIdx.x];
    cluster.sync();  // every CTA's partial of this pass is in its shared memory
    float t = 0.f;
    for (unsigned c = 0; c < nranks; ++c) t += cluster.map_shared_rank(&cpart[pass][0], c)[r];  // rank order: same bits everywhere
    if (pass == 0) mu = t / N;
    else rstd = rsqrtf(t / N + kLnEps);
  }
  if (row < M) {
    float o = (h - mu) * rstd * gamma[col] + beta[col];
    const size_t i = (size_t)row * N + col;
    if (p > 0.f) {
      bool keep;
      if (mask_in != nullptr) {
        keep = mask_in[i] != 0;
      } else {  // the Philox word relu_ln_fwd_kernel uses for (row, col)
        const int j = col >> 5;
        const uint4 rr = rng4(ctrl, rng_stream, (uint32_t)(row * 64 + (col & 31) * 2 + (j >> 2)));
        const uint32_t w = (j & 3) == 0 ? rr.x : (j & 3) == 1 ? rr.y : (j & 3) == 2 ? rr.z : rr.w;
        keep = u01(w) >= p;
      }
      if (mask_out != nullptr) mask_out[i] = keep;
      o = keep ? o / (1.f - p) : 0.f;
    }
    out[i] = o;
    if (h_save != nullptr) h_save[i] = h;
    if (stats != nullptr && blockIdx.x == 0 && warp == 0 && (lane & 7) == 0) {
      stats[2 * row] = mu;
      stats[2 * row + 1] = rstd;
    }
  }
  cluster.sync();  // no CTA may exit while a peer still reads its shared memory
}

// gemm_nt with the PREVIOUS block's epilogue as its prologue: A is not read but computed,
// A[m,:] = dropout(LayerNorm(relu(Yp[m,:] + bias))) for the warp's 4 rows (K = the previous block's width, a multiple
// of 32 and <= 256, so a lane holds its 8 values of each row in registers for the whole k loop), then C = A . B^T as
// above. Every warp of a row group recomputes the same A rows (cheap next to the 8 x K weight loads); the warp with
// blockIdx.x == 0, warp 0 also stores A, the post-ReLU values and the row statistics that the backward needs.
// Dropout keep-masks come from `mask` [M,K] bytes (dropout_mask_kernel, off the dependent chain).
__global__ void __launch_bounds__(128) ln_gemm_nt_kernel(const float* __restrict__ Yp, const float* __restrict__ bias,
                                                         const float* __restrict__ gamma, const float* __restrict__ beta,
                                                         const uint8_t* __restrict__ mask, float p,
                                                         float* __restrict__ Aout, float* __restrict__ Hout,
                                                         float* __restrict__ stats, const float* __restrict__ Bm,
                                                         float* __restrict__ C, int M, int N, int K) {
  PPO_PDL_SYNC();  // programmatic dependent launch: see launch_k
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int m0 = blockIdx.y * 4, n0 = blockIdx.x * 32 + warp * 8;
  if (n0 >= N) return;
  const int vpt = K >> 5;
  const bool writer = blockIdx.x == 0 && warp == 0;
  const float keep_scale = 1.f / (1.f - p);
  float a[4][8];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int row = min(m0 + i, M - 1);
    float h[8], sum = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      h[j] = 0.f;
      if (j < vpt) {
        const int c = lane + 32 * j;
        h[j] = fmaxf(Yp[(size_t)row * K + c] + bias[c], 0.f);
        sum += h[j];
      }
    }
    const float mu = warp_sum(sum) / K;
    float q = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j)
      if (j < vpt) q += (h[j] - mu) * (h[j] - mu);
    const float rstd = rsqrtf(warp_sum(q) / K + kLnEps);
    const bool store = writer && m0 + i < M;
    if (store && lane == 0) {
      stats[2 * row] = mu;
      stats[2 * row + 1] = rstd;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      a[i][j] = 0.f;
      if (j < vpt) {
        const int c = lane + 32 * j;
        float o = (h[j] - mu) * rstd * gamma[c] + beta[c];
        if (p > 0.f) o = mask[(size_t)row * K + c] ? o * keep_scale : 0.f;
        a[i][j] = o;
        if (store) {
          Aout[(size_t)row * K + c] = o;
          Hout[(size_t)row * K + c] = h[j];
        }
      }
    }
  }
  const float* bp[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) bp[j] = Bm + (size_t)min(n0 + j, N - 1) * K;
  float acc[32];
#pragma unroll
  for (int t = 0; t < 32; ++t) acc[t] = 0.f;
#pragma unroll
  for (int kk = 0; kk < 8; ++kk)
    if (kk < vpt) {
      const int k = 32 * kk + lane;
      float wv[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) wv[j] = bp[j][k];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i * 8 + j] = fmaf(a[i][kk], wv[j], acc[i * 8 + j]);
    }
#pragma unroll
  for (int off = 16, n = 16; off >= 1; off >>= 1, n >>= 1) {
    const bool up = (lane & off) != 0;
#pragma unroll
    for (int t = 0; t < 16; ++t)
      if (t < n) {
        const float send = up ? acc[t] : acc[t + n], keep = up ? acc[t + n] : acc[t];
        acc[t] = keep + __shfl_xor_sync(0xffffffffu, send, off);
      }
  }
  const int mi = m0 + (lane >> 3), nj = n0 + (lane & 7);
  if (mi < M && nj < N) C[(size_t)mi * N + nj] = acc[0];
}

// Dropout keep-masks of one network forward (same Philox streams as relu_ln_fwd_kernel: stream id + block index).
__global__ void __launch_bounds__(256) dropout_mask_kernel(uint8_t* __restrict__ mask, int rows, int H, float p,
                                                           const ppo_ctrl* ctrl, uint32_t rng_stream) {
  PPO_PDL_SYNC();  // programmatic dependent launch: see launch_k
  const int lane = threadIdx.x & 31, row = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= rows) return;
  const uint4 r0 = rng4(ctrl, rng_stream, (uint32_t)(row * 64 + lane * 2));
  const uint4 r1 = rng4(ctrl, rng_stream, (uint32_t)(row * 64 + lane * 2 + 1));
  const uint32_t rr[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
  const int vpt = H >> 5;
#pragma unroll
  for (int j = 0; j < 8; ++j)
    if (j < vpt) mask[(size_t)row * H + lane + 32 * j] = u01(rr[j]) >= p;
}


// ---------------------------------------------------------------- gradient all-reduce over NVLink peer memory
// One kernel per network and minibatch instead of NCCL all-reduce + grad_sumsq: every rank's gradient buffer lives in
// symmetric memory (the same allocation mapped into every process; torch.distributed._symmetric_memory does the
// IPC), `grads[p]` is rank p's buffer as seen from this GPU. Each CTA (1) tells the same CTA of every peer that this
// rank's backward kernels are done and waits for theirs, (2) sums its slice over the ranks in rank order - so every
// rank gets bit-identical sums - into the local `out` while accumulating the squared norm, (3) meets the peers again
// before anyone may overwrite a gradient buffer. Flags are monotonically increasing epochs in symmetric memory
// (flags[p] = rank p's flag array, layout [2 barriers][CTA][source rank]); the epoch lives on the device and is
// bumped by the last CTA, so a CUDA-graph replay needs no host value. A wait that does not complete within ~2 s
// sets ctrl-side error word `*error` instead of hanging the GPU.
constexpr int kArCtas = 112, kArMaxWorld = 8;  // 112 CTAs x 256 threads x 2 float4 cover a 0.9 MB gradient
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void peer_barrier(uint32_t* const* flags, int which_barrier, int rank, int world,
                                             uint32_t epoch, uint32_t* error) {
  __syncthreads();
  if (threadIdx.x < world) {
    const int slot = (which_barrier * kArCtas + blockIdx.x) * kArMaxWorld;
    st_release_sys(flags[threadIdx.x] + slot + rank, epoch);           // tell peer threadIdx.x
    const volatile uint32_t* mine = flags[rank] + slot + threadIdx.x;  // hear from peer threadIdx.x
    const long long t0 = clock64();
    while ((int)(*mine - epoch) < 0) {
      if (clock64() - t0 > 4000000000LL) { atomicExch(error, 1u); break; }  // ~2 s at 1.9 GHz
    }
    __threadfence_system();  // acquire: the peers' gradients written before their signal
  }
  __syncthreads();
}

__global__ void __launch_bounds__(256) allreduce_sumsq_kernel(const float* const* grads, uint32_t* const* flags,
                                                              float* __restrict__ out, int n, int rank, int world,
                                                              uint32_t* epoch_dev, uint32_t* error, float* scratch,
                                                              uint32_t* ticket, ppo_ctrl* ctrl, int which) {
  PPO_PDL_SYNC();
  __shared__ float red[8];
  __shared__ bool is_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t epoch = *epoch_dev;
  peer_barrier(flags, 0, rank, world, epoch, error);
  const int n4 = n >> 2;
  float s = 0.f;
  const float4* src[kArMaxWorld];
#pragma unroll
  for (int p = 0; p < kArMaxWorld; ++p) src[p] = reinterpret_cast<const float4*>(grads[p < world ? p : 0]);
  for (int i = blockIdx.x * 256 + threadIdx.x; i < n4; i += gridDim.x * 256) {
    // all ranks' loads of this element are issued before the first add (remote latency paid once per element)
    float4 g[kArMaxWorld];
#pragma unroll
    for (int p = 0; p < kArMaxWorld; ++p)
      if (p < world) g[p] = src[p][i];
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int p = 0; p < kArMaxWorld; ++p)
      if (p < world) { a.x += g[p].x; a.y += g[p].y; a.z += g[p].z; a.w += g[p].w; }
    reinterpret_cast<float4*>(out)[i] = a;
    s += a.x * a.x + a.y * a.y + a.z * a.z + a.w * a.w;
  }
  if (!(which & 256)) peer_barrier(flags, 1, rank, world, epoch, error);
  which &= 255;
  s = warp_sum(s);
  if (lane == 0) red[warp] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < 8; ++w) t += red[w];
    scratch[blockIdx.x] = t;
    __threadfence();
    is_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!is_last || warp != 0) return;
  __threadfence();
  float t = 0.f;
  for (int b = lane; b < (int)gridDim.x; b += 32) t += __ldcg(scratch + b);
  t = warp_sum(t);
  if (lane == 0) {
    ctrl->norm2[which] = t;
    ctrl->adam_t[which] += 1;
    *epoch_dev = epoch + 1;
    *ticket = 0;
  }
}


}  // namespace

extern "C" {

int ppo_linear_relu_ln_fwd(const float* x, const float* W, const float* bias, const float* gamma, const float* beta,
                           float* out, float* h_save, float* stats, int rows, int H, int K, float dropout_p,
                           const uint8_t* mask_in, uint8_t* mask_out, const ppo_ctrl* ctrl, uint32_t rng_stream,
                           void* stream) {
  if (H % 32 != 0 || H > 256 || H <= 0) return invalid("H must be a multiple of 32, <= 256");
  if (rows <= 0 || K <= 0) return invalid("rows / K");
  if (dropout_p > 0.f && mask_in == nullptr && ctrl == nullptr) return invalid("dropout needs ctrl or mask_in");
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(H / 32, (rows + 3) / 4);
  cfg.blockDim = dim3(128);
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = H / 32;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = use_pdl() ? 2 : 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, gemm_nt_ln_cluster_kernel, x, W, bias, gamma, beta, out, h_save, stats,
                                     rows, H, K, dropout_p, mask_in, mask_out, ctrl, rng_stream);
  if (e != cudaSuccess) return fail("ppo_linear_relu_ln_fwd", e);
  PPO_LAUNCH_CHECK("ppo_linear_relu_ln_fwd");
}
int ppo_ln_gemm_nt(const float* y_prev, const float* bias, const float* gamma, const float* beta, const uint8_t* mask,
                   float dropout_p, float* a_out, float* h_out, float* stats, const float* W, float* C, int M, int N,
                   int K, void* stream) {
  if (M <= 0 || N <= 0) return invalid("gemm sizes");
  if (K % 32 != 0 || K > 256 || K <= 0) return invalid("K (the previous block's width) must be a multiple of 32, <= 256");
  if (dropout_p > 0.f && mask == nullptr) return invalid("dropout needs the mask (ppo_dropout_mask)");
  launch_k(ln_gemm_nt_kernel, dim3((N + 31) / 32, (M + 3) / 4), dim3(128), (cudaStream_t)stream, use_pdl(), y_prev,
           bias, gamma, beta, mask, dropout_p, a_out, h_out, stats, W, C, M, N, K);
  PPO_LAUNCH_CHECK("ppo_ln_gemm_nt");
}
int ppo_dropout_mask(uint8_t* mask, int rows, int H, float dropout_p, const ppo_ctrl* ctrl, uint32_t rng_stream,
                     void* stream) {
  if (H % 32 != 0 || H > 256 || H <= 0 || rows <= 0) return invalid("rows / H");
  launch_k(dropout_mask_kernel, dim3((rows + 7) / 8), dim3(256), (cudaStream_t)stream, use_pdl(), mask, rows, H,
           dropout_p, ctrl, rng_stream);
  PPO_LAUNCH_CHECK("ppo_dropout_mask");
}
int ppo_policy_head_fused(const float* a3, const float* W4, const float* b4, const float* log_std,
                          const float* actions, const float* old_logp, const float* adv, float clip_eps,
                          float entropy_coef, float* dz, float* dA3, float* db4, float* dlog_std, float* logp_out,
                          ppo_ctrl* ctrl, float* scratch, uint32_t* ticket, int B, int A, int H3, void* stream) {
  if (A <= 0 || A > 64) return invalid("action dim must be in 1..64");
  if (B <= 0 || H3 <= 0 || H3 > 256) return invalid("B / H3 (<= 256)");
  if (scratch == nullptr || ticket == nullptr || a3 == nullptr || W4 == nullptr || dA3 == nullptr)
    return invalid("null operand");
  launch_k(policy_head_kernel<true>, dim3((B + kHeadRows - 1) / kHeadRows), dim3(256), (cudaStream_t)stream, use_pdl(), 
      nullptr, b4, log_std, actions, old_logp, adv, clip_eps, entropy_coef, dz, db4, dlog_std, logp_out, ctrl, scratch,
      ticket, B, A, a3, W4, dA3, H3);
  PPO_LAUNCH_CHECK("ppo_policy_head_fused");
}
int ppo_pdl_probe(float* v, int n, int mode, void* stream) {
  if (n <= 0) return invalid("n");
  cudaError_t e = launch_k(pdl_probe_kernel, dim3((n + 255) / 256), dim3(256), (cudaStream_t)stream, mode > 0, v, n,
                           mode > 1 ? 1 : 0);
  if (e != cudaSuccess) return fail("ppo_pdl_probe", e);
  return 0;
}
int ppo_allreduce_sumsq(const float* const* grads_dev, uint32_t* const* flags_dev, float* out, int n, int rank,
                        int world, uint32_t* epoch_dev, uint32_t* error_dev, float* scratch, uint32_t* ticket,
                        ppo_ctrl* ctrl, int which, void* stream) {
  if (n <= 0 || (n & 3) != 0 || ((which & 255) != 0 && (which & 255) != 1)) return invalid("n (a multiple of 4) / which");
  if (world < 1 || world > kArMaxWorld || rank < 0 || rank >= world) return invalid("rank / world (<= 8)");
  launch_k(allreduce_sumsq_kernel, dim3(kArCtas), dim3(256), (cudaStream_t)stream, use_pdl(), grads_dev, flags_dev,
           out, n, rank, world, epoch_dev, error_dev, scratch, ticket, ctrl, which);
  PPO_LAUNCH_CHECK("ppo_allreduce_sumsq");
}
int ppo_allreduce_flag_words(void) { return 2 * kArCtas * kArMaxWorld; }

}  // extern "C"
