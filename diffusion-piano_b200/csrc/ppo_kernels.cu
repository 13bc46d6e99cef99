// ppo_kernels.cu - device side of the PPO v2 update (include/ppo_kernels.h). sm_100a.
//
// Everything here is small-batch work (128-row minibatches, 256-wide layers): the kernels are sized so that
// one minibatch update is a short chain of launches inside a CUDA graph, every reduction has a fixed order
// (bit-reproducible runs), and all loop state lives in `ppo_ctrl` on the device.
#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "../../include/ppo_kernels.h"

namespace {

thread_local char g_err[256] = "";

int fail(const char* what, cudaError_t e) {
  snprintf(g_err, sizeof(g_err), "%s: %s", what, cudaGetErrorString(e));
  return -2;
}
int invalid(const char* what) {
  snprintf(g_err, sizeof(g_err), "invalid argument: %s", what);
  return -1;
}
#define PPO_LAUNCH_CHECK(name)                 \
  do {                                         \
    cudaError_t e_ = cudaGetLastError();       \
    if (e_ != cudaSuccess) return fail(name, e_); \
    return 0;                                  \
  } while (0)

// Programmatic dependent launch (sm_90+): a kernel launched with this attribute may start while its predecessor in
// the stream is still running; it must call cudaGridDependencySynchronize() before touching anything the
// predecessor reads or writes. Every kernel launched through launch_k does so as its first statement, so memory
// semantics stay those of a plain stream and only the launch latency of a dependent chain is overlapped.
template <typename... KArgs, typename... Args>
cudaError_t launch_k(void (*kern)(KArgs...), dim3 grid, dim3 block, cudaStream_t s, bool pdl, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = 0;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

#ifdef PPO_EARLY_TRIGGER
#define PPO_PDL_SYNC() do { cudaTriggerProgrammaticLaunchCompletion(); cudaGridDependencySynchronize(); } while (0)
#else
#define PPO_PDL_SYNC() cudaGridDependencySynchronize()
#endif
bool use_pdl() {  // PPO_PDL=0 turns programmatic dependent launch off (A/B measurements)
  static const bool on = [] { const char* e = getenv("PPO_PDL"); return e == nullptr || e[0] != '0'; }();
  return on;
}

constexpr float kLnEps = 1e-5f;
constexpr float kHalfLog2Pi = 0.91893853320467274178f;

// ---------------------------------------------------------------- Philox4x32-10 (counter-based RNG)
__device__ __forceinline__ uint4 philox4x32(uint4 c, uint2 k) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
    uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += 0x9E3779B9u;
    k.y += 0xBB67AE85u;
  }
  return c;
}
__device__ __forceinline__ uint4 rng4(const ppo_ctrl* ctrl, uint32_t stream_id, uint32_t idx) {
  uint64_t ctr = ctrl->rng_counter, seed = ctrl->seed;
  return philox4x32(make_uint4(idx, stream_id, (uint32_t)ctr, (uint32_t)(ctr >> 32)),
                    make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
}
__device__ __forceinline__ float u01(uint32_t x) { return ((x >> 8) + 0.5f) * (1.0f / 16777216.0f); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ---------------------------------------------------------------- Linear epilogue: bias, ReLU, LayerNorm, dropout
// One warp per row, lane l owns columns l, l+32, ... (coalesced 128-byte rows of the GEMM output).
__global__ void __launch_bounds__(256) relu_ln_fwd_kernel(
    const float* y /* may alias h_save */, const float* __restrict__ bias, const float* __restrict__ gamma,
    const float* __restrict__ beta, float* __restrict__ out, float* h_save, float* __restrict__ stats, int rows,
    int H, float p, const uint8_t* __restrict__ mask_in, uint8_t* __restrict__ mask_out, const ppo_ctrl* ctrl,
    uint32_t rng_stream) {
  PPO_PDL_SYNC();  // programmatic dependent launch: see launch_k
  const int lane = threadIdx.x & 31, row = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int vpt = H >> 5;
  float h[8];
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j)
    if (j < vpt) {
      const int c = lane + 32 * j;
      h[j] = fmaxf(y[(size_t)row * H + c] + bias[c], 0.f);
      s += h[j];
    }
  const float mu = warp_sum(s) / H;
  float q = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j)
    if (j < vpt) q += (h[j] - mu) * (h[j] - mu);
  const float rstd = rsqrtf(warp_sum(q) / H + kLnEps);
  if (stats != nullptr && lane == 0) {
    stats[2 * row] = mu;
    stats[2 * row + 1] = rstd;
  }
  uint4 r0 = make_uint4(0, 0, 0, 0), r1 = r0;
  if (p > 0.f && mask_in == nullptr) {
    r0 = rng4(ctrl, rng_stream, (uint32_t)(row * 64 + lane * 2));
    r1 = rng4(ctrl, rng_stream, (uint32_t)(row * 64 + lane * 2 + 1));
  }
  const uint32_t rr[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
  const float keep_scale = 1.f / (1.f - p);
#pragma unroll
  for (int j = 0; j < 8; ++j)
    if (j < vpt) {
      const int c = lane + 32 * j;
      const size_t i = (size_t)row * H + c;
      float o = (h[j] - mu) * rstd * gamma[c] + beta[c];
      if (p > 0.f) {
        const bool keep = mask_in != nullptr ? mask_in[i] != 0 : u01(rr[j]) >= p;
        if (mask_out != nullptr) mask_out[i] = keep;
        o = keep ? o * keep_scale : 0.f;
      }
      out[i] = o;
      if (h_save != nullptr) h_save[i] = h[j];
    }
}

// Backward of the block, row part: dy (what the next dA GEMM waits for). Warp per row, 8 rows per CTA.
__global__ void __launch_bounds__(256) relu_ln_bwd_dy_kernel(
    const float* __restrict__ dout, const float* __restrict__ hbuf, const float* __restrict__ stats,
    const float* __restrict__ gamma, const uint8_t* __restrict__ mask, float p, float* __restrict__ dy, int rows,
    int H) {
  PPO_PDL_SYNC();  // programmatic dependent launch: see launch_k
  const int lane = threadIdx.x & 31, row = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int vpt = H >> 5;
  float dx[8], xh[8], hv[8];
  float s1 = 0.f, s2 = 0.f;
  const float keep_scale = 1.f / (1.f - p), mu = stats[2 * row], rstd = stats[2 * row + 1];
#pragma unroll
  for (int j = 0; j < 8; ++j)
    if (j < vpt) {
      const int c = lane + 32 * j;
      const size_t i = (size_t)row * H + c;
      float g = dout[i];
      if (p > 0.f) g = mask[i] ? g * keep_scale : 0.f;
      hv[j] = hbuf[i];
      xh[j] = (hv[j] - mu) * rstd;
      dx[j] = g * gamma[c];
      s1 += dx[j];
      s2 += dx[j] * xh[j];
    }
  s1 = warp_sum(s1) / H;
  s2 = warp_sum(s2) / H;
#pragma unroll
  for (int j = 0; j < 8; ++j)
    if (j < vpt) {
      const float dh = rstd * (dx[j] - s1 - xh[j] * s2);
      dy[(size_t)row * H + lane + 32 * j] = hv[j] > 0.f ? dh : 0.f;
    }
}

// Backward of the block, column part (off the dependent chain): d bias = sum dy, d gamma = sum dout * xhat,
// d beta = sum dout. One CTA per 32 columns, 8 warps striding the rows, summed through shared memory in warp order.
__global__ void __launch_bounds__(256) relu_ln_colsums_kernel(
    const float* __restrict__ dout, const float* __restrict__ hbuf, const float* __restrict__ stats,
    const uint8_t* __restrict__ mask, float p, const float* __restrict__ dy, float* __restrict__ dbias,
    float* __restrict__ dgamma, float* __restrict__ dbeta, int rows, int H) {
  PPO_PDL_SYNC();  // programmatic dependent launch: see launch_k
  __shared__ float red[3][8][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, c = blockIdx.x * 32 + lane;
  const float keep_scale = 1.f / (1.f - p);
  float sb = 0.f, sg = 0.f, se = 0.f;
  for (int r = warp; r < rows; r += 8) {
    const size_t i = (size_t)r * H + c;
    float g = dout[i];
    if (p > 0.f) g = mask[i] ? g * keep_scale : 0.f;
    sb += dy[i];
    sg += g * ((hbuf[i] - stats[2 * r]) * stats[2 * r + 1]);
    se += g;
  }
  red[0][warp][lane] = sb;
  red[1][warp][lane] = sg;
  red[2][warp][lane] = se;
  __syncthreads();
  if (warp < 3) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += red[warp][w][lane];
    (warp == 0 ? dbias : warp == 1 ? dgamma : dbeta)[c] = t;
  }
}

// ---------------------------------------------------------------- small-M fp32 GEMMs
// The minibatch GEMMs are 128 x 256 x 256..329: cuBLAS answers them with a 32x32x16 SIMT kernel on 32 CTAs of two
// warps, 4.7 - 8 us per call inside the graph (tools/microbench_chain.py) - most of a minibatch's 110 us. These
// three kernels spread the same FFMA work over 100 - 250 CTAs with every global load coalesced and no shared-memory
// staging of the big operand; fixed summation order (bit-reproducible).

// C[M,N] = A[M,K] . B[N,K]^T (forward: activations x weight rows). One warp per 4 x 8 output tile, lanes along k:
// both operands are read along their contiguous dimension; the 32 partial sums per lane are reduced by a
// transposing butterfly (31 shuffles) that leaves output element l of the tile on lane l.
__global__ void __launch_bounds__(128) gemm_nt_kernel(const float* __restrict__ A, const float* __restrict__ Bm,
                                                      float* __restrict__ C, int M, int N, int K) {
  PPO_PDL_SYNC();  // programmatic dependent launch: see launch_k
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int m0 = blockIdx.y * 4, n0 = blockIdx.x * 32 + warp * 8;
  if (n0 >= N) return;
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
  const int mi = m0 + (lane >> 3), nj = n0 + (lane & 7);
  if (mi < M && nj < N) C[(size_t)mi * N + nj] = acc[0];
}

// C[M,K] = A[M,N] . B[N,K] (dA = dY x W). CTA = 8 output rows x 32 output columns; lane = column k (B rows are read
// coalesced), the 8 warps split the contraction index n, so that every warp has all its B loads in flight at once;
// the 8 partial tiles are added through shared memory in warp order.
__global__ void __launch_bounds__(256) gemm_nn_kernel(const float* __restrict__ A, const float* __restrict__ Bm,
                                                      float* __restrict__ C, int M, int N, int K) {
  PPO_PDL_SYNC();  // programmatic dependent launch: see launch_k
  __shared__ __align__(16) float As[8][256];  // reused for the partial tiles [8 warps][8 rows][32 k]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int k = blockIdx.x * 32 + lane, m0 = blockIdx.y * 8;
  const bool kok = k < K;
  float acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
  for (int nb = 0; nb < N; nb += 256) {
    const int nc = min(256, N - nb);
    const int npw = (((nc + 7) >> 3) + 3) & ~3;  // n per warp, a multiple of 4
    __syncthreads();
    for (int t = threadIdx.x; t < 8 * 256; t += 256) {
      const int i = t >> 8, n = t & 255;
      As[i][n] = (n < nc && m0 + i < M) ? A[(size_t)(m0 + i) * N + nb + n] : 0.f;
    }
    __syncthreads();
    const int nlo = warp * npw;
    if (kok && nlo < nc) {
      const float* bcol = Bm + (size_t)(nb + nlo) * K + k;
#pragma unroll 8
      for (int nn = 0; nn < npw; nn += 4) {
        float bv[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) bv[u] = nlo + nn + u < nc ? bcol[(size_t)(nn + u) * K] : 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float4 av = *reinterpret_cast<const float4*>(&As[i][nlo + nn]);
          acc[i] = fmaf(av.x, bv[0], acc[i]);
          acc[i] = fmaf(av.y, bv[1], acc[i]);
          acc[i] = fmaf(av.z, bv[2], acc[i]);
          acc[i] = fmaf(av.w, bv[3], acc[i]);
        }
      }
    }
  }
  __syncthreads();
  float* part = &As[0][0];
#pragma unroll
  for (int i = 0; i < 8; ++i) part[(warp * 8 + i) * 32 + lane] = acc[i];
  __syncthreads();
  {
    const int i = warp;  // 256 threads = 8 rows x 32 columns
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += part[(w * 8 + i) * 32 + lane];
    if (kok && m0 + i < M) C[(size_t)(m0 + i) * K + k] = t;
  }
}

// LayerNorm / ReLU / dropout backward of a block fused with the dA GEMM that consumes it:
// dy[M,H] = block_backward(dout, h, stats, gamma, mask) (as relu_ln_bwd_dy_kernel), C[M,K] = dy . B[H,K].
// Same tiling as gemm_nn_kernel (8 rows x 32 columns per CTA); every CTA of a row block recomputes its 8 dy rows
// straight into shared memory (warp per row), the CTA with blockIdx.x == 0 also stores them for the dW GEMM and
// the column sums. Saves one dependent launch per block on the backward chain.
__global__ void __launch_bounds__(256) ln_bwd_gemm_nn_kernel(
    const float* __restrict__ dout, const float* __restrict__ hbuf, const float* __restrict__ stats,
    const float* __restrict__ gamma, const uint8_t* __restrict__ mask, float p, float* __restrict__ dy,
    const float* __restrict__ Bm, float* __restrict__ C, int M, int H, int K) {
  PPO_PDL_SYNC();  // programmatic dependent launch: see launch_k
  __shared__ __align__(16) float As[8][256];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int k = blockIdx.x * 32 + lane, m0 = blockIdx.y * 8;
  const bool kok = k < K;
  {
    const int row = m0 + warp, vpt = H >> 5;
    const bool live = row < M;
    float dx[8], xh[8], hv[8];
    float s1 = 0.f, s2 = 0.f;
    const float keep_scale = 1.f / (1.f - p);
    const float mu = live ? stats[2 * row] : 0.f, rstd = live ? stats[2 * row + 1] : 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      dx[j] = xh[j] = hv[j] = 0.f;
      if (j < vpt && live) {
        const int c = lane + 32 * j;
        const size_t i = (size_t)row * H + c;
        float g = dout[i];
        if (p > 0.f) g = mask[i] ? g * keep_scale : 0.f;
        hv[j] = hbuf[i];
        xh[j] = (hv[j] - mu) * rstd;
        dx[j] = g * gamma[c];
        s1 += dx[j];
        s2 += dx[j] * xh[j];
      }
    }
    s1 = warp_sum(s1) / H;
    s2 = warp_sum(s2) / H;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = lane + 32 * j;
      float v = 0.f;
      if (j < vpt && live) {
        const float dh = rstd * (dx[j] - s1 - xh[j] * s2);
        v = hv[j] > 0.f ? dh : 0.f;
        if (blockIdx.x == 0) dy[(size_t)row * H + c] = v;
      }
      As[warp][c] = v;  // columns >= H are zero
    }
  }
  __syncthreads();
  float acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
  const int npw = (((H + 7) >> 3) + 3) & ~3;
  const int nlo = warp * npw;
  if (kok && nlo < H) {
    const float* bcol = Bm + (size_t)nlo * K + k;
#pragma unroll 8
    for (int nn = 0; nn < npw; nn += 4) {
      float bv[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) bv[u] = nlo + nn + u < H ? bcol[(size_t)(nn + u) * K] : 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float4 av = *reinterpret_cast<const float4*>(&As[i][nlo + nn]);
        acc[i] = fmaf(av.x, bv[0], acc[i]);
        acc[i] = fmaf(av.y, bv[1], acc[i]);
        acc[i] = fmaf(av.z, bv[2], acc[i]);
        acc[i] = fmaf(av.w, bv[3], acc[i]);
      }
    }
  }
  __syncthreads();
  float* part = &As[0][0];
#pragma unroll
  for (int i = 0; i < 8; ++i) part[(warp * 8 + i) * 32 + lane] = acc[i];
  __syncthreads();
  {
    const int i = warp;
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += part[(w * 8 + i) * 32 + lane];
    if (kok && m0 + i < M) C[(size_t)(m0 + i) * K + k] = t;
  }
}

// C[N,K] = A[M,N]^T . B[M,K] (dW = dY^T x activations). CTA = 8 output rows n x 32 output columns k; lane = k (B rows
// read coalesced), the 8 warps split the contraction index m; partial tiles added through shared memory.
__global__ void __launch_bounds__(256) gemm_tn_kernel(const float* __restrict__ A, const float* __restrict__ Bm,
                                                      float* __restrict__ C, int M, int N, int K) {
  PPO_PDL_SYNC();  // programmatic dependent launch: see launch_k
  __shared__ __align__(16) float As[256][8];  // reused for the partial tiles [8 warps][8 n][32 k]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int k = blockIdx.x * 32 + lane, n0 = blockIdx.y * 8;
  const bool kok = k < K;
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  for (int mb = 0; mb < M; mb += 256) {
    const int mc = min(256, M - mb);
    const int mpw = (mc + 7) >> 3;  // m per warp
    __syncthreads();
    for (int t = threadIdx.x; t < 256 * 8; t += 256) {
      const int m = t >> 3, j = t & 7;
      As[m][j] = (m < mc && n0 + j < N) ? A[(size_t)(mb + m) * N + n0 + j] : 0.f;
    }
    __syncthreads();
    const int mlo = warp * mpw, mhi = min(mc, mlo + mpw);
    if (kok) {
      const float* bcol = Bm + (size_t)mb * K + k;
#pragma unroll 8
      for (int m = mlo; m < mhi; ++m) {
        const float bv = bcol[(size_t)m * K];
        const float4 a0 = *reinterpret_cast<const float4*>(&As[m][0]);
        const float4 a1 = *reinterpret_cast<const float4*>(&As[m][4]);
        acc[0] = fmaf(a0.x, bv, acc[0]);
        acc[1] = fmaf(a0.y, bv, acc[1]);
        acc[2] = fmaf(a0.z, bv, acc[2]);
        acc[3] = fmaf(a0.w, bv, acc[3]);
        acc[4] = fmaf(a1.x, bv, acc[4]);
        acc[5] = fmaf(a1.y, bv, acc[5]);
        acc[6] = fmaf(a1.z, bv, acc[6]);
        acc[7] = fmaf(a1.w, bv, acc[7]);
      }
    }
  }
  __syncthreads();
  float* part = &As[0][0];
#pragma unroll
  for (int j = 0; j < 8; ++j) part[(warp * 8 + j) * 32 + lane] = acc[j];
  __syncthreads();
  {
    const int j = warp;
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += part[(w * 8 + j) * 32 + lane];
    if (kok && n0 + j < N) C[(size_t)(n0 + j) * K + k] = t;
  }
}

// ---------------------------------------------------------------- heads
// Policy head: warp per row, 8 rows per CTA, lane owns action dims lane and lane+32. Column sums (d bias,
// d log_std) and the loss scalars go through per-CTA partials and the last CTA (ticket), in CTA order.
constexpr int kHeadRows = 8, kHeadPart = 130;  // partial record: 64 d-bias, 64 d-log_std, surrogate sum, ratio sum
// FUSED = true is instantiated only by tools/experiments (measured slower, not in this library).
// FUSED: the last Linear and its dA are computed here as well - z = a3 . W4^T with lanes along the H3 inputs and
// a transposing butterfly that leaves z[lane] and z[lane+32] on the lane, dA3 = dz . W4 with dz broadcast by
// shuffles - so that the chain Linear -> head -> dA GEMM is one launch.
template <bool FUSED>
__global__ void __launch_bounds__(256) policy_head_kernel(
    const float* __restrict__ y4, const float* __restrict__ b4, const float* __restrict__ log_std,
    const float* __restrict__ actions, const float* __restrict__ old_logp, const float* __restrict__ adv,
    float clip_eps, float entropy_coef, float* __restrict__ dz, float* __restrict__ db4,
    float* __restrict__ dlog_std, float* __restrict__ logp_out, ppo_ctrl* ctrl, float* scratch, uint32_t* ticket,
    int B, int A, const float* __restrict__ a3, const float* __restrict__ W4, float* __restrict__ dA3, int H3) {
  PPO_PDL_SYNC();  // programmatic dependent launch: see launch_k
  __shared__ float col_b[kHeadRows][64], col_s[kHeadRows][64], red[kHeadRows][2];
  __shared__ bool is_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, row = blockIdx.x * kHeadRows + warp;
  float ls[2], inv_var[2], bb[2];
  bool has[2];
  float ent_row = 0.f;
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int c = lane + 32 * j;
    has[j] = c < A;
    ls[j] = has[j] ? fminf(fmaxf(log_std[c], -20.f), 2.f) : 0.f;
    inv_var[j] = expf(-2.f * ls[j]);
    bb[j] = has[j] ? b4[c] : 0.f;
    if (has[j]) ent_row += 0.5f + kHalfLog2Pi + ls[j];
  }
  const float entropy = warp_sum(ent_row) / A;  // mean over [B, A] of a row-constant
  float acc_b[2] = {0.f, 0.f}, acc_s[2] = {0.f, 0.f}, acc_surr = 0.f, acc_ratio = 0.f;
  if (row < B) {
    float zf[2] = {0.f, 0.f};
    if (FUSED) {
      float av[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) av[i] = lane + 32 * i < H3 ? a3[(size_t)row * H3 + lane + 32 * i] : 0.f;
      float acc[64];  // slot s holds action dim (s >> 1) + 32 * (s & 1): the butterfly ends with slots 2l, 2l+1 on lane l
#pragma unroll
      for (int sl = 0; sl < 64; ++sl) {
        const int d = (sl >> 1) + 32 * (sl & 1);
        float t = 0.f;
        if (d < A) {
#pragma unroll
          for (int i = 0; i < 8; ++i)
            if (lane + 32 * i < H3) t = fmaf(av[i], W4[(size_t)d * H3 + lane + 32 * i], t);
        }
        acc[sl] = t;
      }
#pragma unroll
      for (int off = 16, n = 32; off >= 1; off >>= 1, n >>= 1) {
        const bool up = (lane & off) != 0;
#pragma unroll
        for (int t = 0; t < 32; ++t)
          if (t < n) {
            const float send = up ? acc[t] : acc[t + n], keep = up ? acc[t + n] : acc[t];
            acc[t] = keep + __shfl_xor_sync(0xffffffffu, send, off);
          }
      }
      zf[0] = acc[0];
      zf[1] = acc[1];
    }
    float mu[2], diff[2], gz[2] = {0.f, 0.f}, lp = 0.f;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      mu[j] = diff[j] = 0.f;
      if (has[j]) {
        const size_t i = (size_t)row * A + lane + 32 * j;
        mu[j] = tanhf((FUSED ? zf[j] : y4[i]) + bb[j]);
        diff[j] = actions[i] - mu[j];
        lp += -0.5f * diff[j] * diff[j] * inv_var[j] - ls[j] - kHalfLog2Pi;
      }
    }
    lp = warp_sum(lp);
    const float a = adv[row], ratio = expf(lp - old_logp[row]);
    const float surr1 = ratio * a, surr2 = fminf(fmaxf(ratio, 1.f - clip_eps), 1.f + clip_eps) * a;
    const bool inside = ratio >= 1.f - clip_eps && ratio <= 1.f + clip_eps;
    // torch.min sends the gradient to the smaller argument and splits it on ties
    const float w1 = surr1 < surr2 ? 1.f : surr1 == surr2 ? 0.5f : 0.f;
    const float w2 = surr2 < surr1 ? 1.f : surr1 == surr2 ? 0.5f : 0.f;
    const float dlogp = -(w1 + (inside ? w2 : 0.f)) * a * ratio / B;
    acc_surr = fminf(surr1, surr2);
    acc_ratio = ratio;
    if (lane == 0 && logp_out != nullptr) logp_out[row] = lp;
#pragma unroll
    for (int j = 0; j < 2; ++j)
      if (has[j]) {
        const float dmu = dlogp * diff[j] * inv_var[j];
        const float g = dmu * (1.f - mu[j] * mu[j]);
        dz[(size_t)row * A + lane + 32 * j] = g;
        gz[j] = g;
        acc_b[j] = g;
        acc_s[j] = dlogp * (diff[j] * diff[j] * inv_var[j] - 1.f);
      }
    if (FUSED) {
      float o[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) o[i] = 0.f;
#pragma unroll 5
      for (int d = 0; d < A; ++d) {
        const float g = __shfl_sync(0xffffffffu, d < 32 ? gz[0] : gz[1], d & 31);
#pragma unroll
        for (int i = 0; i < 8; ++i)
          if (lane + 32 * i < H3) o[i] = fmaf(g, W4[(size_t)d * H3 + lane + 32 * i], o[i]);
      }
#pragma unroll
      for (int i = 0; i < 8; ++i)
        if (lane + 32 * i < H3) dA3[(size_t)row * H3 + lane + 32 * i] = o[i];
    }
  }
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    col_b[warp][lane + 32 * j] = acc_b[j];
    col_s[warp][lane + 32 * j] = acc_s[j];
  }
  if (lane == 0) {
    red[warp][0] = acc_surr;
    red[warp][1] = acc_ratio;
  }
  __syncthreads();
  float* part = scratch + (size_t)blockIdx.x * kHeadPart;
  if (threadIdx.x < 64) {
    float sb = 0.f, ss = 0.f;
#pragma unroll
    for (int w = 0; w < kHeadRows; ++w) {
      sb += col_b[w][threadIdx.x];
      ss += col_s[w][threadIdx.x];
    }
    part[threadIdx.x] = sb;
    part[64 + threadIdx.x] = ss;
  } else if (threadIdx.x < 66) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < kHeadRows; ++w) t += red[w][threadIdx.x - 64];
    part[64 + threadIdx.x] = t;  // 128: surrogate sum, 129: ratio sum
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  if (threadIdx.x < kHeadPart) {
    float t = 0.f;
    for (int b = 0; b < (int)gridDim.x; ++b) t += __ldcg(scratch + (size_t)b * kHeadPart + threadIdx.x);
    const int c = threadIdx.x;
    if (c < 64) {
      if (c < A) db4[c] = t;
    } else if (c < 128) {
      if (c - 64 < A) {
        const float raw = log_std[c - 64];
        dlog_std[c - 64] = (raw >= -20.f && raw <= 2.f) ? t - entropy_coef / A : 0.f;
      }
    } else if (c == 128) {
      ctrl->actor_loss = -t / B - entropy_coef * entropy;
      ctrl->entropy = entropy;
    } else {
      ctrl->mean_ratio = t / B;
    }
  }
  if (threadIdx.x == 0) *ticket = 0;
}

__global__ void __launch_bounds__(1024) value_head_kernel(const float* __restrict__ y4, const float* __restrict__ b4,
                                                          const float* __restrict__ returns, float* __restrict__ dv,
                                                          float* __restrict__ db4, ppo_ctrl* ctrl, int B) {
  PPO_PDL_SYNC();  // programmatic dependent launch: see launch_k
  __shared__ float red[32][3];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float sq = 0.f, sv = 0.f, sd = 0.f;
  const float b = b4[0];
  for (int i = threadIdx.x; i < B; i += blockDim.x) {
    const float v = y4[i] + b, e = v - returns[i], g = 2.f * e / B;
    dv[i] = g;
    sq += e * e;
    sv += v;
    sd += g;
  }
  sq = warp_sum(sq);
  sv = warp_sum(sv);
  sd = warp_sum(sd);
  if (lane == 0) red[warp][0] = sq, red[warp][1] = sv, red[warp][2] = sd;
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = 0.f, c = 0.f, d = 0.f;
    for (int w = 0; w < 32; ++w) a += red[w][0], c += red[w][1], d += red[w][2];
    ctrl->critic_loss = a / B;
    ctrl->mean_value = c / B;
    db4[0] = d;
  }
}

// Value head with the critic's last Linear (H3 -> 1) and its dA folded in: warp per row.
__global__ void __launch_bounds__(1024) value_head_fused_kernel(
    const float* __restrict__ a3, const float* __restrict__ w4, const float* __restrict__ b4,
    const float* __restrict__ returns, float* __restrict__ dv, float* __restrict__ dA3, float* __restrict__ db4,
    ppo_ctrl* ctrl, int B, int H3) {
  PPO_PDL_SYNC();  // programmatic dependent launch: see launch_k
  __shared__ float red[32][3];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float wv[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) wv[i] = lane + 32 * i < H3 ? w4[lane + 32 * i] : 0.f;
  const float b = b4[0];
  float sq = 0.f, sv = 0.f, sd = 0.f;
  for (int row = warp; row < B; row += 32) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i)
      if (lane + 32 * i < H3) t = fmaf(a3[(size_t)row * H3 + lane + 32 * i], wv[i], t);
    const float v = warp_sum(t) + b, e = v - returns[row], g = 2.f * e / B;
#pragma unroll
    for (int i = 0; i < 8; ++i)
      if (lane + 32 * i < H3) dA3[(size_t)row * H3 + lane + 32 * i] = g * wv[i];
    if (lane == 0) {
      dv[row] = g;
      sq += e * e;
      sv += v;
      sd += g;
    }
  }
  if (lane == 0) red[warp][0] = sq, red[warp][1] = sv, red[warp][2] = sd;
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = 0.f, c = 0.f, d = 0.f;
    for (int w = 0; w < 32; ++w) a += red[w][0], c += red[w][1], d += red[w][2];
    ctrl->critic_loss = a / B;
    ctrl->mean_value = c / B;
    db4[0] = d;
  }
}

__global__ void add_bias_kernel(const float* __restrict__ y, const float* __restrict__ b, float* __restrict__ out,
                                int n) {
  PPO_PDL_SYNC();  // programmatic dependent launch: see launch_k
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = y[i] + b[0];
}

// select_actions: warp per env.
__global__ void __launch_bounds__(256) sample_actions_kernel(
    const float* __restrict__ y4, const float* __restrict__ b4, const float* __restrict__ log_std,
    const float* __restrict__ noise_in, float* __restrict__ actions, float* __restrict__ logp,
    const ppo_ctrl* ctrl, int N, int A) {
  PPO_PDL_SYNC();  // programmatic dependent launch: see launch_k
  const int lane = threadIdx.x & 31, row = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= N) return;
  float lp = 0.f;
  uint4 r = make_uint4(0, 0, 0, 0);
  if (noise_in == nullptr) r = rng4(ctrl, 0x5a3711u, (uint32_t)(row * 32 + lane));
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int c = lane + 32 * j;
    if (c < A) {
      const size_t i = (size_t)row * A + c;
      float eps;
      if (noise_in != nullptr) {
        eps = noise_in[i];
      } else {  // Box-Muller, one pair of uniforms per normal
        const float u1 = u01(j == 0 ? r.x : r.z), u2 = u01(j == 0 ? r.y : r.w);
        eps = sqrtf(-2.f * logf(u1)) * cospif(2.f * u2);
      }
      const float ls = fminf(fmaxf(log_std[c], -20.f), 2.f);
      actions[i] = tanhf(y4[i] + b4[c]) + expf(ls) * eps;
      lp += -0.5f * eps * eps - ls - kHalfLog2Pi;  // (a - mu)/sigma is eps by construction
    }
  }
  lp = warp_sum(lp);
  if (lane == 0) logp[row] = lp;
}

// ---------------------------------------------------------------- batch preparation (one CTA)
__device__ double block_sum_d(double v, double* sh) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  double t = 0.0;
  for (int w = 0; w < 32; ++w) t += sh[w];
  return t;
}

__global__ void __launch_bounds__(1024) prepare_batch_kernel(
    const float* __restrict__ rewards, const float* __restrict__ dones, const float* __restrict__ values,
    const float* __restrict__ next_values, float gamma, float lam, float* __restrict__ rewards_n,
    float* __restrict__ returns, float* __restrict__ adv_raw, float* __restrict__ adv, ppo_ctrl* ctrl, int N) {
  PPO_PDL_SYNC();  // programmatic dependent launch: see launch_k
  __shared__ double shd[32];
  __shared__ float sA[1024], sB[1024];
  const int tid = threadIdx.x;
  // RunningMeanStd (ppo_v2.py:113-131), float64 as numpy: batch mean, batch var = mean((x - mean)^2)
  double s = 0.0;
  for (int i = tid; i < N; i += 1024) s += (double)rewards[i];
  const double bmean = block_sum_d(s, shd) / N;
  s = 0.0;
  for (int i = tid; i < N; i += 1024) {
    const double d = (double)rewards[i] - bmean;
    s += d * d;
  }
  const double bvar = block_sum_d(s, shd) / N;
  const double mean0 = ctrl->rms_mean, var0 = ctrl->rms_var, cnt0 = ctrl->rms_count;
  const double delta = bmean - mean0, tot = cnt0 + (double)N;
  const double mean1 = mean0 + delta * (double)N / tot;
  const double var1 = (var0 * cnt0 + bvar * (double)N + delta * delta * cnt0 * (double)N / tot) / tot;
  const double inv_std = 1.0 / sqrt(var1 + 1e-8);
  __syncthreads();
  if (tid == 0) {
    ctrl->rms_mean = mean1;
    ctrl->rms_var = var1;
    ctrl->rms_count = tot;
  }
  s = 0.0;
  for (int i = tid; i < N; i += 1024) {
    const float rn = (float)(((double)rewards[i] - mean1) * inv_std);
    rewards_n[i] = rn;
    returns[i] = rn + gamma * (next_values[i] * (1.f - dones[i]));
    s += (double)rn;
  }
  const double mean_rn = block_sum_d(s, shd) / N;
  if (tid == 0) ctrl->mean_reward_n = (float)mean_rn;
  __syncthreads();  // rewards_n visible to the whole CTA
  // the index-wise recursion g_t = delta_t + c_t g_{t+1} (ppo_v2.py:236-246) as a right-to-left scan of the
  // affine maps x -> b + a x: chunk summaries, a 1024-wide suffix scan, then a re-walk of each chunk
  const int chunk = (N + 1023) / 1024, lo = tid * chunk, hi = min(lo + chunk, N);
  float a = 1.f, b = 0.f;
  for (int t = hi - 1; t >= lo; --t) {
    const float nd = 1.f - dones[t], nv = t == N - 1 ? next_values[t] : values[t + 1];
    const float dlt = rewards_n[t] + gamma * nv * nd - values[t], c = gamma * lam * nd;
    b = dlt + c * b;
    a = c * a;
  }
  sA[tid] = a;
  sB[tid] = b;
  __syncthreads();
  for (int off = 1; off < 1024; off <<= 1) {
    float a2 = 1.f, b2 = 0.f;
    const bool on = tid + off < 1024;
    if (on) a2 = sA[tid + off], b2 = sB[tid + off];
    __syncthreads();
    if (on) {  // this chunk's map applied after everything to its right
      sB[tid] = sB[tid] + sA[tid] * b2;
      sA[tid] = sA[tid] * a2;
    }
    __syncthreads();
  }
  float g = tid + 1 < 1024 ? sB[tid + 1] : 0.f;  // gae entering this chunk from the right (0 after the end)
  double sum = 0.0;
  for (int t = hi - 1; t >= lo; --t) {
    const float nd = 1.f - dones[t], nv = t == N - 1 ? next_values[t] : values[t + 1];
    const float dlt = rewards_n[t] + gamma * nv * nd - values[t];
    g = dlt + gamma * lam * nd * g;
    adv[t] = g;
    if (adv_raw != nullptr) adv_raw[t] = g;
    sum += (double)g;
  }
  const double amean = block_sum_d(sum, shd) / N;
  sum = 0.0;
  for (int t = lo; t < hi; ++t) {
    const double d = (double)adv[t] - amean;
    sum += d * d;
  }
  const double astd = sqrt(block_sum_d(sum, shd) / (double)(N - 1));  // torch .std(): unbiased
  const float fm = (float)amean, finv = (float)(1.0 / (astd + 1e-8));
  for (int t = lo; t < hi; ++t) adv[t] = (adv[t] - fm) * finv;
}

// ---------------------------------------------------------------- minibatch selection
__device__ __forceinline__ uint32_t mix32(uint32_t x) {
  x ^= x >> 16;
  x *= 0x85EBCA6Bu;
  x ^= x >> 13;
  x *= 0xC2B2AE35u;
  x ^= x >> 16;
  return x;
}
// Keyed bijection of [0, n): balanced Feistel network on 2*hb bits with cycle walking.
__device__ uint32_t feistel_perm(uint32_t i, uint32_t n, uint32_t key) {
  int bits = 32 - __clz(max(n - 1u, 1u));
  const int hb = (bits + 1) >> 1;
  const uint32_t mask = (1u << hb) - 1u;
  uint32_t x = i;
  do {
    uint32_t l = x >> hb, r = x & mask;
#pragma unroll
    for (int round = 0; round < 6; ++round) {
      const uint32_t f = mix32(r ^ (key + 0x9E3779B9u * (round + 1))) & mask;
      const uint32_t nl = r;
      r = l ^ f;
      l = nl;
    }
    x = (l << hb) | r;
  } while (x >= n);
  return x;
}
__device__ __forceinline__ uint32_t perm_key(const ppo_ctrl* ctrl, uint32_t salt) {
  return mix32((uint32_t)ctrl->seed ^ mix32((uint32_t)ctrl->epoch + 0x632BE5ABu) ^ mix32(salt + 77u));
}

__global__ void __launch_bounds__(128) gather_minibatch_kernel(
    const float* __restrict__ states, const float* __restrict__ actions, const float* __restrict__ old_logp,
    const float* __restrict__ adv, const float* __restrict__ returns, const int32_t* __restrict__ perm,
    int perm_epochs, const ppo_ctrl* ctrl, uint32_t salt, int N, int S, int A, float* __restrict__ b_states,
    float* __restrict__ b_actions, float* __restrict__ b_logp, float* __restrict__ b_adv,
    float* __restrict__ b_returns) {
  PPO_PDL_SYNC();  // programmatic dependent launch: see launch_k
  const int r = blockIdx.x, pos = ctrl->cursor + r;
  int src;
  if (perm != nullptr)
    src = perm[(size_t)(ctrl->epoch % perm_epochs) * N + pos];
  else
    src = (int)feistel_perm((uint32_t)pos, (uint32_t)N, perm_key(ctrl, salt));
  for (int c = threadIdx.x; c < S; c += blockDim.x) b_states[(size_t)r * S + c] = states[(size_t)src * S + c];
  for (int c = threadIdx.x; c < A; c += blockDim.x) b_actions[(size_t)r * A + c] = actions[(size_t)src * A + c];
  if (threadIdx.x == 0) {
    b_logp[r] = old_logp[src];
    b_adv[r] = adv[src];
    b_returns[r] = returns[src];
  }
}
__global__ void permutation_kernel(const ppo_ctrl* ctrl, uint32_t salt, int N, int32_t* out) {
  PPO_PDL_SYNC();  // programmatic dependent launch: see launch_k
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < N) out[i] = (int32_t)feistel_perm((uint32_t)i, (uint32_t)N, perm_key(ctrl, salt));
}

// ---------------------------------------------------------------- clip_grad_norm_ + Adam
__global__ void __launch_bounds__(256) grad_sumsq_kernel(const float* __restrict__ grad, int n, float* scratch,
                                                         uint32_t* ticket, ppo_ctrl* ctrl, int which) {
  PPO_PDL_SYNC();  // programmatic dependent launch: see launch_k
  __shared__ float red[8];
  __shared__ bool is_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float s = 0.f;
  const int n4 = n >> 2;  // tensors are 16-byte aligned and padded with zeros to multiples of 4
  const float4* g4 = reinterpret_cast<const float4*>(grad);
  for (int i = blockIdx.x * 256 + threadIdx.x; i < n4; i += gridDim.x * 256) {
    const float4 g = g4[i];
    s += g.x * g.x + g.y * g.y + g.z * g.z + g.w * g.w;
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) { const float g = grad[(n4 << 2) + threadIdx.x]; s += g * g; }
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
  // fixed order: lane l adds partials l, l+32, ...; then a fixed butterfly
  float t = 0.f;
  for (int b = lane; b < (int)gridDim.x; b += 32) t += __ldcg(scratch + b);
  t = warp_sum(t);
  if (lane == 0) {
    ctrl->norm2[which] = t;
    ctrl->adam_t[which] += 1;
    *ticket = 0;
  }
}

__global__ void __launch_bounds__(256) clip_adam_kernel(float4* __restrict__ theta, const float4* __restrict__ grad,
                                                        float4* __restrict__ m, float4* __restrict__ v, int n4,
                                                        const ppo_ctrl* ctrl, int which, float max_norm,
                                                        float grad_scale, float b1, float b2, float eps) {
  PPO_PDL_SYNC();  // programmatic dependent launch: see launch_k
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n4) return;
  const float total = sqrtf(ctrl->norm2[which]) * grad_scale;
  const float coef = fminf(max_norm / (total + 1e-6f), 1.f) * grad_scale;
  const float t = (float)ctrl->adam_t[which];
  const float bc1 = 1.f - powf(b1, t), inv_sqrt_bc2 = 1.f / sqrtf(1.f - powf(b2, t));
  const float step = ctrl->lr[which] / bc1;
  const float4 g4 = grad[i];
  float4 m4 = m[i], v4 = v[i], p4 = theta[i];
  const float g[4] = {g4.x * coef, g4.y * coef, g4.z * coef, g4.w * coef};
  float* mm = &m4.x;
  float* vv = &v4.x;
  float* pp = &p4.x;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    mm[k] = b1 * mm[k] + (1.f - b1) * g[k];
    vv[k] = b2 * vv[k] + (1.f - b2) * g[k] * g[k];
    pp[k] -= step * mm[k] / (sqrtf(vv[k]) * inv_sqrt_bc2 + eps);
  }
  m[i] = m4;
  v[i] = v4;
  theta[i] = p4;
}

// ---------------------------------------------------------------- control block
__global__ void ctrl_init_kernel(ppo_ctrl* c, uint64_t seed, float lr_a, float lr_c) {
  ppo_ctrl z = {};
  z.seed = seed;
  z.lr[0] = lr_a;
  z.lr[1] = lr_c;
  z.best[0] = -INFINITY;  // actor scheduler: mode "max"
  z.best[1] = INFINITY;   // critic scheduler: mode "min"
  z.rms_mean = 0.0;
  z.rms_var = 1.0;
  z.rms_count = 1e-4;
  *c = z;
}
__global__ void advance_kernel(ppo_ctrl* c, int N, int B) {
  PPO_PDL_SYNC();  // programmatic dependent launch: see launch_k
  c->rng_counter += 1;
  if (B > 0) {
    c->cursor += B;
    if (c->cursor >= N) {
      c->cursor = 0;
      c->epoch += 1;
    }
  }
}
__global__ void begin_update_kernel(ppo_ctrl* c) {
  PPO_PDL_SYNC();  // programmatic dependent launch: see launch_k
  c->cursor = 0;
}

__global__ void plateau_kernel(ppo_ctrl* c, float factor, int patience) {
  PPO_PDL_SYNC();  // programmatic dependent launch: see launch_k
  for (int which = 0; which < 2; ++which) {
    const float metric = which == 0 ? c->mean_reward_n : c->critic_loss;
    const float best = c->best[which];
    // threshold 1e-4, threshold_mode "rel"
    const bool better = which == 0 ? metric > best * (1.f + 1e-4f) : metric < best * (1.f - 1e-4f);
    if (better) {
      c->best[which] = metric;
      c->num_bad[which] = 0;
    } else {
      c->num_bad[which] += 1;
    }
    if (c->num_bad[which] > patience) {
      const float lr = c->lr[which], nlr = lr * factor;
      if (lr - nlr > 1e-8f) c->lr[which] = nlr;
      c->num_bad[which] = 0;
    }
  }
}

}  // namespace

extern "C" {

const char* ppo_last_error(void) { return g_err; }
const char* ppo_version(void) { return "ppo_kernels 0.1 (sm_100a)"; }
int ppo_ctrl_size(void) { return (int)sizeof(ppo_ctrl); }

int ppo_ctrl_init(ppo_ctrl* c, uint64_t seed, float lr_actor, float lr_critic, void* stream) {
  if (c == nullptr) return invalid("ctrl");
  ctrl_init_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(c, seed, lr_actor, lr_critic);
  PPO_LAUNCH_CHECK("ppo_ctrl_init");
}

int ppo_gemm_nt(const float* A, const float* B, float* C, int M, int N, int K, void* stream) {
  if (M <= 0 || N <= 0 || K <= 0) return invalid("gemm sizes");
  launch_k(gemm_nt_kernel, dim3(dim3((N + 31) / 32, (M + 3) / 4)), dim3(128), (cudaStream_t)stream, use_pdl(), A, B, C, M, N, K);
  PPO_LAUNCH_CHECK("ppo_gemm_nt");
}
int ppo_gemm_nn(const float* A, const float* B, float* C, int M, int N, int K, void* stream) {
  if (M <= 0 || N <= 0 || K <= 0) return invalid("gemm sizes");
  launch_k(gemm_nn_kernel, dim3(dim3((K + 31) / 32, (M + 7) / 8)), dim3(256), (cudaStream_t)stream, use_pdl(), A, B, C, M, N, K);
  PPO_LAUNCH_CHECK("ppo_gemm_nn");
}
int ppo_gemm_tn(const float* A, const float* B, float* C, int M, int N, int K, void* stream) {
  if (M <= 0 || N <= 0 || K <= 0) return invalid("gemm sizes");
  launch_k(gemm_tn_kernel, dim3(dim3((K + 31) / 32, (N + 7) / 8)), dim3(256), (cudaStream_t)stream, use_pdl(), A, B, C, M, N, K);
  PPO_LAUNCH_CHECK("ppo_gemm_tn");
}

int ppo_relu_ln_fwd(const float* y, const float* bias, const float* gamma, const float* beta, float* out,
                    float* h_save, float* stats, int rows, int H, float dropout_p, const uint8_t* mask_in,
                    uint8_t* mask_out, const ppo_ctrl* ctrl, uint32_t rng_stream, void* stream) {
  if (H % 32 != 0 || H > 256 || H <= 0) return invalid("H must be a multiple of 32, <= 256");
  if (rows <= 0) return invalid("rows");
  if (dropout_p > 0.f && mask_in == nullptr && ctrl == nullptr) return invalid("dropout needs ctrl or mask_in");
  launch_k(relu_ln_fwd_kernel, dim3((rows + 7) / 8), dim3(256), (cudaStream_t)stream, use_pdl(), y, bias, gamma, beta, out, h_save, stats, rows,
                                                                      H, dropout_p, mask_in, mask_out, ctrl,
                                                                      rng_stream);
  PPO_LAUNCH_CHECK("ppo_relu_ln_fwd");
}

int ppo_relu_ln_bwd_dy(const float* dout, const float* h, const float* stats, const float* gamma, const uint8_t* mask,
                       float dropout_p, float* dy, int rows, int H, void* stream) {
  if (H % 32 != 0 || H > 256 || H <= 0) return invalid("H must be a multiple of 32, <= 256");
  if (rows <= 0) return invalid("rows");
  if (dropout_p > 0.f && mask == nullptr) return invalid("dropout backward needs the mask");
  launch_k(relu_ln_bwd_dy_kernel, dim3((rows + 7) / 8), dim3(256), (cudaStream_t)stream, use_pdl(), dout, h, stats, gamma, mask, dropout_p, dy,
                                                                         rows, H);
  PPO_LAUNCH_CHECK("ppo_relu_ln_bwd_dy");
}

int ppo_ln_bwd_gemm_nn(const float* dout, const float* h, const float* stats, const float* gamma, const uint8_t* mask,
                       float dropout_p, float* dy, const float* W, float* dA, int rows, int H, int K, void* stream) {
  if (H % 32 != 0 || H > 256 || H <= 0) return invalid("H must be a multiple of 32, <= 256");
  if (rows <= 0 || K <= 0) return invalid("rows / K");
  if (dropout_p > 0.f && mask == nullptr) return invalid("dropout backward needs the mask");
  launch_k(ln_bwd_gemm_nn_kernel, dim3(dim3((K + 31) / 32, (rows + 7) / 8)), dim3(256), (cudaStream_t)stream, use_pdl(), 
      dout, h, stats, gamma, mask, dropout_p, dy, W, dA, rows, H, K);
  PPO_LAUNCH_CHECK("ppo_ln_bwd_gemm_nn");
}

int ppo_relu_ln_colsums(const float* dout, const float* h, const float* stats, const uint8_t* mask, float dropout_p,
                        const float* dy, float* dbias, float* dgamma, float* dbeta, int rows, int H, void* stream) {
  if (H % 32 != 0 || H > 256 || H <= 0) return invalid("H must be a multiple of 32, <= 256");
  if (rows <= 0) return invalid("rows");
  if (dropout_p > 0.f && mask == nullptr) return invalid("dropout backward needs the mask");
  launch_k(relu_ln_colsums_kernel, dim3(H / 32), dim3(256), (cudaStream_t)stream, use_pdl(), dout, h, stats, mask, dropout_p, dy, dbias, dgamma,
                                                                  dbeta, rows, H);
  PPO_LAUNCH_CHECK("ppo_relu_ln_colsums");
}

int ppo_relu_ln_bwd(const float* dout, const float* h, const float* stats, const float* gamma, const uint8_t* mask,
                    float dropout_p, float* dy, float* dbias, float* dgamma, float* dbeta, int rows, int H,
                    void* stream) {
  int rc = ppo_relu_ln_bwd_dy(dout, h, stats, gamma, mask, dropout_p, dy, rows, H, stream);
  if (rc != 0) return rc;
  return ppo_relu_ln_colsums(dout, h, stats, mask, dropout_p, dy, dbias, dgamma, dbeta, rows, H, stream);
}

int ppo_policy_head(const float* y4, const float* b4, const float* log_std, const float* actions,
                    const float* old_logp, const float* adv, float clip_eps, float entropy_coef, float* dz,
                    float* db4, float* dlog_std, float* logp_out, ppo_ctrl* ctrl, float* scratch, uint32_t* ticket,
                    int B, int A, void* stream) {
  if (A <= 0 || A > 64) return invalid("action dim must be in 1..64");
  if (B <= 0) return invalid("B");
  if (scratch == nullptr || ticket == nullptr) return invalid("scratch / ticket");
  launch_k(policy_head_kernel<false>, dim3((B + kHeadRows - 1) / kHeadRows), dim3(256), (cudaStream_t)stream, use_pdl(), 
      y4, b4, log_std, actions, old_logp, adv, clip_eps, entropy_coef, dz, db4, dlog_std, logp_out, ctrl, scratch,
      ticket, B, A, nullptr, nullptr, nullptr, 0);
  PPO_LAUNCH_CHECK("ppo_policy_head");
}


int ppo_value_head(const float* y4, const float* b4, const float* returns, float* dv, float* db4, ppo_ctrl* ctrl,
                   int B, void* stream) {
  if (B <= 0) return invalid("B");
  launch_k(value_head_kernel, dim3(1), dim3(1024), (cudaStream_t)stream, use_pdl(), y4, b4, returns, dv, db4, ctrl, B);
  PPO_LAUNCH_CHECK("ppo_value_head");
}

int ppo_value_head_fused(const float* a3, const float* w4, const float* b4, const float* returns, float* dv,
                         float* dA3, float* db4, ppo_ctrl* ctrl, int B, int H3, void* stream) {
  if (B <= 0 || H3 <= 0 || H3 > 256) return invalid("B / H3 (<= 256)");
  launch_k(value_head_fused_kernel, dim3(1), dim3(1024), (cudaStream_t)stream, use_pdl(), a3, w4, b4, returns, dv, dA3, db4, ctrl, B, H3);
  PPO_LAUNCH_CHECK("ppo_value_head_fused");
}

/* measurement probe (tools/microbench_chain.py): v[i] += 1 launched plainly (mode 0), with programmatic dependent
 * launch (1), or with it and an early trigger (2) */

int ppo_add_bias(const float* y4, const float* b4, float* out, int n, void* stream) {
  if (n <= 0) return invalid("n");
  launch_k(add_bias_kernel, dim3((n + 255) / 256), dim3(256), (cudaStream_t)stream, use_pdl(), y4, b4, out, n);
  PPO_LAUNCH_CHECK("ppo_add_bias");
}

int ppo_sample_actions(const float* y4, const float* b4, const float* log_std, const float* noise_in, float* actions,
                       float* logp, const ppo_ctrl* ctrl, int N, int A, void* stream) {
  if (A <= 0 || A > 64) return invalid("action dim must be in 1..64");
  if (N <= 0) return invalid("N");
  launch_k(sample_actions_kernel, dim3((N + 7) / 8), dim3(256), (cudaStream_t)stream, use_pdl(), y4, b4, log_std, noise_in, actions, logp, ctrl,
                                                                      N, A);
  PPO_LAUNCH_CHECK("ppo_sample_actions");
}

int ppo_prepare_batch(const float* rewards, const float* dones, const float* values, const float* next_values,
                      float gamma, float lam, float* rewards_n, float* returns, float* adv_raw, float* adv,
                      ppo_ctrl* ctrl, int N, void* stream) {
  if (N < 2) return invalid("N must be >= 2 (unbiased std)");
  launch_k(prepare_batch_kernel, dim3(1), dim3(1024), (cudaStream_t)stream, use_pdl(), rewards, dones, values, next_values, gamma, lam,
                                                            rewards_n, returns, adv_raw, adv, ctrl, N);
  PPO_LAUNCH_CHECK("ppo_prepare_batch");
}

int ppo_gather_minibatch(const float* states, const float* actions, const float* old_logp, const float* adv,
                         const float* returns, const int32_t* perm, int perm_epochs, const ppo_ctrl* ctrl,
                         uint32_t perm_salt, int N, int S, int A, int B, float* b_states, float* b_actions,
                         float* b_logp, float* b_adv, float* b_returns, void* stream) {
  if (B <= 0 || N <= 0) return invalid("sizes");
  if (perm != nullptr && perm_epochs <= 0) return invalid("perm_epochs");
  launch_k(gather_minibatch_kernel, dim3(B), dim3(128), (cudaStream_t)stream, use_pdl(), states, actions, old_logp, adv, returns, perm,
                                                              perm_epochs, ctrl, perm_salt, N, S, A, b_states,
                                                              b_actions, b_logp, b_adv, b_returns);
  PPO_LAUNCH_CHECK("ppo_gather_minibatch");
}

int ppo_permutation(const ppo_ctrl* ctrl, uint32_t perm_salt, int N, int32_t* out, void* stream) {
  if (N <= 0) return invalid("N");
  launch_k(permutation_kernel, dim3((N + 255) / 256), dim3(256), (cudaStream_t)stream, use_pdl(), ctrl, perm_salt, N, out);
  PPO_LAUNCH_CHECK("ppo_permutation");
}

int ppo_grad_sumsq(const float* grad, int n, float* scratch, uint32_t* ticket, ppo_ctrl* ctrl, int which,
                   void* stream) {
  if (n <= 0 || (which != 0 && which != 1)) return invalid("n / which");
  if ((reinterpret_cast<uintptr_t>(grad) & 15) != 0) return invalid("grad must be 16-byte aligned");
  int blocks = (n / 4 + 2 * 256 - 1) / (2 * 256);  // two float4 per thread
  if (blocks > 592) blocks = 592;                  // 4 CTAs per SM
  if (blocks < 1) blocks = 1;
  launch_k(grad_sumsq_kernel, dim3(blocks), dim3(256), (cudaStream_t)stream, use_pdl(), grad, n, scratch, ticket, ctrl, which);
  PPO_LAUNCH_CHECK("ppo_grad_sumsq");
}


int ppo_clip_adam(float* theta, const float* grad, float* m, float* v, int n, const ppo_ctrl* ctrl, int which,
                  float max_norm, float grad_scale, float beta1, float beta2, float eps, void* stream) {
  if (n <= 0 || (n & 3) != 0 || (which != 0 && which != 1)) return invalid("n (a multiple of 4) / which");
  if (((reinterpret_cast<uintptr_t>(theta) | reinterpret_cast<uintptr_t>(grad) | reinterpret_cast<uintptr_t>(m) |
        reinterpret_cast<uintptr_t>(v)) & 15) != 0)
    return invalid("theta / grad / m / v must be 16-byte aligned");
  launch_k(clip_adam_kernel, dim3((n / 4 + 255) / 256), dim3(256), (cudaStream_t)stream, use_pdl(), 
      reinterpret_cast<float4*>(theta), reinterpret_cast<const float4*>(grad), reinterpret_cast<float4*>(m),
      reinterpret_cast<float4*>(v), n / 4, ctrl, which, max_norm, grad_scale, beta1, beta2, eps);
  PPO_LAUNCH_CHECK("ppo_clip_adam");
}

int ppo_advance(ppo_ctrl* ctrl, int N, int B, void* stream) {
  launch_k(advance_kernel, dim3(1), dim3(1), (cudaStream_t)stream, use_pdl(), ctrl, N, B);
  PPO_LAUNCH_CHECK("ppo_advance");
}

int ppo_begin_update(ppo_ctrl* ctrl, void* stream) {
  launch_k(begin_update_kernel, dim3(1), dim3(1), (cudaStream_t)stream, use_pdl(), ctrl);
  PPO_LAUNCH_CHECK("ppo_begin_update");
}

int ppo_plateau_step(ppo_ctrl* ctrl, float factor, int patience, void* stream) {
  launch_k(plateau_kernel, dim3(1), dim3(1), (cudaStream_t)stream, use_pdl(), ctrl, factor, patience);
  PPO_LAUNCH_CHECK("ppo_plateau_step");
}

}  // extern "C"

// ---------------------------------------------------------------- ppo_grad_sumsq without caller scratch
// No counterpart in the reference: added for callers that cannot allocate
// (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc). The scratch of grad_sumsq_kernel
// (592 block partial sums + the ticket word) lives in this file-scope device array instead of caller buffers:
// 640 words = 2560 bytes on every GPU that loads the library, until the process exits.
// In HOST code the name g_sumsq_static is only a stand-in for the symbol, not a device address: the address is
// resolved once per device, with that device current, by cudaGetSymbolAddress below and that pointer goes to the
// kernel. Every call on a device shares the array: two calls that overlap on the GPU (different streams) race.
constexpr int kStaticPartials = 592;  // the grid cap of ppo_grad_sumsq
constexpr int kStaticWords = 640;     // partial sums [0, 592), ticket at word 592, padding
static __device__ __align__(256) float g_sumsq_static[kStaticWords];

extern "C" int ppo_grad_sumsq_static(const float* grad, int n, ppo_ctrl* ctrl, int which, void* stream) {
  static float* base[64] = {};  // device address of g_sumsq_static, per device ordinal
  if (grad == nullptr || ctrl == nullptr) return invalid("grad / ctrl");
  if (n <= 0 || (which != 0 && which != 1)) return invalid("n / which");
  if (n > PPO_GRAD_SUMSQ_STATIC_MAX_N) {
    snprintf(g_err, sizeof(g_err), "ppo_grad_sumsq_static: n = %d exceeds PPO_GRAD_SUMSQ_STATIC_MAX_N = %d "
             "(use ppo_grad_sumsq with caller scratch)", n, PPO_GRAD_SUMSQ_STATIC_MAX_N);
    return -1;
  }
  if ((reinterpret_cast<uintptr_t>(grad) & 15) != 0) return invalid("grad must be 16-byte aligned");
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail("ppo_grad_sumsq_static: cudaGetDevice", e);
  if (dev < 0 || dev >= 64) return invalid("device ordinal beyond 63");
  if (base[dev] == nullptr) {  // set-up, once per device, with that device current
    float* ptr = nullptr;
    e = cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_sumsq_static);
    if (e != cudaSuccess || ptr == nullptr) return fail("ppo_grad_sumsq_static: cudaGetSymbolAddress", e);
    base[dev] = ptr;
  }
  float* scratch = base[dev];
  uint32_t* ticket = reinterpret_cast<uint32_t*>(base[dev] + kStaticPartials);
  // the array keeps the last call's contents: the ticket must start at zero (an aborted call may have left it
  // counting); the partial sums are overwritten before they are read
  e = cudaMemsetAsync(ticket, 0, sizeof(uint32_t), (cudaStream_t)stream);
  if (e != cudaSuccess) return fail("ppo_grad_sumsq_static: cudaMemsetAsync", e);
  int blocks = (n / 4 + 2 * 256 - 1) / (2 * 256);  // the launch shape of ppo_grad_sumsq: same partials, same order
  if (blocks > kStaticPartials) blocks = kStaticPartials;
  if (blocks < 1) blocks = 1;
  launch_k(grad_sumsq_kernel, dim3(blocks), dim3(256), (cudaStream_t)stream, false, grad, n, scratch, ticket, ctrl, which);
  PPO_LAUNCH_CHECK("ppo_grad_sumsq_static");
}
