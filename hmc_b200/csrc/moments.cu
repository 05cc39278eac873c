// moments.cu -- streamed sample moments.  At 2^20 chains one iteration of q is
// 419 MB, so histories are never stored: each transition's positions are folded
// into per-chain running sums (HBM-bound, 3 reads + 2 writes per element), and
// the cross-chain statistics the diagnostics need (utils.get_ess_true,
// src/utils.py:8-27; R-hat) are reduced to 4*d doubles per rank, which is what
// travels over NCCL.
#include "common.cuh"

namespace hmcb {

template <typename T>
__global__ void __launch_bounds__(256) k_moments_accumulate(const T* __restrict__ q, int64_t sc, int64_t sd, int d,
                                                            int64_t n, T* __restrict__ sum, T* __restrict__ sumsq) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const int64_t c = i / d;
    const int k = (int)(i - c * d);
    const T x = q[c * sc + k * sd];
    sum[i] += x;
    sumsq[i] = fma(x, x, sumsq[i]);
  }
}

// What a sample loop does with one transition's outputs (hmcb_drhmc_sample): running moments as above, the
// transition's work counters added to the running totals (the reference sums state.counts afterwards,
// scripts_old/tests/test_drhmc_funnel.py:258), accepted transitions counted.  One launch per transition and
// chain range; HBM-bound, it overlaps the FP64-bound stage kernels of other ranges.
template <typename T>
__global__ void __launch_bounds__(256) k_post_transition(const T* __restrict__ q, int64_t sc, int64_t sd, int d, int64_t n_chains,
                                                         T* __restrict__ sum, T* __restrict__ sumsq,
                                                         const int32_t* __restrict__ counts_k, int32_t* __restrict__ counts_tot,
                                                         int first, const int32_t* __restrict__ code, int min_accept_code,
                                                         int max_accept_code, int32_t* __restrict__ n_accept) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (sum) {
    const int64_t n = n_chains * d;
    for (int64_t i = i0; i < n; i += stride) {
      const int64_t c = i / d;
      const int k = (int)(i - c * d);
      const T x = q[c * sc + k * sd];
      sum[i] += x;
      sumsq[i] = fma(x, x, sumsq[i]);
    }
  }
  if (counts_tot) {
    for (int64_t i = i0; i < 3 * n_chains; i += stride) counts_tot[i] = first ? counts_k[i] : counts_tot[i] + counts_k[i];
  }
  if (n_accept) {
    for (int64_t i = i0; i < n_chains; i += stride) n_accept[i] += (code[i] >= min_accept_code && code[i] <= max_accept_code) ? 1 : 0;
  }
}

int launch_post_transition(int32_t dtype, int64_t n_chains, int32_t d, const void* q, int64_t sc, int64_t sd, void* sum,
                           void* sumsq, const int32_t* counts_k, int32_t* counts_tot, int first, const int32_t* code,
                           int min_accept_code, int max_accept_code, int32_t* n_accept, cudaStream_t st) {
  if (!sum && !counts_tot && !n_accept) return 0;
  const int64_t n = n_chains * (sum ? d : 3);
  int64_t blocks = (n + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  if (dtype == HMCB_F64)
    k_post_transition<double><<<(int)blocks, 256, 0, st>>>((const double*)q, sc, sd, d, n_chains, (double*)sum, (double*)sumsq,
                                                           counts_k, counts_tot, first, code, min_accept_code, max_accept_code, n_accept);
  else
    k_post_transition<float><<<(int)blocks, 256, 0, st>>>((const float*)q, sc, sd, d, n_chains, (float*)sum, (float*)sumsq,
                                                          counts_k, counts_tot, first, code, min_accept_code, max_accept_code, n_accept);
  HMCB_CUDA(cudaPeekAtLastError());
  return 0;
}

__global__ void __launch_bounds__(256) k_scale_counts(int32_t* __restrict__ counts, int64_t n, int K) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) counts[i] *= K;
}

int launch_scale_counts(int32_t* counts, int64_t n_chains, int K, cudaStream_t st) {
  int64_t blocks = (3 * n_chains + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  k_scale_counts<<<(int)blocks, 256, 0, st>>>(counts, 3 * n_chains, K);
  HMCB_CUDA(cudaPeekAtLastError());
  return 0;
}

// out[0*d+k] = sum_c m_ck, out[1*d+k] = sum_c m_ck^2, out[2*d+k] = sum_c s2_ck (ddof=1),
// out[3*d+k] = sum_c (m_ck - mu_k)^2, with m, s2 the per-chain mean and variance over n_samples.
// Thread t of a block owns dimension k = kbase + t % dw and row offset t / dw, so a
// block reads whole rows contiguously whatever d is (dw = min(d, THREADS)).
template <typename T, int THREADS>
__global__ void __launch_bounds__(THREADS) k_moments_reduce(const T* __restrict__ sum, const T* __restrict__ sumsq,
                                                            int64_t n_chains, int d, double n_samples,
                                                            const double* __restrict__ mu, double* __restrict__ out) {
  const int dw = d < THREADS ? d : THREADS;       // dimensions handled per block row
  const int rows = THREADS / dw;                  // chains per pass
  const int r = threadIdx.x / dw;
  const int k = blockIdx.y * dw + threadIdx.x % dw;
  const bool live = r < rows && k < d;
  double a0 = 0, a1 = 0, a2 = 0, a3 = 0;
  const double inv_n = 1.0 / n_samples;
  const double inv_nm1 = n_samples > 1 ? 1.0 / (n_samples - 1.0) : 0.0;
  const double muk = (mu && live) ? mu[k] : 0.0;
  if (live) {
    for (int64_t c = (int64_t)blockIdx.x * rows + r; c < n_chains; c += (int64_t)gridDim.x * rows) {
      const double s = (double)sum[c * d + k], ss = (double)sumsq[c * d + k];
      const double m = s * inv_n;
      a0 += m;
      a1 += m * m;
      a2 += (ss - s * m) * inv_nm1;
      a3 += (m - muk) * (m - muk);
    }
  }
  __shared__ double sh[4][THREADS];
  sh[0][threadIdx.x] = a0; sh[1][threadIdx.x] = a1; sh[2][threadIdx.x] = a2; sh[3][threadIdx.x] = a3;
  __syncthreads();
  if (r == 0 && k < d) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      double t = 0;
      for (int rr = 0; rr < rows; ++rr) t += sh[j][rr * dw + threadIdx.x];
      atomicAdd(out + (size_t)j * d + k, t);
    }
  }
}

}  // namespace hmcb

using namespace hmcb;

extern "C" int hmcb_moments_accumulate(int32_t dtype, int64_t n_chains, int32_t d, const void* q, int64_t stride_chain,
                                       int64_t stride_dim, void* sum, void* sumsq, void* stream) {
  HMCB_CHECK_ARG(n_chains >= 0 && d >= 1, "bad sizes");
  if (n_chains == 0) return 0;
  HMCB_CHECK_ARG(q && sum && sumsq, "q/sum/sumsq must be device pointers");
  const int64_t n = n_chains * d;
  int64_t blocks = (n + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == HMCB_F64)
    k_moments_accumulate<double><<<(int)blocks, 256, 0, st>>>((const double*)q, stride_chain, stride_dim, d, n,
                                                              (double*)sum, (double*)sumsq);
  else if (dtype == HMCB_F32)
    k_moments_accumulate<float><<<(int)blocks, 256, 0, st>>>((const float*)q, stride_chain, stride_dim, d, n,
                                                             (float*)sum, (float*)sumsq);
  else {
    set_error("unknown dtype %d", dtype);
    return HMCB_E_ARG;
  }
  HMCB_CUDA(cudaPeekAtLastError());
  return 0;
}

extern "C" int hmcb_moments_reduce(int32_t dtype, int64_t n_chains, int32_t d, const void* sum, const void* sumsq,
                                   double n_samples, const double* true_mean, double* out, void* stream) {
  HMCB_CHECK_ARG(n_chains >= 0 && d >= 1 && n_samples >= 1, "bad sizes");
  HMCB_CHECK_ARG(out != nullptr, "out must be a device pointer to 4*d doubles");
  cudaStream_t st = (cudaStream_t)stream;
  HMCB_CUDA(cudaMemsetAsync(out, 0, sizeof(double) * 4 * d, st));
  if (n_chains == 0) return 0;
  HMCB_CHECK_ARG(sum && sumsq, "sum/sumsq must be device pointers");
  constexpr int THREADS = 256;
  const int dw = d < THREADS ? d : THREADS;
  const int rows = THREADS / dw, ky = (d + dw - 1) / dw;
  int64_t bx = (n_chains + rows - 1) / rows;
  const int64_t cap = ((int64_t)sm_count() * 8 + ky - 1) / ky;
  if (bx > cap) bx = cap;
  if (bx < 1) bx = 1;
  dim3 grid((unsigned)bx, (unsigned)ky);
  if (dtype == HMCB_F64)
    k_moments_reduce<double, THREADS><<<grid, THREADS, 0, st>>>((const double*)sum, (const double*)sumsq, n_chains, d,
                                                                n_samples, true_mean, out);
  else if (dtype == HMCB_F32)
    k_moments_reduce<float, THREADS><<<grid, THREADS, 0, st>>>((const float*)sum, (const float*)sumsq, n_chains, d,
                                                               n_samples, true_mean, out);
  else {
    set_error("unknown dtype %d", dtype);
    return HMCB_E_ARG;
  }
  HMCB_CUDA(cudaPeekAtLastError());
  return 0;
}
