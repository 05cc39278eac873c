// probe.cu -- pipe-peak measurement kernel (hmcb_probe_fma): each thread runs
// ILP independent chains of dependent FMAs held in registers, so the launch is
// bound by the FP64 (or FP32) FMA pipe and nothing else.  Its measured rate is
// the denominator for roofline fractions of the leapfrog kernels.
#include "common.cuh"

namespace hmcb {

// The multiplier and addend are compile-time immediates on purpose: with three register
// operands (x*a+b, a and b in registers) the same loop measured 34.1 TFLOP/s on B200 --
// operand fetch, not the pipe -- against 36.8-37.0 with immediates (= 148 SMs x 64 FMA/clk
// x 1.965 GHz).  The roofline denominator is the pipe, so the immediate form is used.
template <typename T, int ILP>
__global__ void __launch_bounds__(256) k_probe_fma(T* sink, long long iters) {
  T x[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) x[i] = (T)(threadIdx.x + i) * (T)1e-3;
  for (long long it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 8; ++u)
#pragma unroll
      for (int i = 0; i < ILP; ++i) x[i] = fma(x[i], (T)0.999999, (T)1e-7);
  }
  T s = T(0);
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += x[i];
  sink[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// The other ceiling: standard normals per second from Philox4x32-10 + Box-Muller with nothing else going on
// (hmcb_probe_normals).  MALA draws d normals per gradient, the K-fused and short-trajectory HMC kernels d per
// L+1 gradients: at small L they are bound by this generator's issue rate (integer multiplies and MUFU), not by
// the FP64 pipe.  ILP independent Philox blocks per thread per iteration, as fill_normals runs them (4).
template <int ILP, bool F64>
__global__ void __launch_bounds__(256) k_probe_normals(double* sink, long long iters, uint64_t seed) {
  const uint64_t chain = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  double acc = 0.0;
  for (long long it = 0; it < iters; ++it) {
    uint32_t w[ILP][4];
#pragma unroll
    for (int j = 0; j < ILP; ++j) Philox::gen(seed, chain, (uint32_t)it, STREAM_MOMENTUM, (uint32_t)j, w[j]);
#pragma unroll
    for (int j = 0; j < ILP; ++j) {
      if (F64) {
        double z[2];
        normal2_f64(w[j], z);
        acc += z[0] + z[1];
      } else {
        float z[4];
        normal4_f32(w[j], z);
        acc += (double)((z[0] + z[1]) + (z[2] + z[3]));
      }
    }
  }
  sink[chain] = acc;
}

}  // namespace hmcb

extern "C" int hmcb_probe_normals(int32_t f64_generator, int32_t blocks, int32_t threads, int64_t iters, void* sink,
                                  double* normal_count, void* stream) {
  using namespace hmcb;
  constexpr int ILP = 4;
  HMCB_CHECK_ARG(blocks > 0 && threads > 0 && threads <= 256 && iters > 0 && sink, "bad probe arguments");
  cudaStream_t st = (cudaStream_t)stream;
  if (f64_generator) k_probe_normals<ILP, true><<<blocks, threads, 0, st>>>((double*)sink, iters, 12345);
  else k_probe_normals<ILP, false><<<blocks, threads, 0, st>>>((double*)sink, iters, 12345);
  HMCB_CUDA(cudaPeekAtLastError());
  if (normal_count) *normal_count = (double)blocks * threads * (double)iters * ILP * (f64_generator ? 2.0 : 4.0);
  return 0;
}

extern "C" int hmcb_probe_fma(int32_t dtype, int32_t blocks, int32_t threads, int64_t iters, void* sink,
                              double* fma_count, void* stream) {
  using namespace hmcb;
  constexpr int ILP = 8;
  HMCB_CHECK_ARG(blocks > 0 && threads > 0 && threads <= 256 && iters > 0 && sink, "bad probe arguments");
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == HMCB_F64)
    k_probe_fma<double, ILP><<<blocks, threads, 0, st>>>((double*)sink, iters);
  else if (dtype == HMCB_F32)
    k_probe_fma<float, ILP><<<blocks, threads, 0, st>>>((float*)sink, iters);
  else {
    set_error("unknown dtype %d", dtype);
    return HMCB_E_ARG;
  }
  HMCB_CUDA(cudaPeekAtLastError());
  if (fma_count) *fma_count = (double)blocks * threads * (double)iters * 8.0 * ILP;
  return 0;
}
