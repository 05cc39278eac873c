// transition_ws.cuh -- warp-specialised form of the single-trajectory transition kernel for
// thread-per-chain shapes (LPC = 1) in float64.
//
// Why: at d = 50 a chain's state fills the register file (255 registers, two warps per
// scheduler), and every transition starts with work that does not touch the FP64 pipe --
// 13 Philox blocks and their Box-Muller transforms per chain.  While a warp is in that
// prologue its scheduler has one warp left to feed the pipe.  Here one block per SM carries two
// compute warpgroups (256 chains, 240 registers each after setmaxnreg.inc) and one helper
// warpgroup (24 registers after setmaxnreg.dec) that generates the NEXT round's momenta into a
// double-buffered shared-memory tile behind a pair of mbarriers; the compute warps' prologue
// shrinks to "read 50 floats from shared memory".  Measured on the d=50 funnel, L=150, 2^20
// chains: 2.14 ms -> 1.89 ms (0.70 -> 0.80 of the 36.7 TFLOP/s pipe).
//
// Draws are the same pure function of (seed, global chain id, iteration, element) as in every
// other kernel, so results are bit-identical to k_transition's.
#pragma once
#include "sync_ptx.cuh"   // mbarrier wrappers
#include "transition.cuh"

namespace hmcb {

constexpr int kWsComputeThreads = 256;
constexpr int kWsThreads = 384;

template <class Model, typename T, int EPL, int MODE>
__global__ void __launch_bounds__(kWsThreads, 1) k_transition_ws(const Model model, const TransArgs<T> a) {
  static_assert(sizeof(T) == 8, "the warp-specialised kernel is float64 only");
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* sP = reinterpret_cast<float*>(smem_raw);                              // [2][EPL][256]
  uint64_t* full = reinterpret_cast<uint64_t*>(sP + 2 * EPL * kWsComputeThreads);
  uint64_t* empty = full + 2;
  const int tid = threadIdx.x, wg = tid >> 7;
  if (tid == 0) {
    for (int s = 0; s < 2; ++s) {
      lr::mbar_init(full + s, kWsThreads - kWsComputeThreads);
      lr::mbar_init(empty + s, kWsComputeThreads);
    }
    lr::fence_barrier_init();
  }
  __syncthreads();
  const int d = a.d;
  const int64_t per_round = (int64_t)gridDim.x * kWsComputeThreads;
  const int rounds = (int)((a.n_chains + per_round - 1) / per_round);
  if (wg == 2) {
    // ---------------- helper warpgroup: momenta for the round after the one being integrated
    asm volatile("setmaxnreg.dec.sync.aligned.u32 24;");
    const int ht = tid - kWsComputeThreads;
    for (int r = 0; r < rounds; ++r) {
      const int s = r & 1;
      lr::mbar_wait(empty + s, ((r >> 1) & 1) ^ 1);
      float* dst = sP + (size_t)s * EPL * kWsComputeThreads;
#pragma unroll 1
      for (int h = 0; h < 2; ++h) {
        const int cth = ht + 128 * h;
        const int64_t chain = ((int64_t)r * gridDim.x + blockIdx.x) * kWsComputeThreads + cth;
        if (chain < a.n_chains) {
          const uint64_t gchain = (uint64_t)(a.chain_offset + chain);
#pragma unroll 1
          for (int b = 0; b < (EPL + 3) / 4; ++b) {
            uint32_t w[4];
            float z[4];
            Philox::gen(a.seed, gchain, a.iteration, STREAM_MOMENTUM, (uint32_t)b, w);
            normal4_f32(w, z);
#pragma unroll
            for (int k = 0; k < 4; ++k)
              if (4 * b + k < EPL) dst[(4 * b + k) * kWsComputeThreads + cth] = (4 * b + k < d) ? z[k] : 0.f;
          }
        }
      }
      lr::mbar_arrive(full + s);
    }
  } else {
    // ---------------- compute warpgroups: one chain per thread
    asm volatile("setmaxnreg.inc.sync.aligned.u32 240;");
    GroupCtx<T> c;
    c.lane = 0; c.d = d; c.mask = 0xffffffffu; c.smem_q = nullptr;
    T minv[1] = {T(1)};
    for (int r = 0; r < rounds; ++r) {
      const int s = r & 1;
      const int64_t chain = ((int64_t)r * gridDim.x + blockIdx.x) * kWsComputeThreads + tid;
      const bool live = chain < a.n_chains;
      const int64_t qoff = (live ? chain : 0) * a.sc;
      T q[EPL], p[EPL];
      load_vec<T, EPL, 1>(q, a.q_in, qoff, a.sd, 0, d, T(1));
      lr::mbar_wait(full + s, (r >> 1) & 1);
      const float* src = sP + (size_t)s * EPL * kWsComputeThreads + tid;
#pragma unroll
      for (int k = 0; k < EPL; ++k) p[k] = (T)src[k * kWsComputeThreads];
      lr::mbar_arrive(empty + s);
      if (live) finish_single<Model, T, EPL, 1, true, MODE>(model, a, c, minv, chain, qoff, q, p);
    }
  }
}

}  // namespace hmcb
