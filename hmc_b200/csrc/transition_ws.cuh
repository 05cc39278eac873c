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
  T* sQ = reinterpret_cast<T*>(smem_raw);                                           // [EPL][256] q0 of the round (optional)
  float* sP = reinterpret_cast<float*>(sQ + (a.tile_io ? (size_t)EPL * kWsComputeThreads : 0));   // [2][EPL][256]
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
      if (a.tile_io) {
#pragma unroll
        for (int k = 0; k < EPL; ++k) sQ[k * kWsComputeThreads + tid] = q[k];
      }
      lr::mbar_wait(full + s, (r >> 1) & 1);
      const float* src = sP + (size_t)s * EPL * kWsComputeThreads + tid;
#pragma unroll
      for (int k = 0; k < EPL; ++k) p[k] = (T)src[k * kWsComputeThreads];
      lr::mbar_arrive(empty + s);
      if (live)
        finish_single<Model, T, EPL, 1, true, MODE>(model, a, c, minv, chain, qoff, q, p, nullptr,
                                                    a.tile_io ? sQ + tid : nullptr, kWsComputeThreads);
    }
  }
}

// ------------------------------------------------------------------ K HMC transitions per launch
// The loop of HMC.sample (src/algorithms.py:244-250; scripts_old/funnel/funnelmpi.py:152-162) inside the
// warp-specialised kernel: a chain's position stays in its thread's registers for all K transitions and the start
// point q0 of the current one -- needed for the no-move test and to come back to on rejection -- in a shared-memory
// tile, so a transition touches global memory only for its accept test's outputs.  At short trajectories (L = 20,
// the reference's usual setting) that is what a transition costs beside its leapfrog steps: one launch per transition
// spends more time on the strided load and store of q than on integrating.  The helper warpgroup produces the momenta
// of (round, transition) pairs one tile ahead, as above.  Every draw is the same function of (seed, global chain id,
// iteration + k) as in k transitions by single launches, so the results are bit-identical to those.
template <class Model, typename T, int EPL>
__global__ void __launch_bounds__(kWsThreads, 1) k_hmc_multi_ws(const Model model, const TransArgs<T> a, const int K,
                                                                int32_t* __restrict__ n_accept) {
  static_assert(sizeof(T) == 8, "the warp-specialised kernel is float64 only");
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* sQ = reinterpret_cast<T*>(smem_raw);                                           // [EPL][256] q0 of the running transition
  float* sP = reinterpret_cast<float*>(sQ + (size_t)EPL * kWsComputeThreads);       // [2][EPL][256] momenta
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
    asm volatile("setmaxnreg.dec.sync.aligned.u32 24;");
    const int ht = tid - kWsComputeThreads;
    uint32_t t = 0;
    for (int r = 0; r < rounds; ++r) {
      for (int k = 0; k < K; ++k, ++t) {
        const int s = t & 1;
        lr::mbar_wait(empty + s, ((t >> 1) & 1) ^ 1);
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
              Philox::gen(a.seed, gchain, a.iteration + (uint32_t)k, STREAM_MOMENTUM, (uint32_t)b, w);
              normal4_f32(w, z);
#pragma unroll
              for (int e = 0; e < 4; ++e)
                if (4 * b + e < EPL) dst[(4 * b + e) * kWsComputeThreads + cth] = (4 * b + e < d) ? z[e] : 0.f;
            }
          }
        }
        lr::mbar_arrive(full + s);
      }
    }
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 240;");
    GroupCtx<T> c;
    c.lane = 0; c.d = d; c.mask = 0xffffffffu; c.smem_q = nullptr;
    T minv[1] = {T(1)};
    const int N = a.L[0] > 1 ? a.L[0] : 1;
    uint32_t t = 0;
    for (int r = 0; r < rounds; ++r) {
      const int64_t chain = ((int64_t)r * gridDim.x + blockIdx.x) * kWsComputeThreads + tid;
      const bool live = chain < a.n_chains;
      const int64_t qoff = (live ? chain : 0) * a.sc;
      T q[EPL], p[EPL];
      load_vec<T, EPL, 1>(q, a.q_in, qoff, a.sd, 0, d, T(1));
#pragma unroll
      for (int e = 0; e < EPL; ++e) sQ[e * kWsComputeThreads + tid] = q[e];
      int nacc = 0;
      for (int k = 0; k < K; ++k, ++t) {
        const int s = t & 1;
        lr::mbar_wait(full + s, (t >> 1) & 1);
        const float* src = sP + (size_t)s * EPL * kWsComputeThreads + tid;
#pragma unroll
        for (int e = 0; e < EPL; ++e) p[e] = (T)src[e * kWsComputeThreads];
        lr::mbar_arrive(empty + s);
        if (live) {
          // one transition, as finish_single<MODE_HMC> (src/pyhmc.py:64-69, src/algorithms.py:192-198)
          const T K0 = kinetic<T, EPL, 1, true>(p, minv, c.mask);
          T V0, V1;
          leapfrog<Model, T, EPL, 1, true>(model, q, p, minv, N, a.eps[0], c, V0, V1);
          const T H0 = V0 + K0;
          const T H1 = V1 + kinetic<T, EPL, 1, true>(p, minv, c.mask);
          const T dq = diff_sum<T, EPL, 1>(q, sQ, tid, kWsComputeThreads, 0, d);   // sum(q0 - q1), src/pyhmc.py:57
          const T prob = t_exp<T>(H0 - H1);
          const bool bad = t_bad(prob) || dq == T(0);
          const double u = draw_uniform<T>(a, chain, 0, (uint32_t)k);
          const T pm = prob < T(1) ? prob : T(1);
          const bool accept = !bad && !(u > (double)pm);
          if (k == K - 1) {
            if (a.code) a.code[chain] = bad ? (a.rule == HMCB_RULE_PYHMC ? 2 : -1) : (accept ? 1 : 0);
            if (a.energies) { a.energies[2 * chain] = H0; a.energies[2 * chain + 1] = H1; }
            if (a.accept_prob) a.accept_prob[chain] = (accept && !t_bad(prob)) ? pm : T(0);
            if (a.p_out) {
              if (!accept) fill_normals<T, EPL, 1>(p, a.seed, (uint64_t)(a.chain_offset + chain), a.iteration + (uint32_t)k, 0, d,
                                                  a.rng_flags);
              store_vec<T, EPL, 1>(p, a.p_out, qoff, a.sd, 0, d);
            }
          }
          if (accept) {
            ++nacc;
#pragma unroll
            for (int e = 0; e < EPL; ++e) sQ[e * kWsComputeThreads + tid] = q[e];
          } else {
#pragma unroll
            for (int e = 0; e < EPL; ++e) q[e] = sQ[e * kWsComputeThreads + tid];
          }
        }
      }
      if (live) {
        store_vec<T, EPL, 1>(q, a.q_out, qoff, a.sd, 0, d);
        if (n_accept) n_accept[chain] += nacc;
        if (a.counts) { a.counts[3 * chain] = 2 * K; a.counts[3 * chain + 1] = K * (N + 1); a.counts[3 * chain + 2] = K; }
      }
    }
  }
}

// ------------------------------------------------------------------ K delayed-rejection transitions, short trajectories
// At the reference's usual trajectory lengths (L = 20) nearly every chain is accepted at the first stage, and a transition
// by single launches costs more in loading and storing q than in integrating (k_hmc_multi_ws above).  This kernel runs a
// chain's consecutive transitions with q in registers AS LONG AS THE FIRST STAGE ACCEPTS; at the first rejection the chain
// hands itself to the delayed-rejection stage kernels exactly as stage 0 of k_transition does (H0, alpha_0, counters, list
// entry, plus the transition it stands in: iter_of) and its lane idles until the round is over.  The stage kernels finish
// that transition and put chains with transitions left on a list; the next pass of this kernel (list != nullptr) takes
// them on from there.  hmcb_drhmc_sample alternates the two until every chain has done its K transitions.  Draws are
// keyed by (chain, iteration + k), so the results are those of K single hmcb_drhmc_step calls, bit for bit.
template <class Model, typename T, int EPL>
__global__ void __launch_bounds__(kWsThreads, 1) k_dr0_multi_ws(const Model model, const TransArgs<T> a, const int max_steps,
                                                                const int32_t* __restrict__ list,
                                                                const int32_t* __restrict__ list_count,
                                                                int32_t* __restrict__ iter_of) {
  static_assert(sizeof(T) == 8, "the warp-specialised kernel is float64 only");
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* sQ = reinterpret_cast<T*>(smem_raw);                                           // [EPL][256] q0 of the running transition
  float* sP = reinterpret_cast<float*>(sQ + (size_t)EPL * kWsComputeThreads);       // [2][EPL][256] momenta
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
  const int d = a.d, K = a.K_total;
  const int64_t count = list ? (int64_t)list_count[0] : a.n_chains;
  const int64_t per_round = (int64_t)gridDim.x * kWsComputeThreads;
  const int rounds = (int)((count + per_round - 1) / per_round);
  if (wg == 2) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 24;");
    const int ht = tid - kWsComputeThreads;
    uint32_t t = 0;
    for (int r = 0; r < rounds; ++r) {
      // the two chains this helper thread draws for, and the transition each of them starts this pass with
      int64_t ch[2];
      int k0[2];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int64_t idx = ((int64_t)r * gridDim.x + blockIdx.x) * kWsComputeThreads + ht + 128 * h;
        ch[h] = idx < count ? (list ? (int64_t)list[idx] : idx) : -1;
        k0[h] = (ch[h] >= 0 && list) ? a.iter_done[ch[h]] : 0;
      }
      for (int j = 0; j < max_steps; ++j, ++t) {
        const int s = t & 1;
        lr::mbar_wait(empty + s, ((t >> 1) & 1) ^ 1);
        float* dst = sP + (size_t)s * EPL * kWsComputeThreads;
#pragma unroll 1
        for (int h = 0; h < 2; ++h) {
          const int cth = ht + 128 * h;
          if (ch[h] >= 0 && k0[h] + j < K) {
            const uint64_t gchain = (uint64_t)(a.chain_offset + ch[h]);
#pragma unroll 1
            for (int b = 0; b < (EPL + 3) / 4; ++b) {
              uint32_t w[4];
              float z[4];
              Philox::gen(a.seed, gchain, a.iteration + (uint32_t)(k0[h] + j), STREAM_MOMENTUM, (uint32_t)b, w);
              normal4_f32(w, z);
#pragma unroll
              for (int e = 0; e < 4; ++e)
                if (4 * b + e < EPL) dst[(4 * b + e) * kWsComputeThreads + cth] = (4 * b + e < d) ? z[e] : 0.f;
            }
          }
        }
        lr::mbar_arrive(full + s);
      }
    }
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 240;");
    GroupCtx<T> c;
    c.lane = 0; c.d = d; c.mask = 0xffffffffu; c.smem_q = nullptr;
    T minv[1] = {T(1)};
    const int N = a.L[0] > 1 ? a.L[0] : 1;
    const int np = a.nprops;
    uint32_t t = 0;
    for (int r = 0; r < rounds; ++r) {
      const int64_t idx = ((int64_t)r * gridDim.x + blockIdx.x) * kWsComputeThreads + tid;
      const bool live = idx < count;
      const int64_t chain = live ? (list ? (int64_t)list[idx] : idx) : 0;
      const int64_t qoff = chain * a.sc;
      int k = (live && list) ? a.iter_done[chain] : 0;
      T q[EPL], p[EPL];
      load_vec<T, EPL, 1>(q, a.q_out, qoff, a.sd, 0, d, T(1));          // the sample call works in place on q_out
#pragma unroll
      for (int e = 0; e < EPL; ++e) sQ[e * kWsComputeThreads + tid] = q[e];
      bool active = live;
      int nacc = 0, ndone = 0, last_code = 0;
      T last_a0 = T(0), last_pt0 = T(1);
      for (int j = 0; j < max_steps; ++j, ++t) {
        const int s = t & 1;
        lr::mbar_wait(full + s, (t >> 1) & 1);
        const float* src = sP + (size_t)s * EPL * kWsComputeThreads + tid;
#pragma unroll
        for (int e = 0; e < EPL; ++e) p[e] = (T)src[e * kWsComputeThreads];
        lr::mbar_arrive(empty + s);
        if (active && k < K) {
          // stage 0 of DRHMC.step (src/algorithms.py:506-529 at j = 0), as finish_single<MODE_DR0>
          const T K0 = kinetic<T, EPL, 1, true>(p, minv, c.mask);
          T V0, V1;
          leapfrog<Model, T, EPL, 1, true>(model, q, p, minv, N, a.eps[0], c, V0, V1);
          const T H0 = V0 + K0;
          const T H1 = V1 + kinetic<T, EPL, 1, true>(p, minv, c.mask);
          const T dq = diff_sum<T, EPL, 1>(q, sQ, tid, kWsComputeThreads, 0, d);
          T prob = t_exp<T>(H0 - H1);
          const T pf_raw = prob;
          if (dq == T(0)) prob = T(0);
          if (t_bad(prob)) prob = T(0);
          const T a0 = prob < T(1) ? prob : T(1);
          T pt0 = T(1);
          if (a.dr_prob && !t_bad(pf_raw)) pt0 = T(1) - (a.dr_prob == 1 ? (pf_raw < T(1) ? pf_raw : T(1)) : a0);
          const double u = draw_uniform<T>(a, chain, 0, (uint32_t)k);
          const bool accept = a.dr_conv ? (u <= (double)a0) : (u < (double)a0);
          bool finished = accept || np == 1;
          int code = a.dr_conv ? (accept ? 0 : -1) : (accept ? 1 : 0);
          if (!finished && a.dr_prob) {
            // the try test of stage 1 (src/algorithms.py:710-711) needs nothing but the uniform: a declined stage ends the
            // transition here (the stage kernel would find the same and return code -2)
            const double u_try = draw_uniform<T>(a, chain, 1, (uint32_t)k);
            if (u_try > (double)pt0) { finished = true; code = -2; }
          }
          if (finished) {
            if (k == K - 1 && a.p_out) {
              if (!accept) fill_normals<T, EPL, 1>(p, a.seed, (uint64_t)(a.chain_offset + chain), a.iteration + (uint32_t)k, 0, d,
                                                  a.rng_flags);
              store_vec<T, EPL, 1>(p, a.p_out, qoff, a.sd, 0, d);
            }
            if (accept) {
              ++nacc;
#pragma unroll
              for (int e = 0; e < EPL; ++e) sQ[e * kWsComputeThreads + tid] = q[e];
            } else {
#pragma unroll
              for (int e = 0; e < EPL; ++e) q[e] = sQ[e * kWsComputeThreads + tid];
            }
            ++ndone; ++k;
            last_code = code; last_a0 = a0; last_pt0 = pt0;
          } else {
            // rejected at stage 0 with stages left: over to the stage kernels (as finish_single does)
            a.ws.H0[chain] = H0;
            a.ws.avec[chain * kMaxProps] = a0;
            if (a.dr_prob) a.ws.ptries[chain * kMaxProps] = pt0;
            a.ws.counts[3 * chain] = 2; a.ws.counts[3 * chain + 1] = N + 1; a.ws.counts[3 * chain + 2] = 1;
            iter_of[chain] = k;
            const int slot = atomicAdd(a.ws.count + 1, 1);
            a.ws.list[(int64_t)1 * a.n_chains + slot] = (int32_t)chain;
#pragma unroll
            for (int e = 0; e < EPL; ++e) q[e] = sQ[e * kWsComputeThreads + tid];
            active = false;
          }
        }
      }
      if (live) {
        store_vec<T, EPL, 1>(q, a.q_out, qoff, a.sd, 0, d);
        a.iter_done[chain] = k;
        if (ndone > 0) {
          if (a.code) a.code[chain] = last_code;
          if (a.avec_out) {
            a.avec_out[chain * np] = last_a0;
            for (int i = 1; i < np; ++i) a.avec_out[chain * np + i] = T(0);
          }
          if (a.ptries_out) {
            a.ptries_out[chain * np] = last_pt0;
            for (int i = 1; i < np; ++i) a.ptries_out[chain * np + i] = T(1);
          }
          if (a.counts) { a.counts[3 * chain] += 2 * ndone; a.counts[3 * chain + 1] += ndone * (N + 1); a.counts[3 * chain + 2] += ndone; }
          if (a.n_accept && nacc) a.n_accept[chain] += nacc;
        }
      }
    }
  }
}

}  // namespace hmcb
