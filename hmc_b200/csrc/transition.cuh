// transition.cuh -- the fused transition kernel: momentum draw, leapfrog
// trajectories, energies, accept test, counters, for a block of chains whose
// state lives in registers.
//
// One kernel template serves three modes:
//   MODE_HMC     vanilla HMC transition   (src/pyhmc.py:64-69, src/algorithms.py:192-198)
//   MODE_DR0     delayed-rejection stage 0 over all chains (src/algorithms.py:500-530, j = 0)
//   MODE_DRN     delayed-rejection stage J >= 1 over the compacted list of chains
//                still rejected, including the ghost recursion get_num
//                (src/algorithms.py:474-497) run as an explicit-stack state machine.
#pragma once
#include <cooperative_groups.h>

#include "common.cuh"
#include "models.cuh"

namespace hmcb {
namespace cg = cooperative_groups;

enum { MODE_HMC = 0, MODE_DR0 = 1, MODE_DRN = 2 };
constexpr int kMaxProps = 6;  // max delayed-rejection stages

// Scratch for delayed rejection (caller-owned; carved by the host, see drhmc.cu).
template <typename T>
struct DrScratch {
  int32_t* count;    // [kMaxProps+1] chains queued per stage
  int32_t* list;     // [kMaxProps][n_chains] local chain ids queued for stage j
  T* H0;             // [n_chains] energy at (q0,p0)
  T* avec;           // [n_chains*kMaxProps] running stage acceptance probabilities
  T* ptries;         // [n_chains*kMaxProps] running try probabilities (probabilistic DR only)
  int32_t* counts;   // [n_chains*3] running Hcount, Vgcount, leapcount
  T* frames;         // [(kMaxProps-1)*2*EPL*frame_threads] per-thread proposal frames
  int64_t frame_threads;
  // continuation of a stage >= 2 (dr_phase): chains whose ghost recursion goes on after its first trajectory
  int32_t* ccount;   // [1] chains parked by the first part of the running stage
  int32_t* clist;    // [ccap] their chain ids
  int32_t* cst_i;    // [3][ccap] Hcount, Vgcount, leapcount so far
  T* cst_f;          // [5][ccap] lHb[1], lpf[0], lav[1][0], lpt[0][J], lpt[1][0]
  T* cframes;        // [2*EPL][ccap] the stage's proposal (q, p), element s of slot k at s*ccap + k
  int32_t ccap;
};

template <typename T>
struct TransArgs {
  int64_t n_chains, chain_offset;
  int32_t d;
  const T* q_in;
  T* q_out;
  int64_t sc, sd;
  uint64_t seed;
  uint32_t iteration;
  int32_t rng_flags;
  const T* p_inject;
  const double* u_inject;
  int32_t u_stride;
  const T* p_prev;   // partial momentum refresh (GHMC family): p = alpha*p_prev + sqrt(1-alpha^2)*draw; nullptr = full refresh
  T alpha, beta;     // beta = sqrt(1 - alpha^2)
  T* p_out;
  int32_t* code;
  T* energies;
  T* avec_out;
  int32_t* counts;
  T* accept_prob;
  const T* minv;     // inverse metric diagonal or nullptr
  int32_t rule;      // HMCB_RULE_*
  int32_t dr_conv;   // 0 algorithms.DRHMC conventions, 1 pyhmc.PyHMC_multistep's (also the probabilistic classes')
  int32_t dr_prob;   // probabilistic DR: 0 off, 1 p_try = 1 - min(1, e^{H0-Hj}), 2 p_try = 1 - alpha_j
  T* ptries_out;     // [n_chains, nprops] or nullptr
  T* p0_out;         // starting momentum of every chain (layout of q) or nullptr
  int32_t nprops, stage;
  int32_t tile_io;   // set by the launcher when the per-warp staging tile was allocated (see warp_tile_load)
  int32_t dr_phase;  // MODE_DRN: 0 the whole stage in one kernel; 1 up to the first point where the chains of a warp part
                     // ways (after the first ghost trajectory), parking those that go on; 2 the parked chains
  // multi-pass sample mode for short trajectories (k_dr0_multi_ws + these stage kernels, see hmcb_drhmc_sample):
  const int32_t* iter_of;   // [n_chains] which transition of the call a chain's running delayed-rejection stages belong to
  int32_t* iter_done;       // [n_chains] transitions a chain has completed
  int32_t* cont_list;       // chains that completed a transition here and have more to do: run on by the next pass
  int32_t* cont_count;      // [1]
  int32_t* n_accept;        // [n_chains] += 1 per accepted transition
  int32_t K_total;          // transitions per chain in the call
  int32_t prefer_plain;  // caller's wish: k_transition rather than the warp-specialised kernel (pipelined sample calls)
  int32_t L[kMaxProps];
  T eps[kMaxProps];
  DrScratch<T> ws;
};

// ------------------------------------------------------------------ state movement
// stride_dim == 1 (row-major [C,d], the API's layout) is special-cased so that element
// addresses are "row pointer + compile-time offset": no per-element 64-bit index math.
template <typename T, int EPL, int LPC>
__device__ __forceinline__ void load_vec(T (&x)[EPL], const T* base, int64_t off, int64_t sd, int lane, int d, T sign) {
  if (sd == 1) {
    const T* row = base + off + 4 * lane;
#pragma unroll
    for (int s = 0; s < EPL; ++s) {
      const int i = Layout<EPL, LPC>::elem(lane, s);
      x[s] = (i < d) ? sign * row[Layout<EPL, LPC>::elem(0, s)] : T(0);
    }
  } else {
#pragma unroll
    for (int s = 0; s < EPL; ++s) {
      const int i = Layout<EPL, LPC>::elem(lane, s);
      x[s] = (i < d) ? sign * base[off + i * sd] : T(0);
    }
  }
}

template <typename T, int EPL, int LPC>
__device__ __forceinline__ void store_vec(const T (&x)[EPL], T* base, int64_t off, int64_t sd, int lane, int d) {
  if (sd == 1) {
    T* row = base + off + 4 * lane;
#pragma unroll
    for (int s = 0; s < EPL; ++s) {
      const int i = Layout<EPL, LPC>::elem(lane, s);
      if (i < d) row[Layout<EPL, LPC>::elem(0, s)] = x[s];
    }
  } else {
#pragma unroll
    for (int s = 0; s < EPL; ++s) {
      const int i = Layout<EPL, LPC>::elem(lane, s);
      if (i < d) base[off + i * sd] = x[s];
    }
  }
}

// sum over the chain of (ref[i] - x[i]) -- the reference's "did not move" test
// (q0-q1).sum()==0, src/pyhmc.py:57 -- with four partial sums (the serial 50-add chain
// behind 50 loads was 20% of the L=1 kernel).  Caller reduces over the group's lanes.
template <typename T, int EPL, int LPC>
__device__ __forceinline__ T diff_sum(const T (&x)[EPL], const T* base, int64_t off, int64_t sd, int lane, int d) {
  T acc[4] = {T(0), T(0), T(0), T(0)};
  const T* row = base + off + (sd == 1 ? 4 * lane : 0);
  // chunks of 8 loads keep the extra live registers small: the state already fills the file
#pragma unroll
  for (int s0 = 0; s0 < EPL; s0 += 8) {
    T r[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int s = s0 + k;
      if (s < EPL) {
        const int i = Layout<EPL, LPC>::elem(lane, s);
        r[k] = (i < d) ? (sd == 1 ? row[Layout<EPL, LPC>::elem(0, s)] : base[off + i * sd]) : T(0);
      }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k)
      if (s0 + k < EPL) acc[k & 3] += r[k] - x[s0 + k];   // padded slots: 0 - 0
  }
  return (acc[0] + acc[1]) + (acc[2] + acc[3]);
}

// Standard normals for every slot of this lane.  The generator runs as a ROLLED loop
// over the lane's 4-element blocks that parks its output in a small local-memory array,
// and only then are the values copied into the (compile-time indexed) register slots.
// Inlining the Philox rounds and log/sincos bodies into every unrolled slot made the
// kernel 26k instructions; calling an out-of-line function per block instead spilled the
// ~100 live state registers around each call and cost 0.55 ms per 2^20-chain transition
// (a quarter of the L=150 kernel).  The loop form has neither problem.
// CNT Philox blocks of one chain side by side (independent ten-round chains), 4 normals each
template <int CNT, int LPC>
__device__ __forceinline__ void normal_group(float* zl, int b0, uint64_t seed, uint64_t gchain, uint32_t iter, int lane) {
  uint32_t w[CNT][4];
  float z4[CNT][4];
#pragma unroll
  for (int j = 0; j < CNT; ++j) Philox::gen(seed, gchain, iter, STREAM_MOMENTUM, (uint32_t)((b0 + j) * LPC + lane), w[j]);
#pragma unroll
  for (int j = 0; j < CNT; ++j) normal4_f32(w[j], z4[j]);
#pragma unroll
  for (int j = 0; j < CNT; ++j) {
    zl[4 * (b0 + j)] = z4[j][0]; zl[4 * (b0 + j) + 1] = z4[j][1];
    zl[4 * (b0 + j) + 2] = z4[j][2]; zl[4 * (b0 + j) + 3] = z4[j][3];
  }
}

template <typename T, int EPL, int LPC>
__device__ __forceinline__ void fill_normals(T (&z)[EPL], uint64_t seed, uint64_t gchain, uint32_t iter, int lane,
                                             int d, int flags) {
  constexpr int NB = Layout<EPL, LPC>::kBlocks;
  if (flags & HMCB_RNG_NORMAL_F64) {
    double zl[4 * NB];
#pragma unroll 1
    for (int b = 0; b < NB; ++b) {
      const uint32_t blk = (uint32_t)(b * LPC + lane);
      uint32_t w[4];
      double z2[2];
      Philox::gen(seed, gchain, iter, STREAM_MOMENTUM, 2 * blk, w);
      normal2_f64(w, z2);
      zl[4 * b] = z2[0]; zl[4 * b + 1] = z2[1];
      Philox::gen(seed, gchain, iter, STREAM_MOMENTUM, 2 * blk + 1, w);
      normal2_f64(w, z2);
      zl[4 * b + 2] = z2[0]; zl[4 * b + 3] = z2[1];
    }
#pragma unroll
    for (int s = 0; s < EPL; ++s) z[s] = (Layout<EPL, LPC>::elem(lane, s) < d) ? (T)zl[s] : T(0);
  } else {
    constexpr int UNR = NB >= 4 ? 4 : NB;          // independent blocks in flight per iteration
    constexpr int NFULL = NB / UNR * UNR;          // blocks covered by whole groups; the rest is one smaller group
    float zl[4 * NB];
#pragma unroll 1
    for (int b0 = 0; b0 < NFULL; b0 += UNR) normal_group<UNR, LPC>(zl, b0, seed, gchain, iter, lane);
    if constexpr (NB > NFULL) normal_group<NB - NFULL, LPC>(zl, NFULL, seed, gchain, iter, lane);
#pragma unroll
    for (int s = 0; s < EPL; ++s) z[s] = (Layout<EPL, LPC>::elem(lane, s) < d) ? (T)zl[s] : T(0);
  }
}

// p = alpha*p_prev + sqrt(1 - alpha**2)*pn (src/algorithms.py:269,603) with numpy's rounding: two products and
// a sum, none of them fused
__device__ __forceinline__ double mix_momentum(double alpha, double pv, double beta, double pn) {
  return __dadd_rn(__dmul_rn(alpha, pv), __dmul_rn(beta, pn));
}
__device__ __forceinline__ float mix_momentum(float alpha, float pv, float beta, float pn) {
  return __fadd_rn(__fmul_rn(alpha, pv), __fmul_rn(beta, pn));
}

// momentum p = z * metricstd, z ~ N(0,1) (src/pyhmc.py:66); z from Philox or injected
template <typename T, int EPL, int LPC, bool UNITM>
__device__ __forceinline__ void draw_momentum(T (&p)[EPL], const TransArgs<T>& a, const T (&minv)[UNITM ? 1 : EPL],
                                              int64_t chain, int lane, int d, uint32_t iter_add = 0) {
  if (a.p_inject) {
    load_vec<T, EPL, LPC>(p, a.p_inject, chain * a.sc, a.sd, lane, d, T(1));
  } else {
    fill_normals<T, EPL, LPC>(p, a.seed, (uint64_t)(a.chain_offset + chain), a.iteration + iter_add, lane, d, a.rng_flags);
  }
  if (!UNITM) {
#pragma unroll
    for (int s = 0; s < EPL; ++s) p[s] *= t_rsqrt<T>(minv[UNITM ? 0 : s]);
  }
  if (a.p_prev) {
    // p = alpha*p + sqrt(1 - alpha**2)*pn (src/algorithms.py:269,603)
    T pv[EPL];
    load_vec<T, EPL, LPC>(pv, a.p_prev, chain * a.sc, a.sd, lane, d, T(1));
#pragma unroll
    for (int s = 0; s < EPL; ++s) p[s] = mix_momentum(a.alpha, pv[s], a.beta, p[s]);
  }
}

template <typename T>
__device__ __forceinline__ double draw_uniform(const TransArgs<T>& a, int64_t chain, int idx, uint32_t iter_add = 0) {
  if (a.u_inject) return a.u_inject[chain * a.u_stride + idx];
  uint32_t w[4];
  Philox::gen(a.seed, (uint64_t)(a.chain_offset + chain), a.iteration + iter_add, STREAM_ACCEPT, (uint32_t)(idx >> 1), w);
  return (idx & 1) ? u53(w[2], w[3]) : u53(w[0], w[1]);
}

// KE = 0.5 * sum(p^2 * invmetric) (src/pyhmc.py:18)
template <typename T, int EPL, int LPC, bool UNITM>
__device__ __forceinline__ T kinetic(const T (&p)[EPL], const T (&minv)[UNITM ? 1 : EPL], unsigned mask) {
  T acc[4] = {T(0), T(0), T(0), T(0)};
#pragma unroll
  for (int s = 0; s < EPL; ++s) acc[s & 3] = fma(UNITM ? p[s] : p[s] * minv[UNITM ? 0 : s], p[s], acc[s & 3]);
  T k = (acc[0] + acc[1]) + (acc[2] + acc[3]);
#pragma unroll
  for (int o = LPC >> 1; o > 0; o >>= 1) k += __shfl_xor_sync(mask, k, o);
  return T(0.5) * k;
}

template <typename T, int EPL, bool UNITM>
__device__ __forceinline__ void drift(T (&q)[EPL], const T (&p)[EPL], const T (&minv)[UNITM ? 1 : EPL], T eps) {
#pragma unroll
  for (int s = 0; s < EPL; ++s) q[s] = fma(eps, UNITM ? p[s] : p[s] * minv[UNITM ? 0 : s], q[s]);
}

// Leapfrog of N steps (src/pyhmc.py:35-48): half kick, N-1 x (drift, kick), drift,
// half kick = N+1 gradient evaluations.  Returns V at the start and end points.
template <class Model, typename T, int EPL, int LPC, bool UNITM>
__device__ __forceinline__ void leapfrog(const Model& model, T (&q)[EPL], T (&p)[EPL],
                                         const T (&minv)[UNITM ? 1 : EPL], int N, T eps, const GroupCtx<T>& c,
                                         T& V_start, T& V_end) {
  V_start = model.template kick<EPL, LPC, true>(q, p, T(0.5) * eps, c);
#pragma unroll 1
  for (int i = 1; i < N; ++i) {
    drift<T, EPL, UNITM>(q, p, minv, eps);
    model.template kick<EPL, LPC, false>(q, p, eps, c);
  }
  drift<T, EPL, UNITM>(q, p, minv, eps);
  V_end = model.template kick<EPL, LPC, true>(q, p, T(0.5) * eps, c);
}

// ------------------------------------------------------------------ coalesced state I/O through a warp tile
// Row-major [C, d] chain state read lane-per-chain touches 32 different sectors per load instruction
// (ncu at L = 20: 27 sectors per request, long_scoreboard 3.1 per issue; one transition of 2^20 chains
// of d = 50 cost 0.51 ms before its first leapfrog step).  The chains of a warp are adjacent rows, so
// their states are one contiguous run of (32/LPC)*d elements: the warp copies that run with
// consecutive lanes on consecutive elements into a shared-memory tile (row stride d|1, so that the
// row-per-lane reads that follow are free of bank conflicts), takes q from there, keeps the tile as
// q0 for the no-move test, overwrites the rows of accepted chains and writes the run back the same
// way.  Rows of rejected (or, in DR stage 0, not yet finished) chains go back unchanged.
template <typename T>
__device__ __forceinline__ void warp_tile_copy(T* tile, int rs, T* __restrict__ gmem, int d, int n_rows, int wlane, bool to_tile) {
  // batches of kU independent global accesses per lane: written as load-then-store per element the
  // compiler kept them in order (it cannot rule out aliasing) and every element paid a full memory latency
  constexpr int kU = 8;
  const int ne = n_rows * d;
  int row = 0, col = wlane;
  while (col >= d) { col -= d; ++row; }
  for (int f0 = wlane; f0 < ne; f0 += 32 * kU) {
    T v[kU];
    if (to_tile) {
#pragma unroll
      for (int u = 0; u < kU; ++u) v[u] = (f0 + 32 * u < ne) ? gmem[f0 + 32 * u] : T(0);
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      if (f0 + 32 * u < ne) {
        if (to_tile) tile[row * rs + col] = v[u];
        else gmem[f0 + 32 * u] = tile[row * rs + col];
      }
      col += 32;
      while (col >= d) { col -= d; ++row; }
    }
  }
}

// Everything of a single-trajectory transition (vanilla HMC step, or DR stage 0) after the
// chain's position q and refreshed momentum p are in registers: energies, trajectory, no-move
// test, accept rule, outputs, and for DR the hand-over of rejected chains to stage 1.  Shared by
// the plain kernel below and the warp-specialised one (k_transition_ws).
template <class Model, typename T, int EPL, int LPC, bool UNITM, int MODE>
__device__ __forceinline__ void finish_single(const Model& model, const TransArgs<T>& a, const GroupCtx<T>& c,
                                              const T (&minv)[UNITM ? 1 : EPL], int64_t chain, int64_t qoff,
                                              T (&q)[EPL], T (&p)[EPL], T* trow = nullptr, const T* q0s = nullptr,
                                              int q0sd = 0) {
  // trow: this chain's row of the warp tile (holds q0; takes the new q on acceptance), or nullptr when
  // the state is read from and written to global memory directly.  q0s: a read-only copy of q0 in shared memory,
  // element i at q0s[i*q0sd] (the warp-specialised kernel keeps one: the no-move test and the copy-through of
  // rejected chains then need no second strided read of q_in)
  using Lay = Layout<EPL, LPC>;
  const int d = a.d;
  int cH, cVg, cLeap;
    if (a.p0_out) store_vec<T, EPL, LPC>(p, a.p0_out, qoff, a.sd, c.lane, d);
    const T K0 = kinetic<T, EPL, LPC, UNITM>(p, minv, c.mask);
    T V0, V1;
    const int N = a.L[0] > 1 ? a.L[0] : 1;
    leapfrog<Model, T, EPL, LPC, UNITM>(model, q, p, minv, N, a.eps[0], c, V0, V1);
    const T H0 = V0 + K0;
    const T H1 = V1 + kinetic<T, EPL, LPC, UNITM>(p, minv, c.mask);
    cH = 2; cVg = N + 1; cLeap = 1;
    // sum(q0 - q1) == 0 is the reference's "did not move" test (src/pyhmc.py:57)
    T dq = trow ? diff_sum<T, EPL, LPC>(q, trow, 0, 1, c.lane, d)
                : (q0s ? diff_sum<T, EPL, LPC>(q, q0s, 0, q0sd, c.lane, d) : diff_sum<T, EPL, LPC>(q, a.q_in, qoff, a.sd, c.lane, d));
    dq = group_sum_masked<LPC, T>(dq, c.mask);
    T prob = t_exp<T>(H0 - H1);

    bool accept;
    int code;
    bool finished = true;
    if (MODE == MODE_HMC) {
      const bool bad = t_bad(prob) || dq == T(0);
      const double u = draw_uniform<T>(a, chain, 0);
      const T pm = prob < T(1) ? prob : T(1);
      accept = !bad && !(u > (double)pm);
      code = bad ? (a.rule == HMCB_RULE_PYHMC ? 2 : -1) : (accept ? 1 : 0);
      if (c.lane == 0) {
        if (a.energies) { a.energies[2 * chain] = H0; a.energies[2 * chain + 1] = H1; }
        if (a.accept_prob) a.accept_prob[chain] = (accept && !t_bad(prob)) ? pm : T(0);
      }
    } else {
      const T pf_raw = prob;
      if (dq == T(0)) prob = T(0);
      if (t_bad(prob)) prob = T(0);
      const T a0 = prob < T(1) ? prob : T(1);
      // probabilistic DR, first stage (src/algorithms.py:702-703, :790-791): from the raw energy ratio
      T pt0 = T(1);
      if (a.dr_prob && !t_bad(pf_raw)) pt0 = T(1) - (a.dr_prob == 1 ? (pf_raw < T(1) ? pf_raw : T(1)) : a0);
      const double u = draw_uniform<T>(a, chain, 0);
      // algorithms.DRHMC: acc < avec[0] (src/algorithms.py:528); PyHMC_multistep's first stage: acc <= avec[0]
      // (src/pyhmc.py:580) and codes j / -1 instead of j+1 / 0
      accept = a.dr_conv ? (u <= (double)a0) : (u < (double)a0);
      code = a.dr_conv ? (accept ? 0 : -1) : (accept ? 1 : 0);
      finished = accept || a.nprops == 1;
      if (c.lane == 0) {
        if (finished) {
          if (a.avec_out) {
            a.avec_out[chain * a.nprops] = a0;
            for (int i = 1; i < a.nprops; ++i) a.avec_out[chain * a.nprops + i] = T(0);
          }
          if (a.ptries_out) {
            a.ptries_out[chain * a.nprops] = pt0;
            for (int i = 1; i < a.nprops; ++i) a.ptries_out[chain * a.nprops + i] = T(1);
          }
        } else {
          a.ws.H0[chain] = H0;
          a.ws.avec[chain * kMaxProps] = a0;
          if (a.dr_prob) a.ws.ptries[chain * kMaxProps] = pt0;
          a.ws.counts[3 * chain] = cH; a.ws.counts[3 * chain + 1] = cVg; a.ws.counts[3 * chain + 2] = cLeap;
          // plain atomicAdd (the compiler aggregates it per warp): a cooperative-groups coalesced_threads()
          // here cost the trajectory loop its register allocation in the warp-specialised kernel
          // (3 spills + 25 moves per leapfrog step at EPL = 50)
          const int slot = atomicAdd(a.ws.count + 1, 1);
          a.ws.list[(int64_t)1 * a.n_chains + slot] = (int32_t)chain;
        }
      }
    }
    if (finished && c.lane == 0) {
      if (a.code) a.code[chain] = code;
      if (a.counts) { a.counts[3 * chain] = cH; a.counts[3 * chain + 1] = cVg; a.counts[3 * chain + 2] = cLeap; }
    }
    if (accept) {
      if (trow) store_vec<T, EPL, LPC>(q, trow, 0, 1, c.lane, d);
      else store_vec<T, EPL, LPC>(q, a.q_out, qoff, a.sd, c.lane, d);
      if (a.p_out) store_vec<T, EPL, LPC>(p, a.p_out, qoff, a.sd, c.lane, d);
    } else if (finished) {
      if (!trow && a.q_out != a.q_in) {
#pragma unroll
        for (int s = 0; s < EPL; ++s) {
          const int i = Lay::elem(c.lane, s);
          if (i < d) a.q_out[qoff + i * a.sd] = q0s ? q0s[i * q0sd] : a.q_in[qoff + i * a.sd];
        }
      }
      if (a.p_out) {
        draw_momentum<T, EPL, LPC, UNITM>(p, a, minv, chain, c.lane, d);
        store_vec<T, EPL, LPC>(p, a.p_out, qoff, a.sd, c.lane, d);
      }
    }
}

// ------------------------------------------------------------------ the kernel
// Threads are grouped LPC lanes per chain; a warp carries 32/LPC chains.  Work
// items (chains, or entries of the stage list) are walked grid-stride so a
// persistent grid covers any count, including one only known on the device.
template <class Model, typename T, int EPL, int LPC, bool UNITM, int MODE, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) k_transition(const Model model, const TransArgs<T> a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  using Lay = Layout<EPL, LPC>;
  constexpr int GPB = THREADS / LPC;
  const int d = a.d;

  model.block_init(smem_raw, THREADS);

  GroupCtx<T> c;
  c.lane = threadIdx.x & (LPC - 1);
  c.d = d;
  const int group_in_warp = (threadIdx.x & 31) / LPC;
  c.mask = (LPC == 32) ? 0xffffffffu : (((1u << LPC) - 1u) << (group_in_warp * LPC));
  c.smem_q = nullptr;
  if (Model::kNeedsSmem) {
    // staging area after the model's own shared block (see GaussianModel)
    T* base = (T*)smem_raw + model.p_smem_elems();
    c.smem_q = base + (size_t)(threadIdx.x / LPC) * d;
  }

  T minv[UNITM ? 1 : EPL];
  if (!UNITM) {
#pragma unroll
    for (int s = 0; s < EPL; ++s) {
      const int i = Lay::elem(c.lane, s);
      minv[UNITM ? 0 : s] = (i < d) ? a.minv[i] : T(1);
    }
  } else {
    minv[0] = T(1);
  }

  const int64_t gthread = (int64_t)blockIdx.x * THREADS + threadIdx.x;
  const int64_t total_groups = (int64_t)gridDim.x * GPB;
  const bool resume = (MODE == MODE_DRN) && a.dr_phase == 2;
  const int64_t count = (MODE == MODE_DRN) ? (int64_t)(resume ? min(a.ws.ccount[0], a.ws.ccap) : a.ws.count[a.stage]) : a.n_chains;
  const int32_t* list = (MODE == MODE_DRN) ? (resume ? a.ws.clist : a.ws.list + (int64_t)a.stage * a.n_chains) : nullptr;
  const int J = (MODE == MODE_DRN) ? a.stage : 0;

  if constexpr (MODE != MODE_DRN && !Model::kNeedsSmem) {
    if (a.tile_io && a.sd == 1 && a.sc == d) {
      // single-trajectory transition with the state staged through a warp tile (see warp_tile_copy)
      constexpr int CPW = 32 / LPC;                         // chains per warp
      const int rs = d | 1;
      const int wlane = threadIdx.x & 31;
      T* tile = reinterpret_cast<T*>(smem_raw) + (size_t)(threadIdx.x >> 5) * CPW * ((EPL * LPC) | 1);
      const int g = wlane / LPC;
      for (int64_t w0 = (gthread - wlane) / LPC; w0 < count; w0 += total_groups) {
        const int nv = (int)(count - w0 < CPW ? count - w0 : CPW);
        warp_tile_copy<T>(tile, rs, const_cast<T*>(a.q_in) + w0 * d, d, nv, wlane, true);
        __syncwarp();
        if (g < nv) {
          const int64_t chain = w0 + g;
          T q[EPL], p[EPL];
          T* trow = tile + g * rs;
          load_vec<T, EPL, LPC>(q, trow, 0, 1, c.lane, d, T(1));
          draw_momentum<T, EPL, LPC, UNITM>(p, a, minv, chain, c.lane, d);
          finish_single<Model, T, EPL, LPC, UNITM, MODE>(model, a, c, minv, chain, chain * a.sc, q, p, trow);
        }
        __syncwarp();
        warp_tile_copy<T>(tile, rs, a.q_out + w0 * d, d, nv, wlane, false);
        __syncwarp();
      }
      return;
    }
  }

  for (int64_t item = gthread / LPC; item < count; item += total_groups) {
    const int64_t chain = (MODE == MODE_DRN) ? (int64_t)list[item] : item;
    const int64_t qoff = chain * a.sc;

    T q[EPL], p[EPL];
    int cH, cVg, cLeap;

    if (MODE != MODE_DRN) {
      // ---------------- single trajectory from (q0, p0): HMC step or DR stage 0
      load_vec<T, EPL, LPC>(q, a.q_in, qoff, a.sd, c.lane, d, T(1));    // loads in flight under the RNG
      draw_momentum<T, EPL, LPC, UNITM>(p, a, minv, chain, c.lane, d);
      finish_single<Model, T, EPL, LPC, UNITM, MODE>(model, a, c, minv, chain, qoff, q, p);
    } else {
      // ---------------- DR stage J >= 1: explicit-stack walk of step/get_num
      // level 0 is DRHMC.step's own loop at j = J; level l >= 1 is a get_num frame
      // whose base state is the proposal stored in frame[l-1] with p flipped.
      int lm[kMaxProps], lj[kMaxProps];
      T lHb[kMaxProps], lpf[kMaxProps], lav[kMaxProps][kMaxProps];
      T lpt[kMaxProps][kMaxProps];   // try probabilities per frame (probabilistic DR; all 1 otherwise)
      const int prob_rule = a.dr_prob;
      int lvl = 0;
      lm[0] = J + 1; lj[0] = J; lHb[0] = a.ws.H0[chain];
#pragma unroll
      for (int i = 0; i < kMaxProps; ++i) {
        lav[0][i] = (i < J) ? a.ws.avec[chain * kMaxProps + i] : T(0);
        lpt[0][i] = (prob_rule && i < J) ? a.ws.ptries[chain * kMaxProps + i] : T(1);
      }
      cH = a.ws.counts[3 * chain]; cVg = a.ws.counts[3 * chain + 1]; cLeap = a.ws.counts[3 * chain + 2];
      const uint32_t it_add = a.iter_of ? (uint32_t)a.iter_of[chain] : 0u;   // the draws of this chain's running transition
      // the end of a chain's transition in the multi-pass sample mode: totals, and on to the next pass if it has more to do
      auto transition_done = [&](bool accepted) {
        if (a.iter_of && c.lane == 0) {
          if (a.n_accept && accepted) a.n_accept[chain] += 1;
          const int done = (int)it_add + 1;
          a.iter_done[chain] = done;
          if (done < a.K_total) a.cont_list[atomicAdd(a.cont_count, 1)] = (int32_t)chain;
        }
      };
      auto put_counts = [&]() {
        if (a.counts && c.lane == 0) {
          if (a.iter_of) { a.counts[3 * chain] += cH; a.counts[3 * chain + 1] += cVg; a.counts[3 * chain + 2] += cLeap; }
          else { a.counts[3 * chain] = cH; a.counts[3 * chain + 1] = cVg; a.counts[3 * chain + 2] = cLeap; }
        }
      };
      if (prob_rule && !resume) {
        // one more uniform decides whether stage J is attempted at all (src/algorithms.py:710-711): code -2 if not
        const double u_try = draw_uniform<T>(a, chain, 2 * J - 1, it_add);
        if (u_try > (double)lpt[0][J - 1]) {
          if (c.lane == 0) {
            for (int i = 0; i < a.nprops; ++i) {
              if (a.avec_out) a.avec_out[chain * a.nprops + i] = lav[0][i];
              if (a.ptries_out) a.ptries_out[chain * a.nprops + i] = lpt[0][i];
            }
            if (a.code) a.code[chain] = -2;
          }
          put_counts();
          transition_done(false);
          if (a.q_out != a.q_in) {
#pragma unroll
            for (int s = 0; s < EPL; ++s) {
              const int i = Lay::elem(c.lane, s);
              if (i < d) a.q_out[qoff + i * a.sd] = a.q_in[qoff + i * a.sd];
            }
          }
          if (a.p_out) {
            draw_momentum<T, EPL, LPC, UNITM>(p, a, minv, chain, c.lane, d, it_add);
            store_vec<T, EPL, LPC>(p, a.p_out, qoff, a.sd, c.lane, d);
          }
          continue;
        }
      }
      const int64_t FT = a.ws.frame_threads;
      T* fr = a.ws.frames + gthread;  // frame f: q at (2f)*EPL*FT, p at (2f+1)*EPL*FT, slot stride FT
      T prob_top = T(0);
      bool have_qp = false, parked = false;
      if (resume) {
        // a chain parked by the first part of this stage: its walk stands in the ghost frame of the stage's proposal
        // (level 1), first ghost trajectory done, the next one (class 1) due; the proposal goes into this thread's frame 0
        const int64_t cc = a.ws.ccap, k = item;
        lvl = 1;
        lm[1] = J; lj[1] = 1;
        cH = a.ws.cst_i[k]; cVg = a.ws.cst_i[cc + k]; cLeap = a.ws.cst_i[2 * cc + k];
        lHb[1] = a.ws.cst_f[k]; lpf[0] = a.ws.cst_f[cc + k];
#pragma unroll
        for (int i = 0; i < kMaxProps; ++i) { lav[1][i] = T(0); lpt[1][i] = T(1); }
        lav[1][0] = a.ws.cst_f[2 * cc + k];
        if (prob_rule) { lpt[0][J] = a.ws.cst_f[3 * cc + k]; lpt[1][0] = a.ws.cst_f[4 * cc + k]; }
#pragma unroll
        for (int s = 0; s < EPL; ++s) { q[s] = a.ws.cframes[(int64_t)s * cc * LPC + k * LPC + c.lane]; p[s] = a.ws.cframes[(int64_t)(EPL + s) * cc * LPC + k * LPC + c.lane]; }
#pragma unroll
        for (int s = 0; s < EPL; ++s) { fr[(int64_t)s * FT] = q[s]; fr[(int64_t)(EPL + s) * FT] = p[s]; p[s] = -p[s]; }
        have_qp = true;
      }

      for (;;) {
        const int j = lj[lvl];
        const int N = a.L[j] > 1 ? a.L[j] : 1;
        if (have_qp) {
          have_qp = false;      // (q, -p) of the proposal are in registers already
        } else if (lvl == 0) {
          load_vec<T, EPL, LPC>(q, a.q_in, qoff, a.sd, c.lane, d, T(1));
          draw_momentum<T, EPL, LPC, UNITM>(p, a, minv, chain, c.lane, d, it_add);
        } else {
          const T* fq = fr + (int64_t)(2 * (lvl - 1)) * EPL * FT;
#pragma unroll
          for (int s = 0; s < EPL; ++s) { q[s] = fq[(int64_t)s * FT]; p[s] = -fq[(int64_t)(EPL + s) * FT]; }
        }
        T Vs, Ve;
        leapfrog<Model, T, EPL, LPC, UNITM>(model, q, p, minv, N, a.eps[j], c, Vs, Ve);
        const T Hn = Ve + kinetic<T, EPL, LPC, UNITM>(p, minv, c.mask);
        cH += 1; cVg += N + 1; cLeap += 1;
        T dq = T(0);
        if (lvl == 0) {
          dq = diff_sum<T, EPL, LPC>(q, a.q_in, qoff, a.sd, c.lane, d);
        } else {
          const T* fq = fr + (int64_t)(2 * (lvl - 1)) * EPL * FT;
#pragma unroll
          for (int s = 0; s < EPL; ++s) dq += fq[(int64_t)s * FT] - q[s];
        }
        dq = group_sum_masked<LPC, T>(dq, c.mask);
        T pf = t_exp<T>(lHb[lvl] - Hn);
        // rule 1 inside a ghost frame takes p_try from the raw ratio (src/algorithms.py:668-669), at level 0
        // from the ratio after the no-move test (:714-719)
        if (prob_rule == 1 && lvl > 0) lpt[lvl][j] = t_bad(pf) ? T(1) : T(1) - (pf < T(1) ? pf : T(1));
        if (dq == T(0)) pf = T(0);
        if (prob_rule == 1 && lvl == 0) lpt[0][j] = t_bad(pf) ? T(1) : T(1) - (pf < T(1) ? pf : T(1));

        if (j > 0) {
          // the ghost chain from the flipped proposal is needed: push a get_num frame
          T* fq = fr + (int64_t)(2 * lvl) * EPL * FT;
#pragma unroll
          for (int s = 0; s < EPL; ++s) { fq[(int64_t)s * FT] = q[s]; fq[(int64_t)(EPL + s) * FT] = p[s]; }
          lpf[lvl] = pf;
          ++lvl;
          if (a.dr_conv) cH += 1;   // PyHMC_multistep.get_num re-evaluates H at its base (src/pyhmc.py:541)
          lm[lvl] = j; lj[lvl] = 0; lHb[lvl] = Hn;
#pragma unroll
          for (int i = 0; i < kMaxProps; ++i) { lav[lvl][i] = T(0); lpt[lvl][i] = T(1); }
          continue;
        }
        // j == 0 can only happen inside a ghost frame here (J >= 1)
        T prob = pf;
        T pf_cur = pf;     // the energy ratio (after the no-move test) of the stage whose alpha is being set
        bool done = false;
        for (;;) {
          T ret;
          bool returning;
          if (t_bad(prob)) {
            ret = T(0);
            returning = true;
          } else {
            const T al = prob < T(1) ? prob : T(1);
            lav[lvl][lj[lvl]] = al;
            if (prob_rule == 2) lpt[lvl][lj[lvl]] = t_bad(pf_cur) ? T(1) : T(1) - al;   // src/algorithms.py:764-765
            T pr = T(1);
            for (int i = 0; i < lm[lvl]; ++i) pr *= (T(1) - lav[lvl][i]);
            if (pr == T(0)) { ret = pr; returning = true; }
            else if (lj[lvl] + 1 == lm[lvl]) {
              if (prob_rule) {
                T pt = T(1);
                for (int i = 0; i < lm[lvl]; ++i) pt *= lpt[lvl][i];
                pr *= pt;                                                              // prod(1-avec) * prod(ptries)
              }
              ret = pr; returning = true;
            }
            else { lj[lvl] += 1; returning = false; }
          }
          if (!returning) break;
          --lvl;
          T den = T(1);
          for (int i = 0; i < lj[lvl]; ++i) den *= (T(1) - lav[lvl][i]);
          if (prob_rule) {
            T pt = T(1);
            for (int i = 0; i < lj[lvl]; ++i) pt *= lpt[lvl][i];
            den *= pt;
          }
          prob = lpf[lvl] * ret / den;
          pf_cur = lpf[lvl];
          if (lvl == 0) { prob_top = prob; done = true; break; }
        }
        if (done) break;
        if (a.dr_phase == 1 && lvl == 1 && lj[1] == 1) {
          // first point where the chains of a warp part ways: some returned to level 0 above (the first ghost stage accepts
          // for sure, or J == 1), this one's ghost recursion goes on.  Park it for the stage's second kernel, where only such
          // chains meet, instead of holding the lanes of those that are done (stage 2 of the d=50 funnel ran 26.9 of 32 lanes).
          int slot = 0;
          if (c.lane == 0) slot = atomicAdd(a.ws.ccount, 1);
          slot = __shfl_sync(c.mask, slot, ((threadIdx.x & 31) / LPC) * LPC);
          if (slot < a.ws.ccap) {
            const int64_t cc = a.ws.ccap;
            if (c.lane == 0) {
              a.ws.clist[slot] = (int32_t)chain;
              a.ws.cst_i[slot] = cH; a.ws.cst_i[cc + slot] = cVg; a.ws.cst_i[2 * cc + slot] = cLeap;
              a.ws.cst_f[slot] = lHb[1]; a.ws.cst_f[cc + slot] = lpf[0]; a.ws.cst_f[2 * cc + slot] = lav[1][0];
              if (prob_rule) { a.ws.cst_f[3 * cc + slot] = lpt[0][J]; a.ws.cst_f[4 * cc + slot] = lpt[1][0]; }
            }
#pragma unroll
            for (int s = 0; s < EPL; ++s) { q[s] = fr[(int64_t)s * FT]; p[s] = fr[(int64_t)(EPL + s) * FT]; }
#pragma unroll
            for (int s = 0; s < EPL; ++s) {
              a.ws.cframes[(int64_t)s * cc * LPC + (int64_t)slot * LPC + c.lane] = q[s];
              a.ws.cframes[(int64_t)(EPL + s) * cc * LPC + (int64_t)slot * LPC + c.lane] = p[s];
            }
            parked = true;
            break;
          }
          // the list is full: this chain finishes its stage here, as without the split
        }
      }
      if (parked) continue;

      // ---------------- DRHMC.step's accept test for stage J (src/algorithms.py:523-529)
      if (t_bad(prob_top)) prob_top = T(0);
      const T aJ = prob_top < T(1) ? prob_top : T(1);
      if (prob_rule == 2) lpt[0][J] = t_bad(lpf[0]) ? T(1) : T(1) - aJ;                 // src/algorithms.py:808-809
      const double u = draw_uniform<T>(a, chain, prob_rule ? 2 * J : J, it_add);
      const bool accept = u < (double)aJ;
      const bool finished = accept || (J == a.nprops - 1);
      if (c.lane == 0) {
        if (finished) {
          if (a.avec_out) {
            for (int i = 0; i < a.nprops; ++i)
              a.avec_out[chain * a.nprops + i] = (i < J) ? lav[0][i] : (i == J ? aJ : T(0));
          }
          if (a.ptries_out) {
            for (int i = 0; i < a.nprops; ++i) a.ptries_out[chain * a.nprops + i] = lpt[0][i];
          }
          if (a.code) a.code[chain] = a.dr_conv ? (accept ? J : -1) : (accept ? J + 1 : 0);
        } else {
          a.ws.avec[chain * kMaxProps + J] = aJ;
          if (prob_rule) a.ws.ptries[chain * kMaxProps + J] = lpt[0][J];
          a.ws.counts[3 * chain] = cH; a.ws.counts[3 * chain + 1] = cVg; a.ws.counts[3 * chain + 2] = cLeap;
          const int slot = atomicAdd(a.ws.count + J + 1, 1);   // aggregated per warp by the compiler
          a.ws.list[(int64_t)(J + 1) * a.n_chains + slot] = (int32_t)chain;
        }
      }
      if (finished) {
        put_counts();
        transition_done(accept);
        if (accept) {
          // the stage-J proposal was parked in frame 0
          const T* fq = fr;
#pragma unroll
          for (int s = 0; s < EPL; ++s) { q[s] = fq[(int64_t)s * FT]; p[s] = fq[(int64_t)(EPL + s) * FT]; }
          store_vec<T, EPL, LPC>(q, a.q_out, qoff, a.sd, c.lane, d);
          if (a.p_out) store_vec<T, EPL, LPC>(p, a.p_out, qoff, a.sd, c.lane, d);
        } else {
          if (a.q_out != a.q_in) {
#pragma unroll
            for (int s = 0; s < EPL; ++s) {
              const int i = Lay::elem(c.lane, s);
              if (i < d) a.q_out[qoff + i * a.sd] = a.q_in[qoff + i * a.sd];
            }
          }
          if (a.p_out) {
            draw_momentum<T, EPL, LPC, UNITM>(p, a, minv, chain, c.lane, d, it_add);
            store_vec<T, EPL, LPC>(p, a.p_out, qoff, a.sd, c.lane, d);
          }
        }
      }
    }
  }
}

// ------------------------------------------------------------------ K HMC transitions per launch, any shape
// The loop of HMC.sample (src/algorithms.py:244-250) for one chain per group of lanes: the position stays in registers
// over the K transitions, the start point of the running transition (no-move test, come-back on rejection) in shared
// memory, slot s of thread t at sQ[s*THREADS + t].  The warp-specialised form of the same loop for float64
// thread-per-chain shapes is k_hmc_multi_ws (transition_ws.cuh).  Transition k draws with iteration a.iteration + k:
// bit-identical to K launches of k_transition<MODE_HMC>.
template <class Model, typename T, int EPL, int LPC, bool UNITM, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) k_hmc_multi(const Model model, const TransArgs<T> a, const int K,
                                                             int32_t* __restrict__ n_accept) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  using Lay = Layout<EPL, LPC>;
  constexpr int GPB = THREADS / LPC;
  const int d = a.d;
  T* sQ = reinterpret_cast<T*>(smem_raw) + threadIdx.x;
  GroupCtx<T> c;
  c.lane = threadIdx.x & (LPC - 1);
  c.d = d;
  const int group_in_warp = (threadIdx.x & 31) / LPC;
  c.mask = (LPC == 32) ? 0xffffffffu : (((1u << LPC) - 1u) << (group_in_warp * LPC));
  c.smem_q = nullptr;
  T minv[UNITM ? 1 : EPL];
  if (!UNITM) {
#pragma unroll
    for (int s = 0; s < EPL; ++s) {
      const int i = Lay::elem(c.lane, s);
      minv[UNITM ? 0 : s] = (i < d) ? a.minv[i] : T(1);
    }
  } else {
    minv[0] = T(1);
  }
  const int N = a.L[0] > 1 ? a.L[0] : 1;
  const int64_t total_groups = (int64_t)gridDim.x * GPB;
  for (int64_t chain = ((int64_t)blockIdx.x * THREADS + threadIdx.x) / LPC; chain < a.n_chains; chain += total_groups) {
    const int64_t qoff = chain * a.sc;
    T q[EPL], p[EPL];
    load_vec<T, EPL, LPC>(q, a.q_in, qoff, a.sd, c.lane, d, T(1));
#pragma unroll
    for (int s = 0; s < EPL; ++s) sQ[s * THREADS] = q[s];
    int nacc = 0;
    for (int k = 0; k < K; ++k) {
      draw_momentum<T, EPL, LPC, UNITM>(p, a, minv, chain, c.lane, d, (uint32_t)k);
      const T K0 = kinetic<T, EPL, LPC, UNITM>(p, minv, c.mask);
      T V0, V1;
      leapfrog<Model, T, EPL, LPC, UNITM>(model, q, p, minv, N, a.eps[0], c, V0, V1);
      const T H0 = V0 + K0;
      const T H1 = V1 + kinetic<T, EPL, LPC, UNITM>(p, minv, c.mask);
      // sum(q0 - q1) with diff_sum's partial sums (padded slots hold 0 on both sides)
      T acc[4] = {T(0), T(0), T(0), T(0)};
#pragma unroll
      for (int s0 = 0; s0 < EPL; s0 += 8) {
#pragma unroll
        for (int e = 0; e < 8; ++e)
          if (s0 + e < EPL) acc[e & 3] += sQ[(s0 + e) * THREADS] - q[s0 + e];
      }
      const T dq = group_sum_masked<LPC, T>((acc[0] + acc[1]) + (acc[2] + acc[3]), c.mask);
      const T prob = t_exp<T>(H0 - H1);
      const bool bad = t_bad(prob) || dq == T(0);
      const double u = draw_uniform<T>(a, chain, 0, (uint32_t)k);
      const T pm = prob < T(1) ? prob : T(1);
      const bool accept = !bad && !(u > (double)pm);
      if (k == K - 1) {
        if (c.lane == 0) {
          if (a.code) a.code[chain] = bad ? (a.rule == HMCB_RULE_PYHMC ? 2 : -1) : (accept ? 1 : 0);
          if (a.energies) { a.energies[2 * chain] = H0; a.energies[2 * chain + 1] = H1; }
          if (a.accept_prob) a.accept_prob[chain] = (accept && !t_bad(prob)) ? pm : T(0);
        }
        if (a.p_out) {
          if (!accept) draw_momentum<T, EPL, LPC, UNITM>(p, a, minv, chain, c.lane, d, (uint32_t)k);
          store_vec<T, EPL, LPC>(p, a.p_out, qoff, a.sd, c.lane, d);
        }
      }
      if (accept) {
        ++nacc;
#pragma unroll
        for (int s = 0; s < EPL; ++s) sQ[s * THREADS] = q[s];
      } else {
#pragma unroll
        for (int s = 0; s < EPL; ++s) q[s] = sQ[s * THREADS];
      }
    }
    store_vec<T, EPL, LPC>(q, a.q_out, qoff, a.sd, c.lane, d);
    if (c.lane == 0) {
      if (n_accept) n_accept[chain] += nacc;
      if (a.counts) { a.counts[3 * chain] = 2 * K; a.counts[3 * chain + 1] = K * (N + 1); a.counts[3 * chain + 2] = K; }
    }
  }
}

}  // namespace hmcb
