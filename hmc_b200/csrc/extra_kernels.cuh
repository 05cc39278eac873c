// extra_kernels.cuh -- MALA (src/mala.py) and AdHMC_eps0 (src/adaptivehmc.py:105-192)
// transitions, built from the same register-resident pieces as transition.cuh.
#pragma once
#include "transition.cuh"

namespace hmcb {

// ------------------------------------------------------------------ MALA
// One step (src/mala.py:45-60,24-42): z ~ N(0,I); q1 = q - eps^2/2 grad V(q) + eps z;
// prob = exp(V(q)-V(q1)) * exp(lognum - logden) with the Gaussian proposal densities;
// code 2 on NaN/inf/no-move, else reject iff u > min(1,prob).  V(q), grad V(q) are
// carried between steps (the reference caches them, :49-52,60): 1 gradient per step.
// n_steps steps are fused in one launch with q, grad V(q), q1, grad V(q1) in registers,
// so the kernel is FP64-bound instead of paying 32d bytes of HBM per gradient.
template <typename T>
struct MalaArgs {
  int64_t n_chains, chain_offset;
  int32_t d;
  const T* q_in;
  T* q_out;
  int64_t sc, sd;
  uint64_t seed;
  uint32_t iteration;
  int32_t rng_flags;
  const T* z_inject;       // [C,d] layout of q (only meaningful for n_steps == 1)
  const double* u_inject;  // [C,u_stride]
  int32_t u_stride;
  T eps;
  int32_t n_steps;
  const T* V_in;           // [C] cached V(q) or nullptr
  const T* Vg_in;          // layout of q, cached grad V(q), or nullptr
  T* V_out;
  T* Vg_out;
  int32_t* code;           // [C] code of the last step
  int32_t* n_accept;       // [C] accepted steps among n_steps
  int32_t* counts;         // [C,3] Hcount(0), Vgcount, leapcount(0)
};

template <class Model, typename T, int EPL, int LPC, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) k_mala(const Model model, const MalaArgs<T> a) {
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
  if (Model::kNeedsSmem) c.smem_q = (T*)smem_raw + model.p_smem_elems() + (size_t)(threadIdx.x / LPC) * d;

  const T half_e2 = T(0.5) * a.eps * a.eps;  // 0.5*eps**2
  const T e2 = a.eps * a.eps;
  const int64_t total_groups = (int64_t)gridDim.x * GPB;
  for (int64_t chain = ((int64_t)blockIdx.x * THREADS + threadIdx.x) / LPC; chain < a.n_chains; chain += total_groups) {
    const int64_t off = chain * a.sc;
    const uint64_t gchain = (uint64_t)(a.chain_offset + chain);
    T q[EPL], g[EPL], q1[EPL], g1[EPL];
    load_vec<T, EPL, LPC>(q, a.q_in, off, a.sd, c.lane, d, T(1));
    int nvg = 0, nacc = 0, code = 0;
    // V_q and V_gq are cached independently (src/mala.py:49-52): either may be supplied without the other
    T Vq = a.V_in ? a.V_in[chain] : T(0);
    if (a.Vg_in) load_vec<T, EPL, LPC>(g, a.Vg_in, off, a.sd, c.lane, d, T(1));
    if (!a.Vg_in || !a.V_in) {
      T tmp[EPL];
#pragma unroll
      for (int s = 0; s < EPL; ++s) tmp[s] = T(0);
      const T V = model.template kick<EPL, LPC, true>(q, tmp, T(-1), c);  // tmp = +grad V
      if (!a.Vg_in) {
#pragma unroll
        for (int s = 0; s < EPL; ++s) g[s] = tmp[s];
        nvg += 1;
      }
      if (!a.V_in) Vq = V;
    }
    for (int k = 0; k < a.n_steps; ++k) {
      // proposal
      if (a.z_inject) {
        load_vec<T, EPL, LPC>(q1, a.z_inject, off, a.sd, c.lane, d, T(1));
      } else {
        fill_normals<T, EPL, LPC>(q1, a.seed, gchain, a.iteration + (uint32_t)k, c.lane, d, a.rng_flags);
      }
#pragma unroll
      for (int s = 0; s < EPL; ++s) {
        q1[s] = (q[s] - half_e2 * g[s]) + a.eps * q1[s];
        g1[s] = T(0);
      }
      const T V1 = model.template kick<EPL, LPC, true>(q1, g1, T(-1), c);
      nvg += 1;
      // Hastings ratio exactly as the reference forms it (src/mala.py:24-30)
      T ln = T(0), ld = T(0), dq = T(0);
#pragma unroll
      for (int s = 0; s < EPL; ++s) {
        const T mu0 = q[s] - half_e2 * g[s];
        const T mu1 = q1[s] - half_e2 * g1[s];
        const T a0 = q1[s] - mu0, a1 = q[s] - mu1;
        ld = fma(a0, a0, ld);
        ln = fma(a1, a1, ln);
        dq += q[s] - q1[s];
      }
      ld = group_sum_masked<LPC, T>(ld, c.mask);
      ln = group_sum_masked<LPC, T>(ln, c.mask);
      dq = group_sum_masked<LPC, T>(dq, c.mask);
      const T logden = T(-0.5) * ld / e2, lognum = T(-0.5) * ln / e2;
      const T prob = t_exp<T>(Vq - V1) * t_exp<T>(lognum - logden);
      const bool bad = t_bad(prob) || dq == T(0);
      bool accept = false;
      if (!bad) {
        double u;
        if (a.u_inject) {
          u = a.u_inject[chain * a.u_stride + k];
        } else {
          uint32_t w[4];
          Philox::gen(a.seed, gchain, a.iteration + (uint32_t)k, STREAM_ACCEPT, 0, w);
          u = u53(w[0], w[1]);
        }
        const T pm = prob < T(1) ? prob : T(1);
        accept = !(u > (double)pm);
      }
      code = bad ? 2 : (accept ? 1 : 0);
      if (accept) {
#pragma unroll
        for (int s = 0; s < EPL; ++s) { q[s] = q1[s]; g[s] = g1[s]; }
        Vq = V1;
        nacc += 1;
      }
    }
    store_vec<T, EPL, LPC>(q, a.q_out, off, a.sd, c.lane, d);
    if (a.Vg_out) store_vec<T, EPL, LPC>(g, a.Vg_out, off, a.sd, c.lane, d);
    if (c.lane == 0) {
      if (a.V_out) a.V_out[chain] = Vq;
      if (a.code) a.code[chain] = code;
      if (a.n_accept) a.n_accept[chain] = nacc;
      if (a.counts) { a.counts[3 * chain] = 0; a.counts[3 * chain + 1] = nvg; a.counts[3 * chain + 2] = 0; }
    }
  }
}

// ------------------------------------------------------------------ GHMC / Persistent GHMC
// algorithms.GHMC.step and Persistent_GHMC.step (src/algorithms.py:265-281, 314-333): partial
// momentum refresh, V and grad V cached between transitions (leapfrog_Vgq, :150-172: the first
// half kick reuses the cached gradient, the last gradient comes with the potential), momentum
// flipped unless accepted; the persistent variant carries its Metropolis uniform.
template <typename T>
struct GhmcArgs {
  TransArgs<T> t;          // chains, rng, p_prev/alpha/beta, minv, L[0], eps[0], code, energies, counts, p_out
  const T* V_in;           // [C] cached V(q) or nullptr
  const T* Vg_in;          // layout of q, cached grad V(q), or nullptr
  T* V_out;
  T* Vg_out;
  const double* pu_in;     // [C] persistent uniform, or nullptr (plain GHMC)
  double* pu_out;
  double delta;
  // hmcb_ghmc_sample: K transitions (k_ghmc_multi), or launch by launch with counts summed (accumulate) -- n_accept[C] += accepted
  int32_t K;
  int32_t accumulate;
  int32_t* n_accept;
  int32_t applied;         // set by the launcher: this shape has k_ghmc_multi
};

template <class Model, typename T, int EPL, int LPC, bool UNITM, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) k_ghmc(const Model model, const GhmcArgs<T> ga) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  using Lay = Layout<EPL, LPC>;
  constexpr int GPB = THREADS / LPC;
  const TransArgs<T>& a = ga.t;
  const int d = a.d;
  model.block_init(smem_raw, THREADS);
  GroupCtx<T> c;
  c.lane = threadIdx.x & (LPC - 1);
  c.d = d;
  const int group_in_warp = (threadIdx.x & 31) / LPC;
  c.mask = (LPC == 32) ? 0xffffffffu : (((1u << LPC) - 1u) << (group_in_warp * LPC));
  c.smem_q = nullptr;
  if (Model::kNeedsSmem) c.smem_q = (T*)smem_raw + model.p_smem_elems() + (size_t)(threadIdx.x / LPC) * d;
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
  const int64_t total_groups = (int64_t)gridDim.x * GPB;
  const bool cached = ga.Vg_in != nullptr;
  const int N = a.L[0] > 1 ? a.L[0] : 1;
  const T eps = a.eps[0], heps = T(0.5) * a.eps[0];
  for (int64_t chain = ((int64_t)blockIdx.x * THREADS + threadIdx.x) / LPC; chain < a.n_chains; chain += total_groups) {
    const int64_t off = chain * a.sc;
    T q[EPL], p[EPL], g[EPL];
    load_vec<T, EPL, LPC>(q, a.q_in, off, a.sd, c.lane, d, T(1));
    draw_momentum<T, EPL, LPC, UNITM>(p, a, minv, chain, c.lane, d);     // includes the alpha-mix with p_prev
    const T K0 = kinetic<T, EPL, LPC, UNITM>(p, minv, c.mask);
    int cH = 0, cVg = 0;
    // Per-chain cache state.  The reference keeps V_q = V_gq = None until a proposal is accepted
    // (src/algorithms.py:275-276); in a batch that is a NaN in V for exactly those chains.
    T V0 = T(0);
    bool has = cached;
    if (cached && ga.V_in) {
      V0 = ga.V_in[chain];
      if (isnan(V0)) has = false;
    }
    const bool had_cache = has;
    if (has) {
      load_vec<T, EPL, LPC>(g, ga.Vg_in, off, a.sd, c.lane, d, T(1));
    }
    if (!has || !ga.V_in) {
      T tmp[EPL];
#pragma unroll
      for (int s = 0; s < EPL; ++s) tmp[s] = T(0);
      const T V = model.template kick<EPL, LPC, true>(q, tmp, T(-1), c);
      if (!has) {
#pragma unroll
        for (int s = 0; s < EPL; ++s) g[s] = tmp[s];
        cVg += 1;
      }
      if (!has || !ga.V_in) { V0 = V; cH += 1; }
    }
#pragma unroll
    for (int s = 0; s < EPL; ++s) p[s] = fma(-heps, g[s], p[s]);
#pragma unroll 1
    for (int i = 1; i < N; ++i) {
      drift<T, EPL, UNITM>(q, p, minv, eps);
      model.template kick<EPL, LPC, false>(q, p, eps, c);
    }
    cVg += N - 1;
    drift<T, EPL, UNITM>(q, p, minv, eps);
#pragma unroll
    for (int s = 0; s < EPL; ++s) g[s] = T(0);
    const T V1 = model.template kick<EPL, LPC, true>(q, g, T(-1), c);      // g = grad V(q1), V1 = V(q1)
    cVg += 1;
#pragma unroll
    for (int s = 0; s < EPL; ++s) p[s] = fma(-heps, g[s], p[s]);
    const T H0 = V0 + K0, H1 = V1 + kinetic<T, EPL, LPC, UNITM>(p, minv, c.mask);
    T dq = diff_sum<T, EPL, LPC>(q, a.q_in, off, a.sd, c.lane, d);
    dq = group_sum_masked<LPC, T>(dq, c.mask);
    const T prob = t_exp<T>(H0 - H1);
    const bool bad = t_bad(prob) || dq == T(0);
    double u;
    if (ga.pu_in) {
      const double w = ga.pu_in[chain] + 1.0 + ga.delta;      // u = (u + 1 + delta) % 2 - 1, Python's sign convention
      double r = fmod(w, 2.0);
      if (r < 0.0) r += 2.0;
      u = r - 1.0;
    } else {
      u = draw_uniform<T>(a, chain, 0);
    }
    const T pm = prob < T(1) ? prob : T(1);
    const bool accept = !bad && !(u > (double)pm);
    if (c.lane == 0) {
      if (a.code) a.code[chain] = bad ? -1 : (accept ? 1 : 0);
      if (a.energies) { a.energies[2 * chain] = H0; a.energies[2 * chain + 1] = H1; }
      if (a.counts) {
        if (ga.accumulate) { a.counts[3 * chain] += cH; a.counts[3 * chain + 1] += cVg; a.counts[3 * chain + 2] += 1; }
        else { a.counts[3 * chain] = cH; a.counts[3 * chain + 1] = cVg; a.counts[3 * chain + 2] = 1; }
      }
      if (ga.n_accept) ga.n_accept[chain] += accept ? 1 : 0;
      if (ga.V_out) ga.V_out[chain] = accept ? V1 : (had_cache ? V0 : (T)NAN);
      if (ga.pu_out) ga.pu_out[chain] = accept ? u / (double)prob : u;
    }
    if (accept) {
      store_vec<T, EPL, LPC>(q, a.q_out, off, a.sd, c.lane, d);
      if (a.p_out) store_vec<T, EPL, LPC>(p, a.p_out, off, a.sd, c.lane, d);
      if (ga.Vg_out) store_vec<T, EPL, LPC>(g, ga.Vg_out, off, a.sd, c.lane, d);
    } else {
      if (a.q_out != a.q_in) {
        load_vec<T, EPL, LPC>(q, a.q_in, off, a.sd, c.lane, d, T(1));
        store_vec<T, EPL, LPC>(q, a.q_out, off, a.sd, c.lane, d);
      }
      if (a.p_out) {
        draw_momentum<T, EPL, LPC, UNITM>(p, a, minv, chain, c.lane, d);   // the refreshed momentum, flipped
#pragma unroll
        for (int s = 0; s < EPL; ++s) p[s] = -p[s];
        store_vec<T, EPL, LPC>(p, a.p_out, off, a.sd, c.lane, d);
      }
      if (ga.Vg_out) {
        if (had_cache) {
          if (ga.Vg_out != ga.Vg_in) {
            load_vec<T, EPL, LPC>(g, ga.Vg_in, off, a.sd, c.lane, d, T(1));
            store_vec<T, EPL, LPC>(g, ga.Vg_out, off, a.sd, c.lane, d);
          }
        } else {
#pragma unroll
          for (int s = 0; s < EPL; ++s) g[s] = (T)NAN;      // still "None" for this chain
          store_vec<T, EPL, LPC>(g, ga.Vg_out, off, a.sd, c.lane, d);
        }
      }
    }
  }
}

// ------------------------------------------------------------------ K GHMC transitions per launch
// The loop of GHMC.sample / Persistent_GHMC.sample (src/algorithms.py:284-303, :336-357) for one chain per group of
// lanes: q, the momentum and the cached gradient stay in registers over the K transitions (with nleap = 1 a single
// launch moves six state arrays per gradient -- HBM-bound like MALA); what a rejection comes back to -- the
// start point, the refreshed momentum (returned flipped) and the cached gradient -- sits in shared memory, slot s of
// thread t at s*THREADS + t.  Transition k draws with iteration a.iteration + k and follows k_ghmc's per-chain cache
// rules (V = grad V = None until the first acceptance: recomputed and counted every time), so the results -- state,
// caches, persistent uniform, the last transition's code and energies, the summed counters -- are bit-identical to K
// launches of k_ghmc.
template <class Model, typename T, int EPL, int LPC, bool UNITM, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) k_ghmc_multi(const Model model, const GhmcArgs<T> ga) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  using Lay = Layout<EPL, LPC>;
  constexpr int GPB = THREADS / LPC;
  const TransArgs<T>& a = ga.t;
  const int d = a.d;
  T* sQ = reinterpret_cast<T*>(smem_raw) + threadIdx.x;
  T* sP = sQ + EPL * THREADS;
  T* sG = sP + EPL * THREADS;
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
  const int64_t total_groups = (int64_t)gridDim.x * GPB;
  const int N = a.L[0] > 1 ? a.L[0] : 1;
  const int K = ga.K;
  const T eps = a.eps[0], heps = T(0.5) * a.eps[0];
  for (int64_t chain = ((int64_t)blockIdx.x * THREADS + threadIdx.x) / LPC; chain < a.n_chains; chain += total_groups) {
    const int64_t off = chain * a.sc;
    T q[EPL], p[EPL], g[EPL];
    load_vec<T, EPL, LPC>(q, a.q_in, off, a.sd, c.lane, d, T(1));
#pragma unroll
    for (int s = 0; s < EPL; ++s) sQ[s * THREADS] = q[s];
    {
      T pv[EPL];
      load_vec<T, EPL, LPC>(pv, a.p_prev, off, a.sd, c.lane, d, T(1));
#pragma unroll
      for (int s = 0; s < EPL; ++s) sP[s * THREADS] = pv[s];      // the carried momentum lives in sP between transitions
    }
    int cH = 0, cVg = 0, nacc = 0;
    T V0 = T(0);
    bool has = ga.Vg_in != nullptr;
    if (has && ga.V_in) {
      V0 = ga.V_in[chain];
      if (isnan(V0)) has = false;
    }
    if (has) {
      load_vec<T, EPL, LPC>(g, ga.Vg_in, off, a.sd, c.lane, d, T(1));
      if (!ga.V_in) {                       // a cached gradient without its potential: one H evaluation, once
        T tmp[EPL];
#pragma unroll
        for (int s = 0; s < EPL; ++s) tmp[s] = T(0);
        V0 = model.template kick<EPL, LPC, true>(q, tmp, T(-1), c);
        cH += 1;
      }
    }
    double pu = ga.pu_in ? ga.pu_in[chain] : 0.0;
#pragma unroll 1
    for (int k = 0; k < K; ++k) {
      // p = alpha * p_carried + beta * pn (draw_momentum's arithmetic, the carried momentum read from sP)
      fill_normals<T, EPL, LPC>(p, a.seed, (uint64_t)(a.chain_offset + chain), a.iteration + (uint32_t)k, c.lane, d, a.rng_flags);
      if (!UNITM) {
#pragma unroll
        for (int s = 0; s < EPL; ++s) p[s] *= t_rsqrt<T>(minv[UNITM ? 0 : s]);
      }
#pragma unroll
      for (int s = 0; s < EPL; ++s) {
        p[s] = mix_momentum(a.alpha, sP[s * THREADS], a.beta, p[s]);
        sP[s * THREADS] = p[s];             // now the refreshed momentum: what a rejection returns, flipped
      }
      const T K0 = kinetic<T, EPL, LPC, UNITM>(p, minv, c.mask);
      const bool had_cache = has;
      if (!has) {
#pragma unroll
        for (int s = 0; s < EPL; ++s) g[s] = T(0);
        V0 = model.template kick<EPL, LPC, true>(q, g, T(-1), c);
        cVg += 1;
        cH += 1;
      } else {
#pragma unroll
        for (int s = 0; s < EPL; ++s) sG[s * THREADS] = g[s];
      }
#pragma unroll
      for (int s = 0; s < EPL; ++s) p[s] = fma(-heps, g[s], p[s]);
#pragma unroll 1
      for (int i = 1; i < N; ++i) {
        drift<T, EPL, UNITM>(q, p, minv, eps);
        model.template kick<EPL, LPC, false>(q, p, eps, c);
      }
      cVg += N - 1;
      drift<T, EPL, UNITM>(q, p, minv, eps);
#pragma unroll
      for (int s = 0; s < EPL; ++s) g[s] = T(0);
      const T V1 = model.template kick<EPL, LPC, true>(q, g, T(-1), c);
      cVg += 1;
#pragma unroll
      for (int s = 0; s < EPL; ++s) p[s] = fma(-heps, g[s], p[s]);
      const T H0 = V0 + K0, H1 = V1 + kinetic<T, EPL, LPC, UNITM>(p, minv, c.mask);
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
      double u;
      if (ga.pu_in) {
        const double w = pu + 1.0 + ga.delta;
        double r = fmod(w, 2.0);
        if (r < 0.0) r += 2.0;
        u = r - 1.0;
      } else {
        u = draw_uniform<T>(a, chain, 0, (uint32_t)k);
      }
      const T pm = prob < T(1) ? prob : T(1);
      const bool accept = !bad && !(u > (double)pm);
      pu = accept ? u / (double)prob : u;
      if (k == K - 1 && c.lane == 0) {
        if (a.code) a.code[chain] = bad ? -1 : (accept ? 1 : 0);
        if (a.energies) { a.energies[2 * chain] = H0; a.energies[2 * chain + 1] = H1; }
      }
      if (accept) {
        ++nacc;
        has = true;
        V0 = V1;
#pragma unroll
        for (int s = 0; s < EPL; ++s) {
          sQ[s * THREADS] = q[s];
          sP[s * THREADS] = p[s];
        }
      } else {
#pragma unroll
        for (int s = 0; s < EPL; ++s) {
          q[s] = sQ[s * THREADS];
          sP[s * THREADS] = -sP[s * THREADS];
        }
        if (had_cache) {
#pragma unroll
          for (int s = 0; s < EPL; ++s) g[s] = sG[s * THREADS];
        }
      }
    }
    store_vec<T, EPL, LPC>(q, a.q_out, off, a.sd, c.lane, d);
    if (a.p_out) {
#pragma unroll
      for (int s = 0; s < EPL; ++s) p[s] = sP[s * THREADS];
      store_vec<T, EPL, LPC>(p, a.p_out, off, a.sd, c.lane, d);
    }
    if (ga.Vg_out) {
      if (!has) {
#pragma unroll
        for (int s = 0; s < EPL; ++s) g[s] = (T)NAN;        // still "None" for this chain
      }
      store_vec<T, EPL, LPC>(g, ga.Vg_out, off, a.sd, c.lane, d);
    }
    if (c.lane == 0) {
      if (ga.V_out) ga.V_out[chain] = has ? V0 : (T)NAN;
      if (ga.pu_out) ga.pu_out[chain] = pu;
      if (ga.n_accept) ga.n_accept[chain] += nacc;
      if (a.counts) { a.counts[3 * chain] = cH; a.counts[3 * chain + 1] = cVg; a.counts[3 * chain + 2] = K; }
    }
  }
}

// ------------------------------------------------------------------ AdHMC_eps0
// Per-transition random step size (src/adaptivehmc.py:111-192): ntry log-spaced trial
// step sizes, nsteps leapfrog steps each from (q0,p0); weights exp(H0-H_i) sqrt(eps_i)
// normalised; inverse-CDF draw of a bin, uniform within the bin; main trajectory of
// n_leap steps; the same sweep from (q1,-p1) gives the reverse probability; Hastings
// factor pb/pf multiplies exp(H0-H1).  Base states live in per-thread scratch frames.
constexpr int kMaxTry = 64;

template <typename T>
struct AdArgs {
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
  const T* minv;
  int32_t n_leap, ntry, nsteps;
  double tint;             // > 0: the main trajectory takes int(tint/eps) + 1 steps (src/adaptivehmc.py:178-179)
  double steps[kMaxTry];   // trial step sizes (host-computed np.logspace)
  double pwts[kMaxTry];    // steps**0.5
  T* p_out;
  int32_t* code;
  T* probs;                // [C,3] prob, prob/hastings, hastings
  T* aux;                  // [C,3] pf_den, pb_num, eps
  int32_t* counts;         // [C,3]
  T* frames;               // [4*EPL*frame_threads]: p0, q1, p1 (+1 spare)
  int64_t frame_threads;
};

template <typename T>
__device__ __forceinline__ double draw_uniform_ad(const AdArgs<T>& a, int64_t chain, uint64_t gchain, int idx) {
  if (a.u_inject) return a.u_inject[chain * a.u_stride + idx];
  uint32_t w[4];
  Philox::gen(a.seed, gchain, a.iteration, STREAM_STEPSIZE, (uint32_t)(idx >> 1), w);
  return (idx & 1) ? u53(w[2], w[3]) : u53(w[0], w[1]);
}

constexpr int kAdRegBaseMaxEpl = 28;   // q, p and the base state (q, p) of the trial sweeps: 4*EPL values per lane

template <class Model, typename T, int EPL, int LPC, bool UNITM, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) k_adhmc_eps0(const Model model, const AdArgs<T> a) {
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
  if (Model::kNeedsSmem) c.smem_q = (T*)smem_raw + model.p_smem_elems() + (size_t)(threadIdx.x / LPC) * d;
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
  const int64_t FT = a.frame_threads;
  T* f_p0 = a.frames + gthread;
  T* f_q1 = f_p0 + (int64_t)EPL * FT;
  T* f_p1 = f_q1 + (int64_t)EPL * FT;
  const int64_t total_groups = (int64_t)gridDim.x * GPB;
  const int ntry = a.ntry;

  for (int64_t chain = gthread / LPC; chain < a.n_chains; chain += total_groups) {
    const int64_t off = chain * a.sc;
    const uint64_t gchain = (uint64_t)(a.chain_offset + chain);
    T q[EPL], p[EPL];
    // the state every trial trajectory of a sweep starts from: in registers when the shape leaves room
    // (4*EPL values per lane), else re-read from global memory / the per-thread scratch for every trial
    constexpr bool REGBASE = EPL <= kAdRegBaseMaxEpl;
    T qb[REGBASE ? EPL : 1], pb[REGBASE ? EPL : 1];
    int cH = 0, cVg = 0, cLeap = 0;
    T pp[kMaxTry];

    // draw p0 and park it
    {
      TransArgs<T> ta;
      ta.chain_offset = a.chain_offset; ta.sc = a.sc; ta.sd = a.sd; ta.seed = a.seed; ta.iteration = a.iteration;
      ta.rng_flags = a.rng_flags; ta.p_inject = a.p_inject; ta.p_prev = nullptr;
      draw_momentum<T, EPL, LPC, UNITM>(p, ta, minv, chain, c.lane, d);
    }
#pragma unroll
    for (int s = 0; s < EPL; ++s) f_p0[(int64_t)s * FT] = p[s];
    if (REGBASE) {
      load_vec<T, EPL, LPC>(q, a.q_in, off, a.sd, c.lane, d, T(1));
#pragma unroll
      for (int s = 0; s < EPL; ++s) { qb[REGBASE ? s : 0] = q[s]; pb[REGBASE ? s : 0] = p[s]; }
    }

    // sweep(from forward base or reverse base): fills pp[] normalised; returns H at the base
    auto sweep = [&](bool reverse, T& Hbase) {
      T sum = T(0);
      for (int i = 0; i <= ntry; ++i) {
        // i == 0 evaluates H at the base through a zero-length "trajectory" (kick with h=0)
        if (REGBASE) {
#pragma unroll
          for (int s = 0; s < EPL; ++s) { q[s] = qb[REGBASE ? s : 0]; p[s] = reverse ? -pb[REGBASE ? s : 0] : pb[REGBASE ? s : 0]; }
        } else if (!reverse) {
          load_vec<T, EPL, LPC>(q, a.q_in, off, a.sd, c.lane, d, T(1));
#pragma unroll
          for (int s = 0; s < EPL; ++s) p[s] = f_p0[(int64_t)s * FT];
        } else {
#pragma unroll
          for (int s = 0; s < EPL; ++s) { q[s] = f_q1[(int64_t)s * FT]; p[s] = -f_p1[(int64_t)s * FT]; }
        }
        if (i == 0) {
          const T K = kinetic<T, EPL, LPC, UNITM>(p, minv, c.mask);
          T tmp[EPL];
#pragma unroll
          for (int s = 0; s < EPL; ++s) tmp[s] = T(0);
          Hbase = model.template kick<EPL, LPC, true>(q, tmp, T(0), c) + K;
          cH += 1;
          continue;
        }
        T Vs, Ve;
        const int N = a.nsteps > 1 ? a.nsteps : 1;
        leapfrog<Model, T, EPL, LPC, UNITM>(model, q, p, minv, N, (T)a.steps[i - 1], c, Vs, Ve);
        const T Hi = Ve + kinetic<T, EPL, LPC, UNITM>(p, minv, c.mask);
        cH += 1; cVg += N + 1; cLeap += 1;
        T w = t_exp<T>(Hbase - Hi) * (T)a.pwts[i - 1];
        if (t_bad(w)) w = T(0);
        pp[i - 1] = w;
        sum += w;
      }
      for (int i = 0; i < ntry; ++i) pp[i] = pp[i] / sum;
    };

    // forward sweep: get_stepsize(q0, p0) (src/adaptivehmc.py:180 -> :111-131).  The
    // reference evaluates H0 twice (hmc_step :163 and get_stepsize :112): count both.
    T H0;
    sweep(false, H0);
    cH += 1;
    T cdf_min = pp[0], run = T(0);
    int isx = -1;
    const double u0 = draw_uniform_ad(a, chain, gchain, 0);
    // cdf = cumsum(pp); sx = U(low=cdf.min()); isx = last index with sx > cdf
    {
      T r2 = T(0);
      for (int i = 0; i < ntry; ++i) { r2 += pp[i]; if (i == 0 || r2 < cdf_min) cdf_min = r2; }
      const T sx = cdf_min + (T(1) - cdf_min) * (T)u0;
      for (int i = 0; i < ntry; ++i) { run += pp[i]; if (sx > run) isx = i; }
    }
    bool broken = (isx < 0) || (isx > ntry - 2);   // the reference raises IndexError here; we reject (code 2)
    if (isx < 0) isx = 0;
    if (isx > ntry - 2) isx = ntry - 2;
    const double u1 = draw_uniform_ad(a, chain, gchain, 1);
    const T ss = (T)(a.steps[isx] + (a.steps[isx + 1] - a.steps[isx]) * u1);
    const T pf_den = pp[isx + 1];

    // main trajectory from (q0, p0)
    if (REGBASE) {
#pragma unroll
      for (int s = 0; s < EPL; ++s) { q[s] = qb[REGBASE ? s : 0]; p[s] = pb[REGBASE ? s : 0]; }
    } else {
      load_vec<T, EPL, LPC>(q, a.q_in, off, a.sd, c.lane, d, T(1));
#pragma unroll
      for (int s = 0; s < EPL; ++s) p[s] = f_p0[(int64_t)s * FT];
    }
    T Vs, Ve;
    int Nm = a.n_leap > 1 ? a.n_leap : 1;
    if (a.tint > 0) { Nm = (int)(a.tint / (double)ss) + 1; if (Nm < 1) Nm = 1; }
    leapfrog<Model, T, EPL, LPC, UNITM>(model, q, p, minv, Nm, ss, c, Vs, Ve);
    const T H1 = Ve + kinetic<T, EPL, LPC, UNITM>(p, minv, c.mask);
    cH += 1; cVg += Nm + 1; cLeap += 1;
    T dq = T(0);
#pragma unroll
    for (int s = 0; s < EPL; ++s) {
      const int i = Lay::elem(c.lane, s);
      if (i < d) dq += (REGBASE ? qb[REGBASE ? s : 0] : a.q_in[off + i * a.sd]) - q[s];
      f_q1[(int64_t)s * FT] = q[s];
      f_p1[(int64_t)s * FT] = p[s];
      if (REGBASE) { qb[REGBASE ? s : 0] = q[s]; pb[REGBASE ? s : 0] = p[s]; }
    }
    dq = group_sum_masked<LPC, T>(dq, c.mask);

    // reverse sweep: get_stepsize(q1, -p1, eps=ss) (:182 -> :132-134)
    T Hb;
    sweep(true, Hb);
    int jb = -1;
    for (int i = ntry - 1; i >= 0; --i) if (a.steps[i] > (double)ss) jb = i;
    if (jb < 0) { jb = ntry - 1; broken = true; }
    const T pb_num = pp[jb];
    const T hast = pb_num / pf_den;
    const T prob = t_exp<T>(H0 - H1) * hast;
    const bool bad = broken || t_bad(prob) || dq == T(0);
    bool accept = false;
    if (!bad) {
      const double u2 = draw_uniform_ad(a, chain, gchain, 2);
      const T pm = prob < T(1) ? prob : T(1);
      accept = !(u2 > (double)pm);
    }
    if (c.lane == 0) {
      if (a.code) a.code[chain] = bad ? 2 : (accept ? 1 : 0);
      if (a.probs) { a.probs[3 * chain] = prob; a.probs[3 * chain + 1] = prob / hast; a.probs[3 * chain + 2] = hast; }
      if (a.aux) { a.aux[3 * chain] = pf_den; a.aux[3 * chain + 1] = pb_num; a.aux[3 * chain + 2] = ss; }
      if (a.counts) { a.counts[3 * chain] = cH; a.counts[3 * chain + 1] = cVg; a.counts[3 * chain + 2] = cLeap; }
    }
    if (accept) {
#pragma unroll
      for (int s = 0; s < EPL; ++s) { q[s] = f_q1[(int64_t)s * FT]; p[s] = f_p1[(int64_t)s * FT]; }
      store_vec<T, EPL, LPC>(q, a.q_out, off, a.sd, c.lane, d);
      if (a.p_out) store_vec<T, EPL, LPC>(p, a.p_out, off, a.sd, c.lane, d);
    } else {
      if (a.q_out != a.q_in) {
#pragma unroll
        for (int s = 0; s < EPL; ++s) {
          const int i = Lay::elem(c.lane, s);
          if (i < d) a.q_out[off + i * a.sd] = a.q_in[off + i * a.sd];
        }
      }
      if (a.p_out) {
#pragma unroll
        for (int s = 0; s < EPL; ++s) p[s] = f_p0[(int64_t)s * FT];
        store_vec<T, EPL, LPC>(p, a.p_out, off, a.sd, c.lane, d);
      }
    }
  }
}

// ------------------------------------------------------------------ AdHMC / AdHMC_tint
// adaptivehmc.AdHMC.hmc_step (src/adaptivehmc.py:227-354) and AdHMC_tint.hmc_step (:394-519): the
// step size walks inside the trajectory.  The trajectory is cut into nseg segments; after each one
// every ratio r_j proposes the step size eps*r_j, a candidate segment is integrated for each, one is
// chosen with probability proportional to w_j e^{H1-H_j}, and the probability that the reverse walk
// (from the same state with the momentum flipped, at the new step size) picks the old step size is
// collected for the Hastings factor.  A segment is `nleap` steps (AdHMC) or max(1, int(tint/eps))
// steps (AdHMC_tint).  The first step size and the last reverse probability come from the same trial
// sweep as AdHMC_eps0 (get_stepsize, :202-224).
//
// Per thread the scratch holds p0, the current base state (q1, p1) and the nr candidate states;
// everything else is scalar.  Step sizes and choice probabilities are double whatever T is.
constexpr int kMaxRatios = 4;

template <typename T>
struct WalkArgs {
  int64_t n_chains, chain_offset;
  int32_t d;
  const T* q_in;
  T* q_out;
  int64_t sc, sd;
  uint64_t seed;
  uint32_t iteration;
  int32_t rng_flags;
  const T* p_inject;
  const double* u_inject;    // [C,u_stride]: sx, sx2, one per later segment (the choice), accept
  int32_t u_stride;
  const T* minv;
  int32_t nseg;              // N = Nleap // nleap (AdHMC) or ntimes (AdHMC_tint)
  int32_t nleap;             // steps per segment (AdHMC)
  int32_t tint_mode;         // 1: steps per segment = max(1, int(tint/eps))
  double tint;
  int32_t nr;
  double ratios[kMaxRatios], pwts0[kMaxRatios];
  double smin, smax;
  int32_t ntry, nsteps;
  double steps[kMaxTry], pwts[kMaxTry];
  T* p_out;
  int32_t* code;             // 1 / 0 / 2, or 100+i when the forward weights of segment i blow up (:296-301)
  T* probs;                  // [C,3] prob, prob/hastings, hastings (NaN on blow-up)
  T* checks;                 // [C,3,nseg] pf_den, pb_num, eps
  int32_t* counts;           // [C,3]
  T* frames;                 // [(3 + 2*nr)*EPL*frame_threads]
  int64_t frame_threads;
};

template <class Model, typename T, int EPL, int LPC, bool UNITM, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) k_adhmc_walk(const Model model, const WalkArgs<T> a) {
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
  if (Model::kNeedsSmem) c.smem_q = (T*)smem_raw + model.p_smem_elems() + (size_t)(threadIdx.x / LPC) * d;
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
  const int64_t FT = a.frame_threads;
  const int64_t slot = (int64_t)EPL * FT;
  T* f_p0 = a.frames + gthread;
  T* f_q1 = f_p0 + slot;
  T* f_p1 = f_q1 + slot;
  T* f_cand = f_p1 + slot;   // candidate j: q at (2j)*slot, p at (2j+1)*slot
  const int64_t total_groups = (int64_t)gridDim.x * GPB;
  const int ntry = a.ntry, nr = a.nr, N = a.nseg;

  for (int64_t chain = gthread / LPC; chain < a.n_chains; chain += total_groups) {
    const int64_t off = chain * a.sc;
    const uint64_t gchain = (uint64_t)(a.chain_offset + chain);
    T q[EPL], p[EPL];
    // the base state (q0, p0) and later (q1, p1) that sweeps, candidate segments and reverse segments all start
    // from: in registers when the shape leaves room, else in the per-thread scratch (k_adhmc_eps0 does the same)
    constexpr bool REGBASE = EPL <= kAdRegBaseMaxEpl;
    T qb[REGBASE ? EPL : 1], pb[REGBASE ? EPL : 1];
    int cH = 0, cVg = 0, cLeap = 0;
    T pp[kMaxTry];

    auto uniform = [&](int idx) -> double {
      if (a.u_inject) return a.u_inject[chain * a.u_stride + idx];
      uint32_t w[4];
      Philox::gen(a.seed, gchain, a.iteration, STREAM_STEPSIZE, (uint32_t)(idx >> 1), w);
      return (idx & 1) ? u53(w[2], w[3]) : u53(w[0], w[1]);
    };
    auto seg_steps = [&](double eps) -> int {
      if (!a.tint_mode) return a.nleap > 1 ? a.nleap : 1;
      const int n = (int)(a.tint / eps);
      return n > 1 ? n : 1;
    };
    auto put_check = [&](int row, int i, double v) {
      if (a.checks && c.lane == 0) a.checks[(chain * 3 + row) * N + i] = (T)v;
    };
    auto load_base = [&](T sign) {
      if (REGBASE) {
#pragma unroll
        for (int s = 0; s < EPL; ++s) { q[s] = qb[REGBASE ? s : 0]; p[s] = sign * pb[REGBASE ? s : 0]; }
      } else {
#pragma unroll
        for (int s = 0; s < EPL; ++s) { q[s] = f_q1[(int64_t)s * FT]; p[s] = sign * f_p1[(int64_t)s * FT]; }
      }
    };
    auto set_base = [&]() {      // (q, p) becomes the base
#pragma unroll
      for (int s = 0; s < EPL; ++s) {
        f_q1[(int64_t)s * FT] = q[s];
        f_p1[(int64_t)s * FT] = p[s];
        if (REGBASE) { qb[REGBASE ? s : 0] = q[s]; pb[REGBASE ? s : 0] = p[s]; }
      }
    };

    {
      TransArgs<T> ta;
      ta.chain_offset = a.chain_offset; ta.sc = a.sc; ta.sd = a.sd; ta.seed = a.seed; ta.iteration = a.iteration;
      ta.rng_flags = a.rng_flags; ta.p_inject = a.p_inject; ta.p_prev = nullptr;
      draw_momentum<T, EPL, LPC, UNITM>(p, ta, minv, chain, c.lane, d);
    }
#pragma unroll
    for (int s = 0; s < EPL; ++s) f_p0[(int64_t)s * FT] = p[s];
    if (REGBASE) {
      load_vec<T, EPL, LPC>(q, a.q_in, off, a.sd, c.lane, d, T(1));
#pragma unroll
      for (int s = 0; s < EPL; ++s) { qb[REGBASE ? s : 0] = q[s]; pb[REGBASE ? s : 0] = p[s]; }
    }
    if (a.checks && c.lane == 0)
      for (int i = 0; i < 3 * N; ++i) a.checks[chain * 3 * N + i] = T(0);

    // get_stepsize's trial sweep from (q0, p0) or from the current base with p flipped
    auto sweep = [&](bool reverse, T& Hbase) {
      T sum = T(0);
      for (int i = 0; i <= ntry; ++i) {
        if (REGBASE) {
          load_base(reverse ? T(-1) : T(1));      // the base registers hold (q0, p0) until the first segment is done
        } else if (!reverse) {
          load_vec<T, EPL, LPC>(q, a.q_in, off, a.sd, c.lane, d, T(1));
#pragma unroll
          for (int s = 0; s < EPL; ++s) p[s] = f_p0[(int64_t)s * FT];
        } else {
          load_base(T(-1));
        }
        if (i == 0) {
          const T K = kinetic<T, EPL, LPC, UNITM>(p, minv, c.mask);
          T tmp[EPL];
#pragma unroll
          for (int s = 0; s < EPL; ++s) tmp[s] = T(0);
          Hbase = model.template kick<EPL, LPC, true>(q, tmp, T(0), c) + K;
          cH += 1;
          continue;
        }
        T Vs, Ve;
        const int Nt = a.nsteps > 1 ? a.nsteps : 1;
        leapfrog<Model, T, EPL, LPC, UNITM>(model, q, p, minv, Nt, (T)a.steps[i - 1], c, Vs, Ve);
        const T Hi = Ve + kinetic<T, EPL, LPC, UNITM>(p, minv, c.mask);
        cH += 1; cVg += Nt + 1; cLeap += 1;
        T w = t_exp<T>(Hbase - Hi) * (T)a.pwts[i - 1];
        if (t_bad(w)) w = T(0);
        pp[i - 1] = w;
        sum += w;
      }
      for (int i = 0; i < ntry; ++i) pp[i] = pp[i] / sum;
    };

    // ---- first step size (src/adaptivehmc.py:268 -> :202-222); H0 is evaluated by hmc_step and by get_stepsize
    T H0;
    sweep(false, H0);
    cH += 1;
    int isx = -1;
    {
      T cdf_min = pp[0], r2 = T(0), run = T(0);
      for (int i = 0; i < ntry; ++i) { r2 += pp[i]; if (i == 0 || r2 < cdf_min) cdf_min = r2; }
      const T sx = cdf_min + (T(1) - cdf_min) * (T)uniform(0);
      for (int i = 0; i < ntry; ++i) { run += pp[i]; if (sx > run) isx = i; }
    }
    bool broken = (isx < 0) || (isx > ntry - 2);   // the reference raises IndexError here; we reject (code 2)
    if (isx < 0) isx = 0;
    if (isx > ntry - 2) isx = ntry - 2;
    double ss = a.steps[isx] + (a.steps[isx + 1] - a.steps[isx]) * uniform(1);
    double prod_pf = (double)pp[isx + 1], prod_pb = 1.0;
    put_check(0, 0, prod_pf);
    put_check(2, 0, ss);

    // ---- first segment from (q0, p0)
    if (REGBASE) {
      load_base(T(1));
    } else {
      load_vec<T, EPL, LPC>(q, a.q_in, off, a.sd, c.lane, d, T(1));
#pragma unroll
      for (int s = 0; s < EPL; ++s) p[s] = f_p0[(int64_t)s * FT];
    }
    T Vs, Ve;
    {
      const int n0 = seg_steps(ss);
      leapfrog<Model, T, EPL, LPC, UNITM>(model, q, p, minv, n0, (T)ss, c, Vs, Ve);
      cH += 1; cVg += n0 + 1; cLeap += 1;
    }
    T H1 = Ve + kinetic<T, EPL, LPC, UNITM>(p, minv, c.mask);
    set_base();
    T Hprev = H0;
    double ssnew = ss;
    int blown = -1, pbind = 0;

    for (int i = 0; i < N - 1; ++i) {
      double pf[kMaxRatios], pb[kMaxRatios];
      T Hs[kMaxRatios];
      // forward candidates (:280-290)
      for (int j = 0; j < nr; ++j) {
        const double ss2 = ss * a.ratios[j];
        const int n = seg_steps(ss2);
        load_base(T(1));
        leapfrog<Model, T, EPL, LPC, UNITM>(model, q, p, minv, n, (T)ss2, c, Vs, Ve);
        Hs[j] = Ve + kinetic<T, EPL, LPC, UNITM>(p, minv, c.mask);
        cH += 1; cVg += n + 1; cLeap += 1;
        T* fq = f_cand + (int64_t)(2 * j) * slot;
#pragma unroll
        for (int s = 0; s < EPL; ++s) { fq[(int64_t)s * FT] = q[s]; fq[slot + (int64_t)s * FT] = p[s]; }
        T pH = t_exp<T>(H1 - Hs[j]);
        if (t_bad(pH)) pH = T(0);
        pf[j] = (double)pH;
      }
      double sum = 0.0;
      for (int j = 0; j < nr; ++j) {
        double w = a.pwts0[j];
        if (j == 0 && a.smin > ss * a.ratios[0]) w = 0.0;
        if (j == nr - 1 && a.smax < ss * a.ratios[nr - 1]) w = 0.0;
        pf[j] *= w;
        sum += pf[j];
      }
      double sum2 = 0.0;
      for (int j = 0; j < nr; ++j) { pf[j] /= sum; sum2 += pf[j]; }
      if (t_bad(sum2)) { blown = i; break; }
      // np.random.choice(nr, p=pf): cdf = cumsum(pf) / cdf[-1]; index = #(cdf <= u)
      int pind = 0;
      {
        double cdf[kMaxRatios], run = 0.0;
        for (int j = 0; j < nr; ++j) { run += pf[j]; cdf[j] = run; }
        const double u = uniform(2 + i);
        for (int j = 0; j < nr; ++j) if (cdf[j] / cdf[nr - 1] <= u) pind = j + 1;
        if (pind > nr - 1) pind = nr - 1;
      }
      ssnew = ss * a.ratios[pind];
      prod_pf *= pf[pind];
      put_check(0, i + 1, pf[pind]);
      const T H2 = Hs[pind];
      // reverse walk from (q1, -p1) at the new step size (:311-329)
      for (int j = 0; j < nr; ++j) {
        const double ss2 = ssnew * a.ratios[j];
        T Hb;
        if (fabs(ss2 - ss) <= 1e-8 + 1e-5 * fabs(ss)) {        // np.allclose(ssnew*ratios[j], ss)
          Hb = Hprev;
          pbind = j;
        } else {
          const int n = seg_steps(ss2);
          load_base(T(-1));
          leapfrog<Model, T, EPL, LPC, UNITM>(model, q, p, minv, n, (T)ss2, c, Vs, Ve);
          Hb = Ve + kinetic<T, EPL, LPC, UNITM>(p, minv, c.mask);
          cH += 1; cVg += n + 1; cLeap += 1;
        }
        T pH = t_exp<T>(H1 - Hb);
        if (t_bad(pH)) pH = T(0);
        pb[j] = (double)pH;
      }
      sum = 0.0;
      for (int j = 0; j < nr; ++j) {
        double w = a.pwts0[j];
        if (j == 0 && a.smin > ssnew * a.ratios[0]) w = 0.0;
        if (j == nr - 1 && a.smax < ssnew * a.ratios[nr - 1]) w = 0.0;
        pb[j] *= w;
        sum += pb[j];
      }
      const double pbn = pb[pbind] / sum;
      prod_pb *= pbn;
      put_check(1, i, pbn);
      put_check(2, i + 1, ssnew);
      // the chosen candidate becomes the base
      ss = ssnew;
      Hprev = H1;
      H1 = H2;
      {
        const T* fq = f_cand + (int64_t)(2 * pind) * slot;
#pragma unroll
        for (int s = 0; s < EPL; ++s) { q[s] = fq[(int64_t)s * FT]; p[s] = fq[slot + (int64_t)s * FT]; }
        set_base();
      }
    }

    if (blown >= 0) {
      // "Something blew up so returning initial position" (:296-301)
      if (c.lane == 0) {
        const T qn = (T)nan("");
        if (a.code) a.code[chain] = 100 + blown;
        if (a.probs) { a.probs[3 * chain] = qn; a.probs[3 * chain + 1] = qn; a.probs[3 * chain + 2] = qn; }
        if (a.counts) { a.counts[3 * chain] = cH; a.counts[3 * chain + 1] = cVg; a.counts[3 * chain + 2] = cLeap; }
      }
    } else {
      // ---- probability that the reverse walk starts with the final step size (:340-342)
      double pb_last = 0.0;
      if (ssnew > a.smin && ssnew < a.smax) {
        T Hb;
        sweep(true, Hb);
        int jb = -1;
        for (int i = ntry - 1; i >= 0; --i) if (a.steps[i] > ssnew) jb = i;
        if (jb < 0) { jb = ntry - 1; broken = true; }
        pb_last = (double)pp[jb];
      }
      prod_pb *= pb_last;
      put_check(1, N - 1, pb_last);
      const double hast = prod_pb / prod_pf;
      const double prob = (double)t_exp<T>(H0 - H1) * hast;
      T dq = T(0);
#pragma unroll
      for (int s = 0; s < EPL; ++s) {
        const int i = Lay::elem(c.lane, s);
        if (i < d) dq += a.q_in[off + i * a.sd] - f_q1[(int64_t)s * FT];
      }
      dq = group_sum_masked<LPC, T>(dq, c.mask);
      const bool bad = broken || t_bad(prob) || dq == T(0);
      bool accept = false;
      if (!bad) {
        const double u = uniform(N + 1);
        accept = !(u > (prob < 1.0 ? prob : 1.0));
      }
      if (c.lane == 0) {
        if (a.code) a.code[chain] = bad ? 2 : (accept ? 1 : 0);
        if (a.probs) { a.probs[3 * chain] = (T)prob; a.probs[3 * chain + 1] = (T)(prob / hast); a.probs[3 * chain + 2] = (T)hast; }
        if (a.counts) { a.counts[3 * chain] = cH; a.counts[3 * chain + 1] = cVg; a.counts[3 * chain + 2] = cLeap; }
      }
      if (accept) {
        load_base(T(1));
        store_vec<T, EPL, LPC>(q, a.q_out, off, a.sd, c.lane, d);
        if (a.p_out) store_vec<T, EPL, LPC>(p, a.p_out, off, a.sd, c.lane, d);
        continue;
      }
    }
    // rejected, no-move or blown up: (q0, p0)
    if (a.q_out != a.q_in) {
#pragma unroll
      for (int s = 0; s < EPL; ++s) {
        const int i = Lay::elem(c.lane, s);
        if (i < d) a.q_out[off + i * a.sd] = a.q_in[off + i * a.sd];
      }
    }
    if (a.p_out) {
#pragma unroll
      for (int s = 0; s < EPL; ++s) p[s] = f_p0[(int64_t)s * FT];
      store_vec<T, EPL, LPC>(p, a.p_out, off, a.sd, c.lane, d);
    }
  }
}

}  // namespace hmcb
