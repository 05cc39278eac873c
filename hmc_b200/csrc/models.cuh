// models.cuh -- device-side model densities with analytic gradients.
//
// A model provides, for the chain whose elements sit in q[EPL] across the LPC
// lanes of a group (see Layout in common.cuh):
//
//   T kick<EPL,LPC,WANT_V>(q, p, h, ctx)   p -= h * grad V(q);  returns V(q) if WANT_V
//
// where V = -log_prob.  The gradient never leaves registers: the momentum update
// is fused into its evaluation.  These are the in-kernel restatements of the
// numpy callables in hmc_b200/models.py; the funnel follows the reference's Stan
// text (scripts/models.py:9-19), the others are builder-defined (the reference
// has no such models).
#pragma once
#include "common.cuh"

namespace hmcb {

template <typename T>
struct GroupCtx {
  int lane;        // lane within the group, 0..LPC-1
  int d;           // true dimension
  unsigned mask;   // warp lanes of this group (shuffle mask)
  T* smem_q;       // gaussian: per-group staging area of d elements (else unused)
};

// -------------------------------------------------------------------- funnel
// V(v,theta) = v^2/(2 s^2) + (n/2) v + exp(-v)/2 * sum(theta^2)
// dV/dv = v/s^2 + n/2 - exp(-v)/2 * S,  dV/dtheta_i = exp(-v) theta_i
template <typename T>
struct FunnelModel {
  T inv_s2, half_n;
  static constexpr bool kNeedsSmem = false;
  __host__ static FunnelModel make(const hmcb_model& m) {
    FunnelModel f;
    f.inv_s2 = (T)(1.0 / (m.par[0] * m.par[0]));
    f.half_n = (T)(0.5 * (m.d - 1));
    return f;
  }
  __host__ __device__ size_t p_smem_elems() const { return 0; }
  __device__ __forceinline__ void block_init(void*, int) const {}

  template <int EPL, int LPC, bool WANT_V>
  __device__ __forceinline__ T kick(const T (&q)[EPL], T (&p)[EPL], T h, const GroupCtx<T>& c) const {
    const bool own_v = (c.lane == 0);
    // exp(-v) first: its dependent chain then overlaps the independent sums below
    const T v = (LPC == 1) ? q[0] : __shfl_sync(c.mask, q[0], 0, LPC);
    const T e = t_exp_loop<T>(-v);
    const T he = h * e;
    // S = sum theta^2: four partial sums so the FMA chain does not serialise
    constexpr int NA = HMCB_FUNNEL_NACC;
    T acc[NA];
#pragma unroll
    for (int i = 0; i < NA; ++i) acc[i] = T(0);
    const T t0 = own_v ? T(0) : q[0];
    acc[0] = t0 * t0;
#pragma unroll
    for (int s = 1; s < EPL; ++s) acc[s % NA] = fma(q[s], q[s], acc[s % NA]);
#pragma unroll
    for (int w = NA >> 1; w > 0; w >>= 1)
#pragma unroll
      for (int i = 0; i < w; ++i) acc[i] += acc[i + w];
    T S = acc[0];
#pragma unroll
    for (int o = LPC >> 1; o > 0; o >>= 1) S += __shfl_xor_sync(c.mask, S, o);
    // every multiply-add below is written as the fma it is meant to be: left to the compiler, whether a product is fused
    // into the sum that follows depends on what the call site makes constant (h = -1 into a zeroed p in the GHMC kernels),
    // and two kernels that must agree bit for bit then differ in the last place for a few chains per step
    const T hE = T(0.5) * e;
    const T gv = fma(-hE, S, fma(v, inv_s2, half_n));
    p[0] = own_v ? fma(-h, gv, p[0]) : fma(-he, q[0], p[0]);
#pragma unroll
    for (int s = 1; s < EPL; ++s) p[s] = fma(-he, q[s], p[s]);
    if (WANT_V) return fma(hE, S, fma(half_n, v, mul_rn(T(0.5) * v * v, inv_s2)));
    return T(0);
  }
};

// -------------------------------------------------------------------- pairwise Rosenbrock
// V = sum_k [(a-u_k)^2 + b (w_k-u_k^2)^2] / T,   (u_k, w_k) = (x[2k], x[2k+1])
template <typename T>
struct RosenbrockModel {
  T a, b, inv_T;
  T b4, a2, bm2;     // 4b, 2a, -2b: operands of the pair gradient, read from the kernel's parameter bank
  static constexpr bool kNeedsSmem = false;
  __host__ static RosenbrockModel make(const hmcb_model& m) {
    RosenbrockModel r;
    r.a = (T)m.par[0];
    r.b = (T)m.par[1];
    r.inv_T = (T)(1.0 / m.par[2]);
    r.b4 = (T)(4.0 * m.par[1]);
    r.a2 = (T)(2.0 * m.par[0]);
    r.bm2 = (T)(-2.0 * m.par[1]);
    return r;
  }
  __host__ __device__ size_t p_smem_elems() const { return 0; }
  __device__ __forceinline__ void block_init(void*, int) const {}

  template <int EPL, int LPC, bool WANT_V>
  __device__ __forceinline__ T kick(const T (&q)[EPL], T (&p)[EPL], T h, const GroupCtx<T>& c) const {
    static_assert(EPL % 2 == 0, "Rosenbrock pairs need an even slot count");
    T Vl = T(0);
    // p -= h grad V with the pair gradient in nested form: with r = w - u^2 and hs = h / T,
    //   -h dV/du = hs (u (4b r - 2) + 2a),   -h dV/dw = (-2b hs) r
    // -- five FP64 instructions per pair (r, 4b r - 2, u (..) + 2a, two fmas into p) where the gradient written out term
    // by term (a-u, b r, 4u, 4u b r, -2(a-u) - .., 2 b r, two fmas) takes nine: the trajectory loop of the d=100 kernels is
    // bound by the FP64 pipe's issue rate (57 % active at 0.37 of the algorithmic peak).  4b, 2a, -2b come from the
    // parameter bank, so the only values live across a kick are hs and -2b hs (four folded per-call factors cost the
    // step-size kernels, which run at their register limit, a quarter of their speed).  Every multiply-add is an explicit
    // fma (see FunnelModel).
    const T hs = h * inv_T;
    const T hsw = hs * bm2;
#pragma unroll
    for (int s = 0; s < EPL; s += 2) {
      const bool valid = Layout<EPL, LPC>::elem(c.lane, s) < c.d;  // padding pairs contribute nothing
      const T u = q[s], w = q[s + 1];
      const T r = fma(-u, u, w);
      const T t = fma(b4, r, T(-2));
      // padding pairs hold q = p = 0: r = 0 leaves p_w alone, and with the constant term masked so does u (4b r - 2) + 2a
      // -- one select on a loop-invariant value instead of one per updated momentum (`valid ? fma(..) : p` put four FSEL
      // per pair into the trajectory loop: as many instructions beside the 98 DFMA of a d=100 lane as the DFMA themselves)
      const T gi = fma(u, t, valid ? a2 : T(0));
      p[s] = fma(hs, gi, p[s]);
      p[s + 1] = fma(hsw, r, p[s + 1]);
      if (WANT_V) {
        const T am = a - u;
        Vl += valid ? fma(am, am, mul_rn(b * r, r)) : T(0);
      }
    }
    if (WANT_V) {
#pragma unroll
      for (int o = LPC >> 1; o > 0; o >>= 1) Vl += __shfl_xor_sync(c.mask, Vl, o);
      return Vl * inv_T;
    }
    return T(0);
  }
};

// -------------------------------------------------------------------- dense Gaussian
// V = x^T P x / 2, grad V = P x.  P (row-major, symmetric) is copied to shared
// memory once per block when it fits, else read through L1/L2.  Each group stages
// its q in shared memory so every lane can read all d elements.
template <typename T>
struct GaussianModel {
  const T* P;      // global
  int d;
  int p_in_smem;   // host decision
  static constexpr bool kNeedsSmem = true;
  __host__ static GaussianModel make(const hmcb_model& m) {
    GaussianModel g;
    g.P = (const T*)m.dev0;
    g.d = m.d;
    g.p_in_smem = 0;
    return g;
  }
  // shared layout: [P (d*d) if p_in_smem][per-group q staging: groups_per_block * d]
  __host__ __device__ size_t p_smem_elems() const { return p_in_smem ? (size_t)d * d : 0; }

  __device__ __forceinline__ void block_init(void* smem, int nthreads) const {
    if (p_in_smem) {
      T* sP = (T*)smem;
      for (int i = threadIdx.x; i < d * d; i += nthreads) sP[i] = P[i];
      __syncthreads();
    }
  }

  template <int EPL, int LPC, bool WANT_V>
  __device__ __forceinline__ T kick(const T (&q)[EPL], T (&p)[EPL], T h, const GroupCtx<T>& c) const {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const T* Pm = p_in_smem ? (const T*)smem_raw : P;
    T* sq = c.smem_q;
    __syncwarp(c.mask);
#pragma unroll
    for (int s = 0; s < EPL; ++s) {
      const int i = Layout<EPL, LPC>::elem(c.lane, s);
      if (i < d) sq[i] = q[s];
    }
    __syncwarp(c.mask);
    T g[EPL];
#pragma unroll
    for (int s = 0; s < EPL; ++s) g[s] = T(0);
    for (int j = 0; j < d; ++j) {
      const T qj = sq[j];
      const T* row = Pm + (size_t)j * d;  // P symmetric: row j == column j
#pragma unroll
      for (int s = 0; s < EPL; ++s) {
        const int i = Layout<EPL, LPC>::elem(c.lane, s);
        const T pij = (i < d) ? row[i] : T(0);
        g[s] = fma(pij, qj, g[s]);
      }
    }
    T Vl = T(0);
#pragma unroll
    for (int s = 0; s < EPL; ++s) {
      if (WANT_V) Vl = fma(q[s], g[s], Vl);
      p[s] = fma(-h, g[s], p[s]);
    }
    if (WANT_V) {
#pragma unroll
      for (int o = LPC >> 1; o > 0; o >>= 1) Vl += __shfl_xor_sync(c.mask, Vl, o);
      return T(0.5) * Vl;
    }
    return T(0);
  }
};

}  // namespace hmcb
