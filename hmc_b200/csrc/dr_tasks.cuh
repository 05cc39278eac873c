// dr_tasks.cuh -- delayed-rejection HMC as resumable per-trajectory tasks in ONE persistent kernel.
//
// The staged form (transition.cuh: one kernel per stage over a compacted list) leaves the FP64 pipe
// idle in three ways: every stage ends in a partial last round of its persistent grid; inside a
// stage >= 2 a warp waits for the lane with the deepest ghost recursion (get_num,
// src/algorithms.py:474-497: 26.9 of 32 lanes active); and nothing overlaps the thin end of one
// stage with the next.  Here the unit of work is ONE leapfrog trajectory of one chain:
//
//   * a chain's walk through DRHMC.step (src/algorithms.py:500-530) and its ghost recursion is a
//     small state machine whose scalars live in per-chain scratch between trajectories (the same
//     explicit stack the staged kernel keeps in registers);
//   * trajectories are queued by length class j (L_j = int(nleap f^j) steps of eps / f^j) in one
//     ring per class, so the 32/LPC trajectories a warp integrates together always have the same
//     trip count -- no divergence in the leapfrog loop, whatever stage or ghost level they belong to;
//   * warps claim work themselves: queued trajectories first (chains already under way hold the
//     longest remaining critical path), fresh chains otherwise -- so first-stage work of other
//     chains fills what used to be the tail of the later stages;
//   * with K > 1 transitions per launch a finished chain is queued again for its next transition
//     (draws are keyed by chain and iteration, so results do not depend on the order in which chains
//     advance), the running moments sum(q), sum(q^2) are folded in, and the only tail left is the
//     one at the end of the K-th transition.
//
// Per-chain results are bit-identical to the staged kernels: each trajectory is integrated by the same
// lanes with the same arithmetic, and the bookkeeping below is the staged kernel's, statement by statement.
//
// Synchronisation: single-producer / single-consumer hand-over per ring entry.  A producer reserves an index
// with atomicAdd on the ring's tail, waits until that entry is empty (-1), fences, and stores the chain id; a
// consumer claims [head, head+n) with a CAS on the head, waits until each entry is filled, reads and empties
// it, and fences before touching the chain's scratch.  A chain sits in at most one ring at a time and every
// ring holds n_chains entries, so a ring never overflows.  All scratch is read with ld.global.cg (L2): the
// previous writer may have been another SM.  The grid is sized to be fully resident (no warp waits on a
// warp that is not running).
#pragma once
#include "transition.cuh"

namespace hmcb {

constexpr int kCtrlStride = 32;   // one counter per 128-byte line
enum { CTRL_FRESH = 0, CTRL_DONE = 1, CTRL_RING0 = 2 };   // ring r: head at CTRL_RING0 + 2r, tail at +1
enum { RING_EXIT = -2, RING_COUNTER = -1 };

template <typename T>
struct TaskArgs {
  TransArgs<T> t;
  int32_t K;             // transitions per chain in this launch
  T* msum;               // [n_chains, d] running sum of q after every transition, or nullptr
  T* msumsq;             // [n_chains, d] running sum of q*q
  int32_t* n_accept;     // [n_chains] += transitions accepted at any stage, or nullptr
  // scratch
  int32_t* ctrl;         // counters, kCtrlStride ints apart; zeroed by the host
  int32_t* rings;        // [nprops + 1][ring_cap]; ring j < nprops: trajectories of class j; ring nprops: chains
                         // starting a later transition (K > 1); all -1 at start
  int32_t ring_cap;
  int32_t* iter;         // [n_chains] transitions completed (K > 1); zeroed by the host
  int32_t* st_i;         // [n_chains][6 + 2 nprops]   J, lvl, Hcount, Vgcount, leapcount, -, lm[], lj[]
  T* st_f;               // [n_chains][2 nprops + 2 nprops^2]   lHb[], lpf[], lav[][], lpt[][]
  T* frames;             // [n_chains][(2 nprops - 1) d]: vector 0 the transition's p0, vectors 1+2l, 2+2l the
                         // proposal (q, p) parked by recursion level l
};

__device__ __forceinline__ int32_t ld_vol(const int32_t* p) { return *reinterpret_cast<const volatile int32_t*>(p); }
__device__ __forceinline__ void st_vol(int32_t* p, int32_t v) { *reinterpret_cast<volatile int32_t*>(p) = v; }

// chain state vectors through L2 (see the header comment); same addressing as load_vec / store_vec
template <typename T, int EPL, int LPC>
__device__ __forceinline__ void load_vec_cg(T (&x)[EPL], const T* base, int64_t off, int64_t sd, int lane, int d, T sign) {
  if (sd == 1) {
    const T* row = base + off + 4 * lane;
#pragma unroll
    for (int s = 0; s < EPL; ++s) {
      const int i = Layout<EPL, LPC>::elem(lane, s);
      x[s] = (i < d) ? sign * __ldcg(row + Layout<EPL, LPC>::elem(0, s)) : T(0);
    }
  } else {
#pragma unroll
    for (int s = 0; s < EPL; ++s) {
      const int i = Layout<EPL, LPC>::elem(lane, s);
      x[s] = (i < d) ? sign * __ldcg(base + off + i * sd) : T(0);
    }
  }
}

// sum over the chain's elements of (ref[i] - x[i]) with the partial sums of diff_sum (transition.cuh): the
// reference's "did not move" test; caller reduces over the group
template <typename T, int EPL, int LPC>
__device__ __forceinline__ T diff_sum_cg(const T (&x)[EPL], const T* base, int64_t off, int64_t sd, int lane, int d) {
  T acc[4] = {T(0), T(0), T(0), T(0)};
  const T* row = base + off + (sd == 1 ? 4 * lane : 0);
#pragma unroll
  for (int s0 = 0; s0 < EPL; s0 += 8) {
    T r[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int s = s0 + k;
      if (s < EPL) {
        const int i = Layout<EPL, LPC>::elem(lane, s);
        r[k] = (i < d) ? (sd == 1 ? __ldcg(row + Layout<EPL, LPC>::elem(0, s)) : __ldcg(base + off + i * sd)) : T(0);
      }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k)
      if (s0 + k < EPL) acc[k & 3] += r[k] - x[s0 + k];
  }
  return (acc[0] + acc[1]) + (acc[2] + acc[3]);
}

// sum += q, sumsq += q*q for one chain (rows of contiguous [C, d] arrays)
template <typename T, int EPL, int LPC>
__device__ __forceinline__ void fold_moments(const T (&q)[EPL], T* msum, T* msumsq, int64_t chain, int lane, int d) {
  T* r1 = msum + chain * d + 4 * lane;
  T* r2 = msumsq + chain * d + 4 * lane;
#pragma unroll
  for (int s = 0; s < EPL; ++s) {
    const int i = Layout<EPL, LPC>::elem(lane, s);
    if (i < d) {
      const int o = Layout<EPL, LPC>::elem(0, s);
      r1[o] = __ldcg(r1 + o) + q[s];
      r2[o] = fma(q[s], q[s], __ldcg(r2 + o));
    }
  }
}

// work counters of one transition: written for a chain's first transition of the launch, added for the later ones
__device__ __forceinline__ void add_counts(int32_t* dst, bool first, int cH, int cVg, int cLeap) {
  if (first) { dst[0] = cH; dst[1] = cVg; dst[2] = cLeap; }
  else { dst[0] = __ldcg(dst) + cH; dst[1] = __ldcg(dst + 1) + cVg; dst[2] = __ldcg(dst + 2) + cLeap; }
}

// lane 0 of a warp looks for work: a full warp of queued trajectories of one class (longest class first), else
// the next fresh chains, else chains starting a later transition, else whatever is queued; RING_EXIT when every
// chain has finished its K transitions.
__device__ __forceinline__ bool try_claim(int32_t* ctrl, int ring, int want, int at_least, uint32_t& base, int& n) {
  int32_t* head = ctrl + (CTRL_RING0 + 2 * ring) * kCtrlStride;
  const int32_t* tail = head + kCtrlStride;
  for (int attempt = 0; attempt < 3; ++attempt) {
    const int32_t h = ld_vol(head);
    const int32_t avail = ld_vol(tail) - h;
    if (avail < at_least) return false;
    const int take = avail < want ? avail : want;
    if (atomicCAS(head, h, h + take) == h) {
      base = (uint32_t)h;
      n = take;
      return true;
    }
  }
  return false;
}

__device__ __forceinline__ void pick_work(int32_t* ctrl, int nprops, int K, int per_warp, int64_t n_chains, int& ring,
                                          uint32_t& base, int& n) {
  for (;;) {
    for (int r = nprops - 1; r >= 0; --r)
      if (try_claim(ctrl, r, per_warp, per_warp, base, n)) { ring = r; return; }
    if ((int64_t)ld_vol(ctrl + CTRL_FRESH * kCtrlStride) < n_chains) {
      const int64_t b = atomicAdd(ctrl + CTRL_FRESH * kCtrlStride, per_warp);
      if (b < n_chains) {
        ring = RING_COUNTER;
        base = (uint32_t)b;
        n = (int)(n_chains - b < per_warp ? n_chains - b : per_warp);
        return;
      }
    }
    if (K > 1 && try_claim(ctrl, nprops, per_warp, per_warp, base, n)) { ring = nprops; return; }
    for (int r = nprops - 1; r >= 0; --r)
      if (try_claim(ctrl, r, per_warp, 1, base, n)) { ring = r; return; }
    if (K > 1 && try_claim(ctrl, nprops, per_warp, 1, base, n)) { ring = nprops; return; }
    if ((int64_t)ld_vol(ctrl + CTRL_DONE * kCtrlStride) >= n_chains) { ring = RING_EXIT; return; }
    __nanosleep(200);
  }
}

template <class Model, typename T, int EPL, int LPC, bool UNITM, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) k_dr_tasks(const Model model, const TaskArgs<T> A) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  using Lay = Layout<EPL, LPC>;
  constexpr int GPW = 32 / LPC;   // chains per warp
  const TransArgs<T>& a = A.t;
  const int d = a.d;
  const int np = a.nprops;

  model.block_init(smem_raw, THREADS);
  GroupCtx<T> c;
  c.lane = threadIdx.x & (LPC - 1);
  c.d = d;
  const int wlane = threadIdx.x & 31;
  const int g = wlane / LPC;
  c.mask = (LPC == 32) ? 0xffffffffu : (((1u << LPC) - 1u) << (g * LPC));
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

  // per-chain scratch: state scalars and parked vectors
  const int si = 6 + 2 * np, sf = 2 * np + 2 * np * np;
  const int64_t fstride = (int64_t)(2 * np - 1) * d;
  const uint32_t cap = (uint32_t)A.ring_cap;
  const int prob_rule = a.dr_prob;

  for (;;) {
    // ------------------------------------------------------------ claim work for this warp
    int ring = 0, n = 0;
    uint32_t base = 0;
    if (wlane == 0) pick_work(A.ctrl, np, A.K, GPW, a.n_chains, ring, base, n);
    ring = __shfl_sync(0xffffffffu, ring, 0);
    base = __shfl_sync(0xffffffffu, base, 0);
    n = __shfl_sync(0xffffffffu, n, 0);
    if (ring == RING_EXIT) break;

    const bool live = g < n;
    int next_ring = -1;       // where this group's chain goes next (lane 0 of the group decides)
    int64_t chain = 0;
    if (live) {
      if (ring == RING_COUNTER) {
        chain = (int64_t)base + g;
      } else {
        int32_t e = 0;
        if (c.lane == 0) {
          int32_t* slot = A.rings + (int64_t)ring * cap + (base + (uint32_t)g) % cap;
          while ((e = ld_vol(slot)) < 0) {}
          st_vol(slot, -1);
        }
        chain = (int64_t)__shfl_sync(c.mask, e, g * LPC);
      }
      __threadfence();   // what the chain's previous task wrote is visible from here on

      const int64_t qoff = chain * a.sc;
      int32_t* SI = A.st_i + chain * si;
      T* SF = A.st_f + chain * sf;
      T* FR = A.frames + chain * fstride;
#define ST_J SI[0]
#define ST_LVL SI[1]
#define ST_LM(l) SI[6 + (l)]
#define ST_LJ(l) SI[6 + np + (l)]
#define ST_LHB(l) SF[(l)]
#define ST_LPF(l) SF[np + (l)]
#define ST_LAV(l, i) SF[2 * np + (l) * np + (i)]
#define ST_LPT(l, i) SF[2 * np + np * np + (l) * np + (i)]

      const bool fresh = (ring == RING_COUNTER) || (ring == np);
      const int it = (fresh && A.K > 1) ? __ldcg(A.iter + chain) : 0;   // transitions this chain has completed
      int lvl = 0, j = 0;
      T q[EPL], p[EPL];
      T K0 = T(0);
      const T* qbase;      // where this trajectory's starting position lives (for the no-move test)
      int64_t qb_off, qb_sd;
      if (fresh) {
        qbase = (it == 0) ? a.q_in : a.q_out;
        qb_off = qoff; qb_sd = a.sd;
        load_vec_cg<T, EPL, LPC>(q, qbase, qoff, a.sd, c.lane, d, T(1));
        draw_momentum<T, EPL, LPC, UNITM>(p, a, minv, chain, c.lane, d, (uint32_t)it);
        if (a.p0_out) store_vec<T, EPL, LPC>(p, a.p0_out, qoff, a.sd, c.lane, d);
        if (np > 1) store_vec<T, EPL, LPC>(p, FR, 0, 1, c.lane, d);     // later stages restart from (q0, p0)
        K0 = kinetic<T, EPL, LPC, UNITM>(p, minv, c.mask);
      } else {
        lvl = __ldcg(&ST_LVL);
        j = __ldcg(&ST_LJ(lvl));          // == ring
        if (lvl == 0) {
          qbase = a.q_out; qb_off = qoff; qb_sd = a.sd;
          load_vec_cg<T, EPL, LPC>(q, a.q_out, qoff, a.sd, c.lane, d, T(1));
          load_vec_cg<T, EPL, LPC>(p, FR, 0, 1, c.lane, d, T(1));
        } else {
          qbase = FR + (int64_t)(1 + 2 * (lvl - 1)) * d; qb_off = 0; qb_sd = 1;
          load_vec_cg<T, EPL, LPC>(q, qbase, 0, 1, c.lane, d, T(1));
          load_vec_cg<T, EPL, LPC>(p, qbase + d, 0, 1, c.lane, d, T(-1));   // the ghost chain starts from the flipped proposal
        }
      }

      const int N = a.L[j] > 1 ? a.L[j] : 1;
      T Vs, Ve;
      leapfrog<Model, T, EPL, LPC, UNITM>(model, q, p, minv, N, a.eps[j], c, Vs, Ve);
      const T Hn = Ve + kinetic<T, EPL, LPC, UNITM>(p, minv, c.mask);
      T dq = diff_sum_cg<T, EPL, LPC>(q, qbase, qb_off, qb_sd, c.lane, d);
      dq = group_sum_masked<LPC, T>(dq, c.mask);

      // what happens to the vectors afterwards is decided by lane 0 of the group
      enum { ACT_NONE = 0, ACT_ACCEPT_REGS = 1, ACT_ACCEPT_FRAME = 2, ACT_REJECT = 3 };
      int action = ACT_NONE;
      int iter_now = it;

      if (fresh) {
        // ---------------- stage 0 of DRHMC.step (src/algorithms.py:506-529 at j = 0), as finish_single
        const T H0 = Vs + K0;
        T prob = t_exp<T>(H0 - Hn);
        const T pf_raw = prob;
        if (dq == T(0)) prob = T(0);
        if (t_bad(prob)) prob = T(0);
        const T a0 = prob < T(1) ? prob : T(1);
        T pt0 = T(1);
        if (prob_rule && !t_bad(pf_raw)) pt0 = T(1) - (prob_rule == 1 ? (pf_raw < T(1) ? pf_raw : T(1)) : a0);
        const double u = draw_uniform<T>(a, chain, 0, (uint32_t)it);
        const bool accept = a.dr_conv ? (u <= (double)a0) : (u < (double)a0);
        bool finished = accept || np == 1;
        int code = a.dr_conv ? (accept ? 0 : -1) : (accept ? 1 : 0);
        const int cH = 2, cVg = N + 1, cLeap = 1;
        if (!finished && prob_rule) {
          // one more uniform decides whether stage 1 is attempted at all (src/algorithms.py:710-711)
          const double u_try = draw_uniform<T>(a, chain, 1, (uint32_t)it);
          if (u_try > (double)pt0) { finished = true; code = -2; }
        }
        if (c.lane == 0) {
          if (finished) {
            if (a.avec_out) {
              a.avec_out[chain * np] = a0;
              for (int i = 1; i < np; ++i) a.avec_out[chain * np + i] = T(0);
            }
            if (a.ptries_out) {
              a.ptries_out[chain * np] = pt0;
              for (int i = 1; i < np; ++i) a.ptries_out[chain * np + i] = T(1);
            }
            if (a.code) a.code[chain] = code;
            if (a.counts) add_counts(a.counts + 3 * chain, it == 0, cH, cVg, cLeap);
          } else {
            ST_J = 1; ST_LVL = 0; SI[2] = cH; SI[3] = cVg; SI[4] = cLeap; SI[5] = it;
            ST_LM(0) = 2; ST_LJ(0) = 1;
            ST_LHB(0) = H0;
            for (int i = 0; i < np; ++i) { ST_LAV(0, i) = T(0); ST_LPT(0, i) = T(1); }
            ST_LAV(0, 0) = a0;
            if (prob_rule) ST_LPT(0, 0) = pt0;
            next_ring = 1;
          }
        }
        if (!accept && it == 0 && a.q_out != a.q_in) {
          // the later stages and the final state read q0 from q_out
#pragma unroll
          for (int s = 0; s < EPL; ++s) {
            const int i = Lay::elem(c.lane, s);
            if (i < d) a.q_out[qoff + i * a.sd] = __ldcg(a.q_in + qoff + i * a.sd);
          }
        }
        if (finished) action = accept ? ACT_ACCEPT_REGS : ACT_REJECT;
      } else {
        // ---------------- a trajectory of stage J >= 1: level 0 is DRHMC.step's own loop at j = J, level l >= 1 a
        // get_num frame (src/algorithms.py:474-497) whose base is the proposal parked by level l-1, flipped
        iter_now = __ldcg(&SI[5]);
        if (j > 0) {
          // this proposal is the base of a ghost chain: park it
          T* fq = FR + (int64_t)(1 + 2 * lvl) * d;
          store_vec<T, EPL, LPC>(q, fq, 0, 1, c.lane, d);
          store_vec<T, EPL, LPC>(p, fq + d, 0, 1, c.lane, d);
        }
        if (c.lane == 0) {
          const int J = __ldcg(&ST_J);
          int cH = __ldcg(&SI[2]) + 1, cVg = __ldcg(&SI[3]) + N + 1, cLeap = __ldcg(&SI[4]) + 1;
          T pf = t_exp<T>(__ldcg(&ST_LHB(lvl)) - Hn);
          // rule 1 inside a ghost frame takes p_try from the raw ratio (src/algorithms.py:668-669), at level 0
          // from the ratio after the no-move test (:714-719)
          if (prob_rule == 1 && lvl > 0) ST_LPT(lvl, j) = t_bad(pf) ? T(1) : T(1) - (pf < T(1) ? pf : T(1));
          if (dq == T(0)) pf = T(0);
          if (prob_rule == 1 && lvl == 0) ST_LPT(0, j) = t_bad(pf) ? T(1) : T(1) - (pf < T(1) ? pf : T(1));

          if (j > 0) {
            ST_LPF(lvl) = pf;
            ++lvl;
            if (a.dr_conv) cH += 1;   // PyHMC_multistep.get_num re-evaluates H at its base (src/pyhmc.py:541)
            ST_LM(lvl) = j; ST_LJ(lvl) = 0; ST_LHB(lvl) = Hn;
            for (int i = 0; i < np; ++i) { ST_LAV(lvl, i) = T(0); ST_LPT(lvl, i) = T(1); }
            ST_LVL = lvl;
            next_ring = 0;
          } else {
            // j == 0 only happens inside a ghost frame here: unwind
            T prob = pf, pf_cur = pf, prob_top = T(0);
            bool done = false;
            for (;;) {
              T ret;
              bool returning;
              const int lm = __ldcg(&ST_LM(lvl));
              int lj = __ldcg(&ST_LJ(lvl));
              if (t_bad(prob)) {
                ret = T(0);
                returning = true;
              } else {
                const T al = prob < T(1) ? prob : T(1);
                ST_LAV(lvl, lj) = al;
                if (prob_rule == 2) ST_LPT(lvl, lj) = t_bad(pf_cur) ? T(1) : T(1) - al;   // src/algorithms.py:764-765
                T pr = T(1);
                for (int i = 0; i < lm; ++i) pr *= (T(1) - (i == lj ? al : __ldcg(&ST_LAV(lvl, i))));
                if (pr == T(0)) { ret = pr; returning = true; }
                else if (lj + 1 == lm) {
                  if (prob_rule) {
                    T pt = T(1);
                    for (int i = 0; i < lm; ++i) pt *= __ldcg(&ST_LPT(lvl, i));
                    pr *= pt;                                                          // prod(1-avec) * prod(ptries)
                  }
                  ret = pr; returning = true;
                }
                else { lj += 1; ST_LJ(lvl) = lj; returning = false; }
              }
              if (!returning) { next_ring = lj; break; }
              --lvl;
              const int ljp = __ldcg(&ST_LJ(lvl));
              T den = T(1);
              for (int i = 0; i < ljp; ++i) den *= (T(1) - __ldcg(&ST_LAV(lvl, i)));
              if (prob_rule) {
                T pt = T(1);
                for (int i = 0; i < ljp; ++i) pt *= __ldcg(&ST_LPT(lvl, i));
                den *= pt;
              }
              pf_cur = __ldcg(&ST_LPF(lvl));
              prob = pf_cur * ret / den;
              if (lvl == 0) { prob_top = prob; done = true; break; }
            }
            ST_LVL = lvl;
            if (done) {
              // ---------------- DRHMC.step's accept test for stage J (src/algorithms.py:523-529)
              if (t_bad(prob_top)) prob_top = T(0);
              const T aJ = prob_top < T(1) ? prob_top : T(1);
              if (prob_rule == 2) ST_LPT(0, J) = t_bad(__ldcg(&ST_LPF(0))) ? T(1) : T(1) - aJ;   // src/algorithms.py:808-809
              const double u = draw_uniform<T>(a, chain, prob_rule ? 2 * J : J, (uint32_t)iter_now);
              const bool accept = u < (double)aJ;
              bool finished = accept || (J == np - 1);
              int code = a.dr_conv ? (accept ? J : -1) : (accept ? J + 1 : 0);
              if (!finished && prob_rule) {
                const T ptJ = __ldcg(&ST_LPT(0, J));
                const double u_try = draw_uniform<T>(a, chain, 2 * J + 1, (uint32_t)iter_now);
                if (u_try > (double)ptJ) { finished = true; code = -2; }
              }
              if (finished) {
                if (a.avec_out) {
                  for (int i = 0; i < np; ++i)
                    a.avec_out[chain * np + i] = (i < J) ? __ldcg(&ST_LAV(0, i)) : (i == J ? aJ : T(0));
                }
                if (a.ptries_out) {
                  for (int i = 0; i < np; ++i) a.ptries_out[chain * np + i] = __ldcg(&ST_LPT(0, i));
                }
                if (a.code) a.code[chain] = code;
                if (a.counts) add_counts(a.counts + 3 * chain, iter_now == 0, cH, cVg, cLeap);
                action = accept ? ACT_ACCEPT_FRAME : ACT_REJECT;
              } else {
                ST_LAV(0, J) = aJ;
                ST_J = J + 1;
                ST_LM(0) = J + 2; ST_LJ(0) = J + 1;
                next_ring = J + 1;
              }
            }
          }
          SI[2] = cH; SI[3] = cVg; SI[4] = cLeap;
        }
        action = __shfl_sync(c.mask, action, g * LPC);
      }

      // ---------------- the end of a transition: new state, returned momentum, running moments, next transition
      if (action != ACT_NONE) {
        if (action == ACT_ACCEPT_FRAME) {
          // the accepted proposal of stage J was parked by level 0
          load_vec_cg<T, EPL, LPC>(q, FR + d, 0, 1, c.lane, d, T(1));
          load_vec_cg<T, EPL, LPC>(p, FR + 2 * d, 0, 1, c.lane, d, T(1));
        }
        if (action != ACT_REJECT) {
          store_vec<T, EPL, LPC>(q, a.q_out, qoff, a.sd, c.lane, d);
          if (a.p_out) store_vec<T, EPL, LPC>(p, a.p_out, qoff, a.sd, c.lane, d);
        } else {
          if (A.msum) load_vec_cg<T, EPL, LPC>(q, a.q_out, qoff, a.sd, c.lane, d, T(1));
          if (a.p_out) {
            // the step returns the momentum it started from (src/algorithms.py:530)
            if (np > 1) {
              load_vec_cg<T, EPL, LPC>(p, FR, 0, 1, c.lane, d, T(1));
            } else {
              draw_momentum<T, EPL, LPC, UNITM>(p, a, minv, chain, c.lane, d, (uint32_t)iter_now);
            }
            store_vec<T, EPL, LPC>(p, a.p_out, qoff, a.sd, c.lane, d);
          }
        }
        if (A.msum) fold_moments<T, EPL, LPC>(q, A.msum, A.msumsq, chain, c.lane, d);
        if (c.lane == 0) {
          if (A.n_accept && action != ACT_REJECT) A.n_accept[chain] = __ldcg(A.n_accept + chain) + 1;
          if (iter_now + 1 < A.K) {
            A.iter[chain] = iter_now + 1;
            next_ring = np;
          } else {
            atomicAdd(A.ctrl + CTRL_DONE * kCtrlStride, 1);
          }
        }
      }
#undef ST_J
#undef ST_LVL
#undef ST_LM
#undef ST_LJ
#undef ST_LHB
#undef ST_LPF
#undef ST_LAV
#undef ST_LPT
    }
    __threadfence();   // every lane's writes to the chain's scratch, before lane 0 of its group hands the chain on
    __syncwarp();

    // ------------------------------------------------------------ hand the chains on
    for (int r = 0; r <= np; ++r) {
      const bool mine = live && c.lane == 0 && next_ring == r;
      const unsigned m = __ballot_sync(0xffffffffu, mine);
      if (m == 0) continue;
      uint32_t t0 = 0;
      const int leader = __ffs(m) - 1;
      if (wlane == leader) t0 = (uint32_t)atomicAdd(A.ctrl + (CTRL_RING0 + 2 * r + 1) * kCtrlStride, __popc(m));
      t0 = __shfl_sync(0xffffffffu, t0, leader);
      if (mine) {
        int32_t* slot = A.rings + (int64_t)r * cap + (t0 + (uint32_t)__popc(m & ((1u << wlane) - 1u))) % cap;
        while (ld_vol(slot) >= 0) {}     // the entry's previous occupant has been claimed; its reader empties it
        st_vol(slot, (int32_t)chain);
      }
    }
    __syncwarp();
  }
}

}  // namespace hmcb
