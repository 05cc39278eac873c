// launch.cuh -- host side: kernel-configuration registry and launch helpers.
//
// Kernels are compiled for a fixed set of (EPL slots per lane, LPC lanes per
// chain) shapes; at run time the smallest shape that holds the model's d is
// picked (least padding, then fewest lanes).  Each inst_*.cu registers the
// shapes it compiled for one (model, dtype).
#pragma once
#include <vector>

#include "extra_kernels.cuh"
#include "transition_ws.cuh"

namespace hmcb {

enum Op { OP_HMC = 0, OP_DR0 = 1, OP_DRN = 2, OP_LEAPFROG = 3, OP_LOGPGRAD = 4, OP_QUERY = 5, OP_MALA = 6, OP_ADHMC = 7,
          OP_QUERY_AD = 8, OP_GHMC = 9, OP_ADWALK = 10, OP_QUERY_WALK = 11, OP_HMC_MULTI = 12, OP_DR0_MULTI = 13, OP_GHMC_MULTI = 14 };
constexpr int kGhmcMaxEpl = 32;  // GHMC keeps q, p and the gradient: 3*EPL values per lane
constexpr int kMalaMaxEpl = 28;  // MALA keeps q, grad, q1, grad1 in registers: 4*EPL values per lane

struct LaunchInfo {
  int threads;        // block size
  int max_blocks;     // resident blocks on the whole device for this kernel
  int epl, lpc;
};

// args is TransArgs<T>* for the transition ops, AuxArgs<T>* for the others, LaunchInfo* for OP_QUERY
using LaunchFn = int (*)(const hmcb_model& hm, int op, void* args, cudaStream_t st);

struct Entry {
  int kind, dtype, unitm, epl, lpc;
  LaunchFn fn;
};

std::vector<Entry>& registry();
struct Registrar {
  explicit Registrar(const Entry& e) { registry().push_back(e); }
};

// pick the compiled shape for this model; nullptr (+ error set) if none fits
const Entry* choose_entry(const hmcb_model& hm, int max_epl = 1 << 30);

// K HMC transitions in one launch (k_hmc_multi_ws); `applied` tells the caller whether this shape has that kernel
template <typename T>
struct MultiHmcArgs {
  TransArgs<T> a;
  int32_t K;
  int32_t* n_accept;
  int32_t applied;
};

// one pass of the multi-pass sample mode for short trajectories (k_dr0_multi_ws); query_only asks whether this shape has it
template <typename T>
struct MultiDrArgs {
  TransArgs<T> a;
  int32_t max_steps;
  const int32_t* list;
  const int32_t* list_count;
  int32_t* iter_of;
  int32_t query_only;
  int32_t applied;
};

// leapfrog-only / logp+grad kernels share the transition kernel's device code
template <typename T>
struct AuxArgs {
  int64_t n_chains;
  int32_t d;
  const T* q_in;
  T* q_out;
  int64_t sc, sd;
  const T* p_in;
  T* p_out;
  T* logp;
  T* grad;
  const T* minv;
  int32_t n_leap;
  T eps;
};

template <class Model, typename T, int EPL, int LPC, bool UNITM, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) k_aux(const Model model, const AuxArgs<T> a, const int op) {
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
  const int64_t total_groups = (int64_t)gridDim.x * GPB;
  for (int64_t chain = ((int64_t)blockIdx.x * THREADS + threadIdx.x) / LPC; chain < a.n_chains; chain += total_groups) {
    const int64_t off = chain * a.sc;
    T q[EPL], p[EPL];
    load_vec<T, EPL, LPC>(q, a.q_in, off, a.sd, c.lane, d, T(1));
    if (op == OP_LEAPFROG) {
      load_vec<T, EPL, LPC>(p, a.p_in, off, a.sd, c.lane, d, T(1));
      T Vs, Ve;
      leapfrog<Model, T, EPL, LPC, UNITM>(model, q, p, minv, a.n_leap > 1 ? a.n_leap : 1, a.eps, c, Vs, Ve);
      store_vec<T, EPL, LPC>(q, a.q_out, off, a.sd, c.lane, d);
      store_vec<T, EPL, LPC>(p, a.p_out, off, a.sd, c.lane, d);
    } else {
      // p = 0 - 1*gradV = grad log_prob; V returned
#pragma unroll
      for (int s = 0; s < EPL; ++s) p[s] = T(0);
      const T V = model.template kick<EPL, LPC, true>(q, p, T(1), c);
      if (a.grad) store_vec<T, EPL, LPC>(p, a.grad, off, a.sd, c.lane, d);
      if (a.logp && c.lane == 0) a.logp[chain] = -V;
    }
  }
}

template <typename T, int EPL, bool UNITM, bool GAUSS>
constexpr int min_blocks_for() {
  // 32-bit registers: q, p (+ metric, + gradient accumulator) plus working set
  constexpr int words = (2 * EPL + (UNITM ? 0 : EPL) + (GAUSS ? EPL : 0)) * (int)(sizeof(T) / 4) + 56;
  constexpr int mb = 65536 / (words * 128);
  return mb < 1 ? 1 : (mb > 8 ? 8 : mb);
}

int max_optin_smem();  // per-block opt-in shared memory limit of the current device

// moments.cu: the per-transition bookkeeping of a sample loop (running moments, summed work counters, accept counts)
int launch_post_transition(int32_t dtype, int64_t n_chains, int32_t d, const void* q, int64_t sc, int64_t sd, void* sum,
                           void* sumsq, const int32_t* counts_k, int32_t* counts_tot, int first, const int32_t* code,
                           int min_accept_code, int max_accept_code, int32_t* n_accept, cudaStream_t st);

int launch_scale_counts(int32_t* counts, int64_t n_chains, int K, cudaStream_t st);

// gauss_dmma.cu: dense-Gaussian HMC transition on the FP64 tensor path; returns 1 when it does not apply
int gauss_dmma_hmc(const hmcb_model& hm, const TransArgs<double>& a, cudaStream_t st);

template <class Model, typename T, int EPL, int LPC, bool UNITM, bool GAUSS>
int launch_any(const hmcb_model& hm, int op, void* args, cudaStream_t st) {
  constexpr int THREADS = 128;
  constexpr int MINB = min_blocks_for<T, EPL, UNITM, GAUSS>();
  // the step-size kernels hold a base state beside the running one and spill it in their trial sweeps at the register budget
  // of the plain transition: one block per SM fewer (measured, 131,072-262,144 chains, AdHMC_eps0 / AdHMC walk in ms --
  // Rosenbrock d=100: 1.20 -> 0.98 / 2.40 -> 2.08, d=20: 0.163 -> 0.143 / 0.344 -> 0.307; funnel d=10: 0.168 -> 0.155 /
  // 0.383 -> 0.330, d=50: 0.859 -> 0.853 / 1.776 -> 1.741, d=100: 0.913 -> 0.937 / 1.821 -> 1.808)
  constexpr int ADMINB = (MINB >= 3) ? MINB - 1 : MINB;
  constexpr int GPB = THREADS / LPC;
  Model model = Model::make(hm);
  size_t smem = 0;
  if constexpr (GAUSS) {
    // per-group q staging, plus P itself when the block can hold it
    const size_t stage = (size_t)GPB * hm.d * sizeof(T);
    const size_t withP = stage + (size_t)hm.d * hm.d * sizeof(T);
    model.p_in_smem = (withP <= (size_t)max_optin_smem()) ? 1 : 0;
    smem = model.p_in_smem ? withP : stage;
  }

  auto prep_smem = [&](auto kernel, int& max_blocks, size_t bytes) -> int {
    if (bytes > 48 * 1024) HMCB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    int per_sm = 0;
    HMCB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, THREADS, bytes));
    if (per_sm < 1) {
      set_error("kernel shape EPL=%d LPC=%d does not fit on an SM (smem %zu B)", EPL, LPC, bytes);
      return HMCB_E_UNSUPPORTED;
    }
    max_blocks = per_sm * sm_count();
    return 0;
  };
  auto prep = [&](auto kernel, int& max_blocks) -> int { return prep_smem(kernel, max_blocks, smem); };
  // per-warp staging tile of the single-trajectory kernels (transition.cuh, warp_tile_copy): one row of d|1
  // elements per chain of the block; only for row-major state and models that use no shared memory themselves
  constexpr size_t kTileBytes = GAUSS ? 0 : (size_t)GPB * ((EPL * LPC) | 1) * sizeof(T);
  // measured on the funnel, 2^20 chains, L=20: d=50 0.81 -> 0.63 ms with the tile, d=10 0.13 -> 0.18 ms without
  // it paying (a 80-byte row per lane already uses most of every sector it touches), hence the lower bound on d
  auto use_tile = [&](const TransArgs<T>& a) -> bool {
    return !GAUSS && a.d >= 24 && a.sd == 1 && a.sc == a.d && !getenv("HMCB_NO_TILE");
  };
  auto grid_for = [&](int64_t items, int max_blocks) -> int {
    int64_t need = (items + GPB - 1) / GPB;
    if (need < 1) need = 1;
    return (int)(need < max_blocks ? need : max_blocks);
  };

  int max_blocks = 0, rc;
  // thread-per-chain float64 shapes with a long RNG prologue: warp-specialised kernel (transition_ws.cuh)
  if constexpr (LPC == 1 && UNITM && sizeof(T) == 8 && EPL >= 20 && !GAUSS) {
    if ((op == OP_HMC || op == OP_DR0) && !getenv("HMCB_NO_WS")) {
      TransArgs<T>& a = *(TransArgs<T>*)args;
      // (not with a partial momentum refresh: at nleap=20 the plain kernel with its staging tile measured
      // 1.25 ms against 1.40 ms here, the strided p_prev / p_out rows being what costs)
      // (nor with the exact f64 normal generator: the helper warpgroup draws with normal4_f32 only, and every
      // other place that re-draws p0 -- the DR stages >= 1, p_out of rejected chains -- would use the f64 one)
      // (nor inside hmcb_drhmc_sample, prefer_plain: a warp-specialised block owns a whole SM's registers and cannot share
      // an SM with the 2-per-SM blocks of another range's later stages -- with 4 ranges pipelined the plain kernel's
      // packing wins: 5.34 -> 5.12 ms per transition at 2^20 chains although stage 0 alone is 8 % slower.  A two-blocks-
      // per-SM form of the warp-specialised kernel (128 compute threads at 232 registers + 128 helpers at 24: no spills in
      // the loop) was tried in its place: 4.939 against 4.945 ms -- inside the pipeline the other ranges' blocks on the
      // same SM already cover a block's momentum prologue, so it is not kept)
      if (!a.p_inject && !a.p_prev && !a.prefer_plain && !(a.rng_flags & HMCB_RNG_NORMAL_F64) && a.n_chains >= 4096) {
        // (the per-warp staging tile of the plain kernel does not pay here: L=150 2.02 -> 2.18 ms.)  tile_io instead
        // selects a shared-memory copy of q0 for the no-move test and the copy-through of rejected chains
        // (funnel d=50, 2^20 chains: HMC L=20 0.748 -> 0.614 ms, L=150 2.055 -> 2.002 ms, DR stage 0 1.975 -> 1.948 ms)
        const size_t q0s_bytes = (size_t)EPL * kWsComputeThreads * sizeof(T);
        const size_t base_smem = 2 * (size_t)EPL * kWsComputeThreads * sizeof(float) + 64;
        // vanilla HMC only: in DR stage 0 the rejected chains are not copied through, and what the no-move test saves
        // is less than what the smaller L1 (204 KB of shared memory per SM) costs the strided row accesses
        // (DRHMC nleap=20: 0.787 -> 0.815 ms with the copy)
        // (measured again with HMCB_WS_Q0S=2: DRHMC nleap=20 0.778 -> 0.801 ms)
        const int q0s_env = getenv("HMCB_WS_Q0S") ? atoi(getenv("HMCB_WS_Q0S")) : 1;     // 0 off, 1 vanilla HMC only, 2 DR stage 0 too
        a.tile_io = ((op == OP_HMC || q0s_env == 2) && q0s_env != 0 && base_smem + q0s_bytes <= (size_t)max_optin_smem()) ? 1 : 0;
        const size_t ws_smem = base_smem + (a.tile_io ? q0s_bytes : 0);
        const int grid = (int)((a.n_chains + kWsComputeThreads - 1) / kWsComputeThreads < sm_count()
                                   ? (a.n_chains + kWsComputeThreads - 1) / kWsComputeThreads
                                   : sm_count());
        if (op == OP_HMC) {
          auto k = k_transition_ws<Model, T, EPL, MODE_HMC>;
          HMCB_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ws_smem));
          k<<<grid, kWsThreads, ws_smem, st>>>(model, a);
        } else {
          auto k = k_transition_ws<Model, T, EPL, MODE_DR0>;
          HMCB_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ws_smem));
          k<<<grid, kWsThreads, ws_smem, st>>>(model, a);
        }
        HMCB_CUDA(cudaPeekAtLastError());
        return 0;
      }
    }
  }
  if (op == OP_DR0_MULTI) {
    MultiDrArgs<T>& m = *(MultiDrArgs<T>*)args;
    m.applied = 0;
    if constexpr (LPC == 1 && UNITM && sizeof(T) == 8 && EPL >= 20 && !GAUSS) {
      const TransArgs<T>& a = m.a;
      const size_t bytes = (size_t)EPL * kWsComputeThreads * (sizeof(T) + 2 * sizeof(float)) + 64;
      if (!a.p_inject && !a.u_inject && !a.p_prev && !a.p0_out && !(a.rng_flags & HMCB_RNG_NORMAL_F64) &&
          bytes <= (size_t)max_optin_smem() && !getenv("HMCB_NO_WS")) {
        m.applied = 1;
        if (!m.query_only) {
          auto k = k_dr0_multi_ws<Model, T, EPL>;
          HMCB_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
          const int64_t need = (a.n_chains + kWsComputeThreads - 1) / kWsComputeThreads;
          const int grid = (int)(need < sm_count() ? need : sm_count());
          k<<<grid, kWsThreads, bytes, st>>>(model, a, m.max_steps, m.list, m.list_count, m.iter_of);
          HMCB_CUDA(cudaPeekAtLastError());
        }
      }
    }
    return 0;
  }
  if (op == OP_HMC_MULTI) {
    MultiHmcArgs<T>& m = *(MultiHmcArgs<T>*)args;
    m.applied = 0;
    if constexpr (LPC == 1 && UNITM && sizeof(T) == 8 && EPL >= 20 && !GAUSS) {
      const TransArgs<T>& a = m.a;
      const size_t bytes = (size_t)EPL * kWsComputeThreads * (sizeof(T) + 2 * sizeof(float)) + 64;
      if (!a.p_inject && !a.u_inject && !a.p_prev && !a.p0_out && !(a.rng_flags & HMCB_RNG_NORMAL_F64) && a.n_chains >= 4096 &&
          bytes <= (size_t)max_optin_smem() && !getenv("HMCB_NO_WS")) {
        auto k = k_hmc_multi_ws<Model, T, EPL>;
        HMCB_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
        const int64_t need = (a.n_chains + kWsComputeThreads - 1) / kWsComputeThreads;
        const int grid = (int)(need < sm_count() ? need : sm_count());
        k<<<grid, kWsThreads, bytes, st>>>(model, a, m.K, m.n_accept);
        HMCB_CUDA(cudaPeekAtLastError());
        m.applied = 1;
      }
    }
    if constexpr (!GAUSS) {
      if (!m.applied) {
        const TransArgs<T>& a = m.a;
        if (!a.p_inject && !a.u_inject && !a.p_prev && !a.p0_out && !getenv("HMCB_NO_MULTI")) {
          auto k = k_hmc_multi<Model, T, EPL, LPC, UNITM, THREADS, MINB>;
          const size_t bytes = (size_t)EPL * THREADS * sizeof(T);
          if ((rc = prep_smem(k, max_blocks, bytes))) return rc;
          k<<<grid_for(a.n_chains, max_blocks), THREADS, bytes, st>>>(model, a, m.K, m.n_accept);
          HMCB_CUDA(cudaPeekAtLastError());
          m.applied = 1;
        }
      }
    }
    return 0;
  }
  switch (op) {
    case OP_HMC: {
      auto k = k_transition<Model, T, EPL, LPC, UNITM, MODE_HMC, THREADS, MINB>;
      TransArgs<T>& a = *(TransArgs<T>*)args;
      a.tile_io = use_tile(a) ? 1 : 0;
      const size_t bytes = smem + (a.tile_io ? kTileBytes : 0);
      if ((rc = prep_smem(k, max_blocks, bytes))) return rc;
      k<<<grid_for(a.n_chains, max_blocks), THREADS, bytes, st>>>(model, a);
      break;
    }
    case OP_DR0: {
      auto k = k_transition<Model, T, EPL, LPC, UNITM, MODE_DR0, THREADS, MINB>;
      TransArgs<T>& a = *(TransArgs<T>*)args;
      a.tile_io = use_tile(a) ? 1 : 0;
      const size_t bytes = smem + (a.tile_io ? kTileBytes : 0);
      if ((rc = prep_smem(k, max_blocks, bytes))) return rc;
      k<<<grid_for(a.n_chains, max_blocks), THREADS, bytes, st>>>(model, a);
      break;
    }
    case OP_DRN: {
      auto k = k_transition<Model, T, EPL, LPC, UNITM, MODE_DRN, THREADS, MINB>;
      if ((rc = prep(k, max_blocks))) return rc;
      TransArgs<T>& a = *(TransArgs<T>*)args;
      // the list length lives on the device: launch the persistent grid the frames were sized for
      const int grid = (int)(a.ws.frame_threads / THREADS);
      k<<<grid, THREADS, smem, st>>>(model, a);
      break;
    }
    case OP_LEAPFROG:
    case OP_LOGPGRAD: {
      auto k = k_aux<Model, T, EPL, LPC, UNITM, THREADS, MINB>;
      if ((rc = prep(k, max_blocks))) return rc;
      AuxArgs<T>& a = *(AuxArgs<T>*)args;
      k<<<grid_for(a.n_chains, max_blocks), THREADS, smem, st>>>(model, a, op);
      break;
    }
    case OP_MALA: {
      if constexpr (EPL <= kMalaMaxEpl) {
        constexpr int MB = min_blocks_for<T, 2 * EPL, true, false>();
        auto k = k_mala<Model, T, EPL, LPC, THREADS, MB>;
        if ((rc = prep(k, max_blocks))) return rc;
        MalaArgs<T>& a = *(MalaArgs<T>*)args;
        k<<<grid_for(a.n_chains, max_blocks), THREADS, smem, st>>>(model, a);
        break;
      } else {
        set_error("internal: MALA is not compiled for EPL=%d", EPL);
        return HMCB_E_UNSUPPORTED;
      }
    }
    case OP_GHMC: {
      if constexpr (EPL <= kGhmcMaxEpl) {
        constexpr int MB = min_blocks_for<T, (3 * EPL + 1) / 2, UNITM, GAUSS>();
        auto k = k_ghmc<Model, T, EPL, LPC, UNITM, THREADS, MB>;
        if ((rc = prep(k, max_blocks))) return rc;
        GhmcArgs<T>& a = *(GhmcArgs<T>*)args;
        k<<<grid_for(a.t.n_chains, max_blocks), THREADS, smem, st>>>(model, a);
        break;
      } else {
        set_error("internal: GHMC is not compiled for EPL=%d", EPL);
        return HMCB_E_UNSUPPORTED;
      }
    }
    case OP_GHMC_MULTI: {
      GhmcArgs<T>& a = *(GhmcArgs<T>*)args;
      a.applied = 0;
      if constexpr (EPL <= kGhmcMaxEpl && !GAUSS) {
        // q0, the refreshed momentum and the cached gradient of the running transition: 3*EPL values per thread
        const size_t bytes = 3 * (size_t)EPL * THREADS * sizeof(T);
        if (bytes <= (size_t)max_optin_smem() && !getenv("HMCB_NO_MULTI")) {
          constexpr int MB = min_blocks_for<T, (3 * EPL + 1) / 2, UNITM, GAUSS>();
          auto k = k_ghmc_multi<Model, T, EPL, LPC, UNITM, THREADS, MB>;
          if ((rc = prep_smem(k, max_blocks, bytes))) return rc;
          k<<<grid_for(a.t.n_chains, max_blocks), THREADS, bytes, st>>>(model, a);
          a.applied = 1;
        }
      }
      break;
    }
    case OP_ADHMC: {
      auto k = k_adhmc_eps0<Model, T, EPL, LPC, UNITM, THREADS, ADMINB>;
      if ((rc = prep(k, max_blocks))) return rc;
      AdArgs<T>& a = *(AdArgs<T>*)args;
      k<<<(int)(a.frame_threads / THREADS), THREADS, smem, st>>>(model, a);
      break;
    }
    case OP_QUERY_AD: {
      auto k = k_adhmc_eps0<Model, T, EPL, LPC, UNITM, THREADS, ADMINB>;
      if ((rc = prep(k, max_blocks))) return rc;
      LaunchInfo* li = (LaunchInfo*)args;
      li->threads = THREADS;
      li->max_blocks = max_blocks;
      li->epl = EPL;
      li->lpc = LPC;
      return 0;
    }
    case OP_ADWALK: {
      auto k = k_adhmc_walk<Model, T, EPL, LPC, UNITM, THREADS, ADMINB>;
      if ((rc = prep(k, max_blocks))) return rc;
      WalkArgs<T>& a = *(WalkArgs<T>*)args;
      k<<<(int)(a.frame_threads / THREADS), THREADS, smem, st>>>(model, a);
      break;
    }
    case OP_QUERY_WALK: {
      auto k = k_adhmc_walk<Model, T, EPL, LPC, UNITM, THREADS, ADMINB>;
      if ((rc = prep(k, max_blocks))) return rc;
      LaunchInfo* li = (LaunchInfo*)args;
      li->threads = THREADS;
      li->max_blocks = max_blocks;
      li->epl = EPL;
      li->lpc = LPC;
      return 0;
    }
    case OP_QUERY: {
      auto k = k_transition<Model, T, EPL, LPC, UNITM, MODE_DRN, THREADS, MINB>;
      if ((rc = prep(k, max_blocks))) return rc;
      LaunchInfo* li = (LaunchInfo*)args;
      li->threads = THREADS;
      li->max_blocks = max_blocks;
      li->epl = EPL;
      li->lpc = LPC;
      return 0;
    }
    default:
      set_error("internal: unknown op %d", op);
      return HMCB_E_ARG;
  }
  HMCB_CUDA(cudaPeekAtLastError());
  return 0;
}

// One line per compiled shape:  HMCB_SHAPE(ModelTmpl, KIND, T, DTYPE, UNITM, EPL, LPC)
#define HMCB_SHAPE(ModelTmpl, KIND, T, DTYPE, UNITM, EPL, LPC)                                              \
  static hmcb::Registrar reg_##KIND##_##DTYPE##_##UNITM##_##EPL##_##LPC(hmcb::Entry{                        \
      KIND, DTYPE, UNITM, EPL, LPC,                                                                         \
      &hmcb::launch_any<hmcb::ModelTmpl<T>, T, EPL, LPC, (UNITM != 0), (KIND == HMCB_MODEL_GAUSSIAN)>});

}  // namespace hmcb
