// api.cu -- C ABI entry points of libhmcb.so (see include/hmcb.h).
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>

#include "launch.cuh"

namespace hmcb {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
  set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
  return HMCB_E_CUDA;
}

static int query_device_attr(cudaDeviceAttr attr) {
  int dev = 0, v = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  if (cudaDeviceGetAttribute(&v, attr, dev) != cudaSuccess) return 0;
  return v;
}

int sm_count() {
  static thread_local int cached_dev = -1, cached = 0;
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev != cached_dev) {
    cached = query_device_attr(cudaDevAttrMultiProcessorCount);
    cached_dev = dev;
  }
  return cached > 0 ? cached : 1;
}

int max_optin_smem() {
  static thread_local int cached_dev = -1, cached = 0;
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev != cached_dev) {
    cached = query_device_attr(cudaDevAttrMaxSharedMemoryPerBlockOptin);
    cached_dev = dev;
  }
  return cached;
}

std::vector<Entry>& registry() {
  static std::vector<Entry> r;
  return r;
}

static const Entry* choose_entry_lanes(const hmcb_model& hm, int max_epl, int min_lpc) {
  const int unitm = hm.invmetric_diag ? 0 : 1;
  int force_lpc = 0;
  if (const char* e = getenv("HMCB_LPC")) force_lpc = atoi(e);
  const Entry* best = nullptr;
  long best_cost = 0;
  // pass 0: the caller's wishes (forced or minimum lanes per chain); pass 1: any shape that holds d
  for (int pass = 0; pass < 2 && !best; ++pass) {
    for (const Entry& e : registry()) {
      if (e.kind != hm.kind || e.dtype != hm.dtype || e.unitm != unitm) continue;
      if (pass == 0 && force_lpc && e.lpc != force_lpc) continue;
      if (pass == 0 && !force_lpc && e.lpc < min_lpc) continue;
      if (slots_needed(hm.d, e.lpc) > e.epl || e.epl > max_epl) continue;
      if (hm.kind == HMCB_MODEL_ROSENBROCK && (e.epl & 1)) continue;
      // least padded slots first, then fewest lanes per chain
      long cost = (long)e.epl * e.lpc * 64 + e.lpc;
      // The pairwise Rosenbrock gradient needs no sum across the lanes of a chain inside the trajectory, so more lanes cost
      // nothing there, and past 50 slots per lane q and p no longer fit the register file beside the gradient's temporaries:
      // d=100, 131,072 chains, HMC L=20: 2 lanes x 52 slots 0.273 ms per launch / 0.192 at 16 transitions per launch,
      // 4 lanes x 28 slots 0.226 / 0.172 (tools/rosen_lanes_probe.py)
      if (hm.kind == HMCB_MODEL_ROSENBROCK && e.epl > 50) cost += cost / 4;
      if (!best || cost < best_cost) {
        best = &e;
        best_cost = cost;
      }
    }
    if (!force_lpc && min_lpc <= 1) break;
  }
  if (!best)
    set_error("no compiled kernel shape for model kind %d, dtype %d, d=%d, %s metric", hm.kind, hm.dtype, hm.d,
              unitm ? "unit" : "diagonal");
  return best;
}

const Entry* choose_entry(const hmcb_model& hm, int max_epl) { return choose_entry_lanes(hm, max_epl, 1); }
static const Entry* choose_entry_quiet(const hmcb_model& hm, int max_epl) {
  const int unitm = hm.invmetric_diag ? 0 : 1;
  for (const Entry& e : registry())
    if (e.kind == hm.kind && e.dtype == hm.dtype && e.unitm == unitm && slots_needed(hm.d, e.lpc) <= e.epl && e.epl <= max_epl &&
        !(hm.kind == HMCB_MODEL_ROSENBROCK && (e.epl & 1)))
      return choose_entry_lanes(hm, max_epl, 1);
  return nullptr;
}

static int check_model(const hmcb_model* m) {
  HMCB_CHECK_ARG(m != nullptr, "model is NULL");
  HMCB_CHECK_ARG(m->d >= 1, "model dimension d=%d must be >= 1", m->d);
  HMCB_CHECK_ARG(m->dtype == HMCB_F64 || m->dtype == HMCB_F32, "unknown dtype %d", m->dtype);
  switch (m->kind) {
    case HMCB_MODEL_FUNNEL:
      HMCB_CHECK_ARG(m->d >= 2, "funnel needs d >= 2");
      HMCB_CHECK_ARG(m->par[0] > 0, "funnel sigma must be > 0");
      break;
    case HMCB_MODEL_GAUSSIAN:
      HMCB_CHECK_ARG(m->dev0 != nullptr, "gaussian needs the precision matrix in dev0");
      break;
    case HMCB_MODEL_ROSENBROCK:
      HMCB_CHECK_ARG(m->d % 2 == 0, "pairwise Rosenbrock needs an even d");
      HMCB_CHECK_ARG(m->par[2] != 0, "rosenbrock temperature T must be non-zero");
      break;
    default:
      set_error("model kind %d has no transition kernel in this build", m->kind);
      return HMCB_E_UNSUPPORTED;
  }
  return 0;
}

static int check_chains(const hmcb_chains* c) {
  HMCB_CHECK_ARG(c != nullptr, "chains is NULL");
  HMCB_CHECK_ARG(c->n_chains >= 0 && c->n_chains < (int64_t)1 << 31, "n_chains=%lld out of range", (long long)c->n_chains);
  if (c->n_chains == 0) return 0;
  HMCB_CHECK_ARG(c->q_in && c->q_out, "q_in/q_out must be device pointers");
  HMCB_CHECK_ARG(c->stride_chain > 0 && c->stride_dim > 0, "strides must be positive");
  return 0;
}

template <typename T>
static void fill_args(TransArgs<T>& a, const hmcb_model* m, const hmcb_chains* c, const hmcb_rng* r,
                      const hmcb_step_out* o) {
  memset(&a, 0, sizeof(a));
  a.n_chains = c->n_chains;
  a.chain_offset = c->chain_offset;
  a.d = m->d;
  a.q_in = (const T*)c->q_in;
  a.q_out = (T*)c->q_out;
  a.sc = c->stride_chain;
  a.sd = c->stride_dim;
  a.seed = r->seed;
  a.iteration = (uint32_t)r->iteration;
  a.rng_flags = r->flags;
  a.p_inject = (const T*)r->p_inject;
  a.u_inject = r->u_inject;
  a.u_stride = r->u_stride;
  a.p_prev = (const T*)r->p_prev;
  a.alpha = (T)(r->p_prev ? r->alpha : 0.0);
  a.beta = (T)(r->p_prev ? sqrt(1.0 - r->alpha * r->alpha) : 1.0);
  if (o) {
    a.p_out = (T*)o->p_out;
    a.code = o->code;
    a.energies = (T*)o->energies;
    a.avec_out = (T*)o->avec;
    a.counts = o->counts;
    a.accept_prob = (T*)o->accept_prob;
    a.ptries_out = (T*)o->ptries;
    a.p0_out = (T*)o->p0_out;
  }
  a.minv = (const T*)m->invmetric_diag;
}

template <typename T>
static int hmc_step_t(const Entry* e, const hmcb_model* m, const hmcb_chains* c, const hmcb_rng* r, int n_leap,
                      double eps, int rule, const hmcb_step_out* o, cudaStream_t st) {
  TransArgs<T> a;
  fill_args(a, m, c, r, o);
  a.rule = rule;
  a.nprops = 1;
  a.L[0] = n_leap;
  a.eps[0] = (T)eps;
  if constexpr (sizeof(T) == 8) {
    // dense Gaussian, f64, unit metric, d <= 104: gradient on the FP64 tensor path
    // (not with a partial momentum refresh: the tensor-path kernel draws a full refresh only)
    if (m->kind == HMCB_MODEL_GAUSSIAN && !r->p_prev && !a.p0_out && !getenv("HMCB_GAUSS_GENERIC")) {
      const int rc = gauss_dmma_hmc(*m, a, st);
      if (rc != 1) return rc;
    }
  }
  return e->fn(*m, OP_HMC, &a, st);
}

// ------------------------------------------------------------------ DR workspace
struct DrLayout {
  size_t off_count, off_list, off_H0, off_avec, off_counts, off_ptries, off_frames, total;
  int64_t frame_threads;
  // continuation buffers of the split stages (ccap == 0: stages run unsplit)
  size_t off_ccount, off_clist, off_csti, off_cstf, off_cframes;
  int64_t ccap;
};

// Stages >= 2 in two kernels (transition.cuh, dr_phase): the second one takes the chains whose ghost recursion goes on
// after its first trajectory.  Its list holds a quarter of the batch (all of a small one); chains beyond that finish
// in the first kernel as before.
static bool dr_split_stages() {
  const char* e = getenv("HMCB_DR_STAGE_SPLIT");
  return !(e && atoi(e) == 0);
}

static size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

static int dr_layout(const Entry* e, const hmcb_model* m, int64_t n_chains, int nprops, DrLayout& L, bool with_cont = true) {
  LaunchInfo li;
  int rc = e->fn(*m, OP_QUERY, &li, nullptr);
  if (rc) return rc;
  const size_t ts = m->dtype == HMCB_F64 ? 8 : 4;
  const int64_t groups_per_block = li.threads / li.lpc;
  int64_t blocks = (n_chains + groups_per_block - 1) / groups_per_block;
  if (blocks < 1) blocks = 1;
  if (blocks > li.max_blocks) blocks = li.max_blocks;
  L.frame_threads = blocks * li.threads;
  size_t off = 0;
  L.off_count = off; off = align_up(off + (kMaxProps + 2) * sizeof(int32_t), 256);
  L.off_list = off;  off = align_up(off + (size_t)kMaxProps * n_chains * sizeof(int32_t), 256);
  L.off_H0 = off;    off = align_up(off + (size_t)n_chains * ts, 256);
  L.off_avec = off;  off = align_up(off + (size_t)n_chains * kMaxProps * ts, 256);
  L.off_counts = off; off = align_up(off + (size_t)n_chains * 3 * sizeof(int32_t), 256);
  L.off_ptries = off; off = align_up(off + (size_t)n_chains * kMaxProps * ts, 256);
  L.off_frames = off;
  const int nframes = nprops > 1 ? nprops - 1 : 0;
  off = align_up(off + (size_t)nframes * 2 * li.epl * L.frame_threads * ts, 256);
  L.ccap = 0;
  if (with_cont && nprops >= 3 && dr_split_stages()) {
    L.ccap = n_chains <= 65536 ? n_chains : n_chains / 4;
    L.off_ccount = off;  off = align_up(off + 64, 256);
    L.off_clist = off;   off = align_up(off + (size_t)L.ccap * sizeof(int32_t), 256);
    L.off_csti = off;    off = align_up(off + (size_t)3 * L.ccap * sizeof(int32_t), 256);
    L.off_cstf = off;    off = align_up(off + (size_t)5 * L.ccap * ts, 256);
    L.off_cframes = off; off = align_up(off + (size_t)2 * li.epl * li.lpc * L.ccap * ts, 256);
  }
  L.total = off;
  return 0;
}

// uniforms one chain can consume in a DR transition: one per stage, plus one "try" draw before each later
// stage under the probabilistic rules (consumption order u_acc0, u_try1, u_acc1, u_try2, ...)
static int dr_uniforms(int nprops, int conv) { return conv >= HMCB_DR_PROB_METROP ? 2 * nprops - 1 : nprops; }

template <typename T>
static int drhmc_step_t(const Entry* e, const hmcb_model* m, const hmcb_chains* c, const hmcb_rng* r, int n_leap,
                        double eps, int nprops, double stepfactor, int conv, const hmcb_step_out* o, void* ws,
                        const DrLayout& L, const hmcb_profile* prof, cudaStream_t st, bool prefer_plain = false) {
  TransArgs<T> a;
  fill_args(a, m, c, r, o);
  a.prefer_plain = prefer_plain ? 1 : 0;
  a.nprops = nprops;
  a.dr_conv = conv != HMCB_DR_ALGORITHMS;   // the probabilistic classes keep PyHMC_multistep's conventions
  a.dr_prob = conv == HMCB_DR_PROB_METROP ? 1 : (conv == HMCB_DR_PROB_HASTINGS ? 2 : 0);
  if (!a.dr_prob) a.ptries_out = nullptr;
  for (int j = 0; j < nprops; ++j) {
    // fac = stepfactor**j; nleapf = int(nleap*fac); step_sizef = step_size/fac (src/algorithms.py:510-511)
    const double fac = pow(stepfactor, (double)j);
    a.L[j] = (int32_t)(n_leap * fac);
    a.eps[j] = (T)(eps / fac);
  }
  char* w = (char*)ws;
  a.ws.count = (int32_t*)(w + L.off_count);
  a.ws.list = (int32_t*)(w + L.off_list);
  a.ws.H0 = (T*)(w + L.off_H0);
  a.ws.avec = (T*)(w + L.off_avec);
  a.ws.counts = (int32_t*)(w + L.off_counts);
  a.ws.ptries = (T*)(w + L.off_ptries);
  a.ws.frames = (T*)(w + L.off_frames);
  a.ws.frame_threads = L.frame_threads;
  if (nprops > 1) HMCB_CUDA(cudaMemsetAsync(a.ws.count, 0, (kMaxProps + 2) * sizeof(int32_t), st));
  a.stage = 0;
  if (prof && prof->events[0]) HMCB_CUDA(cudaEventRecord((cudaEvent_t)prof->events[0], st));
  int rc = e->fn(*m, OP_DR0, &a, st);
  if (prof && prof->events[1] && rc == 0) HMCB_CUDA(cudaEventRecord((cudaEvent_t)prof->events[1], st));
  if (L.ccap > 0) {
    a.ws.ccount = (int32_t*)(w + L.off_ccount);
    a.ws.clist = (int32_t*)(w + L.off_clist);
    a.ws.cst_i = (int32_t*)(w + L.off_csti);
    a.ws.cst_f = (T*)(w + L.off_cstf);
    a.ws.cframes = (T*)(w + L.off_cframes);
    a.ws.ccap = (int32_t)L.ccap;
  }
  for (int j = 1; j < nprops && rc == 0; ++j) {
    a.stage = j;
    // (at short trajectories the second launch and the parking cost more than the idle lanes: DRHMC nleap=20 0.79 -> 0.82 ms)
    if (j >= 2 && L.ccap > 0 && n_leap >= 48) {
      // two kernels: up to the point where the chains of a warp part ways, then those whose ghost recursion goes on
      HMCB_CUDA(cudaMemsetAsync(a.ws.ccount, 0, sizeof(int32_t), st));
      a.dr_phase = 1;
      rc = e->fn(*m, OP_DRN, &a, st);
      a.dr_phase = 2;
      if (rc == 0) rc = e->fn(*m, OP_DRN, &a, st);
      a.dr_phase = 0;
    } else {
      rc = e->fn(*m, OP_DRN, &a, st);
    }
    if (prof && prof->events[j + 1] && rc == 0) HMCB_CUDA(cudaEventRecord((cudaEvent_t)prof->events[j + 1], st));
  }
  return rc;
}

template <typename T>
static int mala_step_t(const Entry* e, const hmcb_model* m, const hmcb_chains* c, const hmcb_rng* r, double eps,
                       int n_steps, const hmcb_mala_io* io, cudaStream_t st) {
  MalaArgs<T> a;
  memset(&a, 0, sizeof(a));
  a.n_chains = c->n_chains; a.chain_offset = c->chain_offset; a.d = m->d;
  a.q_in = (const T*)c->q_in; a.q_out = (T*)c->q_out; a.sc = c->stride_chain; a.sd = c->stride_dim;
  a.seed = r->seed; a.iteration = (uint32_t)r->iteration; a.rng_flags = r->flags;
  a.z_inject = (const T*)r->p_inject; a.u_inject = r->u_inject; a.u_stride = r->u_stride;
  a.eps = (T)eps; a.n_steps = n_steps;
  if (io) {
    a.V_in = (const T*)io->V_in; a.Vg_in = (const T*)io->Vg_in; a.V_out = (T*)io->V_out; a.Vg_out = (T*)io->Vg_out;
    a.code = io->code; a.n_accept = io->n_accept; a.counts = io->counts;
  }
  return e->fn(*m, OP_MALA, &a, st);
}

template <typename T>
static int ghmc_step_t(const Entry* e, const hmcb_model* m, const hmcb_chains* c, const hmcb_rng* r, int n_leap,
                       double eps, const hmcb_ghmc_io* io, cudaStream_t st, int K = 1, int accumulate = 0,
                       int32_t* n_accept = nullptr, int* applied = nullptr) {
  GhmcArgs<T> g;
  memset(&g, 0, sizeof(g));
  fill_args(g.t, m, c, r, nullptr);
  g.t.nprops = 1;
  g.t.L[0] = n_leap;
  g.t.eps[0] = (T)eps;
  g.t.p_out = (T*)io->p_out; g.t.code = io->code; g.t.energies = (T*)io->energies; g.t.counts = io->counts;
  g.V_in = (const T*)io->V_in; g.Vg_in = (const T*)io->Vg_in; g.V_out = (T*)io->V_out; g.Vg_out = (T*)io->Vg_out;
  g.pu_in = io->u_in; g.pu_out = io->u_out; g.delta = io->delta;
  g.K = K; g.accumulate = accumulate; g.n_accept = n_accept;
  if (!applied) return e->fn(*m, OP_GHMC, &g, st);
  const int rc = e->fn(*m, OP_GHMC_MULTI, &g, st);        // the whole loop in one launch, where the shape has the kernel
  *applied = g.applied;
  return rc;
}

static int ad_frame_threads(const Entry* e, const hmcb_model* m, int64_t n_chains, int64_t& ft, int& epl,
                            int query_op = OP_QUERY_AD) {
  LaunchInfo li;
  int rc = e->fn(*m, query_op, &li, nullptr);
  if (rc) return rc;
  const int64_t gpb = li.threads / li.lpc;
  int64_t blocks = (n_chains + gpb - 1) / gpb;
  if (blocks < 1) blocks = 1;
  if (blocks > li.max_blocks) blocks = li.max_blocks;
  ft = blocks * li.threads;
  epl = li.epl;
  return 0;
}

template <typename T>
static int adhmc_step_t(const Entry* e, const hmcb_model* m, const hmcb_chains* c, const hmcb_rng* r, int n_leap,
                        double tint, const double* steps, int ntry, int nsteps, const hmcb_adhmc_out* o, void* ws, int64_t ft,
                        cudaStream_t st) {
  AdArgs<T> a;
  memset(&a, 0, sizeof(a));
  a.n_chains = c->n_chains; a.chain_offset = c->chain_offset; a.d = m->d;
  a.q_in = (const T*)c->q_in; a.q_out = (T*)c->q_out; a.sc = c->stride_chain; a.sd = c->stride_dim;
  a.seed = r->seed; a.iteration = (uint32_t)r->iteration; a.rng_flags = r->flags;
  a.p_inject = (const T*)r->p_inject; a.u_inject = r->u_inject; a.u_stride = r->u_stride;
  a.minv = (const T*)m->invmetric_diag;
  a.n_leap = n_leap; a.tint = tint; a.ntry = ntry; a.nsteps = nsteps;
  for (int i = 0; i < ntry; ++i) { a.steps[i] = steps[i]; a.pwts[i] = pow(steps[i], 0.5); }
  if (o) { a.p_out = (T*)o->p_out; a.code = o->code; a.probs = (T*)o->probs; a.aux = (T*)o->aux; a.counts = o->counts; }
  a.frames = (T*)ws;
  a.frame_threads = ft;
  return e->fn(*m, OP_ADHMC, &a, st);
}

template <typename T>
static int adwalk_step_t(const Entry* e, const hmcb_model* m, const hmcb_chains* c, const hmcb_rng* r,
                         const hmcb_adhmc_walk_par* par, const hmcb_adhmc_walk_out* o, void* ws, int64_t ft,
                         cudaStream_t st) {
  WalkArgs<T> a;
  memset(&a, 0, sizeof(a));
  a.n_chains = c->n_chains; a.chain_offset = c->chain_offset; a.d = m->d;
  a.q_in = (const T*)c->q_in; a.q_out = (T*)c->q_out; a.sc = c->stride_chain; a.sd = c->stride_dim;
  a.seed = r->seed; a.iteration = (uint32_t)r->iteration; a.rng_flags = r->flags;
  a.p_inject = (const T*)r->p_inject; a.u_inject = r->u_inject; a.u_stride = r->u_stride;
  a.minv = (const T*)m->invmetric_diag;
  a.nseg = par->n_segments; a.nleap = par->n_leap_segment;
  a.tint_mode = par->tint > 0 ? 1 : 0; a.tint = par->tint;
  a.nr = par->n_ratios;
  for (int j = 0; j < par->n_ratios; ++j) { a.ratios[j] = par->ratios[j]; a.pwts0[j] = par->pwts0[j]; }
  a.smin = par->smin; a.smax = par->smax;
  a.ntry = par->ntry; a.nsteps = par->nsteps;
  for (int i = 0; i < par->ntry; ++i) { a.steps[i] = par->steps[i]; a.pwts[i] = pow(par->steps[i], 0.5); }
  if (o) { a.p_out = (T*)o->p_out; a.code = o->code; a.probs = (T*)o->probs; a.checks = (T*)o->checks; a.counts = o->counts; }
  a.frames = (T*)ws;
  a.frame_threads = ft;
  return e->fn(*m, OP_ADWALK, &a, st);
}

// ------------------------------------------------------------------ DR step over several streams
// The stage kernels are persistent grids over lists whose lengths are not multiples of the
// resident thread count, and the later stages are short lists of long items: every stage ends
// with SMs running dry (measured at 2^20 chains of the d=50 funnel: 12 % and 16 % of stages 1
// and 2).  Chains are independent, so a large batch is cut into contiguous ranges whose stage
// sequences run on different streams: the tail of one range's stage overlaps the next stage of
// another.  Results are bit-identical (every output is indexed by chain, draws are keyed by
// global chain id).  The helper streams belong to the library (created once per device, outside
// any stream capture); they fork from and join back into the caller's stream through events, so
// from the caller's side all work is still ordered on the stream it passed.
constexpr int kMaxSplit = 16;

static int dr_split_for(int64_t n_chains, const hmcb_profile* prof) {
  if (prof) return 1;                       // per-stage events want the stages one after another
  int s = n_chains >= (1 << 15) ? 2 : 1;    // measured: 2 ranges -20 % at 131k chains, -2.5 % at 2^20; 3 and 4 no better
  if (const char* e = getenv("HMCB_DR_SPLIT")) {
    const int v = atoi(e);
    if (v >= 1 && v <= kMaxSplit) s = v;
  }
  if ((int64_t)s > n_chains) s = 1;
  return s;
}

struct HelperStreams {
  bool ready = false;
  cudaStream_t s[kMaxSplit - 1];
  cudaEvent_t fork, join[kMaxSplit - 1];
};
static std::mutex g_helper_mu;
static std::mutex g_seq_mu;   // serialises the enqueueing of calls that use the per-device fork/join events
static HelperStreams g_helpers[64];

static int helper_streams(HelperStreams** out) {
  int dev = 0;
  HMCB_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) { set_error("device ordinal %d out of range", dev); return HMCB_E_ARG; }
  std::lock_guard<std::mutex> lock(g_helper_mu);
  HelperStreams& h = g_helpers[dev];
  if (!h.ready) {
    for (int i = 0; i < kMaxSplit - 1; ++i) {
      HMCB_CUDA(cudaStreamCreateWithFlags(&h.s[i], cudaStreamNonBlocking));
      HMCB_CUDA(cudaEventCreateWithFlags(&h.join[i], cudaEventDisableTiming));
    }
    HMCB_CUDA(cudaEventCreateWithFlags(&h.fork, cudaEventDisableTiming));
    h.ready = true;
  }
  *out = &h;
  return 0;
}

// chains [lo, hi) of a batch as a batch of its own
static void sub_range(const hmcb_model* m, const hmcb_chains* c, const hmcb_rng* r, const hmcb_step_out* o, int nprops,
                      int64_t lo, int64_t hi, hmcb_chains& cs, hmcb_rng& rs, hmcb_step_out& os) {
  const size_t ts = m->dtype == HMCB_F64 ? 8 : 4;
  auto at = [&](const void* base, int64_t elems) -> void* {
    return base ? (void*)((char*)base + (size_t)elems * ts) : nullptr;
  };
  cs = *c;
  cs.n_chains = hi - lo;
  cs.chain_offset = c->chain_offset + lo;
  cs.q_in = at(c->q_in, lo * c->stride_chain);
  cs.q_out = at(c->q_out, lo * c->stride_chain);
  rs = *r;
  rs.p_inject = at(r->p_inject, lo * c->stride_chain);
  rs.p_prev = at(r->p_prev, lo * c->stride_chain);
  rs.u_inject = r->u_inject ? r->u_inject + lo * r->u_stride : nullptr;
  memset(&os, 0, sizeof(os));
  if (o) {
    os.p_out = at(o->p_out, lo * c->stride_chain);
    os.code = o->code ? o->code + lo : nullptr;
    os.energies = at(o->energies, 2 * lo);
    os.avec = at(o->avec, lo * nprops);
    os.counts = o->counts ? o->counts + 3 * lo : nullptr;
    os.accept_prob = at(o->accept_prob, lo);
    os.ptries = at(o->ptries, lo * nprops);
    os.p0_out = at(o->p0_out, lo * c->stride_chain);
  }
}

static int64_t split_bound(int64_t n, int parts, int i) { return (n * i) / parts; }

// ------------------------------------------------------------------ built-in scratch (hmcb_drhmc_step_static)
// Scratch for callers that cannot supply a workspace.  The name below is only a handle in
// host code: the device address has to be asked for (once per device, with that device
// current) and it is that address which goes into the kernel arguments.
static __device__ __align__(256) unsigned char g_static_scratch[HMCB_STATIC_SCRATCH_BYTES];

// Device code of this file may name the array directly; this clears the stage counters,
// the only part of the scratch that must start from zero.
__global__ void k_static_scratch_clear_counters() {
  if (threadIdx.x < 64) reinterpret_cast<int32_t*>(g_static_scratch)[threadIdx.x] = 0;
}

static std::mutex g_static_mu;
static void* g_static_ptr[128];
static bool g_static_ready[128];

// First use on the current device: resolve and cache the array's device address.
static int static_scratch_ptr(void** out) {
  int dev = 0;
  HMCB_CUDA(cudaGetDevice(&dev));
  HMCB_CHECK_ARG(dev >= 0 && dev < 128, "device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (!g_static_ready[dev]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(g_static_scratch)");
    if (p == nullptr) {
      set_error("built-in scratch has no device address on device %d", dev);
      return HMCB_E_CUDA;
    }
    g_static_ptr[dev] = p;
    g_static_ready[dev] = true;
  }
  *out = g_static_ptr[dev];
  return 0;
}

}  // namespace hmcb

using namespace hmcb;

extern "C" {

int hmcb_version(void) { return HMCB_VERSION; }

const char* hmcb_last_error(void) { return g_err; }

void hmcb_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
  uint32_t c[4] = {ctr[0], ctr[1], ctr[2], ctr[3]};
  uint32_t k0 = key[0], k1 = key[1];
  for (int r = 0; r < 10; ++r) {
    Philox::round(c, k0, k1);
    k0 += Philox::W0; k1 += Philox::W1;
  }
  for (int i = 0; i < 4; ++i) out[i] = c[i];
}

void hmcb_philox_block(uint64_t seed, uint64_t chain, uint32_t iteration, uint32_t stream, uint32_t block,
                       uint32_t out[4]) {
  uint32_t w[4];
  Philox::gen(seed, chain, iteration, stream, block, w);
  for (int i = 0; i < 4; ++i) out[i] = w[i];
}

double hmcb_u53(uint32_t a, uint32_t b) { return u53(a, b); }

int hmcb_hmc_step(const hmcb_model* model, const hmcb_chains* chains, const hmcb_rng* rng, int32_t n_leap,
                  double step_size, int32_t rule, const hmcb_step_out* out, void* stream) {
  int rc;
  if ((rc = check_model(model))) return rc;
  if ((rc = check_chains(chains))) return rc;
  HMCB_CHECK_ARG(rng != nullptr, "rng is NULL");
  HMCB_CHECK_ARG(rule == HMCB_RULE_PYHMC || rule == HMCB_RULE_ALG, "unknown accept rule %d", rule);
  HMCB_CHECK_ARG(n_leap >= 0, "n_leap must be >= 0");
  if (chains->n_chains == 0) return 0;
  const Entry* e = choose_entry(*model);
  if (!e) return HMCB_E_UNSUPPORTED;
  cudaStream_t st = (cudaStream_t)stream;
  if (model->dtype == HMCB_F64) return hmc_step_t<double>(e, model, chains, rng, n_leap, step_size, rule, out, st);
  return hmc_step_t<float>(e, model, chains, rng, n_leap, step_size, rule, out, st);
}

int hmcb_hmc_sample(const hmcb_model* model, const hmcb_chains* chains, const hmcb_rng* rng, int32_t n_leap,
                    double step_size, int32_t rule, const hmcb_sample_io* io, const hmcb_step_out* out, void* stream) {
  int rc;
  if ((rc = check_model(model))) return rc;
  if ((rc = check_chains(chains))) return rc;
  HMCB_CHECK_ARG(rng != nullptr && io != nullptr, "rng/io is NULL");
  HMCB_CHECK_ARG(rule == HMCB_RULE_PYHMC || rule == HMCB_RULE_ALG, "unknown accept rule %d", rule);
  HMCB_CHECK_ARG(n_leap >= 0, "n_leap must be >= 0");
  const int K = io->n_transitions;
  HMCB_CHECK_ARG(K >= 1, "n_transitions must be >= 1");
  HMCB_CHECK_ARG(K == 1 || (!rng->p_inject && !rng->u_inject && !rng->p_prev),
                 "injected draws and a partial momentum refresh are defined for n_transitions == 1 only");
  HMCB_CHECK_ARG((io->moment_sum == nullptr) == (io->moment_sumsq == nullptr), "give both moment_sum and moment_sumsq or neither");
  HMCB_CHECK_ARG(!io->n_accept || (out && out->code), "n_accept needs out->code");
  if (chains->n_chains == 0) return 0;
  const Entry* e = choose_entry(*model);
  if (!e) return HMCB_E_UNSUPPORTED;
  cudaStream_t st = (cudaStream_t)stream;
  // the whole loop in one launch where the shape has the kernel for it and nothing needs the states in between
  if (K > 1 && !io->moment_sum && model->kind != HMCB_MODEL_GAUSSIAN) {
    int applied = 0;
    if (model->dtype == HMCB_F64) {
      MultiHmcArgs<double> m;
      fill_args(m.a, model, chains, rng, out);
      m.a.rule = rule; m.a.nprops = 1; m.a.L[0] = n_leap; m.a.eps[0] = step_size;
      m.K = K; m.n_accept = io->n_accept;
      if ((rc = e->fn(*model, OP_HMC_MULTI, &m, st))) return rc;
      applied = m.applied;
    } else {
      MultiHmcArgs<float> m;
      fill_args(m.a, model, chains, rng, out);
      m.a.rule = rule; m.a.nprops = 1; m.a.L[0] = n_leap; m.a.eps[0] = (float)step_size;
      m.K = K; m.n_accept = io->n_accept;
      if ((rc = e->fn(*model, OP_HMC_MULTI, &m, st))) return rc;
      applied = m.applied;
    }
    if (applied) return 0;
  }
  // otherwise: K launches, later ones in place on the new state; counters are the same for every transition
  for (int k = 0; k < K; ++k) {
    hmcb_chains cs = *chains;
    hmcb_rng rs = *rng;
    if (k > 0) cs.q_in = cs.q_out;
    rs.iteration = rng->iteration + (uint64_t)k;
    if (model->dtype == HMCB_F64) rc = hmc_step_t<double>(e, model, &cs, &rs, n_leap, step_size, rule, out, st);
    else rc = hmc_step_t<float>(e, model, &cs, &rs, n_leap, step_size, rule, out, st);
    if (rc) return rc;
    rc = launch_post_transition(model->dtype, chains->n_chains, model->d, cs.q_out, cs.stride_chain, cs.stride_dim, io->moment_sum,
                                io->moment_sumsq, nullptr, nullptr, 0, out ? out->code : nullptr, 1, 1, io->n_accept, st);
    if (rc) return rc;
  }
  if (K > 1 && out && out->counts) {
    // Hcount, Vgcount, leapcount of one HMC transition are (2, n_leap + 1, 1) whatever happens (src/pyhmc.py:69)
    rc = launch_scale_counts(out->counts, chains->n_chains, K, st);
  }
  return rc;
}

int hmcb_leapfrog(const hmcb_model* model, const hmcb_chains* chains, const void* p_in, void* p_out, int32_t n_leap,
                  double step_size, void* stream) {
  int rc;
  if ((rc = check_model(model))) return rc;
  if ((rc = check_chains(chains))) return rc;
  HMCB_CHECK_ARG(p_in && p_out, "p_in/p_out must be device pointers");
  if (chains->n_chains == 0) return 0;
  const Entry* e = choose_entry(*model);
  if (!e) return HMCB_E_UNSUPPORTED;
  cudaStream_t st = (cudaStream_t)stream;
  if (model->dtype == HMCB_F64) {
    AuxArgs<double> a{chains->n_chains, model->d, (const double*)chains->q_in, (double*)chains->q_out,
                      chains->stride_chain, chains->stride_dim, (const double*)p_in, (double*)p_out, nullptr, nullptr,
                      (const double*)model->invmetric_diag, n_leap, step_size};
    return e->fn(*model, OP_LEAPFROG, &a, st);
  }
  AuxArgs<float> a{chains->n_chains, model->d, (const float*)chains->q_in, (float*)chains->q_out,
                   chains->stride_chain, chains->stride_dim, (const float*)p_in, (float*)p_out, nullptr, nullptr,
                   (const float*)model->invmetric_diag, n_leap, (float)step_size};
  return e->fn(*model, OP_LEAPFROG, &a, st);
}

int hmcb_logp_grad(const hmcb_model* model, const hmcb_chains* chains, void* logp, void* grad, void* stream) {
  int rc;
  if ((rc = check_model(model))) return rc;
  HMCB_CHECK_ARG(chains != nullptr && chains->q_in != nullptr, "chains/q_in is NULL");
  HMCB_CHECK_ARG(chains->stride_chain > 0 && chains->stride_dim > 0, "strides must be positive");
  if (chains->n_chains == 0) return 0;
  const Entry* e = choose_entry(*model);
  if (!e) return HMCB_E_UNSUPPORTED;
  cudaStream_t st = (cudaStream_t)stream;
  if (model->dtype == HMCB_F64) {
    AuxArgs<double> a{chains->n_chains, model->d, (const double*)chains->q_in, nullptr, chains->stride_chain,
                      chains->stride_dim, nullptr, nullptr, (double*)logp, (double*)grad,
                      (const double*)model->invmetric_diag, 0, 0.0};
    return e->fn(*model, OP_LOGPGRAD, &a, st);
  }
  AuxArgs<float> a{chains->n_chains, model->d, (const float*)chains->q_in, nullptr, chains->stride_chain,
                   chains->stride_dim, nullptr, nullptr, (float*)logp, (float*)grad,
                   (const float*)model->invmetric_diag, 0, 0.0f};
  return e->fn(*model, OP_LOGPGRAD, &a, st);
}

size_t hmcb_drhmc_workspace_bytes(const hmcb_model* model, int64_t n_chains, int32_t nprops) {
  if (check_model(model)) return 0;
  if (n_chains < 1 || nprops < 1 || nprops > kMaxProps) {
    set_error("workspace query: n_chains=%lld nprops=%d out of range (nprops <= %d)", (long long)n_chains, nprops, kMaxProps);
    return 0;
  }
  const Entry* e = choose_entry(*model);
  if (!e) return 0;
  DrLayout L;
  const int S = dr_split_for(n_chains, nullptr);
  size_t total = 0;
  for (int i = 0; i < S; ++i) {
    if (dr_layout(e, model, split_bound(n_chains, S, i + 1) - split_bound(n_chains, S, i), nprops, L)) return 0;
    total += L.total;
  }
  return total;
}

int hmcb_drhmc_step(const hmcb_model* model, const hmcb_chains* chains, const hmcb_rng* rng, int32_t n_leap,
                    double step_size, int32_t nprops, double stepfactor, int32_t convention, const hmcb_step_out* out,
                    void* workspace, size_t workspace_bytes, const hmcb_profile* profile, void* stream) {
  int rc;
  if ((rc = check_model(model))) return rc;
  if ((rc = check_chains(chains))) return rc;
  HMCB_CHECK_ARG(rng != nullptr, "rng is NULL");
  HMCB_CHECK_ARG(nprops >= 1 && nprops <= kMaxProps, "nprops=%d must be in 1..%d", nprops, kMaxProps);
  HMCB_CHECK_ARG(stepfactor > 0, "stepfactor must be > 0");
  HMCB_CHECK_ARG(convention >= HMCB_DR_ALGORITHMS && convention <= HMCB_DR_PROB_HASTINGS, "unknown DR convention %d", convention);
  HMCB_CHECK_ARG(n_leap >= 0, "n_leap must be >= 0");
  HMCB_CHECK_ARG(!rng->u_inject || rng->u_stride >= dr_uniforms(nprops, convention),
                 "u_inject needs u_stride >= %d for nprops=%d under convention %d", dr_uniforms(nprops, convention), nprops, convention);
  if (chains->n_chains == 0) return 0;
  const Entry* e = choose_entry(*model);
  if (!e) return HMCB_E_UNSUPPORTED;
  // the workspace is always sized for the un-profiled call (hmcb_drhmc_workspace_bytes), which is the larger one
  const int S = dr_split_for(chains->n_chains, profile);
  DrLayout L[kMaxSplit];
  size_t need = 0;
  for (int i = 0; i < S; ++i) {
    const int64_t lo = split_bound(chains->n_chains, S, i), hi = split_bound(chains->n_chains, S, i + 1);
    if ((rc = dr_layout(e, model, hi - lo, nprops, L[i]))) return rc;
    need += L[i].total;
  }
  if (!workspace || workspace_bytes < need) {
    set_error("drhmc workspace too small: have %zu bytes, need %zu", workspace ? workspace_bytes : (size_t)0, need);
    return HMCB_E_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  auto run = [&](const hmcb_chains* c, const hmcb_rng* r, const hmcb_step_out* o, void* ws, const DrLayout& lay,
                 cudaStream_t on) -> int {
    if (model->dtype == HMCB_F64)
      return drhmc_step_t<double>(e, model, c, r, n_leap, step_size, nprops, stepfactor, convention, o, ws, lay, profile, on);
    return drhmc_step_t<float>(e, model, c, r, n_leap, step_size, nprops, stepfactor, convention, o, ws, lay, profile, on);
  };
  if (S == 1) return run(chains, rng, out, workspace, L[0], st);
  HelperStreams* h;
  if ((rc = helper_streams(&h))) return rc;
  // the fork/join events are shared by every call on this device: enqueue the whole fork - launch - join sequence
  // of one call before another host thread's (only the enqueueing is serialised; the work itself is asynchronous)
  std::lock_guard<std::mutex> seq_lock(g_seq_mu);
  HMCB_CUDA(cudaEventRecord(h->fork, st));
  char* w = (char*)workspace;
  rc = 0;
  for (int i = 0; i < S && rc == 0; ++i) {
    const int64_t lo = split_bound(chains->n_chains, S, i), hi = split_bound(chains->n_chains, S, i + 1);
    hmcb_chains cs;
    hmcb_rng rs;
    hmcb_step_out os;
    sub_range(model, chains, rng, out, nprops, lo, hi, cs, rs, os);
    cudaStream_t on = i == 0 ? st : h->s[i - 1];
    if (i > 0) {
      cudaError_t ce = cudaStreamWaitEvent(on, h->fork, 0);
      if (ce != cudaSuccess) { rc = cuda_fail(ce, "cudaStreamWaitEvent(helper, fork)"); break; }
    }
    rc = run(&cs, &rs, out ? &os : nullptr, w, L[i], on);
    if (i > 0) {
      // a helper stream that was forked is always joined back, whatever happened on it: the caller's stream
      // stays ordered after the library's own streams (and a stream capture stays joined) on the error path too
      cudaError_t ce = cudaEventRecord(h->join[i - 1], on);
      if (ce == cudaSuccess) ce = cudaStreamWaitEvent(st, h->join[i - 1], 0);
      if (ce != cudaSuccess && rc == 0) rc = cuda_fail(ce, "join of the helper stream");
    }
    w += L[i].total;
  }
  return rc;
}

// ------------------------------------------------------------------ K transitions per call, pipelined over chain ranges
// ranges of a sample call: enough that the stage kernels of different ranges and transitions always have
// something to fill each other's tails with, few enough that a range's first stage still covers the GPU
static int sample_split_for(int64_t n_chains) {
  // d=50 funnel, nleap=150, ms per transition by ranges (tools/dr_sample_bench.py, HMCB_SAMPLE_SPLIT):
  //   2^19 chains: 4: 2.533  6: 2.549  8: 2.574      2^18: 4: 1.400  8: 1.357      2^17: 4: 0.892  6: 0.875  8: 0.864  12: 1.36
  //   2^16: 2: 0.734  4: 0.693  8: 0.685             2^15: 2: 0.638  4: 0.624  8: 0.623             2^14: 2: 0.624  4: 0.607  8: 0.624
  // (below ~2^16 chains the transition is its critical path: one trajectory per stage, 0.6 ms)
  int s = n_chains >= (1 << 19) ? 4 : (n_chains >= (1 << 16) ? 8 : (n_chains >= (1 << 14) ? 4 : 1));
  if (const char* e = getenv("HMCB_SAMPLE_SPLIT")) {
    const int v = atoi(e);
    if (v >= 1 && v <= kMaxSplit) s = v;
  }
  if ((int64_t)s > n_chains) s = 1;
  return s;
}

static int sample_layouts(const Entry* e, const hmcb_model* m, int64_t n_chains, int nprops, int S, DrLayout* L, size_t* off_counts,
                          size_t& total) {
  total = 0;
  for (int i = 0; i < S; ++i) {
    const int64_t n = split_bound(n_chains, S, i + 1) - split_bound(n_chains, S, i);
    int rc = dr_layout(e, m, n, nprops, L[i]);
    if (rc) return rc;
    off_counts[i] = total + L[i].total;                              // this range's per-transition counters
    total = off_counts[i] + align_up((size_t)n * 3 * sizeof(int32_t), 256);
  }
  return 0;
}

// the multi-pass form for short trajectories (transition_ws.cuh, k_dr0_multi_ws): one chain range, the stage kernels'
// scratch, then iter_done[C], iter_of[C], two continue lists [C] and their two counters
struct ShortLayout {
  DrLayout L;
  size_t off_done, off_of, off_cl[2], off_cc, total;
};

static int short_layout(const Entry* e, const hmcb_model* m, int64_t n_chains, int nprops, ShortLayout& S) {
  int rc = dr_layout(e, m, n_chains, nprops, S.L, false);
  if (rc) return rc;
  size_t off = S.L.total;
  const size_t cb = align_up((size_t)n_chains * sizeof(int32_t), 256);
  S.off_done = off; off += cb;
  S.off_of = off;   off += cb;
  S.off_cl[0] = off; off += cb;
  S.off_cl[1] = off; off += cb;
  S.off_cc = off;   off += 256;
  S.total = off;
  return 0;
}

size_t hmcb_drhmc_sample_workspace_bytes(const hmcb_model* model, int64_t n_chains, int32_t nprops) {
  if (check_model(model)) return 0;
  if (n_chains < 1 || nprops < 1 || nprops > kMaxProps) {
    set_error("workspace query: n_chains=%lld nprops=%d out of range (nprops <= %d)", (long long)n_chains, nprops, kMaxProps);
    return 0;
  }
  const Entry* e = choose_entry(*model);
  if (!e) return 0;
  DrLayout L[kMaxSplit];
  size_t offc[kMaxSplit], total = 0;
  if (sample_layouts(e, model, n_chains, nprops, sample_split_for(n_chains), L, offc, total)) return 0;
  ShortLayout SL;
  if (short_layout(e, model, n_chains, nprops, SL)) return 0;
  return total > SL.total ? total : SL.total;
}

// Batches from which the short-trajectory form pays: 2^20 chains of the d=50 funnel at nleap=20, 0.79 -> 0.69 ms per
// transition; at 2^17 chains it loses (0.18 -> 0.23 ms: its passes over short lists are mostly barrier latency).
static int64_t short_min_chains() {
  if (const char* ev = getenv("HMCB_DR_SHORT_MIN")) return atoll(ev);
  return (int64_t)1 << 19;
}

// K transitions at short trajectories: chunks of kShortChunk transitions, each a first pass of k_dr0_multi_ws over all
// chains and then, kShortChunk times, the delayed-rejection stage kernels over the chains that stopped and a pass over
// those that go on.  Returns 1 if this shape / call has no such kernel (the caller then runs the pipelined form).
constexpr int kShortChunk = 5;   // measured 2..8 on the d=50 funnel at nleap=20: 0.69-0.71 ms per transition, flat

static int drhmc_sample_short(const Entry* e, const hmcb_model* m, const hmcb_chains* c, const hmcb_rng* r, int n_leap,
                              double eps, int nprops, double stepfactor, int conv, int K, const hmcb_sample_io* io,
                              const hmcb_step_out* o, void* ws, size_t ws_bytes, cudaStream_t st) {
  MultiDrArgs<double> M;
  memset(&M, 0, sizeof(M));
  TransArgs<double>& a = M.a;
  fill_args(a, m, c, r, o);
  M.query_only = 1;
  int rc = e->fn(*m, OP_DR0_MULTI, &M, st);
  if (rc) return rc;
  if (!M.applied) return 1;
  ShortLayout SL;
  if ((rc = short_layout(e, m, c->n_chains, nprops, SL))) return rc;
  if (!ws || ws_bytes < SL.total) {
    set_error("drhmc sample workspace too small: have %zu bytes, need %zu", ws ? ws_bytes : (size_t)0, SL.total);
    return HMCB_E_WORKSPACE;
  }
  const int64_t C = c->n_chains;
  char* w = (char*)ws;
  a.nprops = nprops;
  a.dr_conv = conv != HMCB_DR_ALGORITHMS;
  a.dr_prob = conv == HMCB_DR_PROB_METROP ? 1 : (conv == HMCB_DR_PROB_HASTINGS ? 2 : 0);
  if (!a.dr_prob) a.ptries_out = nullptr;
  for (int j = 0; j < nprops; ++j) {
    const double fac = pow(stepfactor, (double)j);     // src/algorithms.py:510-511
    a.L[j] = (int32_t)(n_leap * fac);
    a.eps[j] = eps / fac;
  }
  const DrLayout& L = SL.L;
  a.ws.count = (int32_t*)(w + L.off_count);
  a.ws.list = (int32_t*)(w + L.off_list);
  a.ws.H0 = (double*)(w + L.off_H0);
  a.ws.avec = (double*)(w + L.off_avec);
  a.ws.counts = (int32_t*)(w + L.off_counts);
  a.ws.ptries = (double*)(w + L.off_ptries);
  a.ws.frames = (double*)(w + L.off_frames);
  a.ws.frame_threads = L.frame_threads;
  int32_t* iter_done = (int32_t*)(w + SL.off_done);
  int32_t* iter_of = (int32_t*)(w + SL.off_of);
  int32_t* CL[2] = {(int32_t*)(w + SL.off_cl[0]), (int32_t*)(w + SL.off_cl[1])};
  int32_t* CC[2] = {(int32_t*)(w + SL.off_cc), (int32_t*)(w + SL.off_cc) + 32};
  a.iter_of = iter_of;
  a.iter_done = iter_done;
  a.n_accept = io->n_accept;
  // in place on q_out; counts are summed over the call
  if (c->q_out != c->q_in) HMCB_CUDA(cudaMemcpyAsync(c->q_out, c->q_in, (size_t)C * m->d * sizeof(double), cudaMemcpyDeviceToDevice, st));
  a.q_in = a.q_out;
  if (a.counts) HMCB_CUDA(cudaMemsetAsync(a.counts, 0, (size_t)C * 3 * sizeof(int32_t), st));
  M.query_only = 0;
  M.iter_of = iter_of;
  const size_t count_bytes = (kMaxProps + 2) * sizeof(int32_t);
  int chunk = kShortChunk;
  if (const char* ev = getenv("HMCB_DR_SHORT_CHUNK")) {
    const int v = atoi(ev);
    if (v >= 1 && v <= 64) chunk = v;
  }
  for (int base = 0; base < K; base += chunk) {
    const int Kc = K - base < chunk ? K - base : chunk;
    a.iteration = (uint32_t)(r->iteration + (uint64_t)base);
    a.K_total = Kc;
    HMCB_CUDA(cudaMemsetAsync(iter_done, 0, (size_t)C * sizeof(int32_t), st));
    HMCB_CUDA(cudaMemsetAsync(a.ws.count, 0, count_bytes, st));
    HMCB_CUDA(cudaMemsetAsync(CC[0], 0, 256, st));
    M.max_steps = Kc; M.list = nullptr; M.list_count = nullptr;
    if ((rc = e->fn(*m, OP_DR0_MULTI, &M, st))) return rc;
    if (nprops == 1) continue;      // nothing is ever handed on
    int cur = 0;
    for (int p = 1; p <= Kc; ++p) {
      a.cont_list = CL[cur];
      a.cont_count = CC[cur];
      for (int j = 1; j < nprops; ++j) {
        a.stage = j;
        if ((rc = e->fn(*m, OP_DRN, &a, st))) return rc;
      }
      if (p == Kc) break;
      HMCB_CUDA(cudaMemsetAsync(a.ws.count, 0, count_bytes, st));
      M.max_steps = Kc - p; M.list = CL[cur]; M.list_count = CC[cur];
      if ((rc = e->fn(*m, OP_DR0_MULTI, &M, st))) return rc;
      cur ^= 1;
      HMCB_CUDA(cudaMemsetAsync(CC[cur], 0, 128, st));
    }
  }
  return 0;
}

int hmcb_drhmc_sample(const hmcb_model* model, const hmcb_chains* chains, const hmcb_rng* rng, int32_t n_leap,
                      double step_size, int32_t nprops, double stepfactor, int32_t convention, const hmcb_sample_io* io,
                      const hmcb_step_out* out, void* workspace, size_t workspace_bytes, void* stream) {
  int rc;
  if ((rc = check_model(model))) return rc;
  if ((rc = check_chains(chains))) return rc;
  HMCB_CHECK_ARG(rng != nullptr && io != nullptr, "rng/io is NULL");
  HMCB_CHECK_ARG(nprops >= 1 && nprops <= kMaxProps, "nprops=%d must be in 1..%d", nprops, kMaxProps);
  HMCB_CHECK_ARG(stepfactor > 0, "stepfactor must be > 0");
  HMCB_CHECK_ARG(convention >= HMCB_DR_ALGORITHMS && convention <= HMCB_DR_PROB_HASTINGS, "unknown DR convention %d", convention);
  HMCB_CHECK_ARG(n_leap >= 0, "n_leap must be >= 0");
  const int K = io->n_transitions;
  HMCB_CHECK_ARG(K >= 1, "n_transitions must be >= 1");
  HMCB_CHECK_ARG(K == 1 || (!rng->p_inject && !rng->u_inject && !rng->p_prev),
                 "injected draws and a partial momentum refresh are defined for n_transitions == 1 only");
  HMCB_CHECK_ARG(!rng->u_inject || rng->u_stride >= dr_uniforms(nprops, convention),
                 "u_inject needs u_stride >= %d for nprops=%d under convention %d", dr_uniforms(nprops, convention), nprops, convention);
  HMCB_CHECK_ARG((io->moment_sum == nullptr) == (io->moment_sumsq == nullptr), "give both moment_sum and moment_sumsq or neither");
  HMCB_CHECK_ARG(!io->n_accept || (out && out->code), "n_accept needs out->code");
  if (chains->n_chains == 0) return 0;
  const Entry* e = choose_entry(*model);
  if (!e) return HMCB_E_UNSUPPORTED;
  // short trajectories: chains run their transitions in registers for as long as the first stage accepts
  // (k_dr0_multi_ws); at nleap=20 a transition by stage launches spends more on loading and storing q than on integrating
  if (K > 1 && n_leap < 48 && !io->moment_sum && model->dtype == HMCB_F64 && !model->invmetric_diag &&
      chains->stride_dim == 1 && chains->stride_chain == model->d && chains->n_chains >= short_min_chains() && !rng->p_inject &&
      !rng->u_inject && !rng->p_prev && !(out && out->p0_out) && !getenv("HMCB_NO_DR_SHORT")) {
    rc = drhmc_sample_short(e, model, chains, rng, n_leap, step_size, nprops, stepfactor, convention, K, io, out, workspace,
                            workspace_bytes, (cudaStream_t)stream);
    if (rc != 1) return rc;
  }
  const int S = sample_split_for(chains->n_chains);
  DrLayout L[kMaxSplit];
  size_t offc[kMaxSplit], need = 0;
  if ((rc = sample_layouts(e, model, chains->n_chains, nprops, S, L, offc, need))) return rc;
  if (!workspace || workspace_bytes < need) {
    set_error("drhmc sample workspace too small: have %zu bytes, need %zu", workspace ? workspace_bytes : (size_t)0, need);
    return HMCB_E_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  HelperStreams* h = nullptr;
  std::unique_lock<std::mutex> seq_lock(g_seq_mu, std::defer_lock);
  if (S > 1) {
    if ((rc = helper_streams(&h))) return rc;
    seq_lock.lock();   // fork/join events are shared per device: one call's enqueueing at a time
    HMCB_CUDA(cudaEventRecord(h->fork, st));
  }
  const size_t ts = model->dtype == HMCB_F64 ? 8 : 4;
  const int min_accept = convention == HMCB_DR_ALGORITHMS ? 1 : 0;
  // transition k of every range is enqueued before transition k+1 of any: all streams have work from the start
  rc = 0;
  int forked = 1;   // streams that wait on the fork event (the caller's included): exactly these are joined below
  for (int i = 1; i < S && rc == 0; ++i) {
    cudaError_t ce = cudaStreamWaitEvent(h->s[i - 1], h->fork, 0);
    if (ce != cudaSuccess) rc = cuda_fail(ce, "cudaStreamWaitEvent(helper, fork)");
    else forked = i + 1;
  }
  for (int k = 0; k < K && rc == 0; ++k) {
    for (int i = 0; i < S && rc == 0; ++i) {
      const int64_t lo = split_bound(chains->n_chains, S, i), hi = split_bound(chains->n_chains, S, i + 1);
      hmcb_chains cs;
      hmcb_rng rs;
      hmcb_step_out os;
      sub_range(model, chains, rng, out, nprops, lo, hi, cs, rs, os);
      if (k > 0) cs.q_in = cs.q_out;                   // later transitions run in place on the new state
      rs.iteration = rng->iteration + (uint64_t)k;
      int32_t* counts_k = (int32_t*)((char*)workspace + offc[i]);
      int32_t* counts_tot = os.counts;
      if (K > 1 && counts_tot) os.counts = counts_k;
      cudaStream_t on = i == 0 ? st : h->s[i - 1];
      void* ws = (char*)workspace + (offc[i] - L[i].total);
      const bool plain = S >= 3;   // see launch.cuh: with several ranges in flight the plain stage-0 kernel packs better
      if (model->dtype == HMCB_F64)
        rc = drhmc_step_t<double>(e, model, &cs, &rs, n_leap, step_size, nprops, stepfactor, convention, &os, ws, L[i], nullptr, on,
                                  plain);
      else
        rc = drhmc_step_t<float>(e, model, &cs, &rs, n_leap, step_size, nprops, stepfactor, convention, &os, ws, L[i], nullptr, on,
                                 plain);
      if (rc) break;
      void* msum = io->moment_sum ? (char*)io->moment_sum + (size_t)lo * model->d * ts : nullptr;
      void* msq = io->moment_sumsq ? (char*)io->moment_sumsq + (size_t)lo * model->d * ts : nullptr;
      rc = launch_post_transition(model->dtype, hi - lo, model->d, cs.q_out, cs.stride_chain, cs.stride_dim, msum, msq, counts_k,
                                  (K > 1) ? counts_tot : nullptr, k == 0, os.code, min_accept, 1 << 30,
                                  io->n_accept ? io->n_accept + lo : nullptr, on);
    }
  }
  for (int i = 1; i < forked; ++i) {
    // a forked helper stream is always joined back, also after an error
    cudaError_t ce = cudaEventRecord(h->join[i - 1], h->s[i - 1]);
    if (ce == cudaSuccess) ce = cudaStreamWaitEvent(st, h->join[i - 1], 0);
    if (ce != cudaSuccess && rc == 0) rc = cuda_fail(ce, "join of the helper stream");
  }
  return rc;
}

int hmcb_drhmc_step_static(const hmcb_model* model, const hmcb_chains* chains, const hmcb_rng* rng, int32_t n_leap,
                           double step_size, int32_t nprops, double stepfactor, int32_t convention,
                           const hmcb_step_out* out, void* stream) {
  int rc;
  if ((rc = check_model(model))) return rc;
  if ((rc = check_chains(chains))) return rc;
  HMCB_CHECK_ARG(rng != nullptr, "rng is NULL");
  HMCB_CHECK_ARG(nprops >= 1 && nprops <= kMaxProps, "nprops=%d must be in 1..%d", nprops, kMaxProps);
  HMCB_CHECK_ARG(stepfactor > 0, "stepfactor must be > 0");
  HMCB_CHECK_ARG(convention >= HMCB_DR_ALGORITHMS && convention <= HMCB_DR_PROB_HASTINGS, "unknown DR convention %d", convention);
  HMCB_CHECK_ARG(n_leap >= 0, "n_leap must be >= 0");
  HMCB_CHECK_ARG(!rng->u_inject || rng->u_stride >= dr_uniforms(nprops, convention),
                 "u_inject needs u_stride >= %d for nprops=%d under convention %d", dr_uniforms(nprops, convention), nprops, convention);
  if (chains->n_chains == 0) return 0;
  const Entry* e = choose_entry(*model);
  if (!e) return HMCB_E_UNSUPPORTED;
  DrLayout L;
  if ((rc = dr_layout(e, model, chains->n_chains, nprops, L, false))) return rc;   // unsplit stages: the scratch it was sized for
  if (L.total > (size_t)HMCB_STATIC_SCRATCH_BYTES) {
    set_error("built-in scratch too small: n_chains=%lld, d=%d, nprops=%d needs %zu bytes, the library holds %u "
              "(HMCB_STATIC_SCRATCH_BYTES); split the batch or call hmcb_drhmc_step with your own workspace",
              (long long)chains->n_chains, model->d, nprops, L.total, (unsigned)HMCB_STATIC_SCRATCH_BYTES);
    return HMCB_E_WORKSPACE;
  }
  void* scratch = nullptr;
  if ((rc = static_scratch_ptr(&scratch))) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  // the previous call's lists and frames are overwritten as they are produced; only the
  // stage counters are read before being written
  k_static_scratch_clear_counters<<<1, 64, 0, st>>>();
  HMCB_CUDA(cudaPeekAtLastError());
  if (model->dtype == HMCB_F64)
    return drhmc_step_t<double>(e, model, chains, rng, n_leap, step_size, nprops, stepfactor, convention, out, scratch,
                                L, nullptr, st);
  return drhmc_step_t<float>(e, model, chains, rng, n_leap, step_size, nprops, stepfactor, convention, out, scratch, L,
                             nullptr, st);
}

int hmcb_mala_step(const hmcb_model* model, const hmcb_chains* chains, const hmcb_rng* rng, double eps,
                   int32_t n_steps, const hmcb_mala_io* io, void* stream) {
  int rc;
  if ((rc = check_model(model))) return rc;
  if ((rc = check_chains(chains))) return rc;
  HMCB_CHECK_ARG(rng != nullptr, "rng is NULL");
  HMCB_CHECK_ARG(n_steps >= 1, "n_steps must be >= 1");
  HMCB_CHECK_ARG(eps > 0, "eps must be > 0");
  HMCB_CHECK_ARG(!rng->p_inject || n_steps == 1, "injected normals are only defined for n_steps == 1");
  HMCB_CHECK_ARG(!rng->u_inject || rng->u_stride >= n_steps, "u_inject needs u_stride >= n_steps");
  if (chains->n_chains == 0) return 0;
  hmcb_model mm = *model;
  mm.invmetric_diag = nullptr;  // the reference's MALA never uses its metric (src/mala.py:45-60)
  const Entry* e = choose_entry(mm, kMalaMaxEpl);
  if (!e) return HMCB_E_UNSUPPORTED;
  cudaStream_t st = (cudaStream_t)stream;
  if (model->dtype == HMCB_F64) return mala_step_t<double>(e, &mm, chains, rng, eps, n_steps, io, st);
  return mala_step_t<float>(e, &mm, chains, rng, eps, n_steps, io, st);
}

int hmcb_ghmc_step(const hmcb_model* model, const hmcb_chains* chains, const hmcb_rng* rng, int32_t n_leap,
                   double step_size, const hmcb_ghmc_io* io, void* stream) {
  int rc;
  if ((rc = check_model(model))) return rc;
  if ((rc = check_chains(chains))) return rc;
  HMCB_CHECK_ARG(rng != nullptr && io != nullptr, "rng/io is NULL");
  HMCB_CHECK_ARG(n_leap >= 0, "n_leap must be >= 0");
  HMCB_CHECK_ARG(!rng->p_prev || (rng->alpha >= -1.0 && rng->alpha <= 1.0), "alpha must be in [-1, 1]");
  if (chains->n_chains == 0) return 0;
  // GHMC moves six state arrays per transition (q, p and the cached gradient, in and out): it is bound by
  // memory, and lanes of a chain that sit side by side read whole lines.  Measured on the d=50 funnel, 2^20
  // chains, nleap=20: 1.61 ms with 2 lanes per chain, 1.34 ms with 4; nleap=1: 1.01 / 0.77 ms.
  const Entry* e = choose_entry_lanes(*model, kGhmcMaxEpl, model->d >= 32 ? 4 : 1);
  if (!e) return HMCB_E_UNSUPPORTED;
  cudaStream_t st = (cudaStream_t)stream;
  if (model->dtype == HMCB_F64) return ghmc_step_t<double>(e, model, chains, rng, n_leap, step_size, io, st);
  return ghmc_step_t<float>(e, model, chains, rng, n_leap, step_size, io, st);
}

int hmcb_ghmc_sample(const hmcb_model* model, const hmcb_chains* chains, const hmcb_rng* rng, int32_t n_leap,
                     double step_size, const hmcb_ghmc_io* io, int32_t n_transitions, int32_t* n_accept, void* stream) {
  int rc;
  if ((rc = check_model(model))) return rc;
  if ((rc = check_chains(chains))) return rc;
  HMCB_CHECK_ARG(rng != nullptr && io != nullptr, "rng/io is NULL");
  HMCB_CHECK_ARG(n_leap >= 0, "n_leap must be >= 0");
  const int K = n_transitions;
  HMCB_CHECK_ARG(K >= 1, "n_transitions must be >= 1");
  HMCB_CHECK_ARG(rng->p_prev != nullptr, "rng->p_prev (the momentum the first transition refreshes) is NULL");
  HMCB_CHECK_ARG(rng->alpha >= -1.0 && rng->alpha <= 1.0, "alpha must be in [-1, 1]");
  HMCB_CHECK_ARG(K == 1 || (!rng->p_inject && !rng->u_inject), "injected draws are defined for n_transitions == 1 only");
  HMCB_CHECK_ARG(io->p_out && io->V_out && io->Vg_out, "p_out, V_out and Vg_out carry the state between transitions: all required");
  HMCB_CHECK_ARG(!io->u_in || io->u_out, "u_in needs u_out");
  if (chains->n_chains == 0) return 0;
  // One launch per transition is bound by its six state arrays and wants hmcb_ghmc_step's 4 lanes per chain; K transitions
  // in one launch are bound by arithmetic and want the fewest lanes that hold d (d=50 funnel, 2^20 chains, nleap=20, per
  // transition: 1.58 ms single; fused 0.71 ms with 4 lanes, 0.58 ms with 2).  HMCB_LPC pins one shape for both.
  const bool fuse = K > 1 && model->kind != HMCB_MODEL_GAUSSIAN && !getenv("HMCB_NO_MULTI");
  const Entry* e = choose_entry_lanes(*model, kGhmcMaxEpl, (model->d >= 32 && !fuse) ? 4 : 1);
  if (!e) return HMCB_E_UNSUPPORTED;
  cudaStream_t st = (cudaStream_t)stream;
  const bool f64 = model->dtype == HMCB_F64;
  if (fuse) {
    int applied = 0;
    rc = f64 ? ghmc_step_t<double>(e, model, chains, rng, n_leap, step_size, io, st, K, 0, n_accept, &applied)
             : ghmc_step_t<float>(e, model, chains, rng, n_leap, step_size, io, st, K, 0, n_accept, &applied);
    if (rc || applied) return rc;
  }
  // otherwise launch by launch, the later ones in place on the outputs of the one before; counters summed by the kernel
  for (int k = 0; k < K; ++k) {
    hmcb_chains cs = *chains;
    hmcb_rng rs = *rng;
    hmcb_ghmc_io is = *io;
    if (k > 0) {
      cs.q_in = cs.q_out;
      rs.p_prev = io->p_out;
      is.V_in = io->V_out;
      is.Vg_in = io->Vg_out;
      if (io->u_in) is.u_in = io->u_out;
    }
    rs.iteration = rng->iteration + (uint64_t)k;
    rc = f64 ? ghmc_step_t<double>(e, model, &cs, &rs, n_leap, step_size, &is, st, 1, k > 0, n_accept)
             : ghmc_step_t<float>(e, model, &cs, &rs, n_leap, step_size, &is, st, 1, k > 0, n_accept);
    if (rc) return rc;
  }
  return 0;
}

// The step-size kernels keep the base state of their trial sweeps in registers when a lane holds at most
// kAdRegBaseMaxEpl slots (extra_kernels.cuh): prefer such a shape (more lanes per chain) when one holds d.
static const Entry* choose_adaptive_entry(const hmcb_model& hm) {
  if (const Entry* e = choose_entry_quiet(hm, kAdRegBaseMaxEpl)) return e;
  return choose_entry(hm);
}

size_t hmcb_adhmc_eps0_workspace_bytes(const hmcb_model* model, int64_t n_chains) {
  if (check_model(model) || n_chains < 1) return 0;
  const Entry* e = choose_adaptive_entry(*model);
  if (!e) return 0;
  int64_t ft;
  int epl;
  if (ad_frame_threads(e, model, n_chains, ft, epl)) return 0;
  return (size_t)4 * epl * ft * (model->dtype == HMCB_F64 ? 8 : 4);
}

int hmcb_adhmc_eps0_step(const hmcb_model* model, const hmcb_chains* chains, const hmcb_rng* rng, int32_t n_leap,
                         double tint, const double* steps, int32_t ntry, int32_t nsteps, const hmcb_adhmc_out* out,
                         void* workspace, size_t workspace_bytes, void* stream) {
  int rc;
  if ((rc = check_model(model))) return rc;
  if ((rc = check_chains(chains))) return rc;
  HMCB_CHECK_ARG(rng != nullptr && steps != nullptr, "rng/steps is NULL");
  HMCB_CHECK_ARG(ntry >= 2 && ntry <= kMaxTry, "ntry=%d must be in 2..%d", ntry, kMaxTry);
  HMCB_CHECK_ARG((n_leap >= 1) != (tint > 0), "give either n_leap >= 1 or tint > 0 (src/adaptivehmc.py:164-171), got n_leap=%d tint=%g", n_leap, tint);
  HMCB_CHECK_ARG(nsteps >= 1 && tint >= 0, "nsteps must be >= 1 and tint >= 0");
  HMCB_CHECK_ARG(!rng->u_inject || rng->u_stride >= 3, "u_inject needs u_stride >= 3");
  if (chains->n_chains == 0) return 0;
  const Entry* e = choose_adaptive_entry(*model);
  if (!e) return HMCB_E_UNSUPPORTED;
  int64_t ft;
  int epl;
  if ((rc = ad_frame_threads(e, model, chains->n_chains, ft, epl))) return rc;
  const size_t need = (size_t)4 * epl * ft * (model->dtype == HMCB_F64 ? 8 : 4);
  if (!workspace || workspace_bytes < need) {
    set_error("adhmc workspace too small: have %zu bytes, need %zu", workspace ? workspace_bytes : (size_t)0, need);
    return HMCB_E_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (model->dtype == HMCB_F64)
    return adhmc_step_t<double>(e, model, chains, rng, n_leap, tint, steps, ntry, nsteps, out, workspace, ft, st);
  return adhmc_step_t<float>(e, model, chains, rng, n_leap, tint, steps, ntry, nsteps, out, workspace, ft, st);
}

size_t hmcb_adhmc_walk_workspace_bytes(const hmcb_model* model, int64_t n_chains, int32_t n_ratios) {
  if (check_model(model) || n_chains < 0 || n_ratios < 1 || n_ratios > kMaxRatios) return 0;
  const Entry* e = choose_adaptive_entry(*model);
  if (!e) return 0;
  int64_t ft;
  int epl;
  if (ad_frame_threads(e, model, n_chains, ft, epl, OP_QUERY_WALK)) return 0;
  return (size_t)(3 + 2 * n_ratios) * epl * ft * (model->dtype == HMCB_F64 ? 8 : 4);
}

int hmcb_adhmc_walk_step(const hmcb_model* model, const hmcb_chains* chains, const hmcb_rng* rng,
                         const hmcb_adhmc_walk_par* par, const hmcb_adhmc_walk_out* out, void* workspace,
                         size_t workspace_bytes, void* stream) {
  int rc;
  if ((rc = check_model(model))) return rc;
  if ((rc = check_chains(chains))) return rc;
  HMCB_CHECK_ARG(rng != nullptr && par != nullptr && par->steps != nullptr, "rng/par/steps is NULL");
  HMCB_CHECK_ARG(par->n_segments >= 2, "n_segments=%d must be >= 2 (the reference needs at least one step-size choice)",
                 par->n_segments);
  HMCB_CHECK_ARG(par->tint > 0 || par->n_leap_segment >= 1, "n_leap_segment must be >= 1 when tint == 0");
  HMCB_CHECK_ARG(par->n_ratios >= 1 && par->n_ratios <= kMaxRatios, "n_ratios=%d must be in 1..%d", par->n_ratios, kMaxRatios);
  HMCB_CHECK_ARG(par->ntry >= 2 && par->ntry <= kMaxTry, "ntry=%d must be in 2..%d", par->ntry, kMaxTry);
  HMCB_CHECK_ARG(par->nsteps >= 1, "nsteps must be >= 1");
  HMCB_CHECK_ARG(par->smin > 0 && par->smax > par->smin, "need 0 < smin < smax");
  for (int j = 0; j < par->n_ratios; ++j) {
    HMCB_CHECK_ARG(par->ratios[j] > 0 && (j == 0 || par->ratios[j] > par->ratios[j - 1]), "ratios must be positive and increasing");
    bool has_inverse = false;
    for (int k = 0; k < par->n_ratios; ++k)
      if (fabs(par->ratios[j] * par->ratios[k] - 1.0) <= 1e-8 + 1e-5) has_inverse = true;
    HMCB_CHECK_ARG(has_inverse, "ratios[%d]=%g has no reciprocal in ratios: the reverse index of the reference is undefined", j,
                   par->ratios[j]);
  }
  HMCB_CHECK_ARG(!rng->u_inject || rng->u_stride >= par->n_segments + 2, "u_inject needs u_stride >= n_segments + 2");
  if (chains->n_chains == 0) return 0;
  const Entry* e = choose_adaptive_entry(*model);
  if (!e) return HMCB_E_UNSUPPORTED;
  int64_t ft;
  int epl;
  if ((rc = ad_frame_threads(e, model, chains->n_chains, ft, epl, OP_QUERY_WALK))) return rc;
  const size_t need = (size_t)(3 + 2 * par->n_ratios) * epl * ft * (model->dtype == HMCB_F64 ? 8 : 4);
  if (!workspace || workspace_bytes < need) {
    set_error("adhmc walk workspace too small: have %zu bytes, need %zu", workspace ? workspace_bytes : (size_t)0, need);
    return HMCB_E_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (model->dtype == HMCB_F64) return adwalk_step_t<double>(e, model, chains, rng, par, out, workspace, ft, st);
  return adwalk_step_t<float>(e, model, chains, rng, par, out, workspace, ft, st);
}

}  // extern "C"
