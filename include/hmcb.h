/* hmcb.h -- C ABI of libhmcb.so, the B200 (sm_100a) batched-chain HMC engine.
 *
 * Drop-in boundary for the sampler hot path of modichirag/hmc.  The reference is
 * pure Python, so there is no FFI in it to bind; each entry point below replaces
 * the per-chain Python call named in its comment (path:line under the reference
 * tree) with ONE call over a whole batch of chains.  hmc_b200/_lib.py is the
 * ctypes binding; INTEGRATION.md shows the stub a reference maintainer would add.
 *
 * Conventions
 *  - Plain pointers and sizes only.  Every buffer is DEVICE memory owned by the
 *    caller; the library allocates no device memory.  What it does keep between
 *    calls, per process: a thread-local error string; per device, created on first
 *    use and never destroyed: the helper streams and fork/join events of the staged
 *    delayed-rejection path (see hmcb_drhmc_step; their enqueueing is serialised by
 *    a process-wide mutex) and the cached device address of the built-in scratch of
 *    hmcb_drhmc_step_static (a fixed array inside the library image).
 *  - All work is enqueued on `stream` (a cudaStream_t passed as void*); no call
 *    synchronises the host.  Errors detected at launch are returned.
 *  - Return value: 0 on success, a negative HMCB_E_* on failure;
 *    hmcb_last_error() then describes it.
 *  - Chain state is addressed as base[chain*stride_chain + dim*stride_dim]
 *    (strides in elements): (d,1) is the row-major [C,d] the Python API takes,
 *    (1,C) is the structure-of-arrays [d][C] the samplers keep between steps.
 *  - Element type of q, p, energies and probabilities is model->dtype.
 */
#ifndef HMCB_H_
#define HMCB_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HMCB_VERSION 1

enum { HMCB_OK = 0, HMCB_E_ARG = -1, HMCB_E_UNSUPPORTED = -2, HMCB_E_CUDA = -3, HMCB_E_WORKSPACE = -4 };

enum { HMCB_F64 = 0, HMCB_F32 = 1 };

/* Model densities with in-kernel analytic gradients.
 * They stand where the reference takes log_prob/grad_log_prob callables
 * (src/algorithms.py:81-89, src/pyhmc.py:7-10). */
enum { HMCB_MODEL_FUNNEL = 0, HMCB_MODEL_GAUSSIAN = 1, HMCB_MODEL_ROSENBROCK = 2, HMCB_MODEL_LOGISTIC = 3 };

typedef struct hmcb_model {
  int32_t kind;               /* HMCB_MODEL_* */
  int32_t d;                  /* dimension of one chain */
  int32_t dtype;              /* HMCB_F64 / HMCB_F32 */
  int32_t n_data;             /* logistic: rows of X; else 0 */
  double par[4];              /* funnel: {sigma}; rosenbrock: {a, b, T}; logistic: {tau} */
  const void* dev0;           /* gaussian: precision P[d*d] row-major (dtype); logistic: X[n_data*d] */
  const void* dev1;           /* logistic: y[n_data] */
  const void* invmetric_diag; /* [d] (dtype) or NULL for the unit metric (src/pyhmc.py:14-19) */
} hmcb_model;

typedef struct hmcb_chains {
  int64_t n_chains;
  int64_t chain_offset;       /* global id of chain 0; keys the RNG so results do not depend on sharding */
  const void* q_in;           /* current positions */
  void* q_out;                /* next positions; may equal q_in */
  int64_t stride_chain;       /* element strides of q_in, q_out, p_out, p_inject */
  int64_t stride_dim;
} hmcb_chains;

/* Random draws.  Production: Philox4x32-10 keyed by `seed`, counter = (global
 * chain id, iteration, stream, block).  Parity ("replay") mode: p_inject and/or
 * u_inject non-NULL feed the draws the reference would have taken from
 * np.random.normal / np.random.uniform (src/pyhmc.py:66,59). */
enum { HMCB_RNG_NORMAL_F64 = 1 };  /* flags: Box-Muller evaluated in f64 instead of f32 */
typedef struct hmcb_rng {
  uint64_t seed;
  uint64_t iteration;
  const void* p_inject;       /* standard normals, same layout as q (dtype), or NULL */
  const double* u_inject;     /* uniforms [n_chains, u_stride] or NULL */
  int32_t u_stride;
  int32_t flags;
  /* Partial momentum refresh of the generalised-HMC family (src/algorithms.py:269,603):
   * momentum = alpha * p_prev + sqrt(1 - alpha^2) * draw.  p_prev NULL = full refresh. */
  const void* p_prev;         /* previous momentum, layout of q, or NULL */
  double alpha;
} hmcb_rng;

/* Per-chain outputs of one transition.  Any pointer may be NULL (not written). */
typedef struct hmcb_step_out {
  void* p_out;                /* momentum returned by the step (p1 if accepted else p0), layout of q */
  int32_t* code;              /* [C] accept code, convention per `rule` */
  void* energies;             /* [C,2] H0,H1 (hmc) */
  void* avec;                 /* [C,nprops] stage acceptance probabilities (drhmc) */
  int32_t* counts;            /* [C,3] Hcount, Vgcount, leapcount (src/pyhmc.py:69) */
  void* accept_prob;          /* [C] min(1, exp(H0-H1)) with NaN->0, 0 if q did not move
                                 (what src/algorithms.py:218-224 feeds dual averaging) */
  void* ptries;               /* [C,nprops] try probabilities, probabilistic DR conventions only
                                 (6th return value of multi_step, src/algorithms.py:706) */
  void* p0_out;               /* the momentum every chain STARTED from (after metric scaling and partial
                                 refresh: the p of src/pyhmc.py:66, p0 of src/algorithms.py:506), layout of q.
                                 Lets a test replay a production (Philox) transition through the reference:
                                 hmcb_hmc_step and hmcb_drhmc_step only; hmcb_logreg_hmc_step rejects it */
} hmcb_step_out;

/* Accept-rule conventions:
 *  PYHMC: codes 1/0/2, uniform consumed only if prob finite and q moved, reject iff u > min(1,prob)
 *         (src/pyhmc.py:50-61; same in src/adaptivehmc.py:63-74)
 *  ALG:   codes 1/0/-1, uniform always consumed (src/algorithms.py:175-189) */
enum { HMCB_RULE_PYHMC = 0, HMCB_RULE_ALG = 1 };

int hmcb_version(void);
const char* hmcb_last_error(void);

/* One vanilla HMC transition for every chain: momentum draw, n_leap leapfrog
 * steps (n_leap+1 gradient evaluations), H0/H1, accept test, counters -- one
 * fused kernel.  Replaces pyhmc.HMC.step / PyHMC.hmc_step (src/pyhmc.py:64-69,
 * 133-138), algorithms.HMC.step (src/algorithms.py:192-198) and
 * adaptivehmc.HMC.hmc_step (src/adaptivehmc.py:77-99) called once per chain. */
int hmcb_hmc_step(const hmcb_model* model, const hmcb_chains* chains, const hmcb_rng* rng,
                  int32_t n_leap, double step_size, int32_t rule, const hmcb_step_out* out, void* stream);

/* Leapfrog only: (q,p) -> (q',p') for every chain.  Replaces HMC.leapfrog
 * (src/pyhmc.py:35-48, src/algorithms.py:133-147).  p_in/p_out use the q layout. */
int hmcb_leapfrog(const hmcb_model* model, const hmcb_chains* chains, const void* p_in, void* p_out,
                  int32_t n_leap, double step_size, void* stream);

/* log_prob and grad_log_prob of every chain (model plug-in check; mirrors the
 * callables at src/algorithms.py:81-89).  logp [C], grad in the q layout. */
int hmcb_logp_grad(const hmcb_model* model, const hmcb_chains* chains, void* logp, void* grad, void* stream);

/* Delayed-rejection HMC transition, algorithms.DRHMC.step (src/algorithms.py:
 * 500-530) with its ghost recursion get_num (:474-497).  Stage j integrates
 * int(n_leap*stepfactor^j) steps of size step_size/stepfactor^j from (q0,p0);
 * chains rejected at stage j are compacted on the device and only they run
 * stage j+1.  code = j+1 (accepted at stage j) or 0.  `workspace` is caller-
 * owned scratch of at least hmcb_drhmc_workspace_bytes(); contents need not be
 * preserved between calls.
 * Large batches (>= 32,768 chains, no `profile`) are cut into 2 contiguous chain ranges whose
 * stage kernels run on helper streams owned by the library, forked from and joined back into
 * `stream` with events: the thin tail of one range's stage overlaps another range's next stage
 * (-20 % at 131k chains, -2.5 % at 2^20).  Results are bit-identical to the single-stream order; the caller still
 * sees all work ordered on `stream`.  The helper streams are created on the first such call per
 * device (make that call outside stream capture); HMCB_DR_SPLIT=1 in the environment turns the
 * splitting off. */
#define HMCB_MAX_PROPS 6
/* Which reference class's bookkeeping a DR step follows (same proposals and probabilities):
 *  HMCB_DR_ALGORITHMS  algorithms.DRHMC.step: accept iff u < alpha_j, code j+1 / 0 (src/algorithms.py:500-530)
 *  HMCB_DR_MULTISTEP   pyhmc.PyHMC_multistep.multi_step (src/pyhmc.py:564-598): first stage accepts iff
 *                      u <= alpha_0, code j / -1, and get_num re-evaluates H at its base (:541) so
 *                      Hcount is larger by one per ghost frame
 *  HMCB_DR_PROB_METROP   probabilistic DR, algorithms.Prob_DRHMC_Metrop.multi_step (src/algorithms.py:687-728;
 *                      same body: pyhmc.PyHMC_multistep_tries, src/pyhmc.py:308-349): MULTISTEP's conventions, and
 *                      stage j >= 1 is attempted only if a further uniform u_try <= ptries[j-1], with
 *                      ptries[j] = 1 - min(1, e^{H0-Hj}); code -2 when a stage is declined; the ghost products
 *                      carry prod(ptries) (get_num, :659-684)
 *  HMCB_DR_PROB_HASTINGS the same with ptries[j] = 1 - alpha_j (algorithms.Prob_DRHMC_Hastings, :744-818;
 *                      pyhmc.PyHMC_multistep_tries3)
 * Injected uniforms are read in the order the reference consumes them: [C, nprops] = u_acc per stage for the first
 * two conventions, [C, 2*nprops-1] = u_acc0, u_try1, u_acc1, u_try2, ... for the probabilistic ones. */
enum { HMCB_DR_ALGORITHMS = 0, HMCB_DR_MULTISTEP = 1, HMCB_DR_PROB_METROP = 2, HMCB_DR_PROB_HASTINGS = 3 };
/* Optional per-stage timing: cudaEvent_t handles (as void*) recorded on `stream`
 * before stage 0, between stages and after the last one (nprops+1 events).  The
 * reference's only timing is time.time() around its sample loop
 * (scripts_old/funnel/funnelmpi.py:148,164-165). */
typedef struct hmcb_profile {
  void* events[HMCB_MAX_PROPS + 1];
} hmcb_profile;
size_t hmcb_drhmc_workspace_bytes(const hmcb_model* model, int64_t n_chains, int32_t nprops);
int hmcb_drhmc_step(const hmcb_model* model, const hmcb_chains* chains, const hmcb_rng* rng,
                    int32_t n_leap, double step_size, int32_t nprops, double stepfactor, int32_t convention,
                    const hmcb_step_out* out, void* workspace, size_t workspace_bytes,
                    const hmcb_profile* profile /* may be NULL */, void* stream);

/* K delayed-rejection transitions of every chain in ONE call: the loop of DRHMC.sample
 * (src/algorithms.py:542-552; scripts_old/funnel/funnelmpi_multistep.py runs the same loop over
 * PyHMC_multistep.multi_step) with the chain state kept on the device.  Each transition is
 * hmcb_drhmc_step's, bit for bit (transition k of a chain draws with iteration rng->iteration + k).
 * Chains are independent, so the batch is cut into up to 8 contiguous chain ranges, each running
 * its K transitions back to back on a stream of its own (the caller's and helper streams the
 * library owns, forked from and joined into `stream`): the thin last round of one range's stage
 * kernel is filled by another range's kernels -- of the same or of a later transition -- so the
 * only tail left is the one at the very end of the call.  The reference keeps every sample
 * (state.samples, :548); at 2^20 chains that does not fit, so the running moments the diagnostics
 * need are folded in instead: after every transition moment_sum += q, moment_sumsq += q*q
 * (contiguous [C,d], model dtype; as hmcb_moments_accumulate).  out: code / avec / ptries / p_out /
 * p0_out of the LAST transition, counts summed over the K transitions; n_accept[c] += transitions
 * accepted at any stage (needs out->code).  Injected draws and p_prev need n_transitions == 1.
 * At short trajectories (n_leap < 48, >= 2^19 chains, float64 thread-per-chain shapes, no moments
 * requested) the call runs differently: a chain's consecutive transitions stay in registers for as
 * long as the first stage accepts (k_dr0_multi_ws), the stage kernels finish the transitions that it
 * does not, pass after pass in chunks of 5 transitions -- same results, 0.79 -> 0.69 ms per transition
 * on the d=50 funnel at n_leap=20.
 * Workspace: hmcb_drhmc_sample_workspace_bytes().  Make the first call per device outside stream
 * capture (helper streams are created then); HMCB_SAMPLE_SPLIT=n in the environment sets the
 * number of ranges (1..16). */
typedef struct hmcb_sample_io {
  int32_t n_transitions;      /* K >= 1 */
  void* moment_sum;           /* [C,d] or NULL (both or neither) */
  void* moment_sumsq;         /* [C,d] or NULL */
  int32_t* n_accept;          /* [C] or NULL */
} hmcb_sample_io;
size_t hmcb_drhmc_sample_workspace_bytes(const hmcb_model* model, int64_t n_chains, int32_t nprops);
int hmcb_drhmc_sample(const hmcb_model* model, const hmcb_chains* chains, const hmcb_rng* rng,
                      int32_t n_leap, double step_size, int32_t nprops, double stepfactor, int32_t convention,
                      const hmcb_sample_io* io, const hmcb_step_out* out, void* workspace, size_t workspace_bytes,
                      void* stream);

/* K vanilla HMC transitions of every chain in ONE call: the loop of HMC.sample (src/algorithms.py:244-250;
 * scripts_old/funnel/funnelmpi.py:152-162 runs it over PyHMC.hmc_step).  Each transition is
 * hmcb_hmc_step's, bit for bit (transition k draws with iteration rng->iteration + k).  For float64
 * thread-per-chain shapes with the unit metric (e.g. the d=50 funnel) and no moments requested the
 * whole loop is one launch: a chain's position stays in registers, the start point of the running
 * transition in shared memory (hmc_b200/csrc/transition_ws.cuh, k_hmc_multi_ws) -- at L = 20 a
 * transition by single launches spends more time loading and storing q than integrating.  Other
 * shapes, and calls that fold the running moments, run K launches inside this one call.  io and out
 * as for hmcb_drhmc_sample: out holds the LAST transition's code / energies / p_out / accept_prob,
 * counts = K x (2, n_leap+1, 1); p0_out is not provided for K > 1. */
int hmcb_hmc_sample(const hmcb_model* model, const hmcb_chains* chains, const hmcb_rng* rng,
                    int32_t n_leap, double step_size, int32_t rule, const hmcb_sample_io* io,
                    const hmcb_step_out* out, void* stream);

/* hmcb_drhmc_step without a caller-supplied workspace.  No counterpart in the
 * reference (its Python never manages device memory); added for hosts that cannot
 * give us scratch: a thin C / ctypes caller that only holds the q and output buffers
 * some other framework handed it and has no device allocator of its own, or a process
 * whose device memory is already fully reserved by the application's pool, or a
 * sampling loop where cudaMalloc is unwelcome because it synchronises the device.
 * The stage lists, per-chain running state and proposal frames then live in an array
 * built into the library, HMCB_STATIC_SCRATCH_BYTES large, which every GPU that loads
 * libhmcb.so holds until the process exits.  This entry point allocates nothing.
 *  - Limit: the call is refused (HMCB_E_WORKSPACE, with the byte counts in
 *    hmcb_last_error()) when hmcb_drhmc_workspace_bytes(model, n_chains, nprops) would
 *    exceed HMCB_STATIC_SCRATCH_BYTES -- e.g. about 19,000 chains of the d=50 funnel at
 *    nprops=3.  Larger batches: split them, or use hmcb_drhmc_step.
 *  - The first call on each device looks up the array's device address and caches it;
 *    make that first call outside any stream capture.
 *  - All calls share the one array and it keeps whatever the previous call left; the
 *    stage counters are cleared at the start of every call, nothing else needs to be.
 *    Two calls whose kernels can overlap on the GPU (different streams, or two host
 *    threads on one device) race on it: serialise them.
 * Same arguments, outputs and results as hmcb_drhmc_step. */
#define HMCB_STATIC_SCRATCH_BYTES (32u << 20)
int hmcb_drhmc_step_static(const hmcb_model* model, const hmcb_chains* chains, const hmcb_rng* rng,
                           int32_t n_leap, double step_size, int32_t nprops, double stepfactor, int32_t convention,
                           const hmcb_step_out* out, void* stream);

/* Streamed sample moments (the reference stores samples[T,C,d] and post-processes,
 * src/utils.py:8-27,207-237; at 2^20 chains that does not fit, so each transition is
 * folded into per-chain running sums).  sum/sumsq are contiguous [C,d] (dtype),
 * updated in place: sum += q, sumsq += q*q. */
int hmcb_moments_accumulate(int32_t dtype, int64_t n_chains, int32_t d, const void* q, int64_t stride_chain,
                            int64_t stride_dim, void* sum, void* sumsq, void* stream);
/* Cross-chain reduction of per-chain mean m and variance s2 (ddof=1) over n_samples:
 * out[4*d] doubles (device) = sum_c m, sum_c m^2, sum_c s2, sum_c (m - true_mean)^2 per
 * dimension -- the quantities utils.get_ess_true (src/utils.py:23-27) and R-hat need;
 * this small vector is what is all-reduced across GPUs.  true_mean [d] may be NULL (0). */
int hmcb_moments_reduce(int32_t dtype, int64_t n_chains, int32_t d, const void* sum, const void* sumsq,
                        double n_samples, const double* true_mean, double* out, void* stream);

/* MALA, mala.MALA.step + metropolis (src/mala.py:45-60,24-42), n_steps >= 1 steps
 * fused per launch.  V(q) and grad V(q) are carried between calls as the reference
 * carries them through its return values (:60): pass V_in/Vg_in = NULL on the first
 * call.  rng->p_inject holds the standard normals z (n_steps == 1 only); codes 1/0/2.
 * The reference's MALA ignores its invmetric_diag; so does this. */
typedef struct hmcb_mala_io {
  const void* V_in;    /* [C] V(q) = -log_prob(q), or NULL */
  const void* Vg_in;   /* grad V(q), layout of q, or NULL */
  void* V_out;         /* [C] */
  void* Vg_out;        /* layout of q */
  int32_t* code;       /* [C] code of the last step */
  int32_t* n_accept;   /* [C] accepted steps among n_steps */
  int32_t* counts;     /* [C,3] Hcount (0), Vgcount, leapcount (0) */
} hmcb_mala_io;
int hmcb_mala_step(const hmcb_model* model, const hmcb_chains* chains, const hmcb_rng* rng, double eps,
                   int32_t n_steps, const hmcb_mala_io* io, void* stream);

/* Generalised HMC, algorithms.GHMC.step (src/algorithms.py:265-281), and its persistent-uniform
 * variant Persistent_GHMC.step (:314-333): momentum partially refreshed from rng->p_prev with
 * rng->alpha, potential and gradient cached between transitions as leapfrog_Vgq does (:150-172),
 * momentum returned flipped unless the proposal is accepted.  Codes 1/0/-1.  Counts follow the
 * reference built with both grad_log_prob and log_prob_and_grad: Vgcount = n_leap (n_leap+1
 * without a cached gradient), Hcount = 0 (1 without a cached potential).  u_in non-NULL selects
 * the persistent variant: u <- (u_in + 1 + delta) mod 2 - 1 is the Metropolis variable and
 * u_out = u / exp(H0-H1) on acceptance, else u. */
typedef struct hmcb_ghmc_io {
  void* p_out;         /* layout of q */
  const void* V_in;    /* [C] or NULL */
  const void* Vg_in;   /* layout of q, or NULL */
  void* V_out;         /* [C] */
  void* Vg_out;        /* layout of q */
  const double* u_in;  /* [C] or NULL */
  double* u_out;       /* [C] */
  double delta;
  int32_t* code;       /* [C] */
  void* energies;      /* [C,2] */
  int32_t* counts;     /* [C,3] */
} hmcb_ghmc_io;
int hmcb_ghmc_step(const hmcb_model* model, const hmcb_chains* chains, const hmcb_rng* rng, int32_t n_leap,
                   double step_size, const hmcb_ghmc_io* io, void* stream);

/* n_transitions consecutive GHMC / Persistent_GHMC transitions, the loop of GHMC.sample and Persistent_GHMC.sample
 * (src/algorithms.py:284-303, :336-357) without the stored history: transition k draws with iteration
 * rng->iteration + k and starts from the state, momentum, cached V / grad V and persistent uniform the one before
 * left.  rng->p_prev is the momentum the first transition refreshes (required); io->V_in / Vg_in / u_in as for
 * hmcb_ghmc_step; io->p_out, V_out, Vg_out (and u_out with u_in) are required and hold the final values; code and
 * energies are those of the LAST transition, counts the sums over all of them, n_accept[C] (or NULL) is INCREMENTED by
 * the number of accepted transitions.  Shapes that have the kernel run all transitions in ONE launch with q, p and the
 * gradient in registers (with a short trajectory a single launch moves six state arrays per gradient); the others
 * (dense Gaussian) launch once per transition, in place.  Either way the result is that of n_transitions calls of
 * hmcb_ghmc_step chained through their outputs: bit-identical when both run the same lanes-per-chain shape (always
 * for d < 32; from d = 32 the step takes 4 lanes per chain for its memory traffic and the fused launch the fewest that
 * hold d, so in-chain sums round differently in the last place -- the environment variable HMCB_LPC pins one shape
 * for both).  Injected draws: n_transitions == 1 only. */
int hmcb_ghmc_sample(const hmcb_model* model, const hmcb_chains* chains, const hmcb_rng* rng, int32_t n_leap,
                     double step_size, const hmcb_ghmc_io* io, int32_t n_transitions, int32_t* n_accept, void* stream);

/* AdHMC_eps0.hmc_step (src/adaptivehmc.py:138-192) and its step-size draw get_stepsize
 * (:111-134).  Give either n_leap >= 1 (and tint = 0) or tint = Tint > 0 (and n_leap = 0): with
 * Tint the trajectory takes int(Tint/eps) + 1 steps of the drawn step size (:178-179).  `steps` is a HOST array of the ntry trial step sizes
 * (np.logspace(log10(smin), log10(smax), ntry), computed by the caller so that the
 * values are bit-identical to numpy's).  Uniform draws, in the reference's order:
 * u[0] bin draw, u[1] within-bin draw, u[2] accept test.  Codes 1/0/2; where the
 * reference would raise IndexError (all trial weights NaN) the chain gets code 2. */
typedef struct hmcb_adhmc_out {
  void* p_out;         /* layout of q */
  int32_t* code;       /* [C] */
  void* probs;         /* [C,3] prob, prob/hastings, hastings */
  void* aux;           /* [C,3] pf_den, pb_num, eps */
  int32_t* counts;     /* [C,3] */
} hmcb_adhmc_out;
size_t hmcb_adhmc_eps0_workspace_bytes(const hmcb_model* model, int64_t n_chains);
int hmcb_adhmc_eps0_step(const hmcb_model* model, const hmcb_chains* chains, const hmcb_rng* rng, int32_t n_leap,
                         double tint, const double* steps, int32_t ntry, int32_t nsteps, const hmcb_adhmc_out* out,
                         void* workspace, size_t workspace_bytes, void* stream);

/* AdHMC.hmc_step (src/adaptivehmc.py:227-354) and AdHMC_tint.hmc_step (:394-519): the step size
 * walks inside the trajectory.  n_segments = Nleap // nleap (AdHMC, every segment n_leap_segment
 * steps) or ntimes (AdHMC_tint, tint = Tint/ntimes > 0 and a segment of step size e is
 * max(1, int(tint/e)) steps).  After each segment every ratio proposes e*ratio, one is chosen with
 * probability ~ pwts0[j] * exp(H1 - H_j) (np.random.choice, :303) and the reverse walk's probability
 * of the old step size enters the Hastings factor; the first step size and the last reverse
 * probability come from get_stepsize's trial sweep (:202-224) over the HOST array `steps`.
 * `ratios` must be increasing and closed under reciprocal (for every j some j' has
 * ratios[j]*ratios[j'] == 1 to np.allclose's tolerance), as the reference's default
 * [1/sqrt2, sqrt2]; otherwise the reference's reverse index is undefined.
 * Uniforms, in the reference's order: u[0] bin draw, u[1] within-bin draw, u[2+i] the choice after
 * segment i (i < n_segments-1), u[n_segments+1] accept test.  Codes 1/0/2, and 100+i when the
 * forward weights of segment i are all zero / NaN (:296-301; (q0,p0) returned, probs NaN). */
#define HMCB_MAX_RATIOS 4
typedef struct hmcb_adhmc_walk_par {
  int32_t n_segments;
  int32_t n_leap_segment;
  double tint;                       /* 0: AdHMC; > 0: AdHMC_tint */
  int32_t n_ratios;
  int32_t ntry, nsteps;              /* trial sweep: ntry_eps0, nsteps_eps0 */
  double ratios[HMCB_MAX_RATIOS], pwts0[HMCB_MAX_RATIOS];
  double smin, smax;
  const double* steps;               /* host, [ntry]: np.logspace / np.linspace(smin, smax, ntry) */
} hmcb_adhmc_walk_par;
typedef struct hmcb_adhmc_walk_out {
  void* p_out;         /* layout of q */
  int32_t* code;       /* [C] */
  void* probs;         /* [C,3] prob, prob/hastings, hastings */
  void* checks;        /* [C,3,n_segments] pf_den, pb_num, eps (np.stack at :347) */
  int32_t* counts;     /* [C,3] */
} hmcb_adhmc_walk_out;
size_t hmcb_adhmc_walk_workspace_bytes(const hmcb_model* model, int64_t n_chains, int32_t n_ratios);
int hmcb_adhmc_walk_step(const hmcb_model* model, const hmcb_chains* chains, const hmcb_rng* rng,
                         const hmcb_adhmc_walk_par* par, const hmcb_adhmc_walk_out* out,
                         void* workspace, size_t workspace_bytes, void* stream);

/* Bayesian logistic regression, log p = sum_i [y_i z_i - softplus(z_i)] - |theta|^2/(2 tau^2),
 * z = X theta (builder-defined model): float32 chains, the
 * gradient X^T(sigmoid(X theta) - y) for all chains at once on the tcgen05 tensor cores
 * (hmc_b200/csrc/logreg.cuh).  model: kind LOGISTIC, dtype F32, d <= 128, n_data rows,
 * par[0] = tau, dev0 = X float32 [n_data, d] row-major, dev1 = y float32 [n_data].
 *  - hmcb_logreg_prepare: once per data set, builds the fp16 hi/lo operand blocks of X in a
 *    caller-owned buffer of hmcb_logreg_prepared_bytes().
 *  - hmcb_logreg_hmc_step: one HMC transition per chain, same semantics and outputs as
 *    hmcb_hmc_step (replaces pyhmc.HMC.step / algorithms.HMC.step, src/pyhmc.py:64-69,
 *    src/algorithms.py:192-198); chain state must be contiguous row-major [C, d] float32;
 *    energies are accumulated in float64 and returned as float32.
 *  - hmcb_logreg_logp_grad: log_prob [C] and grad_log_prob [C, d] (the model plug-in pair,
 *    src/algorithms.py:81-89). */
size_t hmcb_logreg_prepared_bytes(const hmcb_model* model);
int hmcb_logreg_prepare(const hmcb_model* model, void* prepared, size_t prepared_bytes, void* stream);
size_t hmcb_logreg_workspace_bytes(const hmcb_model* model, int64_t n_chains);
int hmcb_logreg_hmc_step(const hmcb_model* model, const void* prepared, const hmcb_chains* chains,
                         const hmcb_rng* rng, int32_t n_leap, double step_size, int32_t rule,
                         const hmcb_step_out* out, void* workspace, size_t workspace_bytes, void* stream);
int hmcb_logreg_logp_grad(const hmcb_model* model, const void* prepared, const void* theta, int64_t n_chains,
                          void* logp, void* grad, void* workspace, size_t workspace_bytes, void* stream);
/*  - hmcb_logreg_mala_step: n_steps MALA steps per chain over the same tensor-core gradient
 *    (mala.MALA.step + metropolis, src/mala.py:45-60,24-42; one gradient sweep per step), semantics
 *    and outputs as hmcb_mala_step with two differences: the cached potential V_in / V_out is
 *    FLOAT64 [C] (the reference carries V(q) as a Python float; float32 would lose the Hastings
 *    ratio's low bits at V ~ 1e5), Vg_in / Vg_out are float32 [C,d].  Workspace:
 *    hmcb_logreg_workspace_bytes(). */
int hmcb_logreg_mala_step(const hmcb_model* model, const void* prepared, const hmcb_chains* chains,
                          const hmcb_rng* rng, double eps, int32_t n_steps, const hmcb_mala_io* io,
                          void* workspace, size_t workspace_bytes, void* stream);

/* The draw contract, on the host (no GPU needed), so that tests can pin it:
 *  - hmcb_philox4x32_10: the raw Philox4x32-10 block function (Salmon et al. 2011), checked
 *    against the published Random123 known-answer vectors;
 *  - hmcb_philox_block: the block a kernel draws for (seed, global chain id, iteration,
 *    stream, block): key = seed, counter = (chain lo, chain hi, iteration, stream<<24 | block).
 *    Streams: 0 momentum normals (block b -> elements 4b..4b+3 with the f32 generator; blocks
 *    2b, 2b+1 -> elements 4b..4b+1, 4b+2..4b+3 with HMCB_RNG_NORMAL_F64), 1 accept uniforms
 *    (uniform i = words (0,1) of block i/2 if i is even, words (2,3) if odd), 2 step-size draws;
 *  - hmcb_u53: the 53-bit uniform in [0,1) made from two words.
 * They stand where the reference calls np.random.normal / np.random.uniform (src/pyhmc.py:66,59). */
void hmcb_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);
void hmcb_philox_block(uint64_t seed, uint64_t chain, uint32_t iteration, uint32_t stream, uint32_t block,
                       uint32_t out[4]);
double hmcb_u53(uint32_t a, uint32_t b);

/* Measurement probe: a register-resident dependent-FMA loop (FP64 or FP32) used
 * to measure the pipe peak that roofline fractions are quoted against.
 * Launches `blocks` x `threads`, each thread doing iters*ilp*unroll FMAs; writes
 * one value per thread to `sink` (>= blocks*threads elements of dtype).  Returns
 * the FMA count of the launch through *fma_count. */
int hmcb_probe_fma(int32_t dtype, int32_t blocks, int32_t threads, int64_t iters, void* sink,
                   double* fma_count, void* stream);

/* Measurement probe: standard normals per second from the library's generator alone (Philox4x32-10 +
 * Box-Muller, 4 independent blocks in flight per thread as the kernels run it; f64_generator != 0 selects the
 * HMCB_RNG_NORMAL_F64 form).  The ceiling MALA (d normals per gradient) and short-trajectory HMC are quoted
 * against.  sink: >= blocks*threads doubles.  The normals generated are returned through *normal_count. */
int hmcb_probe_normals(int32_t f64_generator, int32_t blocks, int32_t threads, int64_t iters, void* sink,
                       double* normal_count, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* HMCB_H_ */
