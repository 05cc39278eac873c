// logreg.cu -- C ABI for the Bayesian logistic-regression model (float32 chains, tensor-core
// gradient).  V(theta) = sum_i [softplus(z_i) - y_i z_i] + |theta|^2/(2 tau^2), z = X theta.
// A leapfrog step here is two launches: the gradient sweep k_logreg_grad_v1 (logreg.cuh)
// and a small per-chain update kernel (fixed-order reduction of the slab partials, prior
// term, kick and drift).  The chain state is C x d float32 (8 MB at 16,384 x 128), so the
// update kernels are noise next to the 2 ms gradient sweep.
#include "logreg.cuh"

namespace hmcb {
namespace lr {

// ------------------------------------------------------------------ prepared model
// [0,256) header: int32 x_scale_log2, uint32 absmax bits | Xh blocks | Xl blocks | y padded
struct PrepLayout {
  int n_blocks;
  size_t off_xh, off_xl, off_y, total;
};
static PrepLayout prep_layout(int n_rows) {
  PrepLayout L;
  L.n_blocks = (n_rows + kTileR - 1) / kTileR;
  L.off_xh = 256;
  L.off_xl = L.off_xh + (size_t)L.n_blocks * kBlkBytes;
  L.off_y = L.off_xl + (size_t)L.n_blocks * kBlkBytes;
  L.total = L.off_y + (size_t)L.n_blocks * kTileR * sizeof(float);
  return L;
}

__global__ void k_absmax(const float* __restrict__ X, size_t n, uint32_t* out_bits) {
  float m = 0.f;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    m = fmaxf(m, fabsf(X[i]));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0) atomicMax(out_bits, __float_as_uint(m));   // non-negative floats order like uints
}
__global__ void k_set_scale(int32_t* hdr) {
  const float m = __uint_as_float(((uint32_t*)hdr)[1]);
  int ex = 0;
  if (m > 0.f && !isinf(m) && !isnan(m)) frexpf(m, &ex);
  hdr[0] = m > 0.f ? 10 - ex : 0;     // max |x'| lands in [2^9, 2^10)
}
__global__ void k_pad_y(const float* __restrict__ y, int n_rows, int n_pad, float* __restrict__ out) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_pad; i += gridDim.x * blockDim.x) out[i] = i < n_rows ? y[i] : 0.f;
}

// ------------------------------------------------------------------ workspace of one HMC step
struct WsLayout {
  int n_slabs, blocks_per_slab, n_tiles;
  size_t off_theta, off_p, off_p0, off_G, off_V, off_H0, total;
};
static size_t al(size_t x) { return (x + 255) / 256 * 256; }
static WsLayout ws_layout(int64_t n_chains, int d, int n_blocks) {
  WsLayout L;
  L.n_tiles = (int)((n_chains + kTileC - 1) / kTileC);
  // slabs: enough CTAs to fill the device several times over with little wave-quantisation loss,
  // each slab at least one flush group long
  const int sms = sm_count();
  int best = 1;
  double best_cost = 1e30;
  const int smax = n_blocks / kFlush > 0 ? (n_blocks / kFlush < 32 ? n_blocks / kFlush : 32) : 1;
  for (int s = 1; s <= smax; ++s) {
    const double ctas = (double)L.n_tiles * s;
    const double waves = ctas / sms;
    double cost = ceil(waves) / waves;                 // idle fraction of the last wave
    if (ctas < 2.0 * sms) cost += 1.0;                  // too few CTAs to balance
    cost += 0.002 * s;                                  // partials traffic grows with s
    if (cost < best_cost) { best_cost = cost; best = s; }
  }
  L.n_slabs = best;
  L.blocks_per_slab = (n_blocks + best - 1) / best;
  size_t off = 0;
  L.off_theta = off; off = al(off + (size_t)n_chains * d * 4);
  L.off_p = off;     off = al(off + (size_t)n_chains * d * 4);
  L.off_p0 = off;    off = al(off + (size_t)n_chains * d * 4);
  L.off_G = off;     off = al(off + (size_t)L.n_slabs * n_chains * kD * 4);
  L.off_V = off;     off = al(off + (size_t)L.n_slabs * n_chains * 8);
  L.off_H0 = off;    off = al(off + (size_t)n_chains * 8);
  L.total = off;
  return L;
}

// ------------------------------------------------------------------ small kernels
struct StepArgs {
  int64_t n_chains, chain_offset;
  int32_t d, n_slabs;
  const float* q_in;
  float* q_out;
  float *theta, *p, *p0;
  const float* Gpart;
  const double* Vpart;
  double* H0;
  uint64_t seed;
  uint32_t iteration;
  const float* p_inject;
  const double* u_inject;
  int32_t u_stride, rule, n_leap;
  float eps, inv_tau2;
  float* p_out;
  int32_t* code;
  float* energies;
  int32_t* counts;
  float* accept_prob;
  float* logp_out;   // grad-only mode
  float* grad_out;
};

// momentum draw (same Philox stream definition as every other kernel) and theta <- q_in
__global__ void k_lr_begin(const StepArgs a) {
  const int nblk = (a.d + 3) / 4;
  const int64_t n = a.n_chains * nblk;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t c = i / nblk;
    const int b = (int)(i - c * nblk);
    float z[4];
    if (a.p_inject) {
#pragma unroll
      for (int k = 0; k < 4; ++k) z[k] = (4 * b + k < a.d) ? a.p_inject[c * a.d + 4 * b + k] : 0.f;
    } else {
      uint32_t w[4];
      Philox::gen(a.seed, (uint64_t)(a.chain_offset + c), a.iteration, STREAM_MOMENTUM, (uint32_t)b, w);
      normal4_f32(w, z);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int col = 4 * b + k;
      if (col < a.d) {
        const int64_t o = c * a.d + col;
        a.p[o] = z[k];
        a.p0[o] = z[k];
        a.theta[o] = a.q_in[o];
      }
    }
  }
}

enum { PH_FIRST = 0, PH_MID = 1, PH_LAST = 2, PH_GRADONLY = 3 };

// one warp per chain: reduce the slab partials in fixed order, add the prior, kick (and drift)
template <int PHASE>
__global__ void __launch_bounds__(256) k_lr_update(const StepArgs a) {
  const int lane = threadIdx.x & 31;
  const int64_t warps = (int64_t)gridDim.x * (blockDim.x >> 5);
  for (int64_t c = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); c < a.n_chains; c += warps) {
    float th[4], g[4], pp[4];
    double th2 = 0.0, ke = 0.0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int col = lane + 32 * k;
      const bool ok = col < a.d;
      th[k] = ok ? a.theta[c * a.d + col] : 0.f;
      float gs = 0.f;
      for (int s = 0; s < a.n_slabs; ++s) gs += ok ? a.Gpart[((size_t)s * a.n_chains + c) * kD + col] : 0.f;
      g[k] = gs + th[k] * a.inv_tau2;                     // grad V = X^T(sigma - y) + theta / tau^2
      if (PHASE != PH_GRADONLY) pp[k] = ok ? a.p[c * a.d + col] : 0.f;
      th2 += (double)th[k] * th[k];
    }
    double V = 0.0;
    if (PHASE != PH_MID) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) th2 += __shfl_xor_sync(0xffffffffu, th2, o);
      for (int s = 0; s < a.n_slabs; ++s) V += a.Vpart[(size_t)s * a.n_chains + c];
      V += 0.5 * th2 * (double)a.inv_tau2;
    }
    if (PHASE == PH_GRADONLY) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int col = lane + 32 * k;
        if (col < a.d && a.grad_out) a.grad_out[c * a.d + col] = -g[k];     // grad log_prob
      }
      if (lane == 0 && a.logp_out) a.logp_out[c] = (float)(-V);
      continue;
    }
    if (PHASE == PH_FIRST) {
#pragma unroll
      for (int k = 0; k < 4; ++k) ke += (double)pp[k] * pp[k];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) ke += __shfl_xor_sync(0xffffffffu, ke, o);
      if (lane == 0) a.H0[c] = V + 0.5 * ke;
    }
    const float h = (PHASE == PH_MID) ? a.eps : 0.5f * a.eps;
#pragma unroll
    for (int k = 0; k < 4; ++k) pp[k] = pp[k] - h * g[k];      // p = p - h * V_g   (src/pyhmc.py:39,42,44)
    if (PHASE != PH_LAST) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int col = lane + 32 * k;
        if (col < a.d) {
          th[k] = th[k] + a.eps * pp[k];                         // q = q + eps * p
          a.theta[c * a.d + col] = th[k];
          a.p[c * a.d + col] = pp[k];
        }
      }
      continue;
    }
    // ---- last phase: energies, no-move test, accept test, outputs (src/pyhmc.py:50-61)
    double dq = 0.0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int col = lane + 32 * k;
      ke += (double)pp[k] * pp[k];
      if (col < a.d) dq += (double)(a.q_in[c * a.d + col] - th[k]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      ke += __shfl_xor_sync(0xffffffffu, ke, o);
      dq += __shfl_xor_sync(0xffffffffu, dq, o);
    }
    const double H0 = a.H0[c], H1 = V + 0.5 * ke;
    const double prob = exp(H0 - H1);
    const bool bad = isnan(prob) || isinf(prob) || dq == 0.0;
    double u;
    if (a.u_inject) {
      u = a.u_inject[c * a.u_stride];
    } else {
      uint32_t w[4];
      Philox::gen(a.seed, (uint64_t)(a.chain_offset + c), a.iteration, STREAM_ACCEPT, 0, w);
      u = u53(w[0], w[1]);
    }
    const double pm = prob < 1.0 ? prob : 1.0;
    const bool accept = !bad && !(u > pm);
    if (lane == 0) {
      if (a.code) a.code[c] = bad ? (a.rule == HMCB_RULE_PYHMC ? 2 : -1) : (accept ? 1 : 0);
      if (a.energies) { a.energies[2 * c] = (float)H0; a.energies[2 * c + 1] = (float)H1; }
      if (a.accept_prob) a.accept_prob[c] = accept ? (float)pm : 0.f;
      if (a.counts) { a.counts[3 * c] = 2; a.counts[3 * c + 1] = (a.n_leap > 1 ? a.n_leap : 1) + 1; a.counts[3 * c + 2] = 1; }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int col = lane + 32 * k;
      if (col < a.d) {
        const int64_t o = c * a.d + col;
        a.q_out[o] = accept ? th[k] : a.q_in[o];
        if (a.p_out) a.p_out[o] = accept ? pp[k] : a.p0[o];
      }
    }
  }
}

// ------------------------------------------------------------------ MALA over the tensor-core gradient
// mala.MALA.step + metropolis (src/mala.py:45-60, 24-42) for the logistic model: per step one gradient sweep at the
// proposal q1 = q - eps^2/2 grad V(q) + eps z, then the Hastings test; V(q) and grad V(q) are carried from step to step
// (and between calls, as the reference carries them through its return values).
struct MalaArgsLR {
  int64_t n_chains, chain_offset;
  int32_t d, n_slabs;
  float* q;            // current state, [C,d] (= q_out)
  float* q1;           // proposal (ws theta)
  float* g0;           // grad V(q) (ws)
  double* V0;          // V(q) (ws)
  const float* Gpart;
  const double* Vpart;
  uint64_t seed;
  uint32_t iteration;
  const float* z_inject;
  const double* u_inject;
  int32_t u_stride, step;
  float eps, inv_tau2;
  int32_t* code;
  int32_t* n_accept;
  int32_t* counts;
  int32_t first;       // first step of the call: n_accept / counts are assigned, not added to
  int32_t nvg0;        // gradient evaluations already spent on V_gq at the start point (0 or 1)
};

__global__ void k_lr_mala_propose(const MalaArgsLR a) {
  const int nblk = (a.d + 3) / 4;
  const int64_t n = a.n_chains * nblk;
  const float half_e2 = 0.5f * a.eps * a.eps;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t c = i / nblk;
    const int b = (int)(i - c * nblk);
    float z[4];
    if (a.z_inject) {
#pragma unroll
      for (int k = 0; k < 4; ++k) z[k] = (4 * b + k < a.d) ? a.z_inject[c * a.d + 4 * b + k] : 0.f;
    } else {
      uint32_t w[4];
      Philox::gen(a.seed, (uint64_t)(a.chain_offset + c), a.iteration + (uint32_t)a.step, STREAM_MOMENTUM, (uint32_t)b, w);
      normal4_f32(w, z);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int col = 4 * b + k;
      if (col < a.d) {
        const int64_t o = c * a.d + col;
        a.q1[o] = (a.q[o] - half_e2 * a.g0[o]) + a.eps * z[k];      // src/mala.py:54
      }
    }
  }
}

// one warp per chain.  INIT: V(q), grad V(q) from the sweep at q.  else: the Hastings test of one step.
template <bool INIT>
__global__ void __launch_bounds__(256) k_lr_mala_update(const MalaArgsLR a) {
  const int lane = threadIdx.x & 31;
  const int64_t warps = (int64_t)gridDim.x * (blockDim.x >> 5);
  const float half_e2 = 0.5f * a.eps * a.eps;
  for (int64_t c = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); c < a.n_chains; c += warps) {
    const float* x = INIT ? a.q : a.q1;       // the point the sweep was taken at
    float th[4], g[4];
    double th2 = 0.0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int col = lane + 32 * k;
      const bool ok = col < a.d;
      th[k] = ok ? x[c * a.d + col] : 0.f;
      float gs = 0.f;
      for (int s = 0; s < a.n_slabs; ++s) gs += ok ? a.Gpart[((size_t)s * a.n_chains + c) * kD + col] : 0.f;
      g[k] = gs + th[k] * a.inv_tau2;
      th2 += (double)th[k] * th[k];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) th2 += __shfl_xor_sync(0xffffffffu, th2, o);
    double V = 0.0;
    for (int s = 0; s < a.n_slabs; ++s) V += a.Vpart[(size_t)s * a.n_chains + c];
    V += 0.5 * th2 * (double)a.inv_tau2;
    if (INIT) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int col = lane + 32 * k;
        if (col < a.d) a.g0[c * a.d + col] = g[k];
      }
      if (lane == 0) a.V0[c] = V;
      continue;
    }
    // Hastings ratio as the reference forms it (src/mala.py:24-30)
    double ld = 0.0, ln = 0.0, dq = 0.0;
    float q0[4], g0[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int col = lane + 32 * k;
      const bool ok = col < a.d;
      q0[k] = ok ? a.q[c * a.d + col] : 0.f;
      g0[k] = ok ? a.g0[c * a.d + col] : 0.f;
      const float mu0 = q0[k] - half_e2 * g0[k], mu1 = th[k] - half_e2 * g[k];
      const double a0 = (double)th[k] - (double)mu0, a1 = (double)q0[k] - (double)mu1;
      ld += a0 * a0;
      ln += a1 * a1;
      dq += (double)q0[k] - (double)th[k];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      ld += __shfl_xor_sync(0xffffffffu, ld, o);
      ln += __shfl_xor_sync(0xffffffffu, ln, o);
      dq += __shfl_xor_sync(0xffffffffu, dq, o);
    }
    const double e2 = (double)a.eps * (double)a.eps;
    const double V0 = a.V0[c];
    const double prob = exp(V0 - V) * exp(-0.5 * ln / e2 + 0.5 * ld / e2);
    const bool bad = isnan(prob) || isinf(prob) || dq == 0.0;
    bool accept = false;
    if (!bad) {
      double u;
      if (a.u_inject) {
        u = a.u_inject[c * a.u_stride + a.step];
      } else {
        uint32_t w[4];
        Philox::gen(a.seed, (uint64_t)(a.chain_offset + c), a.iteration + (uint32_t)a.step, STREAM_ACCEPT, 0, w);
        u = u53(w[0], w[1]);
      }
      accept = !(u > (prob < 1.0 ? prob : 1.0));
    }
    if (accept) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int col = lane + 32 * k;
        if (col < a.d) { a.q[c * a.d + col] = th[k]; a.g0[c * a.d + col] = g[k]; }
      }
    }
    if (lane == 0) {
      if (accept) a.V0[c] = V;
      if (a.code) a.code[c] = bad ? 2 : (accept ? 1 : 0);
      if (a.n_accept) a.n_accept[c] = (a.first ? 0 : a.n_accept[c]) + (accept ? 1 : 0);
      if (a.counts) {
        a.counts[3 * c] = 0;
        a.counts[3 * c + 1] = (a.first ? a.nvg0 : a.counts[3 * c + 1]) + 1;
        a.counts[3 * c + 2] = 0;
      }
    }
  }
}

static int launch_grad(const hmcb_model& m, const unsigned char* prep, const PrepLayout& P, const WsLayout& W,
                       const float* theta, float* Gpart, double* Vpart, int64_t n_chains, bool want_v, cudaStream_t st) {
  GradArgsV1 g{};
  g.Xh_blk = (const __half*)(prep + P.off_xh);
  g.Xl_blk = (const __half*)(prep + P.off_xl);
  g.y = (const float*)(prep + P.off_y);
  g.theta = theta;
  g.Gpart = Gpart;
  g.Vpart = Vpart;
  g.n_chains = n_chains;
  g.d = m.d;
  g.n_rows = m.n_data;
  g.n_blocks = P.n_blocks;
  g.blocks_per_slab = W.blocks_per_slab;
  g.n_slabs = W.n_slabs;
  g.x_scale_dev = (const int32_t*)prep;
  dim3 grid(W.n_tiles, W.n_slabs);
  // the opt-in is a per-device attribute of the function: set on every call (cheap), as launch_any does, so that a
  // second device in the same process gets it too
  if (want_v) {
    HMCB_CUDA(cudaFuncSetAttribute(k_logreg_grad_v1<true, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemV1));
    k_logreg_grad_v1<true, 3><<<grid, kThreadsV1, kSmemV1, st>>>(g);
  } else {
    HMCB_CUDA(cudaFuncSetAttribute(k_logreg_grad_v1<false, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemV1));
    k_logreg_grad_v1<false, 3><<<grid, kThreadsV1, kSmemV1, st>>>(g);
  }
  HMCB_CUDA(cudaPeekAtLastError());
  return 0;
}

static int check_lr_model(const hmcb_model* m) {
  HMCB_CHECK_ARG(m != nullptr && m->kind == HMCB_MODEL_LOGISTIC, "not a logistic-regression model");
  HMCB_CHECK_ARG(m->dtype == HMCB_F32, "logistic regression runs in float32 (tensor-core gradient); got dtype %d", m->dtype);
  HMCB_CHECK_ARG(m->d >= 1 && m->d <= kD, "logistic regression supports 1 <= d <= %d features, got %d", kD, m->d);
  HMCB_CHECK_ARG(m->n_data >= 1, "logistic regression needs n_data >= 1 rows");
  HMCB_CHECK_ARG(m->dev0 && m->dev1, "logistic regression needs X in dev0 and y in dev1");
  HMCB_CHECK_ARG(m->par[0] > 0, "prior scale tau must be > 0");
  HMCB_CHECK_ARG(m->invmetric_diag == nullptr, "logistic regression supports the unit metric only");
  return 0;
}

}  // namespace lr
}  // namespace hmcb

using namespace hmcb;
using namespace hmcb::lr;

extern "C" {

size_t hmcb_logreg_prepared_bytes(const hmcb_model* model) {
  if (check_lr_model(model)) return 0;
  return prep_layout(model->n_data).total;
}

int hmcb_logreg_prepare(const hmcb_model* model, void* prepared, size_t bytes, void* stream) {
  int rc;
  if ((rc = check_lr_model(model))) return rc;
  const PrepLayout P = prep_layout(model->n_data);
  if (!prepared || bytes < P.total) {
    set_error("logreg prepared buffer too small: have %zu bytes, need %zu", prepared ? bytes : (size_t)0, P.total);
    return HMCB_E_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  unsigned char* base = (unsigned char*)prepared;
  HMCB_CUDA(cudaMemsetAsync(base, 0, 256, st));
  const size_t n = (size_t)model->n_data * model->d;
  k_absmax<<<sm_count() * 4, 256, 0, st>>>((const float*)model->dev0, n, (uint32_t*)base + 1);
  k_set_scale<<<1, 1, 0, st>>>((int32_t*)base);
  k_logreg_prep_blk<<<sm_count() * 8, 256, 0, st>>>((const float*)model->dev0, model->n_data, model->d, P.n_blocks, 0,
                                                    (const int32_t*)base, (__half*)(base + P.off_xh),
                                                    (__half*)(base + P.off_xl));
  k_pad_y<<<sm_count(), 256, 0, st>>>((const float*)model->dev1, model->n_data, P.n_blocks * kTileR, (float*)(base + P.off_y));
  HMCB_CUDA(cudaPeekAtLastError());
  return 0;
}

size_t hmcb_logreg_workspace_bytes(const hmcb_model* model, int64_t n_chains) {
  if (check_lr_model(model) || n_chains < 1) return 0;
  return ws_layout(n_chains, model->d, prep_layout(model->n_data).n_blocks).total;
}

int hmcb_logreg_hmc_step(const hmcb_model* model, const void* prepared, const hmcb_chains* chains, const hmcb_rng* rng,
                         int32_t n_leap, double step_size, int32_t rule, const hmcb_step_out* out, void* workspace,
                         size_t workspace_bytes, void* stream) {
  int rc;
  if ((rc = check_lr_model(model))) return rc;
  HMCB_CHECK_ARG(prepared != nullptr, "prepared model buffer is NULL (call hmcb_logreg_prepare first)");
  HMCB_CHECK_ARG(chains != nullptr && rng != nullptr, "chains/rng is NULL");
  HMCB_CHECK_ARG(chains->n_chains >= 0 && chains->n_chains < (int64_t)1 << 31, "n_chains out of range");
  if (chains->n_chains == 0) return 0;
  HMCB_CHECK_ARG(chains->q_in && chains->q_out, "q_in/q_out must be device pointers");
  HMCB_CHECK_ARG(chains->stride_dim == 1 && chains->stride_chain == model->d,
                 "logistic regression needs contiguous row-major [C,d] chain state");
  HMCB_CHECK_ARG(rule == HMCB_RULE_PYHMC || rule == HMCB_RULE_ALG, "unknown accept rule %d", rule);
  HMCB_CHECK_ARG(n_leap >= 0, "n_leap must be >= 0");
  HMCB_CHECK_ARG(!out || !out->p0_out, "p0_out is not provided by the logistic-regression step");
  const int64_t C = chains->n_chains;
  const PrepLayout P = prep_layout(model->n_data);
  const WsLayout W = ws_layout(C, model->d, P.n_blocks);
  if (!workspace || workspace_bytes < W.total) {
    set_error("logreg workspace too small: have %zu bytes, need %zu", workspace ? workspace_bytes : (size_t)0, W.total);
    return HMCB_E_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  unsigned char* ws = (unsigned char*)workspace;
  StepArgs a{};
  a.n_chains = C; a.chain_offset = chains->chain_offset; a.d = model->d; a.n_slabs = W.n_slabs;
  a.q_in = (const float*)chains->q_in; a.q_out = (float*)chains->q_out;
  a.theta = (float*)(ws + W.off_theta); a.p = (float*)(ws + W.off_p); a.p0 = (float*)(ws + W.off_p0);
  a.Gpart = (const float*)(ws + W.off_G); a.Vpart = (const double*)(ws + W.off_V); a.H0 = (double*)(ws + W.off_H0);
  a.seed = rng->seed; a.iteration = (uint32_t)rng->iteration;
  a.p_inject = (const float*)rng->p_inject; a.u_inject = rng->u_inject; a.u_stride = rng->u_stride;
  a.rule = rule; a.n_leap = n_leap; a.eps = (float)step_size;
  a.inv_tau2 = (float)(1.0 / (model->par[0] * model->par[0]));
  if (out) {
    a.p_out = (float*)out->p_out; a.code = out->code; a.energies = (float*)out->energies; a.counts = out->counts;
    a.accept_prob = (float*)out->accept_prob;
  }
  const int N = n_leap > 1 ? n_leap : 1;
  const int ublocks = (int)((C + 7) / 8 < 4096 ? (C + 7) / 8 : 4096);
  k_lr_begin<<<sm_count() * 4, 256, 0, st>>>(a);
  for (int i = 0; i <= N; ++i) {
    const bool ends = (i == 0 || i == N);
    if ((rc = launch_grad(*model, (const unsigned char*)prepared, P, W, a.theta, (float*)(ws + W.off_G),
                          (double*)(ws + W.off_V), C, ends, st)))
      return rc;
    if (i == 0) k_lr_update<PH_FIRST><<<ublocks, 256, 0, st>>>(a);
    else if (i < N) k_lr_update<PH_MID><<<ublocks, 256, 0, st>>>(a);
    else k_lr_update<PH_LAST><<<ublocks, 256, 0, st>>>(a);
  }
  HMCB_CUDA(cudaPeekAtLastError());
  return 0;
}

int hmcb_logreg_mala_step(const hmcb_model* model, const void* prepared, const hmcb_chains* chains, const hmcb_rng* rng,
                          double eps, int32_t n_steps, const hmcb_mala_io* io, void* workspace, size_t workspace_bytes,
                          void* stream) {
  int rc;
  if ((rc = check_lr_model(model))) return rc;
  HMCB_CHECK_ARG(prepared != nullptr, "prepared model buffer is NULL (call hmcb_logreg_prepare first)");
  HMCB_CHECK_ARG(chains != nullptr && rng != nullptr, "chains/rng is NULL");
  HMCB_CHECK_ARG(chains->n_chains >= 0 && chains->n_chains < (int64_t)1 << 31, "n_chains out of range");
  HMCB_CHECK_ARG(n_steps >= 1 && eps > 0, "n_steps must be >= 1 and eps > 0");
  HMCB_CHECK_ARG(!rng->p_inject || n_steps == 1, "injected normals are only defined for n_steps == 1");
  HMCB_CHECK_ARG(!rng->u_inject || rng->u_stride >= n_steps, "u_inject needs u_stride >= n_steps");
  if (chains->n_chains == 0) return 0;
  HMCB_CHECK_ARG(chains->q_in && chains->q_out, "q_in/q_out must be device pointers");
  HMCB_CHECK_ARG(chains->stride_dim == 1 && chains->stride_chain == model->d,
                 "logistic regression needs contiguous row-major [C,d] chain state");
  const int64_t C = chains->n_chains;
  const PrepLayout P = prep_layout(model->n_data);
  const WsLayout W = ws_layout(C, model->d, P.n_blocks);
  if (!workspace || workspace_bytes < W.total) {
    set_error("logreg workspace too small: have %zu bytes, need %zu", workspace ? workspace_bytes : (size_t)0, W.total);
    return HMCB_E_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  unsigned char* ws = (unsigned char*)workspace;
  MalaArgsLR a{};
  a.n_chains = C; a.chain_offset = chains->chain_offset; a.d = model->d; a.n_slabs = W.n_slabs;
  a.q = (float*)chains->q_out; a.q1 = (float*)(ws + W.off_theta); a.g0 = (float*)(ws + W.off_p);
  a.V0 = (double*)(ws + W.off_H0);
  a.Gpart = (const float*)(ws + W.off_G); a.Vpart = (const double*)(ws + W.off_V);
  a.seed = rng->seed; a.iteration = (uint32_t)rng->iteration;
  a.z_inject = (const float*)rng->p_inject; a.u_inject = rng->u_inject; a.u_stride = rng->u_stride;
  a.eps = (float)eps; a.inv_tau2 = (float)(1.0 / (model->par[0] * model->par[0]));
  if (io) { a.code = io->code; a.n_accept = io->n_accept; a.counts = io->counts; }
  const size_t qbytes = (size_t)C * model->d * 4;
  if (chains->q_out != chains->q_in) HMCB_CUDA(cudaMemcpyAsync(a.q, chains->q_in, qbytes, cudaMemcpyDeviceToDevice, st));
  const int ublocks = (int)((C + 7) / 8 < 4096 ? (C + 7) / 8 : 4096);
  const bool have_v = io && io->V_in, have_g = io && io->Vg_in;
  a.nvg0 = have_g ? 0 : 1;
  if (!have_v || !have_g) {
    // V(q), grad V(q) from one sweep at q (src/mala.py:49-52: whichever of the two the caller did not carry over)
    if ((rc = launch_grad(*model, (const unsigned char*)prepared, P, W, a.q, (float*)(ws + W.off_G), (double*)(ws + W.off_V), C,
                          true, st)))
      return rc;
    k_lr_mala_update<true><<<ublocks, 256, 0, st>>>(a);
  }
  if (have_g) HMCB_CUDA(cudaMemcpyAsync(a.g0, io->Vg_in, qbytes, cudaMemcpyDeviceToDevice, st));
  if (have_v) HMCB_CUDA(cudaMemcpyAsync(a.V0, io->V_in, (size_t)C * 8, cudaMemcpyDeviceToDevice, st));
  for (int k = 0; k < n_steps; ++k) {
    a.step = k;
    a.first = k == 0;
    k_lr_mala_propose<<<sm_count() * 4, 256, 0, st>>>(a);
    if ((rc = launch_grad(*model, (const unsigned char*)prepared, P, W, a.q1, (float*)(ws + W.off_G), (double*)(ws + W.off_V), C,
                          true, st)))
      return rc;
    k_lr_mala_update<false><<<ublocks, 256, 0, st>>>(a);
  }
  if (io && io->Vg_out) HMCB_CUDA(cudaMemcpyAsync(io->Vg_out, a.g0, qbytes, cudaMemcpyDeviceToDevice, st));
  if (io && io->V_out) HMCB_CUDA(cudaMemcpyAsync(io->V_out, a.V0, (size_t)C * 8, cudaMemcpyDeviceToDevice, st));
  HMCB_CUDA(cudaPeekAtLastError());
  return 0;
}

int hmcb_logreg_logp_grad(const hmcb_model* model, const void* prepared, const void* theta, int64_t n_chains,
                          void* logp, void* grad, void* workspace, size_t workspace_bytes, void* stream) {
  int rc;
  if ((rc = check_lr_model(model))) return rc;
  HMCB_CHECK_ARG(prepared != nullptr && theta != nullptr, "prepared/theta is NULL");
  if (n_chains <= 0) return 0;
  const PrepLayout P = prep_layout(model->n_data);
  const WsLayout W = ws_layout(n_chains, model->d, P.n_blocks);
  if (!workspace || workspace_bytes < W.total) {
    set_error("logreg workspace too small: have %zu bytes, need %zu", workspace ? workspace_bytes : (size_t)0, W.total);
    return HMCB_E_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  unsigned char* ws = (unsigned char*)workspace;
  StepArgs a{};
  a.n_chains = n_chains; a.d = model->d; a.n_slabs = W.n_slabs;
  a.theta = (float*)(ws + W.off_theta);
  a.Gpart = (const float*)(ws + W.off_G); a.Vpart = (const double*)(ws + W.off_V);
  a.inv_tau2 = (float)(1.0 / (model->par[0] * model->par[0]));
  a.logp_out = (float*)logp; a.grad_out = (float*)grad;
  HMCB_CUDA(cudaMemcpyAsync(a.theta, theta, (size_t)n_chains * model->d * 4, cudaMemcpyDeviceToDevice, st));
  if ((rc = launch_grad(*model, (const unsigned char*)prepared, P, W, a.theta, (float*)(ws + W.off_G),
                        (double*)(ws + W.off_V), n_chains, true, st)))
    return rc;
  const int ublocks = (int)((n_chains + 7) / 8 < 4096 ? (n_chains + 7) / 8 : 4096);
  k_lr_update<PH_GRADONLY><<<ublocks, 256, 0, st>>>(a);
  HMCB_CUDA(cudaPeekAtLastError());
  return 0;
}

}  // extern "C"
