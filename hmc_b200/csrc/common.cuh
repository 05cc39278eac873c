// common.cuh -- shared device/host helpers for libhmcb (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <math.h>
#include <string.h>

#include "../../include/hmcb.h"

#ifndef HMCB_EXP_VARIANT
#define HMCB_EXP_VARIANT 1   // 0 Estrin, 1 even/odd Horner (default), 2 libdevice exp; tuning knob, see tools/kbench.cu
#endif
#ifndef HMCB_FUNNEL_NACC
#define HMCB_FUNNEL_NACC 4   // partial sums for sum(theta^2)
#endif

namespace hmcb {

// ---------------------------------------------------------------- host: errors
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);

#define HMCB_CHECK_ARG(cond, ...)            \
  do {                                       \
    if (!(cond)) {                           \
      hmcb::set_error(__VA_ARGS__);          \
      return HMCB_E_ARG;                     \
    }                                        \
  } while (0)

#define HMCB_CUDA(call)                                        \
  do {                                                         \
    cudaError_t e__ = (call);                                  \
    if (e__ != cudaSuccess) return hmcb::cuda_fail(e__, #call); \
  } while (0)

int sm_count();  // SM count of the current device (cached per device)

// ---------------------------------------------------------------- lane layout
// A chain's d elements are spread over LPC adjacent lanes of a warp ("group"),
// EPL register slots per lane.  Elements are dealt in blocks of 4 (one Philox
// call yields 4 normals): block b = i/4 lives on lane b % LPC, at slots
// 4*(b/LPC) .. +3.  Element 0 (the funnel's v) is always lane 0, slot 0, and a
// Rosenbrock pair (2k, 2k+1) always shares a lane.  Slots whose element index
// is >= d are padding and hold zeros.
template <int EPL, int LPC>
struct Layout {
  static constexpr int kEPL = EPL;
  static constexpr int kLPC = LPC;
  static constexpr int kBlocks = (EPL + 3) / 4;  // element blocks per lane
  __host__ __device__ static inline int elem(int lane, int slot) {
    return (((slot >> 2) * LPC + lane) << 2) + (slot & 3);
  }
};

// slots per lane needed to hold d elements on LPC lanes (host side, for dispatch)
inline int slots_needed(int d, int lpc) {
  int mx = 0;
  for (int i = 0; i < d; ++i) {
    int b = i >> 2;
    int slot = ((b / lpc) << 2) + (i & 3);
    if (slot + 1 > mx) mx = slot + 1;
  }
  return mx;
}

// ---------------------------------------------------------------- group reductions
template <int LPC, typename T>
__device__ __forceinline__ T group_sum(T x) {
#pragma unroll
  for (int o = LPC >> 1; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
  return x;
}

// same, naming only the group's own lanes: safe when groups of one warp diverge
template <int LPC, typename T>
__device__ __forceinline__ T group_sum_masked(T x, unsigned mask) {
#pragma unroll
  for (int o = LPC >> 1; o > 0; o >>= 1) x += __shfl_xor_sync(mask, x, o);
  return x;
}

// ---------------------------------------------------------------- Philox4x32-10
// Counter-based generator (Salmon et al. 2011).  key = seed; counter =
// (chain_lo, chain_hi, iteration, stream<<24 | block).  Stateless: a draw is a
// pure function of (seed, global chain id, iteration, stream, block), so chains
// give the same results however they are sharded over GPUs or lanes.
struct Philox {
  static constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
  __host__ __device__ static inline void round(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
#ifdef __CUDA_ARCH__
    uint32_t hi0 = __umulhi(M0, c[0]), hi1 = __umulhi(M1, c[2]);
#else
    uint32_t hi0 = (uint32_t)(((uint64_t)M0 * c[0]) >> 32), hi1 = (uint32_t)(((uint64_t)M1 * c[2]) >> 32);
#endif
    uint32_t lo0 = M0 * c[0], lo1 = M1 * c[2];
    uint32_t n0 = hi1 ^ c[1] ^ k0, n2 = hi0 ^ c[3] ^ k1;
    c[0] = n0; c[1] = lo1; c[2] = n2; c[3] = lo0;
  }
  __host__ __device__ static inline void gen(uint64_t seed, uint64_t chain, uint32_t iter, uint32_t stream,
                                             uint32_t block, uint32_t (&out)[4]) {
    uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
    out[0] = (uint32_t)chain; out[1] = (uint32_t)(chain >> 32); out[2] = iter; out[3] = (stream << 24) | (block & 0xFFFFFFu);
#pragma unroll
    for (int r = 0; r < 10; ++r) {
      round(out, k0, k1);
      k0 += W0; k1 += W1;
    }
  }
};

enum : uint32_t { STREAM_MOMENTUM = 0, STREAM_ACCEPT = 1, STREAM_STEPSIZE = 2 };

// uniform in [0,1) with 53 bits from two words
__host__ __device__ inline double u53(uint32_t a, uint32_t b) {
  return (double)(((uint64_t)(a >> 5) << 26) | (uint64_t)(b >> 6)) * (1.0 / 9007199254740992.0);
}

// Four standard normals from one Philox block: Box-Muller on two (u1,u2) pairs in f32
// with the hardware approximations (MUFU lg2 / sqrt / sin / cos; about 13 instructions
// per pair against ~70 for the libdevice logf/sincospif bodies -- the momentum draw was
// a third of the L=20 kernel).  Absolute error of a variate ~1e-6, 32-bit radius and
// angle resolution, |z| <= 6.7: far below Monte Carlo resolution for a momentum refresh.
// HMCB_RNG_NORMAL_F64 selects the exact f64 generator below instead.
__device__ __forceinline__ void normal4_f32(const uint32_t (&w)[4], float (&z)[4]) {
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const float u1 = ((float)w[2 * h] + 0.5f) * 2.3283064365386963e-10f;   // (0,1]
    const float r = sqrtf(-1.3862943611198906f * __log2f(u1));              // sqrt(-2 ln u1)
    const float th = (float)w[2 * h + 1] * 1.4629180792671596e-09f;         // 2 pi u2, in [0, 2 pi]
    z[2 * h] = r * __cosf(th);
    z[2 * h + 1] = r * __sinf(th);
  }
}

// Two standard normals in f64 from one Philox block (53-bit uniforms).
__device__ __forceinline__ void normal2_f64(const uint32_t (&w)[4], double (&z)[2]) {
  double u1 = 1.0 - u53(w[0], w[1]);  // (0,1]
  double r = sqrt(-2.0 * log(u1));
  double s, c;
  sincospi(2.0 * u53(w[2], w[3]), &s, &c);
  z[0] = r * c;
  z[1] = r * s;
}

// ---------------------------------------------------------------- small math
template <typename T> __device__ __forceinline__ T t_exp(T x);
template <> __device__ __forceinline__ double t_exp<double>(double x) { return exp(x); }
template <> __device__ __forceinline__ float t_exp<float>(float x) { return expf(x); }

// Branch-free exp for the leapfrog inner loop.  libdevice's exp() carries a slow-path
// branch that splits the loop body into basic blocks, so ptxas cannot interleave its
// ~17-deep dependent FMA chain with the 100 independent drift/sum FMAs of the same
// step (measured: the chain alone was ~40% of the loop's issue time).  This version is
// straight-line: magic-number range reduction, degree-13 Taylor polynomial evaluated as
// two interleaved Horner chains in r^2 (depth 8 instead of 13, few live registers --
// measured 0.75 of DFMA peak vs 0.71 for an Estrin tree and 0.64 for libdevice on the
// d=50 kernel, which sits at the 255-register limit), and a two-factor 2^n scaling that
// overflows to +inf / underflows through denormals to 0 without a branch.  NaN in,
// NaN out.  Relative error < 2 ulp on [-745, 709].
__host__ __device__ __forceinline__ double exp_bf(double x) {
  const double xc0 = x > 800.0 ? 800.0 : x;          // NaN compares false and passes through
  const double xc = xc0 < -800.0 ? -800.0 : xc0;
  const double magic = 6755399441055744.0;            // 1.5 * 2^52: fma rounds x*log2(e) to an integer
  const double t = fma(xc, 1.4426950408889634, magic);
  const double nf = t - magic;
  double r = fma(nf, -6.93147180559945286e-01, xc);
  r = fma(nf, -2.31904681384629956e-17, r);
  const double r2 = r * r;
#if HMCB_EXP_VARIANT == 1
  // even/odd split: two Horner chains in r^2 (fewer live registers than Estrin, depth 8)
  double ev = fma(r2, 1.0 / 479001600.0, 1.0 / 3628800.0);
  double od = fma(r2, 1.0 / 6227020800.0, 1.0 / 39916800.0);
  ev = fma(ev, r2, 1.0 / 40320.0);
  od = fma(od, r2, 1.0 / 362880.0);
  ev = fma(ev, r2, 1.0 / 720.0);
  od = fma(od, r2, 1.0 / 5040.0);
  ev = fma(ev, r2, 1.0 / 24.0);
  od = fma(od, r2, 1.0 / 120.0);
  ev = fma(ev, r2, 0.5);
  od = fma(od, r2, 1.0 / 6.0);
  ev = fma(ev, r2, 1.0);
  od = fma(od, r2, 1.0);
  const double p = fma(od, r, ev);
#else
  const double b0 = fma(r, 1.0, 1.0);
  const double b1 = fma(r, 1.0 / 6.0, 0.5);
  const double b2 = fma(r, 1.0 / 120.0, 1.0 / 24.0);
  const double b3 = fma(r, 1.0 / 5040.0, 1.0 / 720.0);
  const double b4 = fma(r, 1.0 / 362880.0, 1.0 / 40320.0);
  const double b5 = fma(r, 1.0 / 39916800.0, 1.0 / 3628800.0);
  const double b6 = fma(r, 1.0 / 6227020800.0, 1.0 / 479001600.0);
  const double r4 = r2 * r2;
  const double d0 = fma(b1, r2, b0);
  const double d1 = fma(b3, r2, b2);
  const double d2 = fma(b5, r2, b4);
  const double r8 = r4 * r4;
  const double e0 = fma(d1, r4, d0);
  const double e1 = fma(b6, r4, d2);
  const double p = fma(e1, r8, e0);
#endif
  // n sits in the low word of t; split 2^n = 2^n1 * 2^n2 so each factor is a normal double
#ifdef __CUDA_ARCH__
  const int n = __double2loint(t);
  const int n1 = n >> 1, n2 = n - n1;
  const double s1 = __hiloint2double((n1 + 1023) << 20, 0);
  const double s2 = __hiloint2double((n2 + 1023) << 20, 0);
#else
  long long tb;
  memcpy(&tb, &t, 8);
  const int n = (int)(tb & 0xffffffffLL);
  const int n1 = n >> 1, n2 = n - n1;
  const long long b1i = (long long)(n1 + 1023) << 52, b2i = (long long)(n2 + 1023) << 52;
  double s1, s2;
  memcpy(&s1, &b1i, 8);
  memcpy(&s2, &b2i, 8);
#endif
  return (p * s1) * s2;
}

template <typename T> __device__ __forceinline__ T t_exp_loop(T x);
#if HMCB_EXP_VARIANT == 2
template <> __device__ __forceinline__ double t_exp_loop<double>(double x) { return exp(x); }
#else
template <> __device__ __forceinline__ double t_exp_loop<double>(double x) { return exp_bf(x); }
#endif
template <> __device__ __forceinline__ float t_exp_loop<float>(float x) { return expf(x); }

template <typename T> __device__ __forceinline__ T t_rsqrt(T x);
template <> __device__ __forceinline__ double t_rsqrt<double>(double x) { return 1.0 / sqrt(x); }
template <> __device__ __forceinline__ float t_rsqrt<float>(float x) { return 1.0f / sqrtf(x); }

template <typename T> __device__ __forceinline__ bool t_bad(T x) { return isnan(x) || isinf(x); }

// a product the compiler may not fuse into a following sum
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }

}  // namespace hmcb
