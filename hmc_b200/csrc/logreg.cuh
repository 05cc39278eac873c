// logreg.cuh -- Bayesian logistic-regression gradient on the 5th-generation tensor cores
// (tcgen05.mma, accumulators in TMEM).  This is the one model where chains x data is a
// real dense GEMM:   Z = Theta X^T  ->  R = sigmoid(Z) - y  ->  G = R X   (+ prior term),
// shaped like attention with K = V = X and a sigmoid in place of the softmax:
//
//   per CTA: one tile of 128 chains (TMEM lanes), one slab of data rows, walked in blocks
//   of 128 rows.  Per block:
//     GEMM1  Z^T[128 chains x 128 rows] = Theta[128 x D] * X_blk^T      (A, B from smem)
//     epilogue: Z^T TMEM -> registers (thread = chain) -> r = sigmoid(z) - y -> fp16 hi/lo
//               -> back to TMEM as the A operand of GEMM2; log-likelihood summed per thread
//     GEMM2  G[128 chains x D] += R^T[128 x 128 rows] * X_blk           (A from TMEM, B from smem)
//   Z and R never touch shared or global memory.
//
// Precision: operands are split x = hi + lo with hi, lo in fp16 after a power-of-two
// pre-scaling (per matrix for X, per chain for Theta, fixed for R), and each product is
// three MMAs, hi*hi + hi*lo + lo*hi, into one fp32 accumulator: ~2^-21 relative operand
// error at the fp16 tensor rate, half the shared-memory footprint of a 3xTF32 scheme.
//
// Shared-memory operand layout: the no-swizzle canonical "core matrix" tiling of the
// UMMA descriptors -- 8 rows x 16 bytes (8 fp16), 128 B per core matrix.  An X block is
// stored once, as cm(row_blk, dim_blk) at (dim_blk*16 + row_blk)*128 B, and is read by
// GEMM1 as a K-major B operand (MN = rows, K = dims: SBO = 128, LBO = 2048) and by GEMM2
// as an MN-major B operand (MN = dims, K = rows: SBO = 2048, LBO = 128): the 8x8 core
// matrix has the same bytes under both readings.
#pragma once
#include <cuda_fp16.h>

#include "common.cuh"
#include "sync_ptx.cuh"

namespace hmcb {
namespace lr {

constexpr int kD = 128;        // padded feature dimension (K of GEMM1, N of GEMM2)
constexpr int kTileC = 128;    // chains per CTA (M of both GEMMs = TMEM lanes)
constexpr int kTileR = 128;    // data rows per block (N of GEMM1, K of GEMM2)
constexpr int kRScaleLog2 = 10;  // residual r in (-1,1) is stored as r * 2^10

// ------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// D[tmem] (+)= A[smem] * B[smem]
__device__ __forceinline__ void mma_ss(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem]
__device__ __forceinline__ void mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d_tmem),
      "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// 32 lanes x 32 consecutive 32-bit columns: thread t of the warp gets lane (base lane + t)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};" ::
          "r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]),
      "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
      "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// ------------------------------------------------------------------ descriptors
// shared-memory matrix descriptor, no swizzle (layout_type 0), version 1 (Blackwell)
__device__ __forceinline__ uint64_t smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  return d;
}
// instruction descriptor: fp16 x fp16 -> fp32, M x N, majors (0 = K-major, 1 = MN-major)
__host__ __device__ constexpr uint32_t instr_desc(int M, int N, int a_major, int b_major) {
  return (1u << 4)              // c_format = F32
         | (0u << 7)            // a_format = F16
         | (0u << 10)           // b_format = F16
         | ((uint32_t)a_major << 15) | ((uint32_t)b_major << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// byte offset of element (row, col) of a [rows x 128] fp16 operand in the core-matrix tiling
__host__ __device__ __forceinline__ uint32_t cm_offset(int row, int col) {
  return (uint32_t)(((col >> 3) * 16 + (row >> 3)) * 128 + (row & 7) * 16 + (col & 7) * 2);
}

// ------------------------------------------------------------------ pipelined kernel
// Warp-specialised: warp 0 streams X blocks (pre-tiled in global
// memory, one 32 KB bulk copy per matrix per block) into a 2-stage shared-memory ring,
// warp 1 issues the MMAs, 16 epilogue warps (4 per TMEM lane quarter, one 32-row chunk
// each) turn Z into R in place in TMEM.  Z/R is double-buffered in TMEM so that the
// sigmoid of block b overlaps GEMM1 of block b+1.  The tensor core adds into the fp32
// accumulator with truncation (measured: error grows linearly with the number of rows
// accumulated), so G is drained into fp32 registers (round-to-nearest adds) every
// kFlush blocks.
constexpr int kEpiWarps = 16;
constexpr int kThreadsV1 = 32 * (2 + kEpiWarps);
constexpr int kFlush = 8;
constexpr uint32_t kBlkBytes = kTileR * kD * 2;   // one fp16 operand block
// dynamic shared memory of k_logreg_grad_v1: Theta hi/lo, 2 X stages x hi/lo, y ring, V sums, scales, barriers
constexpr size_t kSmemV1 = 6 * (size_t)kBlkBytes + 2 * kTileR * 4 + kTileC * 8 + kTileC * 4 + 16 * 8 + 64;

__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ float fast_sigmoid(float z, float& ez) {
  // ez = exp(-|z|) in (0,1];  sigmoid(|z|) = 1/(1+ez)
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(ez) : "f"(-1.4426950408889634f * fabsf(z)));
  float s;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(s) : "f"(1.0f + ez));
  return z >= 0.f ? s : 1.0f - s;
}

struct GradArgsV1 {
  const __half* Xh_blk;  // [n_blocks][kTileR*kD] fp16, each block already in the core-matrix tiling
  const __half* Xl_blk;
  const float* y;        // [n_blocks*kTileR]
  const float* theta;    // [n_chains][d] fp32
  float* Gpart;          // [n_slabs][n_chains][kD]
  double* Vpart;         // [n_slabs][n_chains] or nullptr (sum of softplus(z) - y z over the slab's rows)
  int64_t n_chains;
  int32_t d, n_rows;
  int32_t blocks_per_slab, n_blocks, n_slabs;
  int32_t x_scale_log2;  // used when x_scale_dev == nullptr
  const int32_t* x_scale_dev;  // device copy of the scale exponent (written by the prepare step)
};

template <bool WANT_V, int NPROD>
__global__ void __launch_bounds__(kThreadsV1, 1) k_logreg_grad_v1(const GradArgsV1 a) {
  extern __shared__ __align__(1024) unsigned char smem[];
  unsigned char* sTh = smem;                         // 32 KB
  unsigned char* sTl = sTh + kBlkBytes;              // 32 KB
  unsigned char* sX = sTl + kBlkBytes;               // 2 stages x (hi 32 KB + lo 32 KB)
  float* sY = reinterpret_cast<float*>(sX + 4 * kBlkBytes);     // 2 x 128 floats
  double* sV = reinterpret_cast<double*>(sY + 2 * kTileR);        // 128 doubles
  int* sTe = reinterpret_cast<int*>(sV + kTileC);                 // 128 ints: per-chain scale exponent
  uint64_t* bars = reinterpret_cast<uint64_t*>(sTe + kTileC);
  uint64_t *x_full = bars, *x_empty = bars + 2, *z_full = bars + 4, *p_full = bars + 6, *zp_empty = bars + 8,
           *g_full = bars + 10, *g_empty = bars + 11;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 12);

  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  const int tile = blockIdx.x, slab = blockIdx.y;
  const int b0 = slab * a.blocks_per_slab;
  const int b1 = min(b0 + a.blocks_per_slab, a.n_blocks);
  const int nb = b1 - b0;

  if (warp == 0) tmem_alloc(tmem_slot, 512);
  if (t == 32) {
    for (int s = 0; s < 2; ++s) {
      mbar_init(x_full + s, 1);
      mbar_init(x_empty + s, 1);
      mbar_init(z_full + s, 1);
      mbar_init(p_full + s, kEpiWarps);
      mbar_init(zp_empty + s, 1);
    }
    mbar_init(g_full, 1);
    mbar_init(g_empty, kEpiWarps);
    fence_barrier_init();
  }
  if (t < kTileC) sV[t] = 0.0;
  const int xs_log2 = a.x_scale_dev ? *a.x_scale_dev : a.x_scale_log2;
  // Theta tile -> fp16 hi/lo in the core-matrix tiling; one warp per chain row, 8 rows per warp
  if (warp >= 2) {
    const int ew = warp - 2;
    for (int r = ew; r < kTileC; r += kEpiWarps) {
      const int64_t chain = (int64_t)tile * kTileC + r;
      const bool live = chain < a.n_chains;
      float v[4];
      float mx = 0.f;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int col = lane * 4 + k;
        v[k] = (live && col < a.d) ? a.theta[chain * a.d + col] : 0.f;
        mx = fmaxf(mx, fabsf(v[k]));
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      int te = 0;
      if (mx > 0.f) {
        int ex;
        frexpf(mx, &ex);
        te = 10 - ex;
      }
      if (lane == 0) sTe[r] = te;
      const float sc = ldexpf(1.0f, te);
      __half h[4], l[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float x = v[k] * sc;
        h[k] = __float2half_rn(x);
        l[k] = __float2half_rn(x - __half2float(h[k]));
      }
      const uint32_t off = cm_offset(r, lane * 4);   // 4 consecutive cols stay inside one 16-byte row of a core matrix
      *reinterpret_cast<uint2*>(sTh + off) = make_uint2((uint32_t)__half_as_ushort(h[0]) | ((uint32_t)__half_as_ushort(h[1]) << 16),
                                                        (uint32_t)__half_as_ushort(h[2]) | ((uint32_t)__half_as_ushort(h[3]) << 16));
      *reinterpret_cast<uint2*>(sTl + off) = make_uint2((uint32_t)__half_as_ushort(l[0]) | ((uint32_t)__half_as_ushort(l[1]) << 16),
                                                        (uint32_t)__half_as_ushort(l[2]) | ((uint32_t)__half_as_ushort(l[3]) << 16));
    }
    fence_proxy_async();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tG = tmem + 256;
  constexpr uint32_t idesc1 = instr_desc(kTileC, kTileR, 0, 0);
  constexpr uint32_t idesc2 = instr_desc(kTileC, kD, 0, 1);

  if (warp == 0) {
    // ---------------- producer: bulk copies of pre-tiled X blocks
    if (lane == 0) {
      for (int i = 0; i < nb; ++i) {
        const int s = i & 1, u = i >> 1;
        mbar_wait(x_empty + s, (u & 1) ^ 1);
        mbar_expect_tx(x_full + s, 2 * kBlkBytes + kTileR * 4);
        const size_t blk = (size_t)(b0 + i);
        bulk_g2s(sX + (2 * s) * kBlkBytes, a.Xh_blk + blk * (kTileR * kD), kBlkBytes, x_full + s);
        bulk_g2s(sX + (2 * s + 1) * kBlkBytes, a.Xl_blk + blk * (kTileR * kD), kBlkBytes, x_full + s);
        bulk_g2s(sY + s * kTileR, a.y + blk * kTileR, kTileR * 4, x_full + s);
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer
    if (lane == 0) {
      const uint32_t aTh = smem_u32(sTh), aTl = smem_u32(sTl);
      auto gemm2 = [&](int i) {   // G (+)= R(i) X(i)
        const int s = i & 1, u = i >> 1;
        mbar_wait(p_full + s, u & 1);
        tc_fence_after();
        const bool fresh = (i % kFlush) == 0;        // first block of a flush group starts a new sum
        if (fresh && i > 0) mbar_wait(g_empty, ((i / kFlush) - 1) & 1);
        tc_fence_after();
        const uint32_t aXh = smem_u32(sX + (2 * s) * kBlkBytes), aXl = smem_u32(sX + (2 * s + 1) * kBlkBytes);
        const uint32_t tP = tmem + s * 128;
#pragma unroll
        for (int kk = 0; kk < kTileR / 16; ++kk) {
          const uint32_t ko = (uint32_t)kk * 256;
          const uint32_t pa = tP + 32 * (kk >> 1) + 8 * (kk & 1);
          mma_ts(tG, pa, smem_desc(aXh + ko, 128, 2048), idesc2, (fresh && kk == 0) ? 0 : 1);
          if (NPROD == 3) {
            mma_ts(tG, pa, smem_desc(aXl + ko, 128, 2048), idesc2, 1);
            mma_ts(tG, pa + 16, smem_desc(aXh + ko, 128, 2048), idesc2, 1);
          }
        }
        tc_commit(x_empty + s);
        tc_commit(zp_empty + s);
        if ((i % kFlush) == kFlush - 1 || i == nb - 1) tc_commit(g_full);
      };
      for (int i = 0; i < nb; ++i) {
        const int s = i & 1, u = i >> 1;
        mbar_wait(x_full + s, u & 1);
        mbar_wait(zp_empty + s, (u & 1) ^ 1);
        tc_fence_after();
        const uint32_t aXh = smem_u32(sX + (2 * s) * kBlkBytes), aXl = smem_u32(sX + (2 * s + 1) * kBlkBytes);
        const uint32_t tZ = tmem + s * 128;
#pragma unroll
        for (int kk = 0; kk < kD / 16; ++kk) {
          const uint32_t ko = (uint32_t)kk * 4096;
          mma_ss(tZ, smem_desc(aTh + ko, 2048, 128), smem_desc(aXh + ko, 2048, 128), idesc1, kk > 0);
          if (NPROD == 3) {
            mma_ss(tZ, smem_desc(aTh + ko, 2048, 128), smem_desc(aXl + ko, 2048, 128), idesc1, 1);
            mma_ss(tZ, smem_desc(aTl + ko, 2048, 128), smem_desc(aXh + ko, 2048, 128), idesc1, 1);
          }
        }
        tc_commit(z_full + s);
        if (i > 0) gemm2(i - 1);
      }
      if (nb > 0) gemm2(nb - 1);
    }
  } else {
    // ---------------- epilogue warps: lane quarter q, 32-row chunk cq
    const int ew = warp - 2;
    const int q = warp & 3, cq = ew >> 2;
    const int r = q * 32 + lane;                       // chain row in the tile = TMEM lane
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    const float zscale = ldexpf(1.0f, -(xs_log2 + sTe[r]));
    const float rscale = ldexpf(1.0f, kRScaleLog2);
    float gacc[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) gacc[j] = 0.f;
    double vsum = 0.0;
    for (int i = 0; i < nb; ++i) {
      const int s = i & 1, u = i >> 1;
      mbar_wait(z_full + s, u & 1);
      tc_fence_after();
      const uint32_t tZ = tmem + s * 128 + lane_base + cq * 32;
      uint32_t zr[32];
      tmem_ld32(tZ, zr);
      tmem_ld_wait();
      uint32_t ph[16], pl[16];
      const int row_base = (b0 + i) * kTileR + cq * 32;
      const float* yy = sY + s * kTileR + cq * 32;
#pragma unroll
      for (int j = 0; j < 32; j += 2) {
        float rr[2];
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const float z = __uint_as_float(zr[j + e]) * zscale;
          const float yv = yy[j + e];
          float ez;
          const float sig = fast_sigmoid(z, ez);
          const bool ok = row_base + j + e < a.n_rows;
          rr[e] = ok ? (sig - yv) * rscale : 0.f;
          if (WANT_V && ok) {
            float l2;
            asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l2) : "f"(1.0f + ez));
            vsum += (double)(fmaf(0.6931471805599453f, l2, fmaxf(z, 0.f)) - yv * z);   // softplus(z) - y z
          }
        }
        const __half2 h2 = __floats2half2_rn(rr[0], rr[1]);
        const float2 hf = __half22float2(h2);
        const __half2 l2 = __floats2half2_rn(rr[0] - hf.x, rr[1] - hf.y);
        ph[j >> 1] = *reinterpret_cast<const uint32_t*>(&h2);
        pl[j >> 1] = *reinterpret_cast<const uint32_t*>(&l2);
      }
      // R overwrites this thread's own 32 Z columns: hi in the first 16, lo in the last 16
      asm volatile(
          "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(tZ),
          "r"(ph[0]), "r"(ph[1]), "r"(ph[2]), "r"(ph[3]), "r"(ph[4]), "r"(ph[5]), "r"(ph[6]), "r"(ph[7]), "r"(ph[8]),
          "r"(ph[9]), "r"(ph[10]), "r"(ph[11]), "r"(ph[12]), "r"(ph[13]), "r"(ph[14]), "r"(ph[15])
          : "memory");
      asm volatile(
          "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(tZ + 16),
          "r"(pl[0]), "r"(pl[1]), "r"(pl[2]), "r"(pl[3]), "r"(pl[4]), "r"(pl[5]), "r"(pl[6]), "r"(pl[7]), "r"(pl[8]),
          "r"(pl[9]), "r"(pl[10]), "r"(pl[11]), "r"(pl[12]), "r"(pl[13]), "r"(pl[14]), "r"(pl[15])
          : "memory");
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_full + s);
      if ((i % kFlush) == kFlush - 1 || i == nb - 1) {
        // drain this flush group's partial G (columns = dims cq*32..+31) into registers
        mbar_wait(g_full, (i / kFlush) & 1);
        tc_fence_after();
        uint32_t gr[32];
        tmem_ld32(tG + lane_base + cq * 32, gr);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) gacc[j] += __uint_as_float(gr[j]);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(g_empty);
      }
    }
    const int64_t chain = (int64_t)tile * kTileC + r;
    if (chain < a.n_chains) {
      const float gscale = ldexpf(1.0f, -(kRScaleLog2 + xs_log2));
      float* gout = a.Gpart + ((size_t)slab * a.n_chains + chain) * kD + cq * 32;
#pragma unroll
      for (int j = 0; j < 32; j += 4)
        *reinterpret_cast<float4*>(gout + j) =
            make_float4(gacc[j] * gscale, gacc[j + 1] * gscale, gacc[j + 2] * gscale, gacc[j + 3] * gscale);
    }
    if (WANT_V) atomicAdd(sV + r, vsum);
  }
  tc_fence_before();
  __syncthreads();
  if (WANT_V && a.Vpart && t < kTileC) {
    const int64_t chain = (int64_t)tile * kTileC + t;
    if (chain < a.n_chains) a.Vpart[(size_t)slab * a.n_chains + chain] = sV[t];
  }
  if (warp == 0) tmem_dealloc(tmem, 512);
}

// X (fp32 [n_rows][d]) -> pre-tiled fp16 hi/lo blocks for the pipelined kernel
__global__ void k_logreg_prep_blk(const float* __restrict__ X, int n_rows, int d, int n_blocks, int x_scale_log2,
                                  const int32_t* __restrict__ x_scale_dev, __half* __restrict__ Xh,
                                  __half* __restrict__ Xl) {
  const float s = ldexpf(1.0f, x_scale_dev ? *x_scale_dev : x_scale_log2);
  const size_t n = (size_t)n_blocks * kTileR * kD;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const size_t blk = i / (kTileR * kD);
    const int w = (int)(i % (kTileR * kD));
    const int r = w / kD, k = w % kD;
    const size_t row = blk * kTileR + r;
    const float v = (row < (size_t)n_rows && k < d) ? X[row * d + k] * s : 0.f;
    const __half h = __float2half_rn(v);
    const size_t o = blk * (kTileR * kD) + cm_offset(r, k) / 2;
    Xh[o] = h;
    Xl[o] = __float2half_rn(v - __half2float(h));
  }
}

}  // namespace lr
}  // namespace hmcb
