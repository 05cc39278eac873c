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
// Warp-specialised: warp 0 streams X blocks (pre-tiled in global memory, one 32 KB bulk copy per
// matrix per block) into a 3-stage shared-memory ring, warp 1 issues the MMAs, 16 epilogue warps
// (4 per TMEM lane quarter, one 32-row chunk each) turn Z into R in place in TMEM.  Theta (the A
// operand of GEMM1) is converted once per CTA and lives in TMEM, like R (the A operand of GEMM2);
// that leaves shared memory for the third X stage.  Z/R is double-buffered in TMEM so that the
// sigmoid of block b overlaps GEMM2 of block b-1 and GEMM1 of block b+1.  tcgen05.mma
// instructions execute in issue order, so a Z buffer needs no barrier between its last reader
// (GEMM2 of block b) and its next writer (GEMM1 of block b+2).
// The tensor core adds into the fp32 accumulator with truncation (measured: error grows linearly
// with the number of rows accumulated), so G is drained into fp32 registers (round-to-nearest
// adds) every kFlush blocks; the drain is done one block late, under the next block's sigmoid.
// Where the time goes was read off clock64 stamps of one CTA (LR_TRACE, tools/logreg_test.cu):
// with two X stages the MMA thread waited a third of every block for the next X block, whose
// load could only start when GEMM2 of the block before had finished (4950 cycles per block);
// with three it is bound by the rate at which one thread can issue 48 MMAs (3720 cycles per
// block, 86 % of the tensor pipe's 3180).
constexpr int kEpiWarps = 16;
constexpr int kThreadsV1 = 32 * (2 + kEpiWarps);
constexpr int kFlush = 8;
constexpr uint32_t kBlkBytes = kTileR * kD * 2;   // one fp16 operand block
constexpr int kXStages = 3;
// dynamic shared memory of k_logreg_grad_v1: kXStages X stages x hi/lo, y ring, V sums, scales, barriers.
// Theta lives in TMEM (the A operand of GEMM1 is read from there), which is what leaves room for the third X
// stage: with two, the load of block i+1 could only start when GEMM2 of block i-1 had finished, and the MMA
// thread sat waiting for it for a third of every block (measured with the clock64 trace of tools/logreg_test.cu).
constexpr size_t kSmemV1 = 2 * kXStages * (size_t)kBlkBytes + kXStages * kTileR * 4 + kTileC * 8 + kTileC * 4 + 24 * 8 + 64;

__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
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
  unsigned long long* trace;   // development only (tools/logreg_test.cu): clock64 stamps of one CTA, 16 per block
  int32_t trace_x, trace_y;    // which CTA
};
#define LR_TRACE(slot, blk)                                                                  \
  do {                                                                                       \
    if (a.trace && blockIdx.x == a.trace_x && blockIdx.y == a.trace_y) a.trace[(size_t)(blk) * 16 + (slot)] = clock64(); \
  } while (0)

template <bool WANT_V, int NPROD>
__global__ void __launch_bounds__(kThreadsV1, 1) k_logreg_grad_v1(const GradArgsV1 a) {
  extern __shared__ __align__(1024) unsigned char smem[];
  unsigned char* sX = smem;                          // kXStages x (hi 32 KB + lo 32 KB)
  float* sY = reinterpret_cast<float*>(sX + 2 * kXStages * kBlkBytes);   // kXStages x 128 floats
  double* sV = reinterpret_cast<double*>(sY + kXStages * kTileR);          // 128 doubles
  int* sTe = reinterpret_cast<int*>(sV + kTileC);                 // 128 ints: per-chain scale exponent
  uint64_t* bars = reinterpret_cast<uint64_t*>(sTe + kTileC);
  uint64_t *x_full = bars, *x_empty = bars + 4, *z_full = bars + 8, *p_full = bars + 10,
           *g_full = bars + 14, *g_empty = bars + 15;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);

  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  const int tile = blockIdx.x, slab = blockIdx.y;
  const int b0 = slab * a.blocks_per_slab;
  const int b1 = min(b0 + a.blocks_per_slab, a.n_blocks);
  const int nb = b1 - b0;

  if (t == 0) LR_TRACE(9, 0);
  if (warp == 0) tmem_alloc(tmem_slot, 512);
  if (t == 32) {
    for (int s = 0; s < kXStages; ++s) {
      mbar_init(x_full + s, 1);
      mbar_init(x_empty + s, 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(z_full + s, 1);
      mbar_init(p_full + s, kEpiWarps);
    }
    mbar_init(g_full, 1);
    mbar_init(g_empty, kEpiWarps);
    fence_barrier_init();
  }
  if (t < kTileC) sV[t] = 0.0;
  const int xs_log2 = a.x_scale_dev ? *a.x_scale_dev : a.x_scale_log2;
  // TMEM columns: Z/R double buffer 0..255, G 256..383, Theta hi 384..447, Theta lo 448..511 (fp16 pairs)
  tc_fence_before();
  __syncthreads();                                   // tmem_slot and the barriers are visible
  tc_fence_after();
  {
    const uint32_t tm0 = *tmem_slot;
    // the chain tile of Theta, read coalesced into a padded staging area that borrows X stages 1 and 2 (the
    // producer starts after the next block-wide barrier); row stride 129 floats keeps the row-per-thread reads
    // below free of bank conflicts
    constexpr int kTS = kD + 1;
    float* sT = reinterpret_cast<float*>(sX + 2 * kBlkBytes);
    static_assert((size_t)kTileC * kTS * 4 <= 4 * (size_t)kBlkBytes, "Theta staging must fit in two X stages");
    for (int idx = t; idx < kTileC * kD; idx += kThreadsV1) {
      const int rr = idx / kD, col = idx - rr * kD;
      const int64_t ch = (int64_t)tile * kTileC + rr;
      sT[rr * kTS + col] = (ch < a.n_chains && col < a.d) ? a.theta[ch * a.d + col] : 0.f;
    }
    __syncthreads();
    if (warp >= 2) {
      // Theta tile -> fp16 hi/lo, scaled per chain, as the TMEM-resident A operand of GEMM1: thread = chain row
      // (TMEM lane), and the four warps of a lane quarter each convert 32 of the 128 dims
      const int ew = warp - 2;
      const int q = warp & 3, cq = ew >> 2;
      const int r = q * 32 + lane;
      const float* th = sT + r * kTS;
      float mx = 0.f;
#pragma unroll 8
      for (int col = 0; col < kD; ++col) mx = fmaxf(mx, fabsf(th[col]));
      int te = 0;
      if (mx > 0.f) {
        int ex;
        frexpf(mx, &ex);
        te = 10 - ex;
      }
      if (cq == 0) sTe[r] = te;
      const float sc = ldexpf(1.0f, te);
      uint32_t ph[16], pl[16];
#pragma unroll
      for (int k = 0; k < 32; k += 2) {
        const int col = cq * 32 + k;
        const float x0 = th[col] * sc;
        const float x1 = th[col + 1] * sc;
        const __half2 h2 = __floats2half2_rn(x0, x1);
        const float2 hf = __half22float2(h2);
        const __half2 l2 = __floats2half2_rn(x0 - hf.x, x1 - hf.y);
        ph[k >> 1] = *reinterpret_cast<const uint32_t*>(&h2);
        pl[k >> 1] = *reinterpret_cast<const uint32_t*>(&l2);
      }
      const uint32_t lane_base = (uint32_t)(q * 32) << 16;
      asm volatile(
          "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(
              tm0 + 384 + lane_base + cq * 16),
          "r"(ph[0]), "r"(ph[1]), "r"(ph[2]), "r"(ph[3]), "r"(ph[4]), "r"(ph[5]), "r"(ph[6]), "r"(ph[7]), "r"(ph[8]),
          "r"(ph[9]), "r"(ph[10]), "r"(ph[11]), "r"(ph[12]), "r"(ph[13]), "r"(ph[14]), "r"(ph[15])
          : "memory");
      asm volatile(
          "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(
              tm0 + 448 + lane_base + cq * 16),
          "r"(pl[0]), "r"(pl[1]), "r"(pl[2]), "r"(pl[3]), "r"(pl[4]), "r"(pl[5]), "r"(pl[6]), "r"(pl[7]), "r"(pl[8]),
          "r"(pl[9]), "r"(pl[10]), "r"(pl[11]), "r"(pl[12]), "r"(pl[13]), "r"(pl[14]), "r"(pl[15])
          : "memory");
      tmem_st_wait();
    }
    fence_proxy_async();       // the staging reads above precede the producer's bulk copies into the same bytes
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  if (t == 0) LR_TRACE(10, 0);
  const uint32_t tG = tmem + 256;
  constexpr uint32_t idesc1 = instr_desc(kTileC, kTileR, 0, 0);
  constexpr uint32_t idesc2 = instr_desc(kTileC, kD, 0, 1);

  if (warp == 0) {
    // ---------------- producer: bulk copies of pre-tiled X blocks
    if (lane == 0) {
      for (int i = 0; i < nb; ++i) {
        const int s = i % kXStages, u = i / kXStages;
        mbar_wait(x_empty + s, (u & 1) ^ 1);
        LR_TRACE(0, i);
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
      const uint32_t tTh = tmem + 384, tTl = tmem + 448;
      auto gemm2 = [&](int i) {   // G (+)= R(i) X(i)
        const int s = i & 1, u = i >> 1;      // Z/R buffer
        const int xs = i % kXStages;          // X stage
        mbar_wait(p_full + s, u & 1);
        LR_TRACE(3, i);
        tc_fence_after();
        const bool fresh = (i % kFlush) == 0;        // first block of a flush group starts a new sum
        if (fresh && i > 0) mbar_wait(g_empty, ((i / kFlush) - 1) & 1);
        tc_fence_after();
        const uint32_t aXh = smem_u32(sX + (2 * xs) * kBlkBytes), aXl = smem_u32(sX + (2 * xs + 1) * kBlkBytes);
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
        tc_commit(x_empty + xs);
        LR_TRACE(4, i);
        if ((i % kFlush) == kFlush - 1 || i == nb - 1) tc_commit(g_full);
      };
      for (int i = 0; i < nb; ++i) {
        const int s = i & 1;
        const int xs = i % kXStages, xu = i / kXStages;
        mbar_wait(x_full + xs, xu & 1);
        LR_TRACE(1, i);
        // No wait for GEMM2(i-2), the last reader of this Z/R buffer: tcgen05.mma instructions of a CTA execute in
        // the order they were issued, and GEMM2(i-2) was issued (after the epilogue had released the buffer through
        // p_full) before this point -- the hardware pipeline orders the overwrite behind it.
        tc_fence_after();
        const uint32_t aXh = smem_u32(sX + (2 * xs) * kBlkBytes), aXl = smem_u32(sX + (2 * xs + 1) * kBlkBytes);
        const uint32_t tZ = tmem + s * 128;
#pragma unroll
        for (int kk = 0; kk < kD / 16; ++kk) {
          const uint32_t ko = (uint32_t)kk * 4096;
          const uint32_t ka = (uint32_t)kk * 8;      // 16 fp16 of K = 8 TMEM columns
          mma_ts(tZ, tTh + ka, smem_desc(aXh + ko, 2048, 128), idesc1, kk > 0);
          if (NPROD == 3) {
            mma_ts(tZ, tTh + ka, smem_desc(aXl + ko, 2048, 128), idesc1, 1);
            mma_ts(tZ, tTl + ka, smem_desc(aXh + ko, 2048, 128), idesc1, 1);
          }
        }
        tc_commit(z_full + s);
        LR_TRACE(2, i);
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
    bool drain_pending = false;
    uint32_t drain_phase = 0;
    for (int i = 0; i < nb; ++i) {
      const int s = i & 1, u = i >> 1;
      mbar_wait(z_full + s, u & 1);
      if (ew == 0 && lane == 0) LR_TRACE(5, i);
      tc_fence_after();
      const uint32_t tZ = tmem + s * 128 + lane_base + cq * 32;
      uint32_t zr[32];
      tmem_ld32(tZ, zr);
      tmem_ld_wait();
      if (ew == 0 && lane == 0) LR_TRACE(6, i);
      uint32_t ph[16], pl[16];
      const int row_base = (b0 + i) * kTileR + cq * 32;
      const float* yy = sY + (i % kXStages) * kTileR + cq * 32;
      {
        // The residual, scaled by 2^10, straight from sigmoid(z) = 1 / (1 + 2^(-z log2 e)):
        //   r' = 1 / ((1 + e) 2^-10) - 2^10 y,   e = exp(-z).
        // ex2 saturates to +inf / 0 at the ends and rcp maps those to 0 / 2^10, so no |z| split is needed; rows
        // past n_rows meet all-zero X rows in GEMM2, so their residual needs no masking.
        // With the energy: softplus(z) - y z = z (1 - y) + ln(1 + e) for either sign of z (z is clamped at -80
        // there so that e stays finite); summed in float over the thread's 32 rows, then added to a double.
        const float nc = -1.4426950408889634f * zscale;
        const float k2 = 1.0f / 1024.0f;
        float vpart = 0.f;
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          const float4 y4 = *reinterpret_cast<const float4*>(yy + j);
          const float yv[4] = {y4.x, y4.y, y4.z, y4.w};
          float rr[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float zraw = __uint_as_float(zr[j + e]);
            float ez, sg;
            if (WANT_V) {
              const float z = fmaxf(zraw * zscale, -80.f);
              asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(ez) : "f"(-1.4426950408889634f * z));
              float l2;
              asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l2) : "f"(1.0f + ez));
              vpart += fmaf(0.6931471805599453f, l2, z * (1.0f - yv[e]));
            } else {
              asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(ez) : "f"(zraw * nc));
            }
            asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(sg) : "f"(fmaf(ez, k2, k2)));
            rr[e] = fmaf(yv[e], -rscale, sg);
          }
#pragma unroll
          for (int e = 0; e < 4; e += 2) {
            const __half2 h2 = __floats2half2_rn(rr[e], rr[e + 1]);
            const float2 hf = __half22float2(h2);
            const __half2 l2 = __floats2half2_rn(rr[e] - hf.x, rr[e + 1] - hf.y);
            ph[(j + e) >> 1] = *reinterpret_cast<const uint32_t*>(&h2);
            pl[(j + e) >> 1] = *reinterpret_cast<const uint32_t*>(&l2);
          }
        }
        if (WANT_V) {
          // padding rows (z = 0, y = 0) each added softplus(0) = ln 2: take them out again
          const int pad = min(32, max(0, row_base + 32 - a.n_rows));
          vsum += (double)vpart - (double)pad * (double)0.6931471805599453f;
        }
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
      if (ew == 0 && lane == 0) LR_TRACE(7, i);
      // drain the previous flush group's partial G (columns = dims cq*32..+31) into registers.  Done here, after
      // this block's residuals are computed and before they are handed over: the wait for GEMM2 of the group's
      // last block is then hidden under that work, and GEMM2 of this block -- the first of the next group, which
      // restarts the sum -- cannot be issued before both p_full and g_empty have been signalled.
      if (drain_pending) {
        mbar_wait(g_full, drain_phase);
        tc_fence_after();
        uint32_t gr[32];
        tmem_ld32(tG + lane_base + cq * 32, gr);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) gacc[j] += __uint_as_float(gr[j]);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(g_empty);
        drain_pending = false;
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_full + s);
      if (ew == 0 && lane == 0) LR_TRACE(8, i);
      if ((i % kFlush) == kFlush - 1 || i == nb - 1) {
        drain_pending = true;
        drain_phase = (uint32_t)((i / kFlush) & 1);
      }
    }
    if (drain_pending) {
      mbar_wait(g_full, drain_phase);
      tc_fence_after();
      uint32_t gr[32];
      tmem_ld32(tG + lane_base + cq * 32, gr);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; ++j) gacc[j] += __uint_as_float(gr[j]);
    }
    if (ew == 0 && lane == 0) LR_TRACE(11, 0);
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
  if (t == 0) LR_TRACE(12, 0);
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
