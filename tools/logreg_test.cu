// logreg_test.cu -- development harness for the tcgen05 logistic-regression gradient:
// runs k_logreg_grad_v1 on random data, checks G and V against a double-precision CPU
// evaluation and times the sweep.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 tools/logreg_test.cu
//   ./a.out N d C slabs reps mode      mode 0: V + 3 products, 1: 3 products, 2: 1 product
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../hmc_b200/csrc/logreg.cuh"

namespace hmcb {
void set_error(const char*, ...) {}
int cuda_fail(cudaError_t e, const char* w) { fprintf(stderr, "CUDA %s at %s\n", cudaGetErrorString(e), w); return -3; }
int sm_count() { return 148; }
}  // namespace hmcb
using namespace hmcb;
using namespace hmcb::lr;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); return 1; } } while (0)

int main(int argc, char** argv) {
  const int N = argc > 1 ? atoi(argv[1]) : 1000, d = argc > 2 ? atoi(argv[2]) : 128, C = argc > 3 ? atoi(argv[3]) : 256;
  const int slabs = argc > 4 ? atoi(argv[4]) : 2, reps = argc > 5 ? atoi(argv[5]) : 1, mode = argc > 6 ? atoi(argv[6]) : 0;
  const int nblk = (N + kTileR - 1) / kTileR;
  srand(3);
  auto rnd = [] { return (double)rand() / RAND_MAX; };
  auto gauss = [&] { return std::sqrt(-2 * std::log(rnd() + 1e-12)) * std::cos(6.283185307 * rnd()); };
  std::vector<float> X((size_t)N * d), y((size_t)nblk * kTileR, 0.f), th((size_t)C * d);
  float xmax = 0;
  for (auto& v : X) { v = (float)(gauss() / std::sqrt((double)d)); xmax = std::fmax(xmax, std::fabs(v)); }
  for (int i = 0; i < N; ++i) y[i] = rnd() < 0.5 ? 1.f : 0.f;
  for (auto& v : th) v = (float)(0.7 * gauss());
  int ex; std::frexp(xmax, &ex);
  const int xs = 10 - ex;

  float *dX, *dy, *dth, *dG; double* dV; __half *dXh, *dXl;
  CK(cudaMalloc(&dX, X.size() * 4)); CK(cudaMalloc(&dy, y.size() * 4)); CK(cudaMalloc(&dth, th.size() * 4));
  CK(cudaMalloc(&dXh, (size_t)nblk * kBlkBytes)); CK(cudaMalloc(&dXl, (size_t)nblk * kBlkBytes));
  CK(cudaMalloc(&dG, (size_t)slabs * C * kD * 4)); CK(cudaMalloc(&dV, (size_t)slabs * C * 8));
  CK(cudaMemcpy(dX, X.data(), X.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dy, y.data(), y.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dth, th.data(), th.size() * 4, cudaMemcpyHostToDevice));
  k_logreg_prep_blk<<<256, 256>>>(dX, N, d, nblk, xs, nullptr, dXh, dXl);
  CK(cudaGetLastError());

  GradArgsV1 b{};
  b.Xh_blk = dXh; b.Xl_blk = dXl; b.y = dy; b.theta = dth; b.Gpart = dG; b.Vpart = dV; b.n_chains = C; b.d = d; b.n_rows = N;
  b.n_blocks = nblk; b.blocks_per_slab = (nblk + slabs - 1) / slabs; b.n_slabs = slabs; b.x_scale_log2 = xs;
  dim3 grid((C + kTileC - 1) / kTileC, slabs);
  CK(cudaFuncSetAttribute(k_logreg_grad_v1<true, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemV1));
  CK(cudaFuncSetAttribute(k_logreg_grad_v1<false, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemV1));
  CK(cudaFuncSetAttribute(k_logreg_grad_v1<false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemV1));
  CK(cudaMemset(dV, 0, (size_t)slabs * C * 8));
  // optional trace of CTA (0,0): argv[7] = 1
  const bool want_trace = argc > 7 && atoi(argv[7]) != 0;
  unsigned long long* dtr = nullptr;
  const int nb0 = b.blocks_per_slab < nblk ? b.blocks_per_slab : nblk;
  if (want_trace) {
    CK(cudaMalloc(&dtr, (size_t)nb0 * 16 * 8));
    CK(cudaMemset(dtr, 0, (size_t)nb0 * 16 * 8));
    b.trace = dtr;
    b.trace_x = argc > 8 ? atoi(argv[8]) : 0;
    b.trace_y = argc > 9 ? atoi(argv[9]) : 0;
  }
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e30f;
  for (int r = 0; r < reps; ++r) {
    cudaEventRecord(e0);
    if (mode == 0) k_logreg_grad_v1<true, 3><<<grid, kThreadsV1, kSmemV1>>>(b);
    else if (mode == 1) k_logreg_grad_v1<false, 3><<<grid, kThreadsV1, kSmemV1>>>(b);
    else k_logreg_grad_v1<false, 1><<<grid, kThreadsV1, kSmemV1>>>(b);
    cudaEventRecord(e1);
    CK(cudaEventSynchronize(e1));
    float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
  }
  CK(cudaGetLastError());
  std::vector<float> G((size_t)slabs * C * kD);
  std::vector<double> V((size_t)slabs * C);
  CK(cudaMemcpy(G.data(), dG, G.size() * 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(V.data(), dV, V.size() * 8, cudaMemcpyDeviceToHost));

  double max_abs = 0, gnorm = 0, vmax = 0;
  const int Ccheck = C < 32 ? C : 32;
  for (int ci = 0; ci < Ccheck; ++ci) {
    const int c = (int)((long long)ci * C / Ccheck);
    std::vector<double> g(d, 0.0);
    double v = 0;
    for (int i = 0; i < N; ++i) {
      double z = 0;
      for (int k = 0; k < d; ++k) z += (double)X[(size_t)i * d + k] * th[(size_t)c * d + k];
      const double r = 1.0 / (1.0 + std::exp(-z)) - y[i];
      v += (z > 0 ? z : 0) + std::log1p(std::exp(-std::fabs(z))) - y[i] * z;
      for (int k = 0; k < d; ++k) g[k] += r * X[(size_t)i * d + k];
    }
    double vg = 0;
    for (int s = 0; s < slabs; ++s) vg += V[(size_t)s * C + c];
    if (mode == 0) vmax = std::fmax(vmax, std::fabs(vg - v) / std::fabs(v));
    for (int k = 0; k < d; ++k) {
      double gg = 0;
      for (int s = 0; s < slabs; ++s) gg += G[((size_t)s * C + c) * kD + k];
      max_abs = std::fmax(max_abs, std::fabs(gg - g[k]));
      gnorm = std::fmax(gnorm, std::fabs(g[k]));
    }
  }
  const double flops = 4.0 * N * (double)kD * C;
  const int nprod = mode == 2 ? 1 : 3;
  if (want_trace) {
    std::vector<unsigned long long> tr((size_t)nb0 * 16);
    cudaMemcpy(tr.data(), dtr, tr.size() * 8, cudaMemcpyDeviceToHost);
    const unsigned long long t0 = tr[(size_t)8 * 16 + 1];
    printf("trace of CTA (0,0), cycles relative to block 8's x_full; per block: prod x_empty | mma x_full, G1 issued, p_full seen, G2 issued | "
           "epi z_full seen, Z loaded, R stored, p_full arrived\n");
    printf("CTA entry %lld, prologue done %lld, first x_full %lld, first z_full seen %lld, last p_full arrived %lld, G drained %lld, exit %lld\n",
           (long long)(tr[9] - t0), (long long)(tr[10] - t0), (long long)(tr[1] - t0), (long long)(tr[5] - t0),
           (long long)(tr[(size_t)(nb0 - 1) * 16 + 8] - t0), (long long)(tr[11] - t0), (long long)(tr[12] - t0));
    for (int i = 8; i < 12 && i < nb0; ++i) {
      printf("blk %2d:", i);
      for (int k = 0; k < 9; ++k) printf(" %7lld", (long long)(tr[(size_t)i * 16 + k] - t0));
      printf("\n");
    }
  }
  printf("mode %d N=%d d=%d C=%d slabs=%d: %.3f ms  algorithmic %.0f TFLOP/s, executed fp16 %.0f TFLOP/s  |G|max %.3g  max abs err %.3g (rel %.3g)  V rel err %.3g\n",
         mode, N, d, C, slabs, best, flops / (best * 1e-3) / 1e12, nprod * flops / (best * 1e-3) / 1e12, gnorm, max_abs,
         max_abs / gnorm, vmax);
  return 0;
}
