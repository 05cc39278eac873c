// logreg_test.cu -- development harness for the tcgen05 logistic-regression gradient:
// runs k_logreg_grad on random data and compares G and V against a double-precision
// CPU evaluation.   nvcc -gencode arch=compute_100a,code=sm_100a tools/logreg_test.cu
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
  const int slabs = argc > 4 ? atoi(argv[4]) : 2, reps = argc > 5 ? atoi(argv[5]) : 1;
  const int Npad = (N + kTileR - 1) / kTileR * kTileR;
  srand(3);
  auto rnd = [] { return (double)rand() / RAND_MAX; };
  auto gauss = [&] { return std::sqrt(-2 * std::log(rnd() + 1e-12)) * std::cos(6.283185307 * rnd()); };
  std::vector<float> X((size_t)N * d), y(Npad, 0.f), th((size_t)C * d);
  float xmax = 0;
  for (auto& v : X) { v = (float)(gauss() / std::sqrt((double)d)); xmax = std::fmax(xmax, std::fabs(v)); }
  for (int i = 0; i < N; ++i) y[i] = rnd() < 0.5 ? 1.f : 0.f;
  for (auto& v : th) v = (float)(0.7 * gauss());
  int ex; std::frexp(xmax, &ex);
  const int xs = 10 - ex;

  float *dX, *dy, *dth, *dG, *dV; __half *dXh, *dXl;
  CK(cudaMalloc(&dX, X.size() * 4)); CK(cudaMalloc(&dy, Npad * 4)); CK(cudaMalloc(&dth, th.size() * 4));
  CK(cudaMalloc(&dXh, (size_t)Npad * kD * 2)); CK(cudaMalloc(&dXl, (size_t)Npad * kD * 2));
  CK(cudaMalloc(&dG, (size_t)slabs * C * kD * 4)); CK(cudaMalloc(&dV, (size_t)slabs * C * 4));
  CK(cudaMemcpy(dX, X.data(), X.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dy, y.data(), Npad * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dth, th.data(), th.size() * 4, cudaMemcpyHostToDevice));
  k_logreg_prep<<<256, 256>>>(dX, N, d, Npad, xs, dXh, dXl);
  CK(cudaGetLastError());

  GradArgs a{};
  a.Xh = dXh; a.Xl = dXl; a.y = dy; a.theta = dth; a.Gpart = dG; a.Vpart = dV; a.n_chains = C; a.d = d; a.n_rows = N;
  const int blocks = Npad / kTileR;
  a.rows_per_slab = ((blocks + slabs - 1) / slabs) * kTileR; a.n_slabs = slabs; a.x_scale_log2 = xs;
  const size_t smem = 4 * 32768 + 512 + 64;
  CK(cudaFuncSetAttribute(k_logreg_grad<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((C + kTileC - 1) / kTileC, slabs);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e30f;
  for (int r = 0; r < reps; ++r) {
    cudaEventRecord(e0);
    k_logreg_grad<true><<<grid, 128, smem>>>(a);
    cudaEventRecord(e1);
    CK(cudaEventSynchronize(e1));
    float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
  }
  CK(cudaGetLastError());
  // ---- pipelined kernel (v1): pre-tiled X
  {
    const int nblk = Npad / kTileR;
    __half *dXhb, *dXlb;
    CK(cudaMalloc(&dXhb, (size_t)nblk * kTileR * kD * 2)); CK(cudaMalloc(&dXlb, (size_t)nblk * kTileR * kD * 2));
    k_logreg_prep_blk<<<256, 256>>>(dX, N, d, nblk, xs, dXhb, dXlb);
    CK(cudaGetLastError());
    GradArgsV1 b{};
    b.Xh_blk = dXhb; b.Xl_blk = dXlb; b.y = dy; b.theta = dth; b.Gpart = dG; b.Vpart = dV; b.n_chains = C; b.d = d; b.n_rows = N;
    b.n_blocks = nblk; b.blocks_per_slab = (nblk + slabs - 1) / slabs; b.n_slabs = slabs; b.x_scale_log2 = xs;
    const size_t smem1 = 6 * 32768 + 2 * 512 + 512 + 512 + 16 * 8 + 64;
    const int mode = argc > 6 ? atoi(argv[6]) : 0;   // 0: V + 3 products, 1: no V + 3 products, 2: no V + 1 product
    auto launch = [&] {
      if (mode == 0) k_logreg_grad_v1<true, 3><<<grid, kThreadsV1, smem1>>>(b);
      else if (mode == 1) k_logreg_grad_v1<false, 3><<<grid, kThreadsV1, smem1>>>(b);
      else k_logreg_grad_v1<false, 1><<<grid, kThreadsV1, smem1>>>(b);
    };
    CK(cudaFuncSetAttribute(k_logreg_grad_v1<true, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem1));
    CK(cudaFuncSetAttribute(k_logreg_grad_v1<false, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem1));
    CK(cudaFuncSetAttribute(k_logreg_grad_v1<false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem1));
    CK(cudaMemset(dG, 0, (size_t)slabs * C * kD * 4));
    float b1 = 1e30f;
    for (int r = 0; r < reps; ++r) {
      cudaEventRecord(e0);
      launch();
      cudaEventRecord(e1);
      CK(cudaEventSynchronize(e1));
      float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < b1) b1 = ms;
    }
    CK(cudaGetLastError());
    printf("v1 mode %d: %.3f ms  (v0 %.3f ms)\n", mode, b1, best);
    best = b1;
  }
  std::vector<float> G((size_t)slabs * C * kD), V((size_t)slabs * C);
  CK(cudaMemcpy(G.data(), dG, G.size() * 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(V.data(), dV, V.size() * 4, cudaMemcpyDeviceToHost));

  // CPU reference in double (a subset of chains when C is large)
  double max_abs = 0, max_rel = 0, gnorm = 0, vmax = 0;
  const int Ccheck = C < 64 ? C : 64;
  for (int ci = 0; ci < Ccheck; ++ci) {
    const int c = (int)((long long)ci * C / Ccheck);
    std::vector<double> g(d, 0.0);
    double v = 0;
    for (int i = 0; i < N; ++i) {
      double z = 0;
      for (int k = 0; k < d; ++k) z += (double)X[(size_t)i * d + k] * th[(size_t)c * d + k];
      const double s = 1.0 / (1.0 + std::exp(-z));
      const double r = s - y[i];
      v += (z > 0 ? z : 0) + std::log1p(std::exp(-std::fabs(z))) - y[i] * z;
      for (int k = 0; k < d; ++k) g[k] += r * X[(size_t)i * d + k];
    }
    double vg = 0;
    for (int s = 0; s < slabs; ++s) vg += V[(size_t)s * C + c];
    vmax = std::fmax(vmax, std::fabs(vg - v) / std::fabs(v));
    for (int k = 0; k < d; ++k) {
      double gg = 0;
      for (int s = 0; s < slabs; ++s) gg += G[((size_t)s * C + c) * kD + k];
      max_abs = std::fmax(max_abs, std::fabs(gg - g[k]));
      gnorm = std::fmax(gnorm, std::fabs(g[k]));
    }
  }
  max_rel = max_abs / gnorm;
  const double flops = 4.0 * N * (double)kD * C;   // two GEMMs, algorithmic (one product each)
  printf("N=%d d=%d C=%d slabs=%d: %.3f ms  %.1f TFLOP/s algorithmic (x3 executed)  |G| max %.3g  max abs err %.3g (rel to max %.3g)  V rel err %.3g\n",
         N, d, C, slabs, best, flops / (best * 1e-3) / 1e12, gnorm, max_abs, max_rel, vmax);
  return 0;
}
