// kbench.cu -- standalone timing harness for one transition-kernel shape (development
// tool, not part of libhmcb.so).  Builds the MODE_HMC kernel for Funnel/f64 at a given
// (EPL, LPC), runs it over C chains for L leapfrog steps and prints ms and the fraction
// of the measured DFMA peak.   nvcc -DKB_EPL=50 -DKB_LPC=1 [-DVARIANT=..] tools/kbench.cu
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../hmc_b200/csrc/launch.cuh"

#ifndef KB_EPL
#define KB_EPL 50
#endif
#ifndef KB_LPC
#define KB_LPC 1
#endif
#ifndef KB_MINB
#define KB_MINB (hmcb::min_blocks_for<double, KB_EPL, true, false>())
#endif

namespace hmcb {
static thread_local char g_err[256];
void set_error(const char* fmt, ...) { (void)fmt; }
int cuda_fail(cudaError_t e, const char* what) { fprintf(stderr, "CUDA %s at %s\n", cudaGetErrorString(e), what); return -3; }
int sm_count() { int v; cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, 0); return v; }
int max_optin_smem() { int v; cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, 0); return v; }
std::vector<Entry>& registry() { static std::vector<Entry> r; return r; }
template <typename T, int ILP>
__global__ void __launch_bounds__(256) k_probe(T* sink, long long iters) {
  T x[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) x[i] = (T)(threadIdx.x + i) * (T)1e-3;
  for (long long it = 0; it < iters; ++it)
#pragma unroll
    for (int u = 0; u < 8; ++u)
#pragma unroll
      for (int i = 0; i < ILP; ++i) x[i] = fma(x[i], (T)0.999999, (T)1e-7);
  T s = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += x[i];
  sink[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}
}  // namespace hmcb

using namespace hmcb;

int main(int argc, char** argv) {
  const int64_t C = argc > 1 ? atoll(argv[1]) : (1 << 20);
  const int L = argc > 2 ? atoi(argv[2]) : 150;
  const int d = argc > 3 ? atoi(argv[3]) : 50;
  const int soa = argc > 4 ? atoi(argv[4]) : 0;   // 1: structure-of-arrays [d][C] strides
  const int reps = 5;
  constexpr int THREADS = 128;
  double *q, *qo, *sink;
  cudaMalloc(&q, C * d * 8); cudaMalloc(&qo, C * d * 8); cudaMalloc(&sink, 148 * 8 * 256 * 8);
  std::vector<double> h(C * d);
  srand(1);
  for (auto& x : h) x = 2.0 * rand() / RAND_MAX - 1.0;
  cudaMemcpy(q, h.data(), C * d * 8, cudaMemcpyHostToDevice);
  hmcb_model hm{}; hm.kind = HMCB_MODEL_FUNNEL; hm.d = d; hm.dtype = HMCB_F64; hm.par[0] = 3.0;
  auto model = FunnelModel<double>::make(hm);
  TransArgs<double> a{};
  a.n_chains = C; a.d = d; a.q_in = q; a.q_out = qo; a.sc = soa ? 1 : d; a.sd = soa ? C : 1; a.seed = 1; a.nprops = 1; a.L[0] = L; a.eps[0] = 0.05;
  auto k = k_transition<FunnelModel<double>, double, KB_EPL, KB_LPC, true, MODE_HMC, THREADS, KB_MINB>;
  int per_sm = 0;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, THREADS, 0);
  cudaFuncAttributes fa; cudaFuncGetAttributes(&fa, k);
  int64_t need = (C * KB_LPC + THREADS - 1) / THREADS;
  int grid = (int)std::min<int64_t>(need, (int64_t)per_sm * sm_count());
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e30f, pk = 1e30f;
  for (int r = 0; r < reps; ++r) {
    a.iteration = r;
    cudaEventRecord(e0);
    k<<<grid, THREADS>>>(model, a);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); if (r && ms < best) best = ms;
    cudaEventRecord(e0);
    k_probe<double, 8><<<148 * 8, 256>>>(sink, 4000);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1); if (r && ms < pk) pk = ms;
  }
  cudaError_t err = cudaDeviceSynchronize();
  const double peak = 2.0 * 148 * 8 * 256 * 4000.0 * 64 / (pk * 1e-3) / 1e12;
  const double ach = (double)C * (L + 1) * 7.0 * d / (best * 1e-3) / 1e12;
  printf("%s EPL=%d LPC=%d regs=%d stack=%zu blocks/SM=%d grid=%d  C=%lld L=%d d=%d  %.3f ms  %.2f TF/s algorithmic = %.3f of measured DFMA peak %.2f TF/s  (%s)\n",
         soa ? "SoA" : "row-major", KB_EPL, KB_LPC, fa.numRegs, fa.localSizeBytes, per_sm, grid, (long long)C, L, d, best, ach, ach / peak, peak,
         cudaGetErrorString(err));
  return 0;
}
