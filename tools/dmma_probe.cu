// dmma_probe.cu -- development microbenchmark: FP64 tensor-core (mma.sync f64) throughput
// on this GPU next to the DFMA pipe, to decide how the dense-Gaussian gradient should run.
#include <cstdio>
#include <cuda_runtime.h>

template <int ILP>
__global__ void __launch_bounds__(256) k_dmma884(double* sink, long long iters) {
  double c[ILP][2];
  double a = threadIdx.x * 1e-3, b = 1.0 + threadIdx.x * 1e-4;
#pragma unroll
  for (int i = 0; i < ILP; ++i) c[i][0] = c[i][1] = i;
  for (long long it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                   : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += c[i][0] + c[i][1];
  sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int ILP>
__global__ void __launch_bounds__(256) k_dmma16816(double* sink, long long iters) {
  double c[ILP][4];
  double a[8], b[4];
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = threadIdx.x * 1e-3 + i;
#pragma unroll
  for (int i = 0; i < 4; ++i) b[i] = 1.0 + threadIdx.x * 1e-4 + i;
#pragma unroll
  for (int i = 0; i < ILP; ++i) c[i][0] = c[i][1] = c[i][2] = c[i][3] = i;
  for (long long it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i)
      asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};"
                   : "+d"(c[i][0]), "+d"(c[i][1]), "+d"(c[i][2]), "+d"(c[i][3])
                   : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]),
                     "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
  sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int ILP>
__global__ void __launch_bounds__(256) k_dfma(double* sink, long long iters) {
  double x[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) x[i] = threadIdx.x * 1e-3 + i;
  for (long long it = 0; it < iters; ++it)
#pragma unroll
    for (int i = 0; i < ILP; ++i) x[i] = fma(x[i], 0.999999, 1e-7);
  double s = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += x[i];
  sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <class F>
float timeit(F f) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e30f;
  for (int r = 0; r < 4; ++r) {
    cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); if (r && ms < best) best = ms;
  }
  return best;
}

int main() {
  double* sink; cudaMalloc(&sink, 148 * 8 * 256 * 8);
  const long long iters = 20000;
  for (int wps = 1; wps <= 8; wps *= 2) {   // blocks of 256 threads per SM
    const int blocks = 148 * wps;
    float t1 = timeit([&] { k_dmma884<8><<<blocks, 256>>>(sink, iters); });
    float t2 = timeit([&] { k_dmma16816<4><<<blocks, 256>>>(sink, iters / 4); });
    float t3 = timeit([&] { k_dfma<8><<<blocks, 256>>>(sink, iters * 4); });
    double f1 = 2.0 * 256 * 8 * iters * (blocks * 8.0) / (t1 * 1e-3) / 1e12;          // 256 FMA per warp-mma
    double f2 = 2.0 * 2048 * 4 * (iters / 4) * (blocks * 8.0) / (t2 * 1e-3) / 1e12;   // 16*8*16 FMA per warp-mma
    double f3 = 2.0 * 8 * (iters * 4) * (blocks * 256.0) / (t3 * 1e-3) / 1e12;
    printf("blocks/SM=%d  dmma.m8n8k4 %.2f TF/s   dmma.m16n8k16 %.2f TF/s   dfma %.2f TF/s   (%s)\n", wps, f1, f2, f3,
           cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}
