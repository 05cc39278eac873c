// cp.reduce.async.bulk (UBLKRED.G.S.ADD.F64) as the running-moments update: every thread stages one chain's row
// (d doubles) and its squares in shared memory and adds both to sum[C,d] / sumsq[C,d] with one bulk reduction each.
// Prints the time per pass over C rows and checks the sums.  nvcc -gencode arch=compute_100a,code=sm_100a
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
__global__ void __launch_bounds__(128) k(const double* __restrict__ q, double* __restrict__ sum, double* __restrict__ sumsq,
                                         int64_t C, int d, int per_lane) {
  extern __shared__ __align__(16) unsigned char smem[];
  double* s = (double*)smem + (size_t)threadIdx.x * d;
  const uint32_t sa = (uint32_t)__cvta_generic_to_shared(s);
  for (int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; c < C; c += (int64_t)gridDim.x * blockDim.x) {
    const double* row = q + c * d;
    for (int pass = 0; pass < 2; ++pass) {
      for (int j = 0; j < d; ++j) { const double x = row[j]; s[j] = pass ? x * x : x; }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      if (per_lane) {
        double* g = (pass ? sumsq : sum) + c * d;
        asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f64 [%0], [%1], %2;" ::"l"(g), "r"(sa), "r"(d * 8) : "memory");
      } else {
        __syncwarp();
        if ((threadIdx.x & 31) == 0) {
          double* g = (pass ? sumsq : sum) + c * d;
          asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f64 [%0], [%1], %2;" ::"l"(g), "r"(sa), "r"(32 * d * 8) : "memory");
        }
      }
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      __syncwarp();
    }
  }
}
int main() {
  const int d = 50;
  const int64_t C = 1 << 20;
  double *q, *s1, *s2;
  cudaMalloc(&q, C * d * 8); cudaMalloc(&s1, C * d * 8); cudaMalloc(&s2, C * d * 8);
  std::vector<double> h(C * d);
  for (size_t i = 0; i < h.size(); ++i) h[i] = (double)(i % 97) * 0.25 - 3.0;
  cudaMemcpy(q, h.data(), C * d * 8, cudaMemcpyHostToDevice);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * d * 8);
  for (int per_lane = 1; per_lane >= 0; --per_lane) {
    for (int grid : {296, 592, 1184}) {
      cudaMemset(s1, 0, C * d * 8); cudaMemset(s2, 0, C * d * 8);
      cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
      k<<<grid, 128, 128 * d * 8>>>(q, s1, s2, C, d, per_lane);
      cudaEventRecord(e0);
      const int reps = 5;
      for (int r = 0; r < reps; ++r) k<<<grid, 128, 128 * d * 8>>>(q, s1, s2, C, d, per_lane);
      cudaEventRecord(e1);
      cudaError_t err = cudaDeviceSynchronize();
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      std::vector<double> a(d), b(d);
      const int64_t c = 777777;
      cudaMemcpy(a.data(), s1 + c * d, d * 8, cudaMemcpyDeviceToHost);
      cudaMemcpy(b.data(), s2 + c * d, d * 8, cudaMemcpyDeviceToHost);
      double x = h[c * d + 7];
      printf("{\"per_lane\": %d, \"grid\": %d, \"ms_per_pass\": %.4f, \"err\": \"%s\", \"sum_ok\": %d, \"sumsq_ok\": %d}\n", per_lane, grid,
             ms / reps, cudaGetErrorString(err), a[7] == 6 * x, b[7] == 6 * x * x);
    }
  }
  return 0;
}
