// Which SM does block b of a full-occupancy persistent grid land on?  296 blocks x 128 threads with
// enough dynamic shared memory that two blocks fit per SM (the footprint of the DR stage kernels).
#include <cstdio>
#include <vector>
#include <map>
__global__ void k(int* smid, long long* t0) {
  extern __shared__ char s[];
  if (threadIdx.x == 0) {
    unsigned id;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(id));
    smid[blockIdx.x] = (int)id;
    t0[blockIdx.x] = clock64();
  }
  // stay resident long enough that all blocks are co-resident
  long long t = clock64();
  while (clock64() - t < 2000000) {}
  if (s[threadIdx.x] == 77) smid[0] = -1;
}
int main() {
  const int G = 296;
  int* d; long long* t;
  cudaMalloc(&d, G * 4); cudaMalloc(&t, G * 8);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  k<<<G, 128, 100 * 1024>>>(d, t);
  std::vector<int> h(G);
  cudaMemcpy(h.data(), d, G * 4, cudaMemcpyDeviceToHost);
  printf("err=%s\n", cudaGetErrorString(cudaGetLastError()));
  for (int b = 0; b < G; ++b) printf("%d%c", h[b], (b % 32 == 31) ? '\n' : ' ');
  printf("\n");
  std::map<int, int> first124;
  for (int b = 0; b < 124; ++b) first124[h[b]]++;
  int two = 0;
  for (auto& kv : first124) two += kv.second >= 2;
  printf("blocks 0..123 cover %zu SMs, %d of them hold two\n", first124.size(), two);
  return 0;
}
