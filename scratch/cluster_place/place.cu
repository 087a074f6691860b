// How many clusters of C CTAs x 1024 threads does a B200 actually run side by side?
// M streams each launch one spin kernel (T us); we record per-CTA smid, start, end (globaltimer).
#include <cuda_runtime.h>
#include <cooperative_groups.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>
struct Rec { unsigned smid; unsigned long long t0, t1; };
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
__global__ void __launch_bounds__(1024, 1) spin(Rec* rec, int base, unsigned long long ns) {
  unsigned long long t0 = gtime();
  while (gtime() - t0 < ns) { __nanosleep(200); }
  if (threadIdx.x == 0) {
    unsigned s; asm("mov.u32 %0, %smid;" : "=r"(s));
    rec[base + blockIdx.x] = Rec{s, t0, gtime()};
  }
}
int main(int argc, char** argv) {
  int C = argc > 1 ? atoi(argv[1]) : 8;      // cluster size; negative: plain grid of -C CTAs (cooperative)
  int M = argc > 2 ? atoi(argv[2]) : 15;
  unsigned long long us = argc > 3 ? atoll(argv[3]) : 1000;
  size_t smem = argc > 4 ? atoll(argv[4]) * 1024 : 0;  // dynamic smem per CTA (forces 1 CTA / SM when > 114 KB)
  if (smem) cudaFuncSetAttribute(spin, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  int n = C < 0 ? -C : C;
  Rec* rec; cudaMallocManaged(&rec, sizeof(Rec) * n * M);
  std::vector<cudaStream_t> st(M);
  for (auto& s : st) cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking);
  if (C > 8) cudaFuncSetAttribute(spin, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
  int maxc = -1;
  if (C > 0) {
    cudaLaunchConfig_t q{}; q.gridDim = dim3(C); q.blockDim = dim3(1024); q.dynamicSmemBytes = smem;
    cudaLaunchAttribute a{}; a.id = cudaLaunchAttributeClusterDimension; a.val.clusterDim = {(unsigned)C, 1, 1};
    q.attrs = &a; q.numAttrs = 1;
    cudaOccupancyMaxActiveClusters(&maxc, spin, &q);
  }
  for (int rep = 0; rep < 2; rep++) {
    cudaDeviceSynchronize();
    for (int i = 0; i < M; i++) {
      cudaLaunchConfig_t q{}; q.gridDim = dim3(n); q.blockDim = dim3(1024); q.stream = st[i]; q.dynamicSmemBytes = smem;
      cudaLaunchAttribute a{};
      if (C > 0) { a.id = cudaLaunchAttributeClusterDimension; a.val.clusterDim = {(unsigned)C, 1, 1}; }
      else { a.id = cudaLaunchAttributeCooperative; a.val.cooperative = 1; }
      q.attrs = &a; q.numAttrs = 1;
      unsigned long long ns = us * 1000;
      int base = i * n;
      cudaError_t e = cudaLaunchKernelEx(&q, spin, rec, base, ns);
      if (e != cudaSuccess) { printf("launch %d: %s\n", i, cudaGetErrorString(e)); return 1; }
    }
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("sync: %s\n", cudaGetErrorString(e)); return 1; }
  }
  unsigned long long lo = ~0ull, hi = 0;
  for (int i = 0; i < n * M; i++) { lo = std::min(lo, rec[i].t0); hi = std::max(hi, rec[i].t1); }
  // concurrency at the first kernel's midpoint
  int first_wave = 0;
  for (int i = 0; i < M; i++) if (rec[i * n].t0 < lo + us * 500) first_wave++;
  printf("C=%d M=%d spin=%llu us: occupancy API says %d clusters; wall %.1f us = %.2f waves; %d kernels started in the first half wave\n",
         C, M, us, maxc, (hi - lo) / 1e3, (hi - lo) / 1e3 / us, first_wave);
  for (int i = 0; i < M; i++) {
    printf("  k%02d start +%7.1f us  sm:", i, (rec[i * n].t0 - lo) / 1e3);
    for (int j = 0; j < n; j++) printf(" %3u", rec[i * n + j].smid);
    printf("\n");
  }
  return 0;
}
