// Does a grid launched with cluster dim 2 and PREFERRED cluster dim 4 fill all 148 SMs (33 clusters of 4 + 8 of 2)?
// Each CTA records its actual cluster size, rank and SM; CTAs spin briefly so that all must be co-resident.
#include <cuda_runtime.h>
#include <stdio.h>
__global__ void __launch_bounds__(256, 1) probe(int* out, int* arrived, int total) {
  extern __shared__ char s[];
  unsigned ncta, rank, smid;
  asm volatile("mov.u32 %0, %%cluster_nctaid.x;" : "=r"(ncta));
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
  if (threadIdx.x == 0) {
    s[0] = 1;
    out[3 * blockIdx.x] = ncta; out[3 * blockIdx.x + 1] = rank; out[3 * blockIdx.x + 2] = smid;
    atomicAdd(arrived, 1);
    long long t0 = clock64();
    while (atomicAdd(arrived, 0) < total && clock64() - t0 < 2000000000ll) {}
    out[3 * gridDim.x + blockIdx.x] = atomicAdd(arrived, 0);   // how many CTAs were co-resident
  }
}
int main() {
  const int smem = 198 * 1024, grid = 148;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  int *out, *arr;
  cudaMalloc(&out, 4 * grid * 4 * sizeof(int)); cudaMalloc(&arr, sizeof(int)); cudaMemset(arr, 0, sizeof(int));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid); cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute a[2];
  a[0].id = cudaLaunchAttributeClusterDimension; a[0].val.clusterDim.x = 2; a[0].val.clusterDim.y = 1; a[0].val.clusterDim.z = 1;
  a[1].id = cudaLaunchAttributePreferredClusterDimension; a[1].val.clusterDim.x = 4; a[1].val.clusterDim.y = 1; a[1].val.clusterDim.z = 1;
  cfg.attrs = a; cfg.numAttrs = 2;
  cudaError_t e = cudaLaunchKernelEx(&cfg, probe, out, arr, grid);
  printf("launch: %s\n", cudaGetErrorString(e));
  e = cudaDeviceSynchronize(); printf("sync: %s\n", cudaGetErrorString(e));
  int h[4 * 148]; cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
  int n4 = 0, n2 = 0, other = 0, minres = 1 << 30;
  for (int b = 0; b < grid; ++b) {
    if (h[3 * b] == 4) n4++; else if (h[3 * b] == 2) n2++; else other++;
    if (h[3 * grid + b] < minres) minres = h[3 * grid + b];
  }
  printf("CTAs in clusters of 4: %d, of 2: %d, other: %d; min co-resident count seen: %d\n", n4, n2, other, minres);
  for (int b = 0; b < 16; ++b) printf("block %d: ncta %d rank %d sm %d\n", b, h[3 * b], h[3 * b + 1], h[3 * b + 2]);
  for (int b = 124; b < 148; ++b) printf("block %d: ncta %d rank %d sm %d\n", b, h[3 * b], h[3 * b + 1], h[3 * b + 2]);
  return 0;
}
