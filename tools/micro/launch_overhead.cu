// Back-to-back launch cost of an (almost) empty kernel as a function of the launch configuration
// used by igemm_conv_kernel: 148 CTAs x 320 threads, up to 225 KiB dynamic shared memory, clusters of
// 2, TMEM allocation.  nvcc -gencode arch=compute_100a,code=sm_100a -o launch_overhead launch_overhead.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstring>

template <int TMEM>
__global__ void __launch_bounds__(320, 1) k(int spin, long long* out) {
  extern __shared__ uint8_t sm[];
  __shared__ uint32_t slot;
  if (TMEM) {
    if (threadIdx.x < 32) {
      uint32_t a = (uint32_t)__cvta_generic_to_shared(&slot);
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(a), "r"(512) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    __syncthreads();
  }
  long long t0 = clock64();
  while (clock64() - t0 < spin) {}
  if (threadIdx.x == 0 && out) sm[0] = 1;
  if (TMEM) {
    __syncthreads();
    if (threadIdx.x < 32)
      asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(slot), "r"(512) : "memory");
  }
}

template <int TMEM>
static float run(int grid, int smem, int cluster, int spin, int reps) {
  auto* kern = k<TMEM>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(grid); cfg.blockDim = dim3(320); cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = cluster; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = cluster > 1 ? 1 : 0;
  long long* out = nullptr;
  for (int i = 0; i < 5; ++i) cudaLaunchKernelEx(&cfg, kern, spin, out);
  cudaDeviceSynchronize();
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  for (int i = 0; i < reps; ++i) cudaLaunchKernelEx(&cfg, kern, spin, out);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) printf("error %s\n", cudaGetErrorString(e));
  return ms * 1e3f / reps;
}

int main() {
  const int reps = 200;
  for (int spin : {0, 20000}) {
    printf("spin %d cycles per CTA\n", spin);
    printf("  148 CTAs, 1 KiB smem, no cluster        : %6.2f us\n", run<0>(148, 1024, 1, spin, reps));
    printf("  148 CTAs, 225 KiB smem, no cluster      : %6.2f us\n", run<0>(148, 230400, 1, spin, reps));
    printf("  148 CTAs, 225 KiB smem, cluster 2       : %6.2f us\n", run<0>(148, 230400, 2, spin, reps));
    printf("  148 CTAs, 225 KiB smem, cluster 2, TMEM : %6.2f us\n", run<1>(148, 230400, 2, spin, reps));
    printf("  148 CTAs, 1 KiB smem, cluster 2, TMEM   : %6.2f us\n", run<1>(148, 1024, 2, spin, reps));
    printf("  148 CTAs, 100 KiB smem, cluster 2       : %6.2f us\n", run<0>(148, 102400, 2, spin, reps));
  }
  return 0;
}
