// Micro-benchmark: issue rate of tcgen05.mma (bf16, K = 16 per instruction) for M=128 and several N,
// back to back from one thread with no pipeline protocol around it.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o mma_rate mma_rate.cu && ./mma_rate
#include <cstdio>
#include <cuda_runtime.h>
#include "../../masterthesis_b200/csrc/ptx.cuh"
using namespace mt;

__global__ void __launch_bounds__(128, 1) rate_kernel(int N, int n_mma, int commit_every, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  for (int i = threadIdx.x; i < 48 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem_raw)[i + 256] = 0;
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); fence_barrier_init(); }
  if (warp == 0) tmem_alloc(smem_u32(&slot), 512);
  fence_proxy_async();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tm = slot;
  if (warp == 1 && lane == 0) {
    const uint32_t idesc = umma_idesc_bf16(128, N, 0, 0);
    const long long t0 = clock64();
    for (int i = 0; i < n_mma; ++i) {
      const int k = i & 3;
      const uint64_t ad = umma_smem_desc(base + k * 32, 16, 1024);
      const uint64_t bd = umma_smem_desc(base + 16384 + k * 32, 16, 1024);
      umma_bf16(tm, ad, bd, idesc, i != 0);
      if (commit_every && (i % commit_every) == commit_every - 1 && i != n_mma - 1) {
        // commit to a barrier nobody waits on would change its phase; use a second barrier-free form:
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
      }
    }
    const long long t1 = clock64();
    umma_commit(smem_u32(&bar));
    // barrier may have flipped several times with commit_every; just spin on time-bounded wait of either parity
    long long tw = clock64();
    while (clock64() - tw < 200000) {}
    const long long t2 = clock64();
    if (blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
  }
  tc_fence_before(); __syncthreads();
  if (warp == 0) tmem_dealloc(tm, 512);
}

// variant that measures completion: waits on the final commit (no intermediate commits)
__global__ void __launch_bounds__(128, 1) done_kernel(int N, int n_mma, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  for (int i = threadIdx.x; i < 48 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem_raw)[i + 256] = 0;
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); fence_barrier_init(); }
  if (warp == 0) tmem_alloc(smem_u32(&slot), 512);
  fence_proxy_async();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tm = slot;
  if (warp == 1) {
    const long long t0 = clock64();
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16(128, N, 0, 0);
      for (int i = 0; i < n_mma; ++i) {
        const int k = i & 3;
        const uint64_t ad = umma_smem_desc(base + k * 32, 16, 1024);
        const uint64_t bd = umma_smem_desc(base + 16384 + k * 32, 16, 1024);
        umma_bf16(tm, ad, bd, idesc, i != 0);
      }
      umma_commit(smem_u32(&bar));
    }
    __syncwarp();
    mbar_wait(smem_u32(&bar), 0);
    const long long t1 = clock64();
    if (blockIdx.x == 0 && lane == 0) out[0] = t1 - t0;
  }
  tc_fence_before(); __syncthreads();
  if (warp == 0) tmem_dealloc(tm, 512);
}

int main() {
  long long* d; cudaMalloc(&d, 64);
  cudaFuncSetAttribute(done_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  for (int grid : {1, 148}) {
    for (int N : {64, 128, 256}) {
      for (int n : {256, 1024}) {
        done_kernel<<<grid, 128, 64 * 1024>>>(N, n, d);
        cudaError_t e = cudaDeviceSynchronize();
        long long h[2] = {0, 0};
        cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
        printf("grid %3d  M=128 N=%3d  %4d MMAs: %8lld clks  -> %.1f clks/MMA  (%s)\n", grid, N, n, h[0], (double)h[0] / n, cudaGetErrorString(e));
      }
    }
  }
  return 0;
}
