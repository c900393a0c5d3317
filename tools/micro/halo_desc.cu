// Micro-experiment: can a K-major SWIZZLE_128B tcgen05.mma A operand start at an arbitrary 128-byte ROW of a
// larger tile that TMA wrote (1024-byte aligned base, address-based swizzle)?  That is what re-using one halo'd
// activation tile for all taps of a 3x3 filter needs: tap (r, s) reads rows shifted by r * pitch + s.
// A tile in shared memory: ROWS rows of 64 bf16 (128 B), row R chunk c stored at base + R*128 + ((c ^ (R&7)) << 4).
// MMA operand rows: m -> R = shift + (m / 8) * pitch + (m % 8)   (8-row groups `pitch` rows apart: SBO = pitch*128).
// Variant 0: descriptor base_offset field (bits 49-51) = 0.  Variant 1: base_offset = (start >> 7) & 7.
// Prints the max |error| against a host reference per (shift, variant).
//   nvcc -gencode arch=compute_100a,code=sm_100a -o halo_desc tools/micro/halo_desc.cu && ./halo_desc
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <cuda_runtime.h>
#include "../../masterthesis_b200/csrc/ptx.cuh"
using namespace mt;

#ifndef PITCH_ROWS
#define PITCH_ROWS 16
#endif
constexpr int PITCH = PITCH_ROWS, ROWS = 18 * PITCH + 40, N = 16;

__global__ void __launch_bounds__(128, 1) halo_kernel(const __nv_bfloat16* X, const __nv_bfloat16* W, int shift,
                                                      int variant, float* out) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_done;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* sm = smem_raw + (base - smem_u32(smem_raw));
  // A: ROWS x 64 bf16, swizzled like a TMA SWIZZLE_128B box at a 1024-aligned base
  for (int i = threadIdx.x; i < ROWS * 8; i += blockDim.x) {
    const int R = i / 8, c = i % 8;
    *reinterpret_cast<uint4*>(sm + R * 128 + ((c ^ (R & 7)) << 4)) = *reinterpret_cast<const uint4*>(X + R * 64 + c * 8);
  }
  const uint32_t b_off = ((ROWS * 128 + 1023) & ~1023);
  for (int i = threadIdx.x; i < N * 8; i += blockDim.x) {
    const int R = i / 8, c = i % 8;
    *reinterpret_cast<uint4*>(sm + b_off + R * 128 + ((c ^ (R & 7)) << 4)) = *reinterpret_cast<const uint4*>(W + R * 64 + c * 8);
  }
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar_done), 1); fence_barrier_init(); }
  if (warp == 0) tmem_alloc(smem_u32(&slot), 32);
  fence_proxy_async();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tm = slot;
  if (warp == 1 && lane == 0) {
    const uint32_t idesc = umma_idesc_bf16(128, N, 0, 0);
    const uint32_t a = base + shift * 128;
    for (int k = 0; k < 4; ++k) {
      uint64_t ad = umma_smem_desc(a + k * 32, 16, PITCH * 128);
      if (variant == 1) ad |= static_cast<uint64_t>((a >> 7) & 7) << 49;
      const uint64_t bd = umma_smem_desc(base + b_off + k * 32, 16, 1024);
      umma_bf16(tm, ad, bd, idesc, k > 0);
    }
    umma_commit(smem_u32(&bar_done));
  }
  __syncwarp();
  mbar_wait(smem_u32(&bar_done), 0);
  tc_fence_after();
  uint32_t v[16];
  tmem_ld16(tm + (static_cast<uint32_t>(warp * 32) << 16), v);
  tmem_ld_wait();
  for (int j = 0; j < 16; ++j) out[(warp * 32 + lane) * 16 + j] = __uint_as_float(v[j]);
  tc_fence_before(); __syncthreads();
  if (warp == 0) tmem_dealloc(tm, 32);
}

int main() {
  std::vector<__nv_bfloat16> hx(ROWS * 64), hw(N * 64);
  std::vector<float> fx(ROWS * 64), fw(N * 64);
  srand(1);
  for (int i = 0; i < ROWS * 64; ++i) { fx[i] = (rand() % 17 - 8) / 8.f; hx[i] = __float2bfloat16(fx[i]); }
  for (int i = 0; i < N * 64; ++i) { fw[i] = (rand() % 9 - 4) / 4.f; hw[i] = __float2bfloat16(fw[i]); }
  __nv_bfloat16 *dx, *dw; float* dout;
  cudaMalloc(&dx, hx.size() * 2); cudaMalloc(&dw, hw.size() * 2); cudaMalloc(&dout, 128 * 16 * 4);
  cudaMemcpy(dx, hx.data(), hx.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dw, hw.data(), hw.size() * 2, cudaMemcpyHostToDevice);
  const int smem = ROWS * 128 + 4096 + 2048;
  cudaFuncSetAttribute(halo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  std::vector<float> ho(128 * 16);
  const int shifts[] = {0, 1, 2, 3, 7, 8, 9, 16, 17, 18, 33, 34};
  for (int variant = 0; variant < 2; ++variant)
    for (int shift : shifts) {
      cudaMemset(dout, 0, 128 * 16 * 4);
      halo_kernel<<<1, 128, smem>>>(dx, dw, shift, variant, dout);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("variant %d shift %d: %s\n", variant, shift, cudaGetErrorString(e)); return 1; }
      cudaMemcpy(ho.data(), dout, 128 * 16 * 4, cudaMemcpyDeviceToHost);
      double worst = 0; int bad_rows = 0;
      for (int m = 0; m < 128; ++m) {
        const int R = shift + (m / 8) * PITCH + (m % 8);
        double row_err = 0;
        for (int n = 0; n < N; ++n) {
          double ref = 0;
          for (int k = 0; k < 64; ++k) ref += (double)fx[R * 64 + k] * fw[n * 64 + k];
          row_err = fmax(row_err, fabs(ref - ho[m * 16 + n]));
        }
        worst = fmax(worst, row_err);
        bad_rows += row_err > 1e-3;
      }
      printf("variant %d (base_offset %s) shift %2d: max err %.4f, wrong rows %d / 128\n", variant,
             variant ? "= (start>>7)&7" : "= 0", shift, worst, bad_rows);
    }
  return 0;
}
