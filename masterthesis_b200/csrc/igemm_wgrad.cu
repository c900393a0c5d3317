// Weight-gradient implicit GEMM for sm_100a.
//   dW[tap][m][n] += sum_{pixels} P[pix, m] * Q[pix (+) tap, n]
// P is the tensor indexed in the un-shifted pixel space (dy for Conv2d, x for ConvTranspose2d),
// Q the one sampled through the tap map (x for Conv2d, dy for ConvTranspose2d).  The reduction
// dimension is the pixel index, so both operands are MN-major in shared memory: one 4-D TMA
// box of (64 channels x 64 pixels) is a [K=64][MN=64] SWIZZLE_128B slab.  Split-K over pixel
// tiles; fp32 partial tiles are combined with vector red.global.add.
// Replaces autograd's Conv2d/ConvTranspose2d weight gradients (reference: the .backward() calls
// of src/models/adain_model.py:219,316,391).
#include "igemm.cuh"
#include "ptx.cuh"

namespace mt {

static constexpr int WG_BM = 128;
static constexpr int WG_BK = 64;           // pixels per stage
static constexpr int WG_CHUNK = 64 * 128;  // one (64 ch x 64 pix) slab in bytes
static constexpr int WG_MAX_STAGES = 8;

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b),
               "f"(c), "f"(d)
               : "memory");
}

__global__ void __launch_bounds__(192, 1)
igemm_wgrad_kernel(const __grid_constant__ CUtensorMap tmP, const __grid_constant__ CUtensorMap tmQ,
                   const WgradParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_full[WG_MAX_STAGES];
  __shared__ __align__(8) uint64_t bar_empty[WG_MAX_STAGES];
  __shared__ __align__(8) uint64_t bar_tmem_full;
  __shared__ uint32_t tmem_base_slot;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int n_chunks = p.BN / 64;
  const uint32_t stage_bytes = (2 + n_chunks) * WG_CHUNK;
  const int S = p.stages;

  // ---- work item
  int item = blockIdx.x;
  const int split = item % p.splits; item /= p.splits;
  const int n_tile = item % p.n_tiles; item /= p.n_tiles;
  const int m_tile = item % p.m_tiles; item /= p.m_tiles;
  const int tap = item;
  const int r = tap / p.kw, s = tap % p.kw;
  const int total_kt = p.tiles_w * p.tiles_h * p.tiles_n;
  const int per = (total_kt + p.splits - 1) / p.splits;
  const int kt_begin = split * per;
  const int kt_end = min(total_kt, kt_begin + per);
  if (kt_begin >= kt_end) return;  // uniform across the CTA
  const int iters = kt_end - kt_begin;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmP);
    prefetch_tmap(&tmQ);
    for (int i = 0; i < S; ++i) {
      mbar_init(smem_u32(&bar_full[i]), 1);
      mbar_init(smem_u32(&bar_empty[i]), 1);
    }
    mbar_init(smem_u32(&bar_tmem_full), 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(smem_u32(&tmem_base_slot), p.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_slot;

  if (warp == 0) {
    for (int it = 0; it < iters; ++it) {
      int kt = kt_begin + it;
      const int tw = kt % p.tiles_w; kt /= p.tiles_w;
      const int th = kt % p.tiles_h;
      const int tn = kt / p.tiles_h;
      const int w0 = tw * p.TW, h0 = th * p.TH, n0 = tn * p.TN;
      const int stage = it % S;
      const uint32_t ph = (it / S) & 1;
      mbar_wait(smem_u32(&bar_empty[stage]), ph ^ 1);
      if (lane == 0) {
        const uint32_t full = smem_u32(&bar_full[stage]);
        const uint32_t a_dst = smem_base + stage * stage_bytes;
        const uint32_t b_dst = a_dst + 2 * WG_CHUNK;
        mbar_expect_tx(full, stage_bytes);
        tma_load_4d(a_dst, &tmP, full, m_tile * WG_BM, w0, h0, n0);
        tma_load_4d(a_dst + WG_CHUNK, &tmP, full, m_tile * WG_BM + 64, w0, h0, n0);
        const int qw = w0 * p.mul_w + p.off_w + s * p.step_w;
        const int qh = h0 * p.mul_h + p.off_h + r * p.step_h;
        for (int c = 0; c < n_chunks; ++c)
          tma_load_4d(b_dst + c * WG_CHUNK, &tmQ, full, n_tile * p.BN + c * 64, qw, qh, n0);
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    const uint32_t idesc = umma_idesc_bf16(WG_BM, p.BN, 1, 1);
    for (int it = 0; it < iters; ++it) {
      const int stage = it % S;
      const uint32_t ph = (it / S) & 1;
      mbar_wait(smem_u32(&bar_full[stage]), ph);
      tc_fence_after();
      if (lane == 0) {
        const uint32_t a_addr = smem_base + stage * stage_bytes;
        const uint32_t b_addr = a_addr + 2 * WG_CHUNK;
#pragma unroll
        for (int k = 0; k < WG_BK / 16; ++k) {
          // 16 pixels = two 8-row groups of 1024 B; chunks of 64 channels are WG_CHUNK apart.
          const uint64_t ad = umma_smem_desc(a_addr + k * 2048, WG_CHUNK, 1024);
          const uint64_t bd = umma_smem_desc(b_addr + k * 2048, WG_CHUNK, 1024);
          umma_bf16(tmem_base, ad, bd, idesc, (it | k) != 0 ? 1u : 0u);
        }
        umma_commit(smem_u32(&bar_empty[stage]));
        if (it == iters - 1) umma_commit(smem_u32(&bar_tmem_full));
      }
      __syncwarp();
    }
  } else {
    const int q = warp & 3;
    const int m = m_tile * WG_BM + q * 32 + lane;
    const bool row_ok = m < p.M_valid;
    float* orow = p.out + (long long)tap * p.out_tap_stride + (long long)m * p.out_m_stride;
    const int n_base = n_tile * p.BN;
    mbar_wait(smem_u32(&bar_tmem_full), 0);
    tc_fence_after();
    for (int c0 = 0; c0 < p.BN; c0 += 16) {
      if (n_base + c0 >= p.N_valid) break;
      uint32_t v[16];
      tmem_ld16(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + c0, v);
      tmem_ld_wait();
      if (row_ok) {
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const int n = n_base + c0 + g * 4;
          if (n < p.N_valid)
            red_add_v4(orow + n, __uint_as_float(v[g * 4 + 0]), __uint_as_float(v[g * 4 + 1]),
                       __uint_as_float(v[g * 4 + 2]), __uint_as_float(v[g * 4 + 3]));
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, p.tmem_cols);
}

int launch_wgrad(const CUtensorMap& tmP, const CUtensorMap& tmQ, const WgradParams& p,
                 cudaStream_t stream) {
  const int stage_bytes = (2 + p.BN / 64) * WG_CHUNK;
  const size_t smem = (size_t)p.stages * stage_bytes + 1024;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(igemm_wgrad_kernel,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024);
    if (e != cudaSuccess) return -(3000 + (int)e);
    attr_set = true;
  }
  const long long grid = (long long)p.kh * p.kw * p.m_tiles * p.n_tiles * p.splits;
  if (grid <= 0 || grid > 0x7fffffffLL) return -1000;
  igemm_wgrad_kernel<<<(unsigned)grid, 192, smem, stream>>>(tmP, tmQ, p);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : -(int)e;
}

}  // namespace mt
