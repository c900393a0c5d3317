// Implicit-GEMM convolution for sm_100a: TMA-staged NHWC bf16 tiles, tcgen05.mma with the
// fp32 accumulator in TMEM, warp-specialised (TMA producer / MMA issuer / 4 epilogue warps).
//
// One launch computes  out[pix, co] = act( sum_{tap, ci} A[pix (+) tap, ci] * W[tap][co][ci] + b[co] )
// for every conv-shaped contraction of the hot path (reference call sites:
// src/models/core/blocks.py:30-42 Conv2d, :69-73 ConvTranspose2d): forward, data-gradient and
// the sub-pixel phases of the transposed convolutions are all expressed through the affine
// tap -> (dh, dw, weight index) maps in IgemmParams.
#include "igemm.cuh"
#include "ptx.cuh"

namespace mt {

static constexpr int BM = 128;
static constexpr int BK = 64;
static constexpr int A_STAGE_BYTES = BM * BK * 2;  // 16 KiB
static constexpr int MAX_STAGES = 8;

__device__ __forceinline__ float apply_act(float v, int act, float slope) {
  if (act == ACT_RELU) return fmaxf(v, 0.f);
  if (act == ACT_LRELU) return v > 0.f ? v : v * slope;
  if (act == ACT_TANH) return tanhf(v);
  return v;
}

__global__ void __launch_bounds__(192, 1)
igemm_conv_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  const IgemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_full[MAX_STAGES];
  __shared__ __align__(8) uint64_t bar_empty[MAX_STAGES];
  __shared__ __align__(8) uint64_t bar_tmem_full;
  __shared__ uint32_t tmem_base_slot;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t stage_bytes = A_STAGE_BYTES + p.BN * BK * 2;
  const int S = p.stages;

  // ---- tile coordinates
  int tile = blockIdx.x;
  const int n_tile = tile % p.n_tiles;
  tile /= p.n_tiles;
  const int tw = tile % p.tiles_w;
  tile /= p.tiles_w;
  const int th = tile % p.tiles_h;
  const int tn = tile / p.tiles_h;
  const int w0 = tw * p.TW, h0 = th * p.TH, n0 = tn * p.TN;
  const int total_iters = p.kh * p.kw * p.k_chunks;

  // ---- one-time setup
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
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
    // ================= TMA producer =================
    int it = 0;
    for (int r = 0; r < p.kh; ++r) {
      for (int s = 0; s < p.kw; ++s) {
        const int ch = h0 * p.mul_h + p.off_h + r * p.step_h;
        const int cw = w0 * p.mul_w + p.off_w + s * p.step_w;
        const int widx = p.w_base + r * p.w_step_r + s * p.w_step_s;
        for (int kc = 0; kc < p.k_chunks; ++kc, ++it) {
          const int stage = it % S;
          const uint32_t ph = (it / S) & 1;
          mbar_wait(smem_u32(&bar_empty[stage]), ph ^ 1);
          if (lane == 0) {
            const uint32_t full = smem_u32(&bar_full[stage]);
            const uint32_t a_dst = smem_base + stage * stage_bytes;
            mbar_expect_tx(full, stage_bytes);
            tma_load_4d(a_dst, &tmA, full, kc * BK, cw, ch, n0);
            tma_load_3d(a_dst + A_STAGE_BYTES, &tmB, full, kc * BK, n_tile * p.BN, widx);
          }
          __syncwarp();
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    const uint32_t idesc = umma_idesc_bf16(BM, p.BN, 0, 0);
    for (int it = 0; it < total_iters; ++it) {
      const int stage = it % S;
      const uint32_t ph = (it / S) & 1;
      mbar_wait(smem_u32(&bar_full[stage]), ph);
      tc_fence_after();
      if (lane == 0) {
        const uint32_t a_addr = smem_base + stage * stage_bytes;
        const uint32_t b_addr = a_addr + A_STAGE_BYTES;
#pragma unroll
        for (int k = 0; k < BK / 16; ++k) {
          const uint64_t ad = umma_smem_desc(a_addr + k * 32, 16, 1024);
          const uint64_t bd = umma_smem_desc(b_addr + k * 32, 16, 1024);
          umma_bf16(tmem_base, ad, bd, idesc, (it | k) != 0 ? 1u : 0u);
        }
        umma_commit(smem_u32(&bar_empty[stage]));
        if (it == total_iters - 1) umma_commit(smem_u32(&bar_tmem_full));
      }
      __syncwarp();
    }
  } else {
    // ================= epilogue: TMEM -> regs -> bias/act -> bf16 -> HBM =================
    const int q = warp & 3;  // TMEM lane quadrant this warp may access
    const int m = q * 32 + lane;
    const int w_i = m % p.TW;
    const int h_i = (m / p.TW) % p.TH;
    const int n_i = m / (p.TW * p.TH);
    const int w = w0 + w_i, h = h0 + h_i, n = n0 + n_i;
    const bool row_ok = (w < p.Wt) && (h < p.Ht) && (n < p.Nt);
    __nv_bfloat16* orow = p.out + p.out_off + (long long)n * p.out_n_stride +
                          (long long)h * p.out_h_stride + (long long)w * p.out_w_stride;
    const int c_base = n_tile * p.BN;

    mbar_wait(smem_u32(&bar_tmem_full), 0);
    tc_fence_after();
    for (int c0 = 0; c0 < p.BN; c0 += 16) {
      if (c_base + c0 >= p.n_store) break;  // warp-uniform
      uint32_t v[16];
      tmem_ld16(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + c0, v);
      tmem_ld_wait();
      float f[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int c = c_base + c0 + j;
        float x = __uint_as_float(v[j]);
        if (c < p.n_bias) x += __ldg(p.bias + c);
        f[j] = apply_act(x, p.act, p.slope);
      }
      if (row_ok) {
#pragma unroll
        for (int g = 0; g < 2; ++g) {
          const int c = c_base + c0 + g * 8;
          if (c < p.n_store) {
            uint4 pk;
            __nv_bfloat162 t0 = __floats2bfloat162_rn(f[g * 8 + 0], f[g * 8 + 1]);
            __nv_bfloat162 t1 = __floats2bfloat162_rn(f[g * 8 + 2], f[g * 8 + 3]);
            __nv_bfloat162 t2 = __floats2bfloat162_rn(f[g * 8 + 4], f[g * 8 + 5]);
            __nv_bfloat162 t3 = __floats2bfloat162_rn(f[g * 8 + 6], f[g * 8 + 7]);
            pk.x = *reinterpret_cast<uint32_t*>(&t0);
            pk.y = *reinterpret_cast<uint32_t*>(&t1);
            pk.z = *reinterpret_cast<uint32_t*>(&t2);
            pk.w = *reinterpret_cast<uint32_t*>(&t3);
            *reinterpret_cast<uint4*>(orow + c) = pk;
          }
        }
      }
    }
  }

  // ---- teardown
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, p.tmem_cols);
}

int launch_igemm(const CUtensorMap& tmA, const CUtensorMap& tmB, const IgemmParams& p,
                 cudaStream_t stream) {
  const int stage_bytes = A_STAGE_BYTES + p.BN * BK * 2;
  const size_t smem = (size_t)p.stages * stage_bytes + 1024;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(igemm_conv_kernel,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024);
    if (e != cudaSuccess) return -(3000 + (int)e);
    attr_set = true;
  }
  const long long grid = (long long)p.tiles_w * p.tiles_h * p.tiles_n * p.n_tiles;
  if (grid <= 0 || grid > 0x7fffffffLL) return -1000;
  igemm_conv_kernel<<<(unsigned)grid, 192, smem, stream>>>(tmA, tmB, p);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : -(int)e;
}

}  // namespace mt
