// Implicit-GEMM convolution for sm_100a: TMA-staged NHWC bf16 tiles, tcgen05.mma with the
// fp32 accumulator in TMEM, warp-specialised (TMA producer / MMA issuer / 4 epilogue warps).
//
// One launch computes  out[pix, co] = act( sum_{tap, ci} A[pix (+) tap, ci] * W[tap][co][ci] + b[co] )
// for every conv-shaped contraction of the hot path (reference call sites:
// src/models/core/blocks.py:30-42 Conv2d, :69-73 ConvTranspose2d): forward, data-gradient and
// the sub-pixel phases of the transposed convolutions are all expressed through the affine
// tap -> (dh, dw, weight index) maps in IgemmParams.
#include <stdlib.h>
#include <stdio.h>
#include "igemm.cuh"
#include "ptx.cuh"

namespace mt {

static constexpr int BM = 128;
static constexpr int BK = 64;
static constexpr int A_STAGE_BYTES = BM * BK * 2;  // 16 KiB
static constexpr int MAX_STAGES = 8;
__device__ long long g_ts[64];   // debug timestamps of CTA 0 (MT_IGEMM_DBG & 32)

__device__ __forceinline__ float apply_act(float v, int act, float slope) {
  if (act == ACT_RELU) return fmaxf(v, 0.f);
  if (act == ACT_LRELU) return v > 0.f ? v : v * slope;
  if (act == ACT_TANH) return tanhf(v);
  return v;
}

static constexpr int NUM_THREADS = 320;   // warp 0 TMA, warp 1 MMA, warps 2-9 epilogue (2 per lane quadrant)

// Epilogue of one warp: TMEM lane quadrant q, columns [col_begin, col_end) of every 128-row
// sub-tile.  Activation / bias are compile-time so the per-element code is cvt + pack only: with a
// lone warp per scheduler the epilogue is issue-bound (measured: 75 k cycles for a 256x256 tile with
// the generic per-element code), hence the slim inner loop, x32 TMEM loads and 8 epilogue warps.
template <int ACT, bool BIAS>
__device__ __forceinline__ void epilogue_warp(const IgemmParams& p, uint32_t tmem_base, int q, int lane,
                                              int col_begin, int col_end, int c_base, int w0, int h0,
                                              int n0) {
  for (int t = 0; t < p.mt; ++t) {
    const int m = t * BM + q * 32 + lane;
    const int w_i = m % p.TW;
    const int h_i = (m / p.TW) % p.TH;
    const int n_i = m / (p.TW * p.TH);
    const int w = w0 + w_i, h = h0 + h_i, n = n0 + n_i;
    const bool row_ok = (w < p.Wt) && (h < p.Ht) && (n < p.Nt);
    __nv_bfloat16* orow = p.out + p.out_off + (long long)n * p.out_n_stride +
                          (long long)h * p.out_h_stride + (long long)w * p.out_w_stride;
    for (int c0 = col_begin; c0 < col_end; c0 += 32) {
      if (c_base + c0 >= p.n_store) break;  // warp-uniform
      uint32_t v[32];
      tmem_ld32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + t * p.BN + c0, v);
      tmem_ld_wait();
      if (!row_ok) continue;
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        const int c = c_base + c0 + g * 8;
        if (c < p.n_store) {
          float f[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) f[j] = __uint_as_float(v[g * 8 + j]);
          if (BIAS) {
            if (c + 8 <= p.n_bias) {
              const float4 b0 = __ldg(reinterpret_cast<const float4*>(p.bias + c));
              const float4 b1 = __ldg(reinterpret_cast<const float4*>(p.bias + c + 4));
              f[0] += b0.x; f[1] += b0.y; f[2] += b0.z; f[3] += b0.w;
              f[4] += b1.x; f[5] += b1.y; f[6] += b1.z; f[7] += b1.w;
            } else {
#pragma unroll
              for (int j = 0; j < 8; ++j)
                if (c + j < p.n_bias) f[j] += __ldg(p.bias + c + j);
            }
          }
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            if (ACT == ACT_RELU) f[j] = fmaxf(f[j], 0.f);
            if (ACT == ACT_LRELU) f[j] = f[j] > 0.f ? f[j] : f[j] * p.slope;
            if (ACT == ACT_TANH) f[j] = tanhf(f[j]);
          }
          uint4 pk;
          __nv_bfloat162 t0 = __floats2bfloat162_rn(f[0], f[1]);
          __nv_bfloat162 t1 = __floats2bfloat162_rn(f[2], f[3]);
          __nv_bfloat162 t2 = __floats2bfloat162_rn(f[4], f[5]);
          __nv_bfloat162 t3 = __floats2bfloat162_rn(f[6], f[7]);
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

__global__ void __launch_bounds__(NUM_THREADS, 1)
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
  const uint32_t a_bytes = A_STAGE_BYTES * p.mt;
  const uint32_t stage_bytes = a_bytes + p.BN * BK * 2;
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

  const long long T0 = clock64();
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
  const long long T1 = clock64();

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
          if (lane == 0 && (p.dbg & 1) && it >= S) {
            mbar_arrive(smem_u32(&bar_full[stage]));
          } else if (lane == 0) {
            const uint32_t full = smem_u32(&bar_full[stage]);
            const uint32_t a_dst = smem_base + stage * stage_bytes;
            mbar_expect_tx(full, stage_bytes);
            tma_load_4d(a_dst, &tmA, full, kc * BK, cw, ch, n0);
            tma_load_3d(a_dst + a_bytes, &tmB, full, kc * BK, n_tile * p.BN, widx);
          }
          __syncwarp();
        }
      }
    }
    if ((p.dbg & 32) && blockIdx.x == 0 && lane == 0) g_ts[6] = clock64() - T0;
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
        const uint32_t b_addr = a_addr + a_bytes;
#pragma unroll
        for (int k = 0; k < BK / 16; ++k) {
          const uint64_t bd = umma_smem_desc(b_addr + k * 32, 16, 1024);
          for (int t = 0; t < ((p.dbg & 2) ? 0 : p.mt); ++t) {   // the 128-row sub-tiles share the B tile
            const uint64_t ad = umma_smem_desc(a_addr + t * A_STAGE_BYTES + k * 32, 16, 1024);
            umma_bf16(tmem_base + t * p.BN, ad, bd, idesc, (it | k) != 0 ? 1u : 0u);
          }
        }
        umma_commit(smem_u32(&bar_empty[stage]));
        if (it == total_iters - 1) umma_commit(smem_u32(&bar_tmem_full));
      }
      __syncwarp();
      if ((p.dbg & 32) && blockIdx.x == 0 && lane == 0 && it < 40) g_ts[8 + it] = clock64() - T0;
    }
  } else {
    // ================= epilogue: TMEM -> regs -> bias/act -> bf16 -> HBM =================
    const int q = warp & 3;            // TMEM lane quadrant this warp may access
    const int grp = (warp - 2) >> 2;   // which half of the columns
    int per = p.BN / 2;
    if (per < 32) per = 32;
    const int col_begin = grp * per, col_end = min(p.BN, col_begin + per);
    const int c_base = n_tile * p.BN;
    mbar_wait(smem_u32(&bar_tmem_full), 0);
    tc_fence_after();
    const long long T2 = clock64();
    if ((p.dbg & 32) && blockIdx.x == 0 && lane == 0 && warp < 6) g_ts[48 + warp] = T2 - T0;
    if (col_begin < col_end) {
      const bool hb = p.n_bias > 0;
#define MT_EPI(A, B) epilogue_warp<A, B>(p, tmem_base, q, lane, col_begin, col_end, c_base, w0, h0, n0)
      switch (p.act) {
        case ACT_RELU:  if (hb) MT_EPI(ACT_RELU, true);  else MT_EPI(ACT_RELU, false);  break;
        case ACT_LRELU: if (hb) MT_EPI(ACT_LRELU, true); else MT_EPI(ACT_LRELU, false); break;
        case ACT_TANH:  if (hb) MT_EPI(ACT_TANH, true);  else MT_EPI(ACT_TANH, false);  break;
        default:        if (hb) MT_EPI(ACT_NONE, true);  else MT_EPI(ACT_NONE, false);  break;
      }
#undef MT_EPI
    }
  }

  // ---- teardown
  const long long T3 = clock64();
  if ((p.dbg & 32) && blockIdx.x == 0 && (threadIdx.x % 32 == 0) && warp < 6) {
    g_ts[warp] = T3 - T0;
    if (warp == 0) g_ts[7] = T1 - T0;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, p.tmem_cols);
}

int launch_igemm(const CUtensorMap& tmA, const CUtensorMap& tmB, const IgemmParams& p,
                 cudaStream_t stream) {
  const int stage_bytes = A_STAGE_BYTES * p.mt + p.BN * BK * 2;
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
  static int dbg = -1;
  if (dbg < 0) { const char* e = getenv("MT_IGEMM_DBG"); dbg = e ? atoi(e) : 0; }
  if (dbg) { IgemmParams q = p; q.dbg = dbg; igemm_conv_kernel<<<(unsigned)grid, NUM_THREADS, smem, stream>>>(tmA, tmB, q); }
  else
  igemm_conv_kernel<<<(unsigned)grid, NUM_THREADS, smem, stream>>>(tmA, tmB, p);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : -(int)e;
}

}  // namespace mt

extern "C" int mt_debug_timestamps(void* out_host) {
  cudaError_t e = cudaMemcpyFromSymbol(out_host, mt::g_ts, sizeof(long long) * 64);
  return e == cudaSuccess ? 0 : -(int)e;
}
