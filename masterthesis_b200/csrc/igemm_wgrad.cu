// Weight-gradient implicit GEMM for sm_100a.
//   dW[tap][m][n] += sum_{pixels} P[pix, m] * Q[pix (+) tap, n]
// P is the tensor indexed in the un-shifted pixel space (dy for Conv2d, x for ConvTranspose2d),
// Q the one sampled through the tap map (x for Conv2d, dy for ConvTranspose2d).  The reduction
// dimension is the pixel index, so both operands are MN-major in shared memory: one 4-D TMA
// box of (64 channels x 64 pixels) is a [K=64][MN=64] SWIZZLE_128B slab.  Split-K over pixel
// tiles; every split stores its fp32 partial tile into its own slab of the output ([splits][taps][M][N],
// plain 16-byte stores: no zeroing, no atomics) and the unpack kernel adds the slabs in split order, so
// weight gradients are reproducible bit for bit.
// Replaces autograd's Conv2d/ConvTranspose2d weight gradients (reference: the .backward() calls
// of src/models/adain_model.py:219,316,391).
#include <stdlib.h>
#include <string.h>
#include "igemm.cuh"
#include "ptx.cuh"

namespace mt {

static constexpr int WG_BM = 128;
static constexpr int WG_BK = 64;           // pixels per stage
static constexpr int WG_CHUNK = 64 * 128;  // one (64 ch x 64 pix) slab in bytes
// Box mode (G = 3 taps of one filter row, stride-1 traversal): the pixel tile is 8 x 8 and ONE (8 + 2) x 8 pixel
// box of the shifted operand serves the three taps: the 8-pixel K groups of the MN-major SWIZZLE_128B operand
// are then `WG_BOX_PITCH` rows apart (descriptor SBO) and tap s starts s rows into the box (the swizzle is a
// function of the shared-memory address, tools/micro/halo_desc.cu).  2.4x fewer bytes of that operand.
static constexpr int WG_BOX_PITCH = 10;
static constexpr int WG_BOX_CHUNK = WG_BOX_PITCH * 8 * 128;   // 10 KiB per 64-channel chunk
template <int CG, int BN, int G, bool BOX>
struct WgStage {
  static constexpr int B_CHUNKS = BN / CG / 64;
  static constexpr uint32_t B_CHUNK_BYTES = BOX ? WG_BOX_CHUNK : WG_CHUNK;
  static constexpr uint32_t BYTES = 2 * WG_CHUNK + (BOX ? 1 : G) * B_CHUNKS * B_CHUNK_BYTES;
};

__device__ __forceinline__ void st_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("st.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d)
               : "memory");
}

// One pass over the ring (k-iterations kb .. kb+S-1) as a compile-time recursion: the stage index and
// with it every shared-memory descriptor is a constant (MMA issue-rate rule, see igemm_fwd.cu).
struct WgCtx {
  uint32_t bar_full, bar_empty, smem_base, tmem_base, phase_bits;
  int iters, kt_begin, m_tile, n_tile, tap0, n_taps, rank, dbg;
};

template <int CG, int BN, int G, bool BOX, int S, int STAGE>
__device__ __forceinline__ void wg_produce_round(const CUtensorMap* tmP, const CUtensorMap* tmQ,
                                                 const WgradParams& p, WgCtx& c, int kb) {
  constexpr int B_CHUNKS = BN / CG / 64;                      // 64-channel slabs of Q staged by this CTA, per tap
  constexpr uint32_t STAGE_BYTES = WgStage<CG, BN, G, BOX>::BYTES;
  if constexpr (STAGE < S) {
    if (kb + STAGE < c.iters) {
      int kt = c.kt_begin + kb + STAGE;
      const int tw = kt % p.tiles_w; kt /= p.tiles_w;
      const int th = kt % p.tiles_h;
      const int tn = kt / p.tiles_h;
      const int w0 = tw * p.TW, h0 = th * p.TH, n0 = tn * p.TN;
      mbar_wait(c.bar_empty + 8 * STAGE, ((c.phase_bits >> STAGE) & 1) ^ 1);
      c.phase_bits ^= 1u << STAGE;
      const uint32_t a_dst = c.smem_base + STAGE * STAGE_BYTES;
      const uint32_t b_dst = a_dst + 2 * WG_CHUNK;
      const uint32_t bytes = BOX ? STAGE_BYTES : (2 + c.n_taps * B_CHUNKS) * WG_CHUNK;
      const uint32_t full = CG == 2 ? ((c.bar_full + 8 * STAGE) & PEER_BIT_MASK) : (c.bar_full + 8 * STAGE);
      // CG == 2: the leader's barrier counts the bytes of both CTAs of the pair
      if ((p.dbg & 1) && kb >= S) {            // debug: handshake only, no TMA traffic (stale operands)
        if (CG == 1 || c.rank == 0) mbar_arrive(full);
        wg_produce_round<CG, BN, G, BOX, S, STAGE + 1>(tmP, tmQ, p, c, kb);
        return;
      }
      if (CG == 1 || c.rank == 0) mbar_expect_tx(full, CG * bytes);
      if (CG == 2) {
        tma_load_4d_2sm(a_dst, tmP, full, c.m_tile * WG_BM, w0, h0, n0);
        tma_load_4d_2sm(a_dst + WG_CHUNK, tmP, full, c.m_tile * WG_BM + 64, w0, h0, n0);
      } else {
        tma_load_4d(a_dst, tmP, full, c.m_tile * WG_BM, w0, h0, n0);
        tma_load_4d(a_dst + WG_CHUNK, tmP, full, c.m_tile * WG_BM + 64, w0, h0, n0);
      }
      const int qw0 = w0 * p.mul_w + p.off_w, qh0 = h0 * p.mul_h + p.off_h;
      const int ch0 = c.n_tile * BN + c.rank * (BN / CG);
      if (BOX) {
        // one box per chunk: origin = tap column 0 of the group's filter row
        const int r = c.tap0 / p.kw;
#pragma unroll
        for (int ch = 0; ch < B_CHUNKS; ++ch) {
          const uint32_t dst = b_dst + ch * WG_BOX_CHUNK;
          if (CG == 2) tma_load_4d_2sm(dst, tmQ, full, ch0 + ch * 64, qw0, qh0 + r * p.step_h, n0);
          else tma_load_4d(dst, tmQ, full, ch0 + ch * 64, qw0, qh0 + r * p.step_h, n0);
        }
      }
#pragma unroll
      for (int g = 0; g < (BOX ? 0 : G); ++g) {
        if (g < c.n_taps) {
          const int tap = c.tap0 + g;
          const int r = tap / p.kw, s = tap - r * p.kw;
          const int qw = qw0 + s * p.step_w, qh = qh0 + r * p.step_h;
#pragma unroll
          for (int ch = 0; ch < B_CHUNKS; ++ch) {
            const uint32_t dst = b_dst + (g * B_CHUNKS + ch) * WG_CHUNK;
            if (CG == 2) tma_load_4d_2sm(dst, tmQ, full, ch0 + ch * 64, qw, qh, n0);
            else tma_load_4d(dst, tmQ, full, ch0 + ch * 64, qw, qh, n0);
          }
        }
      }
    }
    wg_produce_round<CG, BN, G, BOX, S, STAGE + 1>(tmP, tmQ, p, c, kb);
  }
}

template <int CG, int BN, int G, bool BOX, int S, int STAGE>
__device__ __forceinline__ void wg_mma_round(WgCtx& c, int kb) {
  constexpr int B_CHUNKS = BN / CG / 64;
  constexpr uint32_t STAGE_BYTES = WgStage<CG, BN, G, BOX>::BYTES;
  constexpr uint32_t IDESC = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) |
                             (static_cast<uint32_t>(BN >> 3) << 17) |
                             (static_cast<uint32_t>((WG_BM * CG) >> 4) << 24);
  if constexpr (STAGE < S) {
    if (kb + STAGE < c.iters) {
      mbar_wait(c.bar_full + 8 * STAGE, (c.phase_bits >> STAGE) & 1);
      c.phase_bits ^= 1u << STAGE;
      tc_fence_after();
      constexpr uint32_t a_off = SMEM_RING + STAGE * STAGE_BYTES;
      constexpr uint32_t b_off = a_off + 2 * WG_CHUNK;
      // the taps of the group share the P slab of this stage: one accumulator of BN columns per tap
#pragma unroll
      for (int g = 0; g < G; ++g) {
        if (g < c.n_taps && !(c.dbg & 2)) {
#pragma unroll
          for (int k = 0; k < WG_BK / 16; ++k) {
            // 16 pixels = two 8-row groups of 1024 B; chunks of 64 channels are WG_CHUNK apart.
            const uint64_t ad = const_smem_desc(a_off + k * 2048, WG_CHUNK, 1024);
            // box mode: tap g starts g pixel rows into the box, 8-pixel K groups are one box row apart
            const uint64_t bd = BOX ? const_smem_desc(b_off + g * 128 + k * 2 * WG_BOX_PITCH * 128, WG_BOX_CHUNK,
                                                      WG_BOX_PITCH * 128)
                                    : const_smem_desc(b_off + g * B_CHUNKS * WG_CHUNK + k * 2048, WG_CHUNK, 1024);
            const uint32_t acc = (STAGE > 0 || k > 0) ? 1u : (kb != 0 ? 1u : 0u);
            if (CG == 2) umma_bf16_2sm(c.tmem_base + g * BN, ad, bd, IDESC, acc);
            else umma_bf16(c.tmem_base + g * BN, ad, bd, IDESC, acc);
          }
        }
      }
      if (CG == 2) umma_commit_2sm(c.bar_empty + 8 * STAGE);
      else umma_commit(c.bar_empty + 8 * STAGE);
    }
    wg_mma_round<CG, BN, G, BOX, S, STAGE + 1>(c, kb);
  }
}

__device__ long long g_wg_flag;   // < 0: shared-memory ring not at SMEM_RING (checked by the launcher)

// CG = 2: a CTA pair owns two neighbouring 128-row tiles of M (output channels of P) and one N tile; each
// CTA stages its own 128 x 64-pixel slab of P and HALF of the Q slab, the leader issues one cta_group::2
// M = 256 MMA for both (as in igemm_fwd.cu).  Per CTA and 64-pixel block that is 16 + BN/2 * 128 B instead of
// 16 KiB + BN * 128 B through L2 and shared memory: the 256 x 256 weight gradient moves 32 KiB per 512 MMA
// cycles instead of 48.
//
// G > 1: a work item covers G consecutive filter taps.  Their GEMMs share the P operand (dy of the un-shifted
// pixel block) and differ in the shifted Q slab, so a stage holds one P slab + G Q slabs and feeds G
// accumulators of BN columns in TMEM: the P slab crosses L2 -> SM once per G taps.  Used for the narrow
// layers (BN <= 128), which are bound by that stream and not by the tensor pipe.
template <int CG, int BN, int G, bool BOX, int S>
__global__ void __launch_bounds__(192, 1)
igemm_wgrad_kernel(const __grid_constant__ CUtensorMap tmP, const __grid_constant__ CUtensorMap tmQ,
                   const WgradParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_full[S];
  __shared__ __align__(8) uint64_t bar_empty[S];
  __shared__ __align__(8) uint64_t bar_tmem_full;
  __shared__ uint32_t tmem_base_slot;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int rank = (CG == 2) ? (int)cluster_ctarank() : 0;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  if ((smem_base & 0x3FFFFu) != SMEM_RING) {
    if (threadIdx.x == 0) g_wg_flag = -(long long)(smem_base & 0x3FFFFu);
    return;
  }

  // ---- work item (the same for both CTAs of a pair, except the M tile)
  int item = blockIdx.x / CG;
  const int split = item % p.splits; item /= p.splits;
  const int n_tile = item % p.n_tiles; item /= p.n_tiles;
  const int m_groups = p.m_tiles / CG;
  const int m_tile = (item % m_groups) * CG + rank; item /= m_groups;
  const int tap0 = item * G;
  const int n_taps = min(G, p.kh * p.kw - tap0);
  const int total_kt = p.tiles_w * p.tiles_h * p.tiles_n;
  const int per = (total_kt + p.splits - 1) / p.splits;
  const int kt_begin = split * per;
  const int kt_end = min(total_kt, kt_begin + per);
  if (kt_begin >= kt_end) return;  // uniform across the CTA (and the pair)
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
  if (warp == 1) {
    if (CG == 2) tmem_alloc_2sm(smem_u32(&tmem_base_slot), p.tmem_cols);
    else tmem_alloc(smem_u32(&tmem_base_slot), p.tmem_cols);
  }
  tc_fence_before();
  __syncthreads();
  if (CG == 2) cluster_sync_all();     // the peer's barriers exist before anything signals them
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_slot;
  griddep_wait();        // programmatic dependent launch: the set-up above overlapped the previous kernel's tail
  griddep_launch();

  WgCtx c;
  c.bar_full = smem_u32(&bar_full[0]); c.bar_empty = smem_u32(&bar_empty[0]);
  c.smem_base = smem_base; c.tmem_base = tmem_base; c.phase_bits = 0;
  c.iters = iters; c.kt_begin = kt_begin; c.m_tile = m_tile; c.n_tile = n_tile; c.tap0 = tap0; c.n_taps = n_taps; c.rank = rank; c.dbg = p.dbg;
  if (warp == 0) {
    if (lane == 0)
      for (int kb = 0; kb < iters; kb += S) wg_produce_round<CG, BN, G, BOX, S, 0>(&tmP, &tmQ, p, c, kb);
    __syncwarp();
  } else if (warp == 1) {
    if (lane == 0 && rank == 0) {
      for (int kb = 0; kb < iters; kb += S) wg_mma_round<CG, BN, G, BOX, S, 0>(c, kb);
      if (CG == 2) umma_commit_2sm(smem_u32(&bar_tmem_full));
      else umma_commit(smem_u32(&bar_tmem_full));
    }
    __syncwarp();
  } else {
    const int q = warp & 3;
    const int m = m_tile * WG_BM + q * 32 + lane;
    const bool row_ok = m < p.M_valid;
    const int n_base = n_tile * BN;
    mbar_wait(smem_u32(&bar_tmem_full), 0);
    tc_fence_after();
    for (int g = 0; g < n_taps; ++g) {
      float* orow = p.out + (long long)split * p.out_split_stride + (long long)(tap0 + g) * p.out_tap_stride +
                    (long long)m * p.out_m_stride;
      for (int c0 = 0; c0 < BN; c0 += 16) {
        if (n_base + c0 >= p.N_valid) break;
        uint32_t v[16];
        tmem_ld16(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + g * BN + c0, v);
        tmem_ld_wait();
        if (row_ok && !(p.dbg & 4)) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int n = n_base + c0 + j * 4;
            if (n < p.N_valid)
              st_v4(orow + n, __uint_as_float(v[j * 4 + 0]), __uint_as_float(v[j * 4 + 1]),
                    __uint_as_float(v[j * 4 + 2]), __uint_as_float(v[j * 4 + 3]));
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (CG == 2) cluster_sync_all();     // the leader's MMAs read the peer's smem / write its TMEM
  if (warp == 1) {
    if (CG == 2) tmem_dealloc_2sm(tmem_base, p.tmem_cols);
    else tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

template <int CG, int BN, int G, bool BOX = false>
static cudaError_t launch_wgrad_cfg(const CUtensorMap& tmP, const CUtensorMap& tmQ,
                                    const WgradParams& p, cudaStream_t stream) {
  constexpr int kStageBytes = WgStage<CG, BN, G, BOX>::BYTES;
  constexpr int kFit = (200 * 1024) / kStageBytes;
  constexpr int kStages = kFit > 8 ? 8 : kFit;
  constexpr size_t kSmem = (size_t)kStages * kStageBytes + 1024;
  auto* kern = igemm_wgrad_kernel<CG, BN, G, BOX, kStages>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem);
    if (e != cudaSuccess) return e;
    attr_set = true;
  }
  const long long grid = (long long)p.tap_groups * p.m_tiles * p.n_tiles * p.splits;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3((unsigned)grid, 1, 1); cfg.blockDim = dim3(192, 1, 1);
  cfg.dynamicSmemBytes = kSmem; cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CG; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  static int pdl = -1;
  if (pdl < 0) { const char* e = getenv("MT_PDL"); pdl = e ? (atoi(e) != 0) : 1; }
  cfg.attrs = attr; cfg.numAttrs = pdl ? 2 : 1;
  return cudaLaunchKernelEx(&cfg, kern, tmP, tmQ, p);
}

int launch_wgrad(const CUtensorMap& tmP, const CUtensorMap& tmQ, const WgradParams& p_in,
                 cudaStream_t stream) {
  WgradParams p = p_in;
  static int dbg = -1;
  if (dbg < 0) { const char* e = getenv("MT_IGEMM_DBG"); dbg = e ? atoi(e) : 0; }
  p.dbg = dbg;
  const long long grid = (long long)p.tap_groups * p.m_tiles * p.n_tiles * p.splits;
  if (grid <= 0 || grid > 0x7fffffffLL) return -1000;
  if (p.cg == 2 && (p.m_tiles % 2 || p.BN < 128)) return -1010;
  cudaError_t e;
  if (p.tap_group * p.tap_groups < p.kh * p.kw || p.tap_group * p.BN > 512) return -1010;
  if (p.box) {
    if (p.tap_group != 3 || p.TW != 8 || p.TH != 8 || p.TN != 1 || p.kw != 3) return -1010;
    switch (p.cg * 1000 + p.BN) {
      case 2128: e = launch_wgrad_cfg<2, 128, 3, true>(tmP, tmQ, p, stream); break;
      case 1128: e = launch_wgrad_cfg<1, 128, 3, true>(tmP, tmQ, p, stream); break;
      case 1064: e = launch_wgrad_cfg<1, 64, 3, true>(tmP, tmQ, p, stream); break;
      default: return -1010;
    }
  } else
  switch (p.tap_group * 10000 + p.cg * 1000 + p.BN) {
    case 12256: e = launch_wgrad_cfg<2, 256, 1>(tmP, tmQ, p, stream); break;
    case 12128: e = launch_wgrad_cfg<2, 128, 1>(tmP, tmQ, p, stream); break;
    case 11256: e = launch_wgrad_cfg<1, 256, 1>(tmP, tmQ, p, stream); break;
    case 11128: e = launch_wgrad_cfg<1, 128, 1>(tmP, tmQ, p, stream); break;
    case 11064: e = launch_wgrad_cfg<1, 64, 1>(tmP, tmQ, p, stream); break;
    case 32128: e = launch_wgrad_cfg<2, 128, 3>(tmP, tmQ, p, stream); break;
    case 31128: e = launch_wgrad_cfg<1, 128, 3>(tmP, tmQ, p, stream); break;
    case 31064: e = launch_wgrad_cfg<1, 64, 3>(tmP, tmQ, p, stream); break;
    case 41064: e = launch_wgrad_cfg<1, 64, 4>(tmP, tmQ, p, stream); break;
    default: return -1010;
  }
  if (e != cudaSuccess) return -(int)e;
  static bool ring_checked = false;
  if (!ring_checked) {
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(stream, &cs) == cudaSuccess && cs == cudaStreamCaptureStatusNone) {
      long long flag = 0;
      if (cudaStreamSynchronize(stream) != cudaSuccess) return -1011;
      if (cudaMemcpyFromSymbol(&flag, g_wg_flag, sizeof(flag)) != cudaSuccess) return -1011;
      if (flag < 0) return -1011;
      ring_checked = true;
    }
  }
  return 0;
}

}  // namespace mt
