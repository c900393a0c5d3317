// Host side of the convolution entry points of include/mt_b200.h: builds the TMA tensor maps
// and the tap maps that turn conv forward / data-grad / transposed-conv phases into launches of
// the one implicit-GEMM kernel (igemm_fwd.cu), and weight-grad into igemm_wgrad.cu.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cudaTypedefs.h>
#include <stdlib.h>
#include <string.h>
#include <map>
#include <mutex>
#include <utility>

#include "../../include/mt_b200.h"
#include "igemm.cuh"
#include "ew_common.cuh"

namespace mt {

// ------------------------------------------------------------------ tensor maps
static PFN_cuTensorMapEncodeTiled_v12000 g_encode = nullptr;
static std::once_flag g_encode_once;

static void load_encode() {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) ==
          cudaSuccess &&
      qres == cudaDriverEntryPointSuccess)
    g_encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
}

int make_tmap(CUtensorMap* m, const void* base, int rank, const uint64_t* dims,
              const uint64_t* strides_bytes, const uint32_t* box, const uint32_t* estr,
              bool swizzle128) {
  std::call_once(g_encode_once, load_encode);
  if (!g_encode) return -1001;
  CUresult r = g_encode(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank,
                        const_cast<void*>(base), dims, strides_bytes, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE,
                        swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : -(2000 + (int)r);
}

// Generic description of one launch of the conv-shaped GEMM.
struct ConvLike {
  const void* a;            // activation operand, [An][Ah][Aw][Ac_s]
  int An, Ah, Aw, Ac, Ac_s; // logical/stored channels of A (K extent = Ac)
  int a_mul;                // TMA traversal stride on h, w (1 or 2)
  int a_mul_w;              // > 0: traversal stride on w only (paired-pixel data gradient), else a_mul
  bool window;              // row-window stem mode
  int win_stride;           // pixel stride baked into the window map
  int win_cols;             // number of valid output columns in window mode
  const void* b;            // packed weights [taps_total][Bn][Bk_s]
  int taps_total, Bn, Bk, Bk_s;
  int Wt, Ht, Nt;           // pixel space
  int kh, kw, off_h, off_w, step_h, step_w, w_base, w_step_r, w_step_s;
  int n_store, n_bias, bias_wrap;
  long long out_n_stride, out_h_stride, out_w_stride, out_off;
  int act; float slope; const float* bias; void* out;
  float* stats;             // optional fused plane statistics [Nt][stat_parts][n_store][2]
  int stat_parts, stat_part0;
  int accumulate;           // out += result
  long long bias_n_stride;  // > 0: per-image bias rows (needs one image per tile)
  int rowk_kw, rowk_c, rowk_win;   // rows-in-K data gradient (IgemmParams::rowk_*), 0: off
  float* rowk_out;
};

static int run_conv_like(const ConvLike& c, cudaStream_t stream) {
  IgemmParams p;
  memset(&p, 0, sizeof(p));
  const int mul = c.window ? 1 : (c.a_mul_w > 0 ? c.a_mul_w : c.a_mul);
  p.BN = pick_bn(c.Bn);
  static int force_bn = -1;
  if (force_bn < 0) { const char* e = getenv("MT_IGEMM_BN"); force_bn = e ? atoi(e) : 0; }
  if (force_bn > 0 && force_bn < p.BN) p.BN = force_bn;
  p.n_tiles = cdiv(c.Bn, p.BN);
  pick_tile(128, c.Wt, c.Ht, c.Nt, c.a_mul_w > c.a_mul ? c.a_mul_w : c.a_mul, &p.TW, &p.TH, &p.TN);
  p.tiles_w = cdiv(c.Wt, p.TW); p.tiles_h = cdiv(c.Ht, p.TH); p.tiles_n = cdiv(c.Nt, p.TN);
  p.Wt = c.Wt; p.Ht = c.Ht; p.Nt = c.Nt;
  p.kh = c.kh; p.kw = c.kw;
  p.mul_h = c.a_mul; p.mul_w = mul;
  p.off_h = c.off_h; p.off_w = c.off_w; p.step_h = c.step_h; p.step_w = c.step_w;
  p.w_base = c.w_base; p.w_step_r = c.w_step_r; p.w_step_s = c.w_step_s;
  p.k_chunks = c.window ? 1 : cdiv(c.Ac, 64);
  p.n_store = c.n_store; p.n_bias = c.bias ? c.n_bias : 0; p.bias_wrap = c.bias_wrap;
  // CTA pairs (one M = 256 MMA per two pixel tiles, half of the weight tile staged per CTA) whenever
  // the Cout tile is wide enough for the operand fetch to be the limiter and there is work for them.
  const int pix_tiles = p.tiles_w * p.tiles_h * p.tiles_n;
  static int force_cg = -1;
  if (force_cg < 0) { const char* e = getenv("MT_IGEMM_CG"); force_cg = e ? atoi(e) : 0; }
  p.cg = (p.BN >= 128 && pix_tiles >= 8) ? 2 : 1;
  if (force_cg == 1 || (force_cg == 2 && p.BN >= 128)) p.cg = force_cg;
  p.work_items = cdiv(pix_tiles, p.cg) * p.n_tiles;
  p.stages = 0;   // fixed per (cg, BN) instantiation in igemm_fwd.cu
  p.acc_cols = p.BN < 32 ? 32 : p.BN;
  p.tmem_cols = pow2_cols(2 * p.acc_cols);
  p.out_n_stride = c.out_n_stride; p.out_h_stride = c.out_h_stride;
  p.out_w_stride = c.out_w_stride; p.out_off = c.out_off;
  p.act = c.act; p.slope = c.slope; p.bias = c.bias; p.bias_n_stride = c.bias ? c.bias_n_stride : 0;
  p.accumulate = c.accumulate;
  p.out = reinterpret_cast<__nv_bfloat16*>(c.out);
  p.rowk_kw = c.rowk_kw; p.rowk_c = c.rowk_c; p.rowk_win = c.rowk_win; p.rowk_out = c.rowk_out;
  if (p.rowk_kw > 0 && (p.BN != 32 || p.cg != 1 || p.rowk_kw * p.rowk_c > 24)) return -1014;

  // Halo variant (igemm_fwd.cu: igemm_halo_kernel) for narrow stride-1 tap grids: one activation box per
  // K chunk serves all taps, the weight operand stays resident.  MT_HALO=0 keeps the tap-by-tap kernel.
  bool halo = false;
  {
    static int halo_off = -1;
    if (halo_off < 0) { const char* e = getenv("MT_HALO"); halo_off = (e && atoi(e) == 0) ? 1 : 0; }
    const bool unit_steps = (c.kh == 1 || c.step_h == 1 || c.step_h == -1) &&
                            (c.kw == 1 || c.step_w == 1 || c.step_w == -1);
    // (row-window stems qualify too: a "pixel" of their tensor map is a window start, the taps are filter rows)
    if (!halo_off && c.a_mul == 1 && c.a_mul_w <= 1 && unit_steps && p.n_tiles == 1 &&
        p.rowk_kw == 0 && c.Wt >= 8 && c.Ht >= 16 && igemm_halo_supported(2, p.BN, c.kh, c.kw, p.k_chunks)) {
      const int tw = cdiv(c.Wt, 8), th = cdiv(c.Ht, 16);
      // fused statistics: the caller sized its slots for the default tiling (mt_conv_stats_parts)
      if (tw * th * c.Nt >= 8 && (!c.stats || (p.TN == 1 && tw * th == p.tiles_w * p.tiles_h))) halo = true;
    }
  }
  // stride-2 3x3 variant (igemm_halo_s2_kernel): parity-class boxes; K = 64, 128-column CTA pairs, weights resident
  bool halo_s2 = false;
  {
    static int off2 = -1;
    if (off2 < 0) { const char* e = getenv("MT_HALO_S2"); off2 = (e && atoi(e) == 0) ? 1 : 0; }
    if (!halo && !off2 && !c.window && c.a_mul == 2 && c.a_mul_w <= 0 && c.kh == 3 && c.kw == 3 && c.step_h == 1 &&
        c.step_w == 1 && p.k_chunks == 1 && p.BN == 128 && p.n_tiles == 1 && p.rowk_kw == 0 && c.Wt >= 8 &&
        c.Ht >= 16) {
      const int tw = cdiv(c.Wt, 8), th = cdiv(c.Ht, 16);
      if (tw * th * c.Nt >= 8 && (!c.stats || (p.TN == 1 && tw * th == p.tiles_w * p.tiles_h))) halo_s2 = true;
    }
  }
  if (halo_s2) {
    p.TW = 8; p.TH = 16; p.TN = 1;
    p.tiles_w = cdiv(c.Wt, 8); p.tiles_h = cdiv(c.Ht, 16); p.tiles_n = c.Nt;
    p.cg = 2;
    p.work_items = cdiv(p.tiles_w * p.tiles_h * p.tiles_n, 2);
  }
  if (halo) {
    p.TW = 8; p.TH = 16; p.TN = 1;
    p.tiles_w = cdiv(c.Wt, 8); p.tiles_h = cdiv(c.Ht, 16); p.tiles_n = c.Nt;
    p.cg = 2;
    p.work_items = cdiv(p.tiles_w * p.tiles_h * p.tiles_n, 2);
    // box origin = the smallest tap offset; taps are walked in increasing offset order, the weight tap index
    // follows through the (possibly negated) affine map
    const bool neg_h = c.kh > 1 && c.step_h < 0, neg_w = c.kw > 1 && c.step_w < 0;
    p.off_h = c.off_h - (neg_h ? c.kh - 1 : 0);
    p.off_w = c.off_w - (neg_w ? c.kw - 1 : 0);
    p.w_base = c.w_base + (neg_h ? (c.kh - 1) * c.w_step_r : 0) + (neg_w ? (c.kw - 1) * c.w_step_s : 0);
    p.w_step_r = neg_h ? -c.w_step_r : c.w_step_r;
    p.w_step_s = neg_w ? -c.w_step_s : c.w_step_s;
  }

  CUtensorMap tmA, tmB;
  int rc;
  if (halo_s2) {
    uint64_t dims[4] = {(uint64_t)c.Ac, (uint64_t)c.Aw, (uint64_t)c.Ah, (uint64_t)c.An};
    uint64_t str[3] = {(uint64_t)c.Ac_s * 2, (uint64_t)c.Aw * c.Ac_s * 2,
                       (uint64_t)c.Ah * c.Aw * c.Ac_s * 2};
    uint32_t box[4] = {64, 9 * 2, 17 * 2, 1};        // 9 x 17 pixels of one parity class (element strides 2)
    uint32_t es[4] = {1, 2, 2, 1};
    rc = make_tmap(&tmA, c.a, 4, dims, str, box, es);
  } else if (halo) {
    uint64_t dims[4] = {(uint64_t)c.Ac, (uint64_t)c.Aw, (uint64_t)c.Ah, (uint64_t)c.An};
    uint64_t str[3] = {(uint64_t)c.Ac_s * 2, (uint64_t)c.Aw * c.Ac_s * 2,
                       (uint64_t)c.Ah * c.Aw * c.Ac_s * 2};
    if (c.window) {                    // dim0: 64-element window starting at pixel (w * win_stride)
      dims[0] = 64; dims[1] = (uint64_t)c.win_cols;
      str[0] = (uint64_t)c.win_stride * c.Ac_s * 2;
    }
    uint32_t box[4] = {64, (uint32_t)(8 + c.kw - 1), (uint32_t)(16 + c.kh - 1), 1};
    uint32_t es[4] = {1, 1, 1, 1};
    rc = make_tmap(&tmA, c.a, 4, dims, str, box, es);
  } else if (!c.window) {
    uint64_t dims[4] = {(uint64_t)c.Ac, (uint64_t)c.Aw, (uint64_t)c.Ah, (uint64_t)c.An};
    uint64_t str[3] = {(uint64_t)c.Ac_s * 2, (uint64_t)c.Aw * c.Ac_s * 2,
                       (uint64_t)c.Ah * c.Aw * c.Ac_s * 2};
    uint32_t box[4] = {64, (uint32_t)(p.TW * mul), (uint32_t)(p.TH * c.a_mul), (uint32_t)p.TN};
    uint32_t es[4] = {1, (uint32_t)mul, (uint32_t)c.a_mul, 1};
    rc = make_tmap(&tmA, c.a, 4, dims, str, box, es);
  } else {
    // dim0: 64-element window starting at pixel (w*win_stride); dim1: output column.
    uint64_t dims[4] = {64, (uint64_t)c.win_cols, (uint64_t)c.Ah, (uint64_t)c.An};
    uint64_t str[3] = {(uint64_t)c.win_stride * c.Ac_s * 2, (uint64_t)c.Aw * c.Ac_s * 2,
                       (uint64_t)c.Ah * c.Aw * c.Ac_s * 2};
    uint32_t box[4] = {64, (uint32_t)p.TW, (uint32_t)(p.TH * c.a_mul), (uint32_t)p.TN};
    uint32_t es[4] = {1, 1, (uint32_t)c.a_mul, 1};
    rc = make_tmap(&tmA, c.a, 4, dims, str, box, es);
  }
  if (rc) return rc;
  {
    uint64_t dims[3] = {(uint64_t)c.Bk, (uint64_t)c.Bn, (uint64_t)c.taps_total};
    uint64_t str[2] = {(uint64_t)c.Bk_s * 2, (uint64_t)c.Bn * c.Bk_s * 2};
    uint32_t box[3] = {64, (uint32_t)(p.BN / p.cg), 1};   // a CTA of a pair stages half of the tile
    uint32_t es[3] = {1, 1, 1};
    rc = make_tmap(&tmB, c.b, 3, dims, str, box, es);
  }
  if (rc) return rc;
  CUtensorMap tmC;
  {
    // Output box of the epilogue's TMA store: the VALID pixel range of this launch as a strided
    // view (interior of a padded buffer, one sub-pixel phase, ...), so that the store clips tile
    // overhang by itself.  64-channel slabs use the 128B swizzle, narrower tiles dense rows.
    const int sw = p.BN < 64 ? p.BN : 64;
    uint64_t dims[4] = {(uint64_t)c.n_store, (uint64_t)c.Wt, (uint64_t)c.Ht, (uint64_t)c.Nt};
    uint64_t str[3] = {(uint64_t)c.out_w_stride * 2, (uint64_t)c.out_h_stride * 2,
                       (uint64_t)c.out_n_stride * 2};
    uint32_t box[4] = {(uint32_t)sw, (uint32_t)p.TW, (uint32_t)p.TH, (uint32_t)p.TN};
    uint32_t es[4] = {1, 1, 1, 1};
    rc = make_tmap(&tmC, reinterpret_cast<const __nv_bfloat16*>(c.out) + c.out_off, 4, dims, str, box,
                   es, sw == 64);
  }
  if (rc) return rc;
  // fused statistics need one image per tile and the 64-column slab path
  if (c.stats) {
    // the caller sized the buffer with mt_conv_stats_parts(): same tile choice, one image per tile
    if (p.TN != 1 || p.BN < 64 || c.stat_part0 + p.tiles_w * p.tiles_h > c.stat_parts) return -1012;
    p.stats = c.stats; p.n_stat_c = c.n_store;
    p.stat_parts = c.stat_parts; p.stat_part0 = c.stat_part0;
  }
  if (p.bias_n_stride > 0 && p.TN != 1) return -1014;     // a tile would span images
  if (halo_s2) return launch_igemm_halo_s2(tmA, tmB, tmC, p, stream);
  return halo ? launch_igemm_halo(tmA, tmB, tmC, p, stream) : launch_igemm(tmA, tmB, tmC, p, stream);
}

// ---- rows-in-K data gradient of a many-tap stride-1 conv into <= 4 channels (mt_conv2d_dgrad_rowk)
// fp32 OIHW master -> bf16 [kh][32][Cout_s]: row (s * Cin + c) of tap r holds W[:, c, r, s]
__global__ void pack_rowk_kernel(const float* __restrict__ w, int Cout, int Cin, int kh, int kw, int Cout_s,
                                 __nv_bfloat16* __restrict__ out) {
  pdl_wait(); pdl_launch();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= kh * 32 * Cout_s) return;
  const int co = i % Cout_s, row = (i / Cout_s) % 32, r = i / (Cout_s * 32);
  float v = 0.f;
  if (co < Cout && row < kw * Cin) {
    const int s = row / Cin, c = row - s * Cin;
    v = w[(((long long)co * Cin + c) * kh + r) * kw + s];
  }
  out[i] = __float2bfloat16(v);
}

// dx[n, h, w, 0..7] (bf16) = sum of the tile slots that cover column w, lowest tile first
__global__ void rowk_gather_kernel(const float* __restrict__ part, __nv_bfloat16* __restrict__ dx, long long pixels,
                                   int Win, int tiles_w, int TW, int ow) {
  pdl_wait(); pdl_launch();
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= pixels) return;
  const int w = (int)(i % Win);
  const long long row = i / Win;                               // n * Hin + h
  float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
  int t = (w - (ow - 1) + TW - 1) / TW;                        // first tile with w - t * TW <= ow - 1
  if (w - (ow - 1) < 0) t = 0;
  for (; t < tiles_w && t * TW <= w; ++t) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(part + ((row * tiles_w + t) * ow + (w - t * TW)) * 4));
    a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
  }
  bf16x8 o;
  o.v[0] = __floats2bfloat162_rn(a.x, a.y);
  o.v[1] = __floats2bfloat162_rn(a.z, a.w);
  o.v[2] = __floats2bfloat162_rn(0.f, 0.f);
  o.v[3] = o.v[2];
  *reinterpret_cast<bf16x8*>(dx + i * 8) = o;
}

static int check_desc(const MtConv2d* d) {
  if (!d) return -1002;
  if (d->Cin_s % 8 || d->Cout_s % 8 || d->Cin > d->Cin_s || d->Cout > d->Cout_s) return -1003;
  if (d->stride != 1 && d->stride != 2) return -1004;
  if (d->N <= 0 || d->kh <= 0 || d->kw <= 0) return -1005;
  return 0;
}

}  // namespace mt

using namespace mt;

extern "C" int mt_stored_channels(int c) { return round_up(c, 8); }
// Row-window mode is for image stems only (Cin_s <= 16): their buffers come from mt_image_to_nhwc
// callers that append zeroed slack, because a 64-element window over-reads the last pixels.
extern "C" int mt_conv_window_mode(int Cin_s, int kw) {
  return (Cin_s <= 16 && Cin_s * kw <= 64) ? 1 : 0;
}

// ConvTranspose2d(k=3, s=2, p=1[, op=1]) forward runs in the merged-parity mode (see
// mt_conv_transpose2d_fwd_stats): its forward operand is packed by mt_pack_convT3x3s2_weight.
extern "C" int mt_convT_merged_mode(int kh, int kw, int zpad) {
  static int off = -1;
  if (off < 0) { const char* e = getenv("MT_CONVT_MERGED"); off = (e && atoi(e) == 0) ? 1 : 0; }
  return (!off && kh == 3 && kw == 3 && zpad == 1) ? 1 : 0;
}

extern "C" int mt_conv2d_fwd(const MtConv2d* d, const void* x, const void* w_fwd,
                             const float* bias, void* y, void* stream) {
  return mt_conv2d_fwd_stats(d, x, w_fwd, bias, y, nullptr, 0, stream);
}

// Number of per-tile partial-sum slots per image that the fused statistics of this conv produce
// (0: the shape cannot fuse them — several images per tile, or a Cout tile narrower than a slab).
// Row-window stems whose window also covers the NEXT output pixel ((kw + stride) * Cin_s <= 64) compute
// two adjacent output pixels per GEMM row: columns (pixel parity, Cout), window stride 2 * stride.  The
// 64-element windows are the operand stream that bounds these layers (7 x 128 B per output pixel for
// the 7x7 stem): pairing halves it.  w_fwd is then packed by mt_pack_conv_window_pair_weight.
extern "C" int mt_conv_window_pair_mode(int Cin_s, int kw, int stride) {
  static int off = -1;
  if (off < 0) { const char* e = getenv("MT_STEM_PAIR"); off = (e && atoi(e) == 0) ? 1 : 0; }
  return (!off && mt_conv_window_mode(Cin_s, kw) && (kw + stride) * Cin_s <= 64) ? 1 : 0;
}

extern "C" int mt_conv_stats_parts(const MtConv2d* d, int transposed) {
  if (check_desc(d)) return 0;
  static int off = -1;
  if (off < 0) { const char* e = getenv("MT_CONV_NO_STATS"); off = e ? atoi(e) : 0; }
  if (off || pick_bn(d->Cout) < 64) return 0;
  int tw, th, tn;
  if (transposed) {
    if ((d->Hout & 1) || (d->Wout & 1)) return 0;
    pick_tile(128, d->Wout / 2, d->Hout / 2, d->N, 1, &tw, &th, &tn);
    return tn == 1 ? 4 * cdiv(d->Wout / 2, tw) * cdiv(d->Hout / 2, th) : 0;
  }
  if (d->window && !d->single && mt_conv_window_pair_mode(d->Cin_s, d->kw, d->stride) && !(d->Wout & 1)) {
    // paired stem: tiles over Wout / 2 pixel pairs, each slot 2 * Cout_s wide = two of the caller's parts
    pick_tile(128, d->Wout / 2, d->Hout, d->N, d->stride, &tw, &th, &tn);
    return tn == 1 ? 2 * cdiv(d->Wout / 2, tw) * cdiv(d->Hout, th) : 0;
  }
  pick_tile(128, d->Wout, d->Hout, d->N, d->stride, &tw, &th, &tn);
  return tn == 1 ? cdiv(d->Wout, tw) * cdiv(d->Hout, th) : 0;
}

static int conv2d_fwd_impl(const MtConv2d* d, const void* x, const void* w_fwd, const float* bias,
                           long long bias_n_stride, void* y, float* stats, int parts, void* stream);
extern "C" int mt_conv2d_fwd_stats(const MtConv2d* d, const void* x, const void* w_fwd,
                                   const float* bias, void* y, float* stats, int parts,
                                   void* stream) {
  return conv2d_fwd_impl(d, x, w_fwd, bias, 0, y, stats, parts, stream);
}
// 1 if the pixel tile the library picks for this forward launch holds pixels of ONE image (what the per-image
// bias of mt_conv2d_fwd_image_bias needs)
extern "C" int mt_conv2d_image_bias_ok(const MtConv2d* d) {
  if (check_desc(d) || d->window) return 0;
  int tw, th, tn;
  pick_tile(128, d->Wout, d->Hout, d->N, d->stride, &tw, &th, &tn);
  return tn == 1 ? 1 : 0;
}
// per-image bias: bias_n[N][Cout] fp32 (see the header)
extern "C" int mt_conv2d_fwd_image_bias(const MtConv2d* d, const void* x, const void* w_fwd,
                                        const float* bias_n, void* y, void* stream) {
  if (!bias_n) return -1002;
  if (d && d->window) return -1014;
  return conv2d_fwd_impl(d, x, w_fwd, bias_n, d ? d->Cout : 0, y, nullptr, 0, stream);
}
static int conv2d_fwd_impl(const MtConv2d* d, const void* x, const void* w_fwd, const float* bias,
                           long long bias_n_stride, void* y, float* stats, int parts, void* stream) {
  int rc = check_desc(d);
  if (rc) return rc;
  ConvLike c;
  memset(&c, 0, sizeof(c));
  if (d->window && !(mt_conv_window_mode(d->Cin_s, d->kw) && d->zpad == 0)) return -1009;
  const bool win = d->window != 0;
  c.a = x; c.An = d->N; c.Ah = d->Hin; c.Aw = d->Win; c.Ac = d->Cin; c.Ac_s = d->Cin_s;
  c.a_mul = d->stride; c.window = win; c.win_stride = d->stride; c.win_cols = d->Wout;
  c.b = w_fwd;
  c.Wt = d->Wout; c.Ht = d->Hout; c.Nt = d->N;
  if (win) {
    c.taps_total = d->kh; c.Bn = d->Cout; c.Bk = 64; c.Bk_s = 64;
    c.kh = d->kh; c.kw = 1; c.off_h = 0; c.off_w = 0; c.step_h = 1; c.step_w = 0;
    c.w_base = 0; c.w_step_r = 1; c.w_step_s = 0;
  } else {
    c.taps_total = d->kh * d->kw; c.Bn = d->Cout; c.Bk = d->Cin; c.Bk_s = d->Cin_s;
    c.kh = d->kh; c.kw = d->kw; c.off_h = -d->zpad; c.off_w = -d->zpad;
    c.step_h = 1; c.step_w = 1; c.w_base = 0; c.w_step_r = d->kw; c.w_step_s = 1;
  }
  c.n_store = d->Cout_s; c.n_bias = d->Cout;
  c.out_w_stride = d->Cout_s;
  c.out_h_stride = (long long)d->out_W * d->Cout_s;
  c.out_n_stride = (long long)d->out_H * d->out_W * d->Cout_s;
  c.out_off = ((long long)d->out_pad * d->out_W + d->out_pad) * d->Cout_s;
  c.act = d->act; c.slope = d->slope; c.bias = bias; c.out = y;
  c.bias_n_stride = bias_n_stride;
  c.stats = stats; c.stat_parts = parts; c.stat_part0 = 0;
  if (win && !d->single && mt_conv_window_pair_mode(d->Cin_s, d->kw, d->stride)) {
    if (d->Wout & 1) return -1014;
    if (stats && (parts & 1)) return -1012;
    c.Bn = 2 * d->Cout_s; c.n_store = 2 * d->Cout_s; c.bias_wrap = d->Cout_s;
    c.Wt = d->Wout / 2; c.win_cols = d->Wout / 2; c.win_stride = 2 * d->stride;
    c.out_w_stride = 2 * d->Cout_s;
    c.stat_parts = parts / 2;
  }
  return run_conv_like(c, (cudaStream_t)stream);
}

// Sub-pixel phase launcher shared by stride-2 dgrad and ConvTranspose2d forward:
//   out[2a+ph, 2b+pw] = sum_{r == (ph+zp) mod 2, s == (pw+zp) mod 2} A[a + (ph+zp-r)/2, ...] W[r,s]
// The four phases are independent (disjoint output pixels).  Since one phase often cannot fill the GPU
// (small maps: discriminator / style-encoder tails, a persistent launch of 4-64 CTAs on 148 SMs)
// they run side by side on three auxiliary streams forked from and joined back into the caller's
// stream with events — plain stream semantics for the caller, graph edges under capture.
struct PhaseFork {
  cudaStream_t aux[3];
  cudaEvent_t fork, join[3];
  bool ok = false;
};
// One set of auxiliary streams / events per (device, caller stream), created at the first forked call of
// that stream (these are the only CUDA objects the library creates; nothing is ever allocated on the device):
// two caller streams — e.g. the two discriminator updates running side by side — never share fork / join
// events, so neither gets a false dependency on the other.  MT_PHASE_FORK = 0 disables forking; a value
// N > 0 forks only launches of at most N CTAs per phase.
static std::mutex g_fork_mu;
static std::map<std::pair<int, cudaStream_t>, PhaseFork> g_forks;
static int g_fork_max_ctas = -1;
static PhaseFork* phase_fork(cudaStream_t stream) {
  std::lock_guard<std::mutex> lk(g_fork_mu);
  if (g_fork_max_ctas < 0) {
    const char* e = getenv("MT_PHASE_FORK");
    g_fork_max_ctas = e ? atoi(e) : (1 << 30);
    if (g_fork_max_ctas < 0) g_fork_max_ctas = 0;
  }
  if (g_fork_max_ctas == 0) return nullptr;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return nullptr;
  auto key = std::make_pair(dev, stream);
  auto it = g_forks.find(key);
  if (it != g_forks.end()) return it->second.ok ? &it->second : nullptr;
  PhaseFork f;
  bool ok = cudaEventCreateWithFlags(&f.fork, cudaEventDisableTiming) == cudaSuccess;
  for (int i = 0; i < 3 && ok; ++i)
    ok = cudaStreamCreateWithFlags(&f.aux[i], cudaStreamNonBlocking) == cudaSuccess &&
         cudaEventCreateWithFlags(&f.join[i], cudaEventDisableTiming) == cudaSuccess;
  f.ok = ok;
  auto& slot = g_forks[key] = f;
  return ok ? &slot : nullptr;
}
static int phase_work_items(const ConvLike& q) {
  int tw, th, tn;
  pick_tile(128, q.Wt, q.Ht, q.Nt, q.a_mul, &tw, &th, &tn);
  const int BN = pick_bn(q.Bn);
  const int pix = cdiv(q.Wt, tw) * cdiv(q.Ht, th) * cdiv(q.Nt, tn);
  const int cg = (BN >= 128 && pix >= 8) ? 2 : 1;
  return cdiv(pix, cg) * cdiv(q.Bn, BN) * cg;          // CTAs the persistent launch would use
}

static int run_phases(ConvLike c, int kh, int kw, int zp, int OutH, int OutW, long long row_pitch,
                      long long pix_pitch, long long base_off, cudaStream_t stream) {
  ConvLike qs[4];
  int n = 0, max_ctas = 0;
  for (int ph = 0; ph < 2; ++ph)
    for (int pw = 0; pw < 2; ++pw) {
      const int r0 = (ph + zp) & 1, s0 = (pw + zp) & 1;
      const int nr = (kh - r0 + 1) / 2, ns = (kw - s0 + 1) / 2;
      const int Ht = (OutH - ph + 1) / 2, Wt = (OutW - pw + 1) / 2;
      if (Ht <= 0 || Wt <= 0) continue;
      if (nr <= 0 || ns <= 0) return -1006;  // phase without taps: caller must not use k<2
      ConvLike q = c;
      q.Ht = Ht; q.Wt = Wt;
      q.kh = nr; q.kw = ns;
      q.off_h = (ph + zp - r0) / 2; q.off_w = (pw + zp - s0) / 2;
      q.step_h = -1; q.step_w = -1;
      q.w_base = r0 * kw + s0; q.w_step_r = 2 * kw; q.w_step_s = 2;
      q.out_h_stride = 2 * row_pitch; q.out_w_stride = 2 * pix_pitch;
      q.out_off = base_off + ph * row_pitch + pw * pix_pitch;
      q.stat_part0 = (ph * 2 + pw) * (c.stat_parts / 4);
      const int ctas = phase_work_items(q);
      if (ctas > max_ctas) max_ctas = ctas;
      qs[n++] = q;
    }
  PhaseFork* pf = n > 1 ? phase_fork(stream) : nullptr;
  const bool fork = pf != nullptr && max_ctas <= g_fork_max_ctas;
  if (!fork) {
    for (int i = 0; i < n; ++i) {
      int rc = run_conv_like(qs[i], stream);
      if (rc) return rc;
    }
    return 0;
  }
  PhaseFork& g_fork = *pf;
  if (cudaEventRecord(g_fork.fork, stream) != cudaSuccess) return -1013;
  int rc = 0;
  for (int i = 1; i < n && !rc; ++i) {
    if (cudaStreamWaitEvent(g_fork.aux[i - 1], g_fork.fork, 0) != cudaSuccess) rc = -1013;
    if (!rc) rc = run_conv_like(qs[i], g_fork.aux[i - 1]);
    // the join is recorded even after a failed launch so that a capture is never left forked
    if (cudaEventRecord(g_fork.join[i - 1], g_fork.aux[i - 1]) != cudaSuccess && !rc) rc = -1013;
  }
  const int rc0 = run_conv_like(qs[0], stream);
  for (int i = 1; i < n; ++i)
    if (cudaStreamWaitEvent(stream, g_fork.join[i - 1], 0) != cudaSuccess && !rc) rc = -1013;
  return rc ? rc : rc0;
}

// the two row-parity launches of a merged-parity operation, side by side (qs[0] on an auxiliary stream)
static int run_pair(const ConvLike* qs, cudaStream_t st) {
  PhaseFork* pf = phase_fork(st);
  if (!pf) {
    const int rc = run_conv_like(qs[0], st);
    return rc ? rc : run_conv_like(qs[1], st);
  }
  PhaseFork& g_fork = *pf;
  if (cudaEventRecord(g_fork.fork, st) != cudaSuccess) return -1013;
  int rc = 0;
  if (cudaStreamWaitEvent(g_fork.aux[0], g_fork.fork, 0) != cudaSuccess) rc = -1013;
  if (!rc) rc = run_conv_like(qs[0], g_fork.aux[0]);
  // the join is recorded even after a failed launch so that a capture is never left forked
  if (cudaEventRecord(g_fork.join[0], g_fork.aux[0]) != cudaSuccess && !rc) rc = -1013;
  const int rc1 = run_conv_like(qs[1], st);
  if (cudaStreamWaitEvent(st, g_fork.join[0], 0) != cudaSuccess && !rc) rc = -1013;
  return rc ? rc : rc1;
}

// Data gradient of a stride-2, un-padded (reflection-padded input) 3x3 / 4x4 convolution in the
// merged-parity mode: see mt_conv2d_dgrad.  Its w_bwd operand is packed by mt_pack_conv_s2_dgrad_weight.
extern "C" int mt_conv_s2_dgrad_merged_mode(int kh, int kw, int stride, int zpad, int Cin) {
  static int off = -1;
  if (off < 0) { const char* e = getenv("MT_DGRAD_MERGED"); off = (e && atoi(e) == 0) ? 1 : 0; }
  // k = 3 pays 12/9 of the MMA work for the wider tile: only where the layer is operand-bound (<= 128
  // input channels, i.e. a Cout tile below 256 in the single-parity form); k = 4 has no zero blocks
  return (!off && stride == 2 && zpad == 0 && kh == kw && (kh == 4 || (kh == 3 && Cin <= 128))) ? 1 : 0;
}

// Data gradient into <= 8 channels (the image: stems with reflection-padded input, stride 1): the
// single-pixel form uses 8 of the 16 columns of the narrowest MMA and fetches one A tile per filter
// tap.  Paired: columns (pixel parity, ci) = 16, kw + 1 column offsets per filter row, dy traversed
// with stride 2 on w: kh*(kw+1) A tiles per pixel PAIR instead of 2*kh*kw.
extern "C" int mt_conv_s1_dgrad_pair_mode(int kh, int kw, int stride, int zpad, int Cin_s) {
  static int off = -1;
  if (off < 0) { const char* e = getenv("MT_DGRAD_PAIR"); off = (e && atoi(e) == 0) ? 1 : 0; }
  return (!off && stride == 1 && zpad == 0 && kh == kw && Cin_s == 8) ? 1 : 0;
}

// Data gradient of a stride-1, un-padded, many-tap conv into <= 4 channels (the image gradient of the 7x7
// stems, networks.py:30) with the filter ROWS in K and the filter COLUMNS in N: one launch of the implicit
// GEMM over (output row h', dy column w) with kh k-blocks (dy rows h' - r) and columns (s, c); its epilogue
// adds the kw shifted columns (epilogue_rowk).  dy crosses L2 -> SM kh times instead of kh * (kw + 1) / 2.
extern "C" int mt_conv_rowk_dgrad_mode(int kh, int kw, int stride, int zpad, int Cin, int Cout_s) {
  static int off = -1;
  if (off < 0) { const char* e = getenv("MT_DGRAD_ROWK"); off = (e && atoi(e) == 0) ? 1 : 0; }
  return (!off && stride == 1 && zpad == 0 && Cin <= 4 && kw * Cin <= 24 && kh * kw >= 16 && Cout_s % 8 == 0) ? 1 : 0;
}
extern "C" long long mt_packed_conv_rowk_elems(int Cout, int kh) { return (long long)kh * 32 * round_up(Cout, 8); }
extern "C" int mt_pack_conv_rowk_weight(const float* w, int Cout, int Cin, int kh, int kw, void* w_rowk,
                                        void* stream) {
  if (kw * Cin > 24) return -1005;
  const int Cout_s = round_up(Cout, 8), n = kh * 32 * Cout_s;
  cudaError_t e = launch_pdl(pack_rowk_kernel, dim3(cdiv(n, 256)), dim3(256), 0, (cudaStream_t)stream, w, Cout, Cin,
                             kh, kw, Cout_s, reinterpret_cast<__nv_bfloat16*>(w_rowk));
  return e == cudaSuccess ? 0 : -(int)e;
}
static void rowk_tile(const MtConv2d* d, int* TW, int* tiles_w) {
  int TH, TN;
  pick_tile(128, d->Wout, d->Hin, d->N, 1, TW, &TH, &TN);      // the tile run_conv_like picks for this launch
  *tiles_w = cdiv(d->Wout, *TW);
}
extern "C" long long mt_conv_rowk_scratch_elems(const MtConv2d* d) {
  if (!d) return 0;
  int TW, tiles_w;
  rowk_tile(d, &TW, &tiles_w);
  return (long long)d->N * d->Hin * tiles_w * (TW + d->kw - 1) * 4;
}
extern "C" int mt_conv2d_dgrad_rowk(const MtConv2d* d, const void* dy, const void* w_rowk, float* scratch,
                                    void* dx, void* stream) {
  int rc = check_desc(d);
  if (rc) return rc;
  if (!dy || !w_rowk || !scratch || !dx) return -1002;
  if (!mt_conv_rowk_dgrad_mode(d->kh, d->kw, d->stride, d->zpad, d->Cin, d->Cout_s) || d->Cin_s != 8 ||
      d->Hin != d->Hout + d->kh - 1 || d->Win != d->Wout + d->kw - 1)
    return -1014;
  ConvLike c;
  memset(&c, 0, sizeof(c));
  c.a = dy; c.An = d->N; c.Ah = d->Hout; c.Aw = d->Wout; c.Ac = d->Cout; c.Ac_s = d->Cout_s;
  c.a_mul = 1; c.window = false;
  c.b = w_rowk; c.taps_total = d->kh; c.Bn = 32; c.Bk = d->Cout; c.Bk_s = d->Cout_s;
  c.Wt = d->Wout; c.Ht = d->Hin; c.Nt = d->N;
  c.kh = d->kh; c.kw = 1; c.off_h = 0; c.off_w = 0; c.step_h = -1; c.step_w = 0;
  c.w_base = 0; c.w_step_r = 1; c.w_step_s = 0;
  c.n_store = 8; c.n_bias = 0; c.bias = nullptr; c.act = MT_ACT_NONE;
  c.out = dx;                          // only to form a valid (unused) output map
  c.out_w_stride = 8; c.out_h_stride = (long long)d->Win * 8; c.out_n_stride = (long long)d->Hin * d->Win * 8;
  c.rowk_kw = d->kw; c.rowk_c = d->Cin; c.rowk_win = d->Win; c.rowk_out = scratch;
  rc = run_conv_like(c, (cudaStream_t)stream);
  if (rc) return rc;
  int TW, tiles_w;
  rowk_tile(d, &TW, &tiles_w);
  const long long pixels = (long long)d->N * d->Hin * d->Win;
  cudaError_t e = launch_pdl(rowk_gather_kernel, dim3((unsigned)((pixels + 255) / 256)), dim3(256), 0,
                             (cudaStream_t)stream, (const float*)scratch, reinterpret_cast<__nv_bfloat16*>(dx),
                             pixels, d->Win, tiles_w, TW, TW + d->kw - 1);
  return e == cudaSuccess ? 0 : -(int)e;
}

extern "C" int mt_conv2d_dgrad(const MtConv2d* d, const void* dy, const void* w_bwd, void* dx,
                               void* stream) {
  return mt_conv2d_dgrad_acc(d, dy, w_bwd, dx, 0, stream);
}

extern "C" int mt_conv2d_dgrad_acc(const MtConv2d* d, const void* dy, const void* w_bwd, void* dx,
                                   int accumulate, void* stream) {
  int rc = check_desc(d);
  if (rc) return rc;
  ConvLike c;
  memset(&c, 0, sizeof(c));
  c.accumulate = accumulate ? 1 : 0;
  // A = dy [N, Hout, Wout, Cout_s] (dense: the producer of dy writes an unpadded buffer)
  c.a = dy; c.An = d->N; c.Ah = d->Hout; c.Aw = d->Wout; c.Ac = d->Cout; c.Ac_s = d->Cout_s;
  c.a_mul = 1; c.window = false;
  c.b = w_bwd; c.taps_total = d->kh * d->kw; c.Bn = d->Cin; c.Bk = d->Cout; c.Bk_s = d->Cout_s;
  c.Nt = d->N;
  c.n_store = d->Cin_s; c.n_bias = 0; c.bias = nullptr; c.act = MT_ACT_NONE; c.out = dx;
  const long long pix = d->Cin_s, row = (long long)d->Win * d->Cin_s;
  c.out_n_stride = (long long)d->Hin * row;
  if (!d->single && mt_conv_s1_dgrad_pair_mode(d->kh, d->kw, d->stride, d->zpad, d->Cin_s)) {
    // dx[h, 2b+j, ci] = sum_r sum_t dy[h - r, 2b + 1 - t, :] . Wm[r][t][(j, ci), :],  t = 0..kw,
    // filter column s = j - 1 + t (zero block outside 0..kw-1): mt_pack_conv_s1_dgrad_pair_weight
    if (d->Win & 1) return -1014;
    c.taps_total = d->kh * (d->kw + 1); c.Bn = 2 * d->Cin_s;
    c.a_mul_w = 2;
    c.Wt = d->Win / 2; c.Ht = d->Hin;
    c.kh = d->kh; c.kw = d->kw + 1; c.off_h = 0; c.off_w = 1; c.step_h = -1; c.step_w = -1;
    c.w_base = 0; c.w_step_r = d->kw + 1; c.w_step_s = 1;
    c.n_store = 2 * d->Cin_s;
    c.out_h_stride = row; c.out_w_stride = 2 * pix; c.out_off = 0;
    return run_conv_like(c, (cudaStream_t)stream);
  }
  if (d->stride == 1) {
    c.Wt = d->Win; c.Ht = d->Hin;
    c.kh = d->kh; c.kw = d->kw; c.off_h = d->zpad; c.off_w = d->zpad; c.step_h = -1; c.step_w = -1;
    c.w_base = 0; c.w_step_r = d->kw; c.w_step_s = 1;
    c.out_h_stride = row; c.out_w_stride = pix; c.out_off = 0;
    return run_conv_like(c, (cudaStream_t)stream);
  }
  if (!d->single && mt_conv_s2_dgrad_merged_mode(d->kh, d->kw, d->stride, d->zpad, d->Cin)) {
    // dx[2a+ph, 2b+pw, ci] = sum_{dh,dw} dy[a-dh, b-dw, :] . W[:, ci, ph+2dh, pw+2dw]   (taps beyond the
    // filter are zero blocks, dy rows / columns outside the map read as zero).  As for the transposed
    // convolution forward, the two column parities are GEMM columns (pw, ci): 2*Cin_s contiguous
    // channels of the pixel pair (2b, 2b+1); one launch per row parity.  Needs even extents.
    if ((d->Hin & 1) || (d->Win & 1)) return -1014;
    const int k = d->kh, nr0 = (k + 1) / 2, nr1 = k / 2;
    ConvLike qs[2];
    for (int ph = 0; ph < 2; ++ph) {
      ConvLike q = c;
      q.taps_total = 2 * (nr0 + nr1); q.Bn = 2 * d->Cin_s;
      q.Wt = d->Win / 2; q.Ht = d->Hin / 2;
      q.kh = ph ? nr1 : nr0; q.kw = 2; q.off_h = 0; q.off_w = 0; q.step_h = -1; q.step_w = -1;
      q.w_base = ph ? 2 * nr0 : 0; q.w_step_r = 2; q.w_step_s = 1;
      q.n_store = 2 * d->Cin_s;
      q.out_w_stride = 2 * pix; q.out_h_stride = 2 * row; q.out_off = ph * row;
      qs[ph] = q;
    }
    return run_pair(qs, (cudaStream_t)stream);
  }
  return run_phases(c, d->kh, d->kw, d->zpad, d->Hin, d->Win, row, pix, 0, (cudaStream_t)stream);
}

extern "C" int mt_conv_transpose2d_fwd(const MtConv2d* d, const void* x, const void* w_fwd,
                                       const float* bias, void* y, void* stream) {
  return mt_conv_transpose2d_fwd_stats(d, x, w_fwd, bias, y, nullptr, 0, stream);
}

extern "C" int mt_conv_transpose2d_fwd_stats(const MtConv2d* d, const void* x, const void* w_fwd,
                                             const float* bias, void* y, float* stats, int parts,
                                             void* stream) {
  int rc = check_desc(d);
  if (rc) return rc;
  if (d->stride != 2) return -1004;
  ConvLike c;
  memset(&c, 0, sizeof(c));
  c.a = x; c.An = d->N; c.Ah = d->Hin; c.Aw = d->Win; c.Ac = d->Cin; c.Ac_s = d->Cin_s;
  c.a_mul = 1; c.window = false;
  c.b = w_fwd; c.taps_total = d->kh * d->kw; c.Bn = d->Cout; c.Bk = d->Cin; c.Bk_s = d->Cin_s;
  c.Nt = d->N;
  c.n_store = d->Cout_s; c.n_bias = d->Cout; c.bias = bias; c.act = d->act; c.slope = d->slope;
  c.out = y;
  const long long pix = d->Cout_s, row = (long long)d->out_W * d->Cout_s;
  c.out_n_stride = (long long)d->out_H * row;
  const long long base = ((long long)d->out_pad * d->out_W + d->out_pad) * d->Cout_s;
  if (stats && (parts <= 0 || parts % 4)) return -1012;
  if (mt_convT_merged_mode(d->kh, d->kw, d->zpad) && d->Hout == 2 * d->Hin && d->Wout == 2 * d->Win) {
    // ConvTranspose2d(k3, s2, p1, op1) as TWO stride-1 launches over the INPUT pixel grid, one per
    // output row parity ph, each producing both column parities at once: GEMM columns = (pw, co),
    // 2*Cout_s wide, which are contiguous in memory (pixels 2w and 2w+1 of output row 2h+ph).
    //   out[2h+ph, 2w+pw, co] = sum_{dh,dw in {0,1}} x[h+dh, w+dw, :] . Wm[ph][dh][dw][(pw,co), :]
    // with zero blocks where a (parity, offset) pair has no filter tap (mt_pack_convT3x3s2_weight) and
    // x read as zero beyond the bottom / right edge (TMA out-of-bounds fill).  Against the four
    // single-parity launches: each A tile is fetched 6 times instead of 9 and the Cout tile is twice
    // as wide (these layers are bound by the L2 -> SM operand stream, not by the tensor pipe), for
    // 12/9 of the MMA work.
    ConvLike qs[2];
    for (int ph = 0; ph < 2; ++ph) {
      ConvLike q = c;
      q.taps_total = 6; q.Bn = 2 * d->Cout_s;
      q.Wt = d->Win; q.Ht = d->Hin;
      q.kh = ph ? 2 : 1; q.kw = 2; q.off_h = 0; q.off_w = 0; q.step_h = 1; q.step_w = 1;
      q.w_base = ph ? 2 : 0; q.w_step_r = 2; q.w_step_s = 1;
      q.n_store = 2 * d->Cout_s; q.bias_wrap = d->Cout_s;
      q.out_w_stride = 2 * pix; q.out_h_stride = 2 * row; q.out_off = base + ph * row;
      // statistics slots are 2*Cout_s wide here: slot s covers the caller's parts 2s and 2s+1
      q.stats = stats; q.stat_parts = parts / 2; q.stat_part0 = ph * (parts / 4);
      qs[ph] = q;
    }
    return run_pair(qs, (cudaStream_t)stream);
  }
  c.stats = stats; c.stat_parts = parts; c.stat_part0 = 0;   // run_phases: phase i fills quarter i
  return run_phases(c, d->kh, d->kw, d->zpad, d->Hout, d->Wout, row, pix, base, (cudaStream_t)stream);
}

extern "C" int mt_conv_transpose2d_dgrad(const MtConv2d* d, const void* dy, const void* w_bwd,
                                         void* dx, void* stream) {
  int rc = check_desc(d);
  if (rc) return rc;
  // dx[hi] = sum_r dy[2*hi - zpad + r] W[ci, co, r]  == stride-2 zero-padded conv over dy.
  ConvLike c;
  memset(&c, 0, sizeof(c));
  c.a = dy; c.An = d->N; c.Ah = d->Hout; c.Aw = d->Wout; c.Ac = d->Cout; c.Ac_s = d->Cout_s;
  c.a_mul = 2; c.window = false;
  c.b = w_bwd; c.taps_total = d->kh * d->kw; c.Bn = d->Cin; c.Bk = d->Cout; c.Bk_s = d->Cout_s;
  c.Wt = d->Win; c.Ht = d->Hin; c.Nt = d->N;
  c.kh = d->kh; c.kw = d->kw; c.off_h = -d->zpad; c.off_w = -d->zpad; c.step_h = 1; c.step_w = 1;
  c.w_base = 0; c.w_step_r = d->kw; c.w_step_s = 1;
  c.n_store = d->Cin_s; c.n_bias = 0; c.act = MT_ACT_NONE; c.out = dx;
  c.out_w_stride = d->Cin_s; c.out_h_stride = (long long)d->Win * d->Cin_s;
  c.out_n_stride = (long long)d->Hin * d->Win * d->Cin_s; c.out_off = 0;
  return run_conv_like(c, (cudaStream_t)stream);
}
