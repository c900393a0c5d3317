// Weight-gradient entry points and the fp32-master <-> bf16-operand weight (un)packers.
#include <cuda.h>
#include <cuda_runtime.h>
#include <string.h>

#include "../../include/mt_b200.h"
#include "igemm.cuh"
#include "ew_common.cuh"

namespace mt {

struct WgradLike {
  const void* p; int Pn, Ph, Pw, Pc, Pc_s;   // un-shifted tensor -> M
  const void* q; int Qn, Qh, Qw, Qc, Qc_s;   // tap-shifted tensor -> N
  int q_mul; bool window; int win_stride, win_cols;
  int kh, kw, off_h, off_w, step_h, step_w;
  float* out; int N_store;                   // row pitch of out (fp32 elements)
};

// Tile / split-K plan of one weight gradient (shared by the launcher and by the *_wgrad_splits queries
// with which the caller sizes dw_packed).
static void plan_wgrad(const WgradLike& c, WgradParams& p) {
  memset(&p, 0, sizeof(p));
  pick_tile(64, c.Pw, c.Ph, c.Pn, c.q_mul, &p.TW, &p.TH, &p.TN);
  p.tiles_w = cdiv(c.Pw, p.TW); p.tiles_h = cdiv(c.Ph, p.TH); p.tiles_n = cdiv(c.Pn, p.TN);
  p.kh = c.kh; p.kw = c.kw;
  p.mul_h = c.q_mul; p.mul_w = c.window ? 1 : c.q_mul;
  p.off_h = c.off_h; p.off_w = c.off_w; p.step_h = c.step_h; p.step_w = c.step_w;
  const int n_logical = c.window ? 64 : c.Qc;
  p.BN = n_logical >= 256 ? 256 : (n_logical > 64 ? 128 : 64);
  // box mode: stride-1 3x3 (the residual trunk, the style encoder's blocks): one filter row per work item, its
  // three taps read one (8 + 2) x 8 pixel box of the shifted operand; Cin tile of at most 128 so that the
  // three accumulators fit in TMEM.  MT_WGRAD_BOX=0 turns it off.
  static int box_off = -1;
  if (box_off < 0) { const char* e = getenv("MT_WGRAD_BOX"); box_off = (e && atoi(e) == 0) ? 1 : 0; }
  p.box = (!box_off && !c.window && c.q_mul == 1 && c.kh == 3 && c.kw == 3 && c.step_w == 1 && c.step_h == 1 &&
           c.Pw >= 8 && c.Ph >= 8) ? 1 : 0;
  if (p.box) {
    p.TW = 8; p.TH = 8; p.TN = 1;
    p.tiles_w = cdiv(c.Pw, 8); p.tiles_h = cdiv(c.Ph, 8); p.tiles_n = c.Pn;
    if (p.BN > 128) p.BN = 128;
  }
  p.m_tiles = cdiv(c.Pc, 128);
  p.n_tiles = cdiv(n_logical, p.BN);
  // narrow 3x3 layers: three taps (one filter row) per work item share the un-shifted operand's slab
  // (igemm_wgrad.cu); MT_WGRAD_TAPS=1 keeps one tap per work item
  static int force_taps = -1;
  if (force_taps < 0) { const char* e = getenv("MT_WGRAD_TAPS"); force_taps = e ? atoi(e) : 0; }
  p.tap_group = (!c.window && p.kh * p.kw == 9 && p.BN <= 128 && (force_taps != 1 || p.box)) ? 3 : 1;
  // 7x7 row-window stems (taps = filter rows, 64-wide windows): four rows per work item (134 -> 118 us; the
  // 4-row stems would be left with a single group and 148 split-K slabs: 38 -> 46 us, not used there)
  if (c.window && p.BN == 64 && p.kh * p.kw >= 6 && force_taps != 1) p.tap_group = 4;
  p.tap_groups = cdiv(p.kh * p.kw, p.tap_group);
  const int total_kt = p.tiles_w * p.tiles_h * p.tiles_n;
  const int items = p.tap_groups * p.m_tiles * p.n_tiles;
  // one wave of CTAs: every split pays a full fp32 red.add epilogue, so prefer long main loops
  int splits = 148 / items;
  const int max_splits = cdiv(total_kt, 8);   // at least 8 pixel tiles per split
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  const int per = cdiv(total_kt, splits);
  p.splits = cdiv(total_kt, per);
  // CTA pairs (one M = 256 MMA for two M tiles, half of the Q slab staged per CTA) wherever there are two
  // M tiles to pair and the Q slab is wide enough to be worth halving; MT_WGRAD_CG=1 keeps single CTAs
  static int force_cg = -1;
  if (force_cg < 0) { const char* e = getenv("MT_WGRAD_CG"); force_cg = e ? atoi(e) : 0; }
  p.cg = (p.m_tiles % 2 == 0 && p.BN >= 128 && force_cg != 1) ? 2 : 1;
  const int stage_bytes = p.box ? 2 * 8192 + (p.BN / p.cg / 64) * 10240 : (2 + p.tap_group * (p.BN / p.cg / 64)) * 8192;
  p.stages = (200 * 1024) / stage_bytes;
  if (p.stages > 8) p.stages = 8;
  p.tmem_cols = pow2_cols(p.tap_group * p.BN);
  p.M_valid = c.Pc;
  p.N_valid = c.N_store;
  p.out_m_stride = c.N_store;
  p.out_tap_stride = (long long)c.Pc * c.N_store;
  p.out_split_stride = (long long)p.kh * p.kw * p.out_tap_stride;
  p.out = c.out;
}

static int run_wgrad_like(const WgradLike& c, int splits, cudaStream_t stream) {
  WgradParams p;
  plan_wgrad(c, p);
  if (splits != p.splits) return -1015;

  CUtensorMap tmP, tmQ;
  {
    uint64_t dims[4] = {(uint64_t)c.Pc, (uint64_t)c.Pw, (uint64_t)c.Ph, (uint64_t)c.Pn};
    uint64_t str[3] = {(uint64_t)c.Pc_s * 2, (uint64_t)c.Pw * c.Pc_s * 2,
                       (uint64_t)c.Ph * c.Pw * c.Pc_s * 2};
    uint32_t box[4] = {64, (uint32_t)p.TW, (uint32_t)p.TH, (uint32_t)p.TN};
    uint32_t es[4] = {1, 1, 1, 1};
    int rc = make_tmap(&tmP, c.p, 4, dims, str, box, es);
    if (rc) return rc;
  }
  if (!c.window) {
    uint64_t dims[4] = {(uint64_t)c.Qc, (uint64_t)c.Qw, (uint64_t)c.Qh, (uint64_t)c.Qn};
    uint64_t str[3] = {(uint64_t)c.Qc_s * 2, (uint64_t)c.Qw * c.Qc_s * 2,
                       (uint64_t)c.Qh * c.Qw * c.Qc_s * 2};
    uint32_t box[4] = {64, (uint32_t)(p.box ? 10 : p.TW * c.q_mul), (uint32_t)(p.TH * c.q_mul), (uint32_t)p.TN};
    uint32_t es[4] = {1, (uint32_t)c.q_mul, (uint32_t)c.q_mul, 1};
    int rc = make_tmap(&tmQ, c.q, 4, dims, str, box, es);
    if (rc) return rc;
  } else {
    uint64_t dims[4] = {64, (uint64_t)c.win_cols, (uint64_t)c.Qh, (uint64_t)c.Qn};
    uint64_t str[3] = {(uint64_t)c.win_stride * c.Qc_s * 2, (uint64_t)c.Qw * c.Qc_s * 2,
                       (uint64_t)c.Qh * c.Qw * c.Qc_s * 2};
    uint32_t box[4] = {64, (uint32_t)p.TW, (uint32_t)(p.TH * c.q_mul), (uint32_t)p.TN};
    uint32_t es[4] = {1, 1, (uint32_t)c.q_mul, 1};
    int rc = make_tmap(&tmQ, c.q, 4, dims, str, box, es);
    if (rc) return rc;
  }
  return launch_wgrad(tmP, tmQ, p, stream);
}

// ------------------------------------------------------------------ pack / unpack
// Source weight element for (n_idx -> out channel of the packed row, k_idx, r, s):
//   Conv2d          W[co][ci][r][s]   ConvTranspose2d  W[ci][co][r][s]
// 32-bit index arithmetic throughout: weight tensors have < 2^31 elements (checked by the
// launchers) and a 64-bit division costs ~5x a 32-bit one — these kernels are nothing but divisions.
__device__ __forceinline__ unsigned w_index(unsigned co, unsigned ci, unsigned r, unsigned s,
                                            unsigned Cout, unsigned Cin, unsigned kh, unsigned kw,
                                            int transposed) {
  return transposed ? ((ci * Cout + co) * kh + r) * kw + s : ((co * Cin + ci) * kh + r) * kw + s;
}

__global__ void pack_weight_kernel(const float* __restrict__ w, int Cout_, int Cin_, int kh_, int kw_,
                                   int transposed, int window, __nv_bfloat16* __restrict__ wf,
                                   __nv_bfloat16* __restrict__ wb) {
  pdl_wait(); pdl_launch();
  const unsigned Cout = Cout_, Cin = Cin_, kh = kh_, kw = kw_;
  const unsigned Cin_s = (Cin + 7) & ~7u, Cout_s = (Cout + 7) & ~7u;
  const unsigned nf = window ? kh * Cout * 64 : kh * kw * Cout * Cin_s;
  const unsigned nb = kh * kw * Cin * Cout_s;
  const unsigned stride = gridDim.x * blockDim.x;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < nf + nb; i += stride) {
    if (i < nf) {
      if (!wf) continue;
      float v = 0.f;
      if (window) {
        const unsigned e = i % 64, t = i / 64;
        const unsigned r = t / Cout, co = t - r * Cout;
        const unsigned s = e / Cin_s, ci = e - s * Cin_s;
        if (s < kw && ci < Cin) v = w[w_index(co, ci, r, s, Cout, Cin, kh, kw, transposed)];
      } else {
        const unsigned t = i / Cin_s, ci = i - t * Cin_s;
        const unsigned tap = t / Cout, co = t - tap * Cout;
        const unsigned r = tap / kw;
        if (ci < Cin) v = w[w_index(co, ci, r, tap - r * kw, Cout, Cin, kh, kw, transposed)];
      }
      wf[i] = __float2bfloat16(v);
    } else {
      if (!wb) continue;
      const unsigned j = i - nf;
      const unsigned t = j / Cout_s, co = j - t * Cout_s;
      const unsigned tap = t / Cin, ci = t - tap * Cin;
      const unsigned r = tap / kw;
      float v = 0.f;
      if (co < Cout) v = w[w_index(co, ci, r, tap - r * kw, Cout, Cin, kh, kw, transposed)];
      wb[j] = __float2bfloat16(v);
    }
  }
}

// Pair-per-thread variants (non-window layouts): a thread owns one (row, column) pair of a packed
// operand and walks the T filter taps, so the index decomposition (two divisions) is paid once per
// T elements, the master weight is read in runs of T consecutive floats and every packed row is
// written with consecutive lanes on consecutive elements.
//   pass 0: wf[tap][co][ci_s]  thread = (co, ci), ci fastest
//   pass 1: wb[tap][ci][co_s]  thread = (ci, co), co fastest
__global__ void __launch_bounds__(256)
pack_weight_pairs_kernel(const float* __restrict__ w, int Cout_, int Cin_, int T_, int transposed,
                         __nv_bfloat16* __restrict__ wf, __nv_bfloat16* __restrict__ wb) {
  pdl_wait(); pdl_launch();
  const unsigned Cout = Cout_, Cin = Cin_, T = T_;
  const unsigned Cin_s = (Cin + 7) & ~7u, Cout_s = (Cout + 7) & ~7u;
  const unsigned n0 = wf ? Cout * Cin_s : 0, n1 = wb ? Cin * Cout_s : 0;
  const unsigned stride = gridDim.x * blockDim.x;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n0 + n1; i += stride) {
    unsigned co, ci;
    __nv_bfloat16* dst;
    unsigned dst_stride;                       // elements between consecutive taps
    if (i < n0) {
      co = i / Cin_s; ci = i - co * Cin_s;
      dst = wf + i; dst_stride = Cout * Cin_s;
    } else {
      const unsigned j = i - n0;
      ci = j / Cout_s; co = j - ci * Cout_s;
      dst = wb + j; dst_stride = Cin * Cout_s;
    }
    const bool valid = co < Cout && ci < Cin;
    const float* src = w + (size_t)(transposed ? ci * Cout + co : co * Cin + ci) * T;
    for (unsigned t = 0; t < T; ++t)
      dst[(size_t)t * dst_stride] = __float2bfloat16(valid ? __ldg(src + t) : 0.f);
  }
}

// dw_packed -> dw (OIHW / IOHW fp32). Conv2d: [tap][Cout][Cin_s] (window: [kh][Cout][64]);
// ConvTranspose2d: [tap][Cin][Cout_s].
// `g` holds `splits` partial results `split_stride` elements apart (split-K slices of the weight-gradient
// kernel), added here in split order.
__global__ void unpack_wgrad_kernel(const float* __restrict__ g, int splits, unsigned split_stride,
                                    int Cout_, int Cin_, int kh_, int kw_,
                                    int transposed, int window, int accumulate,
                                    float* __restrict__ dw) {
  pdl_wait(); pdl_launch();
  const unsigned Cout = Cout_, Cin = Cin_, kh = kh_, kw = kw_;
  const unsigned Cin_s = (Cin + 7) & ~7u, Cout_s = (Cout + 7) & ~7u;
  const unsigned n = Cout * Cin * kh * kw;
  const unsigned stride = gridDim.x * blockDim.x;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    unsigned t = i / kw;
    const unsigned s = i - t * kw;
    const unsigned t2 = t / kh, r = t - t2 * kh;
    unsigned co, ci;
    if (transposed) { ci = t2 / Cout; co = t2 - ci * Cout; }
    else            { co = t2 / Cin;  ci = t2 - co * Cin; }
    unsigned src;
    if (transposed) src = ((r * kw + s) * Cin + ci) * Cout_s + co;
    else if (window) src = (r * Cout + co) * 64 + s * Cin_s + ci;
    else src = ((r * kw + s) * Cout + co) * Cin_s + ci;
    float v = g[src];
    for (int sp = 1; sp < splits; ++sp) v += g[(size_t)sp * split_stride + src];
    dw[i] = accumulate ? dw[i] + v : v;
  }
}

// Forward operand of the merged-parity ConvTranspose2d(k3, s2, p1): wf[tap 0..5][(pw, co)][ci]
// (rows: 2*Cout_s, columns: Cin_s).  Taps 0,1: output row parity 0 with input offsets (0, dw);
// taps 2 + 2*dh + dw: row parity 1 with offsets (dh, dw).  Filter index per (parity, offset):
// (0,0) -> 1, (1,0) -> 2, (1,1) -> 0, (0,1) -> none (zero block).  Source W[ci][co][r][s].
__global__ void __launch_bounds__(256)
pack_convT3x3s2_kernel(const float* __restrict__ w, int Cout_, int Cin_, __nv_bfloat16* __restrict__ wf) {
  pdl_wait(); pdl_launch();
  const unsigned Cout = Cout_, Cin = Cin_;
  const unsigned Cin_s = (Cin + 7) & ~7u, Cout_s = (Cout + 7) & ~7u;
  const unsigned rows = 2 * Cout_s, per_tap = rows * Cin_s, n = 6 * per_tap;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const unsigned tap = i / per_tap, e = i - tap * per_tap;
    const unsigned row = e / Cin_s, ci = e - row * Cin_s;
    const unsigned pw = row / Cout_s, co = row - pw * Cout_s;
    const unsigned ph = tap >= 2, t = ph ? tap - 2 : tap;
    const unsigned dh = ph ? t >> 1 : 0, dw = t & 1;
    const int r = ph == 0 ? 1 : (dh == 0 ? 2 : 0);
    const int sx = pw == 0 ? (dw == 0 ? 1 : -1) : (dw == 0 ? 2 : 0);
    float v = 0.f;
    if (sx >= 0 && co < Cout && ci < Cin) v = __ldg(w + ((size_t)(ci * Cout + co) * 3 + r) * 3 + sx);
    wf[i] = __float2bfloat16(v);
  }
}

// Data-gradient operand of the merged-parity stride-2 convolution (k = 3 or 4, no zero padding):
// wb[tap][(pw, ci)][co]  (rows: 2*Cin_s, columns: Cout_s).  Taps of row parity 0 first
// (dh < (k+1)/2, dw < 2: index 2*dh + dw), then those of row parity 1 (dh < k/2).  Filter element
// (r, s) = (ph + 2*dh, pw + 2*dw), a zero block when s >= k.  Source W[co][ci][r][s].
__global__ void __launch_bounds__(256)
pack_conv_s2_dgrad_kernel(const float* __restrict__ w, int Cout_, int Cin_, int k_,
                          __nv_bfloat16* __restrict__ wb) {
  pdl_wait(); pdl_launch();
  const unsigned Cout = Cout_, Cin = Cin_, k = k_;
  const unsigned Cin_s = (Cin + 7) & ~7u, Cout_s = (Cout + 7) & ~7u;
  const unsigned nr0 = (k + 1) / 2, nr1 = k / 2;
  const unsigned rows = 2 * Cin_s, per_tap = rows * Cout_s, n = 2 * (nr0 + nr1) * per_tap;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const unsigned tap = i / per_tap, e = i - tap * per_tap;
    const unsigned row = e / Cout_s, co = e - row * Cout_s;
    const unsigned pw = row / Cin_s, ci = row - pw * Cin_s;
    const unsigned ph = tap >= 2 * nr0, t = ph ? tap - 2 * nr0 : tap;
    const unsigned r = ph + 2 * (t >> 1), sx = pw + 2 * (t & 1);
    float v = 0.f;
    if (sx < k && co < Cout && ci < Cin) v = __ldg(w + ((size_t)(co * Cin + ci) * k + r) * k + sx);
    wb[i] = __float2bfloat16(v);
  }
}

// Forward operand of a paired row-window stem: wf[r][(j, co)][e], rows 2*Cout_s, 64 window elements
// e = (px, ci) with Cin_s channels per pixel; output pixel j of the pair sees filter column
// s = px - j*stride (zero outside 0..kw-1).  Source W[co][ci][r][s].
__global__ void __launch_bounds__(256)
pack_window_pair_kernel(const float* __restrict__ w, int Cout_, int Cin_, int kh_, int kw_, int stride_,
                        __nv_bfloat16* __restrict__ wf) {
  pdl_wait(); pdl_launch();
  const unsigned Cout = Cout_, Cin = Cin_, kh = kh_, kw = kw_, stride = stride_;
  const unsigned Cin_s = (Cin + 7) & ~7u, Cout_s = (Cout + 7) & ~7u;
  const unsigned n = kh * 2 * Cout_s * 64;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const unsigned e = i & 63, t = i >> 6;
    const unsigned r = t / (2 * Cout_s), row = t - r * 2 * Cout_s;
    const unsigned j = row / Cout_s, co = row - j * Cout_s;
    const unsigned px = e / Cin_s, ci = e - px * Cin_s;
    const int sx = (int)px - (int)(j * stride);
    float v = 0.f;
    if (sx >= 0 && sx < (int)kw && ci < Cin && co < Cout)
      v = __ldg(w + ((size_t)(co * Cin + ci) * kh + r) * kw + sx);
    wf[i] = __float2bfloat16(v);
  }
}

// Data-gradient operand of the paired-pixel stride-1 mode (Cin_s == 8): wb[r*(kw+1) + t][(j, ci)][co],
// 16 rows x Cout_s columns per tap; filter column s = j - 1 + t, zero block outside 0..kw-1.
__global__ void __launch_bounds__(256)
pack_conv_s1_dgrad_pair_kernel(const float* __restrict__ w, int Cout_, int Cin_, int k_,
                               __nv_bfloat16* __restrict__ wb) {
  pdl_wait(); pdl_launch();
  const unsigned Cout = Cout_, Cin = Cin_, k = k_;
  const unsigned Cout_s = (Cout + 7) & ~7u;
  const unsigned per_tap = 16 * Cout_s, n = k * (k + 1) * per_tap;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const unsigned tap = i / per_tap, e = i - tap * per_tap;
    const unsigned row = e / Cout_s, co = e - row * Cout_s;
    const unsigned j = row >> 3, ci = row & 7;
    const unsigned r = tap / (k + 1), t = tap - r * (k + 1);
    const int sx = (int)j - 1 + (int)t;
    float v = 0.f;
    if (sx >= 0 && sx < (int)k && ci < Cin && co < Cout)
      v = __ldg(w + ((size_t)(co * Cin + ci) * k + r) * k + sx);
    wb[i] = __float2bfloat16(v);
  }
}

// dW[co][ci][r][s] (+)= g[r][co][s*Cin_s + ci] + g[r][Cout_s + co][(s + stride)*Cin_s + ci]
__global__ void unpack_window_pair_wgrad_kernel(const float* __restrict__ g, int splits, unsigned split_stride,
                                                int Cout_, int Cin_, int kh_,
                                                int kw_, int stride_, int accumulate,
                                                float* __restrict__ dw) {
  pdl_wait(); pdl_launch();
  const unsigned Cout = Cout_, Cin = Cin_, kh = kh_, kw = kw_, stride = stride_;
  const unsigned Cin_s = (Cin + 7) & ~7u, Cout_s = (Cout + 7) & ~7u;
  const unsigned n = Cout * Cin * kh * kw;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    unsigned t = i / kw;
    const unsigned s = i - t * kw;
    const unsigned t2 = t / kh, r = t - t2 * kh;
    const unsigned co = t2 / Cin, ci = t2 - co * Cin;
    const size_t s0 = ((size_t)r * 2 * Cout_s + co) * 64 + s * Cin_s + ci;
    const size_t s1 = ((size_t)r * 2 * Cout_s + Cout_s + co) * 64 + (s + stride) * Cin_s + ci;
    float v = 0.f;
    for (int sp = 0; sp < splits; ++sp) v += g[(size_t)sp * split_stride + s0] + g[(size_t)sp * split_stride + s1];
    dw[i] = accumulate ? dw[i] + v : v;
  }
}

}  // namespace mt

using namespace mt;

extern "C" long long mt_packed_conv_s1_dgrad_pair_elems(int Cout, int k) {
  return (long long)k * (k + 1) * 16 * round_up(Cout, 8);
}
extern "C" int mt_pack_conv_s1_dgrad_pair_weight(const float* w, int Cout, int Cin, int k, void* w_bwd,
                                                 void* stream) {
  const long long n = mt_packed_conv_s1_dgrad_pair_elems(Cout, k);
  if (n >= (1ll << 31) || Cin > 8) return -1005;
  int blocks = (int)((n + 255) / 256);
  if (blocks > 148 * 16) blocks = 148 * 16;
  launch_pdl(pack_conv_s1_dgrad_pair_kernel, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, w, Cout, Cin, k,
                                                                           (__nv_bfloat16*)w_bwd);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : -(int)e;
}

extern "C" int mt_unpack_conv_window_pair_wgrad(const float* dw_packed, int splits, int Cout, int Cin, int kh,
                                                int kw, int stride, int accumulate, float* dw, void* stream) {
  const long long n = (long long)Cout * Cin * kh * kw;
  const long long per = mt_packed_conv_window_pair_elems(Cout, kh);
  if (n >= (1ll << 31) || per >= (1ll << 31) || splits < 1) return -1005;
  int blocks = (int)((n + 255) / 256);
  if (blocks > 148 * 16) blocks = 148 * 16;
  launch_plain(unpack_window_pair_wgrad_kernel, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, dw_packed, splits, (unsigned)per,
                                                                            Cout, Cin, kh, kw, stride,
                                                                            accumulate, dw);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : -(int)e;
}

extern "C" long long mt_packed_conv_window_pair_elems(int Cout, int kh) {
  return (long long)kh * 2 * round_up(Cout, 8) * 64;
}
extern "C" int mt_pack_conv_window_pair_weight(const float* w, int Cout, int Cin, int kh, int kw,
                                               int stride, void* w_fwd, void* stream) {
  const long long n = mt_packed_conv_window_pair_elems(Cout, kh);
  if (n >= (1ll << 31)) return -1005;
  int blocks = (int)((n + 255) / 256);
  if (blocks > 148 * 16) blocks = 148 * 16;
  launch_pdl(pack_window_pair_kernel, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, w, Cout, Cin, kh, kw, stride,
                                                                    (__nv_bfloat16*)w_fwd);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : -(int)e;
}

extern "C" long long mt_packed_conv_s2_dgrad_elems(int Cout, int Cin, int k) {
  return 2ll * ((k + 1) / 2 + k / 2) * 2 * round_up(Cin, 8) * round_up(Cout, 8);
}
extern "C" int mt_pack_conv_s2_dgrad_weight(const float* w, int Cout, int Cin, int k, void* w_bwd,
                                            void* stream) {
  const long long n = mt_packed_conv_s2_dgrad_elems(Cout, Cin, k);
  if (n >= (1ll << 31) || (k != 3 && k != 4)) return -1005;
  int blocks = (int)((n + 255) / 256);
  if (blocks > 148 * 16) blocks = 148 * 16;
  launch_pdl(pack_conv_s2_dgrad_kernel, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, w, Cout, Cin, k, (__nv_bfloat16*)w_bwd);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : -(int)e;
}

extern "C" long long mt_packed_convT3x3s2_elems(int Cout, int Cin) {
  return 6ll * 2 * round_up(Cout, 8) * round_up(Cin, 8);
}
extern "C" int mt_pack_convT3x3s2_weight(const float* w, int Cout, int Cin, void* w_fwd, void* stream) {
  const long long n = mt_packed_convT3x3s2_elems(Cout, Cin);
  if (n >= (1ll << 31)) return -1005;
  int blocks = (int)((n + 255) / 256);
  if (blocks > 148 * 16) blocks = 148 * 16;
  launch_pdl(pack_convT3x3s2_kernel, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, w, Cout, Cin, (__nv_bfloat16*)w_fwd);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : -(int)e;
}

extern "C" long long mt_packed_weight_elems(int Cout, int Cin, int kh, int kw, int window,
                                            int bwd) {
  const int Cin_s = round_up(Cin, 8), Cout_s = round_up(Cout, 8);
  if (bwd) return (long long)kh * kw * Cin * Cout_s;
  return window ? (long long)kh * Cout * 64 : (long long)kh * kw * Cout * Cin_s;
}

extern "C" int mt_pack_conv_weight(const float* w, int Cout, int Cin, int kh, int kw,
                                   int transposed, int window, void* w_fwd, void* w_bwd,
                                   void* stream) {
  const long long n = mt_packed_weight_elems(Cout, Cin, kh, kw, window, 0) +
                      mt_packed_weight_elems(Cout, Cin, kh, kw, window, 1);
  if (n >= (1ll << 31)) return -1005;
  if (!window) {
    const long long pairs = (long long)(w_fwd ? Cout * round_up(Cin, 8) : 0) +
                            (long long)(w_bwd ? Cin * round_up(Cout, 8) : 0);
    int pb = (int)((pairs + 255) / 256);
    if (pb > 148 * 16) pb = 148 * 16;
    if (pb < 1) pb = 1;
    launch_pdl(pack_weight_pairs_kernel, dim3(pb), dim3(256), 0, (cudaStream_t)stream, 
        w, Cout, Cin, kh * kw, transposed, (__nv_bfloat16*)w_fwd, (__nv_bfloat16*)w_bwd);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : -(int)e;
  }
  int blocks = (int)((n + 255) / 256);
  if (blocks > 148 * 16) blocks = 148 * 16;
  launch_pdl(pack_weight_kernel, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, 
      w, Cout, Cin, kh, kw, transposed, window, (__nv_bfloat16*)w_fwd, (__nv_bfloat16*)w_bwd);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : -(int)e;
}

extern "C" long long mt_packed_wgrad_elems(int Cout, int Cin, int kh, int kw, int transposed, int window) {
  const int Cin_s = round_up(Cin, 8), Cout_s = round_up(Cout, 8);
  if (transposed) return (long long)kh * kw * Cin * Cout_s;
  return window ? (long long)kh * Cout * 64 : (long long)kh * kw * Cout * Cin_s;
}

extern "C" int mt_unpack_conv_wgrad(const float* dw_packed, int splits, int Cout, int Cin, int kh, int kw,
                                    int transposed, int window, int accumulate, float* dw,
                                    void* stream) {
  const long long n = (long long)Cout * Cin * kh * kw;
  const long long per = mt_packed_wgrad_elems(Cout, Cin, kh, kw, transposed, window);
  if (n >= (1ll << 31) || per >= (1ll << 31) || splits < 1) return -1005;
  int blocks = (int)((n + 255) / 256);
  if (blocks > 148 * 16) blocks = 148 * 16;
  launch_plain(unpack_wgrad_kernel, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, dw_packed, splits, (unsigned)per, Cout, Cin,
                                                                kh, kw, transposed, window, accumulate, dw);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : -(int)e;
}

static int describe_conv_wgrad(const MtConv2d* d, const void* x, const void* dy, float* dw_packed,
                               WgradLike& c) {
  if (!d || d->Cin_s % 8 || d->Cout_s % 8) return -1003;
  memset(&c, 0, sizeof(c));
  if (d->window && !(mt_conv_window_mode(d->Cin_s, d->kw) && d->zpad == 0)) return -1009;
  const bool win = d->window != 0;
  c.p = dy; c.Pn = d->N; c.Ph = d->Hout; c.Pw = d->Wout; c.Pc = d->Cout; c.Pc_s = d->Cout_s;
  c.q = x; c.Qn = d->N; c.Qh = d->Hin; c.Qw = d->Win; c.Qc = d->Cin; c.Qc_s = d->Cin_s;
  c.q_mul = d->stride; c.window = win; c.win_stride = d->stride; c.win_cols = d->Wout;
  if (win) {
    c.kh = d->kh; c.kw = 1; c.off_h = 0; c.off_w = 0; c.step_h = 1; c.step_w = 0;
    c.N_store = 64;
    if (!d->single && mt_conv_window_pair_mode(d->Cin_s, d->kw, d->stride)) {
      // paired stem (see mt_conv_window_pair_mode): dy read as [N, Hout, Wout/2, 2*Cout_s] — the same
      // memory — against windows of stride 2*stride: M = (pixel parity, Cout) fills the 128-row MMA
      // that a 64-channel stem left half empty, and half as many window rows are fetched.
      // dw_packed is then [kh][2*Cout_s][64]; mt_unpack_conv_window_pair_wgrad folds the parities.
      if (d->Wout & 1) return -1014;
      c.Pw = d->Wout / 2; c.Pc = 2 * d->Cout_s; c.Pc_s = 2 * d->Cout_s;
      c.win_stride = 2 * d->stride; c.win_cols = d->Wout / 2;
    }
  } else {
    c.kh = d->kh; c.kw = d->kw; c.off_h = -d->zpad; c.off_w = -d->zpad; c.step_h = 1; c.step_w = 1;
    c.N_store = d->Cin_s;
  }
  c.out = dw_packed;
  return 0;
}

extern "C" int mt_conv2d_wgrad_splits(const MtConv2d* d) {
  WgradLike c;
  if (describe_conv_wgrad(d, nullptr, nullptr, nullptr, c)) return 0;
  WgradParams p;
  plan_wgrad(c, p);
  return p.splits;
}

extern "C" int mt_conv2d_wgrad(const MtConv2d* d, const void* x, const void* dy, float* dw_packed,
                               int splits, void* stream) {
  WgradLike c;
  const int rc = describe_conv_wgrad(d, x, dy, dw_packed, c);
  return rc ? rc : run_wgrad_like(c, splits, (cudaStream_t)stream);
}

static int describe_convT_wgrad(const MtConv2d* d, const void* x, const void* dy, float* dw_packed,
                                WgradLike& c) {
  if (!d || d->Cin_s % 8 || d->Cout_s % 8) return -1003;
  memset(&c, 0, sizeof(c));
  // dW[ci][co][r][s] = sum x[hi, wi, ci] * dy[2hi - zpad + r, 2wi - zpad + s, co]
  c.p = x; c.Pn = d->N; c.Ph = d->Hin; c.Pw = d->Win; c.Pc = d->Cin; c.Pc_s = d->Cin_s;
  c.q = dy; c.Qn = d->N; c.Qh = d->Hout; c.Qw = d->Wout; c.Qc = d->Cout; c.Qc_s = d->Cout_s;
  c.q_mul = 2; c.window = false;
  c.kh = d->kh; c.kw = d->kw; c.off_h = -d->zpad; c.off_w = -d->zpad; c.step_h = 1; c.step_w = 1;
  c.N_store = d->Cout_s;
  c.out = dw_packed;
  return 0;
}

extern "C" int mt_conv_transpose2d_wgrad_splits(const MtConv2d* d) {
  WgradLike c;
  if (describe_convT_wgrad(d, nullptr, nullptr, nullptr, c)) return 0;
  WgradParams p;
  plan_wgrad(c, p);
  return p.splits;
}

extern "C" int mt_conv_transpose2d_wgrad(const MtConv2d* d, const void* x, const void* dy,
                                         float* dw_packed, int splits, void* stream) {
  WgradLike c;
  const int rc = describe_convT_wgrad(d, x, dy, dw_packed, c);
  return rc ? rc : run_wgrad_like(c, splits, (cudaStream_t)stream);
}
