// Host-visible parameter blocks of the implicit-GEMM kernels (igemm_fwd.cu, igemm_wgrad.cu).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

namespace mt {

enum Act : int { ACT_NONE = 0, ACT_RELU = 1, ACT_LRELU = 2, ACT_TANH = 3 };

// D[pixels, Cout] = sum_taps A_tap[pixels, Cin] * W_tap[Cin, Cout]
// A tile: TN x TH x TW pixels (=128 rows) x 64 channels, fetched by one 4-D TMA box.
struct IgemmParams {
  int tiles_w, tiles_h, tiles_n, n_tiles;  // tile grid (pixels) and Cout tiles
  int TW, TH, TN;                          // pixel tile, TW*TH*TN == 128
  int dbg;                                 // debug: 32 = clock64 stamps of CTA 0 (mt_debug_timestamps)
  int cg;                                  // 1: one CTA per 128-pixel tile; 2: CTA pair, M = 256 MMA
  int work_items;                          // ceil(pixel tiles / cg) * n_tiles, walked round-robin
  int acc_cols;                            // TMEM columns between the two accumulator stages
  int Wt, Ht, Nt;                          // valid extents of the pixel space
  int kh, kw;                              // tap grid
  int mul_h, mul_w;                        // TMA coord = pix*mul + off + tap*step
  int off_h, off_w, step_h, step_w;
  int w_base, w_step_r, w_step_s;          // weight tap index = base + r*step_r + s*step_s
  int k_chunks;                            // ceil(K per tap / 64)
  int BN;                                  // Cout tile (16..256, multiple of 16)
  int n_store;                             // channels to store per pixel (multiple of 8)
  int n_bias;                              // number of valid bias entries (0: no bias)
  int bias_wrap;                           // > 0: column c >= bias_wrap reads bias[c - bias_wrap] (two column
                                           // groups sharing one bias: merged ConvTranspose2d phases)
  int stages;
  int tmem_cols;
  long long out_n_stride, out_h_stride, out_w_stride, out_off;  // elements
  int act;
  float slope;
  int accumulate;      // 1: out += result (TMA reduce-add store) instead of out = result
  const float* bias;
  long long bias_n_stride;   // > 0: image n of a tile (TN == 1) reads bias + n * bias_n_stride (per-image bias)
  __nv_bfloat16* out;
  float* stats;        // optional fused plane statistics: [N][stat_parts][n_stat_c][2] fp32 partial
  int n_stat_c;        // (sum, sumsq) per pixel tile; this launch fills slots stat_part0 + tile index
  int stat_parts, stat_part0;
  // "rows in K, columns in N" data gradient into <= 4 channels (BN == 32 only, see epilogue_rowk in igemm_fwd.cu):
  // GEMM column (s, c) holds tap column s of input channel c; the epilogue adds the kw shifted columns and
  // writes fp32 partial rows to rowk_out instead of storing the tile.  rowk_kw == 0: off.
  int rowk_kw, rowk_c, rowk_win;
  float* rowk_out;     // [Nt][Ht][tiles_w][TW + kw - 1][4]
};

// dW_tap[Cout, Cin] = sum_pixels dY[pix, Cout]^T * X_tap[pix, Cin]   (both operands MN-major)
struct WgradParams {
  int tiles_w, tiles_h, tiles_n;  // pixel (K) tiles
  int TW, TH, TN;                 // pixel tile: TW*TH*TN == 64 (one K block)
  int kh, kw;
  int mul_h, mul_w, off_h, off_w, step_h, step_w;  // X coord = pix*mul + off + tap*step
  int m_tiles, n_tiles;           // Cout tiles of 128, Cin tiles of BN
  int BN;
  int cg;                         // 1: one CTA per (tap, M tile, N tile, split); 2: CTA pair over two M tiles
  int dbg;                        // debug (MT_IGEMM_DBG): 1 = no TMA traffic after the first ring round, 2 = no MMAs
  int box;                        // 1: box mode (igemm_wgrad.cu): 8 x 8 pixel tiles, one halo'd box of Q per filter row
  int tap_group, tap_groups;      // filter taps per work item (they share the P slab) and number of such groups
  int splits;                     // split-K factor over pixel tiles
  int stages;
  int tmem_cols;
  int M_valid, N_valid;           // Cout, Cin(stored) extents to write
  long long out_tap_stride, out_m_stride;  // fp32 element strides of dW[tap][m][n]
  long long out_split_stride;     // ... and between the partial results of the split-K slices
  float* out;                     // [splits][taps][M][N]: every split stores its own slab (plain stores, no
                                  // zeroing, no atomics); the unpack kernels add the slabs in split order
};

int make_tmap(CUtensorMap* m, const void* base, int rank, const uint64_t* dims,
              const uint64_t* strides_bytes, const uint32_t* box, const uint32_t* estr,
              bool swizzle128 = true);

int launch_igemm(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC,
                 const IgemmParams& p, cudaStream_t stream);
int igemm_halo_supported(int cg, int BN, int kh, int kw, int k_chunks);
int launch_igemm_halo(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC,
                      const IgemmParams& p, cudaStream_t stream);
int launch_igemm_halo_s2(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC,
                         const IgemmParams& p, cudaStream_t stream);
int launch_wgrad(const CUtensorMap& tmA, const CUtensorMap& tmB, const WgradParams& p,
                 cudaStream_t stream);

static inline int cdiv(int a, int b) { return (a + b - 1) / b; }
static inline int round_up(int a, int b) { return cdiv(a, b) * b; }

// Pick the pixel tile (TW x TH x TN == P) with least padded volume; prefer wide tiles.
static inline void pick_tile(int P, int Wt, int Ht, int Nt, int box_mul, int* TW, int* TH, int* TN) {
  long long best = -1;
  for (int tw = P; tw >= 1; tw >>= 1) {
    if (tw * box_mul > 256) continue;
    for (int th = P / tw; th >= 1; th >>= 1) {
      if (th * box_mul > 256) continue;
      int tn = P / (tw * th);
      if (tn > 256) continue;
      long long vol = (long long)round_up(Wt, tw) * round_up(Ht, th) * round_up(Nt, tn);
      if (best < 0 || vol < best) {
        best = vol;
        *TW = tw; *TH = th; *TN = tn;
      }
    }
  }
}

static inline int pick_bn(int n_logical) {
  if (n_logical >= 256) return 256;
  if (n_logical > 64) return 128;
  if (n_logical > 32) return 64;
  if (n_logical > 16) return 32;
  return 16;
}
static inline int pow2_cols(int bn) {
  int c = 32;
  while (c < bn) c <<= 1;
  return c;
}


}  // namespace mt
