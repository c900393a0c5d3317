/* mt_b200.h — C ABI of the B200-native hot path of kartikkadur/MasterThesis.
 *
 * The reference has no FFI: its hot path is torch.nn modules (src/models/core/blocks.py,
 * norm.py, networks.py). Each entry point below replaces one family of torch.nn calls on
 * that path; the file:line after each declaration is the reference call site it stands in
 * for. All pointers are DEVICE pointers unless a name ends in _host; all activation
 * tensors are NHWC bf16 ("stored" channel counts are multiples of 8); `stream` is a
 * cudaStream_t passed as void*. Every function returns 0 on success or a negative error
 * code (-cudaError_t, or <= -1000 for argument errors) and never allocates device memory: every
 * buffer, scratch included, is the caller's (the *_parts / *_splits / *_elems queries size them).
 * Host-side state the library keeps, all created lazily and documented here:
 *   - per (device, caller stream): three auxiliary streams and four events with which the sub-pixel phases /
 *     row parities of stride-2 data gradients and transposed convolutions run side by side (forked from and
 *     joined back into `stream` before the call returns: plain stream semantics for the caller, graph edges
 *     under capture; MT_PHASE_FORK=0 disables);
 *   - ONE cudaStreamSynchronize + 8-byte cudaMemcpyFromSymbol at the first implicit-GEMM launch of the
 *     process outside stream capture (self check that the kernels' shared-memory ring sits at the offset
 *     their constant descriptors assume; error -1011 otherwise).  No other call synchronises.
 * Device-side state: one 2 KiB `static __device__` array per GPU, the built-in scratch of mt_loss_l1_static.
 * Entry points are re-entrant per stream.  There is no CPU fallback: without a CUDA device every call fails.
 */
#ifndef MT_B200_H
#define MT_B200_H
#ifdef __cplusplus
extern "C" {
#endif

enum { MT_ACT_NONE = 0, MT_ACT_RELU = 1, MT_ACT_LRELU = 2, MT_ACT_TANH = 3 };
enum { MT_NORM_INSTANCE = 0, MT_NORM_ADAIN = 1, MT_NORM_LAYER = 2 };

/* One convolution-shaped contraction.  The input buffer is [N, Hin, Win, Cin_s]; reflection
 * borders are already materialised in it (blocks.py:29-31 folds all padding into
 * ReflectionPad2d), `zpad` is additional logical ZERO padding realised by TMA out-of-bounds
 * fill.  The result is written at offset (out_pad, out_pad) of an [N, out_H, out_W, Cout_s]
 * buffer so that the consumer's reflection border can be filled in place. */
typedef struct MtConv2d {
  int N;
  int Hin, Win, Cin, Cin_s;
  int Hout, Wout, Cout, Cout_s;
  int kh, kw, stride, zpad;
  int out_H, out_W, out_pad;
  int act;      /* MT_ACT_* fused into the epilogue */
  float slope;  /* LeakyReLU slope (functions.py:34 -> 0.01) */
  int window;   /* 1: image-stem row-window mode (requires mt_conv_window_mode(Cin_s,kw) and an
                   input buffer followed by >= 128 B of finite slack, e.g. from the image packer) */
  int single;   /* 1: never use the merged-parity / paired-pixel column modes for this call (they need
                   even extents: the caller sets it for odd image sizes, e.g. the reference's 540x960
                   sampling size whose feature maps are 135 and 67 wide, sample.py:50) and packs the
                   operands in the plain layouts of mt_pack_conv_weight */
} MtConv2d;

/* ---- library --------------------------------------------------------------------- */
int mt_version(void);
const char* mt_error_string(int code);
/* stored channel count for a logical one (multiple of 8) */
int mt_stored_channels(int c);
/* 1 if a (Cin_s, kw) stem runs in row-window mode (K = kh x 64 window elements): Cin_s <= 16 and
 * Cin_s*kw <= 64.  The input buffer of such a conv must be followed by >= 128 B of finite data. */
int mt_conv_window_mode(int Cin_s, int kw);

/* ---- weight packing (fp32 OIHW master -> bf16 GEMM operands) ------------------------
 * nn.Conv2d weight [Cout, Cin, kh, kw] (blocks.py:36) or nn.ConvTranspose2d weight
 * [Cin, Cout, kh, kw] (blocks.py:72), fp32.  `transposed`=1 for ConvTranspose2d.
 *  w_fwd : forward operand  [taps][Cout_s'][Cin_s]   (window mode: [kh][Cout][64])
 *  w_bwd : data-grad operand [taps][Cin_s'][Cout_s]
 * Either output may be NULL. Sizes in elements are returned by mt_packed_weight_elems. */
long long mt_packed_weight_elems(int Cout, int Cin, int kh, int kw, int window, int bwd);
int mt_pack_conv_weight(const float* w, int Cout, int Cin, int kh, int kw, int transposed,
                        int window, void* w_fwd, void* w_bwd, void* stream);

/* ConvTranspose2d(k=3, s=2, p=1, op=1) — blocks.py:69-73 — runs its FORWARD in a merged-parity mode
 * when mt_convT_merged_mode(kh, kw, zpad) is 1: two launches over the input pixel grid, GEMM columns
 * (output column parity, Cout).  Its w_fwd operand is then [6 taps][2*Cout_s][Cin_s] bf16, packed by
 * mt_pack_convT3x3s2_weight (mt_packed_convT3x3s2_elems elements); w_bwd and the weight-gradient
 * layout are unchanged (mt_pack_conv_weight with w_fwd = NULL). */
int mt_convT_merged_mode(int kh, int kw, int zpad);
long long mt_packed_convT3x3s2_elems(int Cout, int Cin);
int mt_pack_convT3x3s2_weight(const float* w, int Cout, int Cin, void* w_fwd, void* stream);

/* The data gradient of a stride-2 convolution without zero padding (reflection-padded input:
 * blocks.py:29-36 with stride 2, k = 3 or 4) runs in the same merged-parity mode when
 * mt_conv_s2_dgrad_merged_mode(kh, kw, stride, zpad, Cin) is 1 (k = 4, or k = 3 with Cin <= 128): mt_conv2d_dgrad then expects w_bwd =
 * [2*((k+1)/2 + k/2) taps][2*Cin_s][Cout_s] bf16 from mt_pack_conv_s2_dgrad_weight, and even Hin / Win
 * (error -1014 otherwise).  w_fwd is unchanged (mt_pack_conv_weight with w_bwd = NULL). */
int mt_conv_s2_dgrad_merged_mode(int kh, int kw, int stride, int zpad, int Cin);
long long mt_packed_conv_s2_dgrad_elems(int Cout, int Cin, int k);
int mt_pack_conv_s2_dgrad_weight(const float* w, int Cout, int Cin, int k, void* w_bwd, void* stream);

/* The data gradient of a stride-1 un-padded conv into <= 8 stored channels (the image gradient of the
 * 7x7 stems, networks.py:30,76) computes a pixel pair per GEMM row when
 * mt_conv_s1_dgrad_pair_mode(kh, kw, stride, zpad, Cin_s) is 1: w_bwd = [kh*(kw+1)][16][Cout_s] bf16 from
 * mt_pack_conv_s1_dgrad_pair_weight, even Win (error -1014 otherwise). */
int mt_conv_s1_dgrad_pair_mode(int kh, int kw, int stride, int zpad, int Cin_s);
long long mt_packed_conv_s1_dgrad_pair_elems(int Cout, int k);
int mt_pack_conv_s1_dgrad_pair_weight(const float* w, int Cout, int Cin, int k, void* w_bwd, void* stream);

/* Rows-in-K data gradient for a stride-1 un-padded conv with many taps into <= 4 channels, i.e. the image
 * gradient of the 7x7 stems (networks.py:30: 49 taps, 3 channels): one implicit-GEMM launch over (output row,
 * dy column) whose K runs over the kh filter rows (dy rows h' - r) and whose columns are (filter column s,
 * channel c); the epilogue adds the kw shifted columns, per-tile partial rows go to `scratch` and a second
 * kernel adds neighbouring tiles in a fixed order (no atomics).  dy crosses L2 kh times instead of
 * kh*(kw+1)/2 times.  When mt_conv_rowk_dgrad_mode(kh, kw, stride, zpad, Cin, Cout_s) is 1 the caller may use
 * mt_conv2d_dgrad_rowk instead of mt_conv2d_dgrad: w_rowk = [kh][32][Cout_s] bf16 from
 * mt_pack_conv_rowk_weight, scratch = mt_conv_rowk_scratch_elems(d) fp32 (caller-owned, needs no zeroing),
 * dx = [N, Hin, Win, 8] bf16 as from mt_conv2d_dgrad.  No even-extent requirement.  MT_DGRAD_ROWK=0 turns the
 * predicate off. */
int mt_conv_rowk_dgrad_mode(int kh, int kw, int stride, int zpad, int Cin, int Cout_s);
long long mt_packed_conv_rowk_elems(int Cout, int kh);
int mt_pack_conv_rowk_weight(const float* w, int Cout, int Cin, int kh, int kw, void* w_rowk, void* stream);
long long mt_conv_rowk_scratch_elems(const MtConv2d* d);
int mt_conv2d_dgrad_rowk(const MtConv2d* d, const void* dy, const void* w_rowk, float* scratch, void* dx,
                         void* stream);

/* Row-window stems (mt_conv_window_mode) whose 64-element window also covers the next output pixel
 * ((kw + stride) * Cin_s <= 64: the 7x7 image stems networks.py:30,76 and the stride-2 stems
 * networks.py:117,361) compute two adjacent output pixels per GEMM row when
 * mt_conv_window_pair_mode(Cin_s, kw, stride) is 1: mt_conv2d_fwd[_stats] then expects w_fwd =
 * [kh][2*Cout_s][64] from mt_pack_conv_window_pair_weight and an even Wout (error -1014 otherwise). */
int mt_conv_window_pair_mode(int Cin_s, int kw, int stride);
long long mt_packed_conv_window_pair_elems(int Cout, int kh);
int mt_pack_conv_window_pair_weight(const float* w, int Cout, int Cin, int kh, int kw, int stride,
                                    void* w_fwd, void* stream);
/* In the same mode mt_conv2d_wgrad writes dw_packed = [splits][kh][2*Cout_s][64] fp32
 * (mt_packed_conv_window_pair_elems elements per split); this unpack adds the splits and folds the two
 * pixel parities into dW (OIHW). */
int mt_unpack_conv_window_pair_wgrad(const float* dw_packed, int splits, int Cout, int Cin, int kh, int kw,
                                     int stride, int accumulate, float* dw, void* stream);

/* ---- convolutions (implicit GEMM, tcgen05/TMEM, TMA) -------------------------------- */
/* nn.Conv2d forward (+bias, +activation)            blocks.py:33-36, networks.py:373-374 */
int mt_conv2d_fwd(const MtConv2d* d, const void* x, const void* w_fwd, const float* bias,
                  void* y, void* stream);
/* Same with a PER-IMAGE bias, bias_n[N][Cout] fp32: the first 1x1 convolution of DecResnetBlock's block1 /
 * block2 (blocks.py:180-190,199-205) reads cat(out, z tiled over the plane); the tiled half of its K is a
 * per-image constant, so  conv1x1(cat(out, z)) = conv1x1(out; W[:, :C]) + (W[:, C:] z_n + b)  and only the
 * first term is a convolution.  Needs one image per pixel tile (Hout * Wout >= 128 or so: -1014 otherwise, the
 * caller then materialises the concatenation). */
int mt_conv2d_image_bias_ok(const MtConv2d* d);   /* 1: the launch has one image per pixel tile */
int mt_conv2d_fwd_image_bias(const MtConv2d* d, const void* x, const void* w_fwd, const float* bias_n,
                             void* y, void* stream);
/* Same, with the statistics pass of the following InstanceNorm / LayerNorm (blocks.py:38-41,
 * networks.py:470-489) fused into the conv epilogue: per-channel sum / sum of squares of the STORED
 * bf16 output of every pixel tile go to stats[N][parts][Cout_s][2] (fp32, every slot written once,
 * no zeroing needed); mt_norm_finalize adds the parts.  `parts` must be mt_conv_stats_parts(d, 0),
 * which is 0 for shapes that cannot fuse (the caller then runs mt_plane_stats). */
int mt_conv_stats_parts(const MtConv2d* d, int transposed);
int mt_conv2d_fwd_stats(const MtConv2d* d, const void* x, const void* w_fwd, const float* bias,
                        void* y, float* stats, int parts, void* stream);
/* d(loss)/d(input buffer), including the reflection border rows/cols (folded later) */
int mt_conv2d_dgrad(const MtConv2d* d, const void* dy, const void* w_bwd, void* dx, void* stream);
/* accumulate != 0: dx += d(loss)/d(input buffer) (the epilogue's TMA store becomes a reduce-add; every element
 * is added exactly once): the residual fan-in `x + block(x)` of ResnetBlock / AdaINResnetBlock
 * (blocks.py:137,166) without a separate add pass — dx then already holds the gradient of the skip path */
int mt_conv2d_dgrad_acc(const MtConv2d* d, const void* dy, const void* w_bwd, void* dx, int accumulate,
                        void* stream);
/* d(loss)/d(weight) as split-K partial results: fp32 dw_packed[splits][taps][Cout][Cin_s] (window mode:
 * [splits][kh][Cout][64]), splits = mt_conv2d_wgrad_splits(d), mt_packed_wgrad_elems() elements per split.
 * Every split stores its own slab (no zeroing by the caller, no atomics); mt_unpack_conv_wgrad adds the
 * slabs in split order and scatters (or adds) the result back to OIHW: weight gradients are
 * reproducible bit for bit.  Error -1015 when `splits` does not match the query. */
long long mt_packed_wgrad_elems(int Cout, int Cin, int kh, int kw, int transposed, int window);
int mt_conv2d_wgrad_splits(const MtConv2d* d);
int mt_conv2d_wgrad(const MtConv2d* d, const void* x, const void* dy, float* dw_packed, int splits,
                    void* stream);
/* accumulate != 0: dw += unpacked (straight into the parameter's .grad, what autograd's
 * AccumulateGrad would do with a returned tensor) */
int mt_unpack_conv_wgrad(const float* dw_packed, int splits, int Cout, int Cin, int kh, int kw,
                         int transposed, int window, int accumulate, float* dw, void* stream);
/* nn.ConvTranspose2d(k=3, s=2, p=1, op=1) forward / dgrad / wgrad     blocks.py:69-73.
 * Descriptor: Hin/Win = input dims, Hout/Wout = 2*Hin, 2*Win; kh=kw=3, stride=2, zpad=1. */
int mt_conv_transpose2d_fwd(const MtConv2d* d, const void* x, const void* w_fwd,
                            const float* bias, void* y, void* stream);
/* with fused output statistics, as mt_conv2d_fwd_stats; parts = mt_conv_stats_parts(d, 1) */
int mt_conv_transpose2d_fwd_stats(const MtConv2d* d, const void* x, const void* w_fwd,
                                  const float* bias, void* y, float* stats, int parts, void* stream);
int mt_conv_transpose2d_dgrad(const MtConv2d* d, const void* dy, const void* w_bwd, void* dx,
                              void* stream);
int mt_conv_transpose2d_wgrad_splits(const MtConv2d* d);
int mt_conv_transpose2d_wgrad(const MtConv2d* d, const void* x, const void* dy, float* dw_packed,
                              int splits, void* stream);

/* ---- layout / padding ------------------------------------------------------------------ */
/* NCHW fp32 image (+ one-hot domain code tiled over H,W: networks.py:136-140) -> NHWC bf16 with
 * the consumer's reflection border (blocks.py:29-31); channels >= Ci+nd are zero. */
int mt_image_to_nhwc(const float* img, const float* onehot, int N, int Ci, int nd, int H, int W,
                     int pad, int Cs, void* out, void* stream);
/* adjoint of the above w.r.t. the image channels: folds border gradients back. Cs must be 8. */
int mt_image_grad_from_nhwc(const void* dpad, int N, int Ci, int H, int W, int pad, int Cs,
                            float* dimg, void* stream);
/* nn.ReflectionPad2d(pad) in place on [N,H+2p,W+2p,C] whose interior is already written */
int mt_reflect_border(void* buf, int N, int H, int W, int C, int pad, void* stream);
/* ReflectionPad2d backward: padded gradient -> dense [N,H,W,C] */
int mt_reflect_fold(const void* dpad, int N, int H, int W, int C, int pad, void* dense,
                    void* stream);
/* dpad[interior] += dense   (residual fan-in, blocks.py:137,166) */
int mt_add_interior(void* dpad, const void* dense, int N, int H, int W, int C, int pad,
                    void* stream);

/* ---- normalisation (InstanceNorm2d / AdaptiveInstanceNorm / LayerNorm) ---------------------
 * functions.py:17, norm.py:5-33.  stats/sums: [N][C][2] fp32; ss: (scale, shift) per (n,c);
 * mr: (mean, rstd); coef: [N][C][3]; gb: fc(s) output [N][2C] (weight | bias).
 * C is the STORED channel count; C_valid <= C the logical one (LayerNorm statistics run over
 * C_valid*HW elements and ln_w / ln_b / dln_* hold C_valid entries; padded channels stay zero). */
/* stats[N][parts][C][2] = per-chunk (sum, sum of squares), parts = mt_plane_stats_parts(N, HW, C);
 * plain stores, every slot written once (no zeroing, no atomics: reproducible forward passes) */
int mt_plane_stats_parts(int N, int HW, int C);
int mt_plane_stats(const void* x, int N, int HW, int C, float* stats, int parts, void* stream);
/* stats: [N][parts][C][2] (from mt_plane_stats or from the conv epilogue, mt_conv2d_fwd_stats) */
/* csum (optional, [N][C]): the per-channel plane sums, needed by mt_norm_bwd_params for the bias gradient of
 * a biased convolution in front of a LayerNorm */
int mt_norm_finalize(const float* stats, int parts, int mode, int N, int C, int C_valid, int HW, float eps,
                     const float* gb, const float* ln_w, const float* ln_b, float* ss, float* mr,
                     float* csum, void* stream);
/* out_pad[interior] = act(x*scale+shift) (+ residual) (+ noise, misc.py:25); ss may be NULL.
 * With out_p > 0 the reflection border of out_pad is written too (no mt_reflect_border needed). */
int mt_affine_act_fwd(const void* x, const float* ss, int N, int H, int W, int C, int act,
                      float slope, const void* res_pad, int res_p, const float* noise,
                      void* out_pad, int out_p, void* stream);
/* g: gradient of the block output, dense (g_pad = 0) or the consumer's PADDED gradient buffer
 * [N, H+2g_pad, W+2g_pad, C], in which case the adjoint of ReflectionPad2d is applied on the fly */
/* sums: [N][parts][C][2] per-chunk (sum g', sum g'*x), parts = mt_norm_bwd_parts(...); plain stores, every
 * slot written once, added in a fixed order by mt_norm_bwd_finalize (reproducible backward passes) */
int mt_norm_bwd_parts(int N, int H, int W, int C, int act, int g_pad);
int mt_norm_bwd_reduce(const void* g, int g_pad, const void* x, const float* ss, int N, int H, int W,
                       int C, int act, float slope, float* sums, int parts, void* stream);
/* red (optional, [N][C][2]): per-sample (sum g'*xhat, sum g') for mt_norm_bwd_params */
int mt_norm_bwd_finalize(const float* sums, int parts, const float* mr, int mode, int N, int C, int C_valid,
                         int HW, const float* gb, const float* ln_w, float* coef, float* dgb, float* red,
                         void* stream);
/* Gradients that sum over the samples (in sample order): the LayerNorm affine parameters dln_w / dln_b
 * [C_valid] (norm.py:19-21; both or neither) and dbias [C_valid] = sum_n sum_hw dx, the gradient of the bias
 * of the convolution in front of the norm (needs coef and the forward csum; exactly zero for InstanceNorm /
 * AdaIN, where the caller skips it).  accumulate != 0: += into the outputs (parameter .grad buffers). */
int mt_norm_bwd_params(const float* red, const float* coef, const float* csum, int N, int C, int C_valid,
                       int HW, float* dln_w, float* dln_b, float* dbias, int accumulate, void* stream);
/* db[c] (+)= sum over `rows` of stats[row][c][0], c < C_valid: column sums of a dense tensor (bias gradients,
 * blocks.py:36) from the partial sums of mt_plane_stats (rows = N * parts), added in row order */
int mt_colsum_into(const float* stats, int rows, int C, int C_valid, float* db, int accumulate, void* stream);
/* the same per image: out[n][c] = sum over the `parts` rows of image n (the per-image column sums of dy that the
 * gradient of a per-image bias needs, mt_conv2d_fwd_image_bias) */
int mt_colsum_images(const float* stats, int N, int parts, int C, int C_valid, float* out, void* stream);
int mt_norm_bwd_apply(const void* g, int g_pad, const void* x, const float* ss, const float* coef,
                      int N, int H, int W, int C, int act, float slope, void* dx, void* stream);
/* backward of conv+bias+activation written into a padded buffer: fold + act' */
int mt_act_bwd_fold(const void* dpad, const void* out_pad, int N, int H, int W, int C, int pad,
                    int act, float slope, void* dy, void* stream);

/* ---- optional generator variants -------------------------------------------- */
/* nn.Upsample(scale_factor=2, mode='nearest') of UpsampleBlock(up_type='nearest')  blocks.py:74-76:
 * dense [N,H,W,C] -> dense [N,2H,2W,C]; bwd sums each 2x2 block of dy [N,2H,2W,C] into dx [N,H,W,C] */
int mt_upsample2_fwd(const void* x, int N, int H, int W, int C, void* out, void* stream);
int mt_upsample2_bwd(const void* dy, int N, int H, int W, int C, void* dx, void* stream);
/* nn.Dropout(0.5) of the decoder blocks (--use_dropout)  blocks.py:133-134,152,190: out = x * m over
 * n bf16 elements, m holding 0 or 1/(1-p); also its own adjoint (dx = dy * m) */
int mt_mul(const void* x, const void* m, long long n, void* out, void* stream);

/* ---- style-encoder pooling algebra, RGB head, content regulariser --------------------------- */
/* nn.AvgPool2d(2,2) over the interior of a padded buffer -> dense [N,H/2,W/2,C]  blocks.py:113 */
int mt_avgpool2_fwd(const void* x_pad, int pad, int N, int H, int W, int C, void* out,
                    void* stream);
/* dpad[interior] += 0.25 * upsample(dpooled) */
int mt_avgpool2_bwd_add(const void* dpooled, void* dpad, int pad, int N, int H, int W, int C,
                        void* stream);
/* out_pad[interior] = act(avgpool2(c) + d): DownResnetBlock sum (blocks.py:118-119) fused with
 * the in-place LeakyReLU that is the only consumer of the sum (blocks.py:107, networks.py:123) */
int mt_pool_add_act_fwd(const void* c, const void* d, int N, int H, int W, int C, int act,
                        float slope, void* out_pad, int out_p, void* stream);
int mt_pool_add_act_bwd(const void* dout_pad, const void* out_pad, int N, int H, int W, int C,
                        int out_p, int act, float slope, void* dd, void* dc, void* stream);
/* nn.AdaptiveAvgPool2d(1)  networks.py:124 */
int mt_gap_fwd(const void* x, int N, int HW, int C, float* out, void* stream);
int mt_gap_bwd(const float* dout, int N, int HW, int C, void* dx, void* stream);
/* ConvTranspose2d(C -> Co<=4, k=1, no bias) + tanh, NHWC bf16 -> NCHW fp32  networks.py:248-251 */
int mt_rgb_head_fwd(const void* x, const float* w, int N, int HW, int C, int Co, float* y,
                    void* stream);
/* dw_parts: scratch [parts][C][4] fp32, parts = mt_rgb_head_bwd_parts(N, HW, C) (per-block partial sums of
 * dw, added in block order: reproducible) */
int mt_rgb_head_bwd_parts(int N, int HW, int C);
int mt_rgb_head_bwd(const float* dy, const float* y, const void* x, const float* w, int N, int HW,
                    int C, int Co, void* dx, float* dw, float* dw_parts, int parts, void* stream);
/* sum(x^2) over the interior (adain_model.py:396-399) and dx = scale * x on the interior;
 * partials: scratch [parts] fp32, parts = mt_sumsq_interior_parts(N, H, W, C) */
int mt_sumsq_interior_parts(int N, int H, int W, int C);
int mt_sumsq_interior(const void* x_pad, int N, int H, int W, int C, int pad, float* out,
                      float* partials, int parts, void* stream);
int mt_scale_interior(const void* x_pad, const float* scale, int N, int H, int W, int C, int pad,
                      void* dx_pad, void* stream);

/* ---- small fp32 layers between the conv stacks and the losses (no library GEMMs in the iteration) ------- */
/* nn.Linear: y[N][Out] = act(x[N][In] w[Out][In]^T + b)  (act: MT_ACT_NONE | MT_ACT_RELU)
 * networks.py:127-128 (fc, fcVar), :256-261 (style MLP), norm.py:27 (AdaIN fc), :77,402-403 (1x1 heads) */
int mt_linear_fwd(const float* x, const float* w, const float* b, int N, int In, int Out, int act,
                  float* y, void* stream);
/* g = dy * act'(y);  dw (+)= g^T x, db (+)= sum_n g (accumulate_params: += into .grad buffers),
 * dx (+)= g w (accumulate_dx); dx / dw / db may be NULL; y only for MT_ACT_RELU.  Fixed-order sums. */
int mt_linear_bwd(const float* dy, const float* y, const float* x, const float* w, int N, int In, int Out,
                  int act, float* dx, float* dw, float* db, int accumulate_params, int accumulate_dx,
                  void* stream);
/* The decoder's style network in one call (networks.py:256-261 + norm.py:27 of every AdaINResnetBlock):
 * h1 = relu(L0(cz)), h2 = relu(L1(h1)), s = L2(h2), gb[k] = fc_k(s) for k < n_heads ([N][head_out] each).
 * mlp_w / mlp_b / head_w / head_b / gb are HOST arrays of device pointers (3 and n_heads entries). */
int mt_style_mlp_fwd(const float* cz, int N, int in_dim, int hid, int style_dim,
                     const float* const* mlp_w, const float* const* mlp_b, int n_heads, int head_out,
                     const float* const* head_w, const float* const* head_b, float* h1, float* h2,
                     float* s, float* const* gb, void* stream);
/* dgb[k] [N][head_out] (host array of device pointers; NULL entries = no gradient) and / or ds_in
 * [N][style_dim] -> every layer's gradients (written, or added when accumulate != 0) and dcz (optional).
 * ds / dh2 / dh1: scratch of N*style_dim / N*hid / N*hid floats. */
int mt_style_mlp_bwd(const float* cz, const float* h1, const float* h2, const float* s, int N,
                     int in_dim, int hid, int style_dim, const float* const* mlp_w, int n_heads,
                     int head_out, const float* const* head_w, const float* const* dgb, const float* ds_in,
                     float* const* d_mlp_w, float* const* d_mlp_b, float* const* d_head_w,
                     float* const* d_head_b, float* ds, float* dh2, float* dh1, float* dcz,
                     int accumulate, void* stream);
/* Discriminator heads (networks.py:373-376,380-383): h [N][S][S][Cs] bf16 (C valid channels) ->
 * out [N][S+2][S+2] fp32 = Conv2d(C,1,1,padding=1,bias=False) (exact-zero border) and
 * cls [N][nd] = Conv2d(C,nd,k=S,bias=False) + AdaptiveAvgPool2d(1); w1 [C], w2 [nd][C][S][S]; nd <= 8 */
int mt_dis_head_fwd(const void* h, const float* w1, const float* w2, int N, int S, int Cs, int C, int nd,
                    float* out, float* cls, void* stream);
/* images n < cls_first carry no class gradient (the fake half of a discriminator batch); dh / (dw1, dw2)
 * optional; accumulate != 0: dw1 / dw2 += (parameter .grad buffers) */
int mt_dis_head_bwd(const float* dout, const float* dcls, const void* h, const float* w1, const float* w2,
                    int N, int S, int Cs, int C, int nd, int cls_first, void* dh, float* dw1, float* dw2,
                    int accumulate, void* stream);
/* ---- loss layer: value and gradient in one pass (loss.py:52-64, adain_model.py:213-216,303-314) ---------- */
/* nn.BCEWithLogitsLoss (mean over n) against target[n] or, when target == NULL, the constant target_const:
 * loss[0] = the mean; dx (optional) = gscale / n * (sigmoid(x) - t) */
int mt_loss_bce_logits(const float* x, const float* target, float target_const, int n, float gscale,
                       float* loss, float* dx, void* stream);
/* nn.L1Loss: loss[0] = weight * mean|a - b|, da (optional) = gscale / n * sign(a - b);
 * partials: scratch [parts], parts = mt_loss_l1_parts(n); a, b, da 16-byte aligned */
int mt_loss_l1_parts(long long n);
int mt_loss_l1(const float* a, const float* b, long long n, float weight, float gscale, float* loss,
               float* da, float* partials, int parts, void* stream);
/* mt_loss_l1 for callers that cannot allocate (code under CUDA graph capture, serving loops, harnesses that
 * forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same
 * arithmetic, same kernels and the same result as mt_loss_l1, but the per-block partial sums live in a
 * file-scope `static __device__` array of 512 floats (2 KiB, held by every GPU that loads the library until
 * exit) instead of a caller buffer; the entry point calls no cudaMalloc / cudaFree / allocating routine.
 * Limit: n <= mt_loss_l1_static_max_n() = 2^30 elements (error -1016 beyond it).  The array is shared by
 * every call on the device and keeps the last call's contents (nothing needs zeroing: every slot read was
 * written by the same call); two calls that OVERLAP on the GPU (streams with no ordering between them) race —
 * order them, or use mt_loss_l1 with separate scratch buffers. */
long long mt_loss_l1_static_max_n(void);
int mt_loss_l1_static(const float* a, const float* b, long long n, float weight, float gscale, float* loss,
                      float* da, void* stream);
/* -0.5 * sum(1 + logvar - mu^2 - exp(logvar)) over n elements (adain_model.py:313-314);
 * dmu = gscale * mu, dlogvar = gscale * 0.5 * (exp(logvar) - 1) (both or neither) */
int mt_loss_kl_normal(const float* mu, const float* logvar, int n, float gscale, float* loss,
                      float* dmu, float* dlogvar, void* stream);

/* ---- optimizer ------------------------------------------------------------------------------ */
/* torch.optim.Adam step (coupled L2 weight decay, adain_model.py:58-61) over a flat fp32 bucket;
 * g is multiplied by gscale first (data-parallel averaging). `step` is the 1-based step count. */
int mt_adam_flat(float* p, const float* g, float* m, float* v, long long n, float lr, float beta1,
                 float beta2, float eps, float wd, int step, float gscale, void* stream);
/* Same step with the scalars read from DEVICE memory, so that the launch can live in a CUDA graph
 * that is replayed with new learning rates / step counts:
 * hyper[0..6] = {lr / (1 - beta1^step), 1 / sqrt(1 - beta2^step), beta1, beta2, eps, wd, gscale}. */
int mt_adam_flat_dev(float* p, const float* g, float* m, float* v, long long n, const float* hyper,
                     void* stream);

/* debug: clock64 stamps of CTA 0 of the last conv launch when MT_IGEMM_DBG&32 (64 x int64, host) */
int mt_debug_timestamps(void* out_host);

#ifdef __cplusplus
}
#endif
#endif /* MT_B200_H */
