// InstanceNorm2d / AdaptiveInstanceNorm / LayerNorm, forward and backward, as a plane-statistics
// pass + per-(n,c) coefficient finalisation + one fused affine/activation/residual/noise pass.
// Reference semantics: functions.py:17 (InstanceNorm2d affine=False, eps 1e-5, biased var),
// norm.py:23-33 ((1+w)*IN(x)+b with w,b = chunk(fc(s),2)), norm.py:5-21 (F.layer_norm over
// (C,H,W), per-channel affine), misc.py:18-26 (x + noise), blocks.py:137,166 (residual add).
#include "../../include/mt_b200.h"
#include <stdlib.h>
#include "ew_common.cuh"

namespace mt {

// ---------------------------------------------------------------- statistics
// x [N, HW, C] -> per-chunk partial sums out[N][chunks][C][2].  One CTA = (n, pixel chunk); thread =
// one 8-channel vector x a pixel lane; warp reads are 512 contiguous bytes.  Every slot is written
// exactly once with a plain store (no zeroing, no atomics): the finalize kernels add the slots in a
// fixed order, so forward AND backward passes are reproducible bit for bit.
// The backward variant reads the gradient through the ADJOINT of the reflection padding when
// gp > 0: g is then the padded buffer [N, H+2gp, W+2gp, C] written by the consumer's data-gradient
// and fold(g) is formed on the fly (only pixels next to the border have extra terms).
constexpr int EW_UNROLL = 4;

// Index arithmetic of the per-plane kernels is 32-bit relative to a per-image base pointer (one
// image has < 2^31 elements, checked by the launchers): with 64-bit multiplies per access the
// kernels spent ~40 of their ~170 instructions per 16-byte vector on addresses and were issue-bound
// at 2 TB/s (ncu: smsp__issue_active 51 %, 148 IMAD in the unrolled loop).
__device__ __forceinline__ bool near_border(int h, int w, int H, int W, int gp) {
  return (unsigned)(h - 1) < (unsigned)gp || (unsigned)(H - 2 - h) < (unsigned)gp ||
         (unsigned)(w - 1) < (unsigned)gp || (unsigned)(W - 2 - w) < (unsigned)gp;
}
// gimg: the padded gradient image of this sample, already offset by the thread's channel vector
__device__ __forceinline__ void add_fold_extras(const __nv_bfloat16* __restrict__ gimg, int h, int w,
                                                int H, int W, int C, int gp, float (&gv)[8]) {
  if (!near_border(h, w, H, W, gp)) return;
  int hs[3], ws[3];
  const int nh = reflect_fold_sources(h, H, gp, hs), nw = reflect_fold_sources(w, W, gp, ws);
  const int Wp = W + 2 * gp;
  for (int a = 0; a < nh; ++a)
    for (int b = 0; b < nw; ++b) {
      if (a == 0 && b == 0) continue;                       // the main term is already loaded
      float f[8];
      load8(gimg + (hs[a] * Wp + ws[b]) * C, f);
#pragma unroll
      for (int j = 0; j < 8; ++j) gv[j] += f[j];
    }
}

template <bool kBackward, int ACT, bool GP>
__global__ void __launch_bounds__(256, 2)
plane_reduce_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ g,
                    const float* __restrict__ ss, int H, int W, FastDiv divW, int C, int gp, int chunks,
                    float slope, float* __restrict__ out) {
  extern __shared__ float red[];  // [lanes][V*16]
  pdl_wait(); pdl_launch();
  const int V = C / 8, HW = H * W;
  const int lanes = blockDim.x / V;
  const int n = blockIdx.x / chunks, chunk = blockIdx.x % chunks;
  const int v = threadIdx.x % V, pl = threadIdx.x / V;
  const int per = (HW + chunks - 1) / chunks;
  const int p0 = chunk * per, p1 = min(HW, p0 + per);
  const int Wp = W + 2 * gp;
  const __nv_bfloat16* ximg = x + (size_t)n * HW * C + v * 8;
  const __nv_bfloat16* gimg = kBackward ? g + (size_t)n * (H + 2 * gp) * Wp * C + v * 8 : nullptr;
  const int goff = (gp * Wp + gp) * C;          // interior origin of the padded gradient
  float a0[8], a1[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) a0[j] = a1[j] = 0.f;
  float sc[8], sh[8];
  if (kBackward) {
    const float4* s4 = reinterpret_cast<const float4*>(ss + ((long long)n * C + v * 8) * 2);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float4 q = __ldg(s4 + j);
      sc[2 * j] = q.x; sh[2 * j] = q.y; sc[2 * j + 1] = q.z; sh[2 * j + 1] = q.w;
    }
  }
  if (pl < lanes) {
    for (int pb = p0 + pl; pb < p1; pb += EW_UNROLL * lanes) {
      uint4 xr[EW_UNROLL], gr[EW_UNROLL];
      int hh[EW_UNROLL];
#pragma unroll
      for (int u = 0; u < EW_UNROLL; ++u) {
        const int p = pb + u * lanes;
        if (p < p1) {
          xr[u] = ldg16(ximg + p * C);
          if (kBackward) {   // (h+gp)*Wp + w+gp == p + 2*gp*h + gp*Wp + gp
            hh[u] = GP ? fast_div(p, divW) : 0;
            gr[u] = ldg16(gimg + goff + (p + 2 * gp * hh[u]) * C);
          }
        }
      }
#pragma unroll
      for (int u = 0; u < EW_UNROLL; ++u) {
        const int p = pb + u * lanes;
        if (p < p1) {
          float xv[8];
          cvt8(xr[u], xv);
          if (!kBackward) {
#pragma unroll
            for (int j = 0; j < 8; ++j) { a0[j] += xv[j]; a1[j] += xv[j] * xv[j]; }
          } else {
            float gv[8];
            cvt8(gr[u], gv);
            if (GP) add_fold_extras(gimg, hh[u], p - hh[u] * W, H, W, C, gp, gv);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float y = xv[j] * sc[j] + sh[j];           // pre-activation (A*xhat + B)
              const float gg = gv[j] * act_grad_pre_t<ACT>(y, slope);
              a0[j] += gg;                                      // sum g
              a1[j] += gg * xv[j];                              // sum g*x  (xhat folded in finalize)
            }
          }
        }
      }
    }
  }
  if (pl < lanes) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      red[(pl * V + v) * 16 + j] = a0[j];
      red[(pl * V + v) * 16 + 8 + j] = a1[j];
    }
  }
  __syncthreads();
  // thread t < V*16 sums over lanes
  for (int t = threadIdx.x; t < V * 16; t += blockDim.x) {
    float s = 0.f;
    for (int l = 0; l < lanes; ++l) s += red[l * V * 16 + t];
    const int vv = t / 16, k = t % 16;
    const int c = vv * 8 + (k & 7);
    out[(((long long)n * chunks + chunk) * C + c) * 2 + (k >> 3)] = s;
  }
}

// ---------------------------------------------------------------- forward finalize
// stats -> ss[n][c] = (scale, shift) with y = x*scale + shift, mr[n][c] = (mean, rstd)
// `stats` is [N][parts][C][2]: parts = 1 from mt_plane_stats, one slot per pixel tile from the
// fused conv epilogue (mt_conv2d_fwd_stats).  Phase A adds the parts in a fixed order (thread =
// channel x part group, FIN_UNROLL loads in flight each: the sum is a chain of L2 round trips, so
// the parallelism across threads and blocks is what bounds it); phase B is per channel.
// grid = (N, channel splits): InstanceNorm / AdaIN blocks own FIN_CSPLIT channels each, LayerNorm
// needs the whole sample in one block.
static constexpr int FIN_UNROLL = 8;
static constexpr int FIN_CSPLIT = 64;
// tot_c[2*c], tot_c[2*c+1] (c < Cb) = sum over the `parts` slots of sp[part*C + c], in a fixed order
// (part groups of FIN_UNROLL loads in flight per thread, then the groups in index order).
__device__ __forceinline__ void sum_parts(const float2* __restrict__ sp, int parts, int C, int Cb,
                                          float* tot_c, float2* part_sm) {
  const int T = blockDim.x, tid = threadIdx.x;
  if (parts == 1) {
    for (int c = tid; c < Cb; c += T) {
      const float2 t = sp[c];
      tot_c[2 * c] = t.x; tot_c[2 * c + 1] = t.y;
    }
  } else {
    const int Cw = Cb < T ? Cb : T, G = T / Cw;
    const int c0 = tid % Cw, pg = tid / Cw;
    if (pg < G)
      for (int c = c0; c < Cb; c += Cw) {
        float2 a[FIN_UNROLL];
#pragma unroll
        for (int u = 0; u < FIN_UNROLL; ++u) a[u] = make_float2(0.f, 0.f);
        int i = pg;
        for (; i + (FIN_UNROLL - 1) * G < parts; i += FIN_UNROLL * G) {
#pragma unroll
          for (int u = 0; u < FIN_UNROLL; ++u) {
            const float2 t = sp[(long long)(i + u * G) * C + c];
            a[u].x += t.x; a[u].y += t.y;
          }
        }
#pragma unroll
        for (int u = 0; u < FIN_UNROLL - 1; ++u)
          if (i + u * G < parts) {
            const float2 t = sp[(long long)(i + u * G) * C + c];
            a[u].x += t.x; a[u].y += t.y;
          }
        float2 r = a[0];
#pragma unroll
        for (int u = 1; u < FIN_UNROLL; ++u) { r.x += a[u].x; r.y += a[u].y; }
        if (G == 1) { tot_c[2 * c] = r.x; tot_c[2 * c + 1] = r.y; }
        else part_sm[tid] = r;                     // G > 1 implies Cw == Cb: one pass
      }
    if (G > 1) {
      __syncthreads();
      if (tid < Cb) {
        float2 r = part_sm[tid];
        for (int g = 1; g < G; ++g) { const float2 t = part_sm[g * Cw + tid]; r.x += t.x; r.y += t.y; }
        tot_c[2 * tid] = r.x; tot_c[2 * tid + 1] = r.y;
      }
    }
  }
  __syncthreads();
}

// `csum` (optional, [N][C]): the per-channel plane sums, kept for the backward pass of a LayerNorm that
// follows a biased convolution (the bias gradient sum_hw dx needs sum_hw x per channel).
__global__ void __launch_bounds__(1024)
norm_finalize_kernel(const float* __restrict__ stats, int parts, int mode, int N, int C, int Cv,
                     int HW, float eps, const float* __restrict__ gb, const float* __restrict__ ln_w,
                     const float* __restrict__ ln_b, float* __restrict__ ss, float* __restrict__ mr,
                     float* __restrict__ csum) {
  extern __shared__ float tot_c[];   // [channels of this block][2]
  __shared__ float2 part_sm[1024];
  pdl_wait(); pdl_launch();
  const int n = blockIdx.x, T = blockDim.x, tid = threadIdx.x;
  const int cb = gridDim.y > 1 ? blockIdx.y * FIN_CSPLIT : 0;           // first channel of the block
  const int Cb = gridDim.y > 1 ? min(FIN_CSPLIT, C - cb) : C;           // channels of the block
  sum_parts(reinterpret_cast<const float2*>(stats) + (long long)n * parts * C + cb, parts, C, Cb, tot_c,
            part_sm);
  if (csum)
    for (int cl = tid; cl < Cb; cl += T) csum[(long long)n * C + cb + cl] = tot_c[2 * cl];
  __shared__ float tot[2];
  if (mode == MT_NORM_LAYER) {                     // gridDim.y == 1
    float s0 = 0.f, s1 = 0.f;
    for (int c = tid; c < Cv; c += T) {
      s0 += tot_c[2 * c];
      s1 += tot_c[2 * c + 1];
    }
    __shared__ float r0[32], r1[32];
    for (int o = 16; o; o >>= 1) {
      s0 += __shfl_xor_sync(0xffffffffu, s0, o);
      s1 += __shfl_xor_sync(0xffffffffu, s1, o);
    }
    if ((tid & 31) == 0) { r0[tid >> 5] = s0; r1[tid >> 5] = s1; }
    __syncthreads();
    if (tid == 0) {
      float t0 = 0.f, t1 = 0.f;
      for (int i = 0; i < (T + 31) / 32; ++i) { t0 += r0[i]; t1 += r1[i]; }
      tot[0] = t0; tot[1] = t1;
    }
    __syncthreads();
  }
  for (int cl = tid; cl < Cb; cl += T) {
    const int c = cb + cl;
    float mean, var, A = 1.f, B = 0.f;
    if (mode == MT_NORM_LAYER) {
      const float cnt = (float)Cv * (float)HW;
      mean = tot[0] / cnt;
      var = tot[1] / cnt - mean * mean;
      if (c >= Cv) { A = 0.f; B = 0.f; }           // padded storage channels stay zero
      else if (ln_w) { A = ln_w[c]; B = ln_b[c]; }
    } else {
      mean = tot_c[2 * cl] / (float)HW;
      var = tot_c[2 * cl + 1] / (float)HW - mean * mean;
      if (mode == MT_NORM_ADAIN) {
        A = 1.f + gb[(long long)n * 2 * C + c];
        B = gb[(long long)n * 2 * C + C + c];
      }
    }
    const float rstd = rsqrtf(fmaxf(var, 0.f) + eps);
    ss[((long long)n * C + c) * 2] = A * rstd;
    ss[((long long)n * C + c) * 2 + 1] = B - A * mean * rstd;
    mr[((long long)n * C + c) * 2] = mean;
    mr[((long long)n * C + c) * 2 + 1] = rstd;
  }
}

// ---------------------------------------------------------------- forward apply
// out_pad[n, h+po, w+po, :] = act(x*scale + shift) (+ res_pad[n, h+pr, w+pr, :]) (+ noise)
// grid = (pixel chunks, N); a thread owns ONE 8-channel vector (its scale/shift live in registers)
// and walks pixels, so a warp reads/writes 512 contiguous bytes per pixel row.
template <int ACT, bool RES, bool NOISE>
__global__ void __launch_bounds__(256, 2)
affine_act_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ ss, int H, int W,
                      FastDiv divW, int C, float slope, const __nv_bfloat16* __restrict__ res, int pr,
                      const float* __restrict__ noise, __nv_bfloat16* __restrict__ out, int po) {
  pdl_wait(); pdl_launch();
  const int V = C / 8, lanes = blockDim.x / V;
  const int v = threadIdx.x % V, pl = threadIdx.x / V;
  if (pl >= lanes) return;
  const int n = blockIdx.y, HW = H * W;
  float sc[8], sh[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { sc[j] = 1.f; sh[j] = 0.f; }
  if (ss) {
    const float4* s4 = reinterpret_cast<const float4*>(ss + ((long long)n * C + v * 8) * 2);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float4 q = __ldg(s4 + j);
      sc[2 * j] = q.x; sh[2 * j] = q.y; sc[2 * j + 1] = q.z; sh[2 * j + 1] = q.w;
    }
  }
  const int Wr = W + 2 * pr, Wo = W + 2 * po;
  // per-image bases (already at the interior origin and at this thread's channel vector)
  const __nv_bfloat16* ximg = x + (size_t)n * HW * C + v * 8;
  const __nv_bfloat16* rimg = RES ? res + ((size_t)n * (H + 2 * pr) * Wr + pr * Wr + pr) * C + v * 8 : nullptr;
  const float* nimg = NOISE ? noise + (size_t)n * HW * C + v * 8 : nullptr;
  __nv_bfloat16* oimg = out + ((size_t)n * (H + 2 * po) * Wo + po * Wo + po) * C + v * 8;
  const bool need_h = (RES && pr) || po;
  const int stride = gridDim.x * lanes;
  for (int pb = blockIdx.x * lanes + pl; pb < HW; pb += EW_UNROLL * stride) {
    uint4 xr[EW_UNROLL], rr[EW_UNROLL];
    int hh[EW_UNROLL];
#pragma unroll
    for (int u = 0; u < EW_UNROLL; ++u) {
      const int p = pb + u * stride;
      if (p < HW) {
        xr[u] = ldg16(ximg + p * C);
        hh[u] = need_h ? fast_div(p, divW) : 0;
        if (RES) rr[u] = ldg16(rimg + (p + 2 * pr * hh[u]) * C);
      }
    }
#pragma unroll
    for (int u = 0; u < EW_UNROLL; ++u) {
      const int p = pb + u * stride;
      if (p < HW) {
        float f[8];
        cvt8(xr[u], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] = act_fwd_t<ACT>(f[j] * sc[j] + sh[j], slope);
        if (RES) {
          float r[8];
          cvt8(rr[u], r);
#pragma unroll
          for (int j = 0; j < 8; ++j) f[j] += r[j];
        }
        if (NOISE) {
          float r[8];
          load8f(nimg + p * C, r);
#pragma unroll
          for (int j = 0; j < 8; ++j) f[j] += r[j];
        }
        store8(oimg + (p + 2 * po * hh[u]) * C, f);
        // the reflection border of the padded output (nn.ReflectionPad2d of the consumer,
        // blocks.py:29-31): every padded position that mirrors this pixel gets the same vector
        if (po && near_border(hh[u], p - hh[u] * W, H, W, po)) {
          int hs[3], ws[3];
          const int w = p - hh[u] * W;
          const int nh = reflect_fold_sources(hh[u], H, po, hs), nw = reflect_fold_sources(w, W, po, ws);
          __nv_bfloat16* opad = oimg - (po * Wo + po) * C;     // origin of the padded image
          for (int a = 0; a < nh; ++a)
            for (int b = 0; b < nw; ++b)
              if (a | b) store8(opad + (hs[a] * Wo + ws[b]) * C, f);
        }
      }
    }
  }
}

// ---------------------------------------------------------------- backward finalize
// sums[n][part][c] = per-chunk (sum g, sum g*x) with g already masked by act'.  Adds the chunks in a
// fixed order (sum_parts), then produces coef[n][c][3] such that dx = c0*g + c1*x + c2, the AdaIN
// modulation gradients, and (optional) red[n][c] = (sum g*xhat, sum g) for norm_bwd_params_kernel.
// grid = (N, channel splits) as in the forward finalize (LayerNorm: the whole sample in one block).
__global__ void __launch_bounds__(1024)
norm_bwd_finalize_kernel(const float* __restrict__ sums, int parts, const float* __restrict__ mr,
                         int mode, int N, int C, int Cv, int HW, const float* __restrict__ gb,
                         const float* __restrict__ ln_w, float* __restrict__ coef,
                         float* __restrict__ dgb, float* __restrict__ red) {
  extern __shared__ float tot_c[];   // [channels of this block][2]
  __shared__ float2 part_sm[1024];
  __shared__ float tot[2];
  __shared__ float r0[32], r1[32];
  pdl_wait(); pdl_launch();
  const int n = blockIdx.x, T = blockDim.x, tid = threadIdx.x;
  const int cb = gridDim.y > 1 ? blockIdx.y * FIN_CSPLIT : 0;
  const int Cb = gridDim.y > 1 ? min(FIN_CSPLIT, C - cb) : C;
  sum_parts(reinterpret_cast<const float2*>(sums) + (long long)n * parts * C + cb, parts, C, Cb, tot_c,
            part_sm);
  if (mode == MT_NORM_LAYER) {                     // gridDim.y == 1
    // mean_n(A g) and mean_n(A g xhat) over all (c, hw)
    float s0 = 0.f, s1 = 0.f;
    for (int c = tid; c < Cv; c += T) {
      const float mean = mr[((long long)n * C + c) * 2], rstd = mr[((long long)n * C + c) * 2 + 1];
      const float sg = tot_c[2 * c], sgx = tot_c[2 * c + 1];
      const float sgxh = (sgx - mean * sg) * rstd;
      const float A = ln_w ? ln_w[c] : 1.f;
      s0 += A * sg; s1 += A * sgxh;
    }
    for (int o = 16; o; o >>= 1) {
      s0 += __shfl_xor_sync(0xffffffffu, s0, o);
      s1 += __shfl_xor_sync(0xffffffffu, s1, o);
    }
    if ((tid & 31) == 0) { r0[tid >> 5] = s0; r1[tid >> 5] = s1; }
    __syncthreads();
    if (tid == 0) {
      float t0 = 0.f, t1 = 0.f;
      for (int i = 0; i < (T + 31) / 32; ++i) { t0 += r0[i]; t1 += r1[i]; }
      const float cnt = (float)Cv * (float)HW;
      tot[0] = t0 / cnt; tot[1] = t1 / cnt;
    }
    __syncthreads();
  }
  for (int cl = tid; cl < Cb; cl += T) {
    const int c = cb + cl;
    const float mean = mr[((long long)n * C + c) * 2], rstd = mr[((long long)n * C + c) * 2 + 1];
    const float sg = tot_c[2 * cl], sgx = tot_c[2 * cl + 1];
    const float sgxh = (sgx - mean * sg) * rstd;   // sum g*xhat
    if (red) { red[((long long)n * C + c) * 2] = sgxh; red[((long long)n * C + c) * 2 + 1] = sg; }
    float A = 1.f, mg, mgx;
    if (mode == MT_NORM_LAYER) {
      A = c >= Cv ? 0.f : (ln_w ? ln_w[c] : 1.f);
      mg = tot[0]; mgx = tot[1];                    // already include A
      // dx = rstd*(A g - mg - xhat*mgx); padded storage channels get no gradient
      const float live = c >= Cv ? 0.f : 1.f;
      coef[((long long)n * C + c) * 3 + 0] = rstd * A;
      coef[((long long)n * C + c) * 3 + 1] = -rstd * rstd * mgx * live;
      coef[((long long)n * C + c) * 3 + 2] = -rstd * (mg - mean * rstd * mgx) * live;
    } else {
      if (mode == MT_NORM_ADAIN) {
        A = 1.f + gb[(long long)n * 2 * C + c];
        dgb[(long long)n * 2 * C + c] = sgxh;       // d(weight) = sum g*xhat
        dgb[(long long)n * 2 * C + C + c] = sg;     // d(bias)   = sum g
      }
      mg = sg / (float)HW; mgx = sgxh / (float)HW;
      // dx = A*rstd*(g - mg - xhat*mgx)
      coef[((long long)n * C + c) * 3 + 0] = A * rstd;
      coef[((long long)n * C + c) * 3 + 1] = -A * rstd * rstd * mgx;
      coef[((long long)n * C + c) * 3 + 2] = -A * rstd * (mg - mean * rstd * mgx);
    }
  }
}

// Parameter gradients that sum over the samples, in sample order (reproducible): thread = channel.
//   dln_w[c] (+)= sum_n sum(g*xhat),  dln_b[c] (+)= sum_n sum(g)          (LayerNorm affine, norm.py:19-21)
//   dbias[c] (+)= sum_n sum_hw dx = sum_n (c0*sum(g) + c1*sum(x) + c2*HW)  (bias of the conv in front of
//   the norm; csum = per-channel plane sums kept by the forward finalize.  For InstanceNorm / AdaIN this
//   is exactly zero and the caller does not ask for it.)
__global__ void norm_bwd_params_kernel(const float* __restrict__ red, const float* __restrict__ coef,
                                       const float* __restrict__ csum, int N, int C, int Cv, int HW,
                                       float* __restrict__ dln_w, float* __restrict__ dln_b,
                                       float* __restrict__ dbias, int accumulate) {
  pdl_wait(); pdl_launch();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= Cv) return;
  float sw = 0.f, sb = 0.f, sd = 0.f;
  for (int n = 0; n < N; ++n) {
    const long long i = (long long)n * C + c;
    const float sgxh = red[i * 2], sg = red[i * 2 + 1];
    sw += sgxh; sb += sg;
    if (dbias) sd += coef[i * 3] * sg + coef[i * 3 + 1] * csum[i] + coef[i * 3 + 2] * (float)HW;
  }
  if (dln_w) { dln_w[c] = accumulate ? dln_w[c] + sw : sw; dln_b[c] = accumulate ? dln_b[c] + sb : sb; }
  if (dbias) dbias[c] = accumulate ? dbias[c] + sd : sd;
}

// dx = c0*(fold(g)*act'(y)) + c1*x + c2      grid = (pixel chunks, N), coefficients in registers;
// g is padded by gp (0: dense) and folded on the fly like in plane_reduce_kernel<true>.
template <int ACT, bool GP>
__global__ void __launch_bounds__(256, 2)
norm_bwd_apply_kernel(const __nv_bfloat16* __restrict__ g, const __nv_bfloat16* __restrict__ x,
                      const float* __restrict__ ss, const float* __restrict__ coef, int H, int W,
                      FastDiv divW, int C, int gp, float slope, __nv_bfloat16* __restrict__ dx) {
  pdl_wait(); pdl_launch();
  const int V = C / 8, lanes = blockDim.x / V, HW = H * W;
  const int v = threadIdx.x % V, pl = threadIdx.x / V;
  if (pl >= lanes) return;
  const int n = blockIdx.y;
  const int Wp = W + 2 * gp;
  const __nv_bfloat16* ximg = x + (size_t)n * HW * C + v * 8;
  const __nv_bfloat16* gimg = g + (size_t)n * (H + 2 * gp) * Wp * C + v * 8;
  __nv_bfloat16* dimg = dx + (size_t)n * HW * C + v * 8;
  const int goff = (gp * Wp + gp) * C;
  float sc[8], sh[8], c0[8], c1[8], c2[8];
  const float* s = ss + ((long long)n * C + v * 8) * 2;
  const float* cf = coef + ((long long)n * C + v * 8) * 3;
  {
    float t[16], u[24];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float4 q = __ldg(reinterpret_cast<const float4*>(s) + j);
      t[4 * j] = q.x; t[4 * j + 1] = q.y; t[4 * j + 2] = q.z; t[4 * j + 3] = q.w;
    }
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      const float4 q = __ldg(reinterpret_cast<const float4*>(cf) + j);
      u[4 * j] = q.x; u[4 * j + 1] = q.y; u[4 * j + 2] = q.z; u[4 * j + 3] = q.w;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      sc[j] = t[2 * j]; sh[j] = t[2 * j + 1];
      c0[j] = u[3 * j]; c1[j] = u[3 * j + 1]; c2[j] = u[3 * j + 2];
    }
  }
  const int stride = gridDim.x * lanes;
  for (int pb = blockIdx.x * lanes + pl; pb < HW; pb += EW_UNROLL * stride) {
    uint4 xr[EW_UNROLL], gr[EW_UNROLL];
    int hh[EW_UNROLL];
#pragma unroll
    for (int u = 0; u < EW_UNROLL; ++u) {
      const int p = pb + u * stride;
      if (p < HW) {
        hh[u] = GP ? fast_div(p, divW) : 0;
        xr[u] = ldg16(ximg + p * C);
        gr[u] = ldg16(gimg + goff + (p + 2 * gp * hh[u]) * C);
      }
    }
#pragma unroll
    for (int u = 0; u < EW_UNROLL; ++u) {
      const int p = pb + u * stride;
      if (p < HW) {
        float xv[8], gv[8], o[8];
        cvt8(xr[u], xv);
        cvt8(gr[u], gv);
        if (GP) add_fold_extras(gimg, hh[u], p - hh[u] * W, H, W, C, gp, gv);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float y = xv[j] * sc[j] + sh[j];
          o[j] = c0[j] * (gv[j] * act_grad_pre_t<ACT>(y, slope)) + c1[j] * xv[j] + c2[j];
        }
        store8(dimg + p * C, o);
      }
    }
  }
}

// dy[n,h,w,:] = fold(dpad)[n,h,w,:] * act'(out_pad interior)  — for conv+bias+act layers whose
// activated output was written straight into the consumer's padded buffer (D, style encoder).
template <int ACT>
__global__ void __launch_bounds__(256, 2)
act_bwd_fold_kernel(const __nv_bfloat16* __restrict__ dpad, const __nv_bfloat16* __restrict__ outp,
                    int H, int W, FastDiv divW, int C, int p, float slope,
                    __nv_bfloat16* __restrict__ dy) {
  // per-plane layout as norm_bwd_apply_kernel: grid (pixel chunks, N), thread = channel vector x lane
  pdl_wait(); pdl_launch();
  const int V = C / 8, lanes = blockDim.x / V, HW = H * W;
  const int v = threadIdx.x % V, pl = threadIdx.x / V;
  if (pl >= lanes) return;
  const int n = blockIdx.y, Wp = W + 2 * p;
  const size_t img = (size_t)n * (H + 2 * p) * Wp * C + v * 8;
  const __nv_bfloat16* gimg = dpad + img;
  const __nv_bfloat16* oimg = outp + img;
  __nv_bfloat16* dimg = dy + (size_t)n * HW * C + v * 8;
  const int off0 = (p * Wp + p) * C;
  const int stride = gridDim.x * lanes;
  for (int pb = blockIdx.x * lanes + pl; pb < HW; pb += EW_UNROLL * stride) {
    uint4 gr[EW_UNROLL], orr[EW_UNROLL];
    int hh[EW_UNROLL];
#pragma unroll
    for (int u = 0; u < EW_UNROLL; ++u) {
      const int q = pb + u * stride;
      if (q < HW) {
        hh[u] = p ? fast_div(q, divW) : 0;
        const int pos = off0 + (q + 2 * p * hh[u]) * C;
        gr[u] = ldg16(gimg + pos);
        orr[u] = ldg16(oimg + pos);
      }
    }
#pragma unroll
    for (int u = 0; u < EW_UNROLL; ++u) {
      const int q = pb + u * stride;
      if (q < HW) {
        float gv[8], o[8];
        cvt8(gr[u], gv);
        cvt8(orr[u], o);
        if (p) add_fold_extras(gimg, hh[u], q - hh[u] * W, H, W, C, p, gv);
#pragma unroll
        for (int j = 0; j < 8; ++j) gv[j] *= act_grad_pre_t<ACT>(o[j], slope);
        store8(dimg + q * C, gv);
      }
    }
  }
}

}  // namespace mt

using namespace mt;
#define ST(s) ((cudaStream_t)(s))

// Grid (pixel chunks, N) of the per-plane kernels: exactly ONE wave — SMs x the number of CTAs
// of THIS kernel instantiation that fit on an SM (register-limited: 2-3 for the 66-100 register
// kernels here).  A fixed "4 CTAs per SM" grid ran 1.33 waves on the 3-resident kernels, i.e. a
// third of the time with a quarter of the machine (ncu: launch__waves_per_multiprocessor).
// MT_EW_CTAS overrides the per-SM count.
static int ew_ctas_override() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("MT_EW_CTAS"); v = e ? atoi(e) : 0; if (v < 0) v = 0; }
  return v;
}
template <typename K>
static int resident_ctas(K kern, size_t smem) {
  if (ew_ctas_override()) return ew_ctas_override();
  int n = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kern, 256, smem) != cudaSuccess || n < 1) n = 2;
  return n;
}
static int plane_chunks(int N, int HW, int C, int ctas_per_sm) {
  const int lanes = 256 / (C / 8);
  int chunks = (ctas_per_sm * NUM_SMS) / N;                 // floor: never spill into a second wave
  const int max_chunks = (HW + lanes * 4 - 1) / (lanes * 4);
  if (chunks > max_chunks) chunks = max_chunks;
  if (chunks < 1) chunks = 1;
  return chunks;
}
static dim3 plane_grid(int N, int HW, int C, int ctas_per_sm) {
  return dim3((unsigned)plane_chunks(N, HW, C, ctas_per_sm), (unsigned)N);
}

// per-image element counts must fit 32-bit offsets (padded extents included)
static bool image_fits(int H, int W, int C, int pad) {
  return (long long)(H + 2 * pad) * (W + 2 * pad) * C < (1ll << 31) && (long long)H * W < (1 << 24);
}

// `parts` != 0: launch (the caller's buffer has that many slots per image; -1012 on a mismatch);
// `parts` == 0: only report the number of slots this instantiation produces.
static int reduce_launch(bool bwd, const void* x, const void* g, const float* ss, int N, int H, int W,
                         int C, int gp, int act, float slope, float* out, int parts, cudaStream_t st) {
  const int HW = H * W;
  if (C % 8 || C / 8 > 256 || N < 1) return -1003;
  const int V = C / 8;
  const int lanes = 256 / V;
  const size_t smem = (size_t)lanes * V * 16 * sizeof(float);
  const FastDiv dw = make_fastdiv(W);
  if (!image_fits(H, W, C, gp)) return -1005;
#define MT_RED(B, A, G)                                                                           \
  do {                                                                                            \
    const int chunks = plane_chunks(N, HW, C, resident_ctas(plane_reduce_kernel<B, A, G>, smem)); \
    if (parts == 0) return chunks;                                                                \
    if (parts != chunks) return -1012;                                                            \
    launch_pdl(plane_reduce_kernel<B, A, G>, dim3(N * chunks), dim3(256), smem, st,              \
               (const __nv_bfloat16*)x, (const __nv_bfloat16*)g, ss, H, W, dw, C, gp, chunks, slope, out); \
  } while (0)
#define MT_RED_G(A) do { if (gp) MT_RED(true, A, true); else MT_RED(true, A, false); } while (0)
  if (!bwd) MT_RED(false, 0, false);
  else if (act == MT_ACT_RELU) MT_RED_G(1);
  else if (act == MT_ACT_LRELU) MT_RED_G(2);
  else MT_RED_G(0);
#undef MT_RED_G
#undef MT_RED
  return cuda_rc();
}

// number of partial-sum slots per image that mt_plane_stats writes for this shape
extern "C" int mt_plane_stats_parts(int N, int HW, int C) {
  const int r = reduce_launch(false, nullptr, nullptr, nullptr, N, HW, 1, C, 0, 0, 0.f, nullptr, 0, nullptr);
  return r < 0 ? 0 : r;
}
extern "C" int mt_plane_stats(const void* x, int N, int HW, int C, float* stats, int parts,
                              void* stream) {
  if (parts < 1) return -1012;
  return reduce_launch(false, x, nullptr, nullptr, N, HW, 1, C, 0, 0, 0.f, stats, parts, ST(stream));
}
extern "C" int mt_norm_finalize(const float* stats, int parts, int mode, int N, int C, int C_valid, int HW, float eps,
                                const float* gb, const float* ln_w, const float* ln_b, float* ss,
                                float* mr, float* csum, void* stream) {
  if (mode == MT_NORM_ADAIN && !gb) return -1008;
  if (C_valid <= 0 || C_valid > C || parts < 1) return -1005;
  const int threads = parts > 1 ? 1024 : 256;
  const int splits = (parts > 1 && mode != MT_NORM_LAYER && C > FIN_CSPLIT) ? (C + FIN_CSPLIT - 1) / FIN_CSPLIT : 1;
  launch_pdl(norm_finalize_kernel, dim3(N, splits), dim3(threads), (size_t)C * 2 * sizeof(float), ST(stream), stats, parts, mode, N, C, C_valid, HW, eps, gb, ln_w, ln_b, ss, mr, csum);
  return cuda_rc();
}
extern "C" int mt_affine_act_fwd(const void* x, const float* ss, int N, int H, int W, int C,
                                 int act, float slope, const void* res_pad, int res_p,
                                 const float* noise, void* out_pad, int out_p, void* stream) {
  if (C % 8 || C / 8 > 256) return -1003;
  if (!image_fits(H, W, C, res_p > out_p ? res_p : out_p)) return -1005;
  const FastDiv dw = make_fastdiv(W);
#define MT_AFF3(A, R, Z)                                                                          \
  launch_pdl(affine_act_fwd_kernel<A, R, Z>, plane_grid(N, H * W, C, resident_ctas(affine_act_fwd_kernel<A, R, Z>, 0)), \
             dim3(256), 0, ST(stream),                                                            \
             (const __nv_bfloat16*)x, ss, H, W, dw, C, slope, (const __nv_bfloat16*)res_pad, res_p, noise, \
             (__nv_bfloat16*)out_pad, out_p)
#define MT_AFF(A)                                                                                 \
  do {                                                                                            \
    if (res_pad && noise) MT_AFF3(A, true, true);                                                 \
    else if (res_pad) MT_AFF3(A, true, false);                                                    \
    else if (noise) MT_AFF3(A, false, true);                                                      \
    else MT_AFF3(A, false, false);                                                                \
  } while (0)
  switch (act) {
    case MT_ACT_RELU: MT_AFF(1); break;
    case MT_ACT_LRELU: MT_AFF(2); break;
    case MT_ACT_TANH: MT_AFF(3); break;
    default: MT_AFF(0); break;
  }
#undef MT_AFF
#undef MT_AFF3
  return cuda_rc();
}
extern "C" int mt_norm_bwd_parts(int N, int H, int W, int C, int act, int g_pad) {
  const int r = reduce_launch(true, nullptr, nullptr, nullptr, N, H, W, C, g_pad, act, 0.f, nullptr, 0, nullptr);
  return r < 0 ? 0 : r;
}
extern "C" int mt_norm_bwd_reduce(const void* g, int g_pad, const void* x, const float* ss, int N, int H,
                                  int W, int C, int act, float slope, float* sums, int parts, void* stream) {
  if (parts < 1) return -1012;
  return reduce_launch(true, x, g, ss, N, H, W, C, g_pad, act, slope, sums, parts, ST(stream));
}
extern "C" int mt_norm_bwd_finalize(const float* sums, int parts, const float* mr, int mode, int N, int C,
                                    int C_valid, int HW, const float* gb, const float* ln_w, float* coef,
                                    float* dgb, float* red, void* stream) {
  if (mode == MT_NORM_ADAIN && (!gb || !dgb)) return -1008;
  if (C_valid <= 0 || C_valid > C || parts < 1) return -1005;
  const int threads = parts > 1 ? 1024 : 256;
  const int splits = (parts > 1 && mode != MT_NORM_LAYER && C > FIN_CSPLIT) ? (C + FIN_CSPLIT - 1) / FIN_CSPLIT : 1;
  launch_pdl(norm_bwd_finalize_kernel, dim3(N, splits), dim3(threads), (size_t)C * 2 * sizeof(float), ST(stream),
             sums, parts, mr, mode, N, C, C_valid, HW, gb, ln_w, coef, dgb, red);
  return cuda_rc();
}
extern "C" int mt_norm_bwd_params(const float* red, const float* coef, const float* csum, int N, int C,
                                  int C_valid, int HW, float* dln_w, float* dln_b, float* dbias,
                                  int accumulate, void* stream) {
  if (C_valid <= 0 || C_valid > C || !red) return -1005;
  if ((dln_w == nullptr) != (dln_b == nullptr)) return -1008;
  if (dbias && (!coef || !csum)) return -1008;
  if (!dln_w && !dbias) return 0;
  launch_pdl(norm_bwd_params_kernel, dim3((C_valid + 127) / 128), dim3(128), 0, ST(stream), red, coef, csum, N, C,
             C_valid, HW, dln_w, dln_b, dbias, accumulate);
  return cuda_rc();
}
extern "C" int mt_norm_bwd_apply(const void* g, int g_pad, const void* x, const float* ss,
                                 const float* coef, int N, int H, int W, int C, int act, float slope,
                                 void* dx, void* stream) {
  if (C % 8 || C / 8 > 256) return -1003;
  if (!image_fits(H, W, C, g_pad)) return -1005;
  const FastDiv dw = make_fastdiv(W);
#define MT_APP2(A, G)                                                                             \
  launch_pdl(norm_bwd_apply_kernel<A, G>, plane_grid(N, H * W, C, resident_ctas(norm_bwd_apply_kernel<A, G>, 0)), \
             dim3(256), 0, ST(stream), (const __nv_bfloat16*)g, (const __nv_bfloat16*)x, ss, coef, H, \
             W, dw, C, g_pad, slope, (__nv_bfloat16*)dx)
#define MT_APP(A) do { if (g_pad) MT_APP2(A, true); else MT_APP2(A, false); } while (0)
  if (act == MT_ACT_RELU) MT_APP(1);
  else if (act == MT_ACT_LRELU) MT_APP(2);
  else MT_APP(0);
#undef MT_APP
#undef MT_APP2
  return cuda_rc();
}
extern "C" int mt_act_bwd_fold(const void* dpad, const void* out_pad, int N, int H, int W, int C,
                               int pad, int act, float slope, void* dy, void* stream) {
  if (C % 8 || C / 8 > 256) return -1003;
  if (!image_fits(H, W, C, pad)) return -1005;
  const FastDiv dw = make_fastdiv(W);
#define MT_ABF(A)                                                                                 \
  launch_pdl(act_bwd_fold_kernel<A>, plane_grid(N, H * W, C, resident_ctas(act_bwd_fold_kernel<A>, 0)), dim3(256), 0, \
             ST(stream), (const __nv_bfloat16*)dpad, (const __nv_bfloat16*)out_pad, H, W, \
             dw, C, pad, slope, (__nv_bfloat16*)dy)
  if (act == MT_ACT_RELU) MT_ABF(1);
  else if (act == MT_ACT_LRELU) MT_ABF(2);
  else MT_ABF(0);
#undef MT_ABF
  return cuda_rc();
}
