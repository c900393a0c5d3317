// Remaining HBM-bound pieces of the path: the style encoder's pooling / shortcut algebra
// (blocks.py:93-119, networks.py:123-124), the RGB head (ConvTranspose2d 1x1 -> tanh,
// networks.py:248-251) and the content-code L2 regulariser (adain_model.py:396-399).
#include "../../include/mt_b200.h"
#include "ew_common.cuh"

namespace mt {

// out[n,h2,w2,:] = mean of the 2x2 window of the interior of x_pad        (nn.AvgPool2d(2,2))
__global__ void __launch_bounds__(256)
avgpool2_fwd_kernel(const __nv_bfloat16* __restrict__ x, int p, int N, int H, int W, int C,
                    __nv_bfloat16* __restrict__ out) {
  pdl_wait(); pdl_launch();
  const int Ho = H / 2, Wo = W / 2, V = C / 8, Hp = H + 2 * p, Wp = W + 2 * p;
  const long long total = (long long)N * Ho * Wo * V;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int v = (int)(i % V); unsigned t = i / V;
    const int w = (int)(t % Wo); t /= Wo;
    const int h = (int)(t % Ho); const int n = (int)(t / Ho);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int dy = 0; dy < 2; ++dy)
#pragma unroll
      for (int dx = 0; dx < 2; ++dx) {
        float f[8];
        load8(x + (((long long)n * Hp + 2 * h + dy + p) * Wp + 2 * w + dx + p) * C + v * 8, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
      }
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] *= 0.25f;
    store8(out + (size_t)i * 8, acc);
  }
}

// dpad[n, h+p, w+p, :] += 0.25 * dpooled[n, h/2, w/2, :]
__global__ void __launch_bounds__(256)
avgpool2_bwd_add_kernel(const __nv_bfloat16* __restrict__ dpo, __nv_bfloat16* __restrict__ dpad,
                        int p, int N, int H, int W, int C) {
  pdl_wait(); pdl_launch();
  const int Ho = H / 2, Wo = W / 2, V = C / 8, Hp = H + 2 * p, Wp = W + 2 * p;
  const long long total = (long long)N * H * W * V;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int v = (int)(i % V); unsigned t = i / V;
    const int w = (int)(t % W); t /= W;
    const int h = (int)(t % H); const int n = (int)(t / H);
    if (h / 2 >= Ho || w / 2 >= Wo) continue;
    float g[8], a[8];
    load8(dpo + (((long long)n * Ho + h / 2) * Wo + w / 2) * C + v * 8, g);
    __nv_bfloat16* d = dpad + (((long long)n * Hp + h + p) * Wp + w + p) * C + v * 8;
    load8(d, a);
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] += 0.25f * g[j];
    store8(d, a);
  }
}

// out_pad[interior] = act(avgpool2(c) + d)          (DownResnetBlock tail + next LeakyReLU)
__global__ void __launch_bounds__(256)
pool_add_act_fwd_kernel(const __nv_bfloat16* __restrict__ c, const __nv_bfloat16* __restrict__ d,
                        int N, int H, int W, int C, int act, float slope,
                        __nv_bfloat16* __restrict__ out, int po) {
  pdl_wait(); pdl_launch();
  const int Ho = H / 2, Wo = W / 2, V = C / 8;
  const long long total = (long long)N * Ho * Wo * V;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int v = (int)(i % V); unsigned t = i / V;
    const int w = (int)(t % Wo); t /= Wo;
    const int h = (int)(t % Ho); const int n = (int)(t / Ho);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int dy = 0; dy < 2; ++dy)
#pragma unroll
      for (int dx = 0; dx < 2; ++dx) {
        float f[8];
        load8(c + (((long long)n * H + 2 * h + dy) * W + 2 * w + dx) * C + v * 8, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
      }
    float dv[8];
    load8(d + (size_t)i * 8, dv);
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = act_fwd(acc[j] * 0.25f + dv[j], act, slope);
    store8(out + (((long long)n * (Ho + 2 * po) + h + po) * (Wo + 2 * po) + w + po) * C + v * 8, acc);
  }
}

// g = fold(dout_pad) * act'(out);  dd = g (half res);  dc = 0.25 * g upsampled (full res)
__global__ void __launch_bounds__(256)
pool_add_act_bwd_kernel(const __nv_bfloat16* __restrict__ dout, const __nv_bfloat16* __restrict__ outp,
                        int N, int H, int W, int C, int po, int act, float slope,
                        __nv_bfloat16* __restrict__ dd, __nv_bfloat16* __restrict__ dc) {
  pdl_wait(); pdl_launch();
  const int Ho = H / 2, Wo = W / 2, V = C / 8, Hp = Ho + 2 * po, Wp = Wo + 2 * po;
  const long long total = (long long)N * Ho * Wo * V;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int v = (int)(i % V); unsigned t = i / V;
    const int w = (int)(t % Wo); t /= Wo;
    const int h = (int)(t % Ho); const int n = (int)(t / Ho);
    int hs[3], ws[3];
    const int nh = reflect_fold_sources(h, Ho, po, hs), nw = reflect_fold_sources(w, Wo, po, ws);
    float g[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int a = 0; a < nh; ++a)
      for (int b = 0; b < nw; ++b) {
        float f[8];
        load8(dout + (((long long)n * Hp + hs[a]) * Wp + ws[b]) * C + v * 8, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] += f[j];
      }
    float o[8];
    load8(outp + (((long long)n * Hp + h + po) * Wp + w + po) * C + v * 8, o);
#pragma unroll
    for (int j = 0; j < 8; ++j) g[j] *= act_grad_pre(o[j], act, slope);
    store8(dd + (size_t)i * 8, g);
#pragma unroll
    for (int j = 0; j < 8; ++j) g[j] *= 0.25f;
#pragma unroll
    for (int dy = 0; dy < 2; ++dy)
#pragma unroll
      for (int dx = 0; dx < 2; ++dx)
        store8(dc + (((long long)n * H + 2 * h + dy) * W + 2 * w + dx) * C + v * 8, g);
  }
}

// nn.AdaptiveAvgPool2d(1): x [N,HW,C] bf16 -> out [N,C] fp32
__global__ void __launch_bounds__(256)
gap_fwd_kernel(const __nv_bfloat16* __restrict__ x, int HW, int C, float* __restrict__ out) {
  pdl_wait(); pdl_launch();
  extern __shared__ float red[];
  const int V = C / 8, lanes = blockDim.x / V, n = blockIdx.x;
  const int v = threadIdx.x % V, pl = threadIdx.x / V;
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (pl < lanes)
    for (int p = pl; p < HW; p += lanes) {
      float f[8];
      load8(x + ((long long)n * HW + p) * C + v * 8, f);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += f[j];
    }
  if (pl < lanes)
#pragma unroll
    for (int j = 0; j < 8; ++j) red[(pl * V + v) * 8 + j] = acc[j];
  __syncthreads();
  for (int t = threadIdx.x; t < C; t += blockDim.x) {
    float s = 0.f;
    for (int l = 0; l < lanes; ++l) s += red[l * C + t];
    out[(long long)n * C + t] = s / (float)HW;
  }
}
__global__ void __launch_bounds__(256)
gap_bwd_kernel(const float* __restrict__ dout, int N, int HW, int C, __nv_bfloat16* __restrict__ dx) {
  pdl_wait(); pdl_launch();
  const int V = C / 8;
  const long long total = (long long)N * HW * V;
  const float inv = 1.f / (float)HW;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int v = (int)(i % V);
    const int n = (int)(i / (unsigned)(HW * V));
    float g[8];
    load8f(dout + (long long)n * C + v * 8, g);
#pragma unroll
    for (int j = 0; j < 8; ++j) g[j] *= inv;
    store8(dx + (size_t)i * 8, g);
  }
}

// RGB head: y[n,o,h,w] = tanh(sum_c x[n,h,w,c] * w[c][o]),  o < Co <= 4, NCHW fp32 output.
__global__ void __launch_bounds__(256)
rgb_head_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ wgt, int N,
                    int HW, int C, int Co, float* __restrict__ y) {
  pdl_wait(); pdl_launch();
  extern __shared__ float ws[];  // [C][4]
  for (int i = threadIdx.x; i < C * 4; i += blockDim.x)
    ws[i] = (i & 3) < Co ? wgt[(i >> 2) * Co + (i & 3)] : 0.f;
  __syncthreads();
  const long long total = (long long)N * HW;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int n = (int)(i / HW);
    const int pix = (int)(i % HW);
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    for (int v = 0; v < C / 8; ++v) {
      float f[8];
      load8(x + i * C + v * 8, f);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float4 q = *reinterpret_cast<const float4*>(ws + (v * 8 + j) * 4);
        a0 += f[j] * q.x; a1 += f[j] * q.y; a2 += f[j] * q.z; a3 += f[j] * q.w;
      }
    }
    const float a[4] = {a0, a1, a2, a3};
    for (int o = 0; o < Co; ++o) y[((long long)n * Co + o) * HW + pix] = tanhf(a[o]);
  }
}

// g_o = dy_o*(1-y_o^2); dx[c] = sum_o g_o w[c][o]; dwp[block][c][o] = this block's sum_pix x[c] g_o
// (plain stores; rgb_head_dw_reduce_kernel adds the blocks in index order: reproducible)
__global__ void __launch_bounds__(256)
rgb_head_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ y,
                    const __nv_bfloat16* __restrict__ x, const float* __restrict__ wgt, int N,
                    int HW, int C, int Co, __nv_bfloat16* __restrict__ dx, float* __restrict__ dwp) {
  pdl_wait(); pdl_launch();
  extern __shared__ float sm[];  // ws[C*4] | red[lanes*V*32]
  float* ws = sm;
  float* red = sm + C * 4;
  for (int i = threadIdx.x; i < C * 4; i += blockDim.x)
    ws[i] = (i & 3) < Co ? wgt[(i >> 2) * Co + (i & 3)] : 0.f;
  __syncthreads();
  const int V = C / 8, lanes = blockDim.x / V;
  const int v = threadIdx.x % V, pl = threadIdx.x / V;
  // this thread's 8 channels x 4 outputs of the weight stay in registers: the same shared-memory
  // words read by all 8 channel groups of a warp were an 8-way bank conflict per pixel
  float4 wq[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) wq[j] = *reinterpret_cast<const float4*>(ws + (v * 8 + j) * 4);
  float acc[8][4];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f;
  const long long total = (long long)N * HW;
  if ((32 % V) == 0 && V >= Co && total * C < (1ll << 31)) {
    // fast path: the V threads of a pixel sit in one warp; thread v < Co fetches channel v of dy / y
    // and the group shares the three products by shuffle (instead of 2*Co scalar loads per thread);
    // two pixels per iteration keep two 16-byte loads in flight.
    const int lane = threadIdx.x & 31, grp = (lane / V) * V;
    // 32-bit index arithmetic (a 64-bit division costs more than the payload of one pixel)
    const int stride = (int)gridDim.x * lanes, tot = (int)total;
    const unsigned uHW = (unsigned)HW;
    // loop bound on the warp's FIRST pixel so that every lane takes part in the shuffles
    for (int i0 = (int)blockIdx.x * lanes + pl; i0 - lane / V < tot; i0 += 2 * stride) {
      const int i1 = i0 + stride;
      const bool ok0 = i0 < tot, ok1 = i1 < tot;
      uint4 x0 = make_uint4(0, 0, 0, 0), x1 = make_uint4(0, 0, 0, 0);
      if (ok0) x0 = ldg16(x + i0 * C + v * 8);
      if (ok1) x1 = ldg16(x + i1 * C + v * 8);
      float own0 = 0.f, own1 = 0.f;
      if (v < Co && ok0) {
        const unsigned n0 = (unsigned)i0 / uHW;
        const unsigned k0 = (n0 * Co + v) * uHW + ((unsigned)i0 - n0 * uHW);
        const float yy = __ldg(y + k0);
        own0 = __ldg(dy + k0) * (1.f - yy * yy);
        if (ok1) {
          const unsigned n1 = (unsigned)i1 / uHW;
          const unsigned k1 = (n1 * Co + v) * uHW + ((unsigned)i1 - n1 * uHW);
          const float y1 = __ldg(y + k1);
          own1 = __ldg(dy + k1) * (1.f - y1 * y1);
        }
      }
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const float own = u ? own1 : own0;
        float g[4];
#pragma unroll
        for (int o = 0; o < 4; ++o) g[o] = __shfl_sync(0xffffffffu, own, grp + (o < V ? o : 0));
#pragma unroll
        for (int o = 0; o < 4; ++o) if (o >= Co) g[o] = 0.f;
        if (u ? !ok1 : !ok0) continue;
        float f[8], d[8];
        cvt8(u ? x1 : x0, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float4 q = wq[j];
          d[j] = g[0] * q.x + g[1] * q.y + g[2] * q.z + g[3] * q.w;
          acc[j][0] += f[j] * g[0]; acc[j][1] += f[j] * g[1];
          acc[j][2] += f[j] * g[2]; acc[j][3] += f[j] * g[3];
        }
        store8(dx + (long long)(u ? i1 : i0) * C + v * 8, d);
      }
    }
  } else if (pl < lanes)
    for (long long i = (long long)blockIdx.x * lanes + pl; i < total; i += (long long)gridDim.x * lanes) {
      const int n = (int)(i / HW);
      const int pix = (int)(i % HW);
      float g[4] = {0, 0, 0, 0};
      for (int o = 0; o < Co; ++o) {
        const long long k = ((long long)n * Co + o) * HW + pix;
        const float yy = __ldg(y + k);
        g[o] = __ldg(dy + k) * (1.f - yy * yy);
      }
      float f[8], d[8];
      load8(x + i * C + v * 8, f);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float4 q = wq[j];
        d[j] = g[0] * q.x + g[1] * q.y + g[2] * q.z + g[3] * q.w;
        acc[j][0] += f[j] * g[0]; acc[j][1] += f[j] * g[1];
        acc[j][2] += f[j] * g[2]; acc[j][3] += f[j] * g[3];
      }
      store8(dx + i * C + v * 8, d);
    }
  if (pl < lanes)
#pragma unroll
    for (int j = 0; j < 8; ++j)
#pragma unroll
      for (int o = 0; o < 4; ++o) red[(pl * V + v) * 32 + j * 4 + o] = acc[j][o];
  __syncthreads();
  for (int t = threadIdx.x; t < V * 32; t += blockDim.x) {
    float s = 0.f;
    for (int l = 0; l < lanes; ++l) s += red[l * V * 32 + t];
    const int c = (t / 32) * 8 + (t % 32) / 4, o = t & 3;
    dwp[((size_t)blockIdx.x * C + c) * 4 + o] = s;
  }
}

// ---- RGB head, warp-tiled form (C / 8 = V in {4, 8, 16}).  A warp owns 32 consecutive pixels and walks them
// in V steps of 32 / V pixels: lane = (pixel of the step, channel vector), so every 16-byte activation load
// and store is part of a contiguous 512-byte warp access, and the NCHW fp32 side (y, dy: 3 planes) is touched
// as one coalesced 128-byte row per plane (lane = pixel), shuffled between the two lane mappings.  The
// thread-per-pixel kernels above moved 134 MB at 1.8 - 2.1 TB/s (16-byte loads 128 bytes apart, scattered
// 4-byte plane accesses); they remain the fallback for other channel counts.
template <int V, int Co>
__global__ void __launch_bounds__(256)
rgb_head_fwd_warp_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ wgt, long long total,
                         int HW, float* __restrict__ y) {
  pdl_wait(); pdl_launch();
  constexpr int C = V * 8, P = 32 / V;
  const int lane = threadIdx.x & 31, v = lane % V, q = lane / V;
  float4 wq[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float* wr = wgt + (v * 8 + j) * Co;
    wq[j] = make_float4(__ldg(wr), Co > 1 ? __ldg(wr + 1) : 0.f, Co > 2 ? __ldg(wr + 2) : 0.f, Co > 3 ? __ldg(wr + 3) : 0.f);
  }
  const long long warps = (long long)gridDim.x * (blockDim.x >> 5);
  for (long long base = ((long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * 32; base < total;
       base += warps * 32) {
    uint4 xr[V];
#pragma unroll
    for (int s = 0; s < V; ++s) {
      const long long pix = base + s * P + q;
      xr[s] = pix < total ? ldg16(x + pix * C + v * 8) : make_uint4(0, 0, 0, 0);
    }
    float r0 = 0.f, r1 = 0.f, r2 = 0.f, r3 = 0.f;
#pragma unroll
    for (int s = 0; s < V; ++s) {
      float f[8];
      cvt8(xr[s], f);
      float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) {        // Co is a template constant: the FMA pipe, not HBM, bounds this loop
        a0 = fmaf(f[j], wq[j].x, a0);
        if (Co > 1) a1 = fmaf(f[j], wq[j].y, a1);
        if (Co > 2) a2 = fmaf(f[j], wq[j].z, a2);
        if (Co > 3) a3 = fmaf(f[j], wq[j].w, a3);
      }
#pragma unroll
      for (int o = 1; o < V; o <<= 1) {
        a0 += __shfl_xor_sync(0xffffffffu, a0, o);
        if (Co > 1) a1 += __shfl_xor_sync(0xffffffffu, a1, o);
        if (Co > 2) a2 += __shfl_xor_sync(0xffffffffu, a2, o);
        if (Co > 3) a3 += __shfl_xor_sync(0xffffffffu, a3, o);
      }
      // pixel (s * P + q) of the warp's 32 belongs to lane s * P + q: fetch it from the group's first lane
      const int src = (lane % P) * V;
      float t0 = __shfl_sync(0xffffffffu, a0, src), t1 = 0.f, t2 = 0.f, t3 = 0.f;
      if (Co > 1) t1 = __shfl_sync(0xffffffffu, a1, src);
      if (Co > 2) t2 = __shfl_sync(0xffffffffu, a2, src);
      if (Co > 3) t3 = __shfl_sync(0xffffffffu, a3, src);
      if (lane / P == s) { r0 = t0; r1 = t1; r2 = t2; r3 = t3; }
    }
    const long long pix = base + lane;
    if (pix < total) {
      const long long n = pix / HW, p = pix - n * HW;
      float* yo = y + n * Co * HW + p;
      yo[0] = tanhf(r0);
      if (Co > 1) yo[HW] = tanhf(r1);
      if (Co > 2) yo[2LL * HW] = tanhf(r2);
      if (Co > 3) yo[3LL * HW] = tanhf(r3);
    }
  }
}

template <int V, int Co>
__global__ void __launch_bounds__(256, 2)      // 2 CTAs / SM: 16 warps to cover the load latency
rgb_head_bwd_warp_kernel(const float* __restrict__ dy, const float* __restrict__ y,
                         const __nv_bfloat16* __restrict__ x, const float* __restrict__ wgt, long long total,
                         int HW, __nv_bfloat16* __restrict__ dx, float* __restrict__ dwp) {
  pdl_wait(); pdl_launch();
  constexpr int C = V * 8, P = 32 / V;
  extern __shared__ float red[];               // [256 / V][V][32]
  const int lane = threadIdx.x & 31, v = lane % V, q = lane / V;
  float wq[8][Co], acc[8][Co];
#pragma unroll
  for (int j = 0; j < 8; ++j)
#pragma unroll
    for (int o = 0; o < Co; ++o) { wq[j][o] = __ldg(wgt + (v * 8 + j) * Co + o); acc[j][o] = 0.f; }
  const long long warps = (long long)gridDim.x * (blockDim.x >> 5);
  for (long long base = ((long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * 32; base < total;
       base += warps * 32) {
    uint4 xr[V];
#pragma unroll
    for (int s = 0; s < V; ++s) {
      const long long pix = base + s * P + q;
      xr[s] = pix < total ? ldg16(x + pix * C + v * 8) : make_uint4(0, 0, 0, 0);
    }
    // lane = pixel: g_o = dy_o * (1 - y_o^2), one coalesced row per plane
    float G[4] = {0.f, 0.f, 0.f, 0.f};
    {
      const long long pix = base + lane;
      if (pix < total) {
        const long long n = pix / HW, p = pix - n * HW;
        const float* yo = y + n * Co * HW + p;
        const float* go = dy + n * Co * HW + p;
#pragma unroll
        for (int o = 0; o < 4; ++o)
          if (o < Co) { const float yy = __ldg(yo + (long long)o * HW); G[o] = __ldg(go + (long long)o * HW) * (1.f - yy * yy); }
      }
    }
#pragma unroll
    for (int s = 0; s < V; ++s) {
      const int src = s * P + q;               // the lane that holds this step's pixel
      float g[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int o = 0; o < Co; ++o) g[o] = __shfl_sync(0xffffffffu, G[o], src);
      const long long pix = base + s * P + q;
      if (pix < total) {
        float f[8], d[8];
        cvt8(xr[s], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) {      // 2 * Co FMAs per element: the FMA pipe, not HBM, bounds this loop
          float dj = g[0] * wq[j][0];
#pragma unroll
          for (int o = 1; o < Co; ++o) dj = fmaf(g[o], wq[j][o], dj);
          d[j] = dj;
#pragma unroll
          for (int o = 0; o < Co; ++o) acc[j][o] = fmaf(f[j], g[o], acc[j][o]);
        }
        store8(dx + pix * C + v * 8, d);
      }
    }
  }
  const int pl = threadIdx.x / V, lanes = blockDim.x / V;
#pragma unroll
  for (int j = 0; j < 8; ++j)
#pragma unroll
    for (int o = 0; o < 4; ++o) red[(pl * V + v) * 32 + j * 4 + o] = o < Co ? acc[j][o < Co ? o : 0] : 0.f;
  __syncthreads();
  for (int t = threadIdx.x; t < V * 32; t += blockDim.x) {
    float s = 0.f;
    for (int l = 0; l < lanes; ++l) s += red[l * V * 32 + t];
    const int c = (t / 32) * 8 + (t % 32) / 4, o = t & 3;
    dwp[((size_t)blockIdx.x * C + c) * 4 + o] = s;
  }
}

// dw[c][o] = sum over the per-block partials, block = 32 (c, o) columns x 8 groups of partials, the groups
// combined in index order (fixed summation order)
__global__ void __launch_bounds__(256)
rgb_head_dw_reduce_kernel(const float* __restrict__ dwp, int parts, int C, int Co, float* __restrict__ dw) {
  pdl_wait(); pdl_launch();
  __shared__ float sm[8][33];
  const int cl = threadIdx.x & 31, g = threadIdx.x >> 5;
  const int t = blockIdx.x * 32 + cl;                      // (c, o) with o < 4
  float a[4] = {0.f, 0.f, 0.f, 0.f};
  if (t < C * 4) {
    int i = g;
    for (; i + 24 < parts; i += 32) {
#pragma unroll
      for (int u = 0; u < 4; ++u) a[u] += dwp[(size_t)(i + 8 * u) * C * 4 + t];
    }
    for (; i < parts; i += 8) a[0] += dwp[(size_t)i * C * 4 + t];
  }
  sm[g][cl] = (a[0] + a[1]) + (a[2] + a[3]);
  __syncthreads();
  if (g == 0 && t < C * 4) {
    float v = sm[0][cl];
#pragma unroll
    for (int k = 1; k < 8; ++k) v += sm[k][cl];
    const int c = t >> 2, o = t & 3;
    if (o < Co) dw[c * Co + o] = v;
  }
}

// sum of squares over the interior of a padded buffer (per-block partial sums, added in block order by
// sum_partials_kernel: reproducible) and its gradient
__global__ void __launch_bounds__(256)
sumsq_interior_kernel(const __nv_bfloat16* __restrict__ x, int N, int H, int W, int C, int p,
                      float* __restrict__ part) {
  pdl_wait(); pdl_launch();
  const int V = C / 8, Hp = H + 2 * p, Wp = W + 2 * p;
  const long long total = (long long)N * H * W * V;
  float s = 0.f;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int v = (int)(i % V); unsigned t = i / V;
    const int w = (int)(t % W); t /= W;
    const int h = (int)(t % H); const int n = (int)(t / H);
    float f[8];
    load8(x + (((long long)n * Hp + h + p) * Wp + w + p) * C + v * 8, f);
#pragma unroll
    for (int j = 0; j < 8; ++j) s += f[j] * f[j];
  }
  for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  __shared__ float r[8];
  if ((threadIdx.x & 31) == 0) r[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < 8; ++i) t += r[i];
    part[blockIdx.x] = t;
  }
}
// out[0] = sum of part[0..n) in a fixed order (one block of 256 threads, strided partial sums, then a
// fixed shuffle / shared-memory tree)
__global__ void __launch_bounds__(256) sum_partials_kernel(const float* __restrict__ part, int n,
                                                           float* __restrict__ out) {
  pdl_wait(); pdl_launch();
  float s = 0.f;
  for (int i = threadIdx.x; i < n; i += 256) s += part[i];
  for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  __shared__ float r[8];
  if ((threadIdx.x & 31) == 0) r[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < 8; ++i) t += r[i];
    out[0] = t;
  }
}
// db[c] (+)= sum over the rows of stats[row][c][0]  (column sums of a dense tensor from the per-chunk
// partial sums of mt_plane_stats).  Block = 32 channels x 8 row groups: a thread adds the rows r = g (mod 8)
// with four loads in flight, the eight groups are then combined in index order (a fixed summation order;
// one thread walking all rows was a chain of ~300 dependent L2 round trips: 36 us for a 2 us job).
__global__ void __launch_bounds__(256) colsum_finalize_kernel(const float* __restrict__ stats, int rows, int C,
                                                              int Cv, float* __restrict__ db, int accumulate) {
  pdl_wait(); pdl_launch();
  __shared__ float sm[8][33];
  const int cl = threadIdx.x & 31, g = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + cl;
  // blockIdx.y: independent group of `rows` rows with its own output row (mt_colsum_images: one per image)
  stats += (size_t)blockIdx.y * rows * C * 2;
  db += (size_t)blockIdx.y * Cv;
  float a[4] = {0.f, 0.f, 0.f, 0.f};
  if (c < Cv) {
    int i = g;
    for (; i + 24 < rows; i += 32) {
#pragma unroll
      for (int u = 0; u < 4; ++u) a[u] += stats[((size_t)(i + 8 * u) * C + c) * 2];
    }
    for (; i < rows; i += 8) a[0] += stats[((size_t)i * C + c) * 2];
  }
  sm[g][cl] = (a[0] + a[1]) + (a[2] + a[3]);
  __syncthreads();
  if (g == 0 && c < Cv) {
    float v = sm[0][cl];
#pragma unroll
    for (int k = 1; k < 8; ++k) v += sm[k][cl];
    db[c] = accumulate ? db[c] + v : v;
  }
}

// dx_pad = scale[0] * x_pad on the interior, 0 on the border
__global__ void __launch_bounds__(256)
scale_interior_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ scale, int N,
                      int H, int W, int C, int p, __nv_bfloat16* __restrict__ dx) {
  pdl_wait(); pdl_launch();
  const int V = C / 8, Hp = H + 2 * p, Wp = W + 2 * p;
  const long long total = (long long)N * Hp * Wp * V;
  const float sc = __ldg(scale);
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    unsigned t = i / V;
    const int wp = (int)(t % Wp); t /= Wp;
    const int hp = (int)(t % Hp);
    float f[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (hp >= p && hp < H + p && wp >= p && wp < W + p) {
      load8(x + (size_t)i * 8, f);
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] *= sc;
    }
    store8(dx + (size_t)i * 8, f);
  }
}

// ---- optional generator variants ------------------------------------------------
// nn.Upsample(scale_factor=2, mode='nearest') (blocks.py:75): out[n, y, x, :] = in[n, y/2, x/2, :].
// One 16-byte vector per thread, indexed by the OUTPUT (coalesced stores; each input vector is read
// by four threads, three of them out of L1/L2).
__global__ void __launch_bounds__(256)
upsample2_fwd_kernel(const __nv_bfloat16* __restrict__ x, int N, int H, int W, int C,
                     __nv_bfloat16* __restrict__ out) {
  pdl_wait(); pdl_launch();
  const int V = C / 8, Ho = 2 * H, Wo = 2 * W;
  const long long total = (long long)N * Ho * Wo * V;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int v = (int)(i % V); unsigned t = i / V;
    const int w = (int)(t % Wo); t /= Wo;
    const int h = (int)(t % Ho); const int n = (int)(t / Ho);
    const uint4 r = *reinterpret_cast<const uint4*>(
        x + (((long long)n * H + h / 2) * W + w / 2) * C + v * 8);
    *reinterpret_cast<uint4*>(out + (size_t)i * 8) = r;
  }
}
// adjoint: dx[n, h, w, :] = sum of the 2x2 block of dy (fp32 sum, one rounding)
__global__ void __launch_bounds__(256)
upsample2_bwd_kernel(const __nv_bfloat16* __restrict__ dy, int N, int H, int W, int C,
                     __nv_bfloat16* __restrict__ dx) {
  pdl_wait(); pdl_launch();
  const int V = C / 8, Wo = 2 * W;
  const long long total = (long long)N * H * W * V;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int v = (int)(i % V); unsigned t = i / V;
    const int w = (int)(t % W); t /= W;
    const int h = (int)(t % H); const int n = (int)(t / H);
    const __nv_bfloat16* p = dy + (((long long)n * 2 * H + 2 * h) * Wo + 2 * w) * C + v * 8;
    const uint4 r0 = ldg16(p), r1 = ldg16(p + C), r2 = ldg16(p + (size_t)Wo * C),
                r3 = ldg16(p + (size_t)Wo * C + C);
    float a[8], b[8], c[8], d[8];
    cvt8(r0, a); cvt8(r1, b); cvt8(r2, c); cvt8(r3, d);
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] = (a[j] + b[j]) + (c[j] + d[j]);
    store8(dx + (size_t)i * 8, a);
  }
}
// nn.Dropout(0.5) (blocks.py:133-134,152,190) with the mask drawn by the host framework:
// out = x * m, m in {0, 1/(1-p)} stored as bf16 (exact); the same kernel is its own adjoint (g * m).
__global__ void __launch_bounds__(256)
mul_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ m, long long nvec,
           __nv_bfloat16* __restrict__ out) {
  pdl_wait(); pdl_launch();
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nvec;
       i += (long long)gridDim.x * blockDim.x) {
    const uint4 rx = ldg16(x + i * 8), rm = ldg16(m + i * 8);
    float a[8], b[8];
    cvt8(rx, a); cvt8(rm, b);
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] *= b[j];
    store8(out + i * 8, a);
  }
}

}  // namespace mt

using namespace mt;
#define ST(s) ((cudaStream_t)(s))
#define BF(p) ((__nv_bfloat16*)(p))
#define CBF(p) ((const __nv_bfloat16*)(p))

extern "C" int mt_upsample2_fwd(const void* x, int N, int H, int W, int C, void* out, void* stream) {
  if (C % 8) return -1003;
  launch_pdl(upsample2_fwd_kernel, dim3(grid_for((long long)N * H * W * 4 * (C / 8), 256)), dim3(256), 0, ST(stream), 
      CBF(x), N, H, W, C, BF(out));
  return cuda_rc();
}
extern "C" int mt_upsample2_bwd(const void* dy, int N, int H, int W, int C, void* dx, void* stream) {
  if (C % 8) return -1003;
  launch_pdl(upsample2_bwd_kernel, dim3(grid_for((long long)N * H * W * (C / 8), 256)), dim3(256), 0, ST(stream), 
      CBF(dy), N, H, W, C, BF(dx));
  return cuda_rc();
}
extern "C" int mt_mul(const void* x, const void* m, long long n, void* out, void* stream) {
  if (n % 8) return -1003;
  launch_pdl(mul_kernel, dim3(grid_for(n / 8, 256)), dim3(256), 0, ST(stream), CBF(x), CBF(m), n / 8, BF(out));
  return cuda_rc();
}
extern "C" int mt_avgpool2_fwd(const void* x_pad, int pad, int N, int H, int W, int C, void* out,
                               void* stream) {
  if (C % 8) return -1003;
  launch_pdl(avgpool2_fwd_kernel, dim3(grid_for((long long)N * (H / 2) * (W / 2) * (C / 8), 256)), dim3(256), 0, ST(stream), CBF(x_pad), pad, N, H, W, C, BF(out));
  return cuda_rc();
}
extern "C" int mt_avgpool2_bwd_add(const void* dpooled, void* dpad, int pad, int N, int H, int W,
                                   int C, void* stream) {
  if (C % 8) return -1003;
  launch_pdl(avgpool2_bwd_add_kernel, dim3(grid_for((long long)N * H * W * (C / 8), 256)), dim3(256), 0, ST(stream), 
      CBF(dpooled), BF(dpad), pad, N, H, W, C);
  return cuda_rc();
}
extern "C" int mt_pool_add_act_fwd(const void* c, const void* d, int N, int H, int W, int C,
                                   int act, float slope, void* out_pad, int out_p, void* stream) {
  if (C % 8) return -1003;
  launch_pdl(pool_add_act_fwd_kernel, dim3(grid_for((long long)N * (H / 2) * (W / 2) * (C / 8), 256)), dim3(256), 0, ST(stream), CBF(c), CBF(d), N, H, W, C, act, slope, BF(out_pad), out_p);
  return cuda_rc();
}
extern "C" int mt_pool_add_act_bwd(const void* dout_pad, const void* out_pad, int N, int H, int W,
                                   int C, int out_p, int act, float slope, void* dd, void* dc,
                                   void* stream) {
  if (C % 8) return -1003;
  launch_pdl(pool_add_act_bwd_kernel, dim3(grid_for((long long)N * (H / 2) * (W / 2) * (C / 8), 256)), dim3(256), 0, ST(stream), CBF(dout_pad), CBF(out_pad), N, H, W, C, out_p, act, slope,
                                          BF(dd), BF(dc));
  return cuda_rc();
}
extern "C" int mt_gap_fwd(const void* x, int N, int HW, int C, float* out, void* stream) {
  if (C % 8 || C / 8 > 256) return -1003;
  const int lanes = 256 / (C / 8);
  launch_pdl(gap_fwd_kernel, dim3(N), dim3(256), (size_t)lanes * C * sizeof(float), ST(stream), CBF(x), HW, C, out);
  return cuda_rc();
}
extern "C" int mt_gap_bwd(const float* dout, int N, int HW, int C, void* dx, void* stream) {
  if (C % 8) return -1003;
  launch_pdl(gap_bwd_kernel, dim3(grid_for((long long)N * HW * (C / 8), 256)), dim3(256), 0, ST(stream), dout, N, HW, C,
                                                                                     BF(dx));
  return cuda_rc();
}
extern "C" int mt_rgb_head_fwd(const void* x, const float* w, int N, int HW, int C, int Co,
                               float* y, void* stream) {
  if (C % 8 || Co > 4 || Co < 1) return -1003;
  const long long total = (long long)N * HW;
  const int wgrid = grid_for((total + 255) / 256, 1, 16);     // 8 warps x 32 pixels per block and pass
#define MT_RGB_F(V_, CO_) launch_pdl(rgb_head_fwd_warp_kernel<V_, CO_>, dim3(wgrid), dim3(256), 0, ST(stream), CBF(x), w, total, HW, y); return cuda_rc()
  switch ((C / 8) * 10 + Co) {
    case 43: MT_RGB_F(4, 3);   case 83: MT_RGB_F(8, 3);   case 163: MT_RGB_F(16, 3);
    case 41: MT_RGB_F(4, 1);   case 81: MT_RGB_F(8, 1);   case 161: MT_RGB_F(16, 1);
    case 44: MT_RGB_F(4, 4);
    case 84: MT_RGB_F(8, 4);
    case 164: MT_RGB_F(16, 4);
    default: break;
  }
#undef MT_RGB_F
  launch_pdl(rgb_head_fwd_kernel, dim3(grid_for((long long)N * HW, 256)), dim3(256), (size_t)C * 4 * sizeof(float), ST(stream), CBF(x), w, N, HW, C, Co, y);
  return cuda_rc();
}
extern "C" int mt_rgb_head_bwd_parts(int N, int HW, int C) {
  if (C % 8 || C / 8 > 256) return 0;
  const int lanes = 256 / (C / 8);
  return grid_for(((long long)N * HW + lanes - 1) / lanes, 1, 4);
}
extern "C" int mt_rgb_head_bwd(const float* dy, const float* y, const void* x, const float* w, int N,
                               int HW, int C, int Co, void* dx, float* dw, float* dw_parts, int parts,
                               void* stream) {
  if (C % 8 || Co > 4 || Co < 1 || C / 8 > 256) return -1003;
  const int V = C / 8, lanes = 256 / V;
  const size_t smem = ((size_t)C * 4 + (size_t)lanes * V * 32) * sizeof(float);
  const int grid = mt_rgb_head_bwd_parts(N, HW, C);
  if (parts != grid || !dw_parts) return -1012;
  const long long total = (long long)N * HW;
  const size_t wsmem = (size_t)256 * 32 * sizeof(float);
#define MT_RGB_B(V_, CO_) launch_pdl(rgb_head_bwd_warp_kernel<V_, CO_>, dim3(grid), dim3(256), wsmem, ST(stream), dy, y, CBF(x), w, total, HW, BF(dx), dw_parts); break
  switch (V * 10 + Co) {
    case 43: MT_RGB_B(4, 3);   case 83: MT_RGB_B(8, 3);   case 163: MT_RGB_B(16, 3);
    case 41: MT_RGB_B(4, 1);   case 81: MT_RGB_B(8, 1);   case 161: MT_RGB_B(16, 1);
    case 44: MT_RGB_B(4, 4);
    case 84: MT_RGB_B(8, 4);
    case 164: MT_RGB_B(16, 4);
    default:
      launch_pdl(rgb_head_bwd_kernel, dim3(grid), dim3(256), smem, ST(stream), dy, y, CBF(x), w, N, HW, C, Co, BF(dx), dw_parts);
  }
#undef MT_RGB_B
  launch_pdl(rgb_head_dw_reduce_kernel, dim3((C * 4 + 31) / 32), dim3(256), 0, ST(stream), dw_parts, parts, C, Co, dw);
  return cuda_rc();
}
extern "C" int mt_sumsq_interior_parts(int N, int H, int W, int C) {
  return C % 8 ? 0 : grid_for((long long)N * H * W * (C / 8), 256, 4);
}
extern "C" int mt_sumsq_interior(const void* x_pad, int N, int H, int W, int C, int pad, float* out,
                                 float* partials, int parts, void* stream) {
  if (C % 8) return -1003;
  if (parts != mt_sumsq_interior_parts(N, H, W, C) || !partials) return -1012;
  launch_pdl(sumsq_interior_kernel, dim3(parts), dim3(256), 0, ST(stream), CBF(x_pad), N, H, W, C, pad, partials);
  launch_pdl(sum_partials_kernel, dim3(1), dim3(256), 0, ST(stream), partials, parts, out);
  return cuda_rc();
}
extern "C" int mt_colsum_into(const float* stats, int rows, int C, int C_valid, float* db, int accumulate,
                              void* stream) {
  if (rows < 1 || C_valid < 1 || C_valid > C) return -1005;
  launch_pdl(colsum_finalize_kernel, dim3((C_valid + 31) / 32), dim3(256), 0, ST(stream), stats, rows, C, C_valid, db, accumulate);
  return cuda_rc();
}
extern "C" int mt_colsum_images(const float* stats, int N, int parts, int C, int C_valid, float* out,
                                void* stream) {
  if (N < 1 || parts < 1 || C_valid < 1 || C_valid > C) return -1005;
  launch_pdl(colsum_finalize_kernel, dim3((C_valid + 31) / 32, N), dim3(256), 0, ST(stream), stats, parts, C, C_valid,
             out, 0);
  return cuda_rc();
}
extern "C" int mt_scale_interior(const void* x_pad, const float* scale, int N, int H, int W, int C,
                                 int pad, void* dx_pad, void* stream) {
  if (C % 8) return -1003;
  launch_pdl(scale_interior_kernel, dim3(grid_for((long long)N * (H + 2 * pad) * (W + 2 * pad) * (C / 8), 256)), dim3(256), 0, ST(stream), CBF(x_pad), scale, N, H, W, C, pad, BF(dx_pad));
  return cuda_rc();
}
