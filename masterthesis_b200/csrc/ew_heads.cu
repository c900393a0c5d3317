// The small fp32 pieces between the convolution stacks and the loss scalars, as kernels of our own (no
// library GEMMs / elementwise launches inside the iteration):
//   * nn.Linear forward / backward for batches of <= a few hundred rows: the decoder's style MLP
//     (networks.py:256-261), the AdaIN modulation layers (norm.py:27), fc / fcVar of the style encoder
//     (networks.py:127-128), the 1x1 heads behind a global average pool (networks.py:77,402-403);
//   * the style MLP as one call: 3 layers + the fc of every AdaIN block (mt_style_mlp_fwd / _bwd);
//   * the discriminator heads (networks.py:373-376: conv1 1x1 with padding 1 -> real/fake map, conv2 k x k
//     -> domain logits) forward and backward;
//   * the loss layer with its gradients in the same pass: BCE-with-logits (loss.py:52-64,
//     adain_model.py:213-216), L1 (adain_model.py:303-306), the reparameterisation KL (adain_model.py:313-314).
// Everything is fixed-order (no atomics): results are reproducible bit for bit.
#include <map>
#include <mutex>
#include "../../include/mt_b200.h"
#include "ew_common.cuh"

namespace mt {

// ---------------------------------------------------------------- nn.Linear
// These layers are a few MFLOP on <= 64 rows: what matters is that no thread walks a long chain of
// dependent global loads.  Shapes: N rows, In <= ~1024, Out <= ~1024.
constexpr int LIN_ROWS = 8;     // rows of x (or g) a block keeps in shared memory / a thread in registers

// y[n][o] = act(sum_i x[n][i] * w[o][i] + b[o]).  Block = 8 warps = 8 output columns x LIN_ROWS rows; the
// rows of x are staged in shared memory once per block, a warp splits the input dimension over its lanes
// (coalesced reads of the weight row) and reduces by shuffles.  grid = (Out / 8, N / LIN_ROWS).
__global__ void __launch_bounds__(256)
linear_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                  int N, int In, int Out, int act, float* __restrict__ y) {
  pdl_wait(); pdl_launch();
  extern __shared__ float xs[];               // [LIN_ROWS][In]
  const int n0 = blockIdx.y * LIN_ROWS, rows = min(LIN_ROWS, N - n0);
  for (int t = threadIdx.x; t < rows * In; t += blockDim.x) xs[t] = __ldg(x + (size_t)n0 * In + t);
  __syncthreads();
  const int o = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (o >= Out) return;
  const float* wr = w + (size_t)o * In;
  float acc[LIN_ROWS];
#pragma unroll
  for (int j = 0; j < LIN_ROWS; ++j) acc[j] = 0.f;
  for (int i = lane; i < In; i += 32) {
    const float wv = __ldg(wr + i);
#pragma unroll
    for (int j = 0; j < LIN_ROWS; ++j)
      if (j < rows) acc[j] = fmaf(xs[j * In + i], wv, acc[j]);
  }
  const float bias = b ? __ldg(b + o) : 0.f;
#pragma unroll
  for (int j = 0; j < LIN_ROWS; ++j) {
    float s = acc[j];
    for (int k = 16; k; k >>= 1) s += __shfl_xor_sync(0xffffffffu, s, k);
    if (lane == 0 && j < rows) {
      s += bias;
      if (act == MT_ACT_RELU) s = fmaxf(s, 0.f);
      y[(size_t)(n0 + j) * Out + o] = s;
    }
  }
}

// Backward, g[n][o] = dy[n][o] * act'(y[n][o]).  Two roles in one launch, both with fixed-order sums:
//   blocks [0, gw): dW[o][i] (+)= sum_n g[n][o] x[n][i], db[o] (+)= sum_n g[n][o]; thread = (o, i), i fastest:
//     the N-long loop reads g[n][o] as a warp broadcast and x[n][i] coalesced (independent loads).
//   blocks [gw, ..): dx[n][i] (+)= sum_o g[n][o] w[o][i] for LIN_ROWS rows per block: the masked gradient rows
//     are staged in shared memory, block = 32 columns i x 8 slices of the o loop (each thread walks Out / 8
//     rows of w, coalesced, four loads in flight), the eight slices are combined in index order.
__global__ void __launch_bounds__(256)
linear_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ y, const float* __restrict__ x,
                  const float* __restrict__ w, int N, int In, int Out, int act, float* __restrict__ dx,
                  float* __restrict__ dw, float* __restrict__ db, int acc_param, int acc_dx, int gw, int gxi) {
  pdl_wait(); pdl_launch();
  extern __shared__ float sm[];
  if ((int)blockIdx.x < gw) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= Out * In) return;
    const int o = t / In, i = t - o * In;
    float s0 = 0.f, s1 = 0.f, sb = 0.f;
    int n = 0;
    for (; n + 1 < N; n += 2) {
      float g0 = __ldg(dy + (size_t)n * Out + o), g1 = __ldg(dy + (size_t)(n + 1) * Out + o);
      if (act == MT_ACT_RELU) {
        if (__ldg(y + (size_t)n * Out + o) <= 0.f) g0 = 0.f;
        if (__ldg(y + (size_t)(n + 1) * Out + o) <= 0.f) g1 = 0.f;
      }
      s0 = fmaf(g0, __ldg(x + (size_t)n * In + i), s0);
      s1 = fmaf(g1, __ldg(x + (size_t)(n + 1) * In + i), s1);
      sb += g0; sb += g1;
    }
    if (n < N) {
      float g0 = __ldg(dy + (size_t)n * Out + o);
      if (act == MT_ACT_RELU && __ldg(y + (size_t)n * Out + o) <= 0.f) g0 = 0.f;
      s0 = fmaf(g0, __ldg(x + (size_t)n * In + i), s0);
      sb += g0;
    }
    const float s = s0 + s1;
    if (dw) dw[t] = acc_param ? dw[t] + s : s;
    if (db && i == 0) db[o] = acc_param ? db[o] + sb : sb;
    return;
  }
  // ---- dx role
  const int bid = blockIdx.x - gw;
  const int it = bid % gxi, nb = bid / gxi;             // column tile, row block
  const int n0 = nb * LIN_ROWS, rows = min(LIN_ROWS, N - n0);
  float* gs = sm;                                        // [LIN_ROWS][Out] masked gradient rows
  float* part = sm + LIN_ROWS * Out;                     // [8][LIN_ROWS][32]
  for (int t = threadIdx.x; t < rows * Out; t += blockDim.x) {
    float g = __ldg(dy + (size_t)n0 * Out + t);
    if (act == MT_ACT_RELU && __ldg(y + (size_t)n0 * Out + t) <= 0.f) g = 0.f;
    gs[t] = g;
  }
  __syncthreads();
  const int il = threadIdx.x & 31, sl = threadIdx.x >> 5;       // 32 columns x 8 slices of the o loop
  const int i = it * 32 + il;
  float acc[LIN_ROWS];
#pragma unroll
  for (int j = 0; j < LIN_ROWS; ++j) acc[j] = 0.f;
  if (i < In) {
#pragma unroll 4
    for (int o = sl; o < Out; o += 8) {
      const float wv = __ldg(w + (size_t)o * In + i);
#pragma unroll
      for (int j = 0; j < LIN_ROWS; ++j)
        if (j < rows) acc[j] = fmaf(gs[j * Out + o], wv, acc[j]);
    }
  }
#pragma unroll
  for (int j = 0; j < LIN_ROWS; ++j) part[(sl * LIN_ROWS + j) * 32 + il] = acc[j];
  __syncthreads();
  if (sl < rows && i < In) {                                    // warp j combines row j, slices in index order
    const int j = sl;
    float s = part[j * 32 + il];
#pragma unroll
    for (int k = 1; k < 8; ++k) s += part[(k * LIN_ROWS + j) * 32 + il];
    float* q = dx + (size_t)(n0 + j) * In + i;
    *q = acc_dx ? *q + s : s;
  }
}

static int launch_linear_fwd(const float* x, const float* w, const float* b, int N, int In, int Out, int act,
                             float* y, cudaStream_t st) {
  if (N < 1 || In < 1 || Out < 1 || (act != MT_ACT_NONE && act != MT_ACT_RELU)) return -1005;
  const size_t smem = (size_t)LIN_ROWS * In * sizeof(float);
  if (smem > 48 * 1024) return -1005;                    // In <= 1536
  launch_pdl(linear_fwd_kernel, dim3(dim3((Out + 7) / 8, (N + LIN_ROWS - 1) / LIN_ROWS)), dim3(256), smem, st, x, w, b, N, In, Out, act, y);
  return cuda_rc();
}
static int launch_linear_bwd(const float* dy, const float* y, const float* x, const float* w, int N, int In,
                             int Out, int act, float* dx, float* dw, float* db, int acc_param, int acc_dx,
                             cudaStream_t st) {
  if (N < 1 || In < 1 || Out < 1 || (act != MT_ACT_NONE && act != MT_ACT_RELU)) return -1005;
  if (act == MT_ACT_RELU && !y) return -1008;
  if ((long long)Out * In >= (1ll << 31) || (long long)N * In >= (1ll << 31)) return -1005;
  const int gw = (dw || db) ? (Out * In + 255) / 256 : 0;
  const int gxi = (In + 31) / 32;
  const int gx = dx ? gxi * ((N + LIN_ROWS - 1) / LIN_ROWS) : 0;
  if (gw + gx == 0) return 0;
  const size_t smem = dx ? ((size_t)LIN_ROWS * Out + 8 * LIN_ROWS * 32) * sizeof(float) : 0;
  if (smem > 48 * 1024) return -1005;                    // Out <= 1280
  launch_pdl(linear_bwd_kernel, dim3(gw + gx), dim3(256), smem, st, dy, y, x, w, N, In, Out, act, dx, dw, db, acc_param, acc_dx, gw, gxi);
  return cuda_rc();
}

// ---------------------------------------------------------------- discriminator heads
// h [N][S][S][Cs] bf16 (C valid channels) -> out [N][S+2][S+2] fp32 (1x1 conv with padding 1: exact-zero
// border) and cls [N][nd] (k x k conv with k == S, i.e. a full contraction, + global average of one value).
// One block per image; thread = channel slice: h is read coalesced over the channels, the thread's
// w2[d][c][0..SS) rows are contiguous; block-level tree reduction in a fixed order.
__global__ void __launch_bounds__(256)
dis_head_fwd_kernel(const __nv_bfloat16* __restrict__ h, const float* __restrict__ w1,
                    const float* __restrict__ w2, int SS, int Cs, int C, int nd, int S,
                    float* __restrict__ out, float* __restrict__ cls) {
  pdl_wait(); pdl_launch();
  extern __shared__ float red[];            // [256][SS + nd]
  const int n = blockIdx.x, tid = threadIdx.x, NV = SS + nd;
  float* mine = red + tid * NV;
  for (int k = 0; k < NV; ++k) mine[k] = 0.f;
  for (int c = tid; c < C; c += blockDim.x) {
    const float a1 = __ldg(w1 + c);
    for (int p0 = 0; p0 < SS; p0 += 16) {             // 16 positions at a time: independent loads first
      float v[16];
#pragma unroll
      for (int k = 0; k < 16; ++k)
        v[k] = p0 + k < SS ? __bfloat162float(h[((size_t)n * SS + p0 + k) * Cs + c]) : 0.f;
#pragma unroll
      for (int k = 0; k < 16; ++k)
        if (p0 + k < SS) mine[p0 + k] = fmaf(v[k], a1, mine[p0 + k]);
      for (int d = 0; d < nd; ++d) {
        const float* wr = w2 + ((size_t)d * C + c) * SS + p0;     // this thread's contiguous filter row
        float acc = 0.f;
#pragma unroll
        for (int k = 0; k < 16; ++k)
          if (p0 + k < SS) acc = fmaf(v[k], __ldg(wr + k), acc);
        mine[SS + d] += acc;
      }
    }
  }
  __syncthreads();
  for (int s = blockDim.x / 2; s > 0; s >>= 1) {
    if (tid < s)
      for (int k = 0; k < NV; ++k) red[tid * NV + k] += red[(tid + s) * NV + k];
    __syncthreads();
  }
  const int So = S + 2;
  for (int k = tid; k < So * So; k += blockDim.x) {
    const int r = k / So, q = k - r * So;
    const bool in = r >= 1 && r <= S && q >= 1 && q <= S;
    out[(size_t)n * So * So + k] = in ? red[(r - 1) * S + (q - 1)] : 0.f;
  }
  for (int d = tid; d < nd; d += blockDim.x) cls[(size_t)n * nd + d] = red[SS + d];
}

// Backward of the heads.  dout [N][S+2][S+2] (border ignored), dcls [N][nd] (images n < cls_first carry no
// class gradient: the fake half of a discriminator batch, adain_model.py:213-216).  Three roles:
//   blocks [0, ga):      dh[n][p][c] = dout[n][p] w1[c] + sum_d dcls[n][d] w2[d][c][p]     thread = (n, p, c)
//   blocks [ga, ga+gb):  dw2[d][c][p] (+)= sum_n dcls[n][d] h[n][p][c]                     thread = (p, c), c fastest
//   blocks [ga+gb, ..):  dw1[c] (+)= sum_{n,p} dout[n][p] h[n][p][c]       block = 32 channels x 8 groups of
//                        (n, p) pairs, combined in index order
__global__ void __launch_bounds__(256)
dis_head_bwd_kernel(const float* __restrict__ dout, const float* __restrict__ dcls,
                    const __nv_bfloat16* __restrict__ h, const float* __restrict__ w1,
                    const float* __restrict__ w2, int N, int SS, int Cs, int C, int nd, int S,
                    __nv_bfloat16* __restrict__ dh, float* __restrict__ dw1, float* __restrict__ dw2,
                    int accumulate, int ga, int gb, int cls_first) {
  pdl_wait(); pdl_launch();
  const int So = S + 2;
  if ((int)blockIdx.x < ga) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (long long)N * SS * Cs) return;
    const int c = (int)(t % Cs);
    const long long r = t / Cs;
    const int p = (int)(r % SS), n = (int)(r / SS);
    float v = 0.f;
    if (c < C) {
      v = __ldg(dout + (size_t)n * So * So + (p / S + 1) * So + (p % S + 1)) * __ldg(w1 + c);
      if (n >= cls_first)
        for (int d = 0; d < nd; ++d)
          v = fmaf(__ldg(dcls + (size_t)n * nd + d), __ldg(w2 + ((size_t)d * C + c) * SS + p), v);
    }
    dh[t] = __float2bfloat16(v);
  } else if ((int)blockIdx.x < ga + gb) {
    const int t = (blockIdx.x - ga) * blockDim.x + threadIdx.x;
    if (t >= SS * C) return;
    const int p = t / C, c = t - p * C;
    float a2[8];
#pragma unroll
    for (int d = 0; d < 8; ++d) a2[d] = 0.f;
    for (int n = cls_first; n < N; ++n) {
      const float v = __bfloat162float(h[((size_t)n * SS + p) * Cs + c]);
#pragma unroll
      for (int d = 0; d < 8; ++d)
        if (d < nd) a2[d] = fmaf(__ldg(dcls + (size_t)n * nd + d), v, a2[d]);
    }
#pragma unroll
    for (int d = 0; d < 8; ++d)
      if (d < nd) {
        float* q = dw2 + ((size_t)d * C + c) * SS + p;
        *q = accumulate ? *q + a2[d] : a2[d];
      }
  } else {
    __shared__ float sm1[8][33];
    const int cl = threadIdx.x & 31, g = threadIdx.x >> 5;
    const int c = (blockIdx.x - ga - gb) * 32 + cl;
    float a = 0.f;
    if (c < C)
      for (int k = g; k < N * SS; k += 8) {
        const int n = k / SS, p = k - n * SS;
        a = fmaf(__ldg(dout + (size_t)n * So * So + (p / S + 1) * So + (p % S + 1)),
                 __bfloat162float(h[(size_t)k * Cs + c]), a);
      }
    sm1[g][cl] = a;
    __syncthreads();
    if (g == 0 && c < C) {
      float v = sm1[0][cl];
#pragma unroll
      for (int k = 1; k < 8; ++k) v += sm1[k][cl];
      dw1[c] = accumulate ? dw1[c] + v : v;
    }
  }
}

// ---------------------------------------------------------------- losses (value + gradient in one pass)
// Single-block kernels for the small tensors (logits, style codes); fixed-order tree reduction.
__device__ __forceinline__ float block_sum_256(float v, float* sm) {
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = 0.f;
  if (threadIdx.x == 0)
    for (int i = 0; i < (int)(blockDim.x >> 5); ++i) t += sm[i];
  __syncthreads();
  return t;    // valid in thread 0
}

// nn.BCEWithLogitsLoss (mean) of x[n] against target[i] (tensor) or the constant tconst:
//   loss = mean( max(x,0) - x t + log1p(exp(-|x|)) ),  dx = gscale / n * (sigmoid(x) - t)
// loss_out[0] = the mean (unweighted, what the reference logs); gscale = the weight of this term in the
// total loss that is back-propagated.
__global__ void __launch_bounds__(256)
bce_logits_kernel(const float* __restrict__ x, const float* __restrict__ target, float tconst, int n,
                  float gscale, float* __restrict__ loss_out, float* __restrict__ dx) {
  pdl_wait(); pdl_launch();
  __shared__ float sm[8];
  float s = 0.f;
  const float inv = 1.f / (float)n;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const float v = x[i], t = target ? target[i] : tconst;
    s += fmaxf(v, 0.f) - v * t + log1pf(expf(-fabsf(v)));
    if (dx) dx[i] = gscale * inv * (1.f / (1.f + expf(-v)) - t);
  }
  const float tot = block_sum_256(s, sm);
  if (threadIdx.x == 0) loss_out[0] = tot * inv;
}

// nn.L1Loss (mean) of a against b, n elements: per-block partial sums + sign gradient, then the finalize
// below adds the partials in block order.  da = gscale / n * sign(a - b).
__global__ void __launch_bounds__(256)
l1_partial_kernel(const float* __restrict__ a, const float* __restrict__ b, long long n, float gscale,
                  float* __restrict__ part, float* __restrict__ da) {
  pdl_wait(); pdl_launch();
  __shared__ float sm[8];
  float s = 0.f;
  const float k = gscale / (float)n;
  const long long n4 = n / 4;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    const float4 u = reinterpret_cast<const float4*>(a)[i], v = reinterpret_cast<const float4*>(b)[i];
    const float d0 = u.x - v.x, d1 = u.y - v.y, d2 = u.z - v.z, d3 = u.w - v.w;
    s += (fabsf(d0) + fabsf(d1)) + (fabsf(d2) + fabsf(d3));
    if (da)
      reinterpret_cast<float4*>(da)[i] = make_float4(d0 > 0.f ? k : (d0 < 0.f ? -k : 0.f), d1 > 0.f ? k : (d1 < 0.f ? -k : 0.f),
                                                     d2 > 0.f ? k : (d2 < 0.f ? -k : 0.f), d3 > 0.f ? k : (d3 < 0.f ? -k : 0.f));
  }
  if (blockIdx.x == 0)
    for (long long i = n4 * 4 + threadIdx.x; i < n; i += blockDim.x) {
      const float d = a[i] - b[i];
      s += fabsf(d);
      if (da) da[i] = d > 0.f ? k : (d < 0.f ? -k : 0.f);
    }
  const float tot = block_sum_256(s, sm);
  if (threadIdx.x == 0) part[blockIdx.x] = tot;
}
__global__ void __launch_bounds__(256)
scaled_sum_kernel(const float* __restrict__ part, int parts, float scale, float* __restrict__ out) {
  pdl_wait(); pdl_launch();
  __shared__ float sm[8];
  float s = 0.f;
  for (int i = threadIdx.x; i < parts; i += blockDim.x) s += part[i];
  const float tot = block_sum_256(s, sm);
  if (threadIdx.x == 0) out[0] = tot * scale;
}

// KL(N(mu, exp(logvar)) || N(0,1)) summed over all n elements (adain_model.py:313-314):
//   loss = -0.5 * sum(1 + logvar - mu^2 - exp(logvar));  dmu = gscale * mu;  dlogvar = gscale * 0.5 (exp(logvar) - 1)
__global__ void __launch_bounds__(256)
kl_normal_kernel(const float* __restrict__ mu, const float* __restrict__ logvar, int n, float gscale,
                 float* __restrict__ loss_out, float* __restrict__ dmu, float* __restrict__ dlogvar) {
  pdl_wait(); pdl_launch();
  __shared__ float sm[8];
  float s = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const float m = mu[i], lv = logvar[i], e = expf(lv);
    s += 1.f + lv - m * m - e;
    if (dmu) { dmu[i] = gscale * m; dlogvar[i] = gscale * 0.5f * (e - 1.f); }
  }
  const float tot = block_sum_256(s, sm);
  if (threadIdx.x == 0) loss_out[0] = -0.5f * tot;
}

}  // namespace mt

using namespace mt;
#define ST(s) ((cudaStream_t)(s))

extern "C" int mt_linear_fwd(const float* x, const float* w, const float* b, int N, int In, int Out, int act,
                             float* y, void* stream) {
  return launch_linear_fwd(x, w, b, N, In, Out, act, y, ST(stream));
}
extern "C" int mt_linear_bwd(const float* dy, const float* y, const float* x, const float* w, int N, int In,
                             int Out, int act, float* dx, float* dw, float* db, int accumulate_params,
                             int accumulate_dx, void* stream) {
  return launch_linear_bwd(dy, y, x, w, N, In, Out, act, dx, dw, db, accumulate_params, accumulate_dx, ST(stream));
}

// The decoder's style network in one call: h1 = relu(L0(cz)), h2 = relu(L1(h1)), s = L2(h2) and, for every
// AdaIN block k < n_heads, gb[k] = fc_k(s)  ([N][head_out], head_out = 2 * channels: weight | bias).
extern "C" int mt_style_mlp_fwd(const float* cz, int N, int in_dim, int hid, int style_dim,
                                const float* const* mlp_w, const float* const* mlp_b, int n_heads, int head_out,
                                const float* const* head_w, const float* const* head_b, float* h1, float* h2,
                                float* s, float* const* gb, void* stream) {
  if (!mlp_w || !mlp_b || (n_heads > 0 && (!head_w || !head_b || !gb))) return -1008;
  int rc = launch_linear_fwd(cz, mlp_w[0], mlp_b[0], N, in_dim, hid, MT_ACT_RELU, h1, ST(stream));
  if (!rc) rc = launch_linear_fwd(h1, mlp_w[1], mlp_b[1], N, hid, hid, MT_ACT_RELU, h2, ST(stream));
  if (!rc) rc = launch_linear_fwd(h2, mlp_w[2], mlp_b[2], N, hid, style_dim, MT_ACT_NONE, s, ST(stream));
  for (int k = 0; k < n_heads && !rc; ++k)
    rc = launch_linear_fwd(s, head_w[k], head_b[k], N, style_dim, head_out, MT_ACT_NONE, gb[k], ST(stream));
  return rc;
}
// Backward: dgb [n_heads][N][head_out] (and / or ds_in [N][style_dim], the gradient reaching s directly) ->
// gradients of every layer, written (accumulate = 0) or added (accumulate = 1: parameter .grad buffers) to
// d_mlp_w / d_mlp_b / d_head_w / d_head_b, and dcz [N][in_dim] (optional).  ds, dh2, dh1: scratch.
extern "C" int mt_style_mlp_bwd(const float* cz, const float* h1, const float* h2, const float* s, int N,
                                int in_dim, int hid, int style_dim, const float* const* mlp_w, int n_heads,
                                int head_out, const float* const* head_w, const float* const* dgb, const float* ds_in,
                                float* const* d_mlp_w, float* const* d_mlp_b, float* const* d_head_w,
                                float* const* d_head_b, float* ds, float* dh2, float* dh1, float* dcz,
                                int accumulate, void* stream) {
  if (!mlp_w || !d_mlp_w || !d_mlp_b || !ds || !dh2 || !dh1) return -1008;
  if (n_heads > 0 && (!head_w || !d_head_w || !d_head_b || !dgb)) return -1008;
  if (n_heads == 0 && !ds_in) return -1008;
  int rc = 0;
  if (ds_in) {
    cudaError_t e = cudaMemcpyAsync(ds, ds_in, (size_t)N * style_dim * sizeof(float), cudaMemcpyDeviceToDevice, ST(stream));
    if (e != cudaSuccess) return -(int)e;
  }
  bool have_ds = ds_in != nullptr;
  for (int k = 0; k < n_heads && !rc; ++k) {
    if (!dgb[k]) continue;                       // a block whose modulation received no gradient
    rc = launch_linear_bwd(dgb[k], nullptr, s, head_w[k], N, style_dim, head_out, MT_ACT_NONE,
                           ds, d_head_w[k], d_head_b[k], accumulate, have_ds ? 1 : 0, ST(stream));
    have_ds = true;
  }
  if (!have_ds) return -1008;
  if (!rc) rc = launch_linear_bwd(ds, nullptr, h2, mlp_w[2], N, hid, style_dim, MT_ACT_NONE, dh2, d_mlp_w[2], d_mlp_b[2], accumulate, 0, ST(stream));
  if (!rc) rc = launch_linear_bwd(dh2, h2, h1, mlp_w[1], N, hid, hid, MT_ACT_RELU, dh1, d_mlp_w[1], d_mlp_b[1], accumulate, 0, ST(stream));
  if (!rc) rc = launch_linear_bwd(dh1, h1, cz, mlp_w[0], N, in_dim, hid, MT_ACT_RELU, dcz, d_mlp_w[0], d_mlp_b[0], accumulate, 0, ST(stream));
  return rc;
}

extern "C" int mt_dis_head_fwd(const void* h, const float* w1, const float* w2, int N, int S, int Cs, int C, int nd,
                               float* out, float* cls, void* stream) {
  if (N < 1 || S < 1 || C < 1 || C > Cs || nd < 1 || nd > 8) return -1005;
  const size_t smem = (size_t)256 * (S * S + nd) * sizeof(float);
  if (smem > 160 * 1024) return -1005;          // S <= 12: crop sizes up to 768 (k = crop / 64, networks.py:374)
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(dis_head_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
    if (e != cudaSuccess) return -(int)e;
    attr_set = true;
  }
  launch_pdl(dis_head_fwd_kernel, dim3(N), dim3(256), smem, ST(stream), (const __nv_bfloat16*)h, w1, w2, S * S, Cs, C, nd, S, out, cls);
  return cuda_rc();
}
extern "C" int mt_dis_head_bwd(const float* dout, const float* dcls, const void* h, const float* w1,
                               const float* w2, int N, int S, int Cs, int C, int nd, int cls_first, void* dh,
                               float* dw1, float* dw2, int accumulate, void* stream) {
  if (N < 1 || S < 1 || C < 1 || C > Cs || nd < 1 || nd > 8) return -1005;
  const long long tot = (long long)N * S * S * Cs;
  if (tot >= (1ll << 31)) return -1005;
  const int ga = dh ? (int)((tot + 255) / 256) : 0;
  const bool params = dw1 && dw2;
  const int gb = params ? (S * S * C + 255) / 256 : 0;
  const int gc = params ? (C + 31) / 32 : 0;
  if (ga + gb + gc == 0) return 0;
  launch_pdl(dis_head_bwd_kernel, dim3(ga + gb + gc), dim3(256), 0, ST(stream), dout, dcls, (const __nv_bfloat16*)h, w1, w2, N, S * S, Cs,
                                                           C, nd, S, (__nv_bfloat16*)dh, dw1, dw2, accumulate, ga, gb,
                                                           cls_first);
  return cuda_rc();
}

extern "C" int mt_loss_bce_logits(const float* x, const float* target, float target_const, int n, float gscale,
                                  float* loss, float* dx, void* stream) {
  if (n < 1 || !loss) return -1005;
  launch_pdl(bce_logits_kernel, dim3(1), dim3(256), 0, ST(stream), x, target, target_const, n, gscale, loss, dx);
  return cuda_rc();
}
extern "C" int mt_loss_l1_parts(long long n) { return n < 1 ? 0 : grid_for((n + 3) / 4, 256, 2); }
extern "C" int mt_loss_l1(const float* a, const float* b, long long n, float weight, float gscale, float* loss,
                          float* da, float* partials, int parts, void* stream) {
  if (n < 1 || !loss) return -1005;
  if (parts != mt_loss_l1_parts(n) || !partials) return -1012;
  if ((((uintptr_t)a | (uintptr_t)b | (uintptr_t)da) & 15) != 0) return -1005;
  launch_pdl(l1_partial_kernel, dim3(parts), dim3(256), 0, ST(stream), a, b, n, gscale, partials, da);
  launch_pdl(scaled_sum_kernel, dim3(1), dim3(256), 0, ST(stream), partials, parts, weight / (float)n, loss);
  return cuda_rc();
}
// ---- mt_loss_l1 without caller scratch: the per-block partial sums live in a file-scope device array.
// No counterpart in the reference; added for callers that cannot
// allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  Every GPU that loads the
// library holds the MT_LOSS_L1_STATIC_MAX_PARTS * 4 = 2 KiB array until exit.  In host code the array's name is
// only a stand-in: the device address comes from cudaGetSymbolAddress, once per device, with that device
// current.  Every slot that the sum kernel reads was written by the partial kernel of the same call, so
// nothing has to start at zero; all calls on one device share the array, so two calls that overlap on the GPU
// (different streams without an ordering between them) race.
#define MT_LOSS_L1_STATIC_MAX_PARTS 512
#define MT_LOSS_L1_STATIC_MAX_N (1ll << 30)
static __device__ __align__(256) float g_l1_static_parts[MT_LOSS_L1_STATIC_MAX_PARTS];

static int l1_static_scratch(float** out) {
  static std::mutex mu;
  static std::map<int, float*> by_device;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return -(int)e;
  std::lock_guard<std::mutex> lock(mu);
  auto it = by_device.find(dev);
  if (it == by_device.end()) {
    float* ptr = nullptr;
    e = cudaGetSymbolAddress((void**)&ptr, g_l1_static_parts);     // `dev` is the current device
    if (e != cudaSuccess) return -(int)e;
    if (!ptr) return -1016;
    it = by_device.emplace(dev, ptr).first;
  }
  *out = it->second;
  return 0;
}

extern "C" long long mt_loss_l1_static_max_n(void) { return MT_LOSS_L1_STATIC_MAX_N; }
extern "C" int mt_loss_l1_static(const float* a, const float* b, long long n, float weight, float gscale,
                                 float* loss, float* da, void* stream) {
  if (n < 1 || !loss || !a || !b) return -1005;
  if (n > MT_LOSS_L1_STATIC_MAX_N) return -1016;
  const int parts = mt_loss_l1_parts(n);
  if (parts < 1 || parts > MT_LOSS_L1_STATIC_MAX_PARTS) return -1016;
  if ((((uintptr_t)a | (uintptr_t)b | (uintptr_t)da) & 15) != 0) return -1005;
  float* partials = nullptr;
  const int rc = l1_static_scratch(&partials);
  if (rc) return rc;
  launch_pdl(l1_partial_kernel, dim3(parts), dim3(256), 0, ST(stream), a, b, n, gscale, partials, da);
  launch_pdl(scaled_sum_kernel, dim3(1), dim3(256), 0, ST(stream), (const float*)partials, parts,
             weight / (float)n, loss);
  return cuda_rc();
}
extern "C" int mt_loss_kl_normal(const float* mu, const float* logvar, int n, float gscale, float* loss,
                                 float* dmu, float* dlogvar, void* stream) {
  if (n < 1 || !loss || ((dmu == nullptr) != (dlogvar == nullptr))) return -1005;
  launch_pdl(kl_normal_kernel, dim3(1), dim3(256), 0, ST(stream), mu, logvar, n, gscale, loss, dmu, dlogvar);
  return cuda_rc();
}
