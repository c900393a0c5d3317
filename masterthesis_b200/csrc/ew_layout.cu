// Layout / padding kernels at the edges of the path: NCHW fp32 images <-> NHWC bf16 buffers,
// reflection borders (nn.ReflectionPad2d, reference blocks.py:29-31) and their adjoint.
#include "../../include/mt_b200.h"
#include "ew_common.cuh"

namespace mt {

// img [N,Ci,H,W] fp32 (+ one-hot c [N,nd] tiled over H,W as in networks.py:136-140)
//   -> out [N,H+2p,W+2p,Cs] bf16, reflection border included, channels >= Ci+nd zero.
__global__ void image_to_nhwc_kernel(const float* __restrict__ img, const float* __restrict__ onehot,
                                     int N, int Ci, int nd, int H, int W, int p, int Cs,
                                     __nv_bfloat16* __restrict__ out) {
  pdl_wait(); pdl_launch();
  const int Hp = H + 2 * p, Wp = W + 2 * p, V = Cs / 8;
  const long long total = (long long)N * Hp * Wp * V;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int v = (int)(i % V); unsigned t = i / V;
    const int wp = (int)(t % Wp); t /= Wp;
    const int hp = (int)(t % Hp); const int n = (int)(t / Hp);
    const int h = reflect_src(hp, H, p), w = reflect_src(wp, W, p);
    float f[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = v * 8 + j;
      float x = 0.f;
      if (c < Ci) x = __ldg(img + (((long long)n * Ci + c) * H + h) * W + w);
      else if (c < Ci + nd) x = __ldg(onehot + (long long)n * nd + (c - Ci));
      f[j] = x;
    }
    store8(out + (size_t)i * 8, f);
  }
}

// Adjoint: dpad [N,H+2p,W+2p,Cs] bf16 -> dimg [N,Ci,H,W] fp32 (border folded onto interior).
__global__ void image_grad_kernel(const __nv_bfloat16* __restrict__ dpad, int N, int Ci, int H,
                                  int W, int p, int Cs, float* __restrict__ dimg) {
  pdl_wait(); pdl_launch();
  const int Wp = W + 2 * p, Hp = H + 2 * p;
  const long long total = (long long)N * H * W;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int w = (int)(i % W); unsigned t = i / W;
    const int h = (int)(t % H); const int n = (int)(t / H);
    int hs[3], ws[3];
    const int nh = reflect_fold_sources(h, H, p, hs), nw = reflect_fold_sources(w, W, p, ws);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int a = 0; a < nh; ++a)
      for (int b = 0; b < nw; ++b) {
        float f[8];
        load8(dpad + (((long long)n * Hp + hs[a]) * Wp + ws[b]) * Cs, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
      }
    for (int c = 0; c < Ci && c < 8; ++c) dimg[(((long long)n * Ci + c) * H + h) * W + w] = acc[c];
  }
}

// Fill the reflection border of buf [N,H+2p,W+2p,C] from its interior, in place.
__global__ void reflect_border_kernel(__nv_bfloat16* __restrict__ buf, int N, int H, int W, int C,
                                      int p) {
  pdl_wait(); pdl_launch();
  const int Hp = H + 2 * p, Wp = W + 2 * p, V = C / 8;
  const int border_px = Hp * Wp - H * W;
  const long long total = (long long)N * border_px * V;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int v = (int)(i % V); unsigned t = i / V;
    int b = (int)(t % border_px); const int n = (int)(t / border_px);
    int hp, wp;
    if (b < 2 * p * Wp) {            // full top and bottom rows
      hp = b / Wp; wp = b % Wp;
      if (hp >= p) hp += H;
    } else {                         // left/right columns of the interior rows
      b -= 2 * p * Wp;
      hp = p + b / (2 * p);
      wp = b % (2 * p);
      if (wp >= p) wp += W;
    }
    const int h = reflect_src(hp, H, p), w = reflect_src(wp, W, p);
    const uint4 val = *reinterpret_cast<const uint4*>(
        buf + (((long long)n * Hp + h + p) * Wp + w + p) * C + v * 8);
    *reinterpret_cast<uint4*>(buf + (((long long)n * Hp + hp) * Wp + wp) * C + v * 8) = val;
  }
}

// dense[n,h,w,:] = sum of dpad over the padded positions that mirrored (h,w)
__global__ void reflect_fold_kernel(const __nv_bfloat16* __restrict__ dpad, int N, int H, int W,
                                    int C, int p, __nv_bfloat16* __restrict__ dense) {
  pdl_wait(); pdl_launch();
  const int Hp = H + 2 * p, Wp = W + 2 * p, V = C / 8;
  const long long total = (long long)N * H * W * V;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int v = (int)(i % V); unsigned t = i / V;
    const int w = (int)(t % W); t /= W;
    const int h = (int)(t % H); const int n = (int)(t / H);
    int hs[3], ws[3];
    const int nh = reflect_fold_sources(h, H, p, hs), nw = reflect_fold_sources(w, W, p, ws);
    float acc[8];
    load8(dpad + (((long long)n * Hp + hs[0]) * Wp + ws[0]) * C + v * 8, acc);
    for (int a = 0; a < nh; ++a)
      for (int b = (a == 0 ? 1 : 0); b < nw; ++b) {
        float f[8];
        load8(dpad + (((long long)n * Hp + hs[a]) * Wp + ws[b]) * C + v * 8, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
      }
    store8(dense + (size_t)i * 8, acc);
  }
}

// dpad[n, h+p, w+p, :] += dense[n,h,w,:]
__global__ void add_interior_kernel(__nv_bfloat16* __restrict__ dpad, const __nv_bfloat16* dense,
                                    int N, int H, int W, int C, int p) {
  pdl_wait(); pdl_launch();
  const int Hp = H + 2 * p, Wp = W + 2 * p, V = C / 8;
  const long long total = (long long)N * H * W * V;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int v = (int)(i % V); unsigned t = i / V;
    const int w = (int)(t % W); t /= W;
    const int h = (int)(t % H); const int n = (int)(t / H);
    __nv_bfloat16* d = dpad + (((long long)n * Hp + h + p) * Wp + w + p) * C + v * 8;
    float a[8], b[8];
    load8(d, a);
    load8(dense + (size_t)i * 8, b);
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] += b[j];
    store8(d, a);
  }
}

}  // namespace mt

using namespace mt;
#define ST(s) ((cudaStream_t)(s))

extern "C" int mt_image_to_nhwc(const float* img, const float* onehot, int N, int Ci, int nd, int H,
                                int W, int pad, int Cs, void* out, void* stream) {
  if (Cs % 8 || Ci + nd > Cs) return -1003;
  const long long total = (long long)N * (H + 2 * pad) * (W + 2 * pad) * (Cs / 8);
  launch_pdl(image_to_nhwc_kernel, dim3(grid_for(total, 256)), dim3(256), 0, ST(stream), 
      img, onehot, N, Ci, onehot ? nd : 0, H, W, pad, Cs, (__nv_bfloat16*)out);
  return cuda_rc();
}
extern "C" int mt_image_grad_from_nhwc(const void* dpad, int N, int Ci, int H, int W, int pad,
                                       int Cs, float* dimg, void* stream) {
  if (Cs != 8 || Ci > 8) return -1003;
  launch_pdl(image_grad_kernel, dim3(grid_for((long long)N * H * W, 256)), dim3(256), 0, ST(stream), 
      (const __nv_bfloat16*)dpad, N, Ci, H, W, pad, Cs, dimg);
  return cuda_rc();
}
extern "C" int mt_reflect_border(void* buf, int N, int H, int W, int C, int pad, void* stream) {
  if (C % 8) return -1003;
  if (pad == 0) return 0;
  if (pad >= H || pad >= W) return -1007;
  const long long total =
      (long long)N * ((H + 2 * pad) * (W + 2 * pad) - H * W) * (C / 8);
  launch_pdl(reflect_border_kernel, dim3(grid_for(total, 256)), dim3(256), 0, ST(stream), (__nv_bfloat16*)buf, N, H,
                                                                       W, C, pad);
  return cuda_rc();
}
extern "C" int mt_reflect_fold(const void* dpad, int N, int H, int W, int C, int pad, void* dense,
                               void* stream) {
  if (C % 8) return -1003;
  const long long total = (long long)N * H * W * (C / 8);
  launch_pdl(reflect_fold_kernel, dim3(grid_for(total, 256)), dim3(256), 0, ST(stream), 
      (const __nv_bfloat16*)dpad, N, H, W, C, pad, (__nv_bfloat16*)dense);
  return cuda_rc();
}
extern "C" int mt_add_interior(void* dpad, const void* dense, int N, int H, int W, int C, int pad,
                               void* stream) {
  if (C % 8) return -1003;
  const long long total = (long long)N * H * W * (C / 8);
  launch_pdl(add_interior_kernel, dim3(grid_for(total, 256)), dim3(256), 0, ST(stream), 
      (__nv_bfloat16*)dpad, (const __nv_bfloat16*)dense, N, H, W, C, pad);
  return cuda_rc();
}
