// Fused Adam over one flat fp32 parameter bucket (torch.optim.Adam semantics with coupled L2
// weight decay, reference adain_model.py:58-61): one launch per optimizer instead of one per
// tensor.  `gscale` folds the data-parallel gradient averaging (1/world) into the same pass.
#include "../../include/mt_b200.h"
#include "ew_common.cuh"

namespace mt {

__global__ void __launch_bounds__(256)
adam_flat_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                 float* __restrict__ v, long long n, float lr_over_bc1, float inv_sqrt_bc2, float b1,
                 float b2, float eps, float wd, float gscale) {
  pdl_wait(); pdl_launch();
  const long long n4 = n / 4;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 pp = reinterpret_cast<float4*>(p)[i];
    const float4 gg = reinterpret_cast<const float4*>(g)[i];
    float4 mm = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    float* P = reinterpret_cast<float*>(&pp);
    const float* G = reinterpret_cast<const float*>(&gg);
    float* M = reinterpret_cast<float*>(&mm);
    float* V = reinterpret_cast<float*>(&vv);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float gr = G[j] * gscale + wd * P[j];
      M[j] = b1 * M[j] + (1.f - b1) * gr;
      V[j] = b2 * V[j] + (1.f - b2) * gr * gr;
      const float denom = sqrtf(V[j]) * inv_sqrt_bc2 + eps;
      P[j] -= lr_over_bc1 * (M[j] / denom);
    }
    reinterpret_cast<float4*>(p)[i] = pp;
    reinterpret_cast<float4*>(m)[i] = mm;
    reinterpret_cast<float4*>(v)[i] = vv;
  }
  // tail
  for (long long i = n4 * 4 + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float gr = g[i] * gscale + wd * p[i];
    const float mi = b1 * m[i] + (1.f - b1) * gr;
    const float vi = b2 * v[i] + (1.f - b2) * gr * gr;
    m[i] = mi; v[i] = vi;
    p[i] -= lr_over_bc1 * (mi / (sqrtf(vi) * inv_sqrt_bc2 + eps));
  }
}

__global__ void __launch_bounds__(256)
adam_flat_dev_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                     float* __restrict__ v, long long n, const float* __restrict__ hyper) {
  pdl_wait(); pdl_launch();
  const float lr_over_bc1 = __ldg(hyper + 0), inv_sqrt_bc2 = __ldg(hyper + 1), b1 = __ldg(hyper + 2),
              b2 = __ldg(hyper + 3), eps = __ldg(hyper + 4), wd = __ldg(hyper + 5),
              gscale = __ldg(hyper + 6);
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long n4 = n / 4;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 pp = reinterpret_cast<float4*>(p)[i];
    const float4 gg = reinterpret_cast<const float4*>(g)[i];
    float4 mm = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    float* P = reinterpret_cast<float*>(&pp);
    const float* G = reinterpret_cast<const float*>(&gg);
    float* M = reinterpret_cast<float*>(&mm);
    float* V = reinterpret_cast<float*>(&vv);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float gr = G[j] * gscale + wd * P[j];
      M[j] = b1 * M[j] + (1.f - b1) * gr;
      V[j] = b2 * V[j] + (1.f - b2) * gr * gr;
      P[j] -= lr_over_bc1 * (M[j] / (sqrtf(V[j]) * inv_sqrt_bc2 + eps));
    }
    reinterpret_cast<float4*>(p)[i] = pp;
    reinterpret_cast<float4*>(m)[i] = mm;
    reinterpret_cast<float4*>(v)[i] = vv;
  }
  for (long long i = n4 * 4 + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float gr = g[i] * gscale + wd * p[i];
    const float mi = b1 * m[i] + (1.f - b1) * gr;
    const float vi = b2 * v[i] + (1.f - b2) * gr * gr;
    m[i] = mi; v[i] = vi;
    p[i] -= lr_over_bc1 * (mi / (sqrtf(vi) * inv_sqrt_bc2 + eps));
  }
}

}  // namespace mt

extern "C" int mt_adam_flat_dev(float* p, const float* g, float* m, float* v, long long n,
                                const float* hyper, void* stream) {
  if (n < 0 || !hyper) return -1005;
  if (n == 0) return 0;
  mt::launch_pdl(mt::adam_flat_dev_kernel, dim3(mt::grid_for(n / 4 + 1, 256)), dim3(256), 0, (cudaStream_t)stream, 
      p, g, m, v, n, hyper);
  return mt::cuda_rc();
}

extern "C" int mt_adam_flat(float* p, const float* g, float* m, float* v, long long n, float lr,
                            float beta1, float beta2, float eps, float wd, int step, float gscale,
                            void* stream) {
  if (step < 1 || n < 0) return -1005;
  if (n == 0) return 0;
  const double bc1 = 1.0 - pow((double)beta1, step);
  const double bc2 = 1.0 - pow((double)beta2, step);
  mt::launch_pdl(mt::adam_flat_kernel, dim3(mt::grid_for(n / 4 + 1, 256)), dim3(256), 0, (cudaStream_t)stream, 
      p, g, m, v, n, (float)(lr / bc1), (float)(1.0 / sqrt(bc2)), beta1, beta2, eps, wd, gscale);
  return mt::cuda_rc();
}
