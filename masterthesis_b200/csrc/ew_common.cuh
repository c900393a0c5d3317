// Shared helpers of the HBM-bound kernels: 16-byte (8 x bf16) vector access, reflection index
// maps, launch sizing.  All activation tensors are NHWC bf16 with C % 8 == 0.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

namespace mt {

// ---- programmatic dependent launch (PDL).  A kernel launched with the attribute may start (prologue, block
// scheduling) while its predecessor in the stream is still draining; it must execute pdl_wait() before its
// first access to global memory (the wait returns once the predecessor has completed and its writes are
// visible).  pdl_launch() lets the NEXT kernel's blocks be scheduled from that point on.  Both are no-ops for
// a kernel launched without the attribute.  MT_PDL=0 launches everything with plain stream ordering.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
static inline bool pdl_enabled() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("MT_PDL"); v = e ? (atoi(e) != 0) : 1; }
  return v != 0;
}
template <bool PDL = true, typename K, typename... Args>
static inline cudaError_t launch_pdl(K kern, dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = (PDL && pdl_enabled()) ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, args...);
}
// plain stream ordering for a kernel that follows a LONG kernel on a side stream (the weight-gradient
// unpack behind igemm_wgrad_kernel): launched early, its grid of up to 2 368 CTAs would sit resident in
// griddepcontrol.wait for the whole weight-gradient kernel and take the thread slots the main stream's
// norm kernels run in (measured: the gain of PDL on the main chain disappeared)
template <typename K, typename... Args>
static inline cudaError_t launch_plain(K kern, dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
  return launch_pdl<false>(kern, grid, block, smem, st, args...);
}

static constexpr int NUM_SMS = 148;

struct alignas(16) bf16x8 {
  __nv_bfloat162 v[4];
};

__device__ __forceinline__ void load8(const __nv_bfloat16* p, float (&f)[8]) {
  const uint4 r = *reinterpret_cast<const uint4*>(p);
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
// raw 16-byte streaming load (read-only path, no L1 allocation) + conversion, kept apart so that a
// thread can issue several loads before it touches the first result (memory-level parallelism:
// ~35 KB must be in flight per SM to saturate HBM3e)
__device__ __forceinline__ uint4 ldg16(const __nv_bfloat16* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void cvt8(const uint4& r, float (&f)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ void store8(__nv_bfloat16* p, const float (&f)[8]) {
  uint4 r;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  *reinterpret_cast<uint4*>(p) = r;
}
__device__ __forceinline__ void load8f(const float* p, float (&f)[8]) {
  const float4 a = *reinterpret_cast<const float4*>(p);
  const float4 b = *reinterpret_cast<const float4*>(p + 4);
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w;
  f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}

// nn.ReflectionPad2d index map: padded coordinate j in [0, L+2p) -> interior index in [0, L)
__device__ __forceinline__ int reflect_src(int j, int L, int p) {
  int i = j - p;
  if (i < 0) i = -i;
  if (i >= L) i = 2 * (L - 1) - i;
  return i;
}
// Adjoint of the map above: padded coordinates whose value came from interior index i.
// Returns the count (1..3) and fills src[] with padded coordinates.
__device__ __forceinline__ int reflect_fold_sources(int i, int L, int p, int (&src)[3]) {
  int n = 0;
  src[n++] = i + p;
  if (i >= 1 && i <= p) src[n++] = p - i;                      // top / left border
  const int k = L - 2 - i;                                     // bottom / right border offset
  if (k >= 0 && k < p) src[n++] = L + p + k;
  return n;
}

// compile-time variants for the hot loops (a run-time `act` costs a branch ladder per element)
template <int ACT>
__device__ __forceinline__ float act_fwd_t(float v, float slope) {
  if (ACT == 1) return fmaxf(v, 0.f);
  if (ACT == 2) return v > 0.f ? v : v * slope;
  if (ACT == 3) return tanhf(v);
  return v;
}
template <int ACT>
__device__ __forceinline__ float act_grad_pre_t(float y, float slope) {
  if (ACT == 1) return y > 0.f ? 1.f : 0.f;
  if (ACT == 2) return y > 0.f ? 1.f : slope;
  return 1.f;
}
// p / W for 0 <= p < 2^24 with a precomputed reciprocal (magic = ceil(2^40 / W))
struct FastDiv {
  unsigned long long magic;
  int d;
};
static inline FastDiv make_fastdiv(int d) {
  FastDiv f;
  f.d = d;
  f.magic = ((1ull << 40) + (unsigned long long)d - 1) / (unsigned long long)d;
  return f;
}
__device__ __forceinline__ int fast_div(int p, const FastDiv& f) {
  return (int)(((unsigned long long)(unsigned)p * f.magic) >> 40);
}

__device__ __forceinline__ float act_fwd(float v, int act, float slope) {
  if (act == 1) return fmaxf(v, 0.f);
  if (act == 2) return v > 0.f ? v : v * slope;
  if (act == 3) return tanhf(v);
  return v;
}
// derivative expressed through the PRE-activation sign (relu / lrelu) — y is the pre-activation
__device__ __forceinline__ float act_grad_pre(float y, int act, float slope) {
  if (act == 1) return y > 0.f ? 1.f : 0.f;
  if (act == 2) return y > 0.f ? 1.f : slope;
  return 1.f;
}

// Flat-index kernels decompose a 32-bit index (a 64-bit division costs ~5x a 32-bit one and these
// kernels do three per element): more than 2^31 work items yields grid 0, i.e. a launch error.
static inline int grid_for(long long work_items, int block, int max_waves = 8) {
  if (work_items >= (1ll << 31)) return 0;
  long long b = (work_items + block - 1) / block;
  const long long cap = (long long)NUM_SMS * max_waves;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}
static inline int cuda_rc() {
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : -(int)e;
}

}  // namespace mt
