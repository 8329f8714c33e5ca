// optim.cu — fused optimizer kernels (sm_100a, HBM-bound: 16-byte vector access,
// one pass over p/g/m/v, optional low-precision parameter write-back fused in).
//
// Replaces the absent csrc behind the reference's builders:
//   FusedAdamBuilder   multi_tensor_adam  (deepspeed/ops/adam/fused_adam.py:176-190)
//   FusedLionBuilder   multi_tensor_lion  (deepspeed/ops/lion/fused_lion.py:118-128)
//   FusedLambBuilder   lamb               (deepspeed/ops/lamb/fused_lamb.py:165)
// and the unscale/clip + fp32->bit16 copy that ZeRO does around the step
// (deepspeed/runtime/zero/stage_1_and_2.py:2117-2177).
#include "sxb_common.cuh"

namespace sxb {
namespace optim {

constexpr int kMaxTensors = 48;
constexpr int kChunk = 2048 * 4;  // elements per block-iteration
constexpr int kThreads = 512;

struct TensorList {
  void* g[kMaxTensors];
  void* p[kMaxTensors];
  void* m[kMaxTensors];
  void* v[kMaxTensors];
  long long numel[kMaxTensors];
  int blk_start[kMaxTensors + 1];
  int n;
};

struct AdamArgs {
  float lr, beta1, beta2, eps, weight_decay;
  float bc1, bc2;      // bias corrections (1 when disabled)
  float grad_scale;    // multiply gradients (1/loss_scale * clip)
  int adamw;           // 1 = decoupled weight decay
};

template <typename T>
__device__ __forceinline__ void load4(const T* p, float (&o)[4]);
template <>
__device__ __forceinline__ void load4<float>(const float* p, float (&o)[4]) {
  float4 t = *reinterpret_cast<const float4*>(p);
  o[0] = t.x; o[1] = t.y; o[2] = t.z; o[3] = t.w;
}
template <>
__device__ __forceinline__ void load4<__nv_bfloat16>(const __nv_bfloat16* p, float (&o)[4]) {
  uint2 t = *reinterpret_cast<const uint2*>(p);
  __nv_bfloat162 a = *reinterpret_cast<__nv_bfloat162*>(&t.x);
  __nv_bfloat162 b = *reinterpret_cast<__nv_bfloat162*>(&t.y);
  o[0] = __low2float(a); o[1] = __high2float(a); o[2] = __low2float(b); o[3] = __high2float(b);
}
template <>
__device__ __forceinline__ void load4<__half>(const __half* p, float (&o)[4]) {
  uint2 t = *reinterpret_cast<const uint2*>(p);
  __half2 a = *reinterpret_cast<__half2*>(&t.x);
  __half2 b = *reinterpret_cast<__half2*>(&t.y);
  o[0] = __low2float(a); o[1] = __high2float(a); o[2] = __low2float(b); o[3] = __high2float(b);
}
template <typename T>
__device__ __forceinline__ void store4(T* p, const float (&o)[4]);
template <>
__device__ __forceinline__ void store4<float>(float* p, const float (&o)[4]) {
  *reinterpret_cast<float4*>(p) = make_float4(o[0], o[1], o[2], o[3]);
}
template <>
__device__ __forceinline__ void store4<__nv_bfloat16>(__nv_bfloat16* p, const float (&o)[4]) {
  uint2 t;
  t.x = pack_bf16(o[0], o[1]);
  t.y = pack_bf16(o[2], o[3]);
  *reinterpret_cast<uint2*>(p) = t;
}
template <>
__device__ __forceinline__ void store4<__half>(__half* p, const float (&o)[4]) {
  __half2 a = __floats2half2_rn(o[0], o[1]);
  __half2 b = __floats2half2_rn(o[2], o[3]);
  uint2 t;
  t.x = *reinterpret_cast<uint32_t*>(&a);
  t.y = *reinterpret_cast<uint32_t*>(&b);
  *reinterpret_cast<uint2*>(p) = t;
}

__device__ __forceinline__ void adam_math(float& p, float g, float& m, float& v,
                                          const AdamArgs& a) {
  g *= a.grad_scale;
  if (!a.adamw) g += a.weight_decay * p;
  m = a.beta1 * m + (1.f - a.beta1) * g;
  v = a.beta2 * v + (1.f - a.beta2) * g * g;
  const float mh = m / a.bc1;
  const float vh = v / a.bc2;
  float upd = mh / (sqrtf(vh) + a.eps);
  if (a.adamw) upd += a.weight_decay * p;
  p -= a.lr * upd;
}

// Flat Adam: fp32 master p/m/v, gradient of type TG, optional low-precision copy TL.
template <typename TG, typename TL>
__global__ void __launch_bounds__(kThreads)
adam_flat_kernel(float* __restrict__ p, const TG* __restrict__ g, float* __restrict__ m,
                 float* __restrict__ v, TL* __restrict__ lp, long long n, AdamArgs a,
                 const int* __restrict__ noop, const float* __restrict__ dyn_scale) {
  if (noop && *noop) return;
  if (dyn_scale) a.grad_scale *= *dyn_scale;
  const long long n4 = n >> 2;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4;
       i += (long long)gridDim.x * blockDim.x) {
    float pp[4], gg[4], mm[4], vv[4];
    load4<float>(p + i * 4, pp);
    load4<TG>(g + i * 4, gg);
    load4<float>(m + i * 4, mm);
    load4<float>(v + i * 4, vv);
#pragma unroll
    for (int k = 0; k < 4; ++k) adam_math(pp[k], gg[k], mm[k], vv[k], a);
    store4<float>(p + i * 4, pp);
    store4<float>(m + i * 4, mm);
    store4<float>(v + i * 4, vv);
    if (lp) store4<TL>(lp + i * 4, pp);
  }
  // tail
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const long long i = (n4 << 2) + threadIdx.x;
    float pp = p[i], mm = m[i], vv = v[i];
    adam_math(pp, to_f32<TG>(g[i]), mm, vv, a);
    p[i] = pp; m[i] = mm; v[i] = vv;
    if (lp) lp[i] = from_f32<TL>(pp);
  }
}

// Multi-tensor Adam where p/m/v share dtype T (apex contract) and g has dtype T too.
template <typename T>
__global__ void __launch_bounds__(kThreads)
adam_multi_kernel(TensorList tl, AdamArgs a, const int* __restrict__ noop) {
  if (noop && *noop) return;
  // locate tensor for this block
  int lo = 0, hi = tl.n;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (tl.blk_start[mid] <= (int)blockIdx.x) lo = mid; else hi = mid;
  }
  const int t = lo;
  const long long base = (long long)(blockIdx.x - tl.blk_start[t]) * kChunk;
  const long long n = tl.numel[t];
  T* p = reinterpret_cast<T*>(tl.p[t]);
  const T* g = reinterpret_cast<const T*>(tl.g[t]);
  T* m = reinterpret_cast<T*>(tl.m[t]);
  T* v = reinterpret_cast<T*>(tl.v[t]);
  const bool aligned = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) |
                         reinterpret_cast<uintptr_t>(m) | reinterpret_cast<uintptr_t>(v)) & 15) == 0;
  const long long end = min(base + (long long)kChunk, n);
  if (aligned) {
    for (long long i = base + threadIdx.x * 4; i < end; i += blockDim.x * 4) {
      if (i + 4 <= end) {
        float pp[4], gg[4], mm[4], vv[4];
        load4<T>(p + i, pp); load4<T>(g + i, gg); load4<T>(m + i, mm); load4<T>(v + i, vv);
#pragma unroll
        for (int k = 0; k < 4; ++k) adam_math(pp[k], gg[k], mm[k], vv[k], a);
        store4<T>(p + i, pp); store4<T>(m + i, mm); store4<T>(v + i, vv);
      } else {
        for (long long j = i; j < end; ++j) {
          float pp = to_f32<T>(p[j]), mm = to_f32<T>(m[j]), vv = to_f32<T>(v[j]);
          adam_math(pp, to_f32<T>(g[j]), mm, vv, a);
          p[j] = from_f32<T>(pp); m[j] = from_f32<T>(mm); v[j] = from_f32<T>(vv);
        }
      }
    }
  } else {
    for (long long j = base + threadIdx.x; j < end; j += blockDim.x) {
      float pp = to_f32<T>(p[j]), mm = to_f32<T>(m[j]), vv = to_f32<T>(v[j]);
      adam_math(pp, to_f32<T>(g[j]), mm, vv, a);
      p[j] = from_f32<T>(pp); m[j] = from_f32<T>(mm); v[j] = from_f32<T>(vv);
    }
  }
}

// Lion (flat): p -= lr * (sign(b1*m + (1-b1)*g) + wd*p) ; m = b2*m + (1-b2)*g
template <typename TG, typename TL>
__global__ void __launch_bounds__(kThreads)
lion_flat_kernel(float* __restrict__ p, const TG* __restrict__ g, float* __restrict__ m,
                 TL* __restrict__ lp, long long n, float lr, float b1, float b2, float wd,
                 float grad_scale, const int* __restrict__ noop) {
  if (noop && *noop) return;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    float pp = p[i];
    const float gg = to_f32<TG>(g[i]) * grad_scale;
    const float mm = m[i];
    const float c = b1 * mm + (1.f - b1) * gg;
    const float s = (c > 0.f) ? 1.f : ((c < 0.f) ? -1.f : 0.f);
    pp = pp * (1.f - lr * wd) - lr * s;
    m[i] = b2 * mm + (1.f - b2) * gg;
    p[i] = pp;
    if (lp) lp[i] = from_f32<TL>(pp);
  }
}

// Adagrad (flat)
template <typename TG, typename TL>
__global__ void __launch_bounds__(kThreads)
adagrad_flat_kernel(float* __restrict__ p, const TG* __restrict__ g, float* __restrict__ h,
                    TL* __restrict__ lp, long long n, float lr, float eps, float wd,
                    float grad_scale) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    float pp = p[i];
    float gg = to_f32<TG>(g[i]) * grad_scale + wd * pp;
    const float hh = h[i] + gg * gg;
    h[i] = hh;
    pp -= lr * gg / (sqrtf(hh) + eps);
    p[i] = pp;
    if (lp) lp[i] = from_f32<TL>(pp);
  }
}

// LAMB stage 1: update m,v, write the raw update into `upd`, accumulate ||p||^2 and ||u||^2.
template <typename TG>
__global__ void __launch_bounds__(kThreads)
lamb_stage1_kernel(const float* __restrict__ p, const TG* __restrict__ g, float* __restrict__ m,
                   float* __restrict__ v, float* __restrict__ upd, long long n, AdamArgs a,
                   float* __restrict__ norms /* [2] zeroed by caller */) {
  __shared__ float red[32];
  float sp = 0.f, su = 0.f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    const float pp = p[i];
    const float gg = to_f32<TG>(g[i]) * a.grad_scale;
    const float mm = a.beta1 * m[i] + (1.f - a.beta1) * gg;
    const float vv = a.beta2 * v[i] + (1.f - a.beta2) * gg * gg;
    m[i] = mm; v[i] = vv;
    const float u = (mm / a.bc1) / (sqrtf(vv / a.bc2) + a.eps) + a.weight_decay * pp;
    upd[i] = u;
    sp += pp * pp;
    su += u * u;
  }
  sp = block_sum(sp, red);
  su = block_sum(su, red);
  if (threadIdx.x == 0) {
    atomicAdd(&norms[0], sp);
    atomicAdd(&norms[1], su);
  }
}
// LAMB stage 2: trust ratio clamp(||p||/||u||, min, max) and apply.
template <typename TL>
__global__ void __launch_bounds__(kThreads)
lamb_stage2_kernel(float* __restrict__ p, const float* __restrict__ upd, TL* __restrict__ lp,
                   long long n, float lr, float min_coeff, float max_coeff,
                   const float* __restrict__ norms, float* __restrict__ coeff_out) {
  const float pn = sqrtf(norms[0]), un = sqrtf(norms[1]);
  float coeff = 1.f;
  if (pn > 0.f && un > 0.f) coeff = fminf(fmaxf(pn / un, min_coeff), max_coeff);
  if (coeff_out && blockIdx.x == 0 && threadIdx.x == 0) *coeff_out = coeff;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    const float pp = p[i] - lr * coeff * upd[i];
    p[i] = pp;
    if (lp) lp[i] = from_f32<TL>(pp);
  }
}

// sum of squares (+ inf/nan flag) over a flat buffer: out[0] += sumsq, out[1] = nonfinite count
template <typename T>
__global__ void __launch_bounds__(kThreads)
sumsq_kernel(const T* __restrict__ x, long long n, float* __restrict__ out) {
  __shared__ float red[32];
  float s = 0.f, bad = 0.f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    const float v = to_f32<T>(x[i]);
    if (!isfinite(v)) bad += 1.f; else s += v * v;
  }
  s = block_sum(s, red);
  bad = block_sum(bad, red);
  if (threadIdx.x == 0) {
    atomicAdd(&out[0], s);
    if (bad > 0.f) atomicAdd(&out[1], bad);
  }
}

static inline int grid_for(long long n, int per_thread = 4) {
  long long b = (n / per_thread + kThreads - 1) / kThreads;
  if (b < 1) b = 1;
  if (b > 148 * 8) b = 148 * 8;
  return (int)b;
}

}  // namespace optim
}  // namespace sxb

using namespace sxb;
using namespace sxb::optim;

static AdamArgs make_args(float lr, float b1, float b2, float eps, float wd, int step, int adamw,
                          int bias_correction, float grad_scale) {
  AdamArgs a;
  a.lr = lr; a.beta1 = b1; a.beta2 = b2; a.eps = eps; a.weight_decay = wd;
  a.bc1 = bias_correction ? 1.f - powf(b1, (float)step) : 1.f;
  a.bc2 = bias_correction ? 1.f - powf(b2, (float)step) : 1.f;
  a.grad_scale = grad_scale; a.adamw = adamw;
  return a;
}

// Flat fused Adam/AdamW on fp32 master weights.
//   g_dtype: kF32 | kBF16 | kF16 ; lp may be NULL ; lp_dtype: kBF16 | kF16
//   noop: device int*, skip everything when *noop != 0 (overflow) — may be NULL
//   dyn_scale: device float* multiplied into grad_scale (e.g. clip coefficient) — may be NULL
SXB_API int sxb_adam_flat(void* p, const void* g, void* m, void* v, void* lp, long long n,
                          int g_dtype, int lp_dtype, float lr, float beta1, float beta2, float eps,
                          float weight_decay, int step, int adamw, int bias_correction,
                          float grad_scale, const void* noop, const void* dyn_scale,
                          void* stream) {
  if (n <= 0) return SXB_OK;
  if ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) |
       reinterpret_cast<uintptr_t>(m) | reinterpret_cast<uintptr_t>(v) |
       reinterpret_cast<uintptr_t>(lp)) & 15)
    return SXB_ERR_UNSUPPORTED;
  AdamArgs a = make_args(lr, beta1, beta2, eps, weight_decay, step, adamw, bias_correction,
                         grad_scale);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const int grid = grid_for(n);
  const int* np = reinterpret_cast<const int*>(noop);
  const float* ds = reinterpret_cast<const float*>(dyn_scale);
#define LAUNCH(TG, TL)                                                                   \
  adam_flat_kernel<TG, TL><<<grid, kThreads, 0, s>>>(                                    \
      (float*)p, (const TG*)g, (float*)m, (float*)v, (TL*)lp, n, a, np, ds)
  if (lp_dtype == kF16) {
    if (g_dtype == kF32) LAUNCH(float, __half);
    else if (g_dtype == kBF16) LAUNCH(__nv_bfloat16, __half);
    else if (g_dtype == kF16) LAUNCH(__half, __half);
    else return SXB_ERR_BAD_ARG;
  } else {
    if (g_dtype == kF32) LAUNCH(float, __nv_bfloat16);
    else if (g_dtype == kBF16) LAUNCH(__nv_bfloat16, __nv_bfloat16);
    else if (g_dtype == kF16) LAUNCH(__half, __nv_bfloat16);
    else return SXB_ERR_BAD_ARG;
  }
#undef LAUNCH
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

// apex-style multi tensor Adam: ptrs = [g0..gn-1, p0.., m0.., v0..] (host array of 4*n pointers)
SXB_API int sxb_multi_tensor_adam(const void* const* ptrs, const long long* numels, int n_tensors,
                                  int dtype, float lr, float beta1, float beta2, float eps,
                                  float weight_decay, int step, int adamw, int bias_correction,
                                  const void* noop, void* stream) {
  AdamArgs a = make_args(lr, beta1, beta2, eps, weight_decay, step, adamw, bias_correction, 1.f);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  for (int base = 0; base < n_tensors; base += kMaxTensors) {
    TensorList tl;
    const int cnt = (n_tensors - base < kMaxTensors) ? (n_tensors - base) : kMaxTensors;
    tl.n = cnt;
    int blocks = 0;
    for (int i = 0; i < cnt; ++i) {
      tl.g[i] = const_cast<void*>(ptrs[0 * n_tensors + base + i]);
      tl.p[i] = const_cast<void*>(ptrs[1 * n_tensors + base + i]);
      tl.m[i] = const_cast<void*>(ptrs[2 * n_tensors + base + i]);
      tl.v[i] = const_cast<void*>(ptrs[3 * n_tensors + base + i]);
      tl.numel[i] = numels[base + i];
      tl.blk_start[i] = blocks;
      blocks += (int)((numels[base + i] + kChunk - 1) / kChunk);
    }
    tl.blk_start[cnt] = blocks;
    if (blocks == 0) continue;
    const int* np = reinterpret_cast<const int*>(noop);
    if (dtype == kF32) adam_multi_kernel<float><<<blocks, kThreads, 0, s>>>(tl, a, np);
    else if (dtype == kBF16) adam_multi_kernel<__nv_bfloat16><<<blocks, kThreads, 0, s>>>(tl, a, np);
    else if (dtype == kF16) adam_multi_kernel<__half><<<blocks, kThreads, 0, s>>>(tl, a, np);
    else return SXB_ERR_BAD_ARG;
    SXB_CHECK_LAUNCH();
  }
  return SXB_OK;
}

SXB_API int sxb_lion_flat(void* p, const void* g, void* m, void* lp, long long n, int g_dtype,
                          int lp_dtype, float lr, float beta1, float beta2, float weight_decay,
                          float grad_scale, const void* noop, void* stream) {
  if (n <= 0) return SXB_OK;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const int grid = grid_for(n, 1);
  const int* np = reinterpret_cast<const int*>(noop);
#define LAUNCH(TG, TL)                                                                          \
  lion_flat_kernel<TG, TL><<<grid, kThreads, 0, s>>>((float*)p, (const TG*)g, (float*)m, (TL*)lp, \
                                                     n, lr, beta1, beta2, weight_decay, grad_scale, np)
  if (lp_dtype == kF16) {
    if (g_dtype == kF32) LAUNCH(float, __half);
    else if (g_dtype == kBF16) LAUNCH(__nv_bfloat16, __half);
    else LAUNCH(__half, __half);
  } else {
    if (g_dtype == kF32) LAUNCH(float, __nv_bfloat16);
    else if (g_dtype == kBF16) LAUNCH(__nv_bfloat16, __nv_bfloat16);
    else LAUNCH(__half, __nv_bfloat16);
  }
#undef LAUNCH
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

SXB_API int sxb_adagrad_flat(void* p, const void* g, void* h, void* lp, long long n, int g_dtype,
                             int lp_dtype, float lr, float eps, float weight_decay,
                             float grad_scale, void* stream) {
  if (n <= 0) return SXB_OK;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const int grid = grid_for(n, 1);
#define LAUNCH(TG, TL)                                                                       \
  adagrad_flat_kernel<TG, TL><<<grid, kThreads, 0, s>>>((float*)p, (const TG*)g, (float*)h,    \
                                                        (TL*)lp, n, lr, eps, weight_decay, grad_scale)
  if (lp_dtype == kF16) {
    if (g_dtype == kF32) LAUNCH(float, __half);
    else if (g_dtype == kBF16) LAUNCH(__nv_bfloat16, __half);
    else LAUNCH(__half, __half);
  } else {
    if (g_dtype == kF32) LAUNCH(float, __nv_bfloat16);
    else if (g_dtype == kBF16) LAUNCH(__nv_bfloat16, __nv_bfloat16);
    else LAUNCH(__half, __nv_bfloat16);
  }
#undef LAUNCH
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

// LAMB on one tensor.  `upd` (n floats) and `norms` (2 floats) are caller-provided scratch.
SXB_API int sxb_lamb(void* p, const void* g, void* m, void* v, void* lp, void* upd, void* norms,
                     void* coeff_out, long long n, int g_dtype, int lp_dtype, float lr, float beta1,
                     float beta2, float eps, float weight_decay, int step, int bias_correction,
                     float min_coeff, float max_coeff, float grad_scale, void* stream) {
  if (n <= 0) return SXB_OK;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  AdamArgs a = make_args(lr, beta1, beta2, eps, weight_decay, step, 1, bias_correction, grad_scale);
  cudaError_t e = cudaMemsetAsync(norms, 0, 2 * sizeof(float), s);
  if (e != cudaSuccess) return (int)e;
  const int grid = grid_for(n, 1);
  if (g_dtype == kF32)
    lamb_stage1_kernel<float><<<grid, kThreads, 0, s>>>((const float*)p, (const float*)g, (float*)m, (float*)v, (float*)upd, n, a, (float*)norms);
  else if (g_dtype == kBF16)
    lamb_stage1_kernel<__nv_bfloat16><<<grid, kThreads, 0, s>>>((const float*)p, (const __nv_bfloat16*)g, (float*)m, (float*)v, (float*)upd, n, a, (float*)norms);
  else
    lamb_stage1_kernel<__half><<<grid, kThreads, 0, s>>>((const float*)p, (const __half*)g, (float*)m, (float*)v, (float*)upd, n, a, (float*)norms);
  SXB_CHECK_LAUNCH();
  if (lp_dtype == kF16)
    lamb_stage2_kernel<__half><<<grid, kThreads, 0, s>>>((float*)p, (const float*)upd, (__half*)lp, n, lr, min_coeff, max_coeff, (const float*)norms, (float*)coeff_out);
  else
    lamb_stage2_kernel<__nv_bfloat16><<<grid, kThreads, 0, s>>>((float*)p, (const float*)upd, (__nv_bfloat16*)lp, n, lr, min_coeff, max_coeff, (const float*)norms, (float*)coeff_out);
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

// out[0] += sum(x^2 over finite x), out[1] += count(non-finite).  Caller zeroes `out`.
SXB_API int sxb_sumsq(const void* x, long long n, int dtype, void* out, void* stream) {
  if (n <= 0) return SXB_OK;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const int grid = grid_for(n, 1);
  if (dtype == kF32) sumsq_kernel<float><<<grid, kThreads, 0, s>>>((const float*)x, n, (float*)out);
  else if (dtype == kBF16) sumsq_kernel<__nv_bfloat16><<<grid, kThreads, 0, s>>>((const __nv_bfloat16*)x, n, (float*)out);
  else if (dtype == kF16) sumsq_kernel<__half><<<grid, kThreads, 0, s>>>((const __half*)x, n, (float*)out);
  else return SXB_ERR_BAD_ARG;
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}
