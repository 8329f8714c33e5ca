// elementwise.cu — RoPE, gated/bias activations, fused softmax-cross-entropy, embedding,
// dtype casts and scaled accumulation (sm_100a; all HBM-bound, 16-byte vector access).
//
// Reference behaviour:
//   gated activation  deepspeed/inference/v2/kernels/core_ops/gated_activations/gated_activation_kernels_cuda.cu:45
//   bias activation   deepspeed/inference/v2/kernels/core_ops/bias_activations/bias_activation_cuda.cu:52-92
//   rotary            deepspeed/sequence/layer.py:84-103, deepspeed/sequence/fpdt_layer.py:26-37
//   vocab-parallel CE deepspeed/sequence/cross_entropy.py:13-58
//   ragged embed      deepspeed/inference/v2/kernels/ragged_ops/embed/embed_cuda.cu:19-90
#include "sxb_common.cuh"

namespace sxb {
namespace ew {

constexpr int kThreads = 256;

template <typename T>
__device__ __forceinline__ void ld8(const T* p, float (&o)[8]) {
  uint4 t = *reinterpret_cast<const uint4*>(p);
  const T* h = reinterpret_cast<const T*>(&t);
#pragma unroll
  for (int i = 0; i < 8; ++i) o[i] = to_f32<T>(h[i]);
}
template <typename T>
__device__ __forceinline__ void st8(T* p, const float (&o)[8]) {
  uint4 t;
  T* h = reinterpret_cast<T*>(&t);
#pragma unroll
  for (int i = 0; i < 8; ++i) h[i] = from_f32<T>(o[i]);
  *reinterpret_cast<uint4*>(p) = t;
}

// ----------------------------------------------------------------------- RoPE
// x: [T, row_stride] with n_heads rotated heads of size D at the start of each row.
// cos/sin: [max_pos, D/2] fp32.  sign = +1 forward, -1 backward (inverse rotation).
template <typename T>
__global__ void __launch_bounds__(kThreads)
rope_kernel(T* __restrict__ x, const long long* __restrict__ pos, const float* __restrict__ cs,
            const float* __restrict__ sn, long long Ttok, int n_heads, int D, long long row_stride,
            float sign) {
  const int half = D >> 1;
  const int vec_per_head = half >> 3;  // 8 elements per thread-iteration
  const long long total = Ttok * n_heads * vec_per_head;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int v = (int)(i % vec_per_head);
    const long long th = i / vec_per_head;
    const int h = (int)(th % n_heads);
    const long long t = th / n_heads;
    const long long p = pos ? pos[t] : t;
    T* base = x + t * row_stride + (long long)h * D + v * 8;
    float a[8], b[8], c[8], s[8];
    ld8<T>(base, a);
    ld8<T>(base + half, b);
    const float4* cp = reinterpret_cast<const float4*>(cs + p * half + v * 8);
    const float4* sp = reinterpret_cast<const float4*>(sn + p * half + v * 8);
    float4 c0 = cp[0], c1 = cp[1], s0 = sp[0], s1 = sp[1];
    c[0] = c0.x; c[1] = c0.y; c[2] = c0.z; c[3] = c0.w; c[4] = c1.x; c[5] = c1.y; c[6] = c1.z; c[7] = c1.w;
    s[0] = s0.x; s[1] = s0.y; s[2] = s0.z; s[3] = s0.w; s[4] = s1.x; s[5] = s1.y; s[6] = s1.z; s[7] = s1.w;
    float oa[8], ob[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const float sk = sign * s[k];
      oa[k] = a[k] * c[k] - b[k] * sk;
      ob[k] = b[k] * c[k] + a[k] * sk;
    }
    st8<T>(base, oa);
    st8<T>(base + half, ob);
  }
}

// ------------------------------------------------------------------ activations
enum Act : int { kIdentity = 0, kRelu = 1, kGelu = 2, kSilu = 3 };

__device__ __forceinline__ float act_fwd(float x, int act) {
  switch (act) {
    case kRelu: return fmaxf(x, 0.f);
    case kGelu: {
      const float k0 = 0.7978845608028654f, k1 = 0.044715f;
      return 0.5f * x * (1.f + tanhf(k0 * (x + k1 * x * x * x)));
    }
    case kSilu: return x / (1.f + __expf(-x));
    default: return x;
  }
}
__device__ __forceinline__ float act_bwd(float x, int act) {
  switch (act) {
    case kRelu: return x > 0.f ? 1.f : 0.f;
    case kGelu: {
      const float k0 = 0.7978845608028654f, k1 = 0.044715f;
      const float u = k0 * (x + k1 * x * x * x);
      const float t = tanhf(u);
      return 0.5f * (1.f + t) + 0.5f * x * (1.f - t * t) * k0 * (1.f + 3.f * k1 * x * x);
    }
    case kSilu: {
      const float s = 1.f / (1.f + __expf(-x));
      return s * (1.f + x * (1.f - s));
    }
    default: return 1.f;
  }
}

// out[T,F] = act(in[:, :F] (+bias)) * (in[:, F:] (+bias))       (concatenated gate|up layout)
template <typename T>
__global__ void __launch_bounds__(kThreads)
gated_fwd_kernel(const T* __restrict__ in, const T* __restrict__ bias, T* __restrict__ out,
                 long long rows, int F, int act) {
  const int vpr = F >> 3;
  const long long total = rows * vpr;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / vpr;
    const int c = (int)(i % vpr) * 8;
    float g[8], u[8], o[8];
    ld8<T>(in + r * 2 * F + c, g);
    ld8<T>(in + r * 2 * F + F + c, u);
    if (bias) {
      float bg[8], bu[8];
      ld8<T>(bias + c, bg);
      ld8<T>(bias + F + c, bu);
#pragma unroll
      for (int k = 0; k < 8; ++k) { g[k] += bg[k]; u[k] += bu[k]; }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) o[k] = act_fwd(g[k], act) * u[k];
    st8<T>(out + r * F + c, o);
  }
}
template <typename T>
__global__ void __launch_bounds__(kThreads)
gated_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ in, T* __restrict__ din,
                 long long rows, int F, int act) {
  const int vpr = F >> 3;
  const long long total = rows * vpr;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / vpr;
    const int c = (int)(i % vpr) * 8;
    float g[8], u[8], d[8], dg[8], du[8];
    ld8<T>(in + r * 2 * F + c, g);
    ld8<T>(in + r * 2 * F + F + c, u);
    ld8<T>(dy + r * F + c, d);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      dg[k] = d[k] * u[k] * act_bwd(g[k], act);
      du[k] = d[k] * act_fwd(g[k], act);
    }
    st8<T>(din + r * 2 * F + c, dg);
    st8<T>(din + r * 2 * F + F + c, du);
  }
}
// y = act(x + bias)  (in place allowed)
template <typename T>
__global__ void __launch_bounds__(kThreads)
bias_act_kernel(const T* __restrict__ x, const T* __restrict__ bias, T* __restrict__ y,
                long long rows, int C, int act) {
  const int vpr = C >> 3;
  const long long total = rows * vpr;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / vpr;
    const int c = (int)(i % vpr) * 8;
    float v[8];
    ld8<T>(x + r * C + c, v);
    if (bias) {
      float b[8];
      ld8<T>(bias + c, b);
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] += b[k];
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = act_fwd(v[k], act);
    st8<T>(y + r * C + c, v);
  }
}
// dx = dy * act'(x + bias)
template <typename T>
__global__ void __launch_bounds__(kThreads)
bias_act_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ x, const T* __restrict__ bias,
                    T* __restrict__ dx, long long rows, int C, int act) {
  const int vpr = C >> 3;
  const long long total = rows * vpr;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / vpr;
    const int c = (int)(i % vpr) * 8;
    float v[8], d[8];
    ld8<T>(x + r * C + c, v);
    ld8<T>(dy + r * C + c, d);
    if (bias) {
      float b[8];
      ld8<T>(bias + c, b);
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] += b[k];
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) d[k] *= act_bwd(v[k], act);
    st8<T>(dx + r * C + c, d);
  }
}

// ------------------------------------------------------------- cross entropy
// One CTA per token row.  loss[row] = lse - logit[label]; logits are overwritten by
// dlogits = (softmax - onehot) * (*grad_scale) so the [T,V] tensor is touched twice
// and never duplicated.  Rows with label == ignore_index get loss 0 / grad 0.
template <typename T>
__global__ void __launch_bounds__(1024)
xent_kernel(T* __restrict__ logits, const long long* __restrict__ labels,
            float* __restrict__ loss, float* __restrict__ lse_out, int V, long long ld,
            long long ignore_index, const float* __restrict__ grad_scale, int write_grad) {
  __shared__ float red[32];
  const long long row = blockIdx.x;
  T* lr = logits + row * ld;
  const long long label = labels[row];
  const int nvec = V >> 3;
  float mx = -INFINITY, sum = 0.f;
  for (int vi = threadIdx.x; vi < nvec; vi += blockDim.x) {
    float v[8];
    ld8<T>(lr + vi * 8, v);
    float lm = v[0];
#pragma unroll
    for (int k = 1; k < 8; ++k) lm = fmaxf(lm, v[k]);
    const float nm = fmaxf(mx, lm);
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) acc += __expf(v[k] - nm);
    sum = sum * __expf(mx - nm) + acc;
    mx = nm;
  }
  for (int j = (nvec << 3) + threadIdx.x; j < V; j += blockDim.x) {
    const float v = to_f32<T>(lr[j]);
    const float nm = fmaxf(mx, v);
    sum = sum * __expf(mx - nm) + __expf(v - nm);
    mx = nm;
  }
  const float gmx = block_max(mx, red);
  sum = (mx == -INFINITY) ? 0.f : sum * __expf(mx - gmx);
  sum = block_sum(sum, red);
  const float lse = gmx + __logf(sum);
  const bool ignored = (label == ignore_index);
  if (threadIdx.x == 0) {
    loss[row] = ignored ? 0.f : (lse - to_f32<T>(lr[label]));
    if (lse_out) lse_out[row] = lse;
  }
  if (!write_grad) return;
  __syncthreads();
  const float gs = ignored ? 0.f : (grad_scale ? *grad_scale : 1.f);
  for (int vi = threadIdx.x; vi < nvec; vi += blockDim.x) {
    float v[8];
    ld8<T>(lr + vi * 8, v);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float pr = __expf(v[k] - lse);
      if ((long long)(vi * 8 + k) == label) pr -= 1.f;
      v[k] = pr * gs;
    }
    st8<T>(lr + vi * 8, v);
  }
  for (int j = (nvec << 3) + threadIdx.x; j < V; j += blockDim.x) {
    float pr = __expf(to_f32<T>(lr[j]) - lse);
    if (j == label) pr -= 1.f;
    lr[j] = from_f32<T>(pr * gs);
  }
}

// ------------------------------------------------------------------ embedding
template <typename T>
__global__ void __launch_bounds__(kThreads)
embed_fwd_kernel(const long long* __restrict__ ids, const T* __restrict__ table,
                 T* __restrict__ out, long long ntok, int H) {
  const int vpr = H >> 3;
  const long long total = ntok * vpr;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long t = i / vpr;
    const int c = (int)(i % vpr) * 8;
    *reinterpret_cast<uint4*>(out + t * H + c) =
        *reinterpret_cast<const uint4*>(table + ids[t] * H + c);
  }
}
// dtable[ids[t]] += dout[t]   (bf16x2 / half2 / float atomics)
template <typename T>
__global__ void __launch_bounds__(kThreads)
embed_bwd_kernel(const long long* __restrict__ ids, const T* __restrict__ dout,
                 T* __restrict__ dtable, long long ntok, int H) {
  const int ppr = H >> 1;
  const long long total = ntok * ppr;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long t = i / ppr;
    const int c = (int)(i % ppr) * 2;
    if constexpr (sizeof(T) == 4) {
      atomicAdd(reinterpret_cast<float*>(dtable) + ids[t] * H + c, to_f32<T>(dout[t * H + c]));
      atomicAdd(reinterpret_cast<float*>(dtable) + ids[t] * H + c + 1, to_f32<T>(dout[t * H + c + 1]));
    } else if constexpr (std::is_same<T, __nv_bfloat16>::value) {
      atomicAdd(reinterpret_cast<__nv_bfloat162*>(dtable + ids[t] * H + c),
                *reinterpret_cast<const __nv_bfloat162*>(dout + t * H + c));
    } else {
      atomicAdd(reinterpret_cast<__half2*>(dtable + ids[t] * H + c),
                *reinterpret_cast<const __half2*>(dout + t * H + c));
    }
  }
}

// ------------------------------------------------------------ cast / axpy
// dst = (accumulate ? dst : 0) + scale * src with independent dtypes.
template <typename TS, typename TD>
__global__ void __launch_bounds__(kThreads)
scale_cast_kernel(const TS* __restrict__ src, TD* __restrict__ dst, long long n, float scale,
                  int accumulate) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    float v = scale * to_f32<TS>(src[i]);
    if (accumulate) v += to_f32<TD>(dst[i]);
    dst[i] = from_f32<TD>(v);
  }
}

static inline int grid_for(long long work) {
  long long b = (work + kThreads - 1) / kThreads;
  if (b < 1) b = 1;
  if (b > 148 * 16) b = 148 * 16;
  return (int)b;
}

}  // namespace ew
}  // namespace sxb

#include <type_traits>
using namespace sxb;
using namespace sxb::ew;

#define DISPATCH_16(dtype, CALL)                          \
  do {                                                    \
    if (dtype == kBF16) { using T = __nv_bfloat16; CALL; } \
    else if (dtype == kF16) { using T = __half; CALL; }   \
    else return SXB_ERR_BAD_ARG;                          \
  } while (0)

SXB_API int sxb_rope(void* x, const void* positions, const void* cos_tab, const void* sin_tab,
                     long long tokens, int n_heads, int head_dim, long long row_stride,
                     int inverse, int dtype, void* stream) {
  if (tokens <= 0 || n_heads <= 0) return SXB_OK;
  if (head_dim % 16) return SXB_ERR_UNSUPPORTED;
  cudaStream_t s = (cudaStream_t)stream;
  const long long work = tokens * n_heads * (head_dim / 16);
  DISPATCH_16(dtype, (rope_kernel<T><<<grid_for(work), kThreads, 0, s>>>(
                         (T*)x, (const long long*)positions, (const float*)cos_tab,
                         (const float*)sin_tab, tokens, n_heads, head_dim, row_stride,
                         inverse ? -1.f : 1.f)));
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

SXB_API int sxb_gated_act_fwd(const void* in, const void* bias, void* out, long long rows, int F,
                              int act, int dtype, void* stream) {
  if (rows <= 0) return SXB_OK;
  if (F % 8) return SXB_ERR_UNSUPPORTED;
  cudaStream_t s = (cudaStream_t)stream;
  DISPATCH_16(dtype, (gated_fwd_kernel<T><<<grid_for(rows * (F / 8)), kThreads, 0, s>>>(
                         (const T*)in, (const T*)bias, (T*)out, rows, F, act)));
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}
SXB_API int sxb_gated_act_bwd(const void* dy, const void* in, void* din, long long rows, int F,
                              int act, int dtype, void* stream) {
  if (rows <= 0) return SXB_OK;
  if (F % 8) return SXB_ERR_UNSUPPORTED;
  cudaStream_t s = (cudaStream_t)stream;
  DISPATCH_16(dtype, (gated_bwd_kernel<T><<<grid_for(rows * (F / 8)), kThreads, 0, s>>>(
                         (const T*)dy, (const T*)in, (T*)din, rows, F, act)));
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}
SXB_API int sxb_bias_act(const void* x, const void* bias, void* y, long long rows, int C, int act,
                         int dtype, void* stream) {
  if (rows <= 0) return SXB_OK;
  if (C % 8) return SXB_ERR_UNSUPPORTED;
  cudaStream_t s = (cudaStream_t)stream;
  DISPATCH_16(dtype, (bias_act_kernel<T><<<grid_for(rows * (C / 8)), kThreads, 0, s>>>(
                         (const T*)x, (const T*)bias, (T*)y, rows, C, act)));
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}
SXB_API int sxb_bias_act_bwd(const void* dy, const void* x, const void* bias, void* dx,
                             long long rows, int C, int act, int dtype, void* stream) {
  if (rows <= 0) return SXB_OK;
  if (C % 8) return SXB_ERR_UNSUPPORTED;
  cudaStream_t s = (cudaStream_t)stream;
  DISPATCH_16(dtype, (bias_act_bwd_kernel<T><<<grid_for(rows * (C / 8)), kThreads, 0, s>>>(
                         (const T*)dy, (const T*)x, (const T*)bias, (T*)dx, rows, C, act)));
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

// logits [rows, V] (row pitch ld) are overwritten with the gradient when write_grad != 0.
SXB_API int sxb_cross_entropy(void* logits, const void* labels, void* loss, void* lse,
                              long long rows, int V, long long ld, long long ignore_index,
                              const void* grad_scale, int write_grad, int dtype, void* stream) {
  if (rows <= 0) return SXB_OK;
  if (ld % 8 || (reinterpret_cast<uintptr_t>(logits) & 15)) return SXB_ERR_UNSUPPORTED;
  cudaStream_t s = (cudaStream_t)stream;
  const int threads = V >= 32768 ? 1024 : 256;
  DISPATCH_16(dtype, (xent_kernel<T><<<(unsigned)rows, threads, 0, s>>>(
                         (T*)logits, (const long long*)labels, (float*)loss, (float*)lse, V, ld,
                         ignore_index, (const float*)grad_scale, write_grad)));
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

SXB_API int sxb_embedding_fwd(const void* ids, const void* table, void* out, long long ntok, int H,
                              int dtype, void* stream) {
  if (ntok <= 0) return SXB_OK;
  if (H % 8) return SXB_ERR_UNSUPPORTED;
  cudaStream_t s = (cudaStream_t)stream;
  DISPATCH_16(dtype, (embed_fwd_kernel<T><<<grid_for(ntok * (H / 8)), kThreads, 0, s>>>(
                         (const long long*)ids, (const T*)table, (T*)out, ntok, H)));
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}
SXB_API int sxb_embedding_bwd(const void* ids, const void* dout, void* dtable, long long ntok,
                              int H, int dtype, void* stream) {
  if (ntok <= 0) return SXB_OK;
  if (H % 2) return SXB_ERR_UNSUPPORTED;
  cudaStream_t s = (cudaStream_t)stream;
  DISPATCH_16(dtype, (embed_bwd_kernel<T><<<grid_for(ntok * (H / 2)), kThreads, 0, s>>>(
                         (const long long*)ids, (const T*)dout, (T*)dtable, ntok, H)));
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

SXB_API int sxb_scale_cast(const void* src, void* dst, long long n, int src_dtype, int dst_dtype,
                           float scale, int accumulate, void* stream) {
  if (n <= 0) return SXB_OK;
  cudaStream_t s = (cudaStream_t)stream;
  const int g = grid_for(n);
#define SC(TS, TD) scale_cast_kernel<TS, TD><<<g, kThreads, 0, s>>>((const TS*)src, (TD*)dst, n, scale, accumulate)
  if (src_dtype == kBF16 && dst_dtype == kF32) SC(__nv_bfloat16, float);
  else if (src_dtype == kF32 && dst_dtype == kBF16) SC(float, __nv_bfloat16);
  else if (src_dtype == kF16 && dst_dtype == kF32) SC(__half, float);
  else if (src_dtype == kF32 && dst_dtype == kF16) SC(float, __half);
  else if (src_dtype == kF32 && dst_dtype == kF32) SC(float, float);
  else if (src_dtype == kBF16 && dst_dtype == kBF16) SC(__nv_bfloat16, __nv_bfloat16);
  else if (src_dtype == kF16 && dst_dtype == kF16) SC(__half, __half);
  else return SXB_ERR_BAD_ARG;
#undef SC
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}
