// norm.cu — RMSNorm / LayerNorm forward + backward with fused residual add (sm_100a).
// One CTA per row in forward (16-byte loads, values kept in registers between the
// reduction and the normalisation so x is read from HBM once); backward is persistent
// over rows with per-CTA weight-gradient partials in shared memory, flushed to a
// [grid, H] fp32 scratch that a second kernel reduces.
//
// Reference behaviour: deepspeed/inference/v2/kernels/core_ops/cuda_rms_norm/rms_norm_cuda.cu:19,84
// (rms_norm / pre_rms_norm) and cuda_layer_norm/layer_norm_cuda.cu:31,220 (fused_ln /
// fused_residual_ln) — forward only there; the training backward is new.
#include "sxb_common.cuh"

namespace sxb {
namespace norm {

constexpr int kThreads = 256;
constexpr int kMaxVec = 8;  // per-thread 8-element vectors kept in registers => H <= 256*8*8 = 16384

template <typename T>
struct Vec8 {
  T v[8];
};
template <typename T>
__device__ __forceinline__ void ld8(const T* p, float (&o)[8]) {
  if constexpr (sizeof(T) == 2) {
    uint4 t = *reinterpret_cast<const uint4*>(p);
    const T* h = reinterpret_cast<const T*>(&t);
#pragma unroll
    for (int i = 0; i < 8; ++i) o[i] = to_f32<T>(h[i]);
  } else {
    float4 a = *reinterpret_cast<const float4*>(p);
    float4 b = *reinterpret_cast<const float4*>(p + 4);
    o[0] = a.x; o[1] = a.y; o[2] = a.z; o[3] = a.w; o[4] = b.x; o[5] = b.y; o[6] = b.z; o[7] = b.w;
  }
}
template <typename T>
__device__ __forceinline__ void st8(T* p, const float (&o)[8]) {
  if constexpr (sizeof(T) == 2) {
    uint4 t;
    T* h = reinterpret_cast<T*>(&t);
#pragma unroll
    for (int i = 0; i < 8; ++i) h[i] = from_f32<T>(o[i]);
    *reinterpret_cast<uint4*>(p) = t;
  } else {
    *reinterpret_cast<float4*>(p) = make_float4(o[0], o[1], o[2], o[3]);
    *reinterpret_cast<float4*>(p + 4) = make_float4(o[4], o[5], o[6], o[7]);
  }
}

// y = norm(x [+ res]) * w (+ b).  If res != null also writes h = x + res to res_out.
// kLayer: LayerNorm (mean subtraction + bias) vs RMSNorm.
template <typename T, bool kLayer, int ITERS>
__global__ void __launch_bounds__(kThreads)
norm_fwd_kernel(const T* __restrict__ x, const T* __restrict__ res, const T* __restrict__ w,
                const T* __restrict__ b, T* __restrict__ y, T* __restrict__ res_out,
                float* __restrict__ mean_out, float* __restrict__ rstd_out, int H, float eps) {
  __shared__ float red[32];
  const long long row = blockIdx.x;
  const T* xr = x + row * H;
  const int nvec = H >> 3;
  float vals[ITERS][8];
  float s = 0.f, ss = 0.f;
#pragma unroll
  for (int it = 0; it < ITERS; ++it) {
    const int vi = threadIdx.x + it * kThreads;
    if (vi < nvec) {
      ld8<T>(xr + vi * 8, vals[it]);
      if (res) {
        float r[8];
        ld8<T>(res + row * H + vi * 8, r);
#pragma unroll
        for (int k = 0; k < 8; ++k) vals[it][k] += r[k];
        if (res_out) {
          // round-trip through T so the normalised value matches what backward re-reads
          st8<T>(res_out + row * H + vi * 8, vals[it]);
          ld8<T>(res_out + row * H + vi * 8, vals[it]);
        }
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        s += vals[it][k];
        ss += vals[it][k] * vals[it][k];
      }
    }
  }
  float mean = 0.f, rstd;
  if (kLayer) {
    s = block_sum(s, red);
    mean = s / H;
    float var = 0.f;
#pragma unroll
    for (int it = 0; it < ITERS; ++it) {
      const int vi = threadIdx.x + it * kThreads;
      if (vi < nvec) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const float d = vals[it][k] - mean;
          var += d * d;
        }
      }
    }
    var = block_sum(var, red);
    rstd = rsqrtf(var / H + eps);
  } else {
    ss = block_sum(ss, red);
    rstd = rsqrtf(ss / H + eps);
  }
  if (threadIdx.x == 0) {
    if (rstd_out) rstd_out[row] = rstd;
    if (kLayer && mean_out) mean_out[row] = mean;
  }
#pragma unroll
  for (int it = 0; it < ITERS; ++it) {
    const int vi = threadIdx.x + it * kThreads;
    if (vi < nvec) {
      float wv[8], o[8];
      ld8<T>(w + vi * 8, wv);
      if (kLayer && b) {
        float bv[8];
        ld8<T>(b + vi * 8, bv);
#pragma unroll
        for (int k = 0; k < 8; ++k) o[k] = (vals[it][k] - mean) * rstd * wv[k] + bv[k];
      } else {
#pragma unroll
        for (int k = 0; k < 8; ++k) o[k] = (vals[it][k] - mean) * rstd * wv[k];
      }
      st8<T>(y + row * H + vi * 8, o);
    }
  }
}

// Backward.  dx = rstd*(g - xhat*mean(g*xhat) [- mean(g)])  with g = dy*w, xhat=(x-mean)*rstd.
// Optionally adds an incoming residual gradient `dres` into dx (pre-norm blocks).
// Weight/bias grads: per-CTA partial rows in `partial` ([gridDim.x, 2, H] fp32).
template <typename T, bool kLayer, int ITERS>
__global__ void __launch_bounds__(kThreads)
norm_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ x, const T* __restrict__ w,
                const float* __restrict__ mean_in, const float* __restrict__ rstd_in,
                const T* __restrict__ dres, T* __restrict__ dx, float* __restrict__ partial,
                long long rows, int H) {
  extern __shared__ float sm[];
  __shared__ float red[32];
  float* dw_acc = sm;       // [H]
  float* db_acc = sm + H;   // [H] (LayerNorm only)
  const int nvec = H >> 3;
  for (int i = threadIdx.x; i < H * (kLayer ? 2 : 1); i += kThreads) sm[i] = 0.f;
  __syncthreads();
  for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
    const float rstd = rstd_in[row];
    const float mean = kLayer ? mean_in[row] : 0.f;
    float g[ITERS][8], xh[ITERS][8];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int it = 0; it < ITERS; ++it) {
      const int vi = threadIdx.x + it * kThreads;
      if (vi < nvec) {
        float dyv[8], wv[8];
        ld8<T>(dy + row * H + vi * 8, dyv);
        ld8<T>(x + row * H + vi * 8, xh[it]);
        ld8<T>(w + vi * 8, wv);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          xh[it][k] = (xh[it][k] - mean) * rstd;
          g[it][k] = dyv[k] * wv[k];
          s1 += g[it][k] * xh[it][k];
          s2 += g[it][k];
          dw_acc[vi * 8 + k] += dyv[k] * xh[it][k];
          if (kLayer) db_acc[vi * 8 + k] += dyv[k];
        }
      }
    }
    s1 = block_sum(s1, red) / H;
    if (kLayer) s2 = block_sum(s2, red) / H; else s2 = 0.f;
#pragma unroll
    for (int it = 0; it < ITERS; ++it) {
      const int vi = threadIdx.x + it * kThreads;
      if (vi < nvec) {
        float o[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) o[k] = rstd * (g[it][k] - xh[it][k] * s1 - s2);
        if (dres) {
          float r[8];
          ld8<T>(dres + row * H + vi * 8, r);
#pragma unroll
          for (int k = 0; k < 8; ++k) o[k] += r[k];
        }
        st8<T>(dx + row * H + vi * 8, o);
      }
    }
  }
  __syncthreads();
  float* prow = partial + (long long)blockIdx.x * (kLayer ? 2 : 1) * H;
  for (int i = threadIdx.x; i < H * (kLayer ? 2 : 1); i += kThreads) prow[i] = sm[i];
}

// out[j] (+)= sum_b partial[b, j]
template <typename T>
__global__ void reduce_partial_kernel(const float* __restrict__ partial, int nblk, int width,
                                      int count, T* __restrict__ out, int accumulate) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= count) return;
  float s = 0.f;
  for (int b = 0; b < nblk; ++b) s += partial[(long long)b * width + j];
  if (accumulate) s += to_f32<T>(out[j]);
  out[j] = from_f32<T>(s);
}

}  // namespace norm
}  // namespace sxb

using namespace sxb;
using namespace sxb::norm;

static bool norm_ok(int H) { return H > 0 && (H % 8) == 0 && H <= kThreads * kMaxVec * 8; }

// Number of fp32 scratch elements sxb_*norm_bwd needs for (rows, H).
SXB_API long long sxb_norm_bwd_scratch_elems(long long rows, int H, int layer) {
  long long grid = rows < 148 * 2 ? rows : 148 * 2;
  return grid * (layer ? 2 : 1) * H;
}

template <bool kLayer>
static int norm_fwd(const void* x, const void* res, const void* w, const void* b, void* y,
                    void* res_out, void* mean, void* rstd, long long rows, int H, float eps,
                    int dtype, cudaStream_t s) {
  if (rows <= 0) return SXB_OK;
  if (!norm_ok(H)) return SXB_ERR_UNSUPPORTED;
  const int iters = (H / 8 + kThreads - 1) / kThreads;
#define LI(T, I)                                                                                  \
  norm_fwd_kernel<T, kLayer, I><<<(unsigned)rows, kThreads, 0, s>>>(                              \
      (const T*)x, (const T*)res, (const T*)w, (const T*)b, (T*)y, (T*)res_out, (float*)mean,      \
      (float*)rstd, H, eps)
#define L(T)                                                                                      \
  do {                                                                                            \
    if (iters <= 1) LI(T, 1); else if (iters <= 2) LI(T, 2); else if (iters <= 4) LI(T, 4);       \
    else LI(T, 8);                                                                                \
  } while (0)
  if (dtype == kBF16) L(__nv_bfloat16);
  else if (dtype == kF16) L(__half);
  else if (dtype == kF32) L(float);
  else return SXB_ERR_BAD_ARG;
#undef L
#undef LI
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

template <bool kLayer>
static int norm_bwd(const void* dy, const void* x, const void* w, const void* mean,
                    const void* rstd, const void* dres, void* dx, void* dw, void* db,
                    void* scratch, long long rows, int H, int dtype, int accumulate_dw,
                    cudaStream_t s) {
  if (rows <= 0) return SXB_OK;
  if (!norm_ok(H)) return SXB_ERR_UNSUPPORTED;
  const int grid = (int)(rows < 148 * 2 ? rows : 148 * 2);
  const int width = (kLayer ? 2 : 1) * H;
  const size_t smem = (size_t)width * sizeof(float);
  const int iters = (H / 8 + kThreads - 1) / kThreads;
#define LI(T, I)                                                                                  \
  do {                                                                                            \
    if (smem > 48 * 1024)                                                                         \
      cudaFuncSetAttribute(norm_bwd_kernel<T, kLayer, I>,                                         \
                           cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);               \
    norm_bwd_kernel<T, kLayer, I><<<grid, kThreads, smem, s>>>(                                   \
        (const T*)dy, (const T*)x, (const T*)w, (const float*)mean, (const float*)rstd,           \
        (const T*)dres, (T*)dx, (float*)scratch, rows, H);                                        \
  } while (0)
#define L(T)                                                                                      \
  do {                                                                                            \
    if (iters <= 1) LI(T, 1); else if (iters <= 2) LI(T, 2); else if (iters <= 4) LI(T, 4);       \
    else LI(T, 8);                                                                                \
    reduce_partial_kernel<T><<<(H + 255) / 256, 256, 0, s>>>((const float*)scratch, grid, width,  \
                                                             H, (T*)dw, accumulate_dw);              \
    if (kLayer && db)                                                                             \
      reduce_partial_kernel<T><<<(H + 255) / 256, 256, 0, s>>>((const float*)scratch + H, grid,   \
                                                               width, H, (T*)db, accumulate_dw);     \
  } while (0)
  if (dtype == kBF16) L(__nv_bfloat16);
  else if (dtype == kF16) L(__half);
  else if (dtype == kF32) L(float);
  else return SXB_ERR_BAD_ARG;
#undef L
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

SXB_API int sxb_rmsnorm_fwd(const void* x, const void* res, const void* w, void* y, void* res_out,
                            void* rstd, long long rows, int H, float eps, int dtype,
                            void* stream) {
  return norm_fwd<false>(x, res, w, nullptr, y, res_out, nullptr, rstd, rows, H, eps, dtype,
                         (cudaStream_t)stream);
}
SXB_API int sxb_rmsnorm_bwd(const void* dy, const void* x, const void* w, const void* rstd,
                            const void* dres, void* dx, void* dw, void* scratch, long long rows,
                            int H, int dtype, int accumulate_dw, void* stream) {
  return norm_bwd<false>(dy, x, w, nullptr, rstd, dres, dx, dw, nullptr, scratch, rows, H, dtype,
                         accumulate_dw, (cudaStream_t)stream);
}
SXB_API int sxb_layernorm_fwd(const void* x, const void* res, const void* w, const void* b,
                              void* y, void* res_out, void* mean, void* rstd, long long rows,
                              int H, float eps, int dtype, void* stream) {
  return norm_fwd<true>(x, res, w, b, y, res_out, mean, rstd, rows, H, eps, dtype,
                        (cudaStream_t)stream);
}
SXB_API int sxb_layernorm_bwd(const void* dy, const void* x, const void* w, const void* mean,
                              const void* rstd, const void* dres, void* dx, void* dw, void* db,
                              void* scratch, long long rows, int H, int dtype, int accumulate_dw,
                              void* stream) {
  return norm_bwd<true>(dy, x, w, mean, rstd, dres, dx, dw, db, scratch, rows, H, dtype,
                        accumulate_dw, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------------
// RMSNorm backward for callers that must not allocate.
//
// Same computation as sxb_rmsnorm_bwd, but the per-CTA weight-gradient partials live in a scratch
// array that is part of the library image instead of a buffer handed in by the caller.  It is meant
// for code that runs when device memory can no longer be asked for: a training job whose memory plan
// has already handed every remaining byte to its parameter / optimizer pools (the 8B single-GPU
// configuration runs within a few GB of the 180 GB limit), a serving process that gave the whole device
// to its KV-cache pool at start-up, or a recovery path that is running precisely because an allocation
// just failed.  Such callers still own their inputs and outputs; only the scratch is ours.
//
// Limit: rows are unlimited, H <= SXB_NOALLOC_NORM_MAX_H (8192); the array holds 296 x 8192 floats
// (about 9.3 MiB) on every GPU that loads the library, for the life of the process.
// Sharing: there is ONE array per device.  Every call overwrites the part it uses before reading it
// (each CTA stores its full partial row, the reduction reads only rows written by this launch), so
// nothing needs to start at zero and stale contents of earlier calls are harmless.  Two calls that
// overlap on the GPU (different streams, or concurrent host threads on one device) race on the array
// and must be serialised by the caller.
// Set-up: the first call on each device resolves the array's device address with
// cudaGetSymbolAddress while that device is current and caches it; make that first call before any
// stream capture begins AND before the process commits the rest of the device's memory: with CUDA's
// lazy module loading (the default) the array only receives its 9.3 MiB of backing store when this
// file's module is first loaded on that device — by that first call or by any other norm kernel — so a
// first call made after memory has run out can itself fail with an out-of-memory error.  Later calls
// only ask which device is current (cudaGetDevice) and launch.
constexpr int SXB_NOALLOC_NORM_MAX_H = 8192;
constexpr int kNoAllocMaxGrid = 148 * 2;
constexpr long long kNoAllocScratchElems = (long long)kNoAllocMaxGrid * SXB_NOALLOC_NORM_MAX_H;
static __device__ __align__(256) float g_noalloc_norm_scratch[kNoAllocScratchElems];

namespace {
constexpr int kMaxDevices = 64;
float* g_noalloc_ptr[kMaxDevices] = {nullptr};
}  // namespace

SXB_API int sxb_rmsnorm_bwd_noalloc(const void* dy, const void* x, const void* w, const void* rstd,
                                    const void* dres, void* dx, void* dw, long long rows, int H, int dtype,
                                    int accumulate_dw, void* stream) {
  if (rows <= 0) return SXB_OK;
  if (!norm_ok(H)) return SXB_ERR_UNSUPPORTED;
  if (H > SXB_NOALLOC_NORM_MAX_H) return SXB_ERR_TOO_LARGE;   // scratch is sized for H <= 8192
  if (sxb_norm_bwd_scratch_elems(rows, H, 0) > kNoAllocScratchElems) return SXB_ERR_TOO_LARGE;
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= kMaxDevices) return SXB_ERR_UNSUPPORTED;
  float* scratch = g_noalloc_ptr[dev];
  if (scratch == nullptr) {
    // In host code `g_noalloc_norm_scratch` is only a handle; the device address has to be asked for,
    // once per device, with that device current.
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_noalloc_norm_scratch);
    if (e != cudaSuccess) return (int)e;
    if (p == nullptr) return SXB_ERR_DRIVER;
    scratch = static_cast<float*>(p);
    g_noalloc_ptr[dev] = scratch;
  }
  return norm_bwd<false>(dy, x, w, nullptr, rstd, dres, dx, dw, nullptr, scratch, rows, H, dtype,
                         accumulate_dw, (cudaStream_t)stream);
}

SXB_API int sxb_rmsnorm_bwd_noalloc_max_h() { return SXB_NOALLOC_NORM_MAX_H; }
