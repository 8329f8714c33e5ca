// softmax_dropout.cu — fused bandwidth-bound pieces of the training transformer layer, the Evoformer attention and
// the block-sparse attention (sm_100a; one HBM round trip per tensor, 16-byte vectors where rows allow it).
//
//   sxb_bias_dropout_residual   out = residual + dropout(x + bias)            (+ 1-byte keep mask)
//   sxb_dropout_bwd             dx  = dy * keep / (1 - p)
//   sxb_biased_softmax_fwd/bwd  softmax(scale * s + bias1 + bias2) over the last dim of [B, H, Sq, Sk], both biases
//                               with arbitrary broadcast strides
//   sxb_bsr_softmax_fwd/bwd     softmax over the non-zero blocks of one block-row of a block-sparse score matrix,
//                               with relative-position bias, key-padding mask and attention mask (add or mul mode)
//
// Reference behaviour:
//   fused bias/dropout/residual + attention softmax of the BERT layer: deepspeed/ops/transformer/transformer.py:149-293
//     (native forward_fp16/backward_fp16 — absent from the reference tree, options :62-84)
//   evoformer biases:  deepspeed/ops/deepspeed4science/evoformer_attn.py:29-106
//   block-sparse softmax: deepspeed/ops/sparse_attention/softmax.py:123-296
#include "sxb_common.cuh"

namespace sxb {
namespace sd {

// ------------------------------------------------------------------ counter-based RNG (Philox-4x32, 10 rounds)
struct Philox {
  uint32_t k0, k1;
  __device__ __forceinline__ Philox(unsigned long long seed) : k0((uint32_t)seed), k1((uint32_t)(seed >> 32)) {}
  __device__ __forceinline__ uint4 operator()(unsigned long long ctr, unsigned long long stream) const {
    uint32_t c0 = (uint32_t)ctr, c1 = (uint32_t)(ctr >> 32), c2 = (uint32_t)stream, c3 = (uint32_t)(stream >> 32);
    uint32_t a = k0, b = k1;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
      const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
      const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
      c0 = hi1 ^ c1 ^ a;
      c1 = lo1;
      c2 = hi0 ^ c3 ^ b;
      c3 = lo0;
      a += 0x9E3779B9u;
      b += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
  }
};

__device__ __forceinline__ float u01(uint32_t r) { return (r >> 8) * (1.0f / 16777216.0f); }

// ------------------------------------------------------------------ bias + dropout + residual
// 4 elements per thread-iteration (one Philox call); cols % 4 == 0.
template <typename T>
__global__ void __launch_bounds__(256)
bias_dropout_residual_kernel(const T* __restrict__ x, const T* __restrict__ bias, const T* __restrict__ residual,
                             T* __restrict__ out, uint8_t* __restrict__ keep, long long n4, int cols, float p,
                             unsigned long long seed, unsigned long long offset) {
  const Philox rng(seed);
  const float inv = p < 1.f ? 1.f / (1.f - p) : 0.f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4;
       i += (long long)gridDim.x * blockDim.x) {
    const long long e = i * 4;
    const int c = (int)(e % cols);
    float v[4], r[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = to_f32<T>(x[e + j]) + (bias ? to_f32<T>(bias[c + j]) : 0.f);
    if (residual) {
#pragma unroll
      for (int j = 0; j < 4; ++j) r[j] = to_f32<T>(residual[e + j]);
    }
    uint32_t m = 0x01010101u;
    if (p > 0.f) {
      const uint4 u = rng((unsigned long long)i, offset);
      const uint32_t rr[4] = {u.x, u.y, u.z, u.w};
      m = 0;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const bool k = u01(rr[j]) >= p;
        v[j] = k ? v[j] * inv : 0.f;
        m |= (k ? 1u : 0u) << (8 * j);
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) out[e + j] = from_f32<T>(v[j] + r[j]);
    if (keep) *reinterpret_cast<uint32_t*>(keep + e) = m;
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
dropout_bwd_kernel(const T* __restrict__ dy, const uint8_t* __restrict__ keep, T* __restrict__ dx, long long n,
                   float inv) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x)
    dx[i] = from_f32<T>(keep[i] ? to_f32<T>(dy[i]) * inv : 0.f);
}

// ------------------------------------------------------------------ biased softmax over [B, H, Sq, Sk]
struct BiasDesc {
  const void* ptr;
  long long sb, sh, sq, sk;  // element strides (0 = broadcast)
};

// One CTA per score row.  Pass 1: online max/sum, pass 2: normalise and store (dropout of the probabilities, when
// wanted, is the bias_dropout_residual kernel with no bias and no residual).
template <typename T>
__global__ void __launch_bounds__(256)
biased_softmax_fwd_kernel(const T* __restrict__ s, T* __restrict__ out, BiasDesc b1, BiasDesc b2, int H, int Sq,
                          int Sk, float scale) {
  __shared__ float red[64];
  const long long row = blockIdx.x;
  const int q = (int)(row % Sq);
  const int h = (int)((row / Sq) % H);
  const long long b = row / ((long long)Sq * H);
  const T* sr = s + row * Sk;
  const T* p1 = b1.ptr ? static_cast<const T*>(b1.ptr) + b * b1.sb + h * b1.sh + q * b1.sq : nullptr;
  const T* p2 = b2.ptr ? static_cast<const T*>(b2.ptr) + b * b2.sb + h * b2.sh + q * b2.sq : nullptr;
  auto val = [&](int k) -> float {
    float v = to_f32<T>(sr[k]) * scale;
    if (p1) v += to_f32<T>(p1[k * b1.sk]);
    if (p2) v += to_f32<T>(p2[k * b2.sk]);
    return v;
  };
  float mx = -INFINITY, sum = 0.f;
  for (int k = threadIdx.x; k < Sk; k += blockDim.x) {
    const float v = val(k);
    if (v > mx) {
      sum = sum * __expf(mx - v) + 1.f;
      mx = v;
    } else if (v > -INFINITY) {
      sum += __expf(v - mx);
    }
  }
  const float gmx = block_max(mx, red);
  __syncthreads();
  sum = (mx > -INFINITY) ? sum * __expf(mx - gmx) : 0.f;
  const float gsum = block_sum(sum, red + 32);
  const float inv_sum = gsum > 0.f ? 1.f / gsum : 0.f;
  T* orow = out + row * Sk;
  for (int k = threadIdx.x; k < Sk; k += blockDim.x) {
    const float v = val(k);
    const float pr = (v > -INFINITY) ? __expf(v - gmx) * inv_sum : 0.f;
    orow[k] = from_f32<T>(pr);
  }
}

// dS = scale * P * (dP - sum_k dP*P).  Optional dbias (same shape as the scores, un-scaled) for additive biases.
template <typename T>
__global__ void __launch_bounds__(256)
biased_softmax_bwd_kernel(const T* __restrict__ y, const T* __restrict__ dy, T* __restrict__ ds,
                          T* __restrict__ dbias, int Sk, float scale) {
  __shared__ float red[32];
  const long long row = blockIdx.x;
  const T* yr = y + row * Sk;
  const T* dr = dy + row * Sk;
  float dot = 0.f;
  for (int k = threadIdx.x; k < Sk; k += blockDim.x) dot += to_f32<T>(yr[k]) * to_f32<T>(dr[k]);
  dot = block_sum(dot, red);
  for (int k = threadIdx.x; k < Sk; k += blockDim.x) {
    const float v = to_f32<T>(yr[k]) * (to_f32<T>(dr[k]) - dot);
    if (dbias) dbias[row * Sk + k] = from_f32<T>(v);
    ds[row * Sk + k] = from_f32<T>(v * scale);
  }
}

// ------------------------------------------------------------------ block-sparse (BSR) softmax
// x: [B, nnz, blk, blk] — the non-zero blocks of a [H, S/blk, S/blk] layout in (head, block-row, block-col) order, so
// the CSR position of a block is its index in x.  One warp per (batch, head, block-row, row-in-block).
struct BsrArgs {
  const int* row_ptr;   // [H * nrb + 1]
  const int* col_idx;   // [nnz] block column
  int H, nrb, ncb, blk, nnz;
  const void* rpe;      // [H, S, S] or null (rpe_sb = batch stride, 0 = shared)
  long long rpe_sb;
  const void* kpm;      // [B, S] or null
  const void* amask;    // [S, S] or null
  int kpm_mode, amask_mode;   // 1 = add, 2 = mul (0 -> -inf, else unchanged)
  float scale;
};

template <typename T>
__global__ void __launch_bounds__(256)
bsr_softmax_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, BsrArgs a, long long n_rows) {
  const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= n_rows) return;
  const int r = (int)(w % a.blk);
  const long long t = w / a.blk;
  const int hr = (int)(t % ((long long)a.H * a.nrb));   // head * nrb + block-row
  const long long b = t / ((long long)a.H * a.nrb);
  const int h = hr / a.nrb, rb = hr % a.nrb;
  const int j0 = a.row_ptr[hr], j1 = a.row_ptr[hr + 1];
  const int S = a.nrb * a.blk, Sk = a.ncb * a.blk;
  const int q = rb * a.blk + r;
  const int ncol = (j1 - j0) * a.blk;
  const T* rpe = a.rpe ? static_cast<const T*>(a.rpe) + b * a.rpe_sb + ((long long)h * S + q) * Sk : nullptr;
  const T* kpm = a.kpm ? static_cast<const T*>(a.kpm) + b * Sk : nullptr;
  const T* am = a.amask ? static_cast<const T*>(a.amask) + (long long)q * Sk : nullptr;
  auto addr = [&](int i) -> long long {
    const int j = j0 + i / a.blk, c = i % a.blk;
    return (((long long)b * a.nnz + j) * a.blk + r) * a.blk + c;
  };
  auto val = [&](int i) -> float {
    const int j = j0 + i / a.blk, c = i % a.blk;
    const int k = a.col_idx[j] * a.blk + c;
    float v = to_f32<T>(x[(((long long)b * a.nnz + j) * a.blk + r) * a.blk + c]) * a.scale;
    if (rpe) v += to_f32<T>(rpe[k]);
    if (kpm) {
      const float m = to_f32<T>(kpm[k]);
      v = a.kpm_mode == 2 ? (m == 0.f ? -INFINITY : v) : v + m;
    }
    if (am) {
      const float m = to_f32<T>(am[k]);
      v = a.amask_mode == 2 ? (m == 0.f ? -INFINITY : v) : v + m;
    }
    return v;
  };
  float mx = -INFINITY;
  for (int i = lane; i < ncol; i += 32) mx = fmaxf(mx, val(i));
  mx = warp_max(mx);
  float sum = 0.f;
  for (int i = lane; i < ncol; i += 32) {
    const float v = val(i);
    sum += (v > -INFINITY) ? __expf(v - mx) : 0.f;
  }
  sum = warp_sum(sum);
  const float inv = sum > 0.f ? 1.f / sum : 0.f;
  for (int i = lane; i < ncol; i += 32) {
    const float v = val(i);
    y[addr(i)] = from_f32<T>((v > -INFINITY) ? __expf(v - mx) * inv : 0.f);
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
bsr_softmax_bwd_kernel(const T* __restrict__ y, const T* __restrict__ dy, T* __restrict__ dx, BsrArgs a,
                       long long n_rows) {
  const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= n_rows) return;
  const int r = (int)(w % a.blk);
  const long long t = w / a.blk;
  const int hr = (int)(t % ((long long)a.H * a.nrb));
  const long long b = t / ((long long)a.H * a.nrb);
  const int j0 = a.row_ptr[hr], j1 = a.row_ptr[hr + 1];
  const int ncol = (j1 - j0) * a.blk;
  auto addr = [&](int i) -> long long {
    const int j = j0 + i / a.blk, c = i % a.blk;
    return (((long long)b * a.nnz + j) * a.blk + r) * a.blk + c;
  };
  float dot = 0.f;
  for (int i = lane; i < ncol; i += 32) {
    const long long o = addr(i);
    dot += to_f32<T>(y[o]) * to_f32<T>(dy[o]);
  }
  dot = warp_sum(dot);
  for (int i = lane; i < ncol; i += 32) {
    const long long o = addr(i);
    dx[o] = from_f32<T>(to_f32<T>(y[o]) * (to_f32<T>(dy[o]) - dot) * a.scale);
  }
}

}  // namespace sd
}  // namespace sxb

using namespace sxb;

#define SD_DISPATCH(dtype, ...)                                    \
  switch (dtype) {                                                 \
    case kBF16: { using T = __nv_bfloat16; __VA_ARGS__; break; }   \
    case kF16: { using T = __half; __VA_ARGS__; break; }           \
    case kF32: { using T = float; __VA_ARGS__; break; }            \
    default: return SXB_ERR_UNSUPPORTED;                           \
  }

static inline unsigned sd_blocks(long long n, int per_block) {
  long long b = (n + per_block - 1) / per_block;
  const long long cap = 148LL * 16;
  return (unsigned)(b < 1 ? 1 : (b > cap ? cap : b));
}

// out = residual + dropout(x + bias); ``keep`` (uint8 per element) may be null when p == 0.  n % 4 == 0, cols % 4 == 0.
SXB_API int sxb_bias_dropout_residual(const void* x, const void* bias, const void* residual, void* out, void* keep,
                                      long long n, int cols, float p, long long seed, long long offset, int dtype,
                                      void* stream) {
  if (n <= 0) return SXB_OK;
  if ((n & 3) || (cols & 3) || p < 0.f || p > 1.f) return SXB_ERR_BAD_ARG;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  SD_DISPATCH(dtype, sd::bias_dropout_residual_kernel<T><<<sd_blocks(n / 4, 256), 256, 0, s>>>(
                         static_cast<const T*>(x), static_cast<const T*>(bias), static_cast<const T*>(residual),
                         static_cast<T*>(out), static_cast<uint8_t*>(keep), n / 4, cols, p,
                         (unsigned long long)seed, (unsigned long long)offset));
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

SXB_API int sxb_dropout_bwd(const void* dy, const void* keep, void* dx, long long n, float p, int dtype,
                            void* stream) {
  if (n <= 0) return SXB_OK;
  if (p < 0.f || p >= 1.f) return SXB_ERR_BAD_ARG;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  SD_DISPATCH(dtype, sd::dropout_bwd_kernel<T><<<sd_blocks(n, 256 * 4), 256, 0, s>>>(
                         static_cast<const T*>(dy), static_cast<const uint8_t*>(keep), static_cast<T*>(dx), n,
                         1.f / (1.f - p)));
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

// scores/out [B, H, Sq, Sk] contiguous; bias strides in elements (b, h, q, k), 0 = broadcast.
SXB_API int sxb_biased_softmax_fwd(const void* scores, void* out, const void* bias1, long long b1b, long long b1h,
                                   long long b1q, long long b1k, const void* bias2, long long b2b, long long b2h,
                                   long long b2q, long long b2k, long long B, int H, int Sq, int Sk, float scale,
                                   int dtype, void* stream) {
  const long long rows = B * H * Sq;
  if (rows <= 0 || Sk <= 0) return SXB_OK;
  if (rows > 0x7fffffffLL) return SXB_ERR_TOO_LARGE;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  sd::BiasDesc d1{bias1, b1b, b1h, b1q, b1k}, d2{bias2, b2b, b2h, b2q, b2k};
  SD_DISPATCH(dtype, sd::biased_softmax_fwd_kernel<T><<<(unsigned)rows, 256, 0, s>>>(
                         static_cast<const T*>(scores), static_cast<T*>(out), d1, d2, H, Sq, Sk, scale));
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

SXB_API int sxb_biased_softmax_bwd(const void* y, const void* dy, void* ds, void* dbias, long long rows, int Sk,
                                   float scale, int dtype, void* stream) {
  if (rows <= 0 || Sk <= 0) return SXB_OK;
  if (rows > 0x7fffffffLL) return SXB_ERR_TOO_LARGE;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  SD_DISPATCH(dtype, sd::biased_softmax_bwd_kernel<T><<<(unsigned)rows, 256, 0, s>>>(
                         static_cast<const T*>(y), static_cast<const T*>(dy), static_cast<T*>(ds),
                         static_cast<T*>(dbias), Sk, scale));
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

SXB_API int sxb_bsr_softmax_fwd(const void* x, void* y, const void* row_ptr, const void* col_idx, long long B, int H,
                                int nrb, int ncb, int blk, int nnz, const void* rpe, long long rpe_sb, const void* kpm,
                                int kpm_mode, const void* amask, int amask_mode, float scale, int dtype,
                                void* stream) {
  const long long n_rows = B * H * nrb * blk;
  if (n_rows <= 0 || nnz <= 0) return SXB_OK;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  sd::BsrArgs a{static_cast<const int*>(row_ptr), static_cast<const int*>(col_idx), H, nrb, ncb, blk, nnz, rpe, rpe_sb,
                kpm, amask, kpm_mode, amask_mode, scale};
  const unsigned blocks = (unsigned)((n_rows + 7) / 8);
  SD_DISPATCH(dtype, sd::bsr_softmax_fwd_kernel<T><<<blocks, 256, 0, s>>>(static_cast<const T*>(x),
                                                                          static_cast<T*>(y), a, n_rows));
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

SXB_API int sxb_bsr_softmax_bwd(const void* y, const void* dy, void* dx, const void* row_ptr, const void* col_idx,
                                long long B, int H, int nrb, int ncb, int blk, int nnz, float scale, int dtype,
                                void* stream) {
  const long long n_rows = B * H * nrb * blk;
  if (n_rows <= 0 || nnz <= 0) return SXB_OK;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  sd::BsrArgs a{static_cast<const int*>(row_ptr), static_cast<const int*>(col_idx), H, nrb, ncb, blk, nnz, nullptr, 0,
                nullptr, nullptr, 0, 0, scale};
  const unsigned blocks = (unsigned)((n_rows + 7) / 8);
  SD_DISPATCH(dtype, sd::bsr_softmax_bwd_kernel<T><<<blocks, 256, 0, s>>>(
                         static_cast<const T*>(y), static_cast<const T*>(dy), static_cast<T*>(dx), a, n_rows));
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}
