// softmax_dropout.cu — fused bandwidth-bound pieces of the training transformer layer, the Evoformer attention and
// the block-sparse attention (sm_100a; one HBM round trip per tensor, 16-byte vectors where rows allow it).
//
//   sxb_bias_dropout_residual   out = residual + dropout(x + bias)            (+ 1-byte keep mask)
//   sxb_dropout_bwd             dx  = dy * keep / (1 - p)
//   sxb_biased_softmax_fwd/bwd  softmax(scale * s + bias1 + bias2) over the last dim of [B, H, Sq, Sk], both biases
//                               with arbitrary broadcast strides, optional dropout of the probabilities
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

// One CTA per score row.  Pass 1: online max/sum, pass 2: normalise (+dropout) and store.
template <typename T>
__global__ void __launch_bounds__(256)
biased_softmax_fwd_kernel(const T* __restrict__ s, T* __restrict__ out, uint8_t* __restrict__ keep, BiasDesc b1,
                          BiasDesc b2, int H, int Sq, int Sk, float scale, float p, unsigned long long seed,
                          unsigned long long offset) {
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
  const Philox rng(seed);
  const float inv_keep = p < 1.f ? 1.f / (1.f - p) : 0.f;
  T* orow = out + row * Sk;
  for (int k = threadIdx.x; k < Sk; k += blockDim.x) {
    const float v = val(k);
    float pr = (v > -INFINITY) ? __expf(v - gmx) * inv_sum : 0.f;
    if (p > 0.f) {
      const unsigned long long idx = (unsigned long long)row * Sk + k;
      const uint4 u = rng(idx >> 2, offset);
      const uint32_t rr = (idx & 3) == 0 ? u.x : (idx & 3) == 1 ? u.y : (idx & 3) == 2 ? u.z : u.w;
      const bool kp = u01(rr) >= p;
      if (keep) keep[idx] = kp ? 1 : 0;
      pr = kp ? pr * inv_keep : 0.f;
    }
    orow[k] = from_f32<T>(pr);
  }
}

// dS = scale * P * (dP - sum_k dP*P); with dropout, ``y`` holds the dropped probabilities and
// dP_eff = dy * keep / (1-p), P = y * (1-p) on kept entries (dropped entries carry no gradient).  Optional
// dbias (same shape as the scores, un-scaled) for the Evoformer pair bias.
template <typename T>
__global__ void __launch_bounds__(256)
biased_softmax_bwd_kernel(const T* __restrict__ y, const T* __restrict__ dy, const uint8_t* __restrict__ keep,
                          T* __restrict__ ds, T* __restrict__ dbias, int Sk, float scale, float p) {
  __shared__ float red[32];
  const long long row = blockIdx.x;
  const T* yr = y + row * Sk;
  const T* dr = dy + row * Sk;
  const uint8_t* kr = keep ? keep + row * Sk : nullptr;
  const float keep_scale = 1.f - p;
  float dot = 0.f;
  for (int k = threadIdx.x; k < Sk; k += blockDim.x) {
    // P*dP_eff = (y*(1-p)) * (dy/(1-p)) = y*dy on kept entries; dropped entries have y == 0
    dot += to_f32<T>(yr[k]) * to_f32<T>(dr[k]);
  }
  dot = block_sum(dot, red);
  for (int k = threadIdx.x; k < Sk; k += blockDim.x) {
    float pr = to_f32<T>(yr[k]), g = to_f32<T>(dr[k]);
    float v;
    if (kr) {
      // dropped entries: P unknown from y (it is 0 there) — recover nothing, their dP_eff is 0 so
      // dS = -P * dot; we need P: callers keep the un-dropped probabilities when p > 0 (see Python).
      const bool kp = kr[k] != 0;
      const float P = pr;                     // un-dropped probability (Python passes it when p > 0)
      const float dP = kp ? g / keep_scale : 0.f;
      v = P * (dP - dot);
    } else {
      v = pr * (g - dot);
    }
    if (dbias) dbias[row * Sk + k] = from_f32<T>(v);
    ds[row * Sk + k] = from_f32<T>(v * scale);
  }
}

// with dropout the dot product must use the un-dropped P and the effective dP
template <typename T>
__global__ void __launch_bounds__(256)
biased_softmax_dropout_bwd_kernel(const T* __restrict__ prob, const T* __restrict__ dy,
                                  const uint8_t* __restrict__ keep, T* __restrict__ ds, T* __restrict__ dbias,
                                  int Sk, float scale, float p) {
  __shared__ float red[32];
  const long long row = blockIdx.x;
  const T* pr = prob + row * Sk;
  const T* dr = dy + row * Sk;
  const uint8_t* kr = keep + row * Sk;
  const float inv = 1.f / (1.f - p);
  float dot = 0.f;
  for (int k = threadIdx.x; k < Sk; k += blockDim.x)
    if (kr[k]) dot += to_f32<T>(pr[k]) * to_f32<T>(dr[k]) * inv;
  dot = block_sum(dot, red);
  for (int k = threadIdx.x; k < Sk; k += blockDim.x) {
    const float P = to_f32<T>(pr[k]);
    const float dP = kr[k] ? to_f32<T>(dr[k]) * inv : 0.f;
    const float v = P * (dP - dot);
    if (dbias) dbias[row * Sk + k] = from_f32<T>(v);
    ds[row * Sk + k] = from_f32<T>(v * scale);
  }
}

}  // namespace sd
}  // namespace sxb
