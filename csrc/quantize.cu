// quantize.cu — grouped integer and floating-point quantisers, sign packing, random-LTD token movers.
//
// Fresh implementations of the native ops the reference loads through op_builder but does not ship
// (SURVEY §2.3): QuantizerBuilder (N8: ds_quantize / quantize / dequantize / swizzle_quant /
// quantized_reduction; call sites deepspeed/ops/quantizer/quantizer.py:15-29,
// runtime/zero/partition_parameters.py:832-862, runtime/comm/coalesced_collectives.py:30-79),
// FPQuantizerBuilder (N9: ops/fp_quantizer/quantize.py:48-165), PackbitsBuilder (N17:
// runtime/comm/compressed.py:46-54) and RandomLTDBuilder (N11: ops/random_ltd/dropping_utils.py:32-114).
//
// All kernels are bandwidth-bound: one CTA per quantisation group, 16-byte global accesses, warp-shuffle
// reductions, no shared-memory staging of the payload (each thread keeps its slice in registers between the
// range pass and the encode pass, so every input byte is read from HBM exactly once).
#include "sxb_common.cuh"

namespace sxb {
namespace quant {

constexpr int kThreads = 256;
// A thread keeps PASSES x 8 elements in registers (PASSES in {1, 2, 4}: group_size <= 2048 / 4096 / 8192).
// The slice loop is fully unrolled so the arrays are never indexed dynamically (no local-memory spill).
#define SXB_FOR_SLICES(p, base)                                    \
  _Pragma("unroll") for (int p = 0; p < PASSES; ++p)               \
    for (int base = (p * kThreads + threadIdx.x) * 8, once_ = 1; once_ && base < group_size; once_ = 0)

__device__ __forceinline__ uint32_t hash_u32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return x;
}
__device__ __forceinline__ float uniform01(uint32_t seed, uint32_t idx) {
  return (hash_u32(seed ^ (idx * 0x9E3779B9u)) >> 8) * (1.0f / 16777216.0f);
}

template <typename T>
__device__ __forceinline__ void load8(const T* p, float (&v)[8]);
template <>
__device__ __forceinline__ void load8<__nv_bfloat16>(const __nv_bfloat16* p, float (&v)[8]) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) { v[2 * i] = __low2float(h[i]); v[2 * i + 1] = __high2float(h[i]); }
}
template <>
__device__ __forceinline__ void load8<__half>(const __half* p, float (&v)[8]) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  const __half2* h = reinterpret_cast<const __half2*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) { v[2 * i] = __low2float(h[i]); v[2 * i + 1] = __high2float(h[i]); }
}
template <>
__device__ __forceinline__ void load8<float>(const float* p, float (&v)[8]) {
  const float4 a = *reinterpret_cast<const float4*>(p);
  const float4 b = *reinterpret_cast<const float4*>(p + 4);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
template <typename T>
__device__ __forceinline__ void store8(T* p, const float (&v)[8]);
template <>
__device__ __forceinline__ void store8<__nv_bfloat16>(__nv_bfloat16* p, const float (&v)[8]) {
  uint4 u;
  u.x = pack_bf16(v[0], v[1]); u.y = pack_bf16(v[2], v[3]); u.z = pack_bf16(v[4], v[5]); u.w = pack_bf16(v[6], v[7]);
  *reinterpret_cast<uint4*>(p) = u;
}
template <>
__device__ __forceinline__ void store8<__half>(__half* p, const float (&v)[8]) {
  __half2 h[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2half2_rn(v[2 * i], v[2 * i + 1]);
  *reinterpret_cast<uint4*>(p) = *reinterpret_cast<uint4*>(h);
}
template <>
__device__ __forceinline__ void store8<float>(float* p, const float (&v)[8]) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
}

__device__ __forceinline__ float block_min(float v, float* sm) { return -block_max(-v, sm); }

// ------------------------------------------------------------------------------- integer
// Group g of the input is written as group `perm(g)` of the output: perm is the identity unless
// (nodes, devs) describe the hierarchical all-to-all of qgZ, where the partition-major order
// (node, dev, r) becomes (dev, node, r) so that each intra-node a2a chunk is contiguous.
__device__ __forceinline__ long long swizzle_group(long long g, long long groups, int nodes, int devs) {
  if (nodes <= 1 || devs <= 1) return g;
  const long long per = groups / ((long long)nodes * devs);
  const long long part = g / per, r = g - part * per;
  const long long node = part / devs, dev = part - node * devs;
  return (dev * nodes + node) * per + r;
}

template <typename T, int BITS, bool SYM, int PASSES>
__global__ void __launch_bounds__(kThreads)
quantize_int_kernel(const T* __restrict__ in, uint8_t* __restrict__ q, float* __restrict__ scales,
                    long long groups, int group_size, int stochastic, uint32_t seed, int nodes, int devs) {
  __shared__ float red[32];
  const long long g = blockIdx.x;
  const T* src = in + g * group_size;
  const long long go = swizzle_group(g, groups, nodes, devs);
  float v[PASSES * 8];
  float lo = INFINITY, hi = -INFINITY, am = 0.f;   // threads without a slice contribute the identities
  SXB_FOR_SLICES(ps, base) {
    float t[8];
    load8<T>(src + base, t);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      v[ps * 8 + i] = t[i];
      lo = fminf(lo, t[i]);
      hi = fmaxf(hi, t[i]);
      am = fmaxf(am, fabsf(t[i]));
    }
  }
  constexpr int QMAX = (1 << (BITS - 1)) - 1;  // 127 / 7
  float scale, offset = 0.f;
  if (SYM) {
    const float amax = block_max(am, red);
    scale = amax > 0.f ? amax / QMAX : 1.f;
    if (threadIdx.x == 0) scales[go] = scale;
  } else {
    const float mx = block_max(hi, red);
    __syncthreads();
    const float mn = block_min(lo, red);
    scale = (mx > mn) ? (mx - mn) / float((1 << BITS) - 1) : 1.f;
    offset = mn;
    if (threadIdx.x == 0) { scales[2 * go] = scale; scales[2 * go + 1] = offset; }
  }
  const float inv = 1.f / scale;
  uint8_t* dst = q + go * ((long long)group_size * BITS / 8);
  SXB_FOR_SLICES(ps, base) {
    int qi[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float x = (v[ps * 8 + i] - offset) * inv;
      float r;
      if (stochastic) r = floorf(x + uniform01(seed, (uint32_t)(g * group_size + base + i)));
      else r = rintf(x);
      if (SYM) r = fminf(fmaxf(r, -float(QMAX) - 1.f), float(QMAX));
      else r = fminf(fmaxf(r, 0.f), float((1 << BITS) - 1));
      qi[i] = (int)r;
    }
    if (BITS == 8) {
      uint2 o;
      o.x = (qi[0] & 0xFF) | ((qi[1] & 0xFF) << 8) | ((qi[2] & 0xFF) << 16) | ((uint32_t)(qi[3] & 0xFF) << 24);
      o.y = (qi[4] & 0xFF) | ((qi[5] & 0xFF) << 8) | ((qi[6] & 0xFF) << 16) | ((uint32_t)(qi[7] & 0xFF) << 24);
      *reinterpret_cast<uint2*>(dst + base) = o;
    } else {
      uint32_t o = 0;
#pragma unroll
      for (int i = 0; i < 8; ++i) o |= (uint32_t)(qi[i] & 0xF) << (4 * i);
      *reinterpret_cast<uint32_t*>(dst + base / 2) = o;
    }
  }
}

template <int BITS, bool SYM>
__device__ __forceinline__ void decode8(const uint8_t* src, int base, float scale, float offset, float (&t)[8]) {
  if (BITS == 8) {
    const uint2 u = *reinterpret_cast<const uint2*>(src + base);
    const uint32_t w[2] = {u.x, u.y};
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const uint32_t b = (w[i >> 2] >> (8 * (i & 3))) & 0xFF;
      t[i] = SYM ? float((int8_t)b) * scale : float(b) * scale + offset;
    }
  } else {
    const uint32_t u = *reinterpret_cast<const uint32_t*>(src + base / 2);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const uint32_t b = (u >> (4 * i)) & 0xF;
      t[i] = SYM ? float(((int)(b << 28)) >> 28) * scale : float(b) * scale + offset;
    }
  }
}

template <typename T, int BITS, bool SYM>
__global__ void __launch_bounds__(kThreads)
dequantize_int_kernel(const uint8_t* __restrict__ q, const float* __restrict__ scales, T* __restrict__ out,
                      long long groups, int group_size) {
  const long long g = blockIdx.x;
  const float scale = SYM ? scales[g] : scales[2 * g];
  const float offset = SYM ? 0.f : scales[2 * g + 1];
  const uint8_t* src = q + g * ((long long)group_size * BITS / 8);
  T* dst = out + g * group_size;
  for (int base = threadIdx.x * 8; base < group_size; base += kThreads * 8) {
    float t[8];
    decode8<BITS, SYM>(src, base, scale, offset, t);
    store8<T>(dst + base, t);
  }
}

// out group g = requant( sum_d dequant(in[d, g]) ): the middle step of qgZ (dequant-reduce-requant).
template <int BITS, bool SYM, int PASSES>
__global__ void __launch_bounds__(kThreads)
quantized_reduction_kernel(const uint8_t* __restrict__ q_in, const float* __restrict__ s_in,
                           uint8_t* __restrict__ q_out, float* __restrict__ s_out, long long out_groups,
                           int group_size, int devs) {
  __shared__ float red[32];
  const long long g = blockIdx.x;
  const long long gbytes = (long long)group_size * BITS / 8;
  float acc[PASSES * 8];
#pragma unroll
  for (int i = 0; i < PASSES * 8; ++i) acc[i] = 0.f;
  for (int d = 0; d < devs; ++d) {
    const long long gi = (long long)d * out_groups + g;
    const float scale = SYM ? s_in[gi] : s_in[2 * gi];
    const float offset = SYM ? 0.f : s_in[2 * gi + 1];
    SXB_FOR_SLICES(ps, base) {
      float t[8];
      decode8<BITS, SYM>(q_in + gi * gbytes, base, scale, offset, t);
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[ps * 8 + i] += t[i];
    }
  }
  float lo = INFINITY, hi = -INFINITY, am = 0.f;
  SXB_FOR_SLICES(ps, base) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      lo = fminf(lo, acc[ps * 8 + i]);
      hi = fmaxf(hi, acc[ps * 8 + i]);
      am = fmaxf(am, fabsf(acc[ps * 8 + i]));
    }
  }
  constexpr int QMAX = (1 << (BITS - 1)) - 1;
  float scale, offset = 0.f;
  if (SYM) {
    const float amax = block_max(am, red);
    scale = amax > 0.f ? amax / QMAX : 1.f;
    if (threadIdx.x == 0) s_out[g] = scale;
  } else {
    const float mx = block_max(hi, red);
    __syncthreads();
    const float mn = block_min(lo, red);
    scale = (mx > mn) ? (mx - mn) / float((1 << BITS) - 1) : 1.f;
    offset = mn;
    if (threadIdx.x == 0) { s_out[2 * g] = scale; s_out[2 * g + 1] = offset; }
  }
  const float inv = 1.f / scale;
  uint8_t* dst = q_out + g * gbytes;
  SXB_FOR_SLICES(ps, base) {
    int qi[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float r = rintf((acc[ps * 8 + i] - offset) * inv);
      if (SYM) r = fminf(fmaxf(r, -float(QMAX) - 1.f), float(QMAX));
      else r = fminf(fmaxf(r, 0.f), float((1 << BITS) - 1));
      qi[i] = (int)r;
    }
    if (BITS == 8) {
      uint2 o;
      o.x = (qi[0] & 0xFF) | ((qi[1] & 0xFF) << 8) | ((qi[2] & 0xFF) << 16) | ((uint32_t)(qi[3] & 0xFF) << 24);
      o.y = (qi[4] & 0xFF) | ((qi[5] & 0xFF) << 8) | ((qi[6] & 0xFF) << 16) | ((uint32_t)(qi[7] & 0xFF) << 24);
      *reinterpret_cast<uint2*>(dst + base) = o;
    } else {
      uint32_t o = 0;
#pragma unroll
      for (int i = 0; i < 8; ++i) o |= (uint32_t)(qi[i] & 0xF) << (4 * i);
      *reinterpret_cast<uint32_t*>(dst + base / 2) = o;
    }
  }
}

}  // namespace quant
}  // namespace sxb

// ======================================================================= floating-point formats
// Group-scaled miniature floats: q_bits in {4, 6, 8, 12} split as 1 sign + E exponent + M mantissa bits
// (e2m1, e3m2, e4m3, e7m4), IEEE-like with subnormals and NO inf/nan codes, so the largest magnitude is
// (2 - 2^-M) * 2^(2^E - 1 - bias): 6, 28, 480, ... (reference formats: ops/fp_quantizer/quantize.py:71-78;
// default FP8 with group 512: linear/config.py:48-53).  Each group stores group_size*q_bits/8 payload bytes
// followed by its fp32 scale (4 bytes), the layout `quantize()` of the reference allocates (:84).
namespace sxb {
namespace quant {

template <int E, int M>
struct MiniFloat {
  static constexpr int kBits = 1 + E + M;
  static constexpr int kBias = (1 << (E - 1)) - 1;
  static constexpr int kMaxExpField = (1 << E) - 1;
  __device__ static float max_value() {
    return (2.0f - exp2f(-float(M))) * exp2f(float(kMaxExpField - kBias));
  }
  // round-to-nearest-even (or stochastic) encode of |x| <= max_value()
  __device__ static uint32_t encode(float x, bool stochastic, float u) {
    const uint32_t sign = (__float_as_uint(x) >> 31) & 1u;
    float a = fabsf(x);
    a = fminf(a, max_value());
    // smallest normal = 2^(1-bias); quantum of the subnormal range = 2^(1-bias-M)
    int e;
    float m = frexpf(a, &e);        // a = m * 2^e, m in [0.5, 1)
    int ef = e - 1 + kBias;         // biased exponent of the normalised value 1.xxx * 2^(e-1)
    float q;                        // value / quantum as integer grid coordinate
    float quantum;
    if (a == 0.f) return sign << (E + M);
    if (ef <= 0) { quantum = exp2f(float(1 - kBias - M)); ef = 0; }
    else quantum = exp2f(float(e - 1 - M));
    q = a / quantum;
    float r = stochastic ? floorf(q + u) : rintf(q);
    // r in [0, 2^(M+1)]; a carry out of the mantissa bumps the exponent
    uint32_t code;
    if (ef == 0) {
      code = (uint32_t)r;           // subnormal (r == 2^M lands exactly on the first normal code)
    } else {
      uint32_t ri = (uint32_t)r;    // in [2^M, 2^(M+1)]
      if (ri >= (2u << M)) { ri >>= 1; ef += 1; }
      if (ef > kMaxExpField) { ef = kMaxExpField; ri = (2u << M) - 1; }
      code = ((uint32_t)ef << M) | (ri & ((1u << M) - 1));
    }
    return (sign << (E + M)) | code;
  }
  __device__ static float decode(uint32_t c) {
    const uint32_t sign = (c >> (E + M)) & 1u;
    const uint32_t ef = (c >> M) & ((1u << E) - 1);
    const uint32_t mant = c & ((1u << M) - 1);
    float v;
    if (ef == 0) v = float(mant) * exp2f(float(1 - kBias - M));
    else v = (1.0f + float(mant) * exp2f(-float(M))) * exp2f(float(int(ef) - kBias));
    return sign ? -v : v;
  }
};

// 8 codes of kBits bits each -> kBits bytes, little-endian bit stream.
template <int BITS>
__device__ __forceinline__ void pack8(const uint32_t (&c)[8], uint8_t* dst) {
  unsigned long long lo = 0, hi = 0;  // up to 96 bits
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int pos = i * BITS;
    if (pos < 64) {
      lo |= (unsigned long long)c[i] << pos;
      if (pos + BITS > 64) hi |= (unsigned long long)c[i] >> (64 - pos);
    } else {
      hi |= (unsigned long long)c[i] << (pos - 64);
    }
  }
#pragma unroll
  for (int b = 0; b < BITS; ++b) dst[b] = (uint8_t)((b < 8 ? (lo >> (8 * b)) : (hi >> (8 * (b - 8)))) & 0xFF);
}
template <int BITS>
__device__ __forceinline__ void unpack8(const uint8_t* src, uint32_t (&c)[8]) {
  unsigned long long lo = 0, hi = 0;
#pragma unroll
  for (int b = 0; b < BITS; ++b) {
    if (b < 8) lo |= (unsigned long long)src[b] << (8 * b);
    else hi |= (unsigned long long)src[b] << (8 * (b - 8));
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int pos = i * BITS;
    unsigned long long v;
    if (pos < 64) {
      v = lo >> pos;
      if (pos + BITS > 64) v |= hi << (64 - pos);
    } else {
      v = hi >> (pos - 64);
    }
    c[i] = (uint32_t)(v & ((1ull << BITS) - 1));
  }
}

template <typename T, int E, int M, int PASSES>
__global__ void __launch_bounds__(kThreads)
fp_quantize_kernel(const T* __restrict__ in, uint8_t* __restrict__ out, long long groups, int group_size,
                   int stochastic, uint32_t seed) {
  using F = MiniFloat<E, M>;
  __shared__ float red[32];
  const long long g = blockIdx.x;
  const T* src = in + g * group_size;
  const long long gbytes = (long long)group_size * F::kBits / 8 + 4;
  uint8_t* dst = out + g * gbytes;
  float v[PASSES * 8];
  float amax = 0.f;
  SXB_FOR_SLICES(ps, base) {
    float t[8];
    load8<T>(src + base, t);
#pragma unroll
    for (int i = 0; i < 8; ++i) { v[ps * 8 + i] = t[i]; amax = fmaxf(amax, fabsf(t[i])); }
  }
  amax = block_max(amax, red);
  const float scale = amax > 0.f ? amax / F::max_value() : 1.f;
  const float inv = 1.f / scale;
  if (threadIdx.x == 0) {
    const uint32_t sb = __float_as_uint(scale);
    uint8_t* sp = dst + gbytes - 4;
    sp[0] = sb & 0xFF; sp[1] = (sb >> 8) & 0xFF; sp[2] = (sb >> 16) & 0xFF; sp[3] = (sb >> 24) & 0xFF;
  }
  SXB_FOR_SLICES(ps, base) {
    uint32_t c[8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
      c[i] = F::encode(v[ps * 8 + i] * inv, stochastic != 0,
                       stochastic ? uniform01(seed, (uint32_t)(g * group_size + base + i)) : 0.f);
    pack8<F::kBits>(c, dst + (base / 8) * F::kBits);
  }
}

// rows == nullptr: dequantise every group.  Otherwise out row r takes input row rows[r] (a "row" is
// groups_per_row consecutive groups): selective_dequantize of the reference (:134-166).
template <typename T, int E, int M>
__global__ void __launch_bounds__(kThreads)
fp_dequantize_kernel(const uint8_t* __restrict__ in, T* __restrict__ out, int group_size,
                     const int* __restrict__ rows, long long groups_per_row) {
  using F = MiniFloat<E, M>;
  const long long go = blockIdx.x;
  long long gi = go;
  if (rows) {
    const long long r = go / groups_per_row;
    gi = (long long)rows[r] * groups_per_row + (go - r * groups_per_row);
  }
  const long long gbytes = (long long)group_size * F::kBits / 8 + 4;
  const uint8_t* src = in + gi * gbytes;
  const uint8_t* sp = src + gbytes - 4;
  const float scale = __uint_as_float((uint32_t)sp[0] | ((uint32_t)sp[1] << 8) | ((uint32_t)sp[2] << 16) |
                                      ((uint32_t)sp[3] << 24));
  T* dst = out + go * group_size;
  for (int base = threadIdx.x * 8; base < group_size; base += kThreads * 8) {
    uint32_t c[8];
    unpack8<F::kBits>(src + (base / 8) * F::kBits, c);
    float t[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) t[i] = F::decode(c[i]) * scale;
    store8<T>(dst + base, t);
  }
}

// ============================================================================ weight-only mini-float GEMV
// y[M, N] = x[M, K] W^T with W [N, K] stored group-quantised (FP6 e3m2 / FP8 e4m3 / FP4 e2m1 payload + fp32 scale per
// `group_size` consecutive elements, K % group_size == 0), for decode-sized M (<= 8): the product is bound by reading
// W once, so the weights stay packed until they are in registers.  One warp owns R = 4 consecutive rows of W and walks
// K 256 elements at a time (lane l: elements 8 l .. 8 l + 7 of the step): 8 codes = kBits bytes per row and lane, the
// matching 8 activations of every batch row are loaded once and used for all R rows.  Decode is a bit trick, not the
// exp2f of the offline dequantiser: the code's exponent|mantissa bits are moved to fp32 position, the exponent rebased
// by an integer add, subnormals (exponent field 0) scaled from the integer mantissa.
// (reference: inference/v2/kernels/core_ops/cuda_linear/fp6_linear.cu — W6A16 on mma.sync; here the
// bandwidth-bound decode case, prefill sizes dequantise into scratch and use the tcgen05 GEMM.)
// Sign to bit 31, exponent|mantissa to fp32 position; multiplying by 2^(127 - bias) rebases the exponent and turns
// the fp32 denormal that a zero exponent field produces into the format's subnormal value (no flush-to-zero: the
// library is built without fast-math).
template <int E, int M>
__device__ __forceinline__ float minifloat_bits(uint32_t c) {
  const uint32_t u = ((c << (31 - (E + M))) & 0x80000000u) | ((c & ((1u << (E + M)) - 1u)) << (23 - M));
  return __uint_as_float(u);
}

// 16 codes of BITS bits from 2*BITS bytes held in BITS/2 words (little-endian bit stream, 8 codes per BITS bytes)
template <int BITS>
__device__ __forceinline__ void unpack16(const uint32_t* __restrict__ wsrc, uint32_t (&c)[16]) {
  constexpr int NW = BITS / 2;                   // words per 16 codes
  uint32_t w[NW];
#pragma unroll
  for (int i = 0; i < NW; ++i) w[i] = wsrc[i];
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const int pos = i * BITS;                    // bit position inside the 2*BITS bytes
    const int wi = pos >> 5, sh = pos & 31;
    uint32_t v = w[wi] >> sh;
    if (sh + BITS > 32 && wi + 1 < NW) v |= w[wi + 1] << (32 - sh);
    c[i] = v & ((1u << BITS) - 1u);
  }
}

template <int E, int M, int MB>
__global__ void __launch_bounds__(256)
fp_gemv_kernel(const __nv_bfloat16* __restrict__ x, const uint8_t* __restrict__ wq, __nv_bfloat16* __restrict__ y,
               int N, int K, int group_size, long long x_stride, long long y_stride) {
  using F = MiniFloat<E, M>;
  constexpr int R = 4;
  constexpr int kBias = (1 << (E - 1)) - 1;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = (blockIdx.x * 8 + warp) * R;
  if (n0 >= N) return;
  const long long gbytes = (long long)group_size * F::kBits / 8 + 4;   // multiple of 4: word loads stay aligned
  const float rebase = __uint_as_float((uint32_t)(127 + 127 - kBias) << 23);
  float acc[R][MB];
#pragma unroll
  for (int r = 0; r < R; ++r)
#pragma unroll
    for (int m = 0; m < MB; ++m) acc[r][m] = 0.f;
  for (int k = lane * 16; k < K; k += 512) {
    float xv[MB][16];
#pragma unroll
    for (int m = 0; m < MB; ++m) {
      float lo[8], hi[8];
      load8<__nv_bfloat16>(x + m * x_stride + k, lo);
      load8<__nv_bfloat16>(x + m * x_stride + k + 8, hi);
#pragma unroll
      for (int i = 0; i < 8; ++i) { xv[m][i] = lo[i]; xv[m][8 + i] = hi[i]; }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
      if (n0 + r >= N) break;
      const long long e0 = (long long)(n0 + r) * K + k;            // element index in the flattened weight
      const long long g = e0 / group_size;
      const int o = (int)(e0 - g * group_size);                     // multiple of 16
      const uint8_t* gp = wq + g * gbytes;
      const float scale = __uint_as_float(*reinterpret_cast<const uint32_t*>(gp + gbytes - 4));
      uint32_t c[16];
      unpack16<F::kBits>(reinterpret_cast<const uint32_t*>(gp + (o / 8) * F::kBits), c);
      float wv[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) wv[i] = (minifloat_bits<E, M>(c[i]) * rebase) * scale;   // rebase first: no overflow
#pragma unroll
      for (int m = 0; m < MB; ++m)
#pragma unroll
        for (int i = 0; i < 16; ++i) acc[r][m] = fmaf(wv[i], xv[m][i], acc[r][m]);
    }
  }
#pragma unroll
  for (int r = 0; r < R; ++r)
#pragma unroll
    for (int m = 0; m < MB; ++m) {
      float v = acc[r][m];
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
      if (lane == 0 && n0 + r < N) y[m * y_stride + n0 + r] = __float2bfloat16_rn(v);
    }
}

// ============================================================================ weight-only integer GEMV
// y[M, N] = x[M, K] W^T for a weight kept as group-wise ASYMMETRIC integer codes in the layout of
// inference/quantization (reference deepspeed/inference/quantization/utils.py): codes uint8 [N, K] (8 bits) or
// [N, K/2] (4 bits, even index in the HIGH nibble), groups of `group_size` along K, scale / min bf16 [N, K/group_size]
// with w = code / scale + min.  Same decomposition as fp_gemv_kernel (one warp: 4 rows of W, 256 columns per step);
// per row and step the code sum-product and the activation sum are accumulated separately inside a group so the
// affine map costs two FMAs per 8 codes instead of one per code:  sum (c/s + m) x = (1/s) sum c x + m sum x.
template <int BITS, int MB>
__global__ void __launch_bounds__(256)
int_gemv_kernel(const __nv_bfloat16* __restrict__ x, const uint8_t* __restrict__ codes,
                const __nv_bfloat16* __restrict__ scale, const __nv_bfloat16* __restrict__ minv,
                __nv_bfloat16* __restrict__ y, int N, int K, int group_size, long long x_stride, long long y_stride) {
  constexpr int R = 4;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = (blockIdx.x * 8 + warp) * R;
  if (n0 >= N) return;
  const int gpr = K / group_size;                       // groups per row
  const long long row_bytes = (long long)K * BITS / 8;
  float acc[R][MB];
#pragma unroll
  for (int r = 0; r < R; ++r)
#pragma unroll
    for (int m = 0; m < MB; ++m) acc[r][m] = 0.f;
  for (int k = lane * 8; k < K; k += 256) {
    float xv[MB][8], xs[MB];
#pragma unroll
    for (int m = 0; m < MB; ++m) {
      load8<__nv_bfloat16>(x + m * x_stride + k, xv[m]);
      xs[m] = ((xv[m][0] + xv[m][1]) + (xv[m][2] + xv[m][3])) + ((xv[m][4] + xv[m][5]) + (xv[m][6] + xv[m][7]));
    }
    const int g = k / group_size;                        // the 8 codes of a lane never straddle a group (gs % 8 == 0)
#pragma unroll
    for (int r = 0; r < R; ++r) {
      if (n0 + r >= N) break;
      const uint8_t* src = codes + (long long)(n0 + r) * row_bytes + (long long)k * BITS / 8;
      float c[8];
      if (BITS == 8) {
        const uint2 u = *reinterpret_cast<const uint2*>(src);
        const uint32_t w[2] = {u.x, u.y};
#pragma unroll
        for (int i = 0; i < 8; ++i) c[i] = float((w[i >> 2] >> (8 * (i & 3))) & 0xFF);
      } else {
        const uint32_t u = *reinterpret_cast<const uint32_t*>(src);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const uint32_t byte = (u >> (8 * (i >> 1))) & 0xFF;          // byte i/2: value 2j in the high nibble
          c[i] = float((i & 1) ? (byte & 0xF) : (byte >> 4));
        }
      }
      const float inv_s = 1.f / __bfloat162float(scale[(long long)(n0 + r) * gpr + g]);
      const float mn = __bfloat162float(minv[(long long)(n0 + r) * gpr + g]);
#pragma unroll
      for (int m = 0; m < MB; ++m) {
        float dot = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) dot = fmaf(c[i], xv[m][i], dot);
        acc[r][m] += fmaf(dot, inv_s, mn * xs[m]);
      }
    }
  }
#pragma unroll
  for (int r = 0; r < R; ++r)
#pragma unroll
    for (int m = 0; m < MB; ++m) {
      float v = acc[r][m];
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
      if (lane == 0 && n0 + r < N) y[m * y_stride + n0 + r] = __float2bfloat16_rn(v);
    }
}

// ================================================================================== sign packing
// 1-bit optimizers: byte i holds the signs of elements 8i..8i+7, most significant bit first (numpy /
// cupy packbits order, which the reference relies on).  unpack writes +1 / -1.
template <typename T>
__global__ void packbits_kernel(const T* __restrict__ in, uint8_t* __restrict__ out, long long n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;  // output byte
  const long long nb = (n + 7) / 8;
  if (i >= nb) return;
  uint32_t b = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const long long e = i * 8 + j;
    const float x = e < n ? to_f32<T>(in[e]) : -1.f;
    b |= (x >= 0.f ? 1u : 0u) << (7 - j);
  }
  out[i] = (uint8_t)b;
}
template <typename T>
__global__ void unpackbits_kernel(const uint8_t* __restrict__ in, T* __restrict__ out, long long n, float scale) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long nb = (n + 7) / 8;
  if (i >= nb) return;
  const uint32_t b = in[i];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const long long e = i * 8 + j;
    if (e < n) out[e] = from_f32<T>(((b >> (7 - j)) & 1u) ? scale : -scale);
  }
}

// =================================================================================== random-LTD
// Sort each row of `idx` [rows, k] ascending (bitonic, k padded to a power of two <= 4096 in smem).
__global__ void token_sort_kernel(int* __restrict__ idx, int k, int kp2) {
  extern __shared__ int sh[];
  int* row = idx + (long long)blockIdx.x * k;
  for (int i = threadIdx.x; i < kp2; i += blockDim.x) sh[i] = i < k ? row[i] : 0x7FFFFFFF;
  __syncthreads();
  for (int size = 2; size <= kp2; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int i = threadIdx.x; i < kp2 / 2; i += blockDim.x) {
        const int lo = 2 * i - (i & (stride - 1));
        const int hi = lo + stride;
        const bool up = ((lo & size) == 0);
        const int a = sh[lo], b = sh[hi];
        if ((a > b) == up) { sh[lo] = b; sh[hi] = a; }
      }
      __syncthreads();
    }
  }
  for (int i = threadIdx.x; i < k; i += blockDim.x) row[i] = sh[i];
}

// gather: out[b, j, :] = act[b, idx[b, j], :]   (batch_first) or out[j, b, :] = act[idx[b, j], b, :].
// scatter is the inverse write into the full activation.  One CTA per (b, j); 16-byte vector copies.
template <bool SCATTER>
__global__ void token_move_kernel(uint4* __restrict__ full, uint4* __restrict__ part, const int* __restrict__ idx,
                                  int batch, int seq, int k, int hidden_vec, int batch_first) {
  const int j = blockIdx.x, b = blockIdx.y;
  const int s = idx[(long long)b * k + j];
  const long long fo = batch_first ? ((long long)b * seq + s) : ((long long)s * batch + b);
  const long long po = batch_first ? ((long long)b * k + j) : ((long long)j * batch + b);
  uint4* f = full + fo * hidden_vec;
  uint4* q = part + po * hidden_vec;
  for (int i = threadIdx.x; i < hidden_vec; i += blockDim.x) {
    if (SCATTER) f[i] = q[i];
    else q[i] = f[i];
  }
}

}  // namespace quant
}  // namespace sxb

// ============================================================================== C entry points
using namespace sxb;
using namespace sxb::quant;

namespace {

template <typename T, int BITS, bool SYM>
int launch_quant_int(const void* in, void* q, float* scales, long long groups, int gs, int stochastic,
                     unsigned seed, int nodes, int devs, cudaStream_t s) {
  const T* ip = static_cast<const T*>(in);
  uint8_t* qp = static_cast<uint8_t*>(q);
  if (gs <= 2048) quantize_int_kernel<T, BITS, SYM, 1><<<(unsigned)groups, kThreads, 0, s>>>(ip, qp, scales, groups, gs, stochastic, seed, nodes, devs);
  else if (gs <= 4096) quantize_int_kernel<T, BITS, SYM, 2><<<(unsigned)groups, kThreads, 0, s>>>(ip, qp, scales, groups, gs, stochastic, seed, nodes, devs);
  else quantize_int_kernel<T, BITS, SYM, 4><<<(unsigned)groups, kThreads, 0, s>>>(ip, qp, scales, groups, gs, stochastic, seed, nodes, devs);
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}
template <typename T>
int dispatch_quant_int(const void* in, void* q, float* scales, long long groups, int gs, int bits, int sym,
                       int stochastic, unsigned seed, int nodes, int devs, cudaStream_t s) {
  if (bits == 8) return sym ? launch_quant_int<T, 8, true>(in, q, scales, groups, gs, stochastic, seed, nodes, devs, s)
                            : launch_quant_int<T, 8, false>(in, q, scales, groups, gs, stochastic, seed, nodes, devs, s);
  return sym ? launch_quant_int<T, 4, true>(in, q, scales, groups, gs, stochastic, seed, nodes, devs, s)
             : launch_quant_int<T, 4, false>(in, q, scales, groups, gs, stochastic, seed, nodes, devs, s);
}
template <typename T>
int dispatch_dequant_int(const void* q, const float* scales, void* out, long long groups, int gs, int bits, int sym,
                         cudaStream_t s) {
  const uint8_t* qp = static_cast<const uint8_t*>(q);
  T* op = static_cast<T*>(out);
  const unsigned g = (unsigned)groups;
  if (bits == 8 && sym) dequantize_int_kernel<T, 8, true><<<g, kThreads, 0, s>>>(qp, scales, op, groups, gs);
  else if (bits == 8) dequantize_int_kernel<T, 8, false><<<g, kThreads, 0, s>>>(qp, scales, op, groups, gs);
  else if (sym) dequantize_int_kernel<T, 4, true><<<g, kThreads, 0, s>>>(qp, scales, op, groups, gs);
  else dequantize_int_kernel<T, 4, false><<<g, kThreads, 0, s>>>(qp, scales, op, groups, gs);
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}
template <int BITS, bool SYM>
int launch_qred(const void* qi, const float* si, void* qo, float* so, long long og, int gs, int devs, cudaStream_t s) {
  const uint8_t* a = static_cast<const uint8_t*>(qi);
  uint8_t* b = static_cast<uint8_t*>(qo);
  if (gs <= 2048) quantized_reduction_kernel<BITS, SYM, 1><<<(unsigned)og, kThreads, 0, s>>>(a, si, b, so, og, gs, devs);
  else if (gs <= 4096) quantized_reduction_kernel<BITS, SYM, 2><<<(unsigned)og, kThreads, 0, s>>>(a, si, b, so, og, gs, devs);
  else quantized_reduction_kernel<BITS, SYM, 4><<<(unsigned)og, kThreads, 0, s>>>(a, si, b, so, og, gs, devs);
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}
template <typename T, int E, int M>
int launch_fpq(const void* in, void* out, long long groups, int gs, int stochastic, unsigned seed, cudaStream_t s) {
  const T* ip = static_cast<const T*>(in);
  uint8_t* op = static_cast<uint8_t*>(out);
  if (gs <= 2048) fp_quantize_kernel<T, E, M, 1><<<(unsigned)groups, kThreads, 0, s>>>(ip, op, groups, gs, stochastic, seed);
  else if (gs <= 4096) fp_quantize_kernel<T, E, M, 2><<<(unsigned)groups, kThreads, 0, s>>>(ip, op, groups, gs, stochastic, seed);
  else fp_quantize_kernel<T, E, M, 4><<<(unsigned)groups, kThreads, 0, s>>>(ip, op, groups, gs, stochastic, seed);
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}
template <typename T>
int dispatch_fpq(const void* in, void* out, long long groups, int gs, int q_bits, int stochastic, unsigned seed,
                 cudaStream_t s) {
  switch (q_bits) {
    case 8: return launch_fpq<T, 4, 3>(in, out, groups, gs, stochastic, seed, s);
    case 6: return launch_fpq<T, 3, 2>(in, out, groups, gs, stochastic, seed, s);
    case 4: return launch_fpq<T, 2, 1>(in, out, groups, gs, stochastic, seed, s);
    case 12: return launch_fpq<T, 7, 4>(in, out, groups, gs, stochastic, seed, s);
  }
  return SXB_ERR_UNSUPPORTED;
}
template <typename T>
int dispatch_fpdq(const void* in, void* out, long long out_groups, int gs, int q_bits, const int* rows,
                  long long gpr, cudaStream_t s) {
  const uint8_t* ip = static_cast<const uint8_t*>(in);
  T* op = static_cast<T*>(out);
  const unsigned g = (unsigned)out_groups;
  switch (q_bits) {
    case 8: fp_dequantize_kernel<T, 4, 3><<<g, kThreads, 0, s>>>(ip, op, gs, rows, gpr); break;
    case 6: fp_dequantize_kernel<T, 3, 2><<<g, kThreads, 0, s>>>(ip, op, gs, rows, gpr); break;
    case 4: fp_dequantize_kernel<T, 2, 1><<<g, kThreads, 0, s>>>(ip, op, gs, rows, gpr); break;
    case 12: fp_dequantize_kernel<T, 7, 4><<<g, kThreads, 0, s>>>(ip, op, gs, rows, gpr); break;
    default: return SXB_ERR_UNSUPPORTED;
  }
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}
bool group_ok(long long groups, int gs) { return groups > 0 && gs > 0 && gs % 8 == 0 && gs <= 8192 && groups < (1ll << 31); }

}  // namespace

// Grouped integer quantisation.  in: [groups, group_size] of `dtype`; q: groups*group_size*bits/8 bytes;
// scales: fp32 [groups] (symmetric) or [groups, 2] = (scale, min) (asymmetric).  nodes/devs > 1 select the
// qgZ swizzled output order.  bits in {4, 8}.
SXB_API int sxb_quantize_int(const void* in, void* q, void* scales, long long groups, int group_size, int bits,
                             int symmetric, int stochastic, unsigned seed, int nodes, int devs, int dtype,
                             void* stream) {
  if (!group_ok(groups, group_size) || (bits != 4 && bits != 8)) return SXB_ERR_BAD_ARG;
  if (nodes > 1 && devs > 1 && groups % ((long long)nodes * devs)) return SXB_ERR_BAD_ARG;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  float* sc = static_cast<float*>(scales);
  switch (dtype) {
    case kBF16: return dispatch_quant_int<__nv_bfloat16>(in, q, sc, groups, group_size, bits, symmetric, stochastic, seed, nodes, devs, s);
    case kF16: return dispatch_quant_int<__half>(in, q, sc, groups, group_size, bits, symmetric, stochastic, seed, nodes, devs, s);
    case kF32: return dispatch_quant_int<float>(in, q, sc, groups, group_size, bits, symmetric, stochastic, seed, nodes, devs, s);
  }
  return SXB_ERR_UNSUPPORTED;
}

SXB_API int sxb_dequantize_int(const void* q, const void* scales, void* out, long long groups, int group_size,
                               int bits, int symmetric, int dtype, void* stream) {
  if (!group_ok(groups, group_size) || (bits != 4 && bits != 8)) return SXB_ERR_BAD_ARG;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const float* sc = static_cast<const float*>(scales);
  switch (dtype) {
    case kBF16: return dispatch_dequant_int<__nv_bfloat16>(q, sc, out, groups, group_size, bits, symmetric, s);
    case kF16: return dispatch_dequant_int<__half>(q, sc, out, groups, group_size, bits, symmetric, s);
    case kF32: return dispatch_dequant_int<float>(q, sc, out, groups, group_size, bits, symmetric, s);
  }
  return SXB_ERR_UNSUPPORTED;
}

// q_in: [devs, out_groups] groups, reduced over devs and re-quantised into q_out [out_groups].
SXB_API int sxb_quantized_reduction(const void* q_in, const void* s_in, void* q_out, void* s_out,
                                    long long out_groups, int group_size, int bits, int symmetric, int devs,
                                    void* stream) {
  if (!group_ok(out_groups, group_size) || (bits != 4 && bits != 8) || devs <= 0) return SXB_ERR_BAD_ARG;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const float* si = static_cast<const float*>(s_in);
  float* so = static_cast<float*>(s_out);
  if (bits == 8) return symmetric ? launch_qred<8, true>(q_in, si, q_out, so, out_groups, group_size, devs, s)
                                  : launch_qred<8, false>(q_in, si, q_out, so, out_groups, group_size, devs, s);
  return symmetric ? launch_qred<4, true>(q_in, si, q_out, so, out_groups, group_size, devs, s)
                   : launch_qred<4, false>(q_in, si, q_out, so, out_groups, group_size, devs, s);
}

// out: [groups, group_size*q_bits/8 + 4] bytes (payload then the fp32 scale).  q_bits in {4, 6, 8, 12}.
SXB_API int sxb_fp_quantize(const void* in, void* out, long long groups, int group_size, int q_bits,
                            int stochastic, unsigned seed, int dtype, void* stream) {
  if (!group_ok(groups, group_size)) return SXB_ERR_BAD_ARG;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  switch (dtype) {
    case kBF16: return dispatch_fpq<__nv_bfloat16>(in, out, groups, group_size, q_bits, stochastic, seed, s);
    case kF16: return dispatch_fpq<__half>(in, out, groups, group_size, q_bits, stochastic, seed, s);
    case kF32: return dispatch_fpq<float>(in, out, groups, group_size, q_bits, stochastic, seed, s);
  }
  return SXB_ERR_UNSUPPORTED;
}

// rows == NULL: dequantise all `out_groups` groups.  Otherwise output row r is input row rows[r], a row
// being groups_per_row consecutive groups (selective dequantisation, e.g. of the experts actually routed to).
SXB_API int sxb_fp_dequantize(const void* in, void* out, long long out_groups, int group_size, int q_bits,
                              const void* rows, long long groups_per_row, int dtype, void* stream) {
  if (!group_ok(out_groups, group_size)) return SXB_ERR_BAD_ARG;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const int* r = static_cast<const int*>(rows);
  switch (dtype) {
    case kBF16: return dispatch_fpdq<__nv_bfloat16>(in, out, out_groups, group_size, q_bits, r, groups_per_row, s);
    case kF16: return dispatch_fpdq<__half>(in, out, out_groups, group_size, q_bits, r, groups_per_row, s);
    case kF32: return dispatch_fpdq<float>(in, out, out_groups, group_size, q_bits, r, groups_per_row, s);
  }
  return SXB_ERR_UNSUPPORTED;
}

SXB_API int sxb_packbits(const void* in, void* out, long long n, int dtype, void* stream) {
  if (n <= 0) return SXB_OK;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const long long nb = (n + 7) / 8;
  const unsigned blocks = (unsigned)((nb + 255) / 256);
  uint8_t* o = static_cast<uint8_t*>(out);
  switch (dtype) {
    case kBF16: packbits_kernel<<<blocks, 256, 0, s>>>(static_cast<const __nv_bfloat16*>(in), o, n); break;
    case kF16: packbits_kernel<<<blocks, 256, 0, s>>>(static_cast<const __half*>(in), o, n); break;
    case kF32: packbits_kernel<<<blocks, 256, 0, s>>>(static_cast<const float*>(in), o, n); break;
    default: return SXB_ERR_UNSUPPORTED;
  }
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}
SXB_API int sxb_unpackbits(const void* in, void* out, long long n, float scale, int dtype, void* stream) {
  if (n <= 0) return SXB_OK;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const long long nb = (n + 7) / 8;
  const unsigned blocks = (unsigned)((nb + 255) / 256);
  const uint8_t* i8 = static_cast<const uint8_t*>(in);
  switch (dtype) {
    case kBF16: unpackbits_kernel<<<blocks, 256, 0, s>>>(i8, static_cast<__nv_bfloat16*>(out), n, scale); break;
    case kF16: unpackbits_kernel<<<blocks, 256, 0, s>>>(i8, static_cast<__half*>(out), n, scale); break;
    case kF32: unpackbits_kernel<<<blocks, 256, 0, s>>>(i8, static_cast<float*>(out), n, scale); break;
    default: return SXB_ERR_UNSUPPORTED;
  }
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

// idx: int32 [rows, k], sorted ascending in place per row (k <= 4096).
SXB_API int sxb_token_sort(void* idx, int rows, int k, void* stream) {
  if (rows <= 0 || k <= 0) return SXB_OK;
  if (k > 4096) return SXB_ERR_UNSUPPORTED;
  int kp2 = 1;
  while (kp2 < k) kp2 <<= 1;
  if (kp2 < 2) kp2 = 2;
  token_sort_kernel<<<rows, 256, kp2 * sizeof(int), reinterpret_cast<cudaStream_t>(stream)>>>(
      static_cast<int*>(idx), k, kp2);
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}
// full: [batch, seq, hidden] (batch_first) or [seq, batch, hidden]; part: same with seq -> k; idx int32 [batch, k].
// scatter = 0: part <- full[idx] ; scatter = 1: full[idx] <- part.  hidden*elem_size must be a multiple of 16.
SXB_API int sxb_token_move(void* full, void* part, const void* idx, int batch, int seq, int k, int hidden,
                           int elem_size, int batch_first, int scatter, void* stream) {
  if (batch <= 0 || k <= 0) return SXB_OK;
  if ((hidden * elem_size) % 16) return SXB_ERR_UNSUPPORTED;
  const int hv = hidden * elem_size / 16;
  dim3 grid(k, batch);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (scatter) token_move_kernel<true><<<grid, 128, 0, s>>>(static_cast<uint4*>(full), static_cast<uint4*>(part), static_cast<const int*>(idx), batch, seq, k, hv, batch_first);
  else token_move_kernel<false><<<grid, 128, 0, s>>>(static_cast<uint4*>(full), static_cast<uint4*>(part), static_cast<const int*>(idx), batch, seq, k, hv, batch_first);
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

// ------------------------------------------------------------------------------------------------
// Row-wise e4m3 quantisation for the tcgen05 fp8 GEMM (gemm_ext.cu): one scale per row = amax / 448, values
// stored as fp8 e4m3 in a [rows, cols] byte matrix of pitch ld_out.  (The reference's FP quantizer stores a
// per-group fp32 scale behind every group of a flat buffer — ops/fp_quantizer/quantize.py:48-95 — a layout a
// tensor-core operand cannot be fetched from; same number format, scales kept in a side vector.)
namespace sxb {
namespace quant {
__device__ __forceinline__ uint32_t f32x4_to_e4m3x4(float a, float b, float c, float d) {
  uint16_t lo, hi;
  asm volatile("cvt.rn.satfinite.e4m3x2.f32 %0, %2, %1;" : "=h"(lo) : "f"(a), "f"(b));
  asm volatile("cvt.rn.satfinite.e4m3x2.f32 %0, %2, %1;" : "=h"(hi) : "f"(c), "f"(d));
  return (uint32_t)lo | ((uint32_t)hi << 16);
}
__global__ void __launch_bounds__(256)
fp8_rowwise_quant_kernel(const __nv_bfloat16* __restrict__ in, uint8_t* __restrict__ out, float* __restrict__ scales,
                         long long rows, int cols, long long ld_in, long long ld_out) {
  __shared__ float red[32];
  for (long long r = blockIdx.x; r < rows; r += gridDim.x) {
    const __nv_bfloat16* x = in + r * ld_in;
    float amax = 0.f;
    for (int c = threadIdx.x * 8; c < cols; c += blockDim.x * 8) {
      const uint4 v = *reinterpret_cast<const uint4*>(x + c);
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
      for (int i = 0; i < 4; ++i)
        amax = fmaxf(amax, fmaxf(fabsf(__low2float(h[i])), fabsf(__high2float(h[i]))));
    }
    amax = block_max(amax, red);
    const float scale = amax > 0.f ? amax * (1.0f / 448.0f) : 1.0f;
    const float inv = 1.0f / scale;
    if (threadIdx.x == 0) scales[r] = scale;
    uint8_t* o = out + r * ld_out;
    for (int c = threadIdx.x * 8; c < cols; c += blockDim.x * 8) {
      const uint4 v = *reinterpret_cast<const uint4*>(x + c);
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
      uint2 q;
      q.x = f32x4_to_e4m3x4(__low2float(h[0]) * inv, __high2float(h[0]) * inv, __low2float(h[1]) * inv,
                            __high2float(h[1]) * inv);
      q.y = f32x4_to_e4m3x4(__low2float(h[2]) * inv, __high2float(h[2]) * inv, __low2float(h[3]) * inv,
                            __high2float(h[3]) * inv);
      *reinterpret_cast<uint2*>(o + c) = q;
    }
    __syncthreads();
  }
}
}  // namespace quant
}  // namespace sxb

// in: bf16 [rows, cols] (pitch ld_in elements, cols % 8 == 0, 16-byte aligned rows); out: e4m3 bytes (pitch ld_out,
// multiple of 8); scales: fp32 [rows].  Dequantised value = e4m3 * scale.
SXB_API int sxb_fp8_rowwise_quant(const void* in, void* out, void* scales, long long rows, int cols, long long ld_in,
                                  long long ld_out, void* stream) {
  if (rows <= 0) return SXB_OK;
  if (cols <= 0 || (cols % 8) || (ld_in % 8) || (ld_out % 8)) return SXB_ERR_UNSUPPORTED;
  const int grid = (int)(rows < 148 * 8 ? rows : 148 * 8);
  sxb::quant::fp8_rowwise_quant_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)in, (uint8_t*)out, (float*)scales, rows, cols, ld_in, ld_out);
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

namespace {
template <int E, int M>
int launch_fp_gemv(const void* x, const void* wq, void* y, int Mrows, int N, int K, int gs, long long xs, long long ys,
                   cudaStream_t s) {
  const __nv_bfloat16* xp = static_cast<const __nv_bfloat16*>(x);
  const uint8_t* wp = static_cast<const uint8_t*>(wq);
  __nv_bfloat16* yp = static_cast<__nv_bfloat16*>(y);
  const unsigned blocks = (unsigned)((N + 31) / 32);      // 8 warps x 4 rows per block
  for (int m0 = 0; m0 < Mrows;) {                         // batch rows in chunks of 4 / 2 / 1 (register budget)
    const int mb = Mrows - m0 >= 4 ? 4 : (Mrows - m0 >= 2 ? 2 : 1);
    const __nv_bfloat16* xc = xp + m0 * xs;
    __nv_bfloat16* yc = yp + m0 * ys;
    if (mb == 4) fp_gemv_kernel<E, M, 4><<<blocks, 256, 0, s>>>(xc, wp, yc, N, K, gs, xs, ys);
    else if (mb == 2) fp_gemv_kernel<E, M, 2><<<blocks, 256, 0, s>>>(xc, wp, yc, N, K, gs, xs, ys);
    else fp_gemv_kernel<E, M, 1><<<blocks, 256, 0, s>>>(xc, wp, yc, N, K, gs, xs, ys);
    m0 += mb;
  }
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}
}  // namespace

// Weight-only quantised linear for small batches: y[M, N] (bf16) = x[M, K] (bf16) W^T, W [N, K] in the layout written
// by sxb_fp_quantize (q_bits 4 / 6 / 8, groups of `group_size` along the flattened weight, K % group_size == 0,
// K % 512 == 0).  Intended for M <= 16 (decode); larger M should dequantise and use the tcgen05 GEMM.
SXB_API int sxb_fp_gemv(const void* x, const void* wq, void* y, int M, int N, int K, int q_bits, int group_size,
                        long long x_stride, long long y_stride, void* stream) {
  if (M <= 0 || N <= 0 || K <= 0) return SXB_ERR_BAD_ARG;
  if (group_size <= 0 || group_size % 16 || K % group_size || K % 512 || x_stride % 8) return SXB_ERR_UNSUPPORTED;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  switch (q_bits) {
    case 8: return launch_fp_gemv<4, 3>(x, wq, y, M, N, K, group_size, x_stride, y_stride, s);
    case 6: return launch_fp_gemv<3, 2>(x, wq, y, M, N, K, group_size, x_stride, y_stride, s);
    case 4: return launch_fp_gemv<2, 1>(x, wq, y, M, N, K, group_size, x_stride, y_stride, s);
  }
  return SXB_ERR_UNSUPPORTED;
}

namespace {
template <int BITS>
int launch_int_gemv(const void* x, const void* codes, const void* scale, const void* minv, void* y, int Mrows, int N,
                    int K, int gs, long long xs, long long ys, cudaStream_t s) {
  const __nv_bfloat16* xp = static_cast<const __nv_bfloat16*>(x);
  const uint8_t* cp = static_cast<const uint8_t*>(codes);
  const __nv_bfloat16* sp = static_cast<const __nv_bfloat16*>(scale);
  const __nv_bfloat16* mp = static_cast<const __nv_bfloat16*>(minv);
  __nv_bfloat16* yp = static_cast<__nv_bfloat16*>(y);
  const unsigned blocks = (unsigned)((N + 31) / 32);
  for (int m0 = 0; m0 < Mrows;) {
    const int mb = Mrows - m0 >= 4 ? 4 : (Mrows - m0 >= 2 ? 2 : 1);
    if (mb == 4) int_gemv_kernel<BITS, 4><<<blocks, 256, 0, s>>>(xp + m0 * xs, cp, sp, mp, yp + m0 * ys, N, K, gs, xs, ys);
    else if (mb == 2) int_gemv_kernel<BITS, 2><<<blocks, 256, 0, s>>>(xp + m0 * xs, cp, sp, mp, yp + m0 * ys, N, K, gs, xs, ys);
    else int_gemv_kernel<BITS, 1><<<blocks, 256, 0, s>>>(xp + m0 * xs, cp, sp, mp, yp + m0 * ys, N, K, gs, xs, ys);
    m0 += mb;
  }
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}
}  // namespace

// W8A16 / W4A16 linear for small batches on the inference/quantization layout (see int_gemv_kernel): x bf16 [M, K],
// codes uint8, scale / min bf16 [N, K / group_size], y bf16 [M, N].  K % 256 == 0, group_size % 8 == 0,
// K % group_size == 0.  (reference: inference/quantization dequantises and calls F.linear for every batch size.)
SXB_API int sxb_int_gemv(const void* x, const void* codes, const void* scale, const void* minv, void* y, int M, int N,
                         int K, int bits, int group_size, long long x_stride, long long y_stride, void* stream) {
  if (M <= 0 || N <= 0 || K <= 0) return SXB_ERR_BAD_ARG;
  if (group_size <= 0 || group_size % 8 || K % group_size || K % 256 || x_stride % 8) return SXB_ERR_UNSUPPORTED;
  if ((reinterpret_cast<uintptr_t>(codes) & 7) != 0) return SXB_ERR_UNSUPPORTED;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (bits == 8) return launch_int_gemv<8>(x, codes, scale, minv, y, M, N, K, group_size, x_stride, y_stride, s);
  if (bits == 4) return launch_int_gemv<4>(x, codes, scale, minv, y, M, N, K, group_size, x_stride, y_stride, s);
  return SXB_ERR_UNSUPPORTED;
}
