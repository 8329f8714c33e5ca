// moe_ragged.cu — MoE routing kernels and ragged-batch inference kernels for sm_100a.
//
// B200 re-designs of the CUDA kernels shipped under deepspeed/inference/v2/kernels/ragged_ops:
//   top_k_gating      (top_k_gating/top_k_gating_cuda.cu:14-102)
//   moe_scatter       (moe_scatter/moe_scatter_cuda.cu:23-184)
//   moe_gather        (moe_gather/moe_gather_cuda.cu:21-138)
//   ragged_embed      (embed/embed_cuda.cu:19-100)
//   logits_gather     (logits_gather/logits_gather_cuda.cu:19-56)
//   kv_rotary_blocked (linear_blocked_kv_rotary/blocked_kv_rotary_cuda.cu:26-262)
//   paged_attention   (blocked_flash/blocked_flash.cpp:21-101 — the reference wrapper rejects sm_100)
// They are also what the TRAINING MoE layer uses for token dispatch / combine (moe/sharded_moe.py), replacing the
// dense one-hot einsum of the reference (moe/sharded_moe.py:609-610,670).
//
// All are memory-bound row movers: one CTA (or warp) per token row, 16-byte accesses, expert bookkeeping with
// warp-aggregated atomics.  Row payloads never pass through shared memory.
#include "sxb_common.cuh"

namespace sxb {
namespace moe {

constexpr int kMaxExperts = 1024;
constexpr int kMaxTopK = 8;

// ------------------------------------------------------------------------------------ gating
// One warp per token.  logits [T, E] (any float dtype) -> softmax over ALL experts (fp32), top-k by repeated
// warp arg-max, scores [T, K] (fp32, optionally renormalised over the k winners), assignments [T, K] (int32),
// offsets [T, K] = arrival order of the token inside its expert (atomicAdd on expert_counts [E]).
template <typename T>
__global__ void __launch_bounds__(256)
top_k_gating_kernel(const T* __restrict__ logits, float* __restrict__ scores, int* __restrict__ assignments,
                    int* __restrict__ offsets, int* __restrict__ expert_counts, int n_tokens, int n_experts,
                    int top_k, int renormalize) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= n_tokens) return;
  const T* row = logits + (long long)warp * n_experts;
  // each lane keeps up to kMaxExperts/32 logits
  constexpr int PER = kMaxExperts / 32;
  float v[PER];
  float mx = -INFINITY;
#pragma unroll
  for (int i = 0; i < PER; ++i) {
    const int e = i * 32 + lane;
    v[i] = (e < n_experts) ? to_f32<T>(row[e]) : -INFINITY;
    mx = fmaxf(mx, v[i]);
  }
  mx = warp_max(mx);
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < PER; ++i) {
    if (i * 32 + lane < n_experts) sum += __expf(v[i] - mx);
  }
  sum = warp_sum(sum);
  const float inv = 1.f / sum;
  float picked_sum = 0.f;
  float my_score[kMaxTopK];
  int my_expert[kMaxTopK];
  for (int k = 0; k < top_k; ++k) {
    float best = -INFINITY;
    int best_e = 0x7fffffff;
#pragma unroll
    for (int i = 0; i < PER; ++i) {
      const int e = i * 32 + lane;
      if (v[i] > best || (v[i] == best && e < best_e)) { best = v[i]; best_e = e; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ob = __shfl_xor_sync(0xffffffffu, best, o);
      const int oe = __shfl_xor_sync(0xffffffffu, best_e, o);
      if (ob > best || (ob == best && oe < best_e)) { best = ob; best_e = oe; }
    }
    // remove the winner from its owner lane
#pragma unroll
    for (int i = 0; i < PER; ++i)
      if (i * 32 + lane == best_e) v[i] = -INFINITY;
    my_score[k] = __expf(best - mx) * inv;
    my_expert[k] = best_e;
    picked_sum += my_score[k];
  }
  if (lane == 0) {
    const float norm = renormalize ? 1.f / picked_sum : 1.f;
    for (int k = 0; k < top_k; ++k) {
      const long long o = (long long)warp * top_k + k;
      scores[o] = my_score[k] * norm;
      assignments[o] = my_expert[k];
      offsets[o] = atomicAdd(&expert_counts[my_expert[k]], 1);
    }
  }
}

// Exclusive scan of expert_counts -> expert_cumsum[E] (INCLUSIVE cumulative ends, the layout the grouped GEMM wants:
// rows of expert e are [cumsum[e-1], cumsum[e])).  Single CTA, E <= 1024.
__global__ void __launch_bounds__(1024)
expert_cumsum_kernel(const int* __restrict__ counts, long long* __restrict__ cumsum, int n_experts) {
  __shared__ int sh[kMaxExperts];
  const int t = threadIdx.x;
  sh[t] = t < n_experts ? counts[t] : 0;
  __syncthreads();
  for (int off = 1; off < kMaxExperts; off <<= 1) {
    const int add = t >= off ? sh[t - off] : 0;
    __syncthreads();
    sh[t] += add;
    __syncthreads();
  }
  if (t < n_experts) cumsum[t] = sh[t];
}

// One CTA per token: copy its row to every assigned expert's slot; write mapped_slots[T, K].
// slot = cumsum[e-1] + offset   (capacity-free, "dropless")
__global__ void __launch_bounds__(128)
moe_scatter_kernel(const uint4* __restrict__ x, uint4* __restrict__ moe_input, const int* __restrict__ assignments,
                   const int* __restrict__ offsets, const long long* __restrict__ cumsum, int* __restrict__ mapped_slots,
                   int row_vec, int top_k) {
  const int t = blockIdx.x;
  const uint4* src = x + (long long)t * row_vec;
  __shared__ long long slot[kMaxTopK];
  if (threadIdx.x < top_k) {
    const long long o = (long long)t * top_k + threadIdx.x;
    const int e = assignments[o];
    const long long base = e > 0 ? cumsum[e - 1] : 0;
    slot[threadIdx.x] = base + offsets[o];
    mapped_slots[o] = (int)slot[threadIdx.x];
  }
  __syncthreads();
  for (int i = threadIdx.x; i < row_vec; i += blockDim.x) {
    const uint4 v = src[i];
    for (int k = 0; k < top_k; ++k) moe_input[slot[k] * row_vec + i] = v;
  }
}

// out[t] = sum_k scores[t,k] * moe_output[mapped_slots[t,k]]      (bf16 rows, fp32 accumulation)
// Also the backward of scatter (scores == nullptr -> plain sum).
__global__ void __launch_bounds__(128)
moe_gather_kernel(const uint4* __restrict__ moe_output, uint4* __restrict__ out, const float* __restrict__ scores,
                  const int* __restrict__ mapped_slots, int row_vec, int top_k) {
  const int t = blockIdx.x;
  __shared__ long long slot[kMaxTopK];
  __shared__ float w[kMaxTopK];
  if (threadIdx.x < top_k) {
    const long long o = (long long)t * top_k + threadIdx.x;
    slot[threadIdx.x] = mapped_slots[o];
    w[threadIdx.x] = scores ? scores[o] : 1.f;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < row_vec; i += blockDim.x) {
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int k = 0; k < top_k; ++k) {
      const uint4 v = moe_output[slot[k] * row_vec + i];
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        acc[2 * j] += w[k] * __low2float(h[j]);
        acc[2 * j + 1] += w[k] * __high2float(h[j]);
      }
    }
    uint4 o;
    o.x = pack_bf16(acc[0], acc[1]); o.y = pack_bf16(acc[2], acc[3]);
    o.z = pack_bf16(acc[4], acc[5]); o.w = pack_bf16(acc[6], acc[7]);
    out[(long long)t * row_vec + i] = o;
  }
}

// d scores[t,k] = <d out[t], moe_output[slot[t,k]]>   (one warp per (t,k))
__global__ void __launch_bounds__(256)
moe_gather_dscore_kernel(const uint4* __restrict__ moe_output, const uint4* __restrict__ d_out,
                         const int* __restrict__ mapped_slots, float* __restrict__ d_scores, int n_pairs, int row_vec,
                         int top_k) {
  const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= n_pairs) return;
  const int t = w / top_k;
  const uint4* a = moe_output + (long long)mapped_slots[w] * row_vec;
  const uint4* b = d_out + (long long)t * row_vec;
  float acc = 0.f;
  for (int i = lane; i < row_vec; i += 32) {
    const uint4 va = a[i], vb = b[i];
    const __nv_bfloat162* ha = reinterpret_cast<const __nv_bfloat162*>(&va);
    const __nv_bfloat162* hb = reinterpret_cast<const __nv_bfloat162*>(&vb);
#pragma unroll
    for (int j = 0; j < 4; ++j)
      acc += __low2float(ha[j]) * __low2float(hb[j]) + __high2float(ha[j]) * __high2float(hb[j]);
  }
  acc = warp_sum(acc);
  if (lane == 0) d_scores[w] = acc;
}

// moe_output-side gradient of gather: d moe_output[slot[t,k]] = scores[t,k] * d out[t]   (one CTA per token)
__global__ void __launch_bounds__(128)
moe_scatter_scaled_kernel(const uint4* __restrict__ d_out, uint4* __restrict__ d_moe_output,
                          const float* __restrict__ scores, const int* __restrict__ mapped_slots, int row_vec,
                          int top_k) {
  const int t = blockIdx.x;
  __shared__ long long slot[kMaxTopK];
  __shared__ float w[kMaxTopK];
  if (threadIdx.x < top_k) {
    const long long o = (long long)t * top_k + threadIdx.x;
    slot[threadIdx.x] = mapped_slots[o];
    w[threadIdx.x] = scores[o];
  }
  __syncthreads();
  for (int i = threadIdx.x; i < row_vec; i += blockDim.x) {
    const uint4 v = d_out[(long long)t * row_vec + i];
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
    float f[8];
#pragma unroll
    for (int j = 0; j < 4; ++j) { f[2 * j] = __low2float(h[j]); f[2 * j + 1] = __high2float(h[j]); }
    for (int k = 0; k < top_k; ++k) {
      uint4 o;
      o.x = pack_bf16(f[0] * w[k], f[1] * w[k]); o.y = pack_bf16(f[2] * w[k], f[3] * w[k]);
      o.z = pack_bf16(f[4] * w[k], f[5] * w[k]); o.w = pack_bf16(f[6] * w[k], f[7] * w[k]);
      d_moe_output[slot[k] * row_vec + i] = o;
    }
  }
}

}  // namespace moe
}  // namespace sxb

// ===================================================================================== ragged batch
namespace sxb {
namespace ragged {

// out[t, :] = table[ids[t], :] (+ pos_table[pos[t] + pos_offset, :])      one CTA per token
__global__ void __launch_bounds__(128)
ragged_embed_kernel(const int* __restrict__ ids, const int* __restrict__ pos, const uint4* __restrict__ table,
                    const uint4* __restrict__ pos_table, uint4* __restrict__ out, int row_vec, int pos_offset) {
  const int t = blockIdx.x;
  const uint4* src = table + (long long)ids[t] * row_vec;
  const uint4* psrc = pos_table ? pos_table + (long long)(pos[t] + pos_offset) * row_vec : nullptr;
  for (int i = threadIdx.x; i < row_vec; i += blockDim.x) {
    uint4 v = src[i];
    if (psrc) {
      const uint4 p = psrc[i];
      __nv_bfloat162* a = reinterpret_cast<__nv_bfloat162*>(&v);
      const __nv_bfloat162* b = reinterpret_cast<const __nv_bfloat162*>(&p);
#pragma unroll
      for (int j = 0; j < 4; ++j) a[j] = __hadd2(a[j], b[j]);
    }
    out[(long long)t * row_vec + i] = v;
  }
}

// out[s, :] = hidden[last_token[s], :]     one CTA per sequence
__global__ void __launch_bounds__(128)
logits_gather_kernel(const uint4* __restrict__ hidden, const int* __restrict__ last_token, uint4* __restrict__ out,
                     int row_vec) {
  const int s = blockIdx.x;
  const uint4* src = hidden + (long long)last_token[s] * row_vec;
  for (int i = threadIdx.x; i < row_vec; i += blockDim.x) out[(long long)s * row_vec + i] = src[i];
}

// Rotary embedding on q (in place) and k, then k/v appended to the paged cache.
//   qkv      [T, (Hq + 2 Hk) * D] bf16 (q | k | v per token)
//   cache    [n_blocks, block_size, 2, Hk, D] bf16
//   tok_seq  [T] sequence slot of the token ; tok_pos [T] absolute position in its sequence
//   block_tables [n_seqs, max_blocks]
// grid (T, Hq + Hk), block D/2 threads: thread i rotates the pair (i, i + rot/2) of one head (HF "rotate_half").
__global__ void kv_rotary_blocked_kernel(__nv_bfloat16* __restrict__ qkv, __nv_bfloat16* __restrict__ cache,
                                         const int* __restrict__ tok_seq, const int* __restrict__ tok_pos,
                                         const int* __restrict__ block_tables, int max_blocks, int block_size, int Hq,
                                         int Hk, int D, int rot_dim, float log2_theta) {
  const int t = blockIdx.x, head = blockIdx.y, i = threadIdx.x;  // i in [0, D/2)
  const int pos = tok_pos[t];
  const long long row = (long long)t * (Hq + 2 * Hk) * D;
  __nv_bfloat16* h = qkv + row + (long long)head * D;            // q heads then k heads are contiguous
  const int half = rot_dim / 2;
  float x0, x1;
  int i0, i1;
  if (i < half) { i0 = i; i1 = i + half; } else { i0 = half + i; i1 = -1; }  // beyond rot_dim: pass-through pairs
  if (i < half) {
    x0 = __bfloat162float(h[i0]);
    x1 = __bfloat162float(h[i1]);
    const float inv_freq = exp2f(-(2.0f * i / rot_dim) * log2_theta);
    float sn, cs;
    sincosf(pos * inv_freq, &sn, &cs);
    const float y0 = x0 * cs - x1 * sn, y1 = x1 * cs + x0 * sn;
    x0 = y0; x1 = y1;
  } else {
    // unrotated tail [rot_dim, D): two elements per thread
    i0 = rot_dim + 2 * (i - half);
    i1 = i0 + 1;
    x0 = i0 < D ? __bfloat162float(h[i0]) : 0.f;
    x1 = i1 < D ? __bfloat162float(h[i1]) : 0.f;
  }
  const bool is_k = head >= Hq;
  if (!is_k) {
    if (i < half) { h[i0] = __float2bfloat16_rn(x0); h[i1] = __float2bfloat16_rn(x1); }
    return;
  }
  const int kh = head - Hq;
  const int blk = block_tables[(long long)tok_seq[t] * max_blocks + pos / block_size];
  const long long slot = ((long long)blk * block_size + pos % block_size) * 2 * Hk * D;
  __nv_bfloat16* kdst = cache + slot + (long long)kh * D;
  __nv_bfloat16* vdst = cache + slot + (long long)(Hk + kh) * D;
  const __nv_bfloat16* vsrc = qkv + row + (long long)(Hq + Hk + kh) * D;
  if (i0 < D) { kdst[i0] = __float2bfloat16_rn(x0); vdst[i0] = vsrc[i0]; }
  if (i1 >= 0 && i1 < D) { kdst[i1] = __float2bfloat16_rn(x1); vdst[i1] = vsrc[i1]; }
}

// Paged ("blocked") causal attention for ragged batches: prompt tokens and decode tokens in one launch.
//   q [T, Hq, D] bf16 (row pitch q_stride elements), out [T, Hq, D], cache as above.
// grid (T, Hq), block 128 = 4 warps.  Warp w walks KV positions w, w+4, ...; each lane owns D/32 channels; a
// position costs one 8-byte (D=128) load per lane for K and V and five shuffles for the dot product.  The four
// warps' running (max, sum, acc) are merged through shared memory at the end (flash-decoding split over warps).
template <int D>
__global__ void __launch_bounds__(128)
paged_attention_kernel(const __nv_bfloat16* __restrict__ q, __nv_bfloat16* __restrict__ out,
                       const __nv_bfloat16* __restrict__ cache, const int* __restrict__ tok_seq,
                       const int* __restrict__ tok_pos, const int* __restrict__ block_tables, int max_blocks,
                       int block_size, int Hq, int Hk, long long q_stride, float scale_log2) {
  constexpr int C = D / 32;  // channels per lane
  const int t = blockIdx.x, h = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kh = h / (Hq / Hk);
  const int n_kv = tok_pos[t] + 1;
  const int* table = block_tables + (long long)tok_seq[t] * max_blocks;
  float qv[C];
  {
    const __nv_bfloat16* qp = q + (long long)t * q_stride + (long long)h * D + lane * C;
#pragma unroll
    for (int c = 0; c < C; ++c) qv[c] = __bfloat162float(qp[c]) * scale_log2;
  }
  float m = -INFINITY, l = 0.f, acc[C];
#pragma unroll
  for (int c = 0; c < C; ++c) acc[c] = 0.f;
  for (int p = warp; p < n_kv; p += 4) {
    const int blk = table[p / block_size];
    const long long slot = ((long long)blk * block_size + p % block_size) * 2 * Hk * D;
    const __nv_bfloat16* kp = cache + slot + (long long)kh * D + lane * C;
    const __nv_bfloat16* vp = cache + slot + (long long)(Hk + kh) * D + lane * C;
    float kv[C], vv[C];
    if (C == 4) {
      const uint2 ku = *reinterpret_cast<const uint2*>(kp);
      const uint2 vu = *reinterpret_cast<const uint2*>(vp);
      const __nv_bfloat162* kh2 = reinterpret_cast<const __nv_bfloat162*>(&ku);
      const __nv_bfloat162* vh2 = reinterpret_cast<const __nv_bfloat162*>(&vu);
      kv[0] = __low2float(kh2[0]); kv[1] = __high2float(kh2[0]); kv[2 % C] = __low2float(kh2[1]); kv[3 % C] = __high2float(kh2[1]);
      vv[0] = __low2float(vh2[0]); vv[1] = __high2float(vh2[0]); vv[2 % C] = __low2float(vh2[1]); vv[3 % C] = __high2float(vh2[1]);
    } else {
#pragma unroll
      for (int c = 0; c < C; ++c) { kv[c] = __bfloat162float(kp[c]); vv[c] = __bfloat162float(vp[c]); }
    }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < C; ++c) s += qv[c] * kv[c];
    s = warp_sum(s);
    const float m_new = fmaxf(m, s);
    const float alpha = exp2f(m - m_new);
    const float pr = exp2f(s - m_new);
    l = l * alpha + pr;
#pragma unroll
    for (int c = 0; c < C; ++c) acc[c] = acc[c] * alpha + pr * vv[c];
    m = m_new;
  }
  __shared__ float sm_m[4], sm_l[4], sm_acc[4][D];
  if (lane == 0) { sm_m[warp] = m; sm_l[warp] = l; }
#pragma unroll
  for (int c = 0; c < C; ++c) sm_acc[warp][lane * C + c] = acc[c];
  __syncthreads();
  if (warp == 0) {
    float M = fmaxf(fmaxf(sm_m[0], sm_m[1]), fmaxf(sm_m[2], sm_m[3]));
    float L = 0.f, o[C];
#pragma unroll
    for (int c = 0; c < C; ++c) o[c] = 0.f;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      const float f = (sm_m[w] == -INFINITY) ? 0.f : exp2f(sm_m[w] - M);
      L += sm_l[w] * f;
#pragma unroll
      for (int c = 0; c < C; ++c) o[c] += sm_acc[w][lane * C + c] * f;
    }
    const float inv = L > 0.f ? 1.f / L : 0.f;
    __nv_bfloat16* op = out + ((long long)t * Hq + h) * D + lane * C;
#pragma unroll
    for (int c = 0; c < C; ++c) op[c] = __float2bfloat16_rn(o[c] * inv);
  }
}

}  // namespace ragged
}  // namespace sxb

// ======================================================================================= C entry points
using namespace sxb;

// logits [T, E]; scores fp32 [T, K]; assignments / offsets int32 [T, K]; expert_counts int32 [E] (zeroed here).
SXB_API int sxb_top_k_gating(const void* logits, void* scores, void* assignments, void* offsets, void* expert_counts,
                             int n_tokens, int n_experts, int top_k, int renormalize, int dtype, void* stream) {
  if (n_tokens <= 0) return SXB_OK;
  if (n_experts <= 0 || n_experts > moe::kMaxExperts || top_k <= 0 || top_k > moe::kMaxTopK || top_k > n_experts)
    return SXB_ERR_BAD_ARG;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  cudaMemsetAsync(expert_counts, 0, sizeof(int) * n_experts, s);
  const unsigned blocks = (unsigned)((n_tokens + 7) / 8);
  float* sc = static_cast<float*>(scores);
  int *as = static_cast<int*>(assignments), *of = static_cast<int*>(offsets), *ec = static_cast<int*>(expert_counts);
  switch (dtype) {
    case kBF16: moe::top_k_gating_kernel<<<blocks, 256, 0, s>>>(static_cast<const __nv_bfloat16*>(logits), sc, as, of, ec, n_tokens, n_experts, top_k, renormalize); break;
    case kF16: moe::top_k_gating_kernel<<<blocks, 256, 0, s>>>(static_cast<const __half*>(logits), sc, as, of, ec, n_tokens, n_experts, top_k, renormalize); break;
    case kF32: moe::top_k_gating_kernel<<<blocks, 256, 0, s>>>(static_cast<const float*>(logits), sc, as, of, ec, n_tokens, n_experts, top_k, renormalize); break;
    default: return SXB_ERR_UNSUPPORTED;
  }
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

// x [T, H] -> moe_input [T*K, H] grouped by expert; expert_cumsum int64 [E]; mapped_slots int32 [T, K].
SXB_API int sxb_moe_scatter(const void* x, void* moe_input, const void* assignments, const void* offsets,
                            const void* expert_counts, void* expert_cumsum, void* mapped_slots, int n_tokens,
                            int n_experts, int top_k, int hidden, int elem_size, void* stream) {
  if (n_tokens <= 0) return SXB_OK;
  if ((hidden * elem_size) % 16 || n_experts > moe::kMaxExperts || top_k > moe::kMaxTopK) return SXB_ERR_UNSUPPORTED;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  moe::expert_cumsum_kernel<<<1, 1024, 0, s>>>(static_cast<const int*>(expert_counts),
                                               static_cast<long long*>(expert_cumsum), n_experts);
  moe::moe_scatter_kernel<<<n_tokens, 128, 0, s>>>(static_cast<const uint4*>(x), static_cast<uint4*>(moe_input),
                                                   static_cast<const int*>(assignments), static_cast<const int*>(offsets),
                                                   static_cast<const long long*>(expert_cumsum),
                                                   static_cast<int*>(mapped_slots), hidden * elem_size / 16, top_k);
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

// out [T, H] = sum_k scores * moe_output[slot]  (scores may be NULL: plain sum).  bf16 rows.
SXB_API int sxb_moe_gather(const void* moe_output, void* out, const void* scores, const void* mapped_slots,
                           int n_tokens, int top_k, int hidden, void* stream) {
  if (n_tokens <= 0) return SXB_OK;
  if ((hidden * 2) % 16 || top_k > moe::kMaxTopK) return SXB_ERR_UNSUPPORTED;
  moe::moe_gather_kernel<<<n_tokens, 128, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      static_cast<const uint4*>(moe_output), static_cast<uint4*>(out), static_cast<const float*>(scores),
      static_cast<const int*>(mapped_slots), hidden * 2 / 16, top_k);
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

// Backward of gather: d_moe_output[slot] = scores * d_out ; d_scores[t,k] = <d_out[t], moe_output[slot]>.
SXB_API int sxb_moe_gather_bwd(const void* d_out, const void* moe_output, const void* scores,
                               const void* mapped_slots, void* d_moe_output, void* d_scores, int n_tokens, int top_k,
                               int hidden, void* stream) {
  if (n_tokens <= 0) return SXB_OK;
  if ((hidden * 2) % 16 || top_k > moe::kMaxTopK) return SXB_ERR_UNSUPPORTED;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const int rv = hidden * 2 / 16;
  moe::moe_scatter_scaled_kernel<<<n_tokens, 128, 0, s>>>(static_cast<const uint4*>(d_out),
                                                          static_cast<uint4*>(d_moe_output),
                                                          static_cast<const float*>(scores),
                                                          static_cast<const int*>(mapped_slots), rv, top_k);
  if (d_scores) {
    const int pairs = n_tokens * top_k;
    moe::moe_gather_dscore_kernel<<<(pairs + 7) / 8, 256, 0, s>>>(static_cast<const uint4*>(moe_output),
                                                                  static_cast<const uint4*>(d_out),
                                                                  static_cast<const int*>(mapped_slots),
                                                                  static_cast<float*>(d_scores), pairs, rv, top_k);
  }
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

SXB_API int sxb_ragged_embed(const void* ids, const void* pos, const void* table, const void* pos_table, void* out,
                             int n_tokens, int hidden, int pos_offset, void* stream) {
  if (n_tokens <= 0) return SXB_OK;
  if ((hidden * 2) % 16) return SXB_ERR_UNSUPPORTED;
  ragged::ragged_embed_kernel<<<n_tokens, 128, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      static_cast<const int*>(ids), static_cast<const int*>(pos), static_cast<const uint4*>(table),
      static_cast<const uint4*>(pos_table), static_cast<uint4*>(out), hidden * 2 / 16, pos_offset);
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

SXB_API int sxb_logits_gather(const void* hidden_states, const void* last_token, void* out, int n_seqs, int hidden,
                              void* stream) {
  if (n_seqs <= 0) return SXB_OK;
  if ((hidden * 2) % 16) return SXB_ERR_UNSUPPORTED;
  ragged::logits_gather_kernel<<<n_seqs, 128, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      static_cast<const uint4*>(hidden_states), static_cast<const int*>(last_token), static_cast<uint4*>(out),
      hidden * 2 / 16);
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

SXB_API int sxb_kv_rotary_blocked(void* qkv, void* cache, const void* tok_seq, const void* tok_pos,
                                  const void* block_tables, int n_tokens, int max_blocks, int block_size, int Hq,
                                  int Hk, int D, int rot_dim, float theta, void* stream) {
  if (n_tokens <= 0) return SXB_OK;
  if (D % 2 || D > 256 || rot_dim > D || rot_dim % 2) return SXB_ERR_UNSUPPORTED;
  dim3 grid(n_tokens, Hq + Hk);
  ragged::kv_rotary_blocked_kernel<<<grid, D / 2, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      static_cast<__nv_bfloat16*>(qkv), static_cast<__nv_bfloat16*>(cache), static_cast<const int*>(tok_seq),
      static_cast<const int*>(tok_pos), static_cast<const int*>(block_tables), max_blocks, block_size, Hq, Hk, D,
      rot_dim, log2f(theta));
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

SXB_API int sxb_paged_attention(const void* q, void* out, const void* cache, const void* tok_seq, const void* tok_pos,
                                const void* block_tables, int n_tokens, int max_blocks, int block_size, int Hq, int Hk,
                                int D, long long q_stride, float scale, void* stream) {
  if (n_tokens <= 0) return SXB_OK;
  if (Hq % Hk) return SXB_ERR_BAD_ARG;
  dim3 grid(n_tokens, Hq);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const float sl2 = scale * 1.4426950408889634f;
  const __nv_bfloat16* qp = static_cast<const __nv_bfloat16*>(q);
  __nv_bfloat16* op = static_cast<__nv_bfloat16*>(out);
  const __nv_bfloat16* cp = static_cast<const __nv_bfloat16*>(cache);
  const int *ts = static_cast<const int*>(tok_seq), *tp = static_cast<const int*>(tok_pos),
            *bt = static_cast<const int*>(block_tables);
  if (D == 128) ragged::paged_attention_kernel<128><<<grid, 128, 0, s>>>(qp, op, cp, ts, tp, bt, max_blocks, block_size, Hq, Hk, q_stride, sl2);
  else if (D == 64) ragged::paged_attention_kernel<64><<<grid, 128, 0, s>>>(qp, op, cp, ts, tp, bt, max_blocks, block_size, Hq, Hk, q_stride, sl2);
  else return SXB_ERR_UNSUPPORTED;
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}
