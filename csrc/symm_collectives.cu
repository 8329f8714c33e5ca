// symm_collectives.cu — collectives fused with their adjacent compute, issued from inside the kernel
// over NVLink 5 / NVSwitch peer mappings (symmetric memory: every rank holds VA mappings of all
// peers' buffers and of their signal pads; optionally an NVLS multicast mapping).
//
//   sxb_symm_rs_adam       K1: gradient reduce-scatter (+1/n, unscale/clip) (+) AdamW on the fp32 master
//                              shard (+) bf16 parameter-shard write-back, one pass, no NCCL.
//                              Replaces reduce_scatter_coalesced + partition_grads + FusedAdam
//                              (runtime/comm/coalesced_collectives.py:159-218, runtime/zero/stage3.py:1388-1540, 2137-2152).
//   sxb_symm_subset_mean   K9: shuffle-exchange replica averaging — the ring membership table is the
//                              kernel's peer-address schedule (runtime/zero/stage_1_and_2.py:2227-2250).
//   sxb_symm_all_gather    K2 (gather half): pull-based all-gather that publishes per-chunk arrival
//                              flags so a concurrently running GEMM can start on arrived weight tiles
//                              (runtime/zero/partition_parameters.py:1221-1261).
//   sxb_symm_a2a_rows      K4-K7 transport: direct peer stores of token rows (MoE dispatch/combine,
//                              Ulysses head-scatter) with per-destination row tables
//                              (moe/sharded_moe.py:96-108, sequence/layer.py:221-254).
//
// Cross-rank ordering uses monotonically increasing epochs written with st.release.sys and polled
// with ld.acquire.sys on per-(channel, peer) 32-bit slots of the signal pad; nothing is ever reset,
// so there is no ABA window.  Every spin is bounded (~2 s) and raises the library abort flag.
#include "sxb_common.cuh"

namespace sxb {
namespace symm {

constexpr int kMaxPeers = 16;
constexpr int kThreads = 512;

struct Peers {
  void* buf[kMaxPeers];        // peer mappings of the payload buffer (index = rank in group)
  uint32_t* sig[kMaxPeers];    // peer mappings of the signal pad
  int rank;
  int world;
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// 16-byte peer load that does not pollute L1 (peer data is never re-read by this SM)
__device__ __forceinline__ uint4 ld_peer16(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ uint4 ld_volatile16(const void* p) {
  uint4 r;
  asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

// slot layout of a signal pad: [channel][kMaxPeers] uint32
__device__ __forceinline__ uint32_t* slot(uint32_t* pad, int channel, int peer) {
  return pad + channel * kMaxPeers + peer;
}

// Tell every peer "rank `me` reached `epoch` on `channel`".
__device__ __forceinline__ void signal_all(const Peers& P, int channel, uint32_t epoch) {
  if (threadIdx.x < P.world) st_release_sys(slot(P.sig[threadIdx.x], channel, P.rank), epoch);
}
// Wait until every peer reached `epoch` on `channel` (threads 0..world-1 poll, then block-sync).
__device__ __forceinline__ bool wait_all(const Peers& P, int channel, uint32_t epoch, uint32_t tag) {
  __shared__ int ok_flag;
  if (threadIdx.x == 0) ok_flag = 1;
  __syncthreads();
  if (threadIdx.x < P.world) {
    const uint32_t* s = slot(P.sig[P.rank], channel, threadIdx.x);
    unsigned long long t0 = 0;
    long long c_next = 0;
    while ((int32_t)(ld_acquire_sys(s) - epoch) < 0) {
      // cross-rank waits absorb start-up skew between processes (allocator warm-up, first-touch page faults):
      // 30 s, not the 2 s of the on-chip pipelines — a false alarm here corrupts the step
      if (hang_should_abort(t0, tag, threadIdx.x, &c_next, 30000000000ull)) { ok_flag = 0; break; }
    }
  }
  __syncthreads();
  return ok_flag != 0;
}

__device__ __forceinline__ void bf16x8_acc(const uint4& v, float (&a)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    a[2 * i] += __low2float(h[i]);
    a[2 * i + 1] += __high2float(h[i]);
  }
}

struct AdamArgs {
  float lr, beta1, beta2, eps, weight_decay, bc1, bc2, grad_scale;
  int adamw;
};

// ------------------------------------------------------------------------------------ K1
// Each rank reduces ITS shard [rank*part, (rank+1)*part) of every peer's flat bf16 gradient buffer and
// immediately applies AdamW.  channel 0 = "gradients ready", channel 1 = "finished reading".
struct RsAdamArgs {
  Peers P;
  long long part;
  float *hp, *m, *v;
  __nv_bfloat16 *lp, *grad_out;
  AdamArgs a;
  uint32_t epoch;
  const void* mc_ptr;
  unsigned int* cta_counter;
  int apply_adam, do_sync;
};
__device__ __forceinline__ void rs_adam_body(const RsAdamArgs& A) {
  const Peers& P = A.P;
  const long long part = A.part;
  float* __restrict__ hp = A.hp; float* __restrict__ m = A.m; float* __restrict__ v = A.v;
  __nv_bfloat16* __restrict__ lp = A.lp; __nv_bfloat16* __restrict__ grad_out = A.grad_out;
  const AdamArgs a = A.a;
  const uint32_t epoch = A.epoch;
  const void* mc_ptr = A.mc_ptr;
  unsigned int* __restrict__ cta_counter = A.cta_counter;
  const int apply_adam = A.apply_adam, do_sync = A.do_sync;
  // do_sync == 0: the "gradients ready" handshake already ran as a one-warp kernel in front of this one on the same
  // stream, so no CTA of this (SM-filling) grid ever spins while it holds an SM next to the backward GEMMs.
  if (do_sync) {
    if (blockIdx.x == 0) signal_all(P, 0, epoch);
    if (!wait_all(P, 0, epoch, 101)) return;
  }
  const long long base = (long long)P.rank * part;
  const long long nvec = part >> 3;
  const float inv_n = 1.0f / (float)P.world;
  constexpr int U = 2;     // independent 16-byte reduce loads in flight per thread (NVLink round trip ~4 us)
  const long long tstride = (long long)gridDim.x * blockDim.x;
  for (long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x; i0 < nvec; i0 += U * tstride) {
    float g[U][8];
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
      for (int k = 0; k < 8; ++k) g[u][k] = 0.f;
    if (mc_ptr) {
      // NVLS: one load returns the in-switch sum over all ranks (fp32 accumulation inside the switch)
      uint4 r[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long i = i0 + u * tstride;
        if (i < nvec) {
          const __nv_bfloat16* src = reinterpret_cast<const __nv_bfloat16*>(mc_ptr) + base + i * 8;
          asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
                       : "=r"(r[u].x), "=r"(r[u].y), "=r"(r[u].z), "=r"(r[u].w)
                       : "l"(src)
                       : "memory");
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (i0 + u * tstride < nvec) bf16x8_acc(r[u], g[u]);
    } else {
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long i = i0 + u * tstride;
        if (i >= nvec) continue;
        uint4 r[kMaxPeers];
#pragma unroll
        for (int pi = 0; pi < kMaxPeers; ++pi) {
          if (pi < P.world) {
            const int peer = (P.rank + pi) % P.world;       // stagger so ranks do not hammer one peer
            r[pi] = ld_peer16(reinterpret_cast<const __nv_bfloat16*>(P.buf[peer]) + base + i * 8);
          }
        }
#pragma unroll
        for (int pi = 0; pi < kMaxPeers; ++pi)
          if (pi < P.world) bf16x8_acc(r[pi], g[u]);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * tstride;
      if (i >= nvec) continue;
      if (apply_adam) {
        float4 p0 = *reinterpret_cast<float4*>(hp + i * 8), p1 = *reinterpret_cast<float4*>(hp + i * 8 + 4);
        float4 m0 = *reinterpret_cast<float4*>(m + i * 8), m1 = *reinterpret_cast<float4*>(m + i * 8 + 4);
        float4 v0 = *reinterpret_cast<float4*>(v + i * 8), v1 = *reinterpret_cast<float4*>(v + i * 8 + 4);
        float pp[8] = {p0.x, p0.y, p0.z, p0.w, p1.x, p1.y, p1.z, p1.w};
        float mm[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
        float vv[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          float gg = g[u][k] * inv_n * a.grad_scale;
          if (!a.adamw) gg += a.weight_decay * pp[k];
          mm[k] = a.beta1 * mm[k] + (1.f - a.beta1) * gg;
          vv[k] = a.beta2 * vv[k] + (1.f - a.beta2) * gg * gg;
          float upd = (mm[k] / a.bc1) / (sqrtf(vv[k] / a.bc2) + a.eps);
          if (a.adamw) upd += a.weight_decay * pp[k];
          pp[k] -= a.lr * upd;
        }
        *reinterpret_cast<float4*>(hp + i * 8) = make_float4(pp[0], pp[1], pp[2], pp[3]);
        *reinterpret_cast<float4*>(hp + i * 8 + 4) = make_float4(pp[4], pp[5], pp[6], pp[7]);
        *reinterpret_cast<float4*>(m + i * 8) = make_float4(mm[0], mm[1], mm[2], mm[3]);
        *reinterpret_cast<float4*>(m + i * 8 + 4) = make_float4(mm[4], mm[5], mm[6], mm[7]);
        *reinterpret_cast<float4*>(v + i * 8) = make_float4(vv[0], vv[1], vv[2], vv[3]);
        *reinterpret_cast<float4*>(v + i * 8 + 4) = make_float4(vv[4], vv[5], vv[6], vv[7]);
        uint4 o;
        o.x = pack_bf16(pp[0], pp[1]); o.y = pack_bf16(pp[2], pp[3]);
        o.z = pack_bf16(pp[4], pp[5]); o.w = pack_bf16(pp[6], pp[7]);
        *reinterpret_cast<uint4*>(lp + i * 8) = o;
      } else {
        uint4 o;
        o.x = pack_bf16(g[u][0] * inv_n, g[u][1] * inv_n); o.y = pack_bf16(g[u][2] * inv_n, g[u][3] * inv_n);
        o.z = pack_bf16(g[u][4] * inv_n, g[u][5] * inv_n); o.w = pack_bf16(g[u][6] * inv_n, g[u][7] * inv_n);
        *reinterpret_cast<uint4*>(grad_out + i * 8) = o;
      }
    }
  }
  // last CTA to finish tells the peers that this rank no longer reads their gradient buffers
  __syncthreads();
  __shared__ unsigned int last;
  if (threadIdx.x == 0) {
    __threadfence();
    last = (atomicAdd(cta_counter, 1u) == gridDim.x - 1) ? 1u : 0u;
  }
  __syncthreads();
  if (last) {
    if (threadIdx.x == 0) *cta_counter = 0u;
    signal_all(P, 1, epoch);
  }
}
__global__ void __launch_bounds__(kThreads) rs_adam_kernel(RsAdamArgs A) { rs_adam_body(A); }
// Single-GPU emulation of `world` ranks as ONE cooperative grid (blockIdx.y = logical rank): the ranks wait on one
// another, so they must be co-resident by construction — separate launches on one GPU give no such guarantee.
__global__ void __launch_bounds__(kThreads) rs_adam_emu_kernel(const RsAdamArgs* per_rank) {
  rs_adam_body(per_rank[blockIdx.y]);
}

// Wait until every peer finished reading (channel 1) — launched before a symmetric buffer is reused.
__global__ void wait_channel_kernel(Peers P, int channel, uint32_t epoch) { wait_all(P, channel, epoch, 102); }

// ------------------------------------------------------------------------------------ K9
// out[i] = mean over `members` (ranks in the column group) of their symmetric shard; fp32 accumulate.
struct Members {
  int idx[kMaxPeers];
  float w[kMaxPeers];   // out = sum_k w[k] * shard[idx[k]]  (mean: 1/count; H-RR: 1/(rings*ring size); push-sum: alpha ratios)
  int count;
};
template <typename T>
struct SubsetMeanArgs {
  Peers P;
  Members M;
  T* out;
  long long n;
  uint32_t epoch;
  unsigned int* cta_counter;
  int me;   // >= 0: two-phase form, this rank's position in M.idx; -1: every rank reads every member's whole shard
};

// Two-phase form for member sets in which EVERY member computes the same weighted sum (RR, H-RR): member k reduces
// slice k of all members' shards into slice k of its OWN symmetric buffer (no rank reads that slice of that buffer in
// this phase), a cross-rank barrier, then every member copies the reduced slices out of their owners.  Per rank
// 2 (m-1)/m shard sizes cross NVLink instead of (m-1): the reduce-scatter + all-gather of a bandwidth-optimal
// all-reduce, in one kernel.  All CTAs wait inside the kernel for the slowest rank, so the launch is cooperative
// (co-residency is checked by the driver).  n must be a multiple of 8.
template <typename T>
__device__ __forceinline__ void subset_mean_two_phase(const SubsetMeanArgs<T>& A) {
  const Peers& P = A.P;
  const Members& M = A.M;
  if (blockIdx.x == 0) signal_all(P, 2, A.epoch);
  if (!wait_all(P, 2, A.epoch, 103)) return;
  const long long nvec = A.n >> 3;
  const long long per = (nvec + M.count - 1) / M.count;
  const long long stride = (long long)gridDim.x * blockDim.x, t0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  {
    const long long lo = per * A.me, hi = min(nvec, lo + per);
    T* mine = reinterpret_cast<T*>(const_cast<void*>(P.buf[P.rank]));
    // two vectors per thread per iteration: 2 x count independent peer loads in flight
    for (long long i = lo + t0; i < hi; i += 2 * stride) {
      const long long i1 = i + stride;
      const bool two = i1 < hi;
      float acc[2][8];
#pragma unroll
      for (int e = 0; e < 8; ++e) { acc[0][e] = 0.f; acc[1][e] = 0.f; }
      for (int k = 0; k < M.count; ++k) {
        const T* src = reinterpret_cast<const T*>(P.buf[M.idx[k]]);
        const uint4 r0 = ld_peer16(src + i * 8);
        uint4 r1 = make_uint4(0u, 0u, 0u, 0u);
        if (two) r1 = ld_peer16(src + i1 * 8);
        const T* h0 = reinterpret_cast<const T*>(&r0);
        const T* h1 = reinterpret_cast<const T*>(&r1);
        const float wk = M.w[k];
#pragma unroll
        for (int e = 0; e < 8; ++e) { acc[0][e] += wk * to_f32<T>(h0[e]); acc[1][e] += wk * to_f32<T>(h1[e]); }
      }
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        if (u == 1 && !two) break;
        uint4 o;
        T* oh = reinterpret_cast<T*>(&o);
#pragma unroll
        for (int e = 0; e < 8; ++e) oh[e] = from_f32<T>(acc[u][e]);
        const long long at = (u == 0 ? i : i1) * 8;
        *reinterpret_cast<uint4*>(mine + at) = o;
        *reinterpret_cast<uint4*>(A.out + at) = o;
      }
    }
  }
  // every CTA of this rank finished its part of the slice -> tell the peers; then wait for theirs
  __shared__ unsigned int last;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    last = (atomicAdd(A.cta_counter, 1u) == gridDim.x - 1) ? 1u : 0u;
  }
  __syncthreads();
  if (last) {
    if (threadIdx.x == 0) *A.cta_counter = 0u;
    signal_all(P, 7, A.epoch);
  }
  if (!wait_all(P, 7, A.epoch, 107)) return;
  for (int k = 0; k < M.count; ++k) {
    if (k == A.me) continue;
    const long long lo = per * k, hi = min(nvec, lo + per);
    const T* src = reinterpret_cast<const T*>(P.buf[M.idx[k]]);
    // four independent 16-byte peer loads in flight per thread (one load per iteration left NVLink at ~370 GB/s)
    for (long long i = lo + t0; i < hi; i += 4 * stride) {
      uint4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (i + u * stride < hi) v[u] = ld_volatile16(src + (i + u * stride) * 8);   // written by the owner in THIS kernel
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (i + u * stride < hi) *reinterpret_cast<uint4*>(A.out + (i + u * stride) * 8) = v[u];
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    last = (atomicAdd(A.cta_counter + 1, 1u) == gridDim.x - 1) ? 1u : 0u;
  }
  __syncthreads();
  if (last) {
    if (threadIdx.x == 0) A.cta_counter[1] = 0u;
    signal_all(P, 3, A.epoch);
  }
}

template <typename T>
__device__ __forceinline__ void subset_mean_body(const SubsetMeanArgs<T>& A) {
  if (A.me >= 0) { subset_mean_two_phase<T>(A); return; }
  const Peers& P = A.P;
  const Members& M = A.M;
  T* __restrict__ out = A.out;
  const long long n = A.n;
  const uint32_t epoch = A.epoch;
  unsigned int* __restrict__ cta_counter = A.cta_counter;
  if (blockIdx.x == 0) signal_all(P, 2, epoch);
  if (!wait_all(P, 2, epoch, 103)) return;
  const long long nvec = n >> 3;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec;
       i += (long long)gridDim.x * blockDim.x) {
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int k = 0; k < M.count; ++k) {
      const uint4 r = ld_peer16(reinterpret_cast<const T*>(P.buf[M.idx[k]]) + i * 8);
      const T* h = reinterpret_cast<const T*>(&r);
      const float wk = M.w[k];
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] += wk * to_f32<T>(h[e]);
    }
    uint4 o;
    T* oh = reinterpret_cast<T*>(&o);
#pragma unroll
    for (int e = 0; e < 8; ++e) oh[e] = from_f32<T>(acc[e]);
    *reinterpret_cast<uint4*>(out + i * 8) = o;
  }
  for (long long i = (nvec << 3) + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    float acc = 0.f;
    for (int k = 0; k < M.count; ++k) acc += M.w[k] * to_f32<T>(reinterpret_cast<const T*>(P.buf[M.idx[k]])[i]);
    out[i] = from_f32<T>(acc);
  }
  __syncthreads();
  __shared__ unsigned int last;
  if (threadIdx.x == 0) {
    __threadfence();
    last = (atomicAdd(cta_counter, 1u) == gridDim.x - 1) ? 1u : 0u;
  }
  __syncthreads();
  if (last) {
    if (threadIdx.x == 0) *cta_counter = 0u;
    signal_all(P, 3, epoch);
  }
}
template <typename T>
__global__ void __launch_bounds__(kThreads) subset_mean_kernel(SubsetMeanArgs<T> A) { subset_mean_body<T>(A); }
template <typename T>
__global__ void __launch_bounds__(kThreads) subset_mean_emu_kernel(const SubsetMeanArgs<T>* per_rank) {
  subset_mean_body<T>(per_rank[blockIdx.y]);
}

// ------------------------------------------------------------------------------------ K2 (gather)
// Pull-based all-gather: dst[(r*part) ...] <- peer r's shard, chunk by chunk; after each chunk the CTA
// that copied it bumps `chunk_flags[chunk]` to `epoch` (release, gpu scope) so a GEMM running on the
// remaining SMs can consume weight tiles as they arrive.  Chunks are visited in flat order.
struct AllGatherArgs {
  Peers P;
  long long part_bytes;
  uint8_t* dst;
  long long chunk_bytes;
  uint32_t* chunk_flags;
  uint32_t epoch;       // handshake epoch (0 = barrier-free pull)
  uint32_t flag_value;  // value published in chunk_flags
};
__device__ __forceinline__ void all_gather_body(const AllGatherArgs& A) {
  const Peers& P = A.P;
  const long long part_bytes = A.part_bytes;
  uint8_t* __restrict__ dst = A.dst;
  const long long chunk_bytes = A.chunk_bytes;
  uint32_t* __restrict__ chunk_flags = A.chunk_flags;
  const uint32_t epoch = A.epoch;
  // epoch == 0: barrier-free pull (the caller established shard readiness once per step with sxb_symm_barrier)
  if (epoch != 0) {
    if (blockIdx.x == 0) signal_all(P, 4, epoch);
    if (!wait_all(P, 4, epoch, 104)) return;
  }
  const long long total = part_bytes * P.world;
  const long long nchunks = (total + chunk_bytes - 1) / chunk_bytes;
  for (long long c = blockIdx.x; c < nchunks; c += gridDim.x) {
    const long long lo = c * chunk_bytes;
    const long long hi = min(lo + chunk_bytes, total);
    // 4 independent 16-byte peer loads in flight per thread (NVLink latency ~2 us: keep >1.5 MB in flight)
    const long long stride = (long long)blockDim.x * 16;
    for (long long off = lo + (long long)threadIdx.x * 16; off < hi; off += 4 * stride) {
      uint4 val[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const long long o = off + u * stride;
        if (o < hi) {
          const int r = (int)(o / part_bytes);
          const long long in_off = o - (long long)r * part_bytes;
          const uint8_t* sp = reinterpret_cast<const uint8_t*>(P.buf[r]) + in_off;
          val[u] = (r == P.rank) ? *reinterpret_cast<const uint4*>(sp) : ld_peer16(sp);
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const long long o = off + u * stride;
        if (o < hi) *reinterpret_cast<uint4*>(dst + o) = val[u];
      }
    }
    __syncthreads();
    if (threadIdx.x == 0 && chunk_flags) {
      __threadfence();
      asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(chunk_flags + c), "r"(A.flag_value) : "memory");
    }
  }
}
__global__ void __launch_bounds__(kThreads) all_gather_kernel(AllGatherArgs A) { all_gather_body(A); }
__global__ void __launch_bounds__(kThreads) all_gather_emu_kernel(const AllGatherArgs* per_rank) {
  all_gather_body(per_rank[blockIdx.y]);
}

// ------------------------------------------------------------------------------------ a2a rows
// Rank-local view of an all-to-all of fixed-size rows: for every destination d, rows
// [send_off[d], send_off[d] + count[d]) of `src` are stored into peer d's symmetric buffer starting at
// row recv_off[d] (offsets precomputed on the host or by a count exchange).  One warp per row.
struct A2A {
  long long send_off[kMaxPeers];
  long long recv_off[kMaxPeers];
  long long count[kMaxPeers];
};
struct A2AArgs {
  Peers P;
  A2A T;
  const uint8_t* src;
  int row_bytes;
  uint32_t epoch;
  unsigned int* cta_counter;
};
__device__ __forceinline__ void a2a_rows_body(const A2AArgs& A) {
  const Peers& P = A.P;
  const A2A& T = A.T;
  const uint8_t* __restrict__ src = A.src;
  const int row_bytes = A.row_bytes;
  const uint32_t epoch = A.epoch;
  unsigned int* __restrict__ cta_counter = A.cta_counter;
  // channel 5: destination buffers free (previous consumer done) ; channel 6: my rows have landed
  if (blockIdx.x == 0) signal_all(P, 5, epoch);
  if (!wait_all(P, 5, epoch, 105)) return;
  long long total = 0;
  for (int d = 0; d < P.world; ++d) total += T.count[d];
  const int warps_per_block = blockDim.x >> 5;
  const int lane = threadIdx.x & 31;
  for (long long w = (long long)blockIdx.x * warps_per_block + (threadIdx.x >> 5); w < total;
       w += (long long)gridDim.x * warps_per_block) {
    long long r = w;
    int d = 0;
    while (r >= T.count[d]) { r -= T.count[d]; ++d; }
    const int peer = (d + P.rank) % P.world;   // not used for addressing; keeps traffic staggered
    (void)peer;
    const uint8_t* s = src + (T.send_off[d] + r) * (long long)row_bytes;
    uint8_t* t = reinterpret_cast<uint8_t*>(P.buf[d]) + (T.recv_off[d] + r) * (long long)row_bytes;
    for (int off = lane * 16; off < row_bytes; off += 32 * 16)
      *reinterpret_cast<uint4*>(t + off) = *reinterpret_cast<const uint4*>(s + off);
  }
  __syncthreads();
  __shared__ unsigned int last;
  if (threadIdx.x == 0) {
    __threadfence_system();
    last = (atomicAdd(cta_counter, 1u) == gridDim.x - 1) ? 1u : 0u;
  }
  __syncthreads();
  if (last) {
    if (threadIdx.x == 0) *cta_counter = 0u;
    __threadfence_system();
    signal_all(P, 6, epoch);
  }
}
__global__ void __launch_bounds__(256) a2a_rows_kernel(A2AArgs A) { a2a_rows_body(A); }
__global__ void __launch_bounds__(256) a2a_rows_emu_kernel(const A2AArgs* per_rank) { a2a_rows_body(per_rank[blockIdx.y]); }

static int fill_peers(Peers& P, const void* const* bufs, const void* const* sigs, int rank, int world) {
  if (world < 1 || world > kMaxPeers || rank < 0 || rank >= world) return SXB_ERR_BAD_ARG;
  for (int i = 0; i < kMaxPeers; ++i) {
    P.buf[i] = i < world ? const_cast<void*>(bufs[i]) : nullptr;
    P.sig[i] = i < world ? reinterpret_cast<uint32_t*>(const_cast<void*>(sigs[i])) : nullptr;
  }
  P.rank = rank;
  P.world = world;
  return SXB_OK;
}

static int grid_cap(long long work_items, int max_ctas) {
  long long g = (work_items + kThreads - 1) / kThreads;
  if (g < 1) g = 1;
  const int cap = max_ctas > 0 ? max_ctas : 148 * 2;
  if (g > cap) g = cap;
  return (int)g;
}

static AdamArgs make_adam(float lr, float beta1, float beta2, float eps, float weight_decay, int step, int adamw,
                          int bias_correction, float grad_scale) {
  AdamArgs a;
  a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.weight_decay = weight_decay;
  a.bc1 = bias_correction ? 1.f - powf(beta1, (float)step) : 1.f;
  a.bc2 = bias_correction ? 1.f - powf(beta2, (float)step) : 1.f;
  a.grad_scale = grad_scale; a.adamw = adamw;
  return a;
}

// Emulation launches: per-rank argument structs are copied to a device staging area (stream-ordered), then ONE
// cooperative grid (x = CTAs per logical rank, y = world) runs every rank's body.
template <typename Args>
static int launch_emu(const void* kernel, const Args* host_args, int world, int ctas, int threads, cudaStream_t s) {
  static void* staging[2] = {nullptr, nullptr};
  static int which = 0;
  const size_t bytes = sizeof(Args) * kMaxPeers;
  if (bytes > (1u << 16)) return SXB_ERR_TOO_LARGE;
  which ^= 1;
  if (!staging[which]) {
    cudaError_t e = cudaMalloc(&staging[which], 1u << 16);
    if (e != cudaSuccess) return (int)e;
  }
  cudaError_t e = cudaMemcpyAsync(staging[which], host_args, sizeof(Args) * world, cudaMemcpyHostToDevice, s);
  if (e != cudaSuccess) return (int)e;
  e = cudaStreamSynchronize(s);     // host_args lives on the caller's stack
  if (e != cudaSuccess) return (int)e;
  const Args* dev_args = reinterpret_cast<const Args*>(staging[which]);
  void* params[] = {(void*)&dev_args};
  e = cudaLaunchCooperativeKernel(kernel, dim3(ctas, world), dim3(threads), params, 0, s);
  return e == cudaSuccess ? SXB_OK : (int)e;
}

}  // namespace symm
}  // namespace sxb

using namespace sxb;
using namespace sxb::symm;

SXB_DEFINE_HANG_ACCESSOR(sxb_symm_hang_info)

// Full barrier on `channel`: tell every peer "I reached `epoch`", then wait until all of them did.
__global__ void barrier_channel_kernel(sxb::symm::Peers P, int channel, uint32_t epoch) {
  sxb::symm::signal_all(P, channel, epoch);
  sxb::symm::wait_all(P, channel, epoch, 110);
}

// One-warp "rank reached `epoch` on `channel`" announcement, stream-ordered after the work it publishes.
__global__ void signal_channel_kernel(sxb::symm::Peers P, int channel, uint32_t epoch) {
  __threadfence_system();
  sxb::symm::signal_all(P, channel, epoch);
}

// K1.  bufs/sigs: host arrays of `world` device pointers (peer mappings of the flat bf16 gradient buffer
// and of the signal pads).  part: elements per shard (multiple of 8).  counter: device uint32 (zeroed once).
// apply_adam == 0 -> plain reduce-scatter(mean) into grad_out (used when the step must be deferred).
// handshake_done != 0: the caller already announced "gradients ready" (sxb_symm_signal on the stream that produced
// them) and waited for the peers (sxb_symm_wait on this stream); no CTA of this grid spins.
SXB_API int sxb_symm_rs_adam(const void* const* bufs, const void* const* sigs, int rank, int world, long long part,
                             void* hp, void* m, void* v, void* lp, void* grad_out, float lr, float beta1,
                             float beta2, float eps, float weight_decay, int step, int adamw, int bias_correction,
                             float grad_scale, unsigned int epoch, const void* multicast_ptr, void* counter,
                             int apply_adam, int max_ctas, int handshake_done, void* stream) {
  RsAdamArgs A;
  int rc = fill_peers(A.P, bufs, sigs, rank, world);
  if (rc) return rc;
  if (part <= 0 || (part & 7)) return SXB_ERR_UNSUPPORTED;
  A.part = part; A.hp = (float*)hp; A.m = (float*)m; A.v = (float*)v;
  A.lp = (__nv_bfloat16*)lp; A.grad_out = (__nv_bfloat16*)grad_out;
  A.a = make_adam(lr, beta1, beta2, eps, weight_decay, step, adamw, bias_correction, grad_scale);
  A.epoch = epoch; A.mc_ptr = multicast_ptr; A.cta_counter = (unsigned int*)counter; A.apply_adam = apply_adam;
  // SXB_RS_SPLIT_SYNC=1 runs the in-kernel handshake as a one-warp kernel in front of the big one
  static const int split_env = [] { const char* e = getenv("SXB_RS_SPLIT_SYNC"); return e ? atoi(e) : 0; }();
  const int split = split_env && !handshake_done;
  if (split) {
    barrier_channel_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(A.P, 0, epoch);
    SXB_CHECK_LAUNCH();
  }
  A.do_sync = (split || handshake_done) ? 0 : 1;
  rs_adam_kernel<<<grid_cap(part / 8, max_ctas), kThreads, 0, (cudaStream_t)stream>>>(A);
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

// K1, `world` logical ranks emulated on ONE GPU by one cooperative grid.  bufs/sigs as above (all local);
// hp/m/v/lp/grad_out/counters: host arrays of `world` per-rank device pointers.  ctas_per_rank * world CTAs of 512
// threads must be co-resident (cudaLaunchCooperativeKernel refuses otherwise).
SXB_API int sxb_symm_rs_adam_emu(const void* const* bufs, const void* const* sigs, int world, long long part,
                                 void* const* hp, void* const* m, void* const* v, void* const* lp,
                                 void* const* grad_out, float lr, float beta1, float beta2, float eps,
                                 float weight_decay, int step, int adamw, int bias_correction, float grad_scale,
                                 unsigned int epoch, void* const* counters, int apply_adam, int ctas_per_rank,
                                 void* stream) {
  if (world < 1 || world > kMaxPeers || part <= 0 || (part & 7) || ctas_per_rank < 1) return SXB_ERR_BAD_ARG;
  RsAdamArgs A[kMaxPeers];
  for (int r = 0; r < world; ++r) {
    int rc = fill_peers(A[r].P, bufs, sigs, r, world);
    if (rc) return rc;
    A[r].part = part;
    A[r].hp = apply_adam ? (float*)hp[r] : nullptr; A[r].m = apply_adam ? (float*)m[r] : nullptr;
    A[r].v = apply_adam ? (float*)v[r] : nullptr; A[r].lp = apply_adam ? (__nv_bfloat16*)lp[r] : nullptr;
    A[r].grad_out = apply_adam ? nullptr : (__nv_bfloat16*)grad_out[r];
    A[r].a = make_adam(lr, beta1, beta2, eps, weight_decay, step, adamw, bias_correction, grad_scale);
    A[r].epoch = epoch; A[r].mc_ptr = nullptr; A[r].cta_counter = (unsigned int*)counters[r];
    A[r].apply_adam = apply_adam; A[r].do_sync = 1;
  }
  return launch_emu((const void*)rs_adam_emu_kernel, A, world, ctas_per_rank, kThreads, (cudaStream_t)stream);
}

SXB_API int sxb_symm_barrier(const void* const* bufs, const void* const* sigs, int rank, int world, int channel,
                             unsigned int epoch, void* stream) {
  Peers P;
  int rc = fill_peers(P, bufs, sigs, rank, world);
  if (rc) return rc;
  barrier_channel_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(P, channel, epoch);
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

SXB_API int sxb_symm_signal(const void* const* bufs, const void* const* sigs, int rank, int world, int channel,
                            unsigned int epoch, void* stream) {
  Peers P;
  int rc = fill_peers(P, bufs, sigs, rank, world);
  if (rc) return rc;
  signal_channel_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(P, channel, epoch);
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

SXB_API int sxb_symm_wait(const void* const* bufs, const void* const* sigs, int rank, int world, int channel,
                          unsigned int epoch, void* stream) {
  Peers P;
  int rc = fill_peers(P, bufs, sigs, rank, world);
  if (rc) return rc;
  wait_channel_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(P, channel, epoch);
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

SXB_API int sxb_symm_subset_wsum2(const void* const* bufs, const void* const* sigs, int rank, int world,
                                  const int* members, const float* weights, int n_members, void* out, long long n,
                                  int dtype, unsigned int epoch, void* counter, int max_ctas, int shared_result,
                                  void* stream);
SXB_API int sxb_symm_subset_wsum(const void* const* bufs, const void* const* sigs, int rank, int world,
                                 const int* members, const float* weights, int n_members, void* out, long long n,
                                 int dtype, unsigned int epoch, void* counter, int max_ctas, void* stream);
// K9.  members: host array of `n_members` group ranks (must contain `rank`); out: local (non-symmetric) result.
static int fill_members(Members& M, const int* members, int n_members, const float* weights = nullptr) {
  if (n_members < 1 || n_members > kMaxPeers) return SXB_ERR_BAD_ARG;
  M.count = n_members;
  for (int i = 0; i < kMaxPeers; ++i) {
    M.idx[i] = i < n_members ? members[i] : 0;
    M.w[i] = i < n_members ? (weights ? weights[i] : 1.0f / (float)n_members) : 0.f;
  }
  return SXB_OK;
}
SXB_API int sxb_symm_subset_mean(const void* const* bufs, const void* const* sigs, int rank, int world,
                                 const int* members, int n_members, void* out, long long n, int dtype,
                                 unsigned int epoch, void* counter, int max_ctas, void* stream) {
  return sxb_symm_subset_wsum(bufs, sigs, rank, world, members, nullptr, n_members, out, n, dtype, epoch, counter,
                              max_ctas, stream);
}
// K9, general form: out = sum_k weights[k] * shard of members[k] (weights == NULL: the mean).  One kernel serves RR
// (all replicas), shuffle (my ring), H-RR (two-level mean as per-member weights) and Gossip push-sum (alpha ratios).
SXB_API int sxb_symm_subset_wsum(const void* const* bufs, const void* const* sigs, int rank, int world,
                                 const int* members, const float* weights, int n_members, void* out, long long n,
                                 int dtype, unsigned int epoch, void* counter, int max_ctas, void* stream) {
  return sxb_symm_subset_wsum2(bufs, sigs, rank, world, members, weights, n_members, out, n, dtype, epoch, counter,
                               max_ctas, 0, stream);
}

template <typename T>
static int launch_subset(const Peers& P, const Members& M, void* out, long long n, unsigned int epoch, void* counter,
                         int max_ctas, int me, cudaStream_t s) {
  SubsetMeanArgs<T> A{P, M, (T*)out, n, epoch, (unsigned int*)counter, me};
  if (me < 0) {
    subset_mean_kernel<T><<<grid_cap(n / 8 + 1, max_ctas), kThreads, 0, s>>>(A);
    return SXB_OK;
  }
  // cooperative: every CTA waits inside the kernel for the other ranks, so all of them must be resident
  int per_sm = 0, dev = 0, sms = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, subset_mean_kernel<T>, kThreads, 0);
  if (e != cudaSuccess) return (int)e;
  if (per_sm < 1) return SXB_ERR_UNSUPPORTED;
  int grid = grid_cap(n / 8 / (M.count > 0 ? M.count : 1) + 1, max_ctas);
  if (grid > per_sm * sms) grid = per_sm * sms;
  void* params[] = {(void*)&A};
  e = cudaLaunchCooperativeKernel((const void*)subset_mean_kernel<T>, dim3(grid), dim3(kThreads), params, 0, s);
  return e == cudaSuccess ? SXB_OK : (int)e;
}

// K9 with the choice of form.  shared_result != 0: every member of `members` calls with the SAME member list and
// weights (RR, H-RR) -> two-phase form (see subset_mean_two_phase): `counter` must point at two zeroed words, n % 8 == 0,
// and the symmetric buffer is overwritten with reduced values.  Falls back to the one-phase form for a single member.
SXB_API int sxb_symm_subset_wsum2(const void* const* bufs, const void* const* sigs, int rank, int world,
                                  const int* members, const float* weights, int n_members, void* out, long long n,
                                  int dtype, unsigned int epoch, void* counter, int max_ctas, int shared_result,
                                  void* stream) {
  Peers P;
  int rc = fill_peers(P, bufs, sigs, rank, world);
  if (rc) return rc;
  Members M;
  if ((rc = fill_members(M, members, n_members, weights))) return rc;
  int me = -1;
  if (shared_result && n_members > 1) {
    if (n % 8) return SXB_ERR_UNSUPPORTED;
    for (int i = 0; i < n_members; ++i)
      if (members[i] == rank) me = i;
    if (me < 0) return SXB_ERR_BAD_ARG;
  }
  cudaStream_t s = (cudaStream_t)stream;
  if (dtype == kBF16) rc = launch_subset<__nv_bfloat16>(P, M, out, n, epoch, counter, max_ctas, me, s);
  else if (dtype == kF16) rc = launch_subset<__half>(P, M, out, n, epoch, counter, max_ctas, me, s);
  else return SXB_ERR_BAD_ARG;
  if (rc) return rc;
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}
// K9 emulated (bf16).  As in the real launch EVERY rank runs the kernel, each over its own ring: members is a
// [world][16] row-major table, n_members[r] the length of row r (row r must contain r).
SXB_API int sxb_symm_subset_mean_emu(const void* const* bufs, const void* const* sigs, int world, const int* members,
                                     const int* n_members, void* const* outs, long long n, unsigned int epoch,
                                     void* const* counters, int ctas_per_rank, int shared_result, void* stream) {
  // shared_result: two-phase form (every rank's row of `members` lists the same ring for all of the ring's ranks);
  // counters[r] then points at two zeroed words
  if (world < 1 || world > kMaxPeers || ctas_per_rank < 1) return SXB_ERR_BAD_ARG;
  if (shared_result && (n % 8)) return SXB_ERR_UNSUPPORTED;
  int rc;
  SubsetMeanArgs<__nv_bfloat16> A[kMaxPeers];
  for (int r = 0; r < world; ++r) {
    if ((rc = fill_peers(A[r].P, bufs, sigs, r, world))) return rc;
    if ((rc = fill_members(A[r].M, members + r * kMaxPeers, n_members[r]))) return rc;
    A[r].out = (__nv_bfloat16*)outs[r]; A[r].n = n; A[r].epoch = epoch;
    A[r].cta_counter = (unsigned int*)counters[r];
    A[r].me = -1;
    if (shared_result && n_members[r] > 1) {
      for (int i = 0; i < n_members[r]; ++i)
        if (members[r * kMaxPeers + i] == r) A[r].me = i;
      if (A[r].me < 0) return SXB_ERR_BAD_ARG;
    }
  }
  return launch_emu((const void*)subset_mean_emu_kernel<__nv_bfloat16>, A, world, ctas_per_rank, kThreads,
                    (cudaStream_t)stream);
}

// K2 gather half.  chunk_flags (optional, device uint32[ceil(world*part_bytes / chunk_bytes)]): chunk c is set to
// `flag_value` (release, gpu scope) once all of its bytes are in dst — the gated GEMM polls these.
SXB_API int sxb_symm_all_gather(const void* const* bufs, const void* const* sigs, int rank, int world,
                                long long part_bytes, void* dst, long long chunk_bytes, void* chunk_flags,
                                unsigned int epoch, int max_ctas, unsigned int flag_value, void* stream) {
  AllGatherArgs A;
  int rc = fill_peers(A.P, bufs, sigs, rank, world);
  if (rc) return rc;
  if ((part_bytes & 15) || (chunk_bytes & 15) || chunk_bytes <= 0) return SXB_ERR_UNSUPPORTED;
  const long long nchunks = (part_bytes * world + chunk_bytes - 1) / chunk_bytes;
  int grid = max_ctas > 0 ? max_ctas : 64;
  if (grid > nchunks) grid = (int)nchunks;
  A.part_bytes = part_bytes; A.dst = (uint8_t*)dst; A.chunk_bytes = chunk_bytes;
  A.chunk_flags = (uint32_t*)chunk_flags; A.epoch = epoch; A.flag_value = flag_value ? flag_value : epoch;
  all_gather_kernel<<<grid, kThreads, 0, (cudaStream_t)stream>>>(A);
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}
SXB_API int sxb_symm_all_gather_emu(const void* const* bufs, const void* const* sigs, int world, long long part_bytes,
                                    void* const* dsts, long long chunk_bytes, void* const* chunk_flags,
                                    unsigned int epoch, unsigned int flag_value, int ctas_per_rank, void* stream) {
  if (world < 1 || world > kMaxPeers || ctas_per_rank < 1) return SXB_ERR_BAD_ARG;
  if ((part_bytes & 15) || (chunk_bytes & 15) || chunk_bytes <= 0) return SXB_ERR_UNSUPPORTED;
  AllGatherArgs A[kMaxPeers];
  for (int r = 0; r < world; ++r) {
    int rc = fill_peers(A[r].P, bufs, sigs, r, world);
    if (rc) return rc;
    A[r].part_bytes = part_bytes; A[r].dst = (uint8_t*)dsts[r]; A[r].chunk_bytes = chunk_bytes;
    A[r].chunk_flags = chunk_flags ? (uint32_t*)chunk_flags[r] : nullptr; A[r].epoch = epoch;
    A[r].flag_value = flag_value ? flag_value : epoch;
  }
  return launch_emu((const void*)all_gather_emu_kernel, A, world, ctas_per_rank, kThreads, (cudaStream_t)stream);
}

// send_off / recv_off / count: host arrays of `world` int64 (rows).
static int fill_a2a(A2A& T, const long long* send_off, const long long* recv_off, const long long* count, int world,
                    long long* total_out) {
  long long total = 0;
  for (int i = 0; i < kMaxPeers; ++i) {
    T.send_off[i] = i < world ? send_off[i] : 0;
    T.recv_off[i] = i < world ? recv_off[i] : 0;
    T.count[i] = i < world ? count[i] : 0;
    if (T.count[i] < 0) return SXB_ERR_BAD_ARG;
    total += T.count[i];
  }
  *total_out = total;
  return SXB_OK;
}
SXB_API int sxb_symm_a2a_rows(const void* const* bufs, const void* const* sigs, int rank, int world,
                              const void* src, const long long* send_off, const long long* recv_off,
                              const long long* count, int row_bytes, unsigned int epoch, void* counter,
                              int max_ctas, void* stream) {
  A2AArgs A;
  int rc = fill_peers(A.P, bufs, sigs, rank, world);
  if (rc) return rc;
  if (row_bytes & 15) return SXB_ERR_UNSUPPORTED;
  long long total = 0;
  if ((rc = fill_a2a(A.T, send_off, recv_off, count, world, &total))) return rc;
  long long g = (total + 7) / 8;
  if (g < 1) g = 1;
  const int cap = max_ctas > 0 ? max_ctas : 148;
  if (g > cap) g = cap;
  A.src = (const uint8_t*)src; A.row_bytes = row_bytes; A.epoch = epoch; A.cta_counter = (unsigned int*)counter;
  a2a_rows_kernel<<<(int)g, 256, 0, (cudaStream_t)stream>>>(A);
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}
// Emulated: srcs[r], and send_off/recv_off/count as [world][world] row-major int64 tables (row r = rank r's tables).
SXB_API int sxb_symm_a2a_rows_emu(const void* const* bufs, const void* const* sigs, int world, const void* const* srcs,
                                  const long long* send_off, const long long* recv_off, const long long* count,
                                  int row_bytes, unsigned int epoch, void* const* counters, int ctas_per_rank,
                                  void* stream) {
  if (world < 1 || world > kMaxPeers || ctas_per_rank < 1 || (row_bytes & 15)) return SXB_ERR_BAD_ARG;
  A2AArgs A[kMaxPeers];
  for (int r = 0; r < world; ++r) {
    int rc = fill_peers(A[r].P, bufs, sigs, r, world);
    if (rc) return rc;
    long long total = 0;
    if ((rc = fill_a2a(A[r].T, send_off + r * world, recv_off + r * world, count + r * world, world, &total))) return rc;
    A[r].src = (const uint8_t*)srcs[r]; A[r].row_bytes = row_bytes; A[r].epoch = epoch;
    A[r].cta_counter = (unsigned int*)counters[r];
  }
  return launch_emu((const void*)a2a_rows_emu_kernel, A, world, ctas_per_rank, 256, (cudaStream_t)stream);
}
