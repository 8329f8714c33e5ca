// moe_ep.cu — expert-parallel MoE dispatch / combine as direct NVLink peer stores and loads (K4 / K5).
//
// The reference moves tokens with a dense one-hot einsum + NCCL all_to_all_single on fixed-capacity [E, C, M]
// blocks, and needs a host sync when tokens are not dropped (moe/sharded_moe.py:587-678, 214-225).  Here:
//
//   sxb_moe_ep_plan   one CTA: publish this rank's per-expert token counts to every peer, wait for theirs, and
//                     derive — on the device, no host round trip — (a) for every global expert the row of the owning
//                     rank's receive buffer at which THIS rank's tokens start, (b) the 256-row-aligned group offsets
//                     of the local experts for the grouped GEMM (csrc/gemm_ext.cu).  The handshake also tells the
//                     peers that this rank's receive / return buffers are free again.
//   sxb_moe_ep_push   fused permute + dispatch: token row t is stored straight into the owning rank's receive
//                     buffer at its final (expert-contiguous) row, optionally scaled by its routing weight
//                     (backward of the combine), optionally producing d(score) = <dy, expert output>; also zeroes
//                     the padding rows of the local buffer.  Ends with a "landed" signal to every peer.
//   sxb_moe_ep_pull   fused combine + un-permute: out[t] = sum_j w[t,j] * (row of the expert output living in the
//                     peer's return buffer), read over NVLink and reduced in registers; optionally keeps the raw rows
//                     for the backward of the routing weights.  Ends with a "done reading" signal.
//
// Cross-rank waits between these kernels are one-warp sxb_symm_wait launches (symm_collectives.cu) on the stream,
// so the row movers never spin.  Signals: monotonically increasing epochs on the return buffer's signal pad.
// `*_emu` entry points run `world` logical ranks as one cooperative grid on one GPU (tests).
#include "sxb_common.cuh"

namespace sxb {
namespace moeep {

constexpr int kMaxPeers = 16;
constexpr int kMaxExperts = 256;
constexpr int kAlign = 256;           // group alignment of the grouped GEMM (one CTA-pair tile)
// signal-pad channels (the pad is [channel][kMaxPeers] uint32; channels 0-6 belong to symm_collectives.cu)
constexpr int CH_PLAN = 8, CH_LANDED = 9, CH_READY = 10, CH_PULLED = 11;

struct Peers {
  void* recv[kMaxPeers];      // receive buffers (rows pushed by peers)
  void* ret[kMaxPeers];       // return buffers (rows pulled by peers)
  int* counts[kMaxPeers];     // [world][E] count matrices
  uint32_t* sig[kMaxPeers];
  int rank, world;
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_peer16(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ uint32_t* slot(uint32_t* pad, int channel, int peer) { return pad + channel * kMaxPeers + peer; }
__device__ __forceinline__ void signal_all(const Peers& P, int channel, uint32_t epoch) {
  if (threadIdx.x < P.world) st_release_sys(slot(P.sig[threadIdx.x], channel, P.rank), epoch);
}
__device__ __forceinline__ bool wait_all(const Peers& P, int channel, uint32_t epoch, uint32_t tag) {
  __shared__ int ok_flag;
  if (threadIdx.x == 0) ok_flag = 1;
  __syncthreads();
  if (threadIdx.x < P.world) {
    const uint32_t* s = slot(P.sig[P.rank], channel, threadIdx.x);
    unsigned long long t0 = 0;
    long long c_next = 0;
    while ((int32_t)(ld_acquire_sys(s) - epoch) < 0) {
      if (hang_should_abort(t0, tag, threadIdx.x, &c_next, 30000000000ull)) { ok_flag = 0; break; }
    }
  }
  __syncthreads();
  return ok_flag != 0;
}

// ------------------------------------------------------------------------------------ plan
// tables (device, int32): [0, E)            send_base: row in the owner's receive buffer where my tokens of expert e start
//                         [E, E+L+1)        group_offsets of my local experts (aligned to 256)
//                         [E+L+1, E+2L+1)   group_counts (real rows per local expert)
//                         [E+2L+1]          0, or the rows the fullest receive buffer would need (> cap_rows)
struct PlanArgs {
  Peers P;
  const int* my_counts;   // [E] tokens of this rank per global expert (from the gating kernel)
  int* tables;
  int E, L, cap_rows;
  uint32_t epoch;
  uint32_t pulled_epoch;  // peers must have finished reading my return buffer up to this epoch (0 = nothing to wait for)
  int exchange;           // 0 = handshake only (backward re-uses the forward tables)
};
__device__ __forceinline__ void plan_body(const PlanArgs& A) {
  const Peers& P = A.P;
  const int E = A.E, L = A.L, W = P.world;
  if (A.exchange) {
    // my row of every peer's count matrix
    for (int i = threadIdx.x; i < W * E; i += blockDim.x) {
      const int d = i / E, e = i - d * E;
      P.counts[d][P.rank * E + e] = A.my_counts[e];
    }
    __threadfence_system();
  }
  __syncthreads();
  signal_all(P, CH_PLAN, A.epoch);
  if (!wait_all(P, CH_PLAN, A.epoch, 201)) return;
  if (A.pulled_epoch != 0 && !wait_all(P, CH_PULLED, A.pulled_epoch, 202)) return;
  if (!A.exchange) return;
  __shared__ int tot[kMaxExperts];        // tokens of expert e over all source ranks
  __shared__ int before_me[kMaxExperts];  // ... from source ranks below mine
  __shared__ int goff[kMaxExperts + 1];   // aligned group offsets inside the owner's buffer (per owner, local index)
  const volatile int* M = P.counts[P.rank];
  for (int e = threadIdx.x; e < E; e += blockDim.x) {
    int t = 0, b = 0;
    for (int s = 0; s < W; ++s) {
      const int c = M[s * E + e];
      t += c;
      if (s < P.rank) b += c;
    }
    tot[e] = t;
    before_me[e] = b;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int overflow = 0;
    for (int d = 0; d < W; ++d) {
      int off = 0;
      for (int le = 0; le < L; ++le) {
        const int e = d * L + le;
        goff[e] = off;
        off += (tot[e] + kAlign - 1) / kAlign * kAlign;
      }
      if (off > A.cap_rows && off > overflow) overflow = off;   // rows the fullest rank would need
      if (d == P.rank) {
        for (int le = 0; le < L; ++le) {
          A.tables[E + le] = goff[d * L + le];
          A.tables[E + L + 1 + le] = tot[d * L + le];
        }
        A.tables[E + L] = off < A.cap_rows ? off : (A.cap_rows / kAlign * kAlign);
      }
    }
    A.tables[E + 2 * L + 1] = overflow;
  }
  __syncthreads();
  for (int e = threadIdx.x; e < E; e += blockDim.x) A.tables[e] = goff[e] + before_me[e];
}
__global__ void __launch_bounds__(256) plan_kernel(PlanArgs A) { plan_body(A); }
__global__ void __launch_bounds__(256) plan_emu_kernel(const PlanArgs* per_rank) { plan_body(per_rank[blockIdx.y]); }

// ------------------------------------------------------------------------------------ push
struct PushArgs {
  Peers P;
  const __nv_bfloat16* x;   // [T, H] source rows (tokens, or dy in the backward of the combine)
  const float* weights;     // [T, K] or null: row is scaled by weights[t, j] before it is stored
  const __nv_bfloat16* pulled;  // [T*K, H] or null: with d_scores, d_scores[t,j] = <x[t], pulled[t*K+j]>
  float* d_scores;
  const int* assign;        // [T, K] global expert
  const int* offsets;       // [T, K] arrival index of the token inside its expert (on this rank)
  const int* tables;
  int T, K, H, E, L, cap_rows;
  uint32_t epoch;
  unsigned int* cta_counter;
};
__device__ __forceinline__ void push_body(const PushArgs& A) {
  const Peers& P = A.P;
  const int lane = threadIdx.x & 31;
  const int warps_per_block = blockDim.x >> 5;
  const long long n_rows = (long long)A.T * A.K;
  const long long n_items = n_rows + (long long)A.L * kAlign;
  const int* goff = A.tables + A.E;
  const int* gcnt = A.tables + A.E + A.L + 1;
  const int vec_per_row = A.H >> 3;   // uint4 per row
  for (long long w = (long long)blockIdx.x * warps_per_block + (threadIdx.x >> 5); w < n_items;
       w += (long long)gridDim.x * warps_per_block) {
    if (w >= n_rows) {
      // zero one padding row of MY receive buffer (between the real rows of a local expert and its aligned end)
      const int le = (int)((w - n_rows) / kAlign), r = (int)((w - n_rows) % kAlign);
      const int row = goff[le] + gcnt[le] + r;
      if (row < goff[le + 1] && row < A.cap_rows) {
        uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(P.recv[P.rank]) + (long long)row * A.H);
        for (int v = lane; v < vec_per_row; v += 32) dst[v] = make_uint4(0u, 0u, 0u, 0u);
      }
      continue;
    }
    const int t = (int)(w / A.K);
    const int e = A.assign[w];
    const int row = A.tables[e] + A.offsets[w];
    const int d = e / A.L;
    const float wt = A.weights ? A.weights[w] : 1.f;
    const uint4* src = reinterpret_cast<const uint4*>(A.x + (long long)t * A.H);
    uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(P.recv[d]) + (long long)row * A.H);
    const uint4* pl = A.d_scores ? reinterpret_cast<const uint4*>(A.pulled + w * A.H) : nullptr;
    const bool store = row < A.cap_rows;
    float dot = 0.f;
    for (int v = lane; v < vec_per_row; v += 32) {
      uint4 val = src[v];
      if (A.weights || pl) {
        __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&val);
        if (pl) {
          const uint4 pv = pl[v];
          const __nv_bfloat162* ph = reinterpret_cast<const __nv_bfloat162*>(&pv);
#pragma unroll
          for (int i = 0; i < 4; ++i)
            dot += __low2float(h[i]) * __low2float(ph[i]) + __high2float(h[i]) * __high2float(ph[i]);
        }
        if (A.weights) {
#pragma unroll
          for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(__low2float(h[i]) * wt, __high2float(h[i]) * wt);
        }
      }
      if (store) dst[v] = val;
    }
    if (A.d_scores) {
      dot = warp_sum(dot);
      if (lane == 0) A.d_scores[w] = dot;
    }
  }
  // last CTA: everything this rank sends has been issued and fenced -> tell every peer
  __syncthreads();
  __shared__ unsigned int last;
  if (threadIdx.x == 0) {
    __threadfence_system();
    last = (atomicAdd(A.cta_counter, 1u) == gridDim.x - 1) ? 1u : 0u;
  }
  __syncthreads();
  if (last) {
    if (threadIdx.x == 0) *A.cta_counter = 0u;
    __threadfence_system();
    signal_all(P, CH_LANDED, A.epoch);
  }
}
__global__ void __launch_bounds__(256) push_kernel(PushArgs A) { push_body(A); }
__global__ void __launch_bounds__(256) push_emu_kernel(const PushArgs* per_rank) { push_body(per_rank[blockIdx.y]); }

// ------------------------------------------------------------------------------------ pull
struct PullArgs {
  Peers P;
  __nv_bfloat16* out;       // [T, H]
  const float* weights;     // [T, K] or null (plain sum: backward of the dispatch)
  __nv_bfloat16* keep;      // [T*K, H] or null: raw pulled rows
  const int* assign;
  const int* offsets;
  const int* tables;
  int T, K, H, E, L, cap_rows;
  uint32_t epoch;
  unsigned int* cta_counter;
};
__device__ __forceinline__ void pull_body(const PullArgs& A) {
  const Peers& P = A.P;
  const int lane = threadIdx.x & 31;
  const int warps_per_block = blockDim.x >> 5;
  const int vec_per_row = A.H >> 3;
  constexpr int KMAX = 8;
  for (long long t = (long long)blockIdx.x * warps_per_block + (threadIdx.x >> 5); t < A.T;
       t += (long long)gridDim.x * warps_per_block) {
    const uint4* rows[KMAX];
    float wt[KMAX];
#pragma unroll
    for (int j = 0; j < KMAX; ++j) {
      if (j < A.K) {
        const long long w = t * A.K + j;
        const int e = A.assign[w];
        int row = A.tables[e] + A.offsets[w];
        if (row >= A.cap_rows) row = A.cap_rows - 1;      // overflow is reported by the plan; stay in bounds
        rows[j] = reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(P.ret[e / A.L]) +
                                                 (long long)row * A.H);
        wt[j] = A.weights ? A.weights[w] : 1.f;
      }
    }
    uint4* dst = reinterpret_cast<uint4*>(A.out + t * A.H);
    for (int v = lane; v < vec_per_row; v += 32) {
      uint4 val[KMAX];
#pragma unroll
      for (int j = 0; j < KMAX; ++j)
        if (j < A.K) val[j] = ld_peer16(rows[j] + v);
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int j = 0; j < KMAX; ++j) {
        if (j < A.K) {
          const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&val[j]);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            acc[2 * i] += wt[j] * __low2float(h[i]);
            acc[2 * i + 1] += wt[j] * __high2float(h[i]);
          }
          if (A.keep) reinterpret_cast<uint4*>(A.keep + (t * A.K + j) * A.H)[v] = val[j];
        }
      }
      uint4 o;
      o.x = pack_bf16(acc[0], acc[1]); o.y = pack_bf16(acc[2], acc[3]);
      o.z = pack_bf16(acc[4], acc[5]); o.w = pack_bf16(acc[6], acc[7]);
      dst[v] = o;
    }
  }
  __syncthreads();
  __shared__ unsigned int last;
  if (threadIdx.x == 0) {
    __threadfence();
    last = (atomicAdd(A.cta_counter, 1u) == gridDim.x - 1) ? 1u : 0u;
  }
  __syncthreads();
  if (last) {
    if (threadIdx.x == 0) *A.cta_counter = 0u;
    signal_all(P, CH_PULLED, A.epoch);
  }
}
__global__ void __launch_bounds__(256) pull_kernel(PullArgs A) { pull_body(A); }
__global__ void __launch_bounds__(256) pull_emu_kernel(const PullArgs* per_rank) { pull_body(per_rank[blockIdx.y]); }

// ------------------------------------------------------------------------------------ host helpers
static int fill_peers(Peers& P, const void* const* recv, const void* const* ret, const void* const* counts,
                      const void* const* sigs, int rank, int world) {
  if (world < 1 || world > kMaxPeers || rank < 0 || rank >= world) return SXB_ERR_BAD_ARG;
  for (int i = 0; i < kMaxPeers; ++i) {
    P.recv[i] = i < world && recv ? const_cast<void*>(recv[i]) : nullptr;
    P.ret[i] = i < world && ret ? const_cast<void*>(ret[i]) : nullptr;
    P.counts[i] = i < world && counts ? reinterpret_cast<int*>(const_cast<void*>(counts[i])) : nullptr;
    P.sig[i] = i < world ? reinterpret_cast<uint32_t*>(const_cast<void*>(sigs[i])) : nullptr;
  }
  P.rank = rank;
  P.world = world;
  return SXB_OK;
}

template <typename Args>
static int launch_emu(const void* kernel, const Args* host_args, int world, int ctas, cudaStream_t s) {
  static void* staging = nullptr;
  constexpr size_t kBytes = sizeof(Args) * kMaxPeers;
  if (!staging) {
    cudaError_t e = cudaMalloc(&staging, kBytes);
    if (e != cudaSuccess) return (int)e;
  }
  cudaError_t e = cudaStreamSynchronize(s);   // previous emulated launch may still read the staging area
  if (e != cudaSuccess) return (int)e;
  e = cudaMemcpyAsync(staging, host_args, sizeof(Args) * world, cudaMemcpyHostToDevice, s);
  if (e != cudaSuccess) return (int)e;
  e = cudaStreamSynchronize(s);
  if (e != cudaSuccess) return (int)e;
  const Args* dev_args = reinterpret_cast<const Args*>(staging);
  void* params[] = {(void*)&dev_args};
  e = cudaLaunchCooperativeKernel(kernel, dim3(ctas, world), dim3(256), params, 0, s);
  return e == cudaSuccess ? SXB_OK : (int)e;
}

static int row_grid(long long items, int max_ctas) {
  long long g = (items + 7) / 8;   // 8 warps per CTA, one row per warp
  if (g < 1) g = 1;
  const int cap = max_ctas > 0 ? max_ctas : 148 * 2;
  return (int)(g > cap ? cap : g);
}

}  // namespace moeep
}  // namespace sxb

using namespace sxb;
using namespace sxb::moeep;

SXB_DEFINE_HANG_ACCESSOR(sxb_moe_ep_hang_info)

// All pointer tables are host arrays of `world` device pointers (peer mappings).  tables: device int32
// [E + 2L + 2] (layout above).  E = world * L.  cap_rows: rows of every receive / return buffer.
SXB_API int sxb_moe_ep_plan(const void* const* counts, const void* const* sigs, int rank, int world,
                            const void* my_counts, void* tables, int E, int L, int cap_rows, unsigned int epoch,
                            unsigned int pulled_epoch, int exchange, void* stream) {
  PlanArgs A;
  int rc = fill_peers(A.P, nullptr, nullptr, counts, sigs, rank, world);
  if (rc) return rc;
  if (E != world * L || E > kMaxExperts || L < 1) return SXB_ERR_BAD_ARG;
  A.my_counts = (const int*)my_counts; A.tables = (int*)tables; A.E = E; A.L = L; A.cap_rows = cap_rows;
  A.epoch = epoch; A.pulled_epoch = pulled_epoch; A.exchange = exchange;
  plan_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(A);
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}
SXB_API int sxb_moe_ep_plan_emu(const void* const* counts, const void* const* sigs, int world,
                                const void* const* my_counts, void* const* tables, int E, int L, int cap_rows,
                                unsigned int epoch, unsigned int pulled_epoch, int exchange, void* stream) {
  if (world < 1 || world > kMaxPeers || E != world * L || E > kMaxExperts) return SXB_ERR_BAD_ARG;
  PlanArgs A[kMaxPeers];
  for (int r = 0; r < world; ++r) {
    int rc = fill_peers(A[r].P, nullptr, nullptr, counts, sigs, r, world);
    if (rc) return rc;
    A[r].my_counts = (const int*)my_counts[r]; A[r].tables = (int*)tables[r]; A[r].E = E; A[r].L = L;
    A[r].cap_rows = cap_rows; A[r].epoch = epoch; A[r].pulled_epoch = pulled_epoch; A[r].exchange = exchange;
  }
  return launch_emu((const void*)plan_emu_kernel, A, world, 1, (cudaStream_t)stream);
}

SXB_API int sxb_moe_ep_push(const void* const* recv, const void* const* sigs, int rank, int world, const void* x,
                            const void* weights, const void* pulled, void* d_scores, const void* assign,
                            const void* offsets, const void* tables, int T, int K, int H, int E, int L,
                            int cap_rows, unsigned int epoch, void* counter, int max_ctas, void* stream) {
  PushArgs A;
  int rc = fill_peers(A.P, recv, nullptr, nullptr, sigs, rank, world);
  if (rc) return rc;
  if ((H & 7) || K < 1 || K > 8 || E != world * L) return SXB_ERR_BAD_ARG;
  A.x = (const __nv_bfloat16*)x; A.weights = (const float*)weights; A.pulled = (const __nv_bfloat16*)pulled;
  A.d_scores = (float*)d_scores; A.assign = (const int*)assign; A.offsets = (const int*)offsets;
  A.tables = (const int*)tables; A.T = T; A.K = K; A.H = H; A.E = E; A.L = L; A.cap_rows = cap_rows;
  A.epoch = epoch; A.cta_counter = (unsigned int*)counter;
  push_kernel<<<row_grid((long long)T * K + (long long)L * kAlign, max_ctas), 256, 0, (cudaStream_t)stream>>>(A);
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}
SXB_API int sxb_moe_ep_push_emu(const void* const* recv, const void* const* sigs, int world, const void* const* x,
                                const void* const* weights, const void* const* pulled, void* const* d_scores,
                                const void* const* assign, const void* const* offsets, const void* const* tables,
                                int T, int K, int H, int E, int L, int cap_rows, unsigned int epoch,
                                void* const* counters, int ctas_per_rank, void* stream) {
  if (world < 1 || world > kMaxPeers || (H & 7) || K < 1 || K > 8 || E != world * L) return SXB_ERR_BAD_ARG;
  PushArgs A[kMaxPeers];
  for (int r = 0; r < world; ++r) {
    int rc = fill_peers(A[r].P, recv, nullptr, nullptr, sigs, r, world);
    if (rc) return rc;
    A[r].x = (const __nv_bfloat16*)x[r]; A[r].weights = weights ? (const float*)weights[r] : nullptr;
    A[r].pulled = pulled ? (const __nv_bfloat16*)pulled[r] : nullptr;
    A[r].d_scores = d_scores ? (float*)d_scores[r] : nullptr;
    A[r].assign = (const int*)assign[r]; A[r].offsets = (const int*)offsets[r]; A[r].tables = (const int*)tables[r];
    A[r].T = T; A[r].K = K; A[r].H = H; A[r].E = E; A[r].L = L; A[r].cap_rows = cap_rows; A[r].epoch = epoch;
    A[r].cta_counter = (unsigned int*)counters[r];
  }
  return launch_emu((const void*)push_emu_kernel, A, world, ctas_per_rank, (cudaStream_t)stream);
}

SXB_API int sxb_moe_ep_pull(const void* const* ret, const void* const* sigs, int rank, int world, void* out,
                            const void* weights, void* keep, const void* assign, const void* offsets,
                            const void* tables, int T, int K, int H, int E, int L, int cap_rows, unsigned int epoch,
                            void* counter, int max_ctas, void* stream) {
  PullArgs A;
  int rc = fill_peers(A.P, nullptr, ret, nullptr, sigs, rank, world);
  if (rc) return rc;
  if ((H & 7) || K < 1 || K > 8 || E != world * L) return SXB_ERR_BAD_ARG;
  A.out = (__nv_bfloat16*)out; A.weights = (const float*)weights; A.keep = (__nv_bfloat16*)keep;
  A.assign = (const int*)assign; A.offsets = (const int*)offsets; A.tables = (const int*)tables;
  A.T = T; A.K = K; A.H = H; A.E = E; A.L = L; A.cap_rows = cap_rows; A.epoch = epoch;
  A.cta_counter = (unsigned int*)counter;
  pull_kernel<<<row_grid(T, max_ctas), 256, 0, (cudaStream_t)stream>>>(A);
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}
SXB_API int sxb_moe_ep_pull_emu(const void* const* ret, const void* const* sigs, int world, void* const* out,
                                const void* const* weights, void* const* keep, const void* const* assign,
                                const void* const* offsets, const void* const* tables, int T, int K, int H, int E,
                                int L, int cap_rows, unsigned int epoch, void* const* counters, int ctas_per_rank,
                                void* stream) {
  if (world < 1 || world > kMaxPeers || (H & 7) || K < 1 || K > 8 || E != world * L) return SXB_ERR_BAD_ARG;
  PullArgs A[kMaxPeers];
  for (int r = 0; r < world; ++r) {
    int rc = fill_peers(A[r].P, nullptr, ret, nullptr, sigs, r, world);
    if (rc) return rc;
    A[r].out = (__nv_bfloat16*)out[r]; A[r].weights = weights ? (const float*)weights[r] : nullptr;
    A[r].keep = keep ? (__nv_bfloat16*)keep[r] : nullptr; A[r].assign = (const int*)assign[r];
    A[r].offsets = (const int*)offsets[r]; A[r].tables = (const int*)tables[r];
    A[r].T = T; A[r].K = K; A[r].H = H; A[r].E = E; A[r].L = L; A[r].cap_rows = cap_rows; A[r].epoch = epoch;
    A[r].cta_counter = (unsigned int*)counters[r];
  }
  return launch_emu((const void*)pull_emu_kernel, A, world, ctas_per_rank, (cudaStream_t)stream);
}
