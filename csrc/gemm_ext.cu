// gemm_ext.cu — the CTA-pair tcgen05 GEMM of gemm_tcgen05.cu with the three extensions the fused paths need:
//
//   * GATED B operand (K2, ZeRO-3 all-gather consumed by the GEMM): the weight matrix lives in a buffer that a
//     gather (copy engines or the pull kernel of symm_collectives.cu) is still filling; the TMA producer polls the
//     gather's per-chunk arrival flags before it loads a B tile, so the GEMM runs on the weight tiles that have
//     arrived instead of waiting for the whole unit  (reference: all_gather_into_tensor followed by F.linear,
//     runtime/zero/partition_parameters.py:1221-1261).  The GEMM itself only reads THIS GPU's memory.
//   * GROUPED problems (K4/K5, MoE experts) in ONE launch over device-side offsets, no host sync:
//       ragged-M  C[rows_g, N] = A[rows_g, K] op(B_g)          forward (NT) and dgrad (NN) of every local expert
//       ragged-K  C_g[M, N]   (+)= A[rows_g, M]^T B[rows_g, N]  wgrad (TN) of every local expert
//     (reference: python loop over experts moe/experts.py:28-38; inference analogue
//     inference/v2/kernels/cutlass_ops/moe_gemm/moe_gemm.cu:26-175 — an external sm80 CUTLASS grouped GEMM).
//     Row offsets are multiples of 256 (ragged-M) / 64 (ragged-K); padding rows are zero by construction of the
//     dispatch kernel, so they contribute nothing.
//   * FP8 operands (K14): e4m3 x e4m3 -> fp32 in TMEM through tcgen05.mma kind::f8f6f4 (K = 32 per instruction),
//     per-row scale of A and per-output-channel scale of B applied by the epilogue
//     (reference: ops/fp_quantizer/fp8_gemm_triton.py:17-165 dequantises to bf16 before the dot).
//
// Same warp roles as the base kernel: warp 0 TMA producer, warp 1 MMA issuer (leader CTA), warps 2-5 epilogue.
#include "sxb_common.cuh"

namespace sxb {
namespace gemmx {

constexpr int kThreads = 192;
constexpr int kStages = 6;
constexpr int kStageC = 128 * 64 * 2;  // one 128x64 bf16 staging tile
constexpr int BN = 256;

enum Layout : int { NT = 0, NN = 1, TN = 2 };
enum GroupMode : int { kPlain = 0, kRaggedM = 1, kRaggedK = 2 };

struct Plan {
  static constexpr int kA = 128 * 128;  // 128 rows x 128 bytes (64 bf16 / 128 fp8 along K)
  static constexpr int kB = 128 * 128;
  static constexpr int kStage = kA + kB;
  static constexpr int kRing = kStage * kStages;
  static constexpr int kCOff = kRing;
  static constexpr int kBarOff = kCOff + 2 * kStageC;
  static constexpr int kTotal = kBarOff + 256 + 1024;
};

struct XParams {
  int M, N, K;  // per-group problem (ragged-M: M unused, ragged-K: K unused)
  int layout, c_dtype, accumulate, direct;
  void* C;
  int ldc;
  float alpha;
  // ---- gating of the B operand
  const uint32_t* gate_flags;
  uint32_t gate_value;
  long long gate_base;   // byte offset of B's first element inside the flagged buffer
  long long gate_chunk;  // bytes per flag
  long long gate_nchunks;
  long long ldb_bytes;
  int b_rows;  // rows of the stored B matrix
  // ---- groups
  int group_mode, num_groups;
  const int* offsets;  // device, num_groups + 1 row offsets
  int b_group_rows;    // rows of one group's matrix inside the stacked B tensor map
  long long c_group_stride;  // ragged-K: elements between the C matrices of consecutive groups
  // ---- fp8 scales (may be null)
  const float* scale_a;
  const float* scale_b;
  int scale_b_group_stride;
  // ---- column scatter to peers (Ulysses K6): 64-column block j of the result goes to peer scatter_table[2j] at
  // column scatter_table[2j+1] of that peer's buffer, row scatter_row0 + (row of the result), pitch scatter_ld
  const int* scatter_table;
  void* scatter_peers[16];
  long long scatter_ld;
  long long scatter_row0;
};

struct Tile {
  int g;         // group
  int m0, n0;    // tile origin (ragged-M: m0 is a global row of A/C; otherwise local to the group)
  int num_kb;
  int k_row0;    // ragged-K: first row of the group in A/B
};

// Maps the persistent-loop index t to a tile.  Every role of the CTA pair walks the same sequence.
template <int BKE>
struct Sched {
  const XParams& p;
  int tiles_n, tiles_m_plain, per_group;
  int g, cum_start, cum_end;  // ragged-M cursor
  __device__ Sched(const XParams& p_) : p(p_) {
    tiles_n = (p.N + BN - 1) / BN;
    tiles_m_plain = (p.M + 255) / 256;
    per_group = tiles_m_plain * tiles_n;
    g = -1;
    cum_start = 0;
    cum_end = 0;
  }
  __device__ __forceinline__ void swizzle(int r, int tiles_m, int& tm, int& tn) const {
    constexpr int GROUP = 8;
    const int group_sz = GROUP * tiles_n;
    const int gi = r / group_sz;
    const int first_m = gi * GROUP;
    const int gm = min(tiles_m - first_m, GROUP);
    const int rr = r - gi * group_sz;
    tm = first_m + (rr % gm);
    tn = rr / gm;
  }
  __device__ bool get(int t, Tile& o) {
    int tm, tn;
    if (p.group_mode == kRaggedM) {
      while (t >= cum_end) {
        ++g;
        if (g >= p.num_groups) return false;
        cum_start = cum_end;
        cum_end += ((p.offsets[g + 1] - p.offsets[g] + 255) / 256) * tiles_n;
      }
      const int tiles_m = (cum_end - cum_start) / tiles_n;
      swizzle(t - cum_start, tiles_m, tm, tn);
      o.g = g;
      o.m0 = p.offsets[g] + tm * 256;
      o.n0 = tn * BN;
      o.num_kb = (p.K + BKE - 1) / BKE;
      o.k_row0 = 0;
      return true;
    }
    if (p.group_mode == kRaggedK) {
      const int gi = t / per_group;
      if (gi >= p.num_groups) return false;
      swizzle(t - gi * per_group, tiles_m_plain, tm, tn);
      o.g = gi;
      o.m0 = tm * 256;
      o.n0 = tn * BN;
      o.k_row0 = p.offsets[gi];
      o.num_kb = (p.offsets[gi + 1] - p.offsets[gi] + BKE - 1) / BKE;
      return true;
    }
    if (t >= per_group) return false;
    swizzle(t, tiles_m_plain, tm, tn);
    o.g = 0;
    o.m0 = tm * 256;
    o.n0 = tn * BN;
    o.num_kb = (p.K + BKE - 1) / BKE;
    o.k_row0 = 0;
    return true;
  }
};

__device__ __forceinline__ uint32_t ld_acquire_sys_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// Whole-warp wait until flags[c] reached `value` for every chunk of [c_lo, c_hi]; [known_lo, known_hi] caches
// what this warp has already seen.  The data the flags cover was written through the generic proxy (pull kernel)
// or by a copy engine; the proxy fence orders the following TMA (async proxy) reads after the acquire.
__device__ __forceinline__ bool gate_wait(const XParams& p, int c_lo, int c_hi, int& known_lo, int& known_hi) {
  for (int c = c_lo; c <= c_hi; ++c) {
    if (c >= known_lo && c <= known_hi) continue;
    unsigned long long t0 = 0;
    long long c_next = 0;
    while ((int32_t)(ld_acquire_sys_u32(p.gate_flags + c) - p.gate_value) < 0) {
      if (hang_should_abort(t0, 31, (uint32_t)c, &c_next, 30000000000ull)) return false;
      __nanosleep(100);
    }
  }
  if (known_hi < known_lo || c_lo > known_hi + 1 || c_hi < known_lo - 1) {
    known_lo = c_lo;
    known_hi = c_hi;
  } else {
    known_lo = min(known_lo, c_lo);
    known_hi = max(known_hi, c_hi);
  }
  asm volatile("fence.proxy.async.global;" ::: "memory");
  return true;
}
// chunk index of a byte offset: 32-bit arithmetic on 256-byte units (a 64-bit division per k-block costs the producer
// more than the k-block's MMA time: measured 2x slowdown of the gated NN GEMM)
__device__ __forceinline__ int gate_chunk_of(long long byte_off, uint32_t chunk256) {
  return (int)((uint32_t)(byte_off >> 8) / chunk256);
}

__device__ __forceinline__ void umma_f8_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t"
      "}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}

template <bool FP8>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
gemm_pair_x_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                   const __grid_constant__ CUtensorMap tmC, const XParams p) {
  constexpr int STAGES = kStages;
  constexpr int BKE = FP8 ? 128 : 64;  // K elements per 128-byte smem row
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Plan::kBarOff);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull_bar = empty_bar + STAGES;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta = cluster_ctarank();
  const bool leader = (cta == 0);
  const bool a_mn = (p.layout == TN);
  const bool b_mn = (p.layout != NT);
  const int pair_id = blockIdx.x >> 1;
  const int num_pairs = gridDim.x >> 1;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmC);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < STAGES; ++s) {
        mbar_init(&full_bar[s], 1);
        mbar_init(&empty_bar[s], 1);
      }
      for (int s = 0; s < 2; ++s) {
        mbar_init(&tfull_bar[s], 1);
        mbar_init(&tempty_bar[s], 8);
      }
      mbar_fence_init();
    }
    __syncwarp();
    tmem_alloc_pair(tmem_slot, 512);
    tmem_relinquish_pair();
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t sbase = __shfl_sync(0xffffffffu, smem_u32(smem), 0);
  const uint32_t tbase = __shfl_sync(0xffffffffu, tmem_base, 0);

  if (warp == 0) {
    // ------------------------------------------------- TMA producer (both CTAs, own halves)
    const bool elected = elect_one();
    int stage = 0;
    uint32_t phase = 0;
    int known_lo = 1, known_hi = 0;
    bool gate_on = (p.gate_flags != nullptr);
    const uint32_t chunk256 = gate_on ? (uint32_t)(p.gate_chunk >> 8) : 1u;
    const int c_first = gate_on ? gate_chunk_of(p.gate_base, chunk256) : 0;
    const int c_last = gate_on ? min(gate_chunk_of(p.gate_base + (long long)p.b_rows * p.ldb_bytes - 1, chunk256),
                                     (int)p.gate_nchunks - 1) : 0;
    Sched<BKE> sched(p);
    Tile ti;
    for (int t = pair_id; sched.get(t, ti); t += num_pairs) {
      const int m0 = ti.m0 + int(cta) * 128;   // this CTA's 128 rows of A
      const int nb = ti.n0 + int(cta) * 128;   // this CTA's 128 rows/columns of B
      const int b_goff = (p.group_mode == kRaggedM) ? ti.g * p.b_group_rows : 0;
      for (int kb = 0; kb < ti.num_kb; ++kb) {
        const int k0 = kb * BKE;
        if (gate_on && (b_mn || kb == 0)) {
          // rows of the stored B matrix this load (b_mn) or this whole tile (K-major) touches
          const long long r_lo = b_mn ? (long long)k0 : (long long)nb;
          long long r_hi = b_mn ? r_lo + BKE : r_lo + 128;
          if (r_hi > p.b_rows) r_hi = p.b_rows;
          if (r_hi > r_lo) {
            const int c_lo = gate_chunk_of(p.gate_base + r_lo * p.ldb_bytes, chunk256);
            int c_hi = min(gate_chunk_of(p.gate_base + r_hi * p.ldb_bytes - 1, chunk256), c_last);
            if (!(c_lo >= known_lo && c_hi <= known_hi)) {
              // not seen yet: wait for the flags of the next 16 k-blocks in one go, so the cross-proxy fence behind
              // a successful wait is paid once per 16 loads
              if (b_mn) {
                long long r_far = r_lo + 16 * BKE;
                if (r_far > p.b_rows) r_far = p.b_rows;
                c_hi = max(c_hi, min(gate_chunk_of(p.gate_base + r_far * p.ldb_bytes - 1, chunk256), c_last));
              }
              if (!gate_wait(p, c_lo, c_hi, known_lo, known_hi)) goto done;
              if (known_lo <= c_first && known_hi >= c_last) gate_on = false;   // the whole matrix has arrived
            }
          }
        }
        const bool ok = mbar_wait_to(&empty_bar[stage], phase ^ 1, 41, kb);
        if (!__all_sync(0xffffffffu, ok)) goto done;
        if (elected) {
          uint8_t* sa = smem + stage * Plan::kStage;
          uint8_t* sb = sa + Plan::kA;
          if (leader) mbar_expect_tx(&full_bar[stage], 2 * Plan::kStage);
          if (!a_mn) {
            tma_load_2d_pair(sa, &tmA, &full_bar[stage], k0, m0);
          } else {
            const int kr = ti.k_row0 + k0;
            tma_load_2d_pair(sa, &tmA, &full_bar[stage], m0, kr);
            tma_load_2d_pair(sa + 64 * 128, &tmA, &full_bar[stage], m0 + 64, kr);
          }
          if (!b_mn) {
            tma_load_2d_pair(sb, &tmB, &full_bar[stage], k0, b_goff + nb);
          } else {
            const int kr = (p.layout == TN ? ti.k_row0 : b_goff) + k0;
            tma_load_2d_pair(sb, &tmB, &full_bar[stage], nb, kr);
            tma_load_2d_pair(sb + 64 * 128, &tmB, &full_bar[stage], nb + 64, kr);
          }
        }
        __syncwarp();
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------- MMA issuer (leader CTA only)
    if (leader) {
      const bool elected = elect_one();
      const uint32_t idesc = FP8 ? make_idesc(256, BN, 0, 0, 0, 0)
                                 : make_idesc(256, BN, 1, 1, a_mn ? 1u : 0u, b_mn ? 1u : 0u);
      constexpr uint32_t kDescHi = (1024u >> 4) | (1u << 14) | (2u << 29);
      const uint32_t a_lo0 = ((sbase >> 4) & 0x3FFFu) | ((a_mn ? (64u * 128) >> 4 : 1u) << 16);
      const uint32_t b_lo0 = (((sbase + Plan::kA) >> 4) & 0x3FFFu) | ((b_mn ? (64u * 128) >> 4 : 1u) << 16);
      const uint32_t a_kstep = a_mn ? (2048u >> 4) : (32u >> 4);
      const uint32_t b_kstep = b_mn ? (2048u >> 4) : (32u >> 4);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      Sched<BKE> sched(p);
      Tile ti;
      for (int t = pair_id; sched.get(t, ti); t += num_pairs) {
        if (ti.num_kb == 0) continue;   // empty group: the epilogue handles it without an accumulator
        const bool ok0 = mbar_wait_to(&tempty_bar[acc], acc_phase ^ 1, 42, t);
        if (!__all_sync(0xffffffffu, ok0)) goto done;
        tc_fence_after();
        const uint32_t d_tmem = tbase + acc * BN;
        for (int kb = 0; kb < ti.num_kb; ++kb) {
          const bool ok = mbar_wait_to(&full_bar[stage], phase, 43, kb);
          if (!__all_sync(0xffffffffu, ok)) goto done;
          tc_fence_after();
          const uint32_t a_lo = a_lo0 + stage * (Plan::kStage >> 4);
          const uint32_t b_lo = b_lo0 + stage * (Plan::kStage >> 4);
          if (elected) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {   // 4 x 32 bytes of K per stage: 16 bf16 or 32 fp8 per instruction
              const uint64_t adesc = (uint64_t(kDescHi) << 32) | (a_lo + k * a_kstep);
              const uint64_t bdesc = (uint64_t(kDescHi) << 32) | (b_lo + k * b_kstep);
              if (FP8) umma_f8_pair(d_tmem, adesc, bdesc, idesc, (kb | k) != 0);
              else umma_f16_pair(d_tmem, adesc, bdesc, idesc, (kb | k) != 0);
            }
            umma_commit_pair(&empty_bar[stage], 3);
          }
          __syncwarp();
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        if (elected) umma_commit_pair(&tfull_bar[acc], 3);
        __syncwarp();
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
  } else {
    // ---------------------------------------------------------------- epilogue (both CTAs)
    const int et = threadIdx.x - 64;
    const int quad = warp & 3;
    const int row = quad * 32 + lane;
    const bool use_tma = !p.direct;
    uint8_t* stage_c = smem + Plan::kCOff;
    int acc = 0;
    uint32_t acc_phase = 0;
    int buf = 0;
    Sched<BKE> sched(p);
    Tile ti;
    for (int t = pair_id; sched.get(t, ti); t += num_pairs) {
      const int m0 = ti.m0 + int(cta) * 128;
      const int n0 = ti.n0;
      // row bound of this tile's matrix and where its C lives
      int m_end;
      if (p.group_mode == kRaggedM) m_end = p.offsets[ti.g + 1];
      else m_end = p.M;
      const long long c_off = (p.group_mode == kRaggedK) ? (long long)ti.g * p.c_group_stride : 0;
      const int c_row_tma = (p.group_mode == kRaggedK) ? ti.g * p.M + m0 : m0;
      const int gr = m0 + row;
      if (ti.num_kb == 0) {
        // empty group (ragged-K): C_g = 0 unless accumulating
        if (!p.accumulate && gr < m_end) {
          for (int gc = n0; gc < min(n0 + BN, p.N); ++gc) {
            if (p.c_dtype == kBF16)
              reinterpret_cast<__nv_bfloat16*>(p.C)[c_off + (size_t)gr * p.ldc + gc] = __float2bfloat16_rn(0.f);
            else
              reinterpret_cast<float*>(p.C)[c_off + (size_t)gr * p.ldc + gc] = 0.f;
          }
        }
        continue;
      }
      if (!mbar_wait_to(&tfull_bar[acc], acc_phase, 44, t)) goto done;
      tc_fence_after();
      float sa = 1.f;
      if (FP8 && p.scale_a != nullptr && gr < m_end) sa = p.scale_a[gr];
      const float* sbp = (FP8 && p.scale_b != nullptr) ? p.scale_b + (long long)ti.g * p.scale_b_group_stride : nullptr;
      const bool half_valid = (m0 < m_end);   // this CTA's 128 rows lie (at least partly) inside the matrix
      const uint32_t t_row = tmem_base + acc * BN + (uint32_t(quad * 32) << 16);
#pragma unroll 1
      for (int c = 0; c < BN / 64; ++c) {
        uint32_t v0[32], v1[32];
        tmem_ld_32x32(t_row + c * 64, v0);
        tmem_ld_32x32(t_row + c * 64 + 32, v1);
        tmem_ld_wait();
        if (c == BN / 64 - 1) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_leader(&tempty_bar[acc]);
        }
        const int gc0 = n0 + c * 64;
        if (FP8) {
#pragma unroll
          for (int j = 0; j < 64; ++j) {
            const int gc = gc0 + j;
            const float sb = (sbp != nullptr && gc < p.N) ? __ldg(sbp + gc) : 1.f;
            uint32_t& w = (j < 32) ? v0[j] : v1[j - 32];
            w = __float_as_uint(__uint_as_float(w) * sa * sb);
          }
        }
        if (use_tma) {
          if (et == 0) tma_store_wait_read<1>();
          asm volatile("bar.sync 1, 128;" ::: "memory");
          uint8_t* dst = stage_c + buf * kStageC + row * 128;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const uint32_t* s = (j < 4) ? &v0[j * 8] : &v1[(j - 4) * 8];
            uint4 o;
            o.x = pack_bf16(__uint_as_float(s[0]), __uint_as_float(s[1]));
            o.y = pack_bf16(__uint_as_float(s[2]), __uint_as_float(s[3]));
            o.z = pack_bf16(__uint_as_float(s[4]), __uint_as_float(s[5]));
            o.w = pack_bf16(__uint_as_float(s[6]), __uint_as_float(s[7]));
            *reinterpret_cast<uint4*>(dst + ((j ^ (row & 7)) << 4)) = o;
          }
          fence_async_smem();
          asm volatile("bar.sync 1, 128;" ::: "memory");
          if (et == 0) {
            if (half_valid && gc0 < p.N) tma_store_2d(&tmC, stage_c + buf * kStageC, gc0, c_row_tma);
            tma_store_commit();
          }
          buf ^= 1;
        } else if (p.scatter_table != nullptr) {
          // direct peer stores: this thread's 64 values are 128 contiguous bytes of one destination row
          if (gr < m_end && gc0 < p.N) {
            const int owner = p.scatter_table[2 * (gc0 >> 6)];
            const int dcol = p.scatter_table[2 * (gc0 >> 6) + 1];
            uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.scatter_peers[owner]) +
                                                  (p.scatter_row0 + gr) * p.scatter_ld + dcol);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const uint32_t* sv = (j < 4) ? &v0[j * 8] : &v1[(j - 4) * 8];
              uint4 o;
              o.x = pack_bf16(p.alpha * __uint_as_float(sv[0]), p.alpha * __uint_as_float(sv[1]));
              o.y = pack_bf16(p.alpha * __uint_as_float(sv[2]), p.alpha * __uint_as_float(sv[3]));
              o.z = pack_bf16(p.alpha * __uint_as_float(sv[4]), p.alpha * __uint_as_float(sv[5]));
              o.w = pack_bf16(p.alpha * __uint_as_float(sv[6]), p.alpha * __uint_as_float(sv[7]));
              dst[j] = o;
            }
          }
        } else if (gr < m_end) {
          if (p.c_dtype == kBF16) {
            __nv_bfloat16* crow = reinterpret_cast<__nv_bfloat16*>(p.C) + c_off + (size_t)gr * p.ldc;
#pragma unroll
            for (int j = 0; j < 64; ++j) {
              const int gc = gc0 + j;
              if (gc < p.N) {
                float r = p.alpha * __uint_as_float(j < 32 ? v0[j] : v1[j - 32]);
                if (p.accumulate) r += __bfloat162float(crow[gc]);
                crow[gc] = __float2bfloat16_rn(r);
              }
            }
          } else {
            float* crow = reinterpret_cast<float*>(p.C) + c_off + (size_t)gr * p.ldc;
            if (gc0 + 64 <= p.N && (p.ldc & 3) == 0 && (c_off & 3) == 0) {
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                const uint32_t* s = (j < 8) ? &v0[j * 4] : &v1[(j - 8) * 4];
                float4 r = make_float4(p.alpha * __uint_as_float(s[0]), p.alpha * __uint_as_float(s[1]),
                                       p.alpha * __uint_as_float(s[2]), p.alpha * __uint_as_float(s[3]));
                float4* dp = reinterpret_cast<float4*>(crow + gc0 + j * 4);
                if (p.accumulate) {
                  float4 o = *dp;
                  r.x += o.x; r.y += o.y; r.z += o.z; r.w += o.w;
                }
                *dp = r;
              }
            } else {
#pragma unroll
              for (int j = 0; j < 64; ++j) {
                const int gc = gc0 + j;
                if (gc < p.N) {
                  float r = p.alpha * __uint_as_float(j < 32 ? v0[j] : v1[j - 32]);
                  if (p.accumulate) r += crow[gc];
                  crow[gc] = r;
                }
              }
            }
          }
        }
      }
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
    if (use_tma && et == 0) tma_store_wait<0>();
  }

done:
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base, 512);
  }
}

// box = [box_rows, 128 bytes]; esize 2 (bf16) or 1 (fp8 as uint8)
static int make_map(CUtensorMap* m, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
                    int esize) {
  const unsigned long long dims[2] = {cols, rows};
  const unsigned long long strides[1] = {ld * (unsigned long long)esize};
  const unsigned int box[2] = {(unsigned int)(128 / esize), box_rows};
  return encode_tensor_map(m, esize == 2 ? kBF16 : kU8, 2, ptr, dims, strides, box, 128);
}

static int num_sms() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

}  // namespace gemmx
}  // namespace sxb

using namespace sxb;
using namespace sxb::gemmx;

SXB_DEFINE_HANG_ACCESSOR(sxb_gemmx_hang_info)

// C (+)= alpha * op(A) op(B) with optional gating / grouping / fp8 operands (see the file header).
//   ab_dtype: kBF16 (2) or kFP8E4M3 (3; NT layout only).  c_dtype: kBF16 or kF32.
//   gate_flags == NULL: no gating.  Else flags[c] >= gate_value (as a wrapping 32-bit sequence number) means bytes
//     [c*gate_chunk, (c+1)*gate_chunk) of the buffer B lives in have arrived; gate_base = byte offset of B in it.
//   group_mode 0: plain.  1 (ragged-M, NT/NN): offsets[num_groups+1] (device, rows, multiples of 256); A and C have
//     total_rows rows; B is the row-wise stack of the groups' matrices (NT: num_groups*N rows; NN: num_groups*K).
//     2 (ragged-K, TN): offsets multiples of 64; A [total_rows, M], B [total_rows, N]; C = num_groups matrices
//     [M, N] spaced c_group_stride elements apart (M multiple of 128).
//   scale_a[row of A] / scale_b[g*N + column]: fp8 dequantisation scales (either may be NULL).
SXB_API int sxb_gemm_ext(const void* A, const void* B, void* C, int M, int N, int K, int layout, long long lda,
                         long long ldb, long long ldc, int ab_dtype, int c_dtype, int accumulate, float alpha,
                         const void* gate_flags, unsigned int gate_value, long long gate_base, long long gate_chunk,
                         long long gate_nchunks, int group_mode, int num_groups, const void* offsets,
                         long long total_rows, long long c_group_stride, const void* scale_a, const void* scale_b,
                         int max_ctas, void* stream) {
  if (layout < 0 || layout > 2) return SXB_ERR_BAD_ARG;
  if (c_dtype != kBF16 && c_dtype != kF32) return SXB_ERR_BAD_ARG;
  if (ab_dtype != kBF16 && ab_dtype != kFP8E4M3) return SXB_ERR_BAD_ARG;
  const bool fp8 = (ab_dtype == kFP8E4M3);
  if (fp8 && layout != NT) return SXB_ERR_UNSUPPORTED;
  const int es = fp8 ? 1 : 2;
  const int align = 16 / es;
  if ((lda % align) || (ldb % align) || (reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15))
    return SXB_ERR_UNSUPPORTED;
  if (group_mode < 0 || group_mode > 2) return SXB_ERR_BAD_ARG;
  if (group_mode != kPlain && (offsets == nullptr || num_groups < 1 || total_rows <= 0)) return SXB_ERR_BAD_ARG;
  if (group_mode == kRaggedM && layout == TN) return SXB_ERR_BAD_ARG;
  if (group_mode == kRaggedK && (layout != TN || (M % 128))) return SXB_ERR_BAD_ARG;
  if (group_mode == kRaggedM && layout == NN && (K % 64)) return SXB_ERR_UNSUPPORTED;
  if (group_mode == kPlain && (M <= 0 || N <= 0 || K <= 0)) return SXB_OK;
  if (N <= 0) return SXB_OK;
  if (gate_flags != nullptr && (gate_chunk <= 0 || (gate_chunk & 255) || gate_nchunks <= 0 || gate_base < 0 ||
                                ((gate_chunk * gate_nchunks) >> 8) > 0xFFFFFFFFll))
    return SXB_ERR_BAD_ARG;   // chunks are multiples of 256 bytes; flagged buffers up to 1 TiB
  const bool tma_c = (c_dtype == kBF16) && !accumulate && alpha == 1.0f && (ldc % 8 == 0) &&
                     ((reinterpret_cast<uintptr_t>(C) & 15) == 0) &&
                     (group_mode != kRaggedK || (c_group_stride == (long long)M * ldc));
  XParams p{};
  p.M = M; p.N = N; p.K = K; p.layout = layout; p.c_dtype = c_dtype; p.accumulate = accumulate ? 1 : 0;
  p.direct = tma_c ? 0 : 1; p.C = C; p.ldc = (int)ldc; p.alpha = alpha;
  p.gate_flags = (const uint32_t*)gate_flags; p.gate_value = gate_value; p.gate_base = gate_base;
  p.gate_chunk = gate_chunk; p.gate_nchunks = gate_nchunks; p.ldb_bytes = ldb * es;
  p.group_mode = group_mode; p.num_groups = num_groups; p.offsets = (const int*)offsets;
  p.c_group_stride = c_group_stride; p.scale_a = (const float*)scale_a; p.scale_b = (const float*)scale_b;
  p.scale_b_group_stride = (group_mode == kRaggedM) ? N : 0;
  // stored shapes of A, B, C
  unsigned long long a_rows, a_cols, b_rows, b_cols, c_rows;
  if (group_mode == kRaggedM) {
    a_rows = total_rows; a_cols = K;
    if (layout == NT) { b_rows = (unsigned long long)num_groups * N; b_cols = K; p.b_group_rows = N; }
    else { b_rows = (unsigned long long)num_groups * K; b_cols = N; p.b_group_rows = K; }
    c_rows = total_rows;
  } else if (group_mode == kRaggedK) {
    a_rows = total_rows; a_cols = M; b_rows = total_rows; b_cols = N; c_rows = (unsigned long long)num_groups * M;
  } else {
    if (layout == TN) { a_rows = K; a_cols = M; } else { a_rows = M; a_cols = K; }
    if (layout == NT) { b_rows = N; b_cols = K; } else { b_rows = K; b_cols = N; }
    c_rows = M;
  }
  p.b_rows = (int)((group_mode == kRaggedM) ? (unsigned long long)p.b_group_rows : b_rows);
  if (gate_flags != nullptr && group_mode != kPlain) return SXB_ERR_UNSUPPORTED;
  CUtensorMap ta, tb, tc;
  int rc;
  if ((rc = make_map(&ta, A, a_rows, a_cols, lda, (layout == TN) ? 64 : 128, es))) return rc;
  if ((rc = make_map(&tb, B, b_rows, b_cols, ldb, (layout == NT) ? 128 : 64, es))) return rc;
  if (tma_c) {
    if ((rc = make_map(&tc, C, c_rows, N, ldc, 128, 2))) return rc;
  } else {
    tc = ta;
  }
  static bool configured[2] = {false, false};
  const void* kfn = fp8 ? (const void*)gemm_pair_x_kernel<true> : (const void*)gemm_pair_x_kernel<false>;
  if (!configured[fp8 ? 1 : 0]) {
    cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, Plan::kTotal);
    if (e != cudaSuccess) return (int)e;
    configured[fp8 ? 1 : 0] = true;
  }
  int grid = num_sms() & ~1;
  if (max_ctas > 0 && max_ctas < grid) grid = max_ctas & ~1;
  if (group_mode == kPlain) {
    const long long tiles = (long long)((M + 255) / 256) * ((N + 255) / 256);
    if (2 * tiles < grid) grid = (int)(2 * tiles);
  }
  if (grid < 2) grid = 2;
  cudaStream_t s = (cudaStream_t)stream;
  if (fp8) gemm_pair_x_kernel<true><<<grid, kThreads, Plan::kTotal, s>>>(ta, tb, tc, p);
  else gemm_pair_x_kernel<false><<<grid, kThreads, Plan::kTotal, s>>>(ta, tb, tc, p);
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

// K6 (Ulysses): C = A op(B) whose epilogue stores every 64-column block of the result straight into the buffer of the
// rank that owns those heads (direct peer stores over NVLink).  table: device int32 [2 * N/64] = (owner, destination
// column) per block; peers: host array of `world` device pointers; the result row r lands in row row0 + r of the
// destination (pitch dst_ld elements).  bf16 only, plain (ungrouped) problem, N % 64 == 0.
// (reference: QKV projection, then three all_to_all_single, sequence/layer.py:388-426)
SXB_API int sxb_gemm_scatter(const void* A, const void* B, int M, int N, int K, int layout, long long lda,
                             long long ldb, const void* table, void* const* peers, int world, long long dst_ld,
                             long long row0, float alpha, int max_ctas, void* stream) {
  if (layout < 0 || layout > 2 || world < 1 || world > 16 || table == nullptr) return SXB_ERR_BAD_ARG;
  if ((N % 64) || (dst_ld % 8) || (lda % 8) || (ldb % 8) || (reinterpret_cast<uintptr_t>(A) & 15) ||
      (reinterpret_cast<uintptr_t>(B) & 15))
    return SXB_ERR_UNSUPPORTED;
  if (M <= 0 || N <= 0 || K <= 0) return SXB_OK;
  XParams p{};
  p.M = M; p.N = N; p.K = K; p.layout = layout; p.c_dtype = kBF16; p.accumulate = 0; p.direct = 1;
  p.C = nullptr; p.ldc = 0; p.alpha = alpha; p.ldb_bytes = ldb * 2;
  p.scatter_table = (const int*)table; p.scatter_ld = dst_ld; p.scatter_row0 = row0;
  for (int i = 0; i < 16; ++i) p.scatter_peers[i] = i < world ? peers[i] : nullptr;
  unsigned long long a_rows, a_cols, b_rows, b_cols;
  if (layout == TN) { a_rows = K; a_cols = M; } else { a_rows = M; a_cols = K; }
  if (layout == NT) { b_rows = N; b_cols = K; } else { b_rows = K; b_cols = N; }
  p.b_rows = (int)b_rows;
  CUtensorMap ta, tb;
  int rc;
  if ((rc = make_map(&ta, A, a_rows, a_cols, lda, (layout == TN) ? 64 : 128, 2))) return rc;
  if ((rc = make_map(&tb, B, b_rows, b_cols, ldb, (layout == NT) ? 128 : 64, 2))) return rc;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(gemm_pair_x_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         Plan::kTotal);
    if (e != cudaSuccess) return (int)e;
    configured = true;
  }
  int grid = num_sms() & ~1;
  if (max_ctas > 0 && max_ctas < grid) grid = max_ctas & ~1;
  const long long tiles = (long long)((M + 255) / 256) * ((N + 255) / 256);
  if (2 * tiles < grid) grid = (int)(2 * tiles);
  if (grid < 2) grid = 2;
  gemm_pair_x_kernel<false><<<grid, kThreads, Plan::kTotal, (cudaStream_t)stream>>>(ta, tb, ta, p);
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

// Stream-ordered 32-bit write issued by the stream front end (no SM, no kernel): used to publish "this piece of the
// gather has landed" behind a copy-engine transfer, for a gated GEMM that is already running on the SMs.
typedef CUresult (*StreamWriteFn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
SXB_API int sxb_stream_write_u32(void* dst, unsigned int value, void* stream) {
  static StreamWriteFn fn = nullptr;
  if (!fn) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuStreamWriteValue32", &ptr, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess || ptr == nullptr)
      return SXB_ERR_DRIVER;
    fn = reinterpret_cast<StreamWriteFn>(ptr);
  }
  if (reinterpret_cast<uintptr_t>(dst) & 3) return SXB_ERR_BAD_ARG;
  CUresult r = fn((CUstream)stream, (CUdeviceptr)dst, value, 0);
  return r == CUDA_SUCCESS ? SXB_OK : SXB_ERR_DRIVER;
}
