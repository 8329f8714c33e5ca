// gemm_tcgen05.cu — persistent, warp-specialised bf16 GEMM for sm_100a.
//
//   C[M,N] (+)= op(A) * op(B)         fp32 accumulation in TMEM
//
//   layout NT : A[M,K] row-major, B[N,K] row-major   (forward  y = x W^T)
//   layout NN : A[M,K] row-major, B[K,N] row-major   (dgrad    dx = dy W)
//   layout TN : A[K,M] row-major, B[K,N] row-major   (wgrad    dW = dy^T x)
//
// Structure (one CTA per SM, 192 threads):
//   warp 0      TMA producer  : cp.async.bulk.tensor -> 128B-swizzled smem ring
//   warp 1      MMA issuer    : one thread issues tcgen05.mma (128 x BN x 16),
//                               accumulators double-buffered in TMEM
//   warps 2..5  epilogue      : tcgen05.ld -> bf16 -> swizzled smem -> TMA store
//                               (or direct global read-modify-write when
//                               accumulating / writing fp32)
// Replaces the cuBLAS calls behind F.linear that the reference relies on
// (reference: deepspeed/inference/v2/kernels/core_ops/blas_kernels/blas_utils.h:31-41,
// and every nn.Linear in the training path).
#include "sxb_common.cuh"

namespace sxb {
namespace gemm {

constexpr int BM = 128;
constexpr int BK = 64;  // 64 bf16 = one 128B swizzle row
constexpr int UMMA_K = 16;
constexpr int kThreads = 192;
constexpr int kEpiThreads = 128;
constexpr int kStageC = 128 * 64 * 2;  // one 128x64 bf16 staging tile (16 KB)

enum Layout : int { NT = 0, NN = 1, TN = 2 };

template <int BN, int STAGES>
struct SmemPlan {
  static constexpr int kA = BM * BK * 2;
  static constexpr int kB = BN * BK * 2;
  static constexpr int kStage = kA + kB;
  static constexpr int kRing = kStage * STAGES;
  static constexpr int kCOff = kRing;
  static constexpr int kBarOff = kCOff + 2 * kStageC;
  static constexpr int kTotal = kBarOff + 256 + 1024;  // + alignment slack
};

struct Params {
  int M, N, K;
  int layout;
  int c_dtype;     // kBF16 or kF32
  int accumulate;  // C += result
  int direct;      // 1 = epilogue writes global memory directly (no TMA store)
  void* C;
  int ldc;
  float alpha;
};

template <int BN, int STAGES>
__global__ void __launch_bounds__(kThreads, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
            const __grid_constant__ CUtensorMap tmC, const Params p) {
  using Plan = SmemPlan<BN, STAGES>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Plan::kBarOff);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull_bar = empty_bar + STAGES;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const bool a_mn = (p.layout == TN);
  const bool b_mn = (p.layout != NT);

  const int tiles_m = (p.M + BM - 1) / BM;
  const int tiles_n = (p.N + BN - 1) / BN;
  const int num_tiles = tiles_m * tiles_n;
  const int num_kb = (p.K + BK - 1) / BK;

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
        mbar_init(&tempty_bar[s], kEpiThreads);
      }
      mbar_fence_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, 2 * BN);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  auto tile_coord = [&](int t, int& tm, int& tn) {
    constexpr int GROUP = 8;
    const int group_sz = GROUP * tiles_n;
    const int g = t / group_sz;
    const int first_m = g * GROUP;
    const int gm = min(tiles_m - first_m, GROUP);
    const int r = t - g * group_sz;
    tm = first_m + (r % gm);
    tn = r / gm;
  };

  // Whole-warp uniform role loops; the async instruction is predicated on one elected lane (see the
  // CTA-pair kernel below for why).
  const uint32_t sbase = __shfl_sync(0xffffffffu, smem_u32(smem), 0);
  const uint32_t tbase = __shfl_sync(0xffffffffu, tmem_base, 0);
  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    const bool elected = elect_one();
    int stage = 0;
    uint32_t phase = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
      int tm, tn;
      tile_coord(t, tm, tn);
      const int m0 = tm * BM, n0 = tn * BN;
      for (int kb = 0; kb < num_kb; ++kb) {
        if (!__all_sync(0xffffffffu, mbar_wait_to(&empty_bar[stage], phase ^ 1, 1, kb))) goto done;
        if (elected) {
          uint8_t* sa = smem + stage * Plan::kStage;
          uint8_t* sb = sa + Plan::kA;
          mbar_expect_tx(&full_bar[stage], Plan::kStage);
          const int k0 = kb * BK;
          if (!a_mn) {
            tma_load_2d(sa, &tmA, &full_bar[stage], k0, m0);
          } else {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j)
              tma_load_2d(sa + j * (64 * BK * 2), &tmA, &full_bar[stage], m0 + j * 64, k0);
          }
          if (!b_mn) {
            tma_load_2d(sb, &tmB, &full_bar[stage], k0, n0);
          } else {
#pragma unroll
            for (int j = 0; j < BN / 64; ++j)
              tma_load_2d(sb + j * (64 * BK * 2), &tmB, &full_bar[stage], n0 + j * 64, k0);
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
    // -------------------------------------------------------------- MMA issuer
    const bool elected = elect_one();
    const uint32_t idesc = make_idesc(BM, BN, 1, 1, a_mn ? 1u : 0u, b_mn ? 1u : 0u);
    constexpr uint32_t kDescHi = (1024u >> 4) | (1u << 14) | (2u << 29);
    const uint32_t a_lo0 = ((sbase >> 4) & 0x3FFFu) | ((a_mn ? (64u * BK * 2) >> 4 : 1u) << 16);
    const uint32_t b_lo0 = (((sbase + Plan::kA) >> 4) & 0x3FFFu) | ((b_mn ? (64u * BK * 2) >> 4 : 1u) << 16);
    const uint32_t a_kstep = a_mn ? (2048u >> 4) : (32u >> 4);
    const uint32_t b_kstep = b_mn ? (2048u >> 4) : (32u >> 4);
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
      if (!__all_sync(0xffffffffu, mbar_wait_to(&tempty_bar[acc], acc_phase ^ 1, 2, t))) goto done;
      tc_fence_after();
      const uint32_t d_tmem = tbase + acc * BN;
      for (int kb = 0; kb < num_kb; ++kb) {
        if (!__all_sync(0xffffffffu, mbar_wait_to(&full_bar[stage], phase, 3, kb))) goto done;
        tc_fence_after();
        const uint32_t a_lo = a_lo0 + stage * (Plan::kStage >> 4);
        const uint32_t b_lo = b_lo0 + stage * (Plan::kStage >> 4);
        if (elected) {
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            const uint64_t adesc = (uint64_t(kDescHi) << 32) | (a_lo + k * a_kstep);
            const uint64_t bdesc = (uint64_t(kDescHi) << 32) | (b_lo + k * b_kstep);
            umma_f16(d_tmem, adesc, bdesc, idesc, (kb | k) != 0);
          }
          umma_commit(&empty_bar[stage]);
        }
        __syncwarp();
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
      if (elected) umma_commit(&tfull_bar[acc]);
      __syncwarp();
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
  } else {
    // ---------------------------------------------------------------- epilogue
    const int et = threadIdx.x - 64;       // 0..127
    const int quad = warp & 3;             // TMEM lane quadrant this warp may read
    const int row = quad * 32 + lane;      // row of the 128-row tile owned by this thread
    const bool use_tma = !p.direct;
    uint8_t* stage_c = smem + Plan::kCOff;
    int acc = 0;
    uint32_t acc_phase = 0;
    int buf = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
      int tm, tn;
      tile_coord(t, tm, tn);
      const int m0 = tm * BM, n0 = tn * BN;
      if (!mbar_wait_to(&tfull_bar[acc], acc_phase, 4, t)) goto done;
      tc_fence_after();
      const uint32_t t_row = tmem_base + acc * BN + (uint32_t(quad * 32) << 16);
#pragma unroll 1
      for (int c = 0; c < BN / 64; ++c) {
        uint32_t v0[32], v1[32];
        tmem_ld_32x32(t_row + c * 64, v0);
        tmem_ld_32x32(t_row + c * 64 + 32, v1);
        tmem_ld_wait();
        if (c == BN / 64 - 1) {
          tc_fence_before();
          mbar_arrive(&tempty_bar[acc]);
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
            tma_store_2d(&tmC, stage_c + buf * kStageC, n0 + c * 64, m0);
            tma_store_commit();
          }
          buf ^= 1;
        } else {
          const int gr = m0 + row;
          if (gr < p.M) {
            const int gc0 = n0 + c * 64;
            if (p.c_dtype == kBF16) {
              __nv_bfloat16* crow = reinterpret_cast<__nv_bfloat16*>(p.C) + (size_t)gr * p.ldc;
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
              float* crow = reinterpret_cast<float*>(p.C) + (size_t)gr * p.ldc;
              if (gc0 + 64 <= p.N && (p.ldc & 3) == 0) {
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                  const uint32_t* s = (j < 8) ? &v0[j * 4] : &v1[(j - 8) * 4];
                  float4 r = make_float4(p.alpha * __uint_as_float(s[0]),
                                         p.alpha * __uint_as_float(s[1]),
                                         p.alpha * __uint_as_float(s[2]),
                                         p.alpha * __uint_as_float(s[3]));
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
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * BN);
  }
}

// ======================================================================= CTA-pair variant
// Two CTAs on one TPC (cluster 2x1x1) own a 256 x 256 output tile.  CTA r stages rows
// [m0 + 128 r, +128) of A and rows [n0 + 128 r, +128) of B (32 KB per stage, 6 stages); the leader
// issues tcgen05.mma.cta_group::2 (M = 256) which reads both CTAs' shared memory, so every byte of
// B is fetched from L2 once per pair instead of once per SM (-33% L2->SM traffic, 2x the prefetch
// depth in time).  Each CTA drains its own 128 TMEM lanes through the same epilogue as above.
constexpr int kPairStages = 6;
struct PairPlan {
  static constexpr int kA = 128 * BK * 2;
  static constexpr int kB = 128 * BK * 2;
  static constexpr int kStage = kA + kB;
  static constexpr int kRing = kStage * kPairStages;
  static constexpr int kCOff = kRing;
  static constexpr int kBarOff = kCOff + 2 * kStageC;
  static constexpr int kTotal = kBarOff + 256 + 1024;
};

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
gemm_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ CUtensorMap tmC, const Params p) {
  using Plan = PairPlan;
  constexpr int STAGES = kPairStages;
  constexpr int BN = 256;
  extern __shared__ uint8_t smem_raw[];
  // identical offset in both CTAs (same kernel, same dynamic smem base) -> peer-bit addressing works
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
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

  const int tiles_m = (p.M + 255) / 256;
  const int tiles_n = (p.N + BN - 1) / BN;
  const int num_tiles = tiles_m * tiles_n;
  const int num_kb = (p.K + BK - 1) / BK;
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
        mbar_init(&full_bar[s], 1);   // leader's producer arrive.expect_tx (bytes of BOTH CTAs)
        mbar_init(&empty_bar[s], 1);  // multicast tcgen05.commit from the leader
      }
      for (int s = 0; s < 2; ++s) {
        mbar_init(&tfull_bar[s], 1);   // multicast commit
        mbar_init(&tempty_bar[s], 8);  // 4 epilogue warps x 2 CTAs arrive on the leader's copy
      }
      mbar_fence_init();
    }
    __syncwarp();
    tmem_alloc_pair(tmem_slot, 512);
    tmem_relinquish_pair();
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();  // peer barriers initialised before any remote arrive / multicast commit
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  auto tile_coord = [&](int t, int& tm, int& tn) {
    constexpr int GROUP = 8;
    const int group_sz = GROUP * tiles_n;
    const int g = t / group_sz;
    const int first_m = g * GROUP;
    const int gm = min(tiles_m - first_m, GROUP);
    const int r = t - g * group_sz;
    tm = first_m + (r % gm);
    tn = r / gm;
  };

  // Producer and MMA roles run with the WHOLE warp in uniform control flow and predicate only the
  // asynchronous instruction on one elected lane: operands then live in uniform registers and the
  // issue path is a handful of instructions per tcgen05.mma (a single-lane region makes the compiler
  // wrap every UTCHMMA/UTMALDG in an R2UR "waterfall" loop, ~175 cycles per MMA -> issue bound).
  const uint32_t sbase = __shfl_sync(0xffffffffu, smem_u32(smem), 0);
  const uint32_t tbase = __shfl_sync(0xffffffffu, tmem_base, 0);
  if (warp == 0) {
    // ------------------------------------------------- TMA producer (both CTAs, own halves)
    const bool elected = elect_one();
    int stage = 0;
    uint32_t phase = 0;
    for (int t = pair_id; t < num_tiles; t += num_pairs) {
      int tm, tn;
      tile_coord(t, tm, tn);
      const int m0 = tm * 256 + int(cta) * 128, n0 = tn * BN + int(cta) * 128;
      for (int kb = 0; kb < num_kb; ++kb) {
        const bool ok = mbar_wait_to(&empty_bar[stage], phase ^ 1, 11, kb);
        if (!__all_sync(0xffffffffu, ok)) goto done;
        if (elected) {
          uint8_t* sa = smem + stage * Plan::kStage;
          uint8_t* sb = sa + Plan::kA;
          if (leader) mbar_expect_tx(&full_bar[stage], 2 * Plan::kStage);
          const int k0 = kb * BK;
          if (!a_mn) {
            tma_load_2d_pair(sa, &tmA, &full_bar[stage], k0, m0);
          } else {
            tma_load_2d_pair(sa, &tmA, &full_bar[stage], m0, k0);
            tma_load_2d_pair(sa + 64 * BK * 2, &tmA, &full_bar[stage], m0 + 64, k0);
          }
          if (!b_mn) {
            tma_load_2d_pair(sb, &tmB, &full_bar[stage], k0, n0);
          } else {
            tma_load_2d_pair(sb, &tmB, &full_bar[stage], n0, k0);
            tma_load_2d_pair(sb + 64 * BK * 2, &tmB, &full_bar[stage], n0 + 64, k0);
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
      const uint32_t idesc = make_idesc(256, BN, 1, 1, a_mn ? 1u : 0u, b_mn ? 1u : 0u);
      // descriptor = {hi: SBO 1024 B | version 1 | 128B swizzle, lo: LBO | (address >> 4)}
      constexpr uint32_t kDescHi = (1024u >> 4) | (1u << 14) | (2u << 29);
      const uint32_t a_lo0 = ((sbase >> 4) & 0x3FFFu) | ((a_mn ? (64u * BK * 2) >> 4 : 1u) << 16);
      const uint32_t b_lo0 = (((sbase + Plan::kA) >> 4) & 0x3FFFu) | ((b_mn ? (64u * BK * 2) >> 4 : 1u) << 16);
      const uint32_t a_kstep = a_mn ? (2048u >> 4) : (32u >> 4);  // 16 K-elements further, in 16 B units
      const uint32_t b_kstep = b_mn ? (2048u >> 4) : (32u >> 4);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int t = pair_id; t < num_tiles; t += num_pairs) {
        const bool ok0 = mbar_wait_to(&tempty_bar[acc], acc_phase ^ 1, 12, t);
        if (!__all_sync(0xffffffffu, ok0)) goto done;
        tc_fence_after();
        const uint32_t d_tmem = tbase + acc * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          const bool ok = mbar_wait_to(&full_bar[stage], phase, 13, kb);
          if (!__all_sync(0xffffffffu, ok)) goto done;
          tc_fence_after();
          const uint32_t a_lo = a_lo0 + stage * (Plan::kStage >> 4);
          const uint32_t b_lo = b_lo0 + stage * (Plan::kStage >> 4);
          if (elected) {
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k) {
              const uint64_t adesc = (uint64_t(kDescHi) << 32) | (a_lo + k * a_kstep);
              const uint64_t bdesc = (uint64_t(kDescHi) << 32) | (b_lo + k * b_kstep);
              umma_f16_pair(d_tmem, adesc, bdesc, idesc, (kb | k) != 0);
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
    for (int t = pair_id; t < num_tiles; t += num_pairs) {
      int tm, tn;
      tile_coord(t, tm, tn);
      const int m0 = tm * 256 + int(cta) * 128, n0 = tn * BN;
      if (!mbar_wait_to(&tfull_bar[acc], acc_phase, 14, t)) goto done;
      tc_fence_after();
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
            tma_store_2d(&tmC, stage_c + buf * kStageC, n0 + c * 64, m0);
            tma_store_commit();
          }
          buf ^= 1;
        } else {
          const int gr = m0 + row;
          if (gr < p.M) {
            const int gc0 = n0 + c * 64;
            if (p.c_dtype == kBF16) {
              __nv_bfloat16* crow = reinterpret_cast<__nv_bfloat16*>(p.C) + (size_t)gr * p.ldc;
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
              float* crow = reinterpret_cast<float*>(p.C) + (size_t)gr * p.ldc;
              if (gc0 + 64 <= p.N && (p.ldc & 3) == 0) {
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                  const uint32_t* s = (j < 8) ? &v0[j * 4] : &v1[(j - 8) * 4];
                  float4 r = make_float4(p.alpha * __uint_as_float(s[0]),
                                         p.alpha * __uint_as_float(s[1]),
                                         p.alpha * __uint_as_float(s[2]),
                                         p.alpha * __uint_as_float(s[3]));
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
  cluster_sync_all();  // no CTA may exit (or free TMEM) while its peer can still touch it
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base, 512);
  }
}

// ------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) ==
            cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
  }
  return fn;
}

// 2-D bf16 tensor map over a row-major [rows, cols] matrix with row pitch `ld` elements;
// box = [box_rows, box_cols] with box_cols * 2 == 128 bytes (128B swizzle).
static int make_map_2d(CUtensorMap* m, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld,
                       uint32_t box_rows, uint32_t box_cols, CUtensorMapDataType dt, int esize) {
  EncodeTiledFn enc = get_encode();
  if (!enc) return SXB_ERR_DRIVER;
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * (uint64_t)esize};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(m, dt, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? SXB_OK : SXB_ERR_DRIVER;
}

static int g_num_sms = 0;
static int num_sms() {
  if (!g_num_sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
    if (g_num_sms <= 0) g_num_sms = 148;
  }
  return g_num_sms;
}

template <int BN, int STAGES>
static int launch(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc,
                  const Params& p, int max_ctas, cudaStream_t stream) {
  using Plan = SmemPlan<BN, STAGES>;
  static bool configured = false;
  auto kfn = gemm_kernel<BN, STAGES>;
  if (!configured) {
    cudaError_t e =
        cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, Plan::kTotal);
    if (e != cudaSuccess) return (int)e;
    configured = true;
  }
  const int tiles = ((p.M + BM - 1) / BM) * ((p.N + BN - 1) / BN);
  int grid = num_sms();
  if (max_ctas > 0 && max_ctas < grid) grid = max_ctas;
  if (tiles < grid) grid = tiles;
  kfn<<<grid, kThreads, Plan::kTotal, stream>>>(ta, tb, tc, p);
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

static int launch_pair(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc,
                       const Params& p, int max_ctas, cudaStream_t stream) {
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(gemm_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         PairPlan::kTotal);
    if (e != cudaSuccess) return (int)e;
    configured = true;
  }
  const int tiles = ((p.M + 255) / 256) * ((p.N + 255) / 256);
  int grid = num_sms() & ~1;
  if (max_ctas > 0 && max_ctas < grid) grid = max_ctas & ~1;
  if (2 * tiles < grid) grid = 2 * tiles;
  if (grid < 2) grid = 2;
  gemm_pair_kernel<<<grid, kThreads, PairPlan::kTotal, stream>>>(ta, tb, tc, p);
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

}  // namespace gemm
}  // namespace sxb

using namespace sxb;
using namespace sxb::gemm;

SXB_DEFINE_HANG_ACCESSOR(sxb_gemm_hang_info)

// C[M,N] (+)= alpha * op(A) op(B).  A/B are bf16; C is bf16 or fp32 (c_dtype).
// lda/ldb/ldc are row pitches in elements of the *stored* matrices.
// max_ctas: 0 = all SMs; otherwise leaves SMs free for a concurrent comm kernel.
// 0 = auto (pair kernel for large problems), 1 = force single-CTA kernel, 2 = force pair kernel
static int g_gemm_mode = 0;
SXB_API int sxb_gemm_set_mode(int mode) {
  if (mode < 0 || mode > 2) return SXB_ERR_BAD_ARG;
  g_gemm_mode = mode;
  return SXB_OK;
}

SXB_API int sxb_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, int layout,
                          int lda, int ldb, int ldc, int c_dtype, int accumulate, float alpha,
                          int max_ctas, void* stream) {
  if (M <= 0 || N <= 0 || K <= 0) return SXB_OK;
  if (layout < 0 || layout > 2) return SXB_ERR_BAD_ARG;
  if (c_dtype != kBF16 && c_dtype != kF32) return SXB_ERR_BAD_ARG;
  if ((lda % 8) || (ldb % 8) || (reinterpret_cast<uintptr_t>(A) & 15) ||
      (reinterpret_cast<uintptr_t>(B) & 15))
    return SXB_ERR_UNSUPPORTED;
  const bool tma_c =
      (c_dtype == kBF16) && !accumulate && alpha == 1.0f && (ldc % 8 == 0) &&
      ((reinterpret_cast<uintptr_t>(C) & 15) == 0);
  Params p;
  p.M = M; p.N = N; p.K = K; p.layout = layout; p.c_dtype = c_dtype;
  p.accumulate = accumulate ? 1 : 0; p.direct = tma_c ? 0 : 1; p.C = C; p.ldc = ldc;
  p.alpha = alpha;
  const int BNsel = (N > 128) ? 256 : 128;
  CUtensorMap ta, tb, tc;
  int rc;
  const CUtensorMapDataType BF = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  const bool big = (M >= 512 && N >= 512 && (long long)((M + 255) / 256) * ((N + 255) / 256) >= 37);
  if (g_gemm_mode == 2 || (g_gemm_mode == 0 && big)) {
    if (layout == TN) rc = make_map_2d(&ta, A, K, M, lda, BK, 64, BF, 2);
    else rc = make_map_2d(&ta, A, M, K, lda, 128, BK, BF, 2);
    if (rc) return rc;
    if (layout == NT) rc = make_map_2d(&tb, B, N, K, ldb, 128, BK, BF, 2);
    else rc = make_map_2d(&tb, B, K, N, ldb, BK, 64, BF, 2);
    if (rc) return rc;
    if (tma_c) {
      rc = make_map_2d(&tc, C, M, N, ldc, 128, 64, BF, 2);
      if (rc) return rc;
    } else {
      tc = ta;
    }
    return launch_pair(ta, tb, tc, p, max_ctas, reinterpret_cast<cudaStream_t>(stream));
  }
  if (layout == TN) rc = make_map_2d(&ta, A, K, M, lda, BK, 64, BF, 2);
  else rc = make_map_2d(&ta, A, M, K, lda, BM, BK, BF, 2);
  if (rc) return rc;
  if (layout == NT) rc = make_map_2d(&tb, B, N, K, ldb, BNsel, BK, BF, 2);
  else rc = make_map_2d(&tb, B, K, N, ldb, BK, 64, BF, 2);
  if (rc) return rc;
  if (tma_c) {
    rc = make_map_2d(&tc, C, M, N, ldc, BM, 64, BF, 2);
    if (rc) return rc;
  } else {
    tc = ta;  // never dereferenced by the direct epilogue
  }
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (BNsel == 256) return launch<256, 4>(ta, tb, tc, p, max_ctas, s);
  return launch<128, 6>(ta, tb, tc, p, max_ctas, s);
}
