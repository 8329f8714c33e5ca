// attention.cu — flash attention forward + backward on tcgen05 / TMEM / TMA (sm_100a).
//
// Layout: q [B, S, Hq, D], k/v [B, S, Hk, D] addressed through 4-D tensor maps
// {D, H, S, B} with arbitrary row pitch, so q/k/v may be views into one fused QKV buffer.
// Causal and full attention, GQA (Hq % Hk == 0), D in {64, 128}.
//
// Three kernels share one skeleton (warp 0 = TMA, warp 1 = one-thread MMA issuer,
// warps 2..5 = 128 "row" threads that own one TMEM lane each):
//   fwd    : S = Q K^T -> online softmax in registers -> P (bf16, swizzled smem) -> O += P V
//   bwd_dq : S, dP = dO V^T -> dS -> dQ += dS K                       (one CTA per 128 queries)
//   bwd_dkv: S^T = K Q^T, dP^T = V dO^T -> P^T, dS^T -> dV += P^T dO, dK += dS^T Q
//                                                                      (one CTA per 128 keys)
// The reference has no attention kernel for this architecture: training uses external
// flash-attn (deepspeed/sequence/fpdt_layer.py:14-20) and the inference wrapper rejects
// anything but sm_8x/sm_90 (deepspeed/inference/v2/kernels/ragged_ops/blocked_flash/blocked_flash.cpp:29-31).
#include "sxb_common.cuh"

namespace sxb {
namespace attn {

constexpr int kThreads = 192;
constexpr int kBwdThreads = 320;  // TMA warp + MMA warp + 8 row-warps
constexpr int BQ = 128;   // rows owned by the 128 row-threads (queries in fwd/dq, keys in dkv)
constexpr int BC = 64;    // inner block (keys in fwd/dq, queries in dkv)
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;
// SXB_ATTN_DEBUG=3: cycle accounting of the heaviest dQ tile (block 0,0,0) — [0] MMA warp total, [1] its wait for
// K/V tiles, [2] its wait for dS, [3] iterations, [4] row warp total, [5] its wait for S/dP, [6] its TMEM load +
// math, [7] its TMEM store + fence
__device__ unsigned long long g_attn_prof[8];
// The counters are compiled in only with -DSXB_ATTN_PROF (build an alternative library and load it through
// SXB_LIB_PATH): even as never-taken branches they sit in the MMA warp's issue loop, which is what bounds the dQ
// kernel (0.74 -> 0.79 ms with them compiled in).
#ifdef SXB_ATTN_PROF
constexpr bool kAttnProf = true;
#else
constexpr bool kAttnProf = false;
#endif

struct Params {
  int B, S, Hq, Hk, D;
  int causal;
  int debug;            // profiling aid (env SXB_ATTN_DEBUG): 1 = row warps skip the exp/dS math, 2 = also skip LDTM
  float scale;          // softmax scale (1/sqrt(D))
  void* out;            // fwd: O [B,S,Hq,D] ; dq: dQ ; dkv: dK
  void* out2;           // dkv: dV
  long long out_stride; // row pitch (elements) of out / out2
  long long out2_stride;
  float* lse;           // [B, Hq, S] natural-log LSE (fwd writes, bwd reads)
  const float* delta;   // [B, Hq, S] rowsum(dO * O) (bwd)
  // Row scatter (Ulysses K7: the kernel's epilogue returns every output row to the rank that owns that piece of the
  // sequence, as direct peer stores): scatter_rows > 0 -> row s of the sequence goes to out_peers[s / scatter_rows],
  // local row b * scatter_rows + s % scatter_rows.  Plain stores only; no TMA is involved on this path.
  int scatter_rows;
  void* out_peers[16];
  void* out2_peers[16];
  // Packed sequences ("varlen"): cu_seqlens[nseq + 1] row offsets into tensors of shape [1, S = total tokens, H, D];
  // blockIdx.y is then the sequence index, tiles past a sequence's end exit, and lse / delta are laid out [Hq, S].
  const int* cu_seqlens;
  // Ragged inference (forward only): one CTA per attention atom (int32 [n_atoms, 6], blockIdx.z), queries from the
  // ragged [T, Hq, D] tensor, K/V from the paged cache through the sequence's block table.
  const int* atoms;
  const int* block_tables;   // [n_seqs, max_blocks]
  int max_blocks, block_size;
  // Additive bias per (batch, key) in RAW-score units (the host divides the caller's bias by the softmax scale), e.g.
  // a key-padding mask with -inf on padded keys: fp32 [B, S] (dense batches only).  Added to S before the maximum, so
  // softmax, lse and the backward recomputation all see score * scale + bias.
  const float* kbias;
  long long kbias_bstride, kbias_hstride;   // element strides over batch and q head (0: the same bias for every head)
  // pack > 1 (atoms of at most 128 / pack queries, e.g. decode): the `pack` q heads that share one KV head are packed
  // into the tile's rows (row r = query r / pack, head hk * pack + r % pack), one CTA per KV head: K/V are read once
  // per KV head instead of once per q head.
  int pack;
};

// The rows of the tensors one CTA works on: sequence b of a dense batch, or sequence b of a packed one.
struct Win { int off, len, tb; };   // row offset inside TMA batch index `tb`, length
__device__ __forceinline__ Win window(const Params& p, int b) {
  Win w;
  if (p.cu_seqlens != nullptr) {
    w.off = p.cu_seqlens[b];
    w.len = p.cu_seqlens[b + 1] - w.off;
    w.tb = 0;
  } else {
    w.off = 0; w.len = p.S; w.tb = b;
  }
  return w;
}
__device__ __forceinline__ long long stat_base(const Params& p, const Win& w, int b, int h) {
  return p.cu_seqlens != nullptr ? (long long)h * p.S + w.off : ((long long)b * p.Hq + h) * p.S;
}

// Destination row pointer of output row (b, s): local buffer, or the owning peer's buffer when scattering.
__device__ __forceinline__ __nv_bfloat16* out_row_ptr(const Params& p, bool second, int b, int s) {
  if (p.scatter_rows > 0) {
    const int peer = s / p.scatter_rows;
    const long long r = (long long)b * p.scatter_rows + (s - peer * p.scatter_rows);
    return second ? reinterpret_cast<__nv_bfloat16*>(p.out2_peers[peer]) + r * p.out2_stride
                  : reinterpret_cast<__nv_bfloat16*>(p.out_peers[peer]) + r * p.out_stride;
  }
  const long long r = (long long)b * p.S + s;
  return second ? reinterpret_cast<__nv_bfloat16*>(p.out2) + r * p.out2_stride
                : reinterpret_cast<__nv_bfloat16*>(p.out) + r * p.out_stride;
}

// Write one row-thread's 64 bf16 values (128 B) into a [128 x 128B] K-major SW128 tile.
__device__ __forceinline__ void store_row_sw128(uint8_t* tile, int row, const float (&v)[64]) {
  uint8_t* dst = tile + row * 128;
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    uint4 o;
    o.x = pack_bf16(v[c * 8 + 0], v[c * 8 + 1]);
    o.y = pack_bf16(v[c * 8 + 2], v[c * 8 + 3]);
    o.z = pack_bf16(v[c * 8 + 4], v[c * 8 + 5]);
    o.w = pack_bf16(v[c * 8 + 6], v[c * 8 + 7]);
    *reinterpret_cast<uint4*>(dst + ((c ^ (row & 7)) << 4)) = o;
  }
}

// Half-row variants for the backward kernels, where TWO warps share each TMEM lane quadrant and split
// the 64 columns of a block between them (8 row-warps per CTA -> 2 per scheduler instead of 1).
__device__ __forceinline__ void store_half_sw128(uint8_t* tile, int row, int half, const float (&v)[32]) {
  uint8_t* dst = tile + row * 128;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    uint4 o;
    o.x = pack_bf16(v[c * 8 + 0], v[c * 8 + 1]);
    o.y = pack_bf16(v[c * 8 + 2], v[c * 8 + 3]);
    o.z = pack_bf16(v[c * 8 + 4], v[c * 8 + 5]);
    o.w = pack_bf16(v[c * 8 + 6], v[c * 8 + 7]);
    *reinterpret_cast<uint4*>(dst + (((half * 4 + c) ^ (row & 7)) << 4)) = o;
  }
}
__device__ __forceinline__ void ld_row32x2(uint32_t taddr_a, uint32_t taddr_b, float (&a)[32], float (&b)[32]) {
  uint32_t x[32], y[32];
  tmem_ld_32x32(taddr_a, x);
  tmem_ld_32x32(taddr_b, y);
  tmem_ld_wait();
#pragma unroll
  for (int i = 0; i < 32; ++i) {
    a[i] = __uint_as_float(x[i]);
    b[i] = __uint_as_float(y[i]);
  }
}

__device__ __forceinline__ float4 lds128(uint32_t saddr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(saddr));
  return v;
}

__device__ __forceinline__ void ld_row64(uint32_t taddr, float (&v)[64]) {
  uint32_t a[32], b[32];
  tmem_ld_32x32(taddr, a);
  tmem_ld_32x32(taddr + 32, b);
  tmem_ld_wait();
#pragma unroll
  for (int i = 0; i < 32; ++i) {
    v[i] = __uint_as_float(a[i]);
    v[32 + i] = __uint_as_float(b[i]);
  }
}

// K-major operand whose K extent is D (Q, K, V, dO tiles): `rows` rows, D/64 halves of
// [rows x 128B]; k-step kk covers d in [16kk, 16kk+16).
__device__ __forceinline__ uint64_t desc_kmajor_d(uint32_t base, int rows, int kk) {
  return make_smem_desc(base + (kk >> 2) * (rows * 128) + (kk & 3) * 32, 16, 1024, 2);
}
// K-major operand with a single 128B-wide tile (P / dS tiles): k-step kk covers 16 columns.
__device__ __forceinline__ uint64_t desc_kmajor_1(uint32_t base, int kk) {
  return make_smem_desc(base + kk * 32, 16, 1024, 2);
}
// MN-major operand: tile stored as [krows x D] (D contiguous) in D/64 halves of
// [krows x 128B]; k-step kk covers krows [16kk, 16kk+16).
__device__ __forceinline__ uint64_t desc_mnmajor(uint32_t base, int krows, int kk) {
  return make_smem_desc(base + kk * 2048, krows * 128, 1024, 2);
}

// ============================================================================ forward
template <int D>
struct FwdSmem {
  static constexpr int kQ = BQ * D * 2;
  static constexpr int kKV = BC * D * 2;
  static constexpr int kStages = 2;
  static constexpr int oQ = 0;
  static constexpr int oK = oQ + kQ;
  static constexpr int oV = oK + kStages * kKV;
  static constexpr int oBar = oV + kStages * kKV;   // 96 KB + barriers at D = 128: two CTAs per SM
  static constexpr int kTotal = oBar + 128 + 1024;
};

// MC = 2: launched as 2-CTA clusters over adjacent q heads of one GQA group.  Both CTAs walk the same K/V blocks, so
// each loads one 64-wide half of every K/V tile and multicasts it into both CTAs' rings: L2 -> SM traffic (the
// measured limiter of this kernel: 5.4 TB/s of the ~6.9 TB/s L2 can deliver) is halved.
template <int D, int MC>
__global__ void __launch_bounds__(kThreads, 2)
attn_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                const __grid_constant__ CUtensorMap tmV, const Params p) {
  using L = FwdSmem<D>;
  constexpr int NH = D / 64;  // 128-byte halves along D
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::oBar);
  uint64_t* q_full = bars + 0;
  uint64_t* k_full = bars + 1;    // [2]
  uint64_t* v_full = bars + 3;    // [2]
  uint64_t* k_empty = bars + 5;   // [2]  released as soon as S = Q K^T of that block finished
  uint64_t* v_empty = bars + 7;   // [2]  released when O += P V finished
  uint64_t* s_full = bars + 9;    // [2]
  uint64_t* p_full = bars + 11;
  uint64_t* o_done = bars + 12;
  uint64_t* s_free = bars + 13;   // [2]  row warps have copied S(j) out of TMEM (one arrive per warp)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 15);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // tile index is the SLOWEST grid dimension: CTAs are handed out in linear block order, so every (head, batch)
  // of the heaviest causal tile starts before any lighter tile (longest-first packing over the 148 SMs)
  const int pack = p.pack > 1 ? p.pack : 1;
  const int b = blockIdx.y;
  const int hk = pack > 1 ? (int)blockIdx.x : (int)blockIdx.x / (p.Hq / p.Hk);
  const int h = pack > 1 ? hk * pack : (int)blockIdx.x;      // (first) q head of this CTA
  // The tile: TMA row of its first query, that query's position inside its sequence, valid rows, keys it attends to.
  Win w;
  int q_row0, q_pos0, q_rows, kv_len;
  const int* pages = nullptr;           // paged K/V: this sequence's block table
  if (p.atoms != nullptr) {
    // attention atom (csrc/host/ragged_host.cpp): (seq, q_start_idx, q_len, global_q_idx, kv_blocks, total_extent)
    const int* a = p.atoms + 6 * (int)blockIdx.z;
    w.off = 0; w.len = 0; w.tb = 0;
    q_row0 = a[1]; q_rows = a[2]; q_pos0 = a[3]; kv_len = a[5];
    pages = p.block_tables + (long long)a[0] * p.max_blocks;
  } else {
    const int q0 = (gridDim.z - 1 - blockIdx.z) * BQ;
    w = window(p, b);
    if (q0 >= w.len) return;   // packed mode: tile beyond this sequence (uniform over the CTA and its cluster peer)
    q_row0 = w.off + q0; q_pos0 = q0; q_rows = min(BQ, w.len - q0);
    kv_len = p.causal ? min(w.len, q0 + BQ) : w.len;
  }
  const int nblk = (kv_len + BC - 1) / BC;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV);
  }
  if (warp == 1) {
    if (lane == 0) {
      mbar_init(q_full, 1);
      for (int i = 0; i < 2; ++i) {
        mbar_init(&k_full[i], 1); mbar_init(&v_full[i], 1); mbar_init(&k_empty[i], MC);
        mbar_init(&v_empty[i], MC); mbar_init(&s_full[i], 1);
      }
      mbar_init(p_full, 4);     // one arrival per row warp
      mbar_init(o_done, 1);
      mbar_init(&s_free[0], 4);
      mbar_init(&s_free[1], 4);
      mbar_fence_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, 256);
  }
  tc_fence_before();
  if (MC > 1) cluster_sync_all();   // peer barriers exist before any multicast byte or remote arrive lands
  else __syncthreads();
  tc_fence_after();
  const uint32_t crank = (MC > 1) ? cluster_ctarank() : 0;
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem;          // 2 x 64 columns
  const uint32_t tO = tmem + 128;    // D columns

  // Producer / MMA warps run whole-warp in uniform control flow; only the asynchronous instruction is
  // predicated on one elected lane (keeps operands in uniform registers: no R2UR waterfall per MMA).
  if (warp == 0) {
    const bool elected = elect_one();
    if (elected) {
      mbar_expect_tx(q_full, L::kQ);
      for (int hh = 0; hh < NH; ++hh)
        tma_load_4d(smem + L::oQ + hh * (BQ * 128), &tmQ, q_full, hh * 64, h, q_row0, w.tb);
    }
    __syncwarp();
    // One 64-key tile of K or V into a ring slot: contiguous rows, or the rows of the page(s) the block table names
    // (tensor map {D, Hk, block_size, n_blocks}; pages smaller than the tile take several boxes; a tile that reaches
    // past the sequence re-reads its last page, the rows are masked).
    auto load_kv = [&](const CUtensorMap* tm, uint8_t* dst, uint64_t* bar, int j) {
      if (pages == nullptr) {
        if (MC > 1) {
          tma_load_4d_mc(dst + crank * (BC * 128), tm, bar, crank * 64, hk, w.off + j * BC, w.tb, 3);
        } else {
          for (int hh = 0; hh < NH; ++hh)
            tma_load_4d(dst + hh * (BC * 128), tm, bar, hh * 64, hk, w.off + j * BC, w.tb);
        }
        return;
      }
      const int bs = p.block_size, sub = min(bs, BC), last = (kv_len - 1) / bs;
      for (int r = 0; r < BC; r += sub) {
        const int pos = j * BC + r;
        const int blk = min(pos / bs, last);
        const int page = pages[blk], in_page = (blk == pos / bs) ? pos % bs : 0;
        if (MC > 1) {
          tma_load_4d_mc(dst + crank * (BC * 128) + r * 128, tm, bar, crank * 64, hk, in_page, page, 3);
        } else {
          for (int hh = 0; hh < NH; ++hh)
            tma_load_4d(dst + hh * (BC * 128) + r * 128, tm, bar, hh * 64, hk, in_page, page);
        }
      }
    };
    // K and V rings advance independently (polling state machine): a K slot is recycled right
    // after its QK^T, so the next-but-one K tile is in flight ~2 iterations before it is needed.
    int jk = 0, jv = 0;
    unsigned long long t0 = 0;
    long long c_next = 0;
    while (jk < nblk || jv < nblk) {
      bool progressed = false;
      const bool k_ok = jk < nblk &&
          __shfl_sync(0xffffffffu, (int)mbar_try_wait(&k_empty[jk & 1], ((jk >> 1) & 1) ^ 1), 0);
      if (k_ok) {
        const int st = jk & 1;
        if (elected) {
          mbar_expect_tx(&k_full[st], L::kKV);
          load_kv(&tmK, smem + L::oK + st * L::kKV, &k_full[st], jk);
        }
        __syncwarp();
        ++jk;
        progressed = true;
      }
      const bool v_ok = jv < nblk &&
          __shfl_sync(0xffffffffu, (int)mbar_try_wait(&v_empty[jv & 1], ((jv >> 1) & 1) ^ 1), 0);
      if (v_ok) {
        const int st = jv & 1;
        if (elected) {
          mbar_expect_tx(&v_full[st], L::kKV);
          load_kv(&tmV, smem + L::oV + st * L::kKV, &v_full[st], jv);
        }
        __syncwarp();
        ++jv;
        progressed = true;
      }
      if (progressed) { t0 = 0; c_next = 0; }
      else {
        const bool abort_ = hang_should_abort(t0, 21, jk * 1000 + jv, &c_next);
        if (__any_sync(0xffffffffu, abort_)) goto done;
      }
    }
  } else if (warp == 1) {
    const bool elected = elect_one();
    const uint32_t sbase = __shfl_sync(0xffffffffu, smem_u32(smem), 0);
    const uint32_t tS_u = __shfl_sync(0xffffffffu, tS, 0);
    const uint32_t tO_u = tS_u + 128;
    const uint32_t idesc_s = make_idesc(BQ, BC, 1, 1, 0, 0);
    const uint32_t idesc_o = make_idesc(BQ, D, 1, 1, 0, 1);
    const uint32_t sQ = sbase + L::oQ;
    auto issue_qk = [&](int j) -> bool {
      const int st = j & 1;
      if (!__all_sync(0xffffffffu, mbar_wait_to(&k_full[st], (j >> 1) & 1, 11, j))) return false;
      tc_fence_after();
      const uint32_t sK = sbase + L::oK + st * L::kKV;
      if (elected) {
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk)
          umma_f16(tS_u + (j & 1) * BC, desc_kmajor_d(sQ, BQ, kk), desc_kmajor_d(sK, BC, kk),
                   idesc_s, kk != 0);
        umma_commit(&s_full[j & 1]);
        if (MC > 1) umma_commit_mc(&k_empty[st], 3);
        else umma_commit(&k_empty[st]);
      }
      __syncwarp();
      return true;
    };
    if (!__all_sync(0xffffffffu, mbar_wait_to(q_full, 0, 10))) goto done;
    if (!issue_qk(0)) goto done;
    // Measured: issuing S(j+2) as soon as the row warps have copied S(j) out of TMEM (an "early
    // release" of the S buffer) is SLOWER (0.98-1.24 ms vs 0.83 ms at S=8192) than keeping S one block
    // ahead, so QK(j+1) is issued right behind PV(j-1).
    for (int j = 0; j < nblk; ++j) {
      const int st = j & 1;
      if (j + 1 < nblk && !issue_qk(j + 1)) goto done;
      if (!__all_sync(0xffffffffu, mbar_wait_to(p_full, j & 1, 22))) goto done;
      if (!__all_sync(0xffffffffu, mbar_wait_to(&v_full[st], (j >> 1) & 1, 23))) goto done;
      tc_fence_after();
      const uint32_t sV = sbase + L::oV + st * L::kKV;
      if (elected) {
#pragma unroll
        for (int kk = 0; kk < BC / 16; ++kk)   // A = P(j), bf16-packed over the first 32 columns of S(j) in TMEM
          umma_f16_ts(tO_u, tS_u + (j & 1) * BC + kk * 8, desc_mnmajor(sV, BC, kk), idesc_o, (j | kk) != 0);
        if (MC > 1) umma_commit_mc(&v_empty[st], 3);
        else umma_commit(&v_empty[st]);
        umma_commit(o_done);
      }
      __syncwarp();
    }
  } else {
    const int quad = warp & 3;
    const int row = quad * 32 + lane;
    const uint32_t lane_off = uint32_t(quad * 32) << 16;
    const int qtok = pack > 1 ? row / pack : row;               // query (token) index inside the tile
    const int qhead = pack > 1 ? h + (row - qtok * pack) : h;
    const int qpos = q_pos0 + qtok;
    const float sc = p.scale * kLog2e;
    float m_run = -INFINITY, l_run = 0.f;
    for (int j = 0; j < nblk; ++j) {
      if (!mbar_wait_to(&s_full[j & 1], (j >> 1) & 1, 24)) goto done;
      tc_fence_after();
      float s[64];
      ld_row64(tS + (j & 1) * BC + lane_off, s);
      const int k0 = j * BC;
      if (p.kbias != nullptr) {
        const float4* kb = reinterpret_cast<const float4*>(p.kbias + b * p.kbias_bstride + h * p.kbias_hstride + k0);   // warp-uniform
        const int nv = min(16, (p.S - k0 + 3) >> 2);
#pragma unroll
        for (int c4 = 0; c4 < 16; ++c4) {
          if (c4 < nv) {
            const float4 t = __ldg(kb + c4);
            s[c4 * 4 + 0] += t.x; s[c4 * 4 + 1] += t.y; s[c4 * 4 + 2] += t.z; s[c4 * 4 + 3] += t.w;
          }
        }
      }
      const bool need_mask = (p.causal && (k0 + BC - 1 > q_pos0)) || (k0 + BC > kv_len);
      // row max on raw scores (scale > 0 commutes with max); 4 independent chains
      float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
      if (need_mask) {
#pragma unroll
        for (int c = 0; c < 64; ++c) {
          const int kp = k0 + c;
          const bool ok = (kp < kv_len) && (!p.causal || kp <= qpos);
          s[c] = ok ? s[c] : -INFINITY;
          mx4[c & 3] = fmaxf(mx4[c & 3], s[c]);
        }
      } else {
#pragma unroll
        for (int c = 0; c < 64; ++c) mx4[c & 3] = fmaxf(mx4[c & 3], s[c]);
      }
      const float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3])) * sc;
      // Lazy rescale: keep the running max stale unless it grew by more than 2^8.
      float m_new = m_run;
      const bool grow = mx > m_run + 8.f;
      const bool any_grow = __any_sync(0xffffffffu, grow) != 0;
      if (any_grow) m_new = fmaxf(m_run, mx);
      const float m_use = (m_new == -INFINITY) ? 0.f : m_new;
      float rs4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int c = 0; c < 64; ++c) {
        s[c] = fast_exp2(fmaf(s[c], sc, -m_use));
        rs4[c & 3] += s[c];
      }
      const float rs = (rs4[0] + rs4[1]) + (rs4[2] + rs4[3]);
      if (j > 0) {
        if (!mbar_wait_to(o_done, (j - 1) & 1, 25)) goto done;
        tc_fence_after();
        if (any_grow) {
          const float alpha = (m_run == -INFINITY) ? 0.f : fast_exp2(m_run - m_use);
          l_run *= alpha;
#pragma unroll 1
          for (int c = 0; c < D / 32; ++c) {
            uint32_t o[32];
            tmem_ld_32x32(tO + lane_off + c * 32, o);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            tmem_st_32x32(tO + lane_off + c * 32, o);
          }
          tmem_st_wait();
        }
      }
      m_run = m_new;
      l_run += rs;
      // P(j) goes back into TMEM over S(j).  S(j+1) shares the 128-column group: a tcgen05.st into a group the tensor
      // core is still accumulating into was measured to corrupt data, so QK(j+1) must have landed first (it was
      // issued a whole iteration ago; this wait is normally already satisfied).
      if (j + 1 < nblk) {
        if (!mbar_wait_to(&s_full[(j + 1) & 1], ((j + 1) >> 1) & 1, 27)) goto done;
      }
      {
        uint32_t pk[32];
#pragma unroll
        for (int c = 0; c < 32; ++c) pk[c] = pack_bf16(s[2 * c], s[2 * c + 1]);
        tmem_st_32x32(tS + (j & 1) * BC + lane_off, pk);
        tmem_st_wait();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_full);
    }
    if (!mbar_wait_to(o_done, (nblk - 1) & 1, 26)) goto done;
    tc_fence_after();
    const float inv_l = (l_run > 0.f) ? 1.f / l_run : 0.f;
    {
      // tcgen05.ld is warp-collective (.sync.aligned): every lane issues it, only the global
      // stores are predicated on the row being inside the sequence.
      const bool valid = qtok < q_rows;
      const int grow = valid ? q_row0 + qtok : q_row0;   // row of the [rows, H, D] tensors
      __nv_bfloat16* orow = out_row_ptr(p, false, w.tb, grow) + (long long)qhead * D;
      // scattering forward keeps a LOCAL copy of O as well (p.out, pitch out2_stride): the backward needs
      // rowsum(dO * O) where the heads are, not where the sequence is
      __nv_bfloat16* orow_keep = (p.scatter_rows > 0 && p.out != nullptr)
          ? reinterpret_cast<__nv_bfloat16*>(p.out) + ((long long)w.tb * p.S + grow) * p.out2_stride +
                (long long)qhead * D
          : nullptr;
#pragma unroll 1
      for (int c = 0; c < D / 32; ++c) {
        uint32_t o[32];
        tmem_ld_32x32(tO + lane_off + c * 32, o);
        tmem_ld_wait();
        if (valid) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            uint4 w;
            w.x = pack_bf16(__uint_as_float(o[i * 8 + 0]) * inv_l, __uint_as_float(o[i * 8 + 1]) * inv_l);
            w.y = pack_bf16(__uint_as_float(o[i * 8 + 2]) * inv_l, __uint_as_float(o[i * 8 + 3]) * inv_l);
            w.z = pack_bf16(__uint_as_float(o[i * 8 + 4]) * inv_l, __uint_as_float(o[i * 8 + 5]) * inv_l);
            w.w = pack_bf16(__uint_as_float(o[i * 8 + 6]) * inv_l, __uint_as_float(o[i * 8 + 7]) * inv_l);
            *reinterpret_cast<uint4*>(orow + c * 32 + i * 8) = w;
            if (orow_keep) *reinterpret_cast<uint4*>(orow_keep + c * 32 + i * 8) = w;
          }
        }
      }
      if (valid && p.lse)
        p.lse[stat_base(p, w, b, qhead) - w.off + grow] =
            (l_run > 0.f) ? (m_run + log2f(l_run)) * kLn2 : -INFINITY;
    }
  }
done:
  tc_fence_before();
  if (MC > 1) cluster_sync_all();   // neither CTA leaves while the peer can still write its ring or barriers
  else __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem, 256);
  }
}

// ============================================================================ forward, 128-key blocks (D = 128)
// Same roles and interfaces as attn_fwd_kernel; what differs is the block shape.  Measured on the 64-key kernel: both
// CTAs of an SM together take ~2050 cycles per 64-key step although their MMAs need 1024 — S = Q K^T is an SS-form MMA
// with N = 64 that re-reads the 4 KB Q slice for every 2 KB K slice, 128 KB of shared-memory operand traffic per
// step and SM at the ~70 B/clk the tensor core sustains from shared memory.  With 128-key blocks the same Q slice
// serves twice the keys (96 KB per 128 keys instead of 128 KB).  The price is TMEM: S is 128 fp32 columns, so there
// is ONE S buffer next to O (256 columns, still two CTAs per SM) and inside a CTA the steps run strictly one after
// the other — QK^T(j), softmax(j) (P overwrites S), O += P V(j) — the second CTA of the SM fills the gaps.  K and V
// tiles are single slots of 32 KB each: K(j+1) is requested when QK^T(j) retires, V(j+1) when PV(j) does.
// Eight row warps: two per TMEM lane quadrant, each owning 64 of the block's 128 columns (and 64 of O's), so every
// scheduler has four softmax warps of the SM's two CTAs to interleave (with four per CTA the exponentials, FMAs and
// TMEM traffic of one warp per scheduler ran back to back: XU and tensor pipe both ~50 % busy).  The two warps of a
// row exchange their partial row maxima through shared memory once per block; partial row sums meet at the end.
constexpr int FBC = 128;
constexpr int kFwd128Threads = 320;
struct Fwd128Smem {
  static constexpr int kQ = BQ * 128 * 2;
  static constexpr int kKV = FBC * 128 * 2;
  static constexpr int oQ = 0;
  static constexpr int oK = oQ + kQ;
  static constexpr int oV = oK + kKV;
  static constexpr int oBar = oV + kKV;          // 96 KB + barriers: two CTAs per SM
  static constexpr int oXch = oBar + 128;        // float [2][128]: partial row max / row sum of the two column halves
  static constexpr int kTotal = oXch + 1024 + 1024;
};

template <int MC>
__global__ void __launch_bounds__(kFwd128Threads, 2)
attn_fwd128_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                   const __grid_constant__ CUtensorMap tmV, const Params p) {
  using L = Fwd128Smem;
  constexpr int D = 128, NH = 2;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::oBar);
  uint64_t* q_full = bars + 0;
  uint64_t* k_full = bars + 1;
  uint64_t* v_full = bars + 2;
  uint64_t* k_empty = bars + 3;   // QK^T of the block retired (every CTA of the cluster)
  uint64_t* v_empty = bars + 4;   // O += P V of the block retired
  uint64_t* s_full = bars + 5;
  uint64_t* p_full = bars + 6;    // one arrival per row warp
  uint64_t* o_done = bars + 7;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int pack = p.pack > 1 ? p.pack : 1;
  const int b = blockIdx.y;
  const int hk = pack > 1 ? (int)blockIdx.x : (int)blockIdx.x / (p.Hq / p.Hk);
  const int h = pack > 1 ? hk * pack : (int)blockIdx.x;
  Win w;
  int q_row0, q_pos0, q_rows, kv_len;
  const int* pages = nullptr;
  if (p.atoms != nullptr) {
    const int* a = p.atoms + 6 * (int)blockIdx.z;
    w.off = 0; w.len = 0; w.tb = 0;
    q_row0 = a[1]; q_rows = a[2]; q_pos0 = a[3]; kv_len = a[5];
    pages = p.block_tables + (long long)a[0] * p.max_blocks;
  } else {
    const int q0 = (gridDim.z - 1 - blockIdx.z) * BQ;
    w = window(p, b);
    if (q0 >= w.len) return;
    q_row0 = w.off + q0; q_pos0 = q0; q_rows = min(BQ, w.len - q0);
    kv_len = p.causal ? min(w.len, q0 + BQ) : w.len;
  }
  const int nblk = (kv_len + FBC - 1) / FBC;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV);
  }
  if (warp == 1) {
    if (lane == 0) {
      mbar_init(q_full, 1); mbar_init(k_full, 1); mbar_init(v_full, 1);
      mbar_init(k_empty, MC); mbar_init(v_empty, MC);
      mbar_init(s_full, 1); mbar_init(p_full, 8); mbar_init(o_done, 1);
      mbar_fence_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, 256);
  }
  tc_fence_before();
  if (MC > 1) cluster_sync_all();
  else __syncthreads();
  tc_fence_after();
  const uint32_t crank = (MC > 1) ? cluster_ctarank() : 0;
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem;          // 128 columns: S, then P (bf16 pairs) over its first 64
  const uint32_t tO = tmem + 128;    // D columns

  if (warp == 0) {
    const bool elected = elect_one();
    if (elected) {
      mbar_expect_tx(q_full, L::kQ);
      for (int hh = 0; hh < NH; ++hh)
        tma_load_4d(smem + L::oQ + hh * (BQ * 128), &tmQ, q_full, hh * 64, h, q_row0, w.tb);
    }
    __syncwarp();
    auto load_kv = [&](const CUtensorMap* tm, uint8_t* dst, uint64_t* bar, int j) {
      if (pages == nullptr) {
        if (MC > 1) {
          tma_load_4d_mc(dst + crank * (FBC * 128), tm, bar, crank * 64, hk, w.off + j * FBC, w.tb, 3);
        } else {
          for (int hh = 0; hh < NH; ++hh)
            tma_load_4d(dst + hh * (FBC * 128), tm, bar, hh * 64, hk, w.off + j * FBC, w.tb);
        }
        return;
      }
      const int bs = p.block_size, sub = min(bs, FBC), last = (kv_len - 1) / bs;
      for (int r = 0; r < FBC; r += sub) {
        const int pos = j * FBC + r;
        const int blk = min(pos / bs, last);
        const int page = pages[blk], in_page = (blk == pos / bs) ? pos % bs : 0;
        if (MC > 1) {
          tma_load_4d_mc(dst + crank * (FBC * 128) + r * 128, tm, bar, crank * 64, hk, in_page, page, 3);
        } else {
          for (int hh = 0; hh < NH; ++hh)
            tma_load_4d(dst + hh * (FBC * 128) + r * 128, tm, bar, hh * 64, hk, in_page, page);
        }
      }
    };
    for (int j = 0; j < nblk; ++j) {
      if (!__all_sync(0xffffffffu, mbar_wait_to(k_empty, (j & 1) ^ 1, 50, j))) goto done;
      if (elected) {
        mbar_expect_tx(k_full, L::kKV);
        load_kv(&tmK, smem + L::oK, k_full, j);
      }
      __syncwarp();
      if (!__all_sync(0xffffffffu, mbar_wait_to(v_empty, (j & 1) ^ 1, 51, j))) goto done;
      if (elected) {
        mbar_expect_tx(v_full, L::kKV);
        load_kv(&tmV, smem + L::oV, v_full, j);
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    const bool elected = elect_one();
    const uint32_t sbase = __shfl_sync(0xffffffffu, smem_u32(smem), 0);
    const uint32_t tS_u = __shfl_sync(0xffffffffu, tS, 0);
    const uint32_t tO_u = tS_u + 128;
    const uint32_t idesc_s = make_idesc(BQ, FBC, 1, 1, 0, 0);
    const uint32_t idesc_o = make_idesc(BQ, D, 1, 1, 0, 1);
    const uint32_t sQ = sbase + L::oQ, sK = sbase + L::oK, sV = sbase + L::oV;
    if (!__all_sync(0xffffffffu, mbar_wait_to(q_full, 0, 52))) goto done;
    const bool prof = kAttnProf && p.debug == 3 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0;
    long long pr_t0 = prof ? clock64() : 0, pr_k = 0, pr_p = 0, pr_v = 0;
    for (int j = 0; j < nblk; ++j) {
      long long c0 = prof ? clock64() : 0;
      if (!__all_sync(0xffffffffu, mbar_wait_to(k_full, j & 1, 53, j))) goto done;
      if (prof) pr_k += clock64() - c0;
      tc_fence_after();
      if (elected) {
        // S(j) overwrites P(j-1): issued behind PV(j-1) in the same thread's MMA stream, which executes in order
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk)
          umma_f16(tS_u, desc_kmajor_d(sQ, BQ, kk), desc_kmajor_d(sK, FBC, kk), idesc_s, kk != 0);
        umma_commit(s_full);
        if (MC > 1) umma_commit_mc(k_empty, 3);
        else umma_commit(k_empty);
      }
      __syncwarp();
      c0 = prof ? clock64() : 0;
      if (!__all_sync(0xffffffffu, mbar_wait_to(p_full, j & 1, 54, j))) goto done;
      if (prof) { pr_p += clock64() - c0; c0 = clock64(); }
      if (!__all_sync(0xffffffffu, mbar_wait_to(v_full, j & 1, 55, j))) goto done;
      if (prof) pr_v += clock64() - c0;
      tc_fence_after();
      if (elected) {
#pragma unroll
        for (int kk = 0; kk < FBC / 16; ++kk)
          // P of keys [64 hf, 64 hf + 64) sits bf16-packed on the first 32 of that half's own S columns
          umma_f16_ts(tO_u, tS_u + (kk >> 2) * 64 + (kk & 3) * 8, desc_mnmajor(sV, FBC, kk), idesc_o, (j | kk) != 0);
        if (MC > 1) umma_commit_mc(v_empty, 3);
        else umma_commit(v_empty);
        if (j == nblk - 1) umma_commit(o_done);
      }
      __syncwarp();
    }
    if (prof && lane == 0) {
      g_attn_prof[0] = clock64() - pr_t0; g_attn_prof[1] = pr_k; g_attn_prof[2] = pr_p; g_attn_prof[3] = nblk;
      g_attn_prof[7] = pr_v;
    }
  } else {
    const int quad = warp & 3;                 // TMEM lane quadrant (hardware: warp id mod 4)
    const int hf = (warp - 2) >> 2;            // which 64 of the block's 128 key columns (and of O's 128)
    const int row = quad * 32 + lane;
    const uint32_t lane_off = uint32_t(quad * 32) << 16;
    const int qtok = pack > 1 ? row / pack : row;
    const int qhead = pack > 1 ? h + (row - qtok * pack) : h;
    const int qpos = q_pos0 + qtok;
    const float sc = p.scale * kLog2e;
    float* xch = reinterpret_cast<float*>(smem + L::oXch);     // [2][128]
    const uint32_t tSh = tS + hf * 64 + lane_off;               // this warp's 64 S columns
    float m_run = -INFINITY, l_run = 0.f;                       // l_run: partial sum over this warp's columns
    const bool rprof = kAttnProf && p.debug == 3 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && warp == 2;
    long long rp_t0 = rprof ? clock64() : 0, rp_wait = 0, rp_work = 0;
    for (int j = 0; j < nblk; ++j) {
      // s_full(j) is a commit behind QK^T(j): every earlier MMA of this CTA, PV(j-1) included, has retired, so O may
      // be rescaled and S may be overwritten with P without further waits
      long long rc0 = rprof ? clock64() : 0;
      if (!mbar_wait_to(s_full, j & 1, 56, j)) goto done;
      long long rc1 = rprof ? clock64() : 0;
      rp_wait += rc1 - rc0;
      tc_fence_after();
      const int kh = j * FBC + hf * 64;
      const bool need_mask = (p.causal && (kh + 63 > q_pos0)) || (kh + 64 > kv_len);
      // pass 1: maximum over this warp's 64 raw scores (scale > 0 commutes with max)
      float mx = -INFINITY;
#pragma unroll 1
      for (int q4 = 0; q4 < 2; ++q4) {
        uint32_t x[32];
        tmem_ld_32x32(tSh + q4 * 32, x);
        tmem_ld_wait();
        float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
        if (need_mask) {
#pragma unroll
          for (int c = 0; c < 32; ++c) {
            const int kp = kh + q4 * 32 + c;
            const bool ok = (kp < kv_len) && (!p.causal || kp <= qpos);
            mx4[c & 3] = fmaxf(mx4[c & 3], ok ? __uint_as_float(x[c]) : -INFINITY);
          }
        } else {
#pragma unroll
          for (int c = 0; c < 32; ++c) mx4[c & 3] = fmaxf(mx4[c & 3], __uint_as_float(x[c]));
        }
        mx = fmaxf(mx, fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3])));
      }
      // the row's other 64 columns belong to the partner warp (same quadrant): swap partial maxima
      xch[hf * 128 + row] = mx;
      named_bar_sync(1 + quad, 64);
      mx = fmaxf(mx, xch[(hf ^ 1) * 128 + row]) * sc;
      // lazy rescale: keep the running max stale unless it grew by more than 2^8 (both warps of a row see the same mx)
      float m_new = m_run;
      const bool grow = mx > m_run + 8.f;
      const bool any_grow = __any_sync(0xffffffffu, grow) != 0;
      if (any_grow) m_new = fmaxf(m_run, mx);
      const float m_use = (m_new == -INFINITY) ? 0.f : m_new;
      if (j > 0 && any_grow) {
        const float alpha = (m_run == -INFINITY) ? 0.f : fast_exp2(m_run - m_use);
        l_run *= alpha;
#pragma unroll 1
        for (int c = hf * 2; c < hf * 2 + 2; ++c) {               // this warp's 64 of O's 128 columns
          uint32_t o[32];
          tmem_ld_32x32(tO + lane_off + c * 32, o);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
          tmem_st_32x32(tO + lane_off + c * 32, o);
        }
        tmem_st_wait();
      }
      m_run = m_new;
      // pass 2: P = exp2(s * sc - m), bf16-packed over the first 32 of this warp's own 64 S columns.  The first 32
      // scores are loaded before they are overwritten; the second 32 land on columns 16..31, already consumed.
      float rs = 0.f;
#pragma unroll 1
      for (int q4 = 0; q4 < 2; ++q4) {
        uint32_t x[32];
        tmem_ld_32x32(tSh + q4 * 32, x);
        tmem_ld_wait();
        float e[32];
        float rs4[4] = {0.f, 0.f, 0.f, 0.f};
        if (need_mask) {
#pragma unroll
          for (int c = 0; c < 32; ++c) {
            const int kp = kh + q4 * 32 + c;
            const bool ok = (kp < kv_len) && (!p.causal || kp <= qpos);
            e[c] = ok ? fast_exp2(fmaf(__uint_as_float(x[c]), sc, -m_use)) : 0.f;
            rs4[c & 3] += e[c];
          }
        } else {
#pragma unroll
          for (int c = 0; c < 32; ++c) {
            e[c] = fast_exp2(fmaf(__uint_as_float(x[c]), sc, -m_use));
            rs4[c & 3] += e[c];
          }
        }
        rs += (rs4[0] + rs4[1]) + (rs4[2] + rs4[3]);
        uint32_t pk[16];
#pragma unroll
        for (int c = 0; c < 16; ++c) pk[c] = pack_bf16(e[2 * c], e[2 * c + 1]);
        tmem_st_32x16(tSh + q4 * 16, pk);
      }
      tmem_st_wait();
      l_run += rs;
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_full);
      if (rprof) rp_work += clock64() - rc1;
    }
    if (rprof && lane == 0) { g_attn_prof[4] = clock64() - rp_t0; g_attn_prof[5] = rp_wait; g_attn_prof[6] = rp_work; }
    if (!mbar_wait_to(o_done, 0, 57)) goto done;
    tc_fence_after();
    xch[hf * 128 + row] = l_run;
    named_bar_sync(1 + quad, 64);
    l_run += xch[(hf ^ 1) * 128 + row];
    const float inv_l = (l_run > 0.f) ? 1.f / l_run : 0.f;
    {
      const bool valid = qtok < q_rows;
      const int grow = valid ? q_row0 + qtok : q_row0;
      __nv_bfloat16* orow = out_row_ptr(p, false, w.tb, grow) + (long long)qhead * D;
      __nv_bfloat16* orow_keep = (p.scatter_rows > 0 && p.out != nullptr)
          ? reinterpret_cast<__nv_bfloat16*>(p.out) + ((long long)w.tb * p.S + grow) * p.out2_stride +
                (long long)qhead * D
          : nullptr;
#pragma unroll 1
      for (int c = hf * 2; c < hf * 2 + 2; ++c) {
        uint32_t o[32];
        tmem_ld_32x32(tO + lane_off + c * 32, o);
        tmem_ld_wait();
        if (valid) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            uint4 w4;
            w4.x = pack_bf16(__uint_as_float(o[i * 8 + 0]) * inv_l, __uint_as_float(o[i * 8 + 1]) * inv_l);
            w4.y = pack_bf16(__uint_as_float(o[i * 8 + 2]) * inv_l, __uint_as_float(o[i * 8 + 3]) * inv_l);
            w4.z = pack_bf16(__uint_as_float(o[i * 8 + 4]) * inv_l, __uint_as_float(o[i * 8 + 5]) * inv_l);
            w4.w = pack_bf16(__uint_as_float(o[i * 8 + 6]) * inv_l, __uint_as_float(o[i * 8 + 7]) * inv_l);
            *reinterpret_cast<uint4*>(orow + c * 32 + i * 8) = w4;
            if (orow_keep) *reinterpret_cast<uint4*>(orow_keep + c * 32 + i * 8) = w4;
          }
        }
      }
      if (valid && p.lse && hf == 0)
        p.lse[stat_base(p, w, b, qhead) - w.off + grow] =
            (l_run > 0.f) ? (m_run + log2f(l_run)) * kLn2 : -INFINITY;
    }
  }
done:
  tc_fence_before();
  if (MC > 1) cluster_sync_all();
  else __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem, 256);
  }
}

// ============================================================================ forward, two heads per CTA (D = 128)
// Cycle accounting of attn_fwd128_kernel (SXB_ATTN_DEBUG=3) showed why two independent CTAs per SM do not give the
// ping-pong one hopes for: per 128-key block a CTA's row warps work 2145 cycles and wait 1646, its MMA warp issues for
// 1150 and waits 2500 — the two CTAs of an SM run in LOCKSTEP, both in their softmax phase (sharing the MUFU, 16
// exp2/clk/SM: 2 x 128 x 128 / 16 = 2048 cycles) and then both in their MMA phase (2 x 1024 cycles).  This kernel puts
// the two tiles into ONE CTA — two q heads of the same GQA group, same query rows, so they share every K/V tile — and
// lets the single MMA warp order the work: QK_A, QK_B | PV_A, QK_A' | PV_B, QK_B' | ...  While the rows of tile A
// turn S_A into P_A the tensor pipe does tile B's two MMAs, and vice versa: MUFU and tensor pipe each see 2048 cycles
// of work per pair of blocks and run concurrently.  K/V tiles are fetched once for both heads.
// TMEM: S_A 0..127, O_A 128..255, S_B 256..383, O_B 384..511 (a tile's row warps only touch its own two 128-column
// groups, and only while no MMA writes them).  18 warps: TMA, MMA, 8 row warps per tile (two per lane quadrant, 64 of
// the 128 key columns each).
constexpr int kFwdPairThreads = 576;
struct FwdPairSmem {
  static constexpr int kQ = BQ * 128 * 2;
  static constexpr int kKV = FBC * 128 * 2;
  static constexpr int kStages = 2;
  static constexpr int oQ = 0;                        // two Q tiles
  static constexpr int oK = oQ + 2 * kQ;
  static constexpr int oV = oK + kStages * kKV;
  static constexpr int oBar = oV + kStages * kKV;     // 192 KB
  static constexpr int oXch = oBar + 256;             // float [2 tiles][2 halves][128]
  static constexpr int kTotal = oXch + 2048 + 1024;
};

__global__ void __launch_bounds__(kFwdPairThreads, 1)
attn_fwd_pair_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                     const __grid_constant__ CUtensorMap tmV, const Params p) {
  using L = FwdPairSmem;
  constexpr int D = 128, NH = 2, NS = L::kStages;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::oBar);
  uint64_t* q_full = bars + 0;
  uint64_t* k_full = bars + 1;    // [2]
  uint64_t* v_full = bars + 3;    // [2]
  uint64_t* k_empty = bars + 5;   // [2] both tiles' QK^T of the block retired
  uint64_t* v_empty = bars + 7;   // [2] both tiles' PV of the block retired
  uint64_t* s_full = bars + 9;    // [2 tiles]
  uint64_t* p_full = bars + 11;   // [2 tiles] one arrival per row warp of the tile
  uint64_t* o_done = bars + 13;   // [2 tiles]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 15);
  // turn[t]: tile t's row warps may run their exponentials for their next block (8 arrivals: the other tile's warps).
  // Left alone the two tiles' softmax phases drift into lockstep, share the MUFU (each taking twice as long) and leave
  // the tensor pipe idle meanwhile; taking turns keeps one tile in its exponentials while the other's MMAs run.
  uint64_t* turn = bars + 17;     // [2]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.y;
  const int h0 = 2 * (int)blockIdx.x;              // heads h0 and h0 + 1: same K/V head (host checks the group is even)
  const int hk = h0 / (p.Hq / p.Hk);
  Win w;
  int q_row0, q_pos0, q_rows, kv_len;
  const int* pages = nullptr;
  if (p.atoms != nullptr) {
    const int* a = p.atoms + 6 * (int)blockIdx.z;
    w.off = 0; w.len = 0; w.tb = 0;
    q_row0 = a[1]; q_rows = a[2]; q_pos0 = a[3]; kv_len = a[5];
    pages = p.block_tables + (long long)a[0] * p.max_blocks;
  } else {
    const int q0 = (gridDim.z - 1 - blockIdx.z) * BQ;
    w = window(p, b);
    if (q0 >= w.len) return;
    q_row0 = w.off + q0; q_pos0 = q0; q_rows = min(BQ, w.len - q0);
    kv_len = p.causal ? min(w.len, q0 + BQ) : w.len;
  }
  const int nblk = (kv_len + FBC - 1) / FBC;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV);
  }
  if (warp == 1) {
    if (lane == 0) {
      mbar_init(q_full, 1);
      for (int i = 0; i < 2; ++i) {
        mbar_init(&k_full[i], 1); mbar_init(&v_full[i], 1); mbar_init(&k_empty[i], 1); mbar_init(&v_empty[i], 1);
        mbar_init(&s_full[i], 1); mbar_init(&p_full[i], 8); mbar_init(&o_done[i], 1);
        mbar_init(&turn[i], 8);
      }
      mbar_fence_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, 512);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    const bool elected = elect_one();
    if (elected) {
      mbar_expect_tx(q_full, 2 * L::kQ);
      for (int t = 0; t < 2; ++t)
        for (int hh = 0; hh < NH; ++hh)
          tma_load_4d(smem + L::oQ + t * L::kQ + hh * (BQ * 128), &tmQ, q_full, hh * 64, h0 + t, q_row0, w.tb);
    }
    __syncwarp();
    auto load_kv = [&](const CUtensorMap* tm, uint8_t* dst, uint64_t* bar, int j) {
      if (pages == nullptr) {
        for (int hh = 0; hh < NH; ++hh)
          tma_load_4d(dst + hh * (FBC * 128), tm, bar, hh * 64, hk, w.off + j * FBC, w.tb);
        return;
      }
      const int bs = p.block_size, sub = min(bs, FBC), last = (kv_len - 1) / bs;
      for (int r = 0; r < FBC; r += sub) {
        const int pos = j * FBC + r;
        const int blk = min(pos / bs, last);
        const int page = pages[blk], in_page = (blk == pos / bs) ? pos % bs : 0;
        for (int hh = 0; hh < NH; ++hh)
          tma_load_4d(dst + hh * (FBC * 128) + r * 128, tm, bar, hh * 64, hk, in_page, page);
      }
    };
    for (int j = 0; j < nblk; ++j) {
      const int st = j % NS;
      const uint32_t ph = ((j / NS) & 1) ^ 1;
      if (!__all_sync(0xffffffffu, mbar_wait_to(&k_empty[st], ph, 60, j))) goto done;
      if (elected) {
        mbar_expect_tx(&k_full[st], L::kKV);
        load_kv(&tmK, smem + L::oK + st * L::kKV, &k_full[st], j);
      }
      __syncwarp();
      if (!__all_sync(0xffffffffu, mbar_wait_to(&v_empty[st], ph, 61, j))) goto done;
      if (elected) {
        mbar_expect_tx(&v_full[st], L::kKV);
        load_kv(&tmV, smem + L::oV + st * L::kKV, &v_full[st], j);
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    const bool elected = elect_one();
    const uint32_t sbase = __shfl_sync(0xffffffffu, smem_u32(smem), 0);
    const uint32_t tm_u = __shfl_sync(0xffffffffu, tmem, 0);
    const uint32_t idesc_s = make_idesc(BQ, FBC, 1, 1, 0, 0);
    const uint32_t idesc_o = make_idesc(BQ, D, 1, 1, 0, 1);
    // S(t, j) = Q_t K(j)^T into the tile's S columns; overwrites P(t, j-1), which PV(t, j-1) — issued earlier by this
    // thread, MMAs execute in issue order — has consumed
    auto issue_qk = [&](int t, int j) {
      const uint32_t sQ = sbase + L::oQ + t * L::kQ, sK = sbase + L::oK + (j % NS) * L::kKV;
      if (elected) {
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk)
          umma_f16(tm_u + t * 256, desc_kmajor_d(sQ, BQ, kk), desc_kmajor_d(sK, FBC, kk), idesc_s, kk != 0);
        umma_commit(&s_full[t]);
      }
      __syncwarp();
    };
    auto issue_pv = [&](int t, int j) {
      const uint32_t sV = sbase + L::oV + (j % NS) * L::kKV;
      if (elected) {
#pragma unroll
        for (int kk = 0; kk < FBC / 16; ++kk)
          umma_f16_ts(tm_u + t * 256 + 128, tm_u + t * 256 + (kk >> 2) * 64 + (kk & 3) * 8, desc_mnmajor(sV, FBC, kk),
                      idesc_o, (j | kk) != 0);
        if (j == nblk - 1) umma_commit(&o_done[t]);
      }
      __syncwarp();
    };
    if (!__all_sync(0xffffffffu, mbar_wait_to(q_full, 0, 62))) goto done;
    if (!__all_sync(0xffffffffu, mbar_wait_to(&k_full[0], 0, 63, 0))) goto done;
    tc_fence_after();
    issue_qk(0, 0);
    issue_qk(1, 0);
    if (elected) umma_commit(&k_empty[0]);
    __syncwarp();
    const bool prof = kAttnProf && p.debug == 3 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0;
    long long pr_t0 = prof ? clock64() : 0, pr_kv = 0, pr_p0 = 0, pr_p1 = 0;
    for (int j = 0; j < nblk; ++j) {
      const int st = j % NS, stn = (j + 1) % NS;
      const bool more = j + 1 < nblk;
      long long c0 = prof ? clock64() : 0;
      if (!__all_sync(0xffffffffu, mbar_wait_to(&v_full[st], (j / NS) & 1, 64, j))) goto done;
      if (prof) { pr_kv += clock64() - c0; c0 = clock64(); }
      if (!__all_sync(0xffffffffu, mbar_wait_to(&p_full[0], j & 1, 65, j))) goto done;
      if (prof) pr_p0 += clock64() - c0;
      tc_fence_after();
      issue_pv(0, j);
      if (more) {
        c0 = prof ? clock64() : 0;
        if (!__all_sync(0xffffffffu, mbar_wait_to(&k_full[stn], ((j + 1) / NS) & 1, 63, j + 1))) goto done;
        if (prof) pr_kv += clock64() - c0;
        tc_fence_after();
        issue_qk(0, j + 1);
      }
      c0 = prof ? clock64() : 0;
      if (!__all_sync(0xffffffffu, mbar_wait_to(&p_full[1], j & 1, 66, j))) goto done;
      if (prof) pr_p1 += clock64() - c0;
      tc_fence_after();
      issue_pv(1, j);
      if (elected) umma_commit(&v_empty[st]);
      __syncwarp();
      if (more) {
        issue_qk(1, j + 1);
        if (elected) umma_commit(&k_empty[stn]);
        __syncwarp();
      }
    }
    if (prof && lane == 0) {
      g_attn_prof[0] = clock64() - pr_t0; g_attn_prof[1] = pr_kv; g_attn_prof[2] = pr_p0; g_attn_prof[3] = nblk;
      g_attn_prof[7] = pr_p1;
    }
  } else {
    const int t = (warp - 2) >> 3;               // tile (head h0 + t)
    const int quad = warp & 3;                   // TMEM lane quadrant (hardware: warp id mod 4)
    const int hf = ((warp - 2) & 7) >> 2;        // which 64 of the block's 128 key columns (and of O's 128)
    const int row = quad * 32 + lane;
    const uint32_t lane_off = uint32_t(quad * 32) << 16;
    const int qhead = h0 + t;
    const int qpos = q_pos0 + row;
    const float sc = p.scale * kLog2e;
    float* xch = reinterpret_cast<float*>(smem + L::oXch) + t * 256;     // [2][128]
    const uint32_t tS = tmem + t * 256, tO = tS + 128;
    const uint32_t tSh = tS + hf * 64 + lane_off;
    const int bar_id = 1 + t * 4 + quad;
    float m_run = -INFINITY, l_run = 0.f;
    const bool rprof = kAttnProf && p.debug == 3 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && warp == 2;
    long long rp_t0 = rprof ? clock64() : 0, rp_wait = 0, rp_work = 0;
    for (int j = 0; j < nblk; ++j) {
      // s_full[t](j) is a commit behind QK(t, j): PV(t, j-1) has retired, so O_t may be rescaled and S_t overwritten
      long long rc0 = rprof ? clock64() : 0;
      if (!mbar_wait_to(&s_full[t], j & 1, 67, j)) goto done;
      long long rc1 = rprof ? clock64() : 0;
      rp_wait += rc1 - rc0;
      tc_fence_after();
      const int kh = j * FBC + hf * 64;
      const bool need_mask = (p.causal && (kh + 63 > q_pos0)) || (kh + 64 > kv_len);
      float mx = -INFINITY;
#pragma unroll 1
      for (int q4 = 0; q4 < 2; ++q4) {
        uint32_t x[32];
        tmem_ld_32x32(tSh + q4 * 32, x);
        tmem_ld_wait();
        float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
        if (need_mask) {
#pragma unroll
          for (int c = 0; c < 32; ++c) {
            const int kp = kh + q4 * 32 + c;
            const bool ok = (kp < kv_len) && (!p.causal || kp <= qpos);
            mx4[c & 3] = fmaxf(mx4[c & 3], ok ? __uint_as_float(x[c]) : -INFINITY);
          }
        } else {
#pragma unroll
          for (int c = 0; c < 32; ++c) mx4[c & 3] = fmaxf(mx4[c & 3], __uint_as_float(x[c]));
        }
        mx = fmaxf(mx, fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3])));
      }
      xch[hf * 128 + row] = mx;
      named_bar_sync8(bar_id, 64);
      mx = fmaxf(mx, xch[(hf ^ 1) * 128 + row]) * sc;
      float m_new = m_run;
      const bool grow = mx > m_run + 8.f;
      const bool any_grow = __any_sync(0xffffffffu, grow) != 0;
      if (any_grow) m_new = fmaxf(m_run, mx);
      const float m_use = (m_new == -INFINITY) ? 0.f : m_new;
      if (j > 0 && any_grow) {
        const float alpha = (m_run == -INFINITY) ? 0.f : fast_exp2(m_run - m_use);
        l_run *= alpha;
#pragma unroll 1
        for (int c = hf * 2; c < hf * 2 + 2; ++c) {
          uint32_t o[32];
          tmem_ld_32x32(tO + lane_off + c * 32, o);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
          tmem_st_32x32(tO + lane_off + c * 32, o);
        }
        tmem_st_wait();
      }
      m_run = m_new;
      // my turn at the MUFU: tile 0 goes first; tile t's block j follows the other tile's block j - (t == 0)
      if (p.debug != 4 && (t == 1 || j > 0)) {
        if (!mbar_wait_to(&turn[t], (t == 1 ? j : j - 1) & 1, 69, j)) goto done;
      }
      float rs = 0.f;
#pragma unroll 1
      for (int q4 = 0; q4 < 2; ++q4) {
        uint32_t x[32];
        tmem_ld_32x32(tSh + q4 * 32, x);
        tmem_ld_wait();
        float e[32];
        float rs4[4] = {0.f, 0.f, 0.f, 0.f};
        if (need_mask) {
#pragma unroll
          for (int c = 0; c < 32; ++c) {
            const int kp = kh + q4 * 32 + c;
            const bool ok = (kp < kv_len) && (!p.causal || kp <= qpos);
            e[c] = ok ? fast_exp2(fmaf(__uint_as_float(x[c]), sc, -m_use)) : 0.f;
            rs4[c & 3] += e[c];
          }
        } else {
#pragma unroll
          for (int c = 0; c < 32; ++c) {
            e[c] = fast_exp2(fmaf(__uint_as_float(x[c]), sc, -m_use));
            rs4[c & 3] += e[c];
          }
        }
        rs += (rs4[0] + rs4[1]) + (rs4[2] + rs4[3]);
        uint32_t pk[16];
#pragma unroll
        for (int c = 0; c < 16; ++c) pk[c] = pack_bf16(e[2 * c], e[2 * c + 1]);
        tmem_st_32x16(tSh + q4 * 16, pk);
      }
      if (lane == 0) mbar_arrive(&turn[t ^ 1]);      // exponentials issued: the other tile may start its own
      tmem_st_wait();
      l_run += rs;
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_full[t]);
      if (rprof) rp_work += clock64() - rc1;
    }
    if (rprof && lane == 0) { g_attn_prof[4] = clock64() - rp_t0; g_attn_prof[5] = rp_wait; g_attn_prof[6] = rp_work; }
    if (!mbar_wait_to(&o_done[t], 0, 68)) goto done;
    tc_fence_after();
    xch[hf * 128 + row] = l_run;
    named_bar_sync8(bar_id, 64);
    l_run += xch[(hf ^ 1) * 128 + row];
    const float inv_l = (l_run > 0.f) ? 1.f / l_run : 0.f;
    {
      const bool valid = row < q_rows;
      const int grow = valid ? q_row0 + row : q_row0;
      __nv_bfloat16* orow = out_row_ptr(p, false, w.tb, grow) + (long long)qhead * D;
      __nv_bfloat16* orow_keep = (p.scatter_rows > 0 && p.out != nullptr)
          ? reinterpret_cast<__nv_bfloat16*>(p.out) + ((long long)w.tb * p.S + grow) * p.out2_stride +
                (long long)qhead * D
          : nullptr;
#pragma unroll 1
      for (int c = hf * 2; c < hf * 2 + 2; ++c) {
        uint32_t o[32];
        tmem_ld_32x32(tO + lane_off + c * 32, o);
        tmem_ld_wait();
        if (valid) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            uint4 w4;
            w4.x = pack_bf16(__uint_as_float(o[i * 8 + 0]) * inv_l, __uint_as_float(o[i * 8 + 1]) * inv_l);
            w4.y = pack_bf16(__uint_as_float(o[i * 8 + 2]) * inv_l, __uint_as_float(o[i * 8 + 3]) * inv_l);
            w4.z = pack_bf16(__uint_as_float(o[i * 8 + 4]) * inv_l, __uint_as_float(o[i * 8 + 5]) * inv_l);
            w4.w = pack_bf16(__uint_as_float(o[i * 8 + 6]) * inv_l, __uint_as_float(o[i * 8 + 7]) * inv_l);
            *reinterpret_cast<uint4*>(orow + c * 32 + i * 8) = w4;
            if (orow_keep) *reinterpret_cast<uint4*>(orow_keep + c * 32 + i * 8) = w4;
          }
        }
      }
      if (valid && p.lse && hf == 0)
        p.lse[stat_base(p, w, b, qhead) - w.off + grow] =
            (l_run > 0.f) ? (m_run + log2f(l_run)) * kLn2 : -INFINITY;
    }
  }
done:
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem, 512);
  }
}

// ===================================================================== backward: dQ
// One CTA per (128-query tile, head, batch); loops over 64-key blocks.  S / dP accumulators are
// double-buffered in TMEM so the tensor core computes block j+1 while the row threads turn block j
// into dS; K/V tiles ride a 3-deep TMA ring.
template <int D>
struct DqSmem {
  static constexpr int kQ = BQ * D * 2;      // Q and dO tiles
  static constexpr int kKV = BC * D * 2;
  // K/V ring.  Measured: deepening it to 6 stages (two reusing the Q / dO staging tiles once those are parked in TMEM)
  // changed nothing (0.74 -> 0.76 ms at S=8192): the MMA warp waits ~120 cycles per iteration for K/V.  What bounds
  // this kernel is the issue of the TS-form MMAs themselves: cycle accounting of one CTA (SXB_ATTN_DEBUG=3,
  // tools/attn_prof.py) gives ~1500 cycles per 64-key iteration = 20 MMAs x ~64 cycles + ~230 cycles of waits — an
  // A operand read from TMEM costs ~64 cycles per K=16 step whatever N is, so the N=64 S / dP MMAs run at half rate.
  static constexpr int kOwnStages = 4;
  static constexpr int kStages = 4;
  static constexpr int oQ = 0;
  static constexpr int oDO = oQ + kQ;
  static constexpr int oK = oDO + kQ;
  static constexpr int oV = oK + kOwnStages * kKV;
  static constexpr int oBar = oV + kOwnStages * kKV;
  static constexpr int kTotal = oBar + 256 + 1024;
  __host__ __device__ static constexpr int offK(int st) {
    return st < kOwnStages ? oK + st * kKV : (st == kOwnStages ? oQ : oDO);
  }
  __host__ __device__ static constexpr int offV(int st) {
    return st < kOwnStages ? oV + st * kKV : (st == kOwnStages ? oQ + kKV : oDO + kKV);
  }
};

template <int D, int MC>   // MC = 2: K/V ring filled by multicast from a 2-CTA cluster over a q-head pair (see forward)
__global__ void __launch_bounds__(kBwdThreads, 1)
attn_bwd_dq_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                   const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmDO,
                   const Params p) {
  using L = DqSmem<D>;
  constexpr int NH = D / 64;
  constexpr int NS = L::kStages;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::oBar);
  uint64_t* q_full = bars + 0;
  uint64_t* kv_full = bars + 17;  // [NS]
  uint64_t* kv_empty = bars + 23; // [NS]
  uint64_t* s_full = bars + 9;    // [2] S and dP of buffer b ready
  uint64_t* dq_done = bars + 12;  // committed once, after the last block's dQ MMAs
  // [2] dS of buffer b parked in TMEM (one arrival per row warp).  Row warps can finish block j+1 before the MMA warp
  // has observed block j, so the two in-flight blocks need distinct barriers (a single one would alias phases).
  uint64_t* ds_full = bars + 13;
  uint64_t* qa_full = bars + 15;  // Q / dO copied into TMEM as A operands (one arrival per row warp)
  uint64_t* qa_all = bars + 29;   // ... in every CTA of the cluster (one commit-arrival per CTA's MMA warp)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);
  static_assert(23 + NS <= 29 && NS >= L::kOwnStages, "barrier area is 256 bytes");

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int qt = gridDim.z - 1 - blockIdx.z;   // heaviest causal tiles first across all heads
  const int h = blockIdx.x, b = blockIdx.y;
  const Win w = window(p, b);
  const int S_ = w.len;
  const int hk = h / (p.Hq / p.Hk);
  const int q0 = qt * BQ;
  if (q0 >= S_) return;
  const int kv_len = p.causal ? min(S_, q0 + BQ) : S_;
  const int nblk = (kv_len + BC - 1) / BC;

  if (warp == 1) {
    if (lane == 0) {
      mbar_init(q_full, 1);
      for (int i = 0; i < NS; ++i) { mbar_init(&kv_full[i], 1); mbar_init(&kv_empty[i], MC); }
      mbar_init(&s_full[0], 1);
      mbar_init(&s_full[1], 1);
      mbar_init(&ds_full[0], 8);   // one arrival per row warp
      mbar_init(&ds_full[1], 8);
      mbar_init(qa_full, 8);
      mbar_init(qa_all, MC);
      mbar_init(dq_done, 1);
      mbar_fence_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, 512);
  }
  tc_fence_before();
  if (MC > 1) cluster_sync_all();
  else __syncthreads();
  tc_fence_after();
  const uint32_t crank = (MC > 1) ? cluster_ctarank() : 0;
  const uint32_t tmem = *tmem_slot;
  // buffer b: S at 128*b, dP at 128*b + 64 ; dQ at 256 ; Q as A operand at 384, dO at 448 (bf16 pairs, D/2 columns)
  const uint32_t tDQ = tmem + 256;

  if (warp == 0) {
    const bool elected = elect_one();
    if (elected) {
      mbar_expect_tx(q_full, 2 * L::kQ);
      for (int hh = 0; hh < NH; ++hh) {
        tma_load_4d(smem + L::oQ + hh * (BQ * 128), &tmQ, q_full, hh * 64, h, w.off + q0, w.tb);
        tma_load_4d(smem + L::oDO + hh * (BQ * 128), &tmDO, q_full, hh * 64, h, w.off + q0, w.tb);
      }
    }
    __syncwarp();
    for (int j = 0; j < nblk; ++j) {
      const int st = j % NS;
      if (!__all_sync(0xffffffffu, mbar_wait_to(&kv_empty[st], ((j / NS) & 1) ^ 1, 27, j))) goto done;
      // stages 4 and 5 live where Q and dO were staged: free once the row warps have copied them into TMEM — in
      // BOTH CTAs of a multicast pair, since each CTA's load lands in both (qa_all counts the peer's row warps too)
      if (NS > L::kOwnStages && j == L::kOwnStages && !__all_sync(0xffffffffu, mbar_wait_to(qa_all, 0, 44))) goto done;
      if (elected) {
        mbar_expect_tx(&kv_full[st], 2 * L::kKV);
        if (MC > 1) {
          tma_load_4d_mc(smem + L::offK(st) + crank * (BC * 128), &tmK, &kv_full[st], crank * 64, hk,
                         w.off + j * BC, w.tb, 3);
          tma_load_4d_mc(smem + L::offV(st) + crank * (BC * 128), &tmV, &kv_full[st], crank * 64, hk,
                         w.off + j * BC, w.tb, 3);
        } else {
          for (int hh = 0; hh < NH; ++hh) {
            tma_load_4d(smem + L::offK(st) + hh * (BC * 128), &tmK, &kv_full[st], hh * 64, hk,
                        w.off + j * BC, w.tb);
            tma_load_4d(smem + L::offV(st) + hh * (BC * 128), &tmV, &kv_full[st], hh * 64, hk,
                        w.off + j * BC, w.tb);
          }
        }
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    const bool elected = elect_one();
    const uint32_t sbase = __shfl_sync(0xffffffffu, smem_u32(smem), 0);
    const uint32_t tm_u = __shfl_sync(0xffffffffu, tmem, 0);
    const uint32_t tDQ_u = tm_u + 256;
    const uint32_t idesc_s = make_idesc(BQ, BC, 1, 1, 0, 0);
    const uint32_t idesc_dq = make_idesc(BQ, D, 1, 1, 0, 1);
    auto issue_sdp = [&](int j) -> bool {
      const int st = j % NS;
      if (!__all_sync(0xffffffffu, mbar_wait_to(&kv_full[st], (j / NS) & 1, 29, j))) return false;
      tc_fence_after();
      const uint32_t sK = sbase + L::offK(st);
      const uint32_t sV = sbase + L::offV(st);
      const uint32_t tS = tm_u + (j & 1) * 128, tDP = tS + 64;
      if (elected) {
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk)
          umma_f16_ts(tS, tm_u + 384 + kk * 8, desc_kmajor_d(sK, BC, kk), idesc_s, kk != 0);
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk)
          umma_f16_ts(tDP, tm_u + 448 + kk * 8, desc_kmajor_d(sV, BC, kk), idesc_s, kk != 0);
        umma_commit(&s_full[j & 1]);
      }
      __syncwarp();
      return true;
    };
    // Q and dO are the A operands of every S / dP MMA of this CTA: the row warps park them in TMEM once, so those
    // MMAs read only the K / V tile from shared memory (operand fetch from smem, not the tensor pipe, was the
    // measured limiter: ~70 B/clk/SM)
    if (!__all_sync(0xffffffffu, mbar_wait_to(qa_full, 0, 28))) goto done;
    tc_fence_after();
    if (elected) {
      // tell the TMA warps of the whole cluster that THIS CTA's Q / dO staging tiles may be overwritten
      if (MC > 1) umma_commit_mc(qa_all, 3);
      else umma_commit(qa_all);
    }
    __syncwarp();
    if (!issue_sdp(0)) goto done;
    const bool prof = kAttnProf && p.debug == 3 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0;
    long long pr_t0 = prof ? clock64() : 0, pr_kv = 0, pr_ds = 0;
    for (int j = 0; j < nblk; ++j) {
      const int st = j % NS;
      // buffer (j+1)&1 was last read for block j-1, whose ds_full this warp already waited on
      long long c0 = prof ? clock64() : 0;
      if (j + 1 < nblk && prof) { mbar_wait_to(&kv_full[(j + 1) % NS], ((j + 1) / NS) & 1, 29, j); pr_kv += clock64() - c0; }
      if (j + 1 < nblk && !issue_sdp(j + 1)) goto done;
      c0 = prof ? clock64() : 0;
      if (!__all_sync(0xffffffffu, mbar_wait_to(&ds_full[j & 1], (j >> 1) & 1, 30, j))) goto done;
      if (prof) pr_ds += clock64() - c0;
      tc_fence_after();
      const uint32_t sK = sbase + L::offK(st);
      if (elected) {
#pragma unroll
        for (int kk = 0; kk < BC / 16; ++kk)
          umma_f16_ts(tDQ_u, tm_u + (j & 1) * 128 + (kk >> 1) * 32 + (kk & 1) * 8, desc_mnmajor(sK, BC, kk),
                      idesc_dq, (j | kk) != 0);
        if (MC > 1) umma_commit_mc(&kv_empty[st], 3);
        else umma_commit(&kv_empty[st]);
        if (j == nblk - 1) umma_commit(dq_done);
      }
      __syncwarp();
    }
    if (prof && lane == 0) {
      g_attn_prof[0] = clock64() - pr_t0; g_attn_prof[1] = pr_kv; g_attn_prof[2] = pr_ds; g_attn_prof[3] = nblk;
    }
  } else {
    const int quad = warp & 3;                 // TMEM lane quadrant (hardware: warp id mod 4)
    const int half = (warp - 2) >> 2;          // which 32 of the block's 64 key columns
    const int row = quad * 32 + lane;
    const uint32_t lane_off = uint32_t(quad * 32) << 16;
    const int qpos = q0 + row;
    const float sc = p.scale * kLog2e;
    const long long stat_idx = stat_base(p, w, b, h) + min(qpos, S_ - 1);
    const float lse2 = p.lse[stat_idx] * kLog2e;
    const float dlt = p.delta[stat_idx];
    {
      // half 0 copies its Q rows, half 1 its dO rows: undo the 128B swizzle, 16 bytes = 4 packed columns at a time
      if (!mbar_wait_to(q_full, 0, 43)) goto done;
      const uint8_t* tile = smem + (half == 0 ? L::oQ : L::oDO) + row * 128;
      const uint32_t tA = tmem + 384 + half * 64 + lane_off;
#pragma unroll
      for (int hh = 0; hh < NH; ++hh) {
        uint32_t w[32];
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const uint4 v = *reinterpret_cast<const uint4*>(tile + hh * (BQ * 128) + ((c ^ (row & 7)) << 4));
          w[c * 4 + 0] = v.x; w[c * 4 + 1] = v.y; w[c * 4 + 2] = v.z; w[c * 4 + 3] = v.w;
        }
        tmem_st_32x32(tA + hh * 32, w);
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(qa_full);
    }
    const bool rprof = kAttnProf && p.debug == 3 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && warp == 2;
    long long rp_t0 = rprof ? clock64() : 0, rp_wait = 0, rp_math = 0, rp_st = 0;
    for (int j = 0; j < nblk; ++j) {
      long long rc0 = rprof ? clock64() : 0;
      if (!mbar_wait_to(&s_full[j & 1], (j >> 1) & 1, 31, j)) goto done;
      tc_fence_after();
      long long rc1 = rprof ? clock64() : 0;
      rp_wait += rc1 - rc0;
      float s[32], dp[32];
      const uint32_t tS = tmem + (j & 1) * 128 + lane_off + half * 32;
      if (p.debug < 2 || p.debug == 3) {
        ld_row32x2(tS, tS + 64, s, dp);
      } else {
#pragma unroll
        for (int c = 0; c < 32; ++c) { s[c] = 0.f; dp[c] = 0.f; }
      }
      const int k0 = j * BC + half * 32;
      if (p.kbias != nullptr) {
        const float4* kb = reinterpret_cast<const float4*>(p.kbias + b * p.kbias_bstride + h * p.kbias_hstride + k0);
        const int nv = min(8, (S_ - k0 + 3) >> 2);
#pragma unroll
        for (int c4 = 0; c4 < 8; ++c4) {
          if (c4 < nv) {
            const float4 t = __ldg(kb + c4);
            s[c4 * 4 + 0] += t.x; s[c4 * 4 + 1] += t.y; s[c4 * 4 + 2] += t.z; s[c4 * 4 + 3] += t.w;
          }
        }
      }
      const bool need_mask = (p.causal && (k0 + 31 > q0)) || (k0 + 32 > S_) || (q0 + BQ > S_);
      if (p.debug == 1 || p.debug == 2) {
        // profiling aid (SXB_ATTN_DEBUG): no exponentials, results are meaningless
#pragma unroll
        for (int c = 0; c < 32; ++c) s[c] = s[c] * dp[c];
      } else if (need_mask) {
#pragma unroll
        for (int c = 0; c < 32; ++c) {
          const int kp = k0 + c;
          const bool ok = (kp < S_) && (!p.causal || kp <= qpos) && (qpos < S_);
          const float pr = ok ? fast_exp2(fmaf(s[c], sc, -lse2)) : 0.f;
          s[c] = pr * (dp[c] - dlt) * p.scale;
        }
      } else {
#pragma unroll
        for (int c = 0; c < 32; ++c) {
          const float pr = fast_exp2(fmaf(s[c], sc, -lse2));
          s[c] = pr * (dp[c] - dlt) * p.scale;
        }
      }
      {
        // dS (bf16) is parked in TMEM over the first half of this warp's own S columns and consumed from there as
        // the A operand of dQ += dS K (no smem round trip, no async-proxy fence, one arrival per warp)
        uint32_t pk[16];
#pragma unroll
        for (int c = 0; c < 16; ++c) pk[c] = pack_bf16(s[2 * c], s[2 * c + 1]);
        long long rc2 = rprof ? clock64() : 0;
        rp_math += rc2 - rc1;
        tmem_st_32x16(tS, pk);
        tmem_st_wait();
        tc_fence_before();
        if (rprof) rp_st += clock64() - rc2;
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&ds_full[j & 1]);
    }
    if (rprof && lane == 0) {
      g_attn_prof[4] = clock64() - rp_t0; g_attn_prof[5] = rp_wait; g_attn_prof[6] = rp_math; g_attn_prof[7] = rp_st;
    }
    if (!mbar_wait_to(dq_done, 0, 33)) goto done;
    tc_fence_after();
    __nv_bfloat16* orow = out_row_ptr(p, false, w.tb, w.off + min(qpos, S_ - 1)) + (long long)h * D;
#pragma unroll 1
    for (int c = half * (D / 64); c < (half + 1) * (D / 64); ++c) {
      uint32_t o[32];
      tmem_ld_32x32(tDQ + lane_off + c * 32, o);
      tmem_ld_wait();
      if (qpos < S_) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          uint4 w;
          w.x = pack_bf16(__uint_as_float(o[i * 8 + 0]), __uint_as_float(o[i * 8 + 1]));
          w.y = pack_bf16(__uint_as_float(o[i * 8 + 2]), __uint_as_float(o[i * 8 + 3]));
          w.z = pack_bf16(__uint_as_float(o[i * 8 + 4]), __uint_as_float(o[i * 8 + 5]));
          w.w = pack_bf16(__uint_as_float(o[i * 8 + 6]), __uint_as_float(o[i * 8 + 7]));
          *reinterpret_cast<uint4*>(orow + c * 32 + i * 8) = w;
        }
      }
    }
  }
done:
  tc_fence_before();
  if (MC > 1) cluster_sync_all();
  else __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem, 512);
  }
}

// ================================================================== backward: dK, dV
// One CTA per (128-key tile, kv head, batch); loops over the q heads of the GQA group and over
// 64-query blocks.  Row threads own KEY rows (transposed formulation).  S^T / dP^T are
// double-buffered in TMEM, Q/dO tiles ride a 4-deep TMA ring, per-query statistics are prefetched
// one iteration ahead.
template <int D>
struct DkvSmem {
  static constexpr int kKV = BQ * D * 2;   // K and V tiles (128 keys)
  static constexpr int kQ = BC * D * 2;    // Q_i / dO_i tiles (64 queries)
  static constexpr int kStages = 4;         // P^T / dS^T live in TMEM, so the ring takes the smem they used
  static constexpr int oK = 0;
  static constexpr int oV = oK + kKV;
  static constexpr int oQ = oV + kKV;
  static constexpr int oDO = oQ + kStages * kQ;
  static constexpr int oStat = oDO + kStages * kQ;       // [stages][2][64] floats
  static constexpr int oBar = oStat + kStages * 2 * BC * 4;
  static constexpr int kTotal = oBar + 256 + 1024;
};

template <int D>
__global__ void __launch_bounds__(kBwdThreads, 1)
attn_bwd_dkv_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                    const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmDO,
                    const Params p) {
  using L = DkvSmem<D>;
  constexpr int NH = D / 64;
  constexpr int NS = L::kStages;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::oBar);
  uint64_t* kv_full = bars + 0;
  uint64_t* q_full = bars + 1;    // [NS]
  uint64_t* q_empty = bars + 5;   // [NS]
  uint64_t* s_full = bars + 9;    // [2] S^T(it) in TMEM
  uint64_t* dp_full = bars + 11;  // [2] dP^T(it) in TMEM
  uint64_t* acc_done = bars + 15;  // committed once, after the last iteration's dK MMAs
  uint64_t* stat_full = bars + 16;  // [NS] lse / delta of the stage's 64 queries staged in smem
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 20);
  // [2] each: P^T / dS^T of buffer b packed into TMEM (one arrival per row warp).  Row warps can finish iteration
  // it+1 before the MMA warp has observed iteration it, so consecutive iterations use distinct barriers.
  uint64_t* p_full = bars + 22;
  uint64_t* ds_full = bars + 24;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kt = blockIdx.z;                   // key tile 0 sees every query block: heaviest first
  const int hk = blockIdx.x, b = blockIdx.y;
  const Win w = window(p, b);
  const int S_ = w.len;
  const int group = p.Hq / p.Hk;
  const int n0 = kt * BQ;
  if (n0 >= S_) return;
  const int nq_blocks = (S_ + BC - 1) / BC;
  const int i_start = p.causal ? (n0 / BC) : 0;
  const int per_head = nq_blocks - i_start;
  const int niter = per_head * group;

  if (warp == 1) {
    if (lane == 0) {
      mbar_init(kv_full, 1);
      for (int i = 0; i < NS; ++i) { mbar_init(&q_full[i], 1); mbar_init(&q_empty[i], 1); }
      for (int i = 0; i < NS; ++i) mbar_init(&stat_full[i], 1);
      for (int i = 0; i < 2; ++i) { mbar_init(&s_full[i], 1); mbar_init(&dp_full[i], 1); }
      for (int i = 0; i < 2; ++i) { mbar_init(&p_full[i], 8); mbar_init(&ds_full[i], 8); }   // one arrival per row warp
      mbar_init(acc_done, 1);
      mbar_fence_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, 512);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  // buffer b: S^T at 128*b, dP^T at 128*b + 64 ; dV at 256 ; dK at 384
  const uint32_t tDV = tmem + 256, tDK = tmem + 384;

  if (warp == 0) {
    const bool elected = elect_one();
    if (elected) {
      mbar_expect_tx(kv_full, 2 * L::kKV);
      for (int hh = 0; hh < NH; ++hh) {
        tma_load_4d(smem + L::oK + hh * (BQ * 128), &tmK, kv_full, hh * 64, hk, w.off + n0, w.tb);
        tma_load_4d(smem + L::oV + hh * (BQ * 128), &tmV, kv_full, hh * 64, hk, w.off + n0, w.tb);
      }
    }
    __syncwarp();
    // The otherwise idle lanes of this warp stage the per-query statistics of each block (-lse*log2e and delta, 64
    // floats each) next to the Q/dO tiles: loaded one iteration ahead, written once the ring slot is free.
    float* stat = reinterpret_cast<float*>(smem + L::oStat);
    int st = 0, hrel = 0, qrel = 0;
    uint32_t ph = 1;
    auto stat_load = [&](int hh, int qr, float (&v)[4]) {
      const int qb = (i_start + qr) * BC;
      const long long base = stat_base(p, w, b, hk * group + hh);
      const int q0i = min(qb + lane, S_ - 1), q1i = min(qb + 32 + lane, S_ - 1);
      v[0] = -p.lse[base + q0i] * kLog2e;
      v[1] = -p.lse[base + q1i] * kLog2e;
      v[2] = p.delta[base + q0i];
      v[3] = p.delta[base + q1i];
    };
    float nxt[4] = {0.f, 0.f, 0.f, 0.f};
    if (niter > 0) stat_load(0, 0, nxt);
    for (int it = 0; it < niter; ++it) {
      const int h = hk * group + hrel;
      const int qi = i_start + qrel;
      if (!__all_sync(0xffffffffu, mbar_wait_to(&q_empty[st], ph, 34, it))) goto done;
      if (elected) {
        mbar_expect_tx(&q_full[st], 2 * L::kQ);
        for (int hh = 0; hh < NH; ++hh) {
          tma_load_4d(smem + L::oQ + st * L::kQ + hh * (BC * 128), &tmQ, &q_full[st], hh * 64, h,
                      w.off + qi * BC, w.tb);
          tma_load_4d(smem + L::oDO + st * L::kQ + hh * (BC * 128), &tmDO, &q_full[st], hh * 64, h,
                      w.off + qi * BC, w.tb);
        }
      }
      float* sp = stat + st * 128;
      sp[lane] = nxt[0];
      sp[32 + lane] = nxt[1];
      sp[64 + lane] = nxt[2];
      sp[96 + lane] = nxt[3];
      __syncwarp();
      if (lane == 0) mbar_arrive(&stat_full[st]);
      if (++qrel == per_head) { qrel = 0; ++hrel; }
      if (++st == NS) { st = 0; ph ^= 1; }
      if (it + 1 < niter) stat_load(hrel, qrel, nxt);
    }
  } else if (warp == 1) {
    const bool elected = elect_one();
    const uint32_t sbase = __shfl_sync(0xffffffffu, smem_u32(smem), 0);
    const uint32_t tm_u = __shfl_sync(0xffffffffu, tmem, 0);
    const uint32_t tDV_u = tm_u + 256, tDK_u = tm_u + 384;
    const uint32_t idesc_s = make_idesc(BQ, BC, 1, 1, 0, 0);
    const uint32_t idesc_acc = make_idesc(BQ, D, 1, 1, 0, 1);
    const uint32_t sK = sbase + L::oK, sV = sbase + L::oV;
    auto issue_first = [&](int it) -> bool {
      const int st = it % NS;
      if (!__all_sync(0xffffffffu, mbar_wait_to(&q_full[st], (it / NS) & 1, 36, it))) return false;
      tc_fence_after();
      const uint32_t sQ = sbase + L::oQ + st * L::kQ;
      const uint32_t sDO = sbase + L::oDO + st * L::kQ;
      const uint32_t tST = tm_u + (it & 1) * 128, tDPT = tST + 64;
      if (elected) {
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk)
          umma_f16(tST, desc_kmajor_d(sK, BQ, kk), desc_kmajor_d(sQ, BC, kk), idesc_s, kk != 0);
        umma_commit(&s_full[it & 1]);
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk)
          umma_f16(tDPT, desc_kmajor_d(sV, BQ, kk), desc_kmajor_d(sDO, BC, kk), idesc_s, kk != 0);
        umma_commit(&dp_full[it & 1]);
      }
      __syncwarp();
      return true;
    };
    if (!__all_sync(0xffffffffu, mbar_wait_to(kv_full, 0, 35))) goto done;
    if (niter > 0 && !issue_first(0)) goto done;
    for (int it = 0; it < niter; ++it) {
      const int st = it % NS;
      // S^T/dP^T(it+1) overwrite the TMEM buffer that held P^T/dS^T(it-1): safe without a barrier because the
      // dV/dK MMAs that read them were issued earlier by this same thread and the tensor pipe runs in order.
      if (it + 1 < niter && !issue_first(it + 1)) goto done;
      if (!__all_sync(0xffffffffu, mbar_wait_to(&p_full[it & 1], (it >> 1) & 1, 37, it))) goto done;
      tc_fence_after();
      const uint32_t sQ = sbase + L::oQ + st * L::kQ;
      const uint32_t sDO = sbase + L::oDO + st * L::kQ;
      // P^T (bf16, packed 2 per column) lives in the first 16 columns of each half of the S^T buffer, dS^T
      // likewise inside the dP^T buffer: k-step kk (16 queries) = 8 columns.
      const uint32_t tPT = tm_u + (it & 1) * 128, tDST = tPT + 64;
      if (elected) {
#pragma unroll
        for (int kk = 0; kk < BC / 16; ++kk)
          umma_f16_ts(tDV_u, tPT + (kk >> 1) * 32 + (kk & 1) * 8, desc_mnmajor(sDO, BC, kk), idesc_acc,
                      (it | kk) != 0);
      }
      __syncwarp();
      if (!__all_sync(0xffffffffu, mbar_wait_to(&ds_full[it & 1], (it >> 1) & 1, 39, it))) goto done;
      tc_fence_after();
      if (elected) {
#pragma unroll
        for (int kk = 0; kk < BC / 16; ++kk)
          umma_f16_ts(tDK_u, tDST + (kk >> 1) * 32 + (kk & 1) * 8, desc_mnmajor(sQ, BC, kk), idesc_acc,
                      (it | kk) != 0);
        umma_commit(&q_empty[st]);
        if (it == niter - 1) umma_commit(acc_done);
      }
      __syncwarp();
    }
  } else {
    const int quad = warp & 3;
    const int half = (warp - 2) >> 2;          // which 32 of the block's 64 query columns
    const int row = quad * 32 + lane;
    const uint32_t lane_off = uint32_t(quad * 32) << 16;
    const int kpos = n0 + row;
    const float sc = p.scale * kLog2e;
    // additive key bias of this thread's key for the q head being processed (raw-score units -> log2 domain)
    const float* kb_row = (p.kbias != nullptr && kpos < S_) ? p.kbias + b * p.kbias_bstride + kpos : nullptr;
    int hrel_r = 0;
    float kb2 = kb_row ? kb_row[(long long)(hk * group) * p.kbias_hstride] * sc : 0.f;
    const uint32_t stat_u = smem_u32(smem + L::oStat) + half * 128;   // this warp's 32 of the 64 queries
    int st = 0, qrel = 0;
    uint32_t ph = 0;
    for (int it = 0; it < niter; ++it) {
      const int qb = (i_start + qrel) * BC + half * 32;
      const uint32_t tST = tmem + (it & 1) * 128 + lane_off + half * 32;
      const uint32_t sstat = stat_u + st * 512;
      const bool need_mask = (p.causal && (n0 + BQ - 1 > qb)) || (qb + 32 > S_) || (n0 + BQ > S_);
      if (!mbar_wait_to(&stat_full[st], ph, 41, it)) goto done;
      if (!mbar_wait_to(&s_full[it & 1], (it >> 1) & 1, 38, it)) goto done;
      // Measured: starting on S^T while the dP^T MMAs of the same buffer are still in flight corrupts results
      // (tcgen05.st into a TMEM buffer the tensor core is accumulating into), so wait for both.
      if (!mbar_wait_to(&dp_full[it & 1], (it >> 1) & 1, 42, it)) goto done;
      tc_fence_after();
      float s[32];
      {
        uint32_t x[32];
        tmem_ld_32x32(tST, x);
        tmem_ld_wait();
        if (p.kbias == nullptr) {
#pragma unroll
          for (int c4 = 0; c4 < 8; ++c4) {
            const float4 nl = lds128(sstat + c4 * 16);          // -lse * log2e, warp-broadcast
            s[c4 * 4 + 0] = fast_exp2(fmaf(__uint_as_float(x[c4 * 4 + 0]), sc, nl.x));
            s[c4 * 4 + 1] = fast_exp2(fmaf(__uint_as_float(x[c4 * 4 + 1]), sc, nl.y));
            s[c4 * 4 + 2] = fast_exp2(fmaf(__uint_as_float(x[c4 * 4 + 2]), sc, nl.z));
            s[c4 * 4 + 3] = fast_exp2(fmaf(__uint_as_float(x[c4 * 4 + 3]), sc, nl.w));
          }
        } else {
#pragma unroll
          for (int c4 = 0; c4 < 8; ++c4) {                      // key bias: one more add per element
            const float4 nl = lds128(sstat + c4 * 16);
            s[c4 * 4 + 0] = fast_exp2(fmaf(__uint_as_float(x[c4 * 4 + 0]), sc, nl.x + kb2));
            s[c4 * 4 + 1] = fast_exp2(fmaf(__uint_as_float(x[c4 * 4 + 1]), sc, nl.y + kb2));
            s[c4 * 4 + 2] = fast_exp2(fmaf(__uint_as_float(x[c4 * 4 + 2]), sc, nl.z + kb2));
            s[c4 * 4 + 3] = fast_exp2(fmaf(__uint_as_float(x[c4 * 4 + 3]), sc, nl.w + kb2));
          }
        }
      }
      if (need_mask) {
#pragma unroll
        for (int c = 0; c < 32; ++c) {
          const int qp = qb + c;
          const bool ok = (qp < S_) && (kpos < S_) && (!p.causal || kpos <= qp);
          s[c] = ok ? s[c] : 0.f;
        }
      }
      {
        // bf16-pack this thread's 32 values of P^T and park them in TMEM over the first half of the very columns
        // they were computed from (own lanes, own column range: no cross-warp hazard)
        uint32_t pk[16];
#pragma unroll
        for (int c = 0; c < 16; ++c) pk[c] = pack_bf16(s[2 * c], s[2 * c + 1]);
        tmem_st_32x16(tST, pk);
        tmem_st_wait();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_full[it & 1]);             // dV += P^T dO can start
      {
        uint32_t x[32], dk[16];
        tmem_ld_32x32(tST + 64, x);
        tmem_ld_wait();
#pragma unroll
        for (int c4 = 0; c4 < 8; ++c4) {
          const float4 dl = lds128(sstat + 256 + c4 * 16);     // delta
          const float d0 = s[c4 * 4 + 0] * p.scale * (__uint_as_float(x[c4 * 4 + 0]) - dl.x);
          const float d1 = s[c4 * 4 + 1] * p.scale * (__uint_as_float(x[c4 * 4 + 1]) - dl.y);
          const float d2 = s[c4 * 4 + 2] * p.scale * (__uint_as_float(x[c4 * 4 + 2]) - dl.z);
          const float d3 = s[c4 * 4 + 3] * p.scale * (__uint_as_float(x[c4 * 4 + 3]) - dl.w);
          dk[c4 * 2 + 0] = pack_bf16(d0, d1);
          dk[c4 * 2 + 1] = pack_bf16(d2, d3);
        }
        tmem_st_32x16(tST + 64, dk);
        tmem_st_wait();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&ds_full[it & 1]);
      if (++qrel == per_head) {
        qrel = 0;
        ++hrel_r;                                         // next q head of the group: its own bias row
        if (kb_row && p.kbias_hstride != 0 && hrel_r < group)
          kb2 = kb_row[(long long)(hk * group + hrel_r) * p.kbias_hstride] * sc;
      }
      if (++st == NS) { st = 0; ph ^= 1; }
    }
    if (niter > 0) {
      if (!mbar_wait_to(acc_done, 0, 40)) goto done;
      tc_fence_after();
    }
    // half 0 drains dV, half 1 drains dK
    const bool is_dv = (half == 0);
    __nv_bfloat16* dst_row = out_row_ptr(p, is_dv, w.tb, w.off + min(kpos, S_ - 1)) + (long long)hk * D;
    const uint32_t t_acc = (is_dv ? tDV : tDK) + lane_off;
#pragma unroll 1
    for (int cc = 0; cc < D / 32; ++cc) {
      uint32_t o[32];
      if (niter > 0) {
        tmem_ld_32x32(t_acc + cc * 32, o);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) o[i] = 0u;
      }
      if (kpos < S_) {
        __nv_bfloat16* dst = dst_row + cc * 32;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          uint4 w;
          w.x = pack_bf16(__uint_as_float(o[i * 8 + 0]), __uint_as_float(o[i * 8 + 1]));
          w.y = pack_bf16(__uint_as_float(o[i * 8 + 2]), __uint_as_float(o[i * 8 + 3]));
          w.z = pack_bf16(__uint_as_float(o[i * 8 + 4]), __uint_as_float(o[i * 8 + 5]));
          w.w = pack_bf16(__uint_as_float(o[i * 8 + 6]), __uint_as_float(o[i * 8 + 7]));
          *reinterpret_cast<uint4*>(dst + i * 8) = w;
        }
      }
    }
  }
done:
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem, 512);
  }
}

// delta[b,h,s] = sum_d dO[b,s,h,d] * O[b,s,h,d]   (one warp per (token, head))
__global__ void __launch_bounds__(256)
attn_delta_kernel(const __nv_bfloat16* __restrict__ o, const __nv_bfloat16* __restrict__ d_o,
                  float* __restrict__ delta, int B, int S, int H, int D, long long o_stride,
                  long long do_stride) {
  const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const long long total = (long long)B * S * H;
  if (w >= total) return;
  const int h = (int)(w % H);
  const long long t = w / H;  // b*S + s
  const __nv_bfloat16* op = o + t * o_stride + (long long)h * D;
  const __nv_bfloat16* dp = d_o + t * do_stride + (long long)h * D;
  float acc = 0.f;
  for (int d = lane * 2; d < D; d += 64) {
    const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(op + d);
    const __nv_bfloat162 g = *reinterpret_cast<const __nv_bfloat162*>(dp + d);
    acc += __low2float(a) * __low2float(g) + __high2float(a) * __high2float(g);
  }
  acc = warp_sum(acc);
  if (lane == 0) {
    const long long bb = t / S, s = t % S;
    delta[(bb * H + h) * S + s] = acc;
  }
}

static int make_qkv_map(CUtensorMap* m, const void* ptr, int B, int S, int H, int D,
                        long long row_stride, int box_rows, int box_heads = 1) {
  unsigned long long dims[4] = {(unsigned long long)D, (unsigned long long)H,
                                (unsigned long long)S, (unsigned long long)B};
  unsigned long long strides[3] = {(unsigned long long)D * 2, (unsigned long long)row_stride * 2,
                                   (unsigned long long)row_stride * 2 * (unsigned long long)S};
  unsigned int box[4] = {64, (unsigned int)box_heads, (unsigned int)box_rows, 1};
  return encode_tensor_map(m, kBF16, 4, ptr, dims, strides, box, 128);
}

template <typename K>
static int set_smem(K kernel, int bytes) {
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  return e == cudaSuccess ? SXB_OK : (int)e;
}

// Launch `kernel` as 2x1x1 thread-block clusters (grid.x even).
template <typename K, typename... Args>
static int launch_cluster2(K kernel, dim3 grid, int threads, int smem_bytes, cudaStream_t s, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = dim3(threads);
  cfg.dynamicSmemBytes = smem_bytes;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 2;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kernel, args...);
  return e == cudaSuccess ? SXB_OK : (int)e;
}

// Head-pair multicast needs D = 128 (two 64-wide halves to split) and an even GQA group (pair shares its K/V head).
// SXB_ATTN_MC=0 forces the single-CTA kernels (A/B measurements).
// TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from another
// GPU's memory hung the GPU: K/V that are not plain allocations of the current device take the single-CTA kernels.
// The test itself (runtime attributes + driver allocation range + foreign-range registry) is range_is_local().
static bool use_multicast(int Hq, int Hk, int D, const void* k, long long k_bytes, const void* v, long long v_bytes) {
  static const int env = [] { const char* e = getenv("SXB_ATTN_MC"); return e ? atoi(e) : 1; }();
  return env != 0 && D == 128 && Hk > 0 && ((Hq / Hk) % 2 == 0) && (Hq % 2 == 0) &&
         range_is_local(k, k_bytes) && range_is_local(v, v_bytes);
}
// bytes spanned by a [B, S, H, D] bf16 tensor with row pitch `stride` elements
static long long span_bytes(int B, int S, int H, int D, long long stride) {
  return (((long long)B * S - 1) * stride + (long long)H * D) * 2;
}

}  // namespace attn
}  // namespace sxb

using namespace sxb;
using namespace sxb::attn;

SXB_DEFINE_HANG_ACCESSOR(sxb_attn_hang_info)

// SXB_ATTN_DEBUG=3 cycle accounting of the last dQ launch (8 x uint64, host pointer); see g_attn_prof.
SXB_API int sxb_attn_prof(unsigned long long* out_host) {
  cudaError_t e = cudaMemcpyFromSymbol(out_host, sxb::attn::g_attn_prof, sizeof(unsigned long long) * 8);
  return e == cudaSuccess ? SXB_OK : (int)e;
}

// Which kernel family sxb_attn_fwd / the dQ pass would pick for these K/V: 1 = 2-CTA multicast, 0 = single CTA.
SXB_API int sxb_attn_would_multicast(const void* k, const void* v, int B, int S, int Hq, int Hk, int D,
                                     long long k_stride, long long v_stride) {
  return use_multicast(Hq, Hk, D, k, span_bytes(B, S, Hk, D, k_stride), v, span_bytes(B, S, Hk, D, v_stride)) ? 1 : 0;
}

// SXB_ATTN_FWD128: forward kernel family for head_dim 128 — 0: 64-key blocks (attn_fwd_kernel), 1: 128-key blocks,
// one head per CTA (attn_fwd128_kernel), 2: 128-key blocks, two heads of a GQA group per CTA (attn_fwd_pair_kernel).
static int fwd_mode() {
  static const int env = [] { const char* e = getenv("SXB_ATTN_FWD128"); return e ? atoi(e) : 0; }();
  return env;
}
static bool use_fwd128(int D) { return fwd_mode() != 0 && D == 128; }
static bool use_fwd_pair(int D, int Hq, int Hk) {
  return fwd_mode() == 2 && D == 128 && Hk > 0 && (Hq % 2) == 0 && ((Hq / Hk) % 2) == 0;
}

static bool attn_args_ok(int B, int S, int Hq, int Hk, int D, long long qs, long long ks,
                         long long vs) {
  if (B <= 0 || S <= 0 || Hq <= 0 || Hk <= 0 || Hq % Hk) return false;
  if (D != 64 && D != 128) return false;
  if ((qs % 8) || (ks % 8) || (vs % 8)) return false;
  return true;
}

// Scatter configuration of the next launches on THIS host thread (set by the *_scatter entry points, cleared after).
namespace {
struct ScatterCfg {
  int rows = 0;
  void* out[16] = {};
  void* out2[16] = {};
};
thread_local ScatterCfg g_scatter;
thread_local long long g_dv_delta = 0;
thread_local void* g_keep = nullptr;
thread_local long long g_keep_stride = 0;
// Packed-sequence configuration of the next launches on this host thread (set by the *_varlen entry points).
struct VarlenCfg {
  const int* cu = nullptr;
  int nseq = 0, max_len = 0;
};
thread_local VarlenCfg g_varlen;
thread_local const float* g_kbias = nullptr;   // set by the *_kbias entry points for the launches they make
thread_local long long g_kbias_bstride = 0, g_kbias_hstride = 0;
// grid over (heads, sequences, tiles of the longest sequence)
dim3 attn_grid(int heads, int B, int S) {
  if (g_varlen.cu != nullptr) return dim3(heads, g_varlen.nseq, (g_varlen.max_len + sxb::attn::BQ - 1) / sxb::attn::BQ);
  return dim3(heads, B, (S + sxb::attn::BQ - 1) / sxb::attn::BQ);
}
void apply_scatter(sxb::attn::Params& p, bool second_too) {
  p.scatter_rows = g_scatter.rows;
  for (int i = 0; i < 16; ++i) {
    p.out_peers[i] = g_scatter.out[i];
    p.out2_peers[i] = second_too ? g_scatter.out2[i] : nullptr;
  }
}
}  // namespace

// bf16 flash attention forward.  q/k/v/out: [B, S, H, D] with row pitches (elements) q_stride...
// lse: [B, Hq, S] fp32 (may be NULL).
SXB_API int sxb_attn_fwd(const void* q, const void* k, const void* v, void* out, void* lse, int B,
                         int S, int Hq, int Hk, int D, long long q_stride, long long k_stride,
                         long long v_stride, long long o_stride, int causal, float scale,
                         void* stream) {
  if (!attn_args_ok(B, S, Hq, Hk, D, q_stride, k_stride, v_stride) || (o_stride % 8))
    return SXB_ERR_UNSUPPORTED;
  CUtensorMap tq, tk, tv;
  int rc;
  const bool wide = use_fwd128(D) && g_kbias == nullptr;       // 128-key blocks (attn_fwd128_kernel); no key bias there
  const int kv_rows = wide ? FBC : BC;
  if ((rc = make_qkv_map(&tq, q, B, S, Hq, D, q_stride, BQ))) return rc;
  if ((rc = make_qkv_map(&tk, k, B, S, Hk, D, k_stride, kv_rows))) return rc;
  if ((rc = make_qkv_map(&tv, v, B, S, Hk, D, v_stride, kv_rows))) return rc;
  Params p{};
  p.B = B; p.S = S; p.Hq = Hq; p.Hk = Hk; p.D = D; p.causal = causal; p.scale = scale;
  p.out = out; p.out_stride = o_stride; p.lse = (float*)lse;
  p.cu_seqlens = g_varlen.cu;
  p.kbias = g_kbias; p.kbias_bstride = g_kbias_bstride; p.kbias_hstride = g_kbias_hstride;
  {
    static const int dbg = [] { const char* e = getenv("SXB_ATTN_DEBUG"); return e ? atoi(e) : 0; }();
    p.debug = dbg;
  }
  apply_scatter(p, false);
  if (p.scatter_rows > 0) { p.out = g_keep; p.out2_stride = g_keep_stride; }
  dim3 grid = attn_grid(Hq, B, S);
  cudaStream_t s = (cudaStream_t)stream;
  if (wide && use_fwd_pair(D, Hq, Hk)) {
    grid.x = Hq / 2;
    if ((rc = set_smem(attn_fwd_pair_kernel, FwdPairSmem::kTotal))) return rc;
    attn_fwd_pair_kernel<<<grid, kFwdPairThreads, FwdPairSmem::kTotal, s>>>(tq, tk, tv, p);
    SXB_CHECK_LAUNCH();
    return SXB_OK;
  }
  if (wide) {
    if (use_multicast(Hq, Hk, D, k, span_bytes(B, S, Hk, D, k_stride), v, span_bytes(B, S, Hk, D, v_stride))) {
      if ((rc = set_smem(attn_fwd128_kernel<2>, Fwd128Smem::kTotal))) return rc;
      if ((rc = launch_cluster2(attn_fwd128_kernel<2>, grid, kFwd128Threads, Fwd128Smem::kTotal, s, tq, tk, tv, p))) return rc;
    } else {
      if ((rc = set_smem(attn_fwd128_kernel<1>, Fwd128Smem::kTotal))) return rc;
      attn_fwd128_kernel<1><<<grid, kFwd128Threads, Fwd128Smem::kTotal, s>>>(tq, tk, tv, p);
    }
    SXB_CHECK_LAUNCH();
    return SXB_OK;
  }
  if (use_multicast(Hq, Hk, D, k, span_bytes(B, S, Hk, D, k_stride), v, span_bytes(B, S, Hk, D, v_stride))) {
    if ((rc = set_smem(attn_fwd_kernel<128, 2>, FwdSmem<128>::kTotal))) return rc;
    if ((rc = launch_cluster2(attn_fwd_kernel<128, 2>, grid, kThreads, FwdSmem<128>::kTotal, s, tq, tk, tv, p)))
      return rc;
  } else if (D == 128) {
    if ((rc = set_smem(attn_fwd_kernel<128, 1>, FwdSmem<128>::kTotal))) return rc;
    attn_fwd_kernel<128, 1><<<grid, kThreads, FwdSmem<128>::kTotal, s>>>(tq, tk, tv, p);
  } else {
    if ((rc = set_smem(attn_fwd_kernel<64, 1>, FwdSmem<64>::kTotal))) return rc;
    attn_fwd_kernel<64, 1><<<grid, kThreads, FwdSmem<64>::kTotal, s>>>(tq, tk, tv, p);
  }
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

// Backward.  delta: [B, Hq, S] fp32 scratch written here; dq/dk/dv: [B, S, H, D] bf16.
SXB_API int sxb_attn_bwd(const void* q, const void* k, const void* v, const void* out,
                         const void* d_out, const void* lse, void* delta, void* dq, void* dk,
                         void* dv, int B, int S, int Hq, int Hk, int D, long long q_stride,
                         long long k_stride, long long v_stride, long long o_stride,
                         long long do_stride, long long dq_stride, long long dk_stride,
                         long long dv_stride, int causal, float scale, void* stream) {
  if (!attn_args_ok(B, S, Hq, Hk, D, q_stride, k_stride, v_stride) || (o_stride % 8) ||
      (do_stride % 8) || (dq_stride % 8) || (dk_stride % 8) || (dv_stride % 8))
    return SXB_ERR_UNSUPPORTED;
  cudaStream_t s = (cudaStream_t)stream;
  {
    const long long warps = (long long)B * S * Hq;
    const long long blocks = (warps * 32 + 255) / 256;
    attn_delta_kernel<<<(unsigned)blocks, 256, 0, s>>>((const __nv_bfloat16*)out,
                                                       (const __nv_bfloat16*)d_out, (float*)delta,
                                                       B, S, Hq, D, o_stride, do_stride);
    SXB_CHECK_LAUNCH();
  }
  int rc;
  Params p{};
  p.B = B; p.S = S; p.Hq = Hq; p.Hk = Hk; p.D = D; p.causal = causal; p.scale = scale;
  p.lse = (float*)const_cast<void*>(lse); p.delta = (const float*)delta;
  p.cu_seqlens = g_varlen.cu;
  p.kbias = g_kbias; p.kbias_bstride = g_kbias_bstride; p.kbias_hstride = g_kbias_hstride;
  {
    static int dbg = -1;
    if (dbg < 0) { const char* e = getenv("SXB_ATTN_DEBUG"); dbg = e ? atoi(e) : 0; }
    p.debug = dbg;
  }
  // ---- dQ
  {
    CUtensorMap tq, tk, tv, tdo;
    if ((rc = make_qkv_map(&tq, q, B, S, Hq, D, q_stride, BQ))) return rc;
    if ((rc = make_qkv_map(&tdo, d_out, B, S, Hq, D, do_stride, BQ))) return rc;
    if ((rc = make_qkv_map(&tk, k, B, S, Hk, D, k_stride, BC))) return rc;
    if ((rc = make_qkv_map(&tv, v, B, S, Hk, D, v_stride, BC))) return rc;
    p.out = dq; p.out_stride = dq_stride;
    apply_scatter(p, false);
    dim3 grid = attn_grid(Hq, B, S);
    if (use_multicast(Hq, Hk, D, k, span_bytes(B, S, Hk, D, k_stride), v, span_bytes(B, S, Hk, D, v_stride))) {
      if ((rc = set_smem(attn_bwd_dq_kernel<128, 2>, DqSmem<128>::kTotal))) return rc;
      if ((rc = launch_cluster2(attn_bwd_dq_kernel<128, 2>, grid, kBwdThreads, DqSmem<128>::kTotal, s, tq, tk, tv,
                                tdo, p)))
        return rc;
    } else if (D == 128) {
      if ((rc = set_smem(attn_bwd_dq_kernel<128, 1>, DqSmem<128>::kTotal))) return rc;
      attn_bwd_dq_kernel<128, 1><<<grid, kBwdThreads, DqSmem<128>::kTotal, s>>>(tq, tk, tv, tdo, p);
    } else {
      if ((rc = set_smem(attn_bwd_dq_kernel<64, 1>, DqSmem<64>::kTotal))) return rc;
      attn_bwd_dq_kernel<64, 1><<<grid, kBwdThreads, DqSmem<64>::kTotal, s>>>(tq, tk, tv, tdo, p);
    }
    SXB_CHECK_LAUNCH();
  }
  // ---- dK, dV
  {
    CUtensorMap tq, tk, tv, tdo;
    if ((rc = make_qkv_map(&tq, q, B, S, Hq, D, q_stride, BC))) return rc;
    if ((rc = make_qkv_map(&tdo, d_out, B, S, Hq, D, do_stride, BC))) return rc;
    if ((rc = make_qkv_map(&tk, k, B, S, Hk, D, k_stride, BQ))) return rc;
    if ((rc = make_qkv_map(&tv, v, B, S, Hk, D, v_stride, BQ))) return rc;
    p.out = dk; p.out_stride = dk_stride; p.out2 = dv; p.out2_stride = dv_stride;
    if (g_scatter.rows > 0) {
      // dK / dV go to the same peers as dQ, at their own column offsets: the caller passes the dK bases in out2[]
      // and the dV bases are dK + dv_col_delta (see sxb_attn_bwd_scatter)
      p.scatter_rows = g_scatter.rows;
      for (int i = 0; i < 16; ++i) {
        p.out_peers[i] = g_scatter.out2[i];
        p.out2_peers[i] = g_scatter.out2[i] ? (void*)((__nv_bfloat16*)g_scatter.out2[i] + g_dv_delta) : nullptr;
      }
    }
    dim3 grid = attn_grid(Hk, B, S);
    if (D == 128) {
      if ((rc = set_smem(attn_bwd_dkv_kernel<128>, DkvSmem<128>::kTotal))) return rc;
      attn_bwd_dkv_kernel<128><<<grid, kBwdThreads, DkvSmem<128>::kTotal, s>>>(tq, tk, tv, tdo, p);
    } else {
      if ((rc = set_smem(attn_bwd_dkv_kernel<64>, DkvSmem<64>::kTotal))) return rc;
      attn_bwd_dkv_kernel<64><<<grid, kBwdThreads, DkvSmem<64>::kTotal, s>>>(tq, tk, tv, tdo, p);
    }
    SXB_CHECK_LAUNCH();
  }
  return SXB_OK;
}


// K7 (Ulysses): forward attention whose epilogue stores each output row straight into the buffer of the rank that
// owns that part of the sequence.  out_peers: host array of `world` device pointers — peer r's destination for the
// rows [r*rows_per_peer, (r+1)*rows_per_peer) of this rank's heads (column offset already applied), row pitch
// o_stride.  `out` is unused.  (reference: all_to_all_single after the local attention, sequence/layer.py:434-437)
SXB_API int sxb_attn_fwd_scatter(const void* q, const void* k, const void* v, void* const* out_peers, int world,
                                 int rows_per_peer, void* out_keep, long long keep_stride, void* lse, int B, int S,
                                 int Hq, int Hk, int D, long long q_stride, long long k_stride, long long v_stride,
                                 long long o_stride, int causal, float scale, void* stream) {
  // out_keep (may be NULL): local [B, S, Hq, D] copy of O with row pitch keep_stride, for the backward pass
  if (world < 1 || world > 16 || rows_per_peer <= 0 || (long long)rows_per_peer * world < S) return SXB_ERR_BAD_ARG;
  if (keep_stride % 8) return SXB_ERR_UNSUPPORTED;
  g_scatter = ScatterCfg{};
  g_scatter.rows = rows_per_peer;
  for (int i = 0; i < world; ++i) g_scatter.out[i] = out_peers[i];
  g_keep = out_keep;
  g_keep_stride = keep_stride;
  const int rc = sxb_attn_fwd(q, k, v, out_peers[0], lse, B, S, Hq, Hk, D, q_stride, k_stride, v_stride, o_stride,
                              causal, scale, stream);
  g_scatter = ScatterCfg{};
  g_keep = nullptr;
  return rc;
}

// Backward with scattered dQ / dK / dV: dq_peers[r] / dk_peers[r] are peer r's destinations (column offsets of this
// rank's heads applied) with row pitches dq_stride / dk_stride; dV lands dv_col_delta elements after dK with pitch
// dv_stride.
SXB_API int sxb_attn_bwd_scatter(const void* q, const void* k, const void* v, const void* out, const void* d_out,
                                 const void* lse, void* delta, void* const* dq_peers, void* const* dk_peers,
                                 long long dv_col_delta, int world, int rows_per_peer, int B, int S, int Hq, int Hk,
                                 int D, long long q_stride, long long k_stride, long long v_stride, long long o_stride,
                                 long long do_stride, long long dq_stride, long long dk_stride, long long dv_stride,
                                 int causal, float scale, void* stream) {
  if (world < 1 || world > 16 || rows_per_peer <= 0 || (long long)rows_per_peer * world < S) return SXB_ERR_BAD_ARG;
  g_scatter = ScatterCfg{};
  g_scatter.rows = rows_per_peer;
  for (int i = 0; i < world; ++i) { g_scatter.out[i] = dq_peers[i]; g_scatter.out2[i] = dk_peers[i]; }
  g_dv_delta = dv_col_delta;
  const int rc = sxb_attn_bwd(q, k, v, out, d_out, lse, delta, dq_peers[0], dk_peers[0],
                              (__nv_bfloat16*)dk_peers[0] + dv_col_delta, B, S, Hq, Hk, D, q_stride, k_stride, v_stride,
                              o_stride, do_stride, dq_stride, dk_stride, dv_stride, causal, scale, stream);
  g_scatter = ScatterCfg{};
  g_dv_delta = 0;
  return rc;
}

// Packed sequences (cu_seqlens int32 [nseq + 1] on the device; q/k/v/out [total, H, D]; lse / delta [Hq, total]):
// every sequence attends only to itself.  The reference reaches this through flash-attn's varlen entry points
// (sequence/fpdt_layer.py:14-20, inference/v2 blocked_flash); here it is the same three kernels with a per-CTA
// row window.
SXB_API int sxb_attn_fwd_varlen(const void* q, const void* k, const void* v, void* out, void* lse,
                                const void* cu_seqlens, int nseq, int max_seqlen, int total, int Hq, int Hk, int D,
                                long long q_stride, long long k_stride, long long v_stride, long long o_stride,
                                int causal, float scale, void* stream) {
  if (cu_seqlens == nullptr || nseq <= 0 || max_seqlen <= 0 || max_seqlen > total) return SXB_ERR_BAD_ARG;
  g_varlen.cu = static_cast<const int*>(cu_seqlens);
  g_varlen.nseq = nseq;
  g_varlen.max_len = max_seqlen;
  const int rc = sxb_attn_fwd(q, k, v, out, lse, 1, total, Hq, Hk, D, q_stride, k_stride, v_stride, o_stride, causal,
                              scale, stream);
  g_varlen = VarlenCfg{};
  return rc;
}

SXB_API int sxb_attn_bwd_varlen(const void* q, const void* k, const void* v, const void* out, const void* d_out,
                                const void* lse, void* delta, void* dq, void* dk, void* dv, const void* cu_seqlens,
                                int nseq, int max_seqlen, int total, int Hq, int Hk, int D, long long q_stride,
                                long long k_stride, long long v_stride, long long o_stride, long long do_stride,
                                long long dq_stride, long long dk_stride, long long dv_stride, int causal, float scale,
                                void* stream) {
  if (cu_seqlens == nullptr || nseq <= 0 || max_seqlen <= 0 || max_seqlen > total) return SXB_ERR_BAD_ARG;
  g_varlen.cu = static_cast<const int*>(cu_seqlens);
  g_varlen.nseq = nseq;
  g_varlen.max_len = max_seqlen;
  const int rc = sxb_attn_bwd(q, k, v, out, d_out, lse, delta, dq, dk, dv, 1, total, Hq, Hk, D, q_stride, k_stride,
                              v_stride, o_stride, do_stride, dq_stride, dk_stride, dv_stride, causal, scale, stream);
  g_varlen = VarlenCfg{};
  return rc;
}

// Blocked ("paged") flash attention over a ragged batch on the tensor cores: the forward kernel above, one CTA per
// (attention atom, q head).  q/out [T, Hq, D]; cache [n_blocks, block_size, 2, Hk, D] (K then V per slot); atoms
// int32 [n_atoms, 6] and block_tables int32 [n_seqs, max_blocks] on the device.  block_size: a multiple of 8 that
// divides 64, or a multiple of 64.  Causal by position: query i of an atom sits at global_q_idx + i.
// (reference: inference/v2/kernels/ragged_ops/blocked_flash/blocked_flash.cpp:21-101 over flash-attn's sm_80 kernels)
// n_packed: the LAST n_packed atoms of the table have at most 128 / (Hq / Hk) queries each and run with the GQA group
// packed into the tile (one CTA per KV head); 0 = none.
SXB_API int sxb_attn_fwd_atoms(const void* q, const void* cache, void* out, const void* atoms, int n_atoms,
                               int n_packed, const void* block_tables, int max_blocks, int n_blocks, int block_size,
                               int T, int Hq, int Hk, int D, long long q_stride, long long o_stride, float scale,
                               void* stream) {
  if (n_atoms <= 0) return SXB_OK;
  const int group = Hk > 0 ? Hq / Hk : 0;
  if (n_packed < 0 || n_packed > n_atoms || (n_packed > 0 && (group < 2 || BQ % group))) return SXB_ERR_BAD_ARG;
  if (!attn_args_ok(1, T, Hq, Hk, D, q_stride, 8, 8) || (o_stride % 8)) return SXB_ERR_UNSUPPORTED;
  if (block_size < 8 || (block_size % 8) || (block_size < BC ? (BC % block_size) : (block_size % BC)))
    return SXB_ERR_UNSUPPORTED;
  CUtensorMap tq, tk, tv;
  int rc;
  if ((rc = make_qkv_map(&tq, q, 1, T, Hq, D, q_stride, BQ))) return rc;
  const long long slot = 2LL * Hk * D;   // elements per cached token: K heads then V heads
  // 128-key tiles need pages that tile them exactly; other page sizes keep the 64-key kernel
  const bool wide = use_fwd128(D) && (block_size <= FBC ? (FBC % block_size) == 0 : (block_size % FBC) == 0);
  const int tile_rows = wide ? FBC : BC;
  const int box_rows = block_size < tile_rows ? block_size : tile_rows;
  const __nv_bfloat16* kbase = static_cast<const __nv_bfloat16*>(cache);
  if ((rc = make_qkv_map(&tk, kbase, n_blocks, block_size, Hk, D, slot, box_rows))) return rc;
  if ((rc = make_qkv_map(&tv, kbase + (long long)Hk * D, n_blocks, block_size, Hk, D, slot, box_rows))) return rc;
  Params p{};
  p.B = 1; p.S = T; p.Hq = Hq; p.Hk = Hk; p.D = D; p.causal = 1; p.scale = scale;
  p.out = out; p.out_stride = o_stride;
  p.atoms = static_cast<const int*>(atoms);
  p.block_tables = static_cast<const int*>(block_tables);
  p.max_blocks = max_blocks; p.block_size = block_size;
  cudaStream_t s = (cudaStream_t)stream;
  if (n_packed > 0) {
    // Q box = {64, group heads, 128 / group queries}: the TMA writes rows in (query, head) order
    CUtensorMap tqp;
    if ((rc = make_qkv_map(&tqp, q, 1, T, Hq, D, q_stride, BQ / group, group))) return rc;
    Params pp = p;
    pp.pack = group;
    pp.atoms = p.atoms + 6LL * (n_atoms - n_packed);
    dim3 gridp(Hk, 1, n_packed);
    if (wide) {
      if ((rc = set_smem(attn_fwd128_kernel<1>, Fwd128Smem::kTotal))) return rc;
      attn_fwd128_kernel<1><<<gridp, kFwd128Threads, Fwd128Smem::kTotal, s>>>(tqp, tk, tv, pp);
    } else if (D == 128) {
      if ((rc = set_smem(attn_fwd_kernel<128, 1>, FwdSmem<128>::kTotal))) return rc;
      attn_fwd_kernel<128, 1><<<gridp, kThreads, FwdSmem<128>::kTotal, s>>>(tqp, tk, tv, pp);
    } else {
      if ((rc = set_smem(attn_fwd_kernel<64, 1>, FwdSmem<64>::kTotal))) return rc;
      attn_fwd_kernel<64, 1><<<gridp, kThreads, FwdSmem<64>::kTotal, s>>>(tqp, tk, tv, pp);
    }
    SXB_CHECK_LAUNCH();
    n_atoms -= n_packed;
    if (n_atoms == 0) return SXB_OK;
  }
  dim3 grid(Hq, 1, n_atoms);
  const long long cache_bytes = (long long)n_blocks * block_size * slot * 2;
  if (wide && use_fwd_pair(D, Hq, Hk)) {
    grid.x = Hq / 2;
    if ((rc = set_smem(attn_fwd_pair_kernel, FwdPairSmem::kTotal))) return rc;
    attn_fwd_pair_kernel<<<grid, kFwdPairThreads, FwdPairSmem::kTotal, s>>>(tq, tk, tv, p);
    SXB_CHECK_LAUNCH();
    return SXB_OK;
  }
  if (wide) {
    if (use_multicast(Hq, Hk, D, cache, cache_bytes, cache, cache_bytes)) {
      if ((rc = set_smem(attn_fwd128_kernel<2>, Fwd128Smem::kTotal))) return rc;
      if ((rc = launch_cluster2(attn_fwd128_kernel<2>, grid, kFwd128Threads, Fwd128Smem::kTotal, s, tq, tk, tv, p))) return rc;
    } else {
      if ((rc = set_smem(attn_fwd128_kernel<1>, Fwd128Smem::kTotal))) return rc;
      attn_fwd128_kernel<1><<<grid, kFwd128Threads, Fwd128Smem::kTotal, s>>>(tq, tk, tv, p);
    }
    SXB_CHECK_LAUNCH();
    return SXB_OK;
  }
  if (use_multicast(Hq, Hk, D, cache, cache_bytes, cache, cache_bytes)) {
    if ((rc = set_smem(attn_fwd_kernel<128, 2>, FwdSmem<128>::kTotal))) return rc;
    if ((rc = launch_cluster2(attn_fwd_kernel<128, 2>, grid, kThreads, FwdSmem<128>::kTotal, s, tq, tk, tv, p)))
      return rc;
  } else if (D == 128) {
    if ((rc = set_smem(attn_fwd_kernel<128, 1>, FwdSmem<128>::kTotal))) return rc;
    attn_fwd_kernel<128, 1><<<grid, kThreads, FwdSmem<128>::kTotal, s>>>(tq, tk, tv, p);
  } else {
    if ((rc = set_smem(attn_fwd_kernel<64, 1>, FwdSmem<64>::kTotal))) return rc;
    attn_fwd_kernel<64, 1><<<grid, kThreads, FwdSmem<64>::kTotal, s>>>(tq, tk, tv, p);
  }
  SXB_CHECK_LAUNCH();
  return SXB_OK;
}

// Dense attention with an additive per-key bias (key-padding masks of BERT-style encoders and the v1 inference path;
// reference: the attn_mask argument of the fused softmax in csrc/transformer/softmax_kernels.cu and
// ops/transformer/inference softmax_context).  key_bias: fp32 [B, S] in RAW-score units, i.e. the caller's additive
// bias divided by `scale`; -inf masks a key.  S must be a multiple of 4 (16-byte loads of the bias).
// per_head != 0: key_bias is [B, Hq, S] (ALiBi-style slopes x key position); else [B, S].
SXB_API int sxb_attn_fwd_kbias(const void* q, const void* k, const void* v, void* out, void* lse, const void* key_bias,
                               int per_head, int B, int S, int Hq, int Hk, int D, long long q_stride,
                               long long k_stride, long long v_stride, long long o_stride, int causal, float scale,
                               void* stream) {
  if (key_bias == nullptr || (S % 4) || (reinterpret_cast<uintptr_t>(key_bias) & 15)) return SXB_ERR_UNSUPPORTED;
  g_kbias = static_cast<const float*>(key_bias);
  g_kbias_hstride = per_head ? S : 0;
  g_kbias_bstride = per_head ? (long long)Hq * S : S;
  const int rc = sxb_attn_fwd(q, k, v, out, lse, B, S, Hq, Hk, D, q_stride, k_stride, v_stride, o_stride, causal, scale,
                              stream);
  g_kbias = nullptr;
  return rc;
}

SXB_API int sxb_attn_bwd_kbias(const void* q, const void* k, const void* v, const void* out, const void* d_out,
                               const void* lse, void* delta, void* dq, void* dk, void* dv, const void* key_bias,
                               int per_head, int B, int S, int Hq, int Hk, int D, long long q_stride,
                               long long k_stride, long long v_stride, long long o_stride, long long do_stride,
                               long long dq_stride, long long dk_stride, long long dv_stride, int causal, float scale,
                               void* stream) {
  if (key_bias == nullptr || (S % 4) || (reinterpret_cast<uintptr_t>(key_bias) & 15)) return SXB_ERR_UNSUPPORTED;
  g_kbias = static_cast<const float*>(key_bias);
  g_kbias_hstride = per_head ? S : 0;
  g_kbias_bstride = per_head ? (long long)Hq * S : S;
  const int rc = sxb_attn_bwd(q, k, v, out, d_out, lse, delta, dq, dk, dv, B, S, Hq, Hk, D, q_stride, k_stride,
                              v_stride, o_stride, do_stride, dq_stride, dk_stride, dv_stride, causal, scale, stream);
  g_kbias = nullptr;
  return rc;
}
