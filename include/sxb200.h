/* sxb200.h — C ABI of libsxb200.so (sm_100a kernels) and libsxb200_host.so (host-side natives).
 *
 * Conventions
 *   - every entry point returns int: 0 = ok, >0 = cudaError_t, <0 = argument error
 *     (-1 bad argument, -2 unsupported shape/alignment, -3 size beyond a stated limit,
 *      -4 driver entry point unavailable);
 *   - `stream` is a cudaStream_t passed as void*; all launches are asynchronous on it;
 *   - dtype codes: 0 f32, 1 f16, 2 bf16, 3 fp8-e4m3, 4 i8, 5 u8, 6 i32, 7 i64;
 *   - pointers are device pointers unless stated otherwise; nothing here allocates device memory
 *     unless stated otherwise: scratch buffers are passed in by the caller.
 * Each declaration cites the reference code whose role it takes (paths under deepspeed/).
 */
#ifndef SXB200_H
#define SXB200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* ---- library ------------------------------------------------------------------------------- */
int sxb_abi_version(void);
int sxb_device_info(int* sm_count, int* cc_major, int* cc_minor, long long* total_mem);
/* pipeline watchdog records (8 x uint32, host pointer): see csrc/sxb_common.cuh mbar_wait_to */
int sxb_gemm_hang_info(unsigned int* out_host);
/* Kernel selection for sxb_gemm_bf16: 0 = auto (CTA-pair cta_group::2 kernel, 256x256 tiles, for large
 * problems; single-CTA 128x256 kernel otherwise), 1 = force single-CTA, 2 = force CTA-pair. */
int sxb_gemm_set_mode(int mode);
int sxb_attn_hang_info(unsigned int* out_host);
/* Local-ownership registry for TMA multicast (no counterpart in the reference).  Kernels that multicast a tile
 * across a cluster only do so when the tensor map covers this GPU's memory alone: ordinary allocation of the
 * current device, whole extent inside one allocation, not registered here.  comm/symm.py registers every peer /
 * multicast mapping of symmetric memory.  sxb_range_is_local returns 1/0; sxb_attn_would_multicast reports which
 * kernel family (1 = 2-CTA multicast, 0 = single CTA) sxb_attn_fwd / the dQ pass would take for these K/V. */
int sxb_register_foreign_range(const void* p, long long bytes);
int sxb_range_is_local(const void* p, long long bytes);
int sxb_attn_would_multicast(const void* k, const void* v, int B, int S, int Hq, int Hk, int D, long long k_stride,
                             long long v_stride);

/* ---- tcgen05 GEMM: C[M,N] (+)= alpha * op(A) op(B); layout 0 NT, 1 NN, 2 TN.
 *      Replaces cuBLAS behind F.linear (inference/v2/kernels/core_ops/blas_kernels/blas_utils.h:31-41). */
int sxb_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, int layout, int lda, int ldb,
                  int ldc, int c_dtype, int accumulate, float alpha, int max_ctas, void* stream);

/* ---- fused optimizers (ops/adam/fused_adam.py:176-190, ops/lion/fused_lion.py:118-128,
 *      ops/lamb/fused_lamb.py:165; unscale/clip + fp32->bit16 copy of runtime/zero/stage_1_and_2.py:2117-2177) */
int sxb_adam_flat(void* p, const void* g, void* m, void* v, void* lp, long long n, int g_dtype, int lp_dtype,
                  float lr, float beta1, float beta2, float eps, float weight_decay, int step, int adamw,
                  int bias_correction, float grad_scale, const void* noop, const void* dyn_scale, void* stream);
int sxb_multi_tensor_adam(const void* const* ptrs_host, const long long* numels_host, int n_tensors, int dtype,
                          float lr, float beta1, float beta2, float eps, float weight_decay, int step, int adamw,
                          int bias_correction, const void* noop, void* stream);
int sxb_lion_flat(void* p, const void* g, void* m, void* lp, long long n, int g_dtype, int lp_dtype, float lr,
                  float beta1, float beta2, float weight_decay, float grad_scale, const void* noop, void* stream);
int sxb_adagrad_flat(void* p, const void* g, void* h, void* lp, long long n, int g_dtype, int lp_dtype, float lr,
                     float eps, float weight_decay, float grad_scale, void* stream);
int sxb_lamb(void* p, const void* g, void* m, void* v, void* lp, void* upd_scratch, void* norms_scratch,
             void* coeff_out, long long n, int g_dtype, int lp_dtype, float lr, float beta1, float beta2, float eps,
             float weight_decay, int step, int bias_correction, float min_coeff, float max_coeff,
             float grad_scale, void* stream);
int sxb_sumsq(const void* x, long long n, int dtype, void* out2, void* stream);

/* ---- normalisation (inference/v2/kernels/core_ops/cuda_rms_norm/rms_norm_cuda.cu:19,84;
 *      cuda_layer_norm/layer_norm_cuda.cu:31,220) — backward needs fp32 scratch of
 *      sxb_norm_bwd_scratch_elems(rows, H, layer) elements */
long long sxb_norm_bwd_scratch_elems(long long rows, int H, int layer);
int sxb_rmsnorm_fwd(const void* x, const void* res, const void* w, void* y, void* res_out, void* rstd,
                    long long rows, int H, float eps, int dtype, void* stream);
int sxb_rmsnorm_bwd(const void* dy, const void* x, const void* w, const void* rstd, const void* dres, void* dx,
                    void* dw, void* scratch, long long rows, int H, int dtype, int accumulate_dw, void* stream);
/* RMSNorm backward for callers that cannot allocate device memory when they need it (a trainer whose
 * memory plan already owns the whole GPU, a server that gave the device to its KV-cache pool, a
 * recovery path running because an allocation failed).  No counterpart in the reference: the reference
 * always allocates through torch.  Identical to sxb_rmsnorm_bwd except that the weight-gradient
 * partials live in a ~9.3 MiB scratch array inside the library image (every GPU on which the norm
 * module of the library has been loaded holds it until the process exits; with CUDA's default lazy
 * module loading that happens at the first norm kernel or the first call of this function).  Limit: H <= sxb_rmsnorm_bwd_noalloc_max_h() (8192), otherwise returns -3.  One array
 * per device, shared by all calls: calls that overlap on the GPU race and must be serialised by the
 * caller; nothing needs zeroing between calls.  The first call on a device looks up the array's device
 * address (cudaGetSymbolAddress) — make it before any stream capture and before the rest of the device
 * memory is committed (the lazily loaded array needs its 9.3 MiB then).  Never calls cudaMalloc/cudaFree. */
int sxb_rmsnorm_bwd_noalloc(const void* dy, const void* x, const void* w, const void* rstd, const void* dres,
                            void* dx, void* dw, long long rows, int H, int dtype, int accumulate_dw, void* stream);
int sxb_rmsnorm_bwd_noalloc_max_h(void);
int sxb_layernorm_fwd(const void* x, const void* res, const void* w, const void* b, void* y, void* res_out,
                      void* mean, void* rstd, long long rows, int H, float eps, int dtype, void* stream);
int sxb_layernorm_bwd(const void* dy, const void* x, const void* w, const void* mean, const void* rstd,
                      const void* dres, void* dx, void* dw, void* db, void* scratch, long long rows, int H,
                      int dtype, int accumulate_dw, void* stream);

/* ---- elementwise (sequence/layer.py:84-103 rotary; core_ops/gated_activations/gated_activation_kernels_cuda.cu:45;
 *      core_ops/bias_activations/bias_activation_cuda.cu:52-92; sequence/cross_entropy.py:13-58;
 *      ragged_ops/embed/embed_cuda.cu:19-90) */
int sxb_rope(void* x, const void* positions, const void* cos_tab, const void* sin_tab, long long tokens,
             int n_heads, int head_dim, long long row_stride, int inverse, int dtype, void* stream);
int sxb_gated_act_fwd(const void* in, const void* bias, void* out, long long rows, int F, int act, int dtype, void* stream);
int sxb_gated_act_bwd(const void* dy, const void* in, void* din, long long rows, int F, int act, int dtype, void* stream);
int sxb_bias_act(const void* x, const void* bias, void* y, long long rows, int C, int act, int dtype, void* stream);
int sxb_bias_act_bwd(const void* dy, const void* x, const void* bias, void* dx, long long rows, int C, int act,
                     int dtype, void* stream);
int sxb_cross_entropy(void* logits, const void* labels, void* loss, void* lse, long long rows, int V, long long ld,
                      long long ignore_index, const void* grad_scale, int write_grad, int dtype, void* stream);
int sxb_embedding_fwd(const void* ids, const void* table, void* out, long long ntok, int H, int dtype, void* stream);
int sxb_embedding_bwd(const void* ids, const void* dout, void* dtable, long long ntok, int H, int dtype, void* stream);
int sxb_scale_cast(const void* src, void* dst, long long n, int src_dtype, int dst_dtype, float scale,
                   int accumulate, void* stream);

/* ---- flash attention on tcgen05 (sequence/fpdt_layer.py:14-20 external flash-attn;
 *      inference/v2/kernels/ragged_ops/blocked_flash/blocked_flash.cpp:21-101) */
int sxb_attn_fwd(const void* q, const void* k, const void* v, void* out, void* lse, int B, int S, int Hq, int Hk,
                 int D, long long q_stride, long long k_stride, long long v_stride, long long o_stride, int causal,
                 float scale, void* stream);
int sxb_attn_bwd(const void* q, const void* k, const void* v, const void* out, const void* d_out, const void* lse,
                 void* delta_scratch, void* dq, void* dk, void* dv, int B, int S, int Hq, int Hk, int D,
                 long long q_stride, long long k_stride, long long v_stride, long long o_stride,
                 long long do_stride, long long dq_stride, long long dk_stride, long long dv_stride, int causal,
                 float scale, void* stream);

/* ---- host library (libsxb200_host.so): CPU optimizers for ZeRO-Offload
 *      (ops/adam/cpu_adam.py:94-102,163-165; ops/adagrad/cpu_adagrad.py:22-29,90-96; ops/lion/cpu_lion.py:57-64,125) */
int sxbh_create_adam(int id, float lr, float b1, float b2, float eps, float wd, int adamw, int log);
int sxbh_destroy_adam(int id);
int sxbh_adam_update(int id, long long step, float lr, float b1, float b2, float eps, float wd, int bias_correction,
                     float* p, const float* g, float* m, float* v, long long n, void* lp, int lp_kind, float grad_scale);
int sxbh_create_adagrad(int id, float lr, float eps, float wd, int log);
int sxbh_destroy_adagrad(int id);
int sxbh_adagrad_update(int id, long long step, float lr, float eps, float wd, float* p, const float* g, float* h,
                        long long n, void* lp, int lp_kind);
int sxbh_create_lion(int id, float lr, float b1, float b2, float wd, int log);
int sxbh_destroy_lion(int id);
int sxbh_lion_update(int id, long long step, float lr, float b1, float b2, float wd, float* p, const float* g,
                     float* m, long long n, void* lp, int lp_kind);

/* ---- host library: shared-memory all-reduce among the processes of one host, for CPU tensors
 *      (reference: ShareMemCommBuilder / torch.ops.deepspeed.inference_all_reduce_, comm/torch.py:29-35,166-171).
 *      init is collective (rank 0 creates /dev/shm/<name>); slot_bytes: multiple of 64, >= 4096; larger buffers are
 *      reduced in pieces.  dtype: 0 fp32, 1 bf16, 2 fp16; accumulation in fp32.  Returns 0 or -errno. */
int sxbh_shm_init(const char* name, int world, int rank, long long slot_bytes);
int sxbh_shm_allreduce(void* buf, long long numel, int dtype);
int sxbh_shm_barrier(void);
int sxbh_shm_finalize(void);

/* ---- host library: schedule-compiler core (reference: the C++ runtime behind DeepCompileBuilder, compile/util.py,
 *      compile/list_schedule.py:171-209).  A plan is parallel arrays per node: kind (0 gather, 1 compute, 2 release,
 *      3 reduce, 4 offload, 5 reload), unit, phase (0 fwd, 1 bwd), bytes, time_us.
 *      simulate: out[0] peak bytes, out[1] time (us), out[2] exposed communication (us).
 *      hoist_gathers: perm[i] = input index of the node placed at position i after memory-bounded hoisting. */
int sxbh_plan_simulate(int n, const int* kind, const int* unit, const int* phase, const long long* bytes,
                       const double* time_us, const int* persistent, int n_persistent, long long base_bytes,
                       double* out);
int sxbh_plan_hoist_gathers(int n, const int* kind, const int* unit, const int* phase, const long long* bytes,
                            const double* time_us, const int* persistent, int n_persistent, long long budget,
                            int* perm);

/* ==== declarations below are kept in step with csrc/ by tools/gen_header.py ==== */

/* ---- csrc/attention.cu */
/* K7 (Ulysses): forward attention whose epilogue stores each output row straight into the buffer of the rank that
 * owns that part of the sequence.  out_peers: host array of `world` device pointers — peer r's destination for the
 * rows [r*rows_per_peer, (r+1)*rows_per_peer) of this rank's heads (column offset already applied), row pitch
 * o_stride.  `out` is unused.  (reference: all_to_all_single after the local attention, sequence/layer.py:434-437) */
int sxb_attn_fwd_scatter(const void* q, const void* k, const void* v, void* const* out_peers, int world,
                         int rows_per_peer, void* out_keep, long long keep_stride, void* lse, int B, int S, int Hq,
                         int Hk, int D, long long q_stride, long long k_stride, long long v_stride,
                         long long o_stride, int causal, float scale, void* stream);
/* Backward with scattered dQ / dK / dV: dq_peers[r] / dk_peers[r] are peer r's destinations (column offsets of this
 * rank's heads applied) with row pitches dq_stride / dk_stride; dV lands dv_col_delta elements after dK with pitch
 * dv_stride. */
int sxb_attn_bwd_scatter(const void* q, const void* k, const void* v, const void* out, const void* d_out,
                         const void* lse, void* delta, void* const* dq_peers, void* const* dk_peers,
                         long long dv_col_delta, int world, int rows_per_peer, int B, int S, int Hq, int Hk, int D,
                         long long q_stride, long long k_stride, long long v_stride, long long o_stride,
                         long long do_stride, long long dq_stride, long long dk_stride, long long dv_stride,
                         int causal, float scale, void* stream);
/* Packed sequences (cu_seqlens int32 [nseq + 1] on the device; q/k/v/out [total, H, D]; lse / delta [Hq, total]):
 * every sequence attends only to itself.  The reference reaches this through flash-attn's varlen entry points
 * (sequence/fpdt_layer.py:14-20, inference/v2 blocked_flash); here it is the same three kernels with a per-CTA
 * row window. */
int sxb_attn_fwd_varlen(const void* q, const void* k, const void* v, void* out, void* lse, const void* cu_seqlens,
                        int nseq, int max_seqlen, int total, int Hq, int Hk, int D, long long q_stride,
                        long long k_stride, long long v_stride, long long o_stride, int causal, float scale,
                        void* stream);
int sxb_attn_bwd_varlen(const void* q, const void* k, const void* v, const void* out, const void* d_out,
                        const void* lse, void* delta, void* dq, void* dk, void* dv, const void* cu_seqlens,
                        int nseq, int max_seqlen, int total, int Hq, int Hk, int D, long long q_stride,
                        long long k_stride, long long v_stride, long long o_stride, long long do_stride,
                        long long dq_stride, long long dk_stride, long long dv_stride, int causal, float scale,
                        void* stream);
/* Blocked ("paged") flash attention over a ragged batch on the tensor cores: the forward kernel above, one CTA per
 * (attention atom, q head).  q/out [T, Hq, D]; cache [n_blocks, block_size, 2, Hk, D] (K then V per slot); atoms
 * int32 [n_atoms, 6] and block_tables int32 [n_seqs, max_blocks] on the device.  block_size: a multiple of 8 that
 * divides 64, or a multiple of 64.  Causal by position: query i of an atom sits at global_q_idx + i.
 * (reference: inference/v2/kernels/ragged_ops/blocked_flash/blocked_flash.cpp:21-101 over flash-attn's sm_80 kernels) */
int sxb_attn_fwd_atoms(const void* q, const void* cache, void* out, const void* atoms, int n_atoms, int n_packed,
                       const void* block_tables, int max_blocks, int n_blocks, int block_size, int T, int Hq,
                       int Hk, int D, long long q_stride, long long o_stride, float scale, void* stream);

/* ---- csrc/gemm_ext.cu */
/* C (+)= alpha * op(A) op(B) with optional gating / grouping / fp8 operands (see the file header).
 *   ab_dtype: kBF16 (2) or kFP8E4M3 (3; NT layout only).  c_dtype: kBF16 or kF32.
 *   gate_flags == NULL: no gating.  Else flags[c] >= gate_value (as a wrapping 32-bit sequence number) means bytes
 *     [c*gate_chunk, (c+1)*gate_chunk) of the buffer B lives in have arrived; gate_base = byte offset of B in it.
 *   group_mode 0: plain.  1 (ragged-M, NT/NN): offsets[num_groups+1] (device, rows, multiples of 256); A and C have
 *     total_rows rows; B is the row-wise stack of the groups' matrices (NT: num_groups*N rows; NN: num_groups*K).
 *     2 (ragged-K, TN): offsets multiples of 64; A [total_rows, M], B [total_rows, N]; C = num_groups matrices
 *     [M, N] spaced c_group_stride elements apart (M multiple of 128).
 *   scale_a[row of A] / scale_b[g*N + column]: fp8 dequantisation scales (either may be NULL). */
int sxb_gemm_ext(const void* A, const void* B, void* C, int M, int N, int K, int layout, long long lda,
                 long long ldb, long long ldc, int ab_dtype, int c_dtype, int accumulate, float alpha,
                 const void* gate_flags, unsigned int gate_value, long long gate_base, long long gate_chunk,
                 long long gate_nchunks, int group_mode, int num_groups, const void* offsets, long long total_rows,
                 long long c_group_stride, const void* scale_a, const void* scale_b, int max_ctas, void* stream);
/* K6 (Ulysses): C = A op(B) whose epilogue stores every 64-column block of the result straight into the buffer of the
 * rank that owns those heads (direct peer stores over NVLink).  table: device int32 [2 * N/64] = (owner, destination
 * column) per block; peers: host array of `world` device pointers; the result row r lands in row row0 + r of the
 * destination (pitch dst_ld elements).  bf16 only, plain (ungrouped) problem, N % 64 == 0.
 * (reference: QKV projection, then three all_to_all_single, sequence/layer.py:388-426) */
int sxb_gemm_scatter(const void* A, const void* B, int M, int N, int K, int layout, long long lda, long long ldb,
                     const void* table, void* const* peers, int world, long long dst_ld, long long row0,
                     float alpha, int max_ctas, void* stream);
int sxb_stream_write_u32(void* dst, unsigned int value, void* stream);

/* ---- csrc/moe_ep.cu */
/* All pointer tables are host arrays of `world` device pointers (peer mappings).  tables: device int32
 * [E + 2L + 2] (layout above).  E = world * L.  cap_rows: rows of every receive / return buffer. */
int sxb_moe_ep_plan(const void* const* counts, const void* const* sigs, int rank, int world, const void* my_counts,
                    void* tables, int E, int L, int cap_rows, unsigned int epoch, unsigned int pulled_epoch,
                    int exchange, void* stream);
int sxb_moe_ep_plan_emu(const void* const* counts, const void* const* sigs, int world,
                        const void* const* my_counts, void* const* tables, int E, int L, int cap_rows,
                        unsigned int epoch, unsigned int pulled_epoch, int exchange, void* stream);
int sxb_moe_ep_push(const void* const* recv, const void* const* sigs, int rank, int world, const void* x,
                    const void* weights, const void* pulled, void* d_scores, const void* assign,
                    const void* offsets, const void* tables, int T, int K, int H, int E, int L, int cap_rows,
                    unsigned int epoch, void* counter, int max_ctas, void* stream);
int sxb_moe_ep_push_emu(const void* const* recv, const void* const* sigs, int world, const void* const* x,
                        const void* const* weights, const void* const* pulled, void* const* d_scores,
                        const void* const* assign, const void* const* offsets, const void* const* tables, int T,
                        int K, int H, int E, int L, int cap_rows, unsigned int epoch, void* const* counters,
                        int ctas_per_rank, void* stream);
int sxb_moe_ep_pull(const void* const* ret, const void* const* sigs, int rank, int world, void* out,
                    const void* weights, void* keep, const void* assign, const void* offsets, const void* tables,
                    int T, int K, int H, int E, int L, int cap_rows, unsigned int epoch, void* counter,
                    int max_ctas, void* stream);
int sxb_moe_ep_pull_emu(const void* const* ret, const void* const* sigs, int world, void* const* out,
                        const void* const* weights, void* const* keep, const void* const* assign,
                        const void* const* offsets, const void* const* tables, int T, int K, int H, int E, int L,
                        int cap_rows, unsigned int epoch, void* const* counters, int ctas_per_rank, void* stream);

/* ---- csrc/moe_ragged.cu */
/* logits [T, E]; scores fp32 [T, K]; assignments / offsets int32 [T, K]; expert_counts int32 [E] (zeroed here). */
int sxb_top_k_gating(const void* logits, void* scores, void* assignments, void* offsets, void* expert_counts,
                     int n_tokens, int n_experts, int top_k, int renormalize, int dtype, void* stream);
/* x [T, H] -> moe_input [T*K, H] grouped by expert; expert_cumsum int64 [E]; mapped_slots int32 [T, K]. */
int sxb_moe_scatter(const void* x, void* moe_input, const void* assignments, const void* offsets,
                    const void* expert_counts, void* expert_cumsum, void* mapped_slots, int n_tokens,
                    int n_experts, int top_k, int hidden, int elem_size, void* stream);
/* out [T, H] = sum_k scores * moe_output[slot]  (scores may be NULL: plain sum).  bf16 rows. */
int sxb_moe_gather(const void* moe_output, void* out, const void* scores, const void* mapped_slots, int n_tokens,
                   int top_k, int hidden, void* stream);
/* Backward of gather: d_moe_output[slot] = scores * d_out ; d_scores[t,k] = <d_out[t], moe_output[slot]>. */
int sxb_moe_gather_bwd(const void* d_out, const void* moe_output, const void* scores, const void* mapped_slots,
                       void* d_moe_output, void* d_scores, int n_tokens, int top_k, int hidden, void* stream);
int sxb_ragged_embed(const void* ids, const void* pos, const void* table, const void* pos_table, void* out,
                     int n_tokens, int hidden, int pos_offset, void* stream);
int sxb_logits_gather(const void* hidden_states, const void* last_token, void* out, int n_seqs, int hidden,
                      void* stream);
int sxb_kv_rotary_blocked(void* qkv, void* cache, const void* tok_seq, const void* tok_pos,
                          const void* block_tables, int n_tokens, int max_blocks, int block_size, int Hq, int Hk,
                          int D, int rot_dim, float theta, void* stream);
int sxb_paged_attention(const void* q, void* out, const void* cache, const void* tok_seq, const void* tok_pos,
                        const void* block_tables, int n_tokens, int max_blocks, int block_size, int Hq, int Hk,
                        int D, long long q_stride, float scale, void* stream);

/* ---- csrc/quantize.cu */
/* Grouped integer quantisation.  in: [groups, group_size] of `dtype`; q: groups*group_size*bits/8 bytes;
 * scales: fp32 [groups] (symmetric) or [groups, 2] = (scale, min) (asymmetric).  nodes/devs > 1 select the
 * qgZ swizzled output order.  bits in {4, 8}. */
int sxb_quantize_int(const void* in, void* q, void* scales, long long groups, int group_size, int bits,
                     int symmetric, int stochastic, unsigned seed, int nodes, int devs, int dtype, void* stream);
int sxb_dequantize_int(const void* q, const void* scales, void* out, long long groups, int group_size, int bits,
                       int symmetric, int dtype, void* stream);
/* q_in: [devs, out_groups] groups, reduced over devs and re-quantised into q_out [out_groups]. */
int sxb_quantized_reduction(const void* q_in, const void* s_in, void* q_out, void* s_out, long long out_groups,
                            int group_size, int bits, int symmetric, int devs, void* stream);
/* out: [groups, group_size*q_bits/8 + 4] bytes (payload then the fp32 scale).  q_bits in {4, 6, 8, 12}. */
int sxb_fp_quantize(const void* in, void* out, long long groups, int group_size, int q_bits, int stochastic,
                    unsigned seed, int dtype, void* stream);
/* rows == NULL: dequantise all `out_groups` groups.  Otherwise output row r is input row rows[r], a row
 * being groups_per_row consecutive groups (selective dequantisation, e.g. of the experts actually routed to). */
int sxb_fp_dequantize(const void* in, void* out, long long out_groups, int group_size, int q_bits,
                      const void* rows, long long groups_per_row, int dtype, void* stream);
int sxb_packbits(const void* in, void* out, long long n, int dtype, void* stream);
int sxb_unpackbits(const void* in, void* out, long long n, float scale, int dtype, void* stream);
/* idx: int32 [rows, k], sorted ascending in place per row (k <= 4096). */
int sxb_token_sort(void* idx, int rows, int k, void* stream);
/* full: [batch, seq, hidden] (batch_first) or [seq, batch, hidden]; part: same with seq -> k; idx int32 [batch, k].
 * scatter = 0: part <- full[idx] ; scatter = 1: full[idx] <- part.  hidden*elem_size must be a multiple of 16. */
int sxb_token_move(void* full, void* part, const void* idx, int batch, int seq, int k, int hidden, int elem_size,
                   int batch_first, int scatter, void* stream);
/* in: bf16 [rows, cols] (pitch ld_in elements, cols % 8 == 0, 16-byte aligned rows); out: e4m3 bytes (pitch ld_out,
 * multiple of 8); scales: fp32 [rows].  Dequantised value = e4m3 * scale. */
int sxb_fp8_rowwise_quant(const void* in, void* out, void* scales, long long rows, int cols, long long ld_in,
                          long long ld_out, void* stream);

/* ---- csrc/softmax_dropout.cu */
/* out = residual + dropout(x + bias); ``keep`` (uint8 per element) may be null when p == 0.  n % 4 == 0, cols % 4 == 0. */
int sxb_bias_dropout_residual(const void* x, const void* bias, const void* residual, void* out, void* keep,
                              long long n, int cols, float p, long long seed, long long offset, int dtype,
                              void* stream);
int sxb_dropout_bwd(const void* dy, const void* keep, void* dx, long long n, float p, int dtype, void* stream);
/* scores/out [B, H, Sq, Sk] contiguous; bias strides in elements (b, h, q, k), 0 = broadcast. */
int sxb_biased_softmax_fwd(const void* scores, void* out, const void* bias1, long long b1b, long long b1h,
                           long long b1q, long long b1k, const void* bias2, long long b2b, long long b2h,
                           long long b2q, long long b2k, long long B, int H, int Sq, int Sk, float scale,
                           int dtype, void* stream);
int sxb_biased_softmax_bwd(const void* y, const void* dy, void* ds, void* dbias, long long rows, int Sk,
                           float scale, int dtype, void* stream);
int sxb_bsr_softmax_fwd(const void* x, void* y, const void* row_ptr, const void* col_idx, long long B, int H,
                        int nrb, int ncb, int blk, int nnz, const void* rpe, long long rpe_sb, const void* kpm,
                        int kpm_mode, const void* amask, int amask_mode, float scale, int dtype, void* stream);
int sxb_bsr_softmax_bwd(const void* y, const void* dy, void* dx, const void* row_ptr, const void* col_idx,
                        long long B, int H, int nrb, int ncb, int blk, int nnz, float scale, int dtype,
                        void* stream);

/* ---- csrc/symm_collectives.cu */
/* K1.  bufs/sigs: host arrays of `world` device pointers (peer mappings of the flat bf16 gradient buffer
 * and of the signal pads).  part: elements per shard (multiple of 8).  counter: device uint32 (zeroed once).
 * apply_adam == 0 -> plain reduce-scatter(mean) into grad_out (used when the step must be deferred).
 * handshake_done != 0: the caller already announced "gradients ready" (sxb_symm_signal on the stream that produced
 * them) and waited for the peers (sxb_symm_wait on this stream); no CTA of this grid spins. */
int sxb_symm_rs_adam(const void* const* bufs, const void* const* sigs, int rank, int world, long long part,
                     void* hp, void* m, void* v, void* lp, void* grad_out, float lr, float beta1, float beta2,
                     float eps, float weight_decay, int step, int adamw, int bias_correction, float grad_scale,
                     unsigned int epoch, const void* multicast_ptr, void* counter, int apply_adam, int max_ctas,
                     int handshake_done, void* stream);
/* K1, `world` logical ranks emulated on ONE GPU by one cooperative grid.  bufs/sigs as above (all local);
 * hp/m/v/lp/grad_out/counters: host arrays of `world` per-rank device pointers.  ctas_per_rank * world CTAs of 512
 * threads must be co-resident (cudaLaunchCooperativeKernel refuses otherwise). */
int sxb_symm_rs_adam_emu(const void* const* bufs, const void* const* sigs, int world, long long part,
                         void* const* hp, void* const* m, void* const* v, void* const* lp, void* const* grad_out,
                         float lr, float beta1, float beta2, float eps, float weight_decay, int step, int adamw,
                         int bias_correction, float grad_scale, unsigned int epoch, void* const* counters,
                         int apply_adam, int ctas_per_rank, void* stream);
int sxb_symm_barrier(const void* const* bufs, const void* const* sigs, int rank, int world, int channel,
                     unsigned int epoch, void* stream);
int sxb_symm_signal(const void* const* bufs, const void* const* sigs, int rank, int world, int channel,
                    unsigned int epoch, void* stream);
int sxb_symm_wait(const void* const* bufs, const void* const* sigs, int rank, int world, int channel,
                  unsigned int epoch, void* stream);
int sxb_symm_subset_mean(const void* const* bufs, const void* const* sigs, int rank, int world, const int* members,
                         int n_members, void* out, long long n, int dtype, unsigned int epoch, void* counter,
                         int max_ctas, void* stream);
/* K9, general form: out = sum_k weights[k] * shard of members[k] (weights == NULL: the mean).  One kernel serves RR
 * (all replicas), shuffle (my ring), H-RR (two-level mean as per-member weights) and Gossip push-sum (alpha ratios). */
int sxb_symm_subset_wsum(const void* const* bufs, const void* const* sigs, int rank, int world, const int* members,
                         const float* weights, int n_members, void* out, long long n, int dtype,
                         unsigned int epoch, void* counter, int max_ctas, void* stream);
/* K9 emulated (bf16).  As in the real launch EVERY rank runs the kernel, each over its own ring: members is a
 * [world][16] row-major table, n_members[r] the length of row r (row r must contain r). */
int sxb_symm_subset_mean_emu(const void* const* bufs, const void* const* sigs, int world, const int* members,
                             const int* n_members, void* const* outs, long long n, unsigned int epoch,
                             void* const* counters, int ctas_per_rank, int shared_result, void* stream);
/* K2 gather half.  chunk_flags (optional, device uint32[ceil(world*part_bytes / chunk_bytes)]): chunk c is set to
 * `flag_value` (release, gpu scope) once all of its bytes are in dst — the gated GEMM polls these. */
int sxb_symm_all_gather(const void* const* bufs, const void* const* sigs, int rank, int world,
                        long long part_bytes, void* dst, long long chunk_bytes, void* chunk_flags,
                        unsigned int epoch, int max_ctas, unsigned int flag_value, void* stream);
int sxb_symm_all_gather_emu(const void* const* bufs, const void* const* sigs, int world, long long part_bytes,
                            void* const* dsts, long long chunk_bytes, void* const* chunk_flags, unsigned int epoch,
                            unsigned int flag_value, int ctas_per_rank, void* stream);
int sxb_symm_a2a_rows(const void* const* bufs, const void* const* sigs, int rank, int world, const void* src,
                      const long long* send_off, const long long* recv_off, const long long* count, int row_bytes,
                      unsigned int epoch, void* counter, int max_ctas, void* stream);
/* Emulated: srcs[r], and send_off/recv_off/count as [world][world] row-major int64 tables (row r = rank r's tables). */
int sxb_symm_a2a_rows_emu(const void* const* bufs, const void* const* sigs, int world, const void* const* srcs,
                          const long long* send_off, const long long* recv_off, const long long* count,
                          int row_bytes, unsigned int epoch, void* const* counters, int ctas_per_rank,
                          void* stream);

/* ---- csrc/symm_collectives.cu */
/* K9 with the choice of form.  shared_result != 0: every member of `members` calls with the SAME member list and
 * weights (RR, H-RR) -> two-phase form (see subset_mean_two_phase): `counter` must point at two zeroed words, n % 8 == 0,
 * and the symmetric buffer is overwritten with reduced values.  Falls back to the one-phase form for a single member. */
int sxb_symm_subset_wsum2(const void* const* bufs, const void* const* sigs, int rank, int world,
                          const int* members, const float* weights, int n_members, void* out, long long n,
                          int dtype, unsigned int epoch, void* counter, int max_ctas, int shared_result,
                          void* stream);

/* ---- csrc/attention.cu */
/* SXB_ATTN_DEBUG=3 cycle accounting of the last dQ launch (8 x uint64, host pointer); see g_attn_prof. */
int sxb_attn_prof(unsigned long long* out_host);

/* ---- csrc/quantize.cu */
/* Weight-only quantised linear for small batches: y[M, N] (bf16) = x[M, K] (bf16) W^T, W [N, K] in the layout written
 * by sxb_fp_quantize (q_bits 4 / 6 / 8, groups of `group_size` along the flattened weight, K % group_size == 0,
 * K % 256 == 0).  Intended for M <= 16 (decode); larger M should dequantise and use the tcgen05 GEMM. */
int sxb_fp_gemv(const void* x, const void* wq, void* y, int M, int N, int K, int q_bits, int group_size,
                long long x_stride, long long y_stride, void* stream);

/* ---- csrc/attention.cu */
/* Dense attention with an additive per-key bias (key-padding masks of BERT-style encoders and the v1 inference path;
 * reference: the attn_mask argument of the fused softmax in csrc/transformer/softmax_kernels.cu and
 * ops/transformer/inference softmax_context).  key_bias: fp32 [B, S] in RAW-score units, i.e. the caller's additive
 * bias divided by `scale`; -inf masks a key.  S must be a multiple of 4 (16-byte loads of the bias). */
int sxb_attn_fwd_kbias(const void* q, const void* k, const void* v, void* out, void* lse, const void* key_bias,
                       int per_head, int B, int S, int Hq, int Hk, int D, long long q_stride, long long k_stride,
                       long long v_stride, long long o_stride, int causal, float scale, void* stream);
int sxb_attn_bwd_kbias(const void* q, const void* k, const void* v, const void* out, const void* d_out,
                       const void* lse, void* delta, void* dq, void* dk, void* dv, const void* key_bias,
                       int per_head, int B, int S, int Hq, int Hk, int D, long long q_stride, long long k_stride,
                       long long v_stride, long long o_stride, long long do_stride, long long dq_stride,
                       long long dk_stride, long long dv_stride, int causal, float scale, void* stream);

/* ---- csrc/host/aio.cpp */
int sxbh_aio_create(long long block_size, int queue_depth, int single_submit, int overlap_events,
                    int intra_op_parallelism);
int sxbh_aio_destroy(int id);
/* async = 0: returns when the bytes are transferred (0 or -errno).  async = 1: returns immediately; sxbh_aio_wait
 * completes it. */
int sxbh_aio_pread(int id, void* buffer, long long nbytes, const char* path, long long file_offset, int async);
int sxbh_aio_pwrite(int id, const void* buffer, long long nbytes, const char* path, long long file_offset,
                    int async);
int sxbh_aio_wait(int id);
long long sxbh_aio_info(int id, int what);
long long sxbh_file_size(const char* path);

/* ---- csrc/host/cpu_optim.cpp */
/* bf16 <-> fp32 bulk converters used by the offload paths (gradients arrive from the GPU in bf16). */
int sxbh_bf16_to_f32(const uint16_t* src, float* dst, long long n, float scale, int accumulate);
int sxbh_num_threads();
int sxbh_simd_width();

/* ---- csrc/host/ragged_host.cpp */
/* atoms: int32 [max_atoms, 6] = (seq_index, q_start_idx, q_len, global_q_idx, kv_blocks, total_extent)
 * returns the number of atoms written, or -1 when max_atoms is too small. */
int sxbh_build_atoms(int32_t* atoms, int max_atoms, const int32_t* seq_start, const int32_t* seq_tokens,
                     const int32_t* seq_seen, int n_seqs, int q_block, int kv_block);
/* tok_seq / tok_pos: int32 [total tokens]; last_tok: int32 [n_seqs]; returns total tokens. */
int sxbh_fill_ragged_batch(int32_t* tok_seq, int32_t* tok_pos, int32_t* last_tok, const int32_t* seq_tokens,
                           const int32_t* seq_seen, int n_seqs);

/* ---- csrc/quantize.cu */
/* W8A16 / W4A16 linear for small batches on the inference/quantization layout (see int_gemv_kernel): x bf16 [M, K],
 * codes uint8, scale / min bf16 [N, K / group_size], y bf16 [M, N].  K % 256 == 0, group_size % 8 == 0,
 * K % group_size == 0.  (reference: inference/quantization dequantises and calls F.linear for every batch size.) */
int sxb_int_gemv(const void* x, const void* codes, const void* scale, const void* minv, void* y, int M, int N,
                 int K, int bits, int group_size, long long x_stride, long long y_stride, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SXB200_H */
