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

#ifdef __cplusplus
}
#endif
#endif /* SXB200_H */
