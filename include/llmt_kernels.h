/* C ABI of the sm_100a kernel library (llm_training_b200/_C/libllmt_kernels.so).
 *
 * Conventions
 *   - every function returns 0 on success; on failure llmt_last_error() describes the problem
 *   - pointers are device pointers unless stated otherwise; bf16 tensors are passed as void*
 *   - the caller owns all memory, including scratch buffers (sizes are documented per function)
 *   - work is enqueued on `stream`; nothing synchronises the device
 * Citations name the reference code path each kernel replaces (paths relative to the reference repo).
 */
#ifndef LLMT_KERNELS_H_
#define LLMT_KERNELS_H_

#include <cuda_runtime.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* llmt_last_error(void);
int llmt_abi_version(void);
unsigned long long llmt_launch_count(void); /* kernels launched by this library since load */
/* CTAs per SM for the optimizer-update kernels (1..8, default 8).  2 lets them run in the background of the next
 * forward pass: tensor-core CTAs of other streams still fit on every SM. */
void llmt_set_update_ctas_per_sm(int ctas);

/* ---- GEMM: C[M,N] = op(A)[M,K] * op(B)[K,N] (+ R), bf16 in, fp32 accumulate, tcgen05/TMEM/TMA ----------
 * replaces cuBLAS behind nn.Linear (src/llm_training/models/llama/llama_model.py:459-462,421-423,55).
 * a_mn_major=0: A is [M,K]; 1: A is [K,M].  b_mn_major=0: B is [N,K]; 1: B is [K,N].  out_fp32: C/R fp32.
 * R may be NULL; R == C accumulates in place.  cta_group 1|2, max_ctas<=0 -> all SMs. */
int llmt_gemm_bf16(const void* A, const void* B, void* C, const void* R, int M, int N, int K, int64_t lda,
                   int64_t ldb, int64_t ldc, int64_t ldr, int a_mn_major, int b_mn_major, int out_fp32,
                   int cta_group, int max_ctas, cudaStream_t stream);

/* Fused collective GEMMs for tensor/sequence parallelism over symmetric (peer-mapped) memory; replaces the
 * DTensor all-gather / reduce-scatter around the column-/row-parallel linears
 * (src/llm_training/models/llama/llama_model.py:223-238).  a_ptrs / c_ptrs are HOST arrays of `world` device
 * pointers (this rank's view of every peer's buffer).
 *   mode 1: C[M,N] = concat_r(A_r)[M,K] * op(B)   (A tiles are TMA-loaded from their owner rank over NVLink)
 *   mode 2: rows r*rows_per_rank.. of A[M,K] * op(B) are TMA-reduce-added into c_ptrs[r] (must be initialised;
 *           barrier before and after). */
int llmt_gemm_bf16_dist(int mode, const void* A, const void* const* a_ptrs, const void* B, void* C,
                        void* const* c_ptrs, int world, int rows_per_rank, int M, int N, int K, int64_t lda,
                        int64_t ldb, int64_t ldc, int b_mn_major, int cta_group, int max_ctas, cudaStream_t stream);

/* All-gather -> GEMM, "land once": extra warps of the GEMM kernel copy each remote [rows_per_rank, K] block over
 * NVLink into a_full [M, K] exactly once while the tensor cores consume the blocks already present; the TMA
 * producer waits on per-block arrival counters (flags: int32 [M / (128 * cta_group)], reset by the call).  a_full
 * is left holding the gathered activation for the backward pass. */
int llmt_gemm_bf16_ag(const void* const* a_ptrs, void* a_full, const void* B, void* C, int* flags, int rank,
                      int world, int rows_per_rank, int M, int N, int K, int64_t lda, int64_t ld_full, int64_t ldb,
                      int64_t ldc, int b_mn_major, int cta_group, int max_ctas, cudaStream_t stream);

/* ZeRO-3 parameter all-gather -> first consumer GEMM: C = A * W^T with W = elements [w_off, w_off + N*K) of a
 * sharding unit's flat parameter buffer whose slices live on their owners (shard_ptrs: HOST array of `world`
 * pointers into the symmetric shard arena, shard_numel elements each).  Extra warps of the GEMM pull W over NVLink
 * into full_base + w_off n-block by n-block; tiles are walked n-major and the TMA producer of an n-block waits for
 * that block's arrival counter (flags: int32 [ceil(N/256)], reset by the call).  Replaces FSDP2's all-gather before
 * the layer's first linear (src/llm_training/models/llama/llama_model.py:263-268). */
int llmt_gemm_bf16_wgather(const void* A, const void* const* shard_ptrs, void* full_base, int64_t w_off,
                           int64_t shard_numel, void* C, int* flags, int rank, int world, int M, int N, int K,
                           int64_t lda, int64_t ldc, int cta_group, int max_ctas, cudaStream_t stream);

/* ---- NVLink collectives for the sharding engine (csrc/kernels/comm.cu); replace NCCL all-gather /
 *      reduce-scatter issued by FSDP2 / ZeRO (llama_model.py:263-268, deepspeed_strategy.py:39-44) ------------ */
int llmt_comm_barrier(const void* const* signal_pads, int rank, int world, int slot_base, uint32_t epoch,
                      cudaStream_t stream);
int llmt_comm_allgather_pull(const void* const* full_ptrs, int rank, int world, int64_t shard_bytes,
                             cudaStream_t stream);
/* ZeRO-3 parameter gather: full[r * shard_bytes ..] <- rank r's slice, pulled from the symmetric shard arena */
int llmt_comm_gather_shards(const void* const* shard_ptrs, void* dst_full, int rank, int world, int64_t shard_bytes,
                            cudaStream_t stream);
int llmt_comm_gather_range(const void* const* shard_ptrs, void* dst_full, int world, int64_t shard_bytes, int64_t lo,
                           int64_t hi, int64_t skip_lo, int64_t skip_hi, cudaStream_t stream);
int llmt_comm_rs_num_partials(void);
int llmt_comm_rs_sumsq(const void* const* grad_ptrs, int rank, int world, int64_t shard_elems, int is_fp32,
                       void* out_shard, float* partial, float* norm_sq, int accumulate_norm, cudaStream_t stream);
/* reduce-scatter by in-switch reduction (multimem.ld_reduce): mc_grads = multicast address of the unit's gradients */
int llmt_comm_rs_sumsq_mc(const void* mc_grads, int rank, int world, int64_t shard_elems, void* out_shard,
                          float* partial, float* norm_sq, int accumulate_norm, cudaStream_t stream);
/* buf_ptrs: every rank's symmetric fp32 buffer of 2 * world * 32 floats (double-buffered by epoch parity) */
int llmt_comm_allreduce_f32(const void* const* buf_ptrs, const void* const* signal_pads, int rank, int world,
                            int slot_base, uint32_t epoch, float* value, int n, cudaStream_t stream);
int llmt_comm_adamw_push(float* master, float* m, float* v, const void* grad_shard, const void* const* param_ptrs,
                         int rank, int world, int64_t shard_elems, int grad_is_fp32, float lr, float beta1,
                         float beta2, float eps, float weight_decay, int step, float grad_scale, float max_norm,
                         const float* norm_sq, cudaStream_t stream);

/* Same with NVSwitch multicast: mc_params is the multicast address of the arena range (NULL -> unicast stores); one
 * multimem.st per 16 bytes leaves the GPU and the switch writes every rank's copy. */
int llmt_comm_adamw_push_mc(float* master, float* m, float* v, const void* grad_shard, const void* const* param_ptrs,
                            void* mc_params, int rank, int world, int64_t shard_elems, int grad_is_fp32, float lr,
                            float beta1, float beta2, float eps, float weight_decay, int step, float grad_scale,
                            float max_norm, const float* norm_sq, cudaStream_t stream);

/* ---- packed causal attention (src/llm_training/ops/attention_op.py:538-654, flash-attn varlen) -----------
 * qkv: bf16 [T, (Hq+2Hkv)*D]; out: bf16 [T, Hq*D]; lse: fp32 [Hq, T] (log2 units); seg_start/seg_end int32 [T].
 * bwd scratch: delta fp32 [Hq, T]; dq_acc fp32 [T, Hq*D], must be zero on entry. */
int llmt_attn_fwd(const void* qkv, void* out, float* lse, const int* seg_start, int tokens, int hq, int hkv,
                  int head_dim, int64_t qkv_stride, float scale, int window, cudaStream_t stream);
int llmt_attn_bwd(const void* qkv, const void* out, const void* dout, const float* lse, float* delta, float* dq_acc,
                  void* dqkv, const int* seg_start, const int* seg_end, int tokens, int hq, int hkv, int head_dim,
                  int64_t qkv_stride, float scale, int window, cudaStream_t stream);
/* Same with the remaining arguments of the reference's dispatcher (ops/attention_op.py:545-551):
 * softcap > 0: scores become softcap * tanh(score * scale / softcap); dropout_thr in [0, 255]: probabilities are
 * dropped with p = dropout_thr / 256 and the kept ones scaled by 1 / (1 - p).  The mask is a stateless hash of
 * (seed, head, query, key): pass the SAME softcap / dropout_thr / seed to the backward call. */
int llmt_attn_fwd_ex(const void* qkv, void* out, float* lse, const int* seg_start, int tokens, int hq, int hkv,
                     int head_dim, int64_t qkv_stride, float scale, int window, float softcap, int dropout_thr,
                     uint32_t seed, cudaStream_t stream);
int llmt_attn_bwd_ex(const void* qkv, const void* out, const void* dout, const float* lse, float* delta, float* dq_acc,
                     void* dqkv, const int* seg_start, const int* seg_end, int tokens, int hq, int hkv, int head_dim,
                     int64_t qkv_stride, float scale, int window, float softcap, int dropout_thr, uint32_t seed,
                     int* dq_order, float* dkv_acc, cudaStream_t stream);
/* dkv_acc: NULL, or an fp32 scratch [tokens, 2*hkv*head_dim] when llmt_attn_bwd_splits(tokens, hkv) > 1: with few kv
 * heads on a GPU (tensor-parallel shards) several CTAs then share each kv block and add their dK / dV tiles there. */
int llmt_attn_bwd_splits(int tokens, int hkv);
/* dq_order: NULL, or an int scratch of hq * ceil(tokens / 64) entries: the kv blocks then add their dQ contributions
 * in ascending order, which makes the backward bit-reproducible (the reference's `deterministic` flag,
 * src/llm_training/ops/attention_op.py:551,590-592). */

/* ---- RMSNorm (src/llm_training/ops/liger_kernel/rms_norm_op.py:15-19) -------------------------------------
 * bwd scratch: dw_partial fp32 [llmt_rmsnorm_bwd_num_partials(rows), cols]. */
int llmt_rmsnorm_fwd(const void* x, const void* w, void* y, float* rstd, int rows, int cols, float eps,
                     cudaStream_t stream);
int llmt_rmsnorm_bwd_num_partials(int rows);
int llmt_rmsnorm_bwd(const void* dy, const void* x, const void* w, const float* rstd, void* dx, void* dw,
                     float* dw_partial, int rows, int cols, int dw_is_fp32, int accumulate_dw, cudaStream_t stream);
/* Same, plus dres (nullable, bf16 [rows, cols]): dx = norm backward + dres (the gradient of the residual connection
 * around a pre-norm block), added in fp32 inside the kernel. */
int llmt_rmsnorm_bwd_res(const void* dy, const void* x, const void* w, const float* rstd, const void* dres, void* dx,
                         void* dw, float* dw_partial, int rows, int cols, int dw_is_fp32, int accumulate_dw,
                         cudaStream_t stream);

/* RMSNorm backward without a scratch argument.  This entry point has no counterpart in the reference.
 * It is for hosts that cannot allocate device memory or lend a buffer (all device memory pre-planned by
 * another owner, loops that cannot take cudaMalloc's implicit synchronisation, plain C callers): the dw
 * partial sums go to an array built into the library (4.625 MiB per GPU, held until process exit), so
 * cols must be <= llmt_rmsnorm_bwd_noalloc_max_cols() (8192) — larger sizes fail with an error.  The call
 * makes no cudaMalloc / cudaFree.  All calls on a device share that array: calls that overlap on the GPU
 * (different streams) race, same-stream calls are safe.  The first call on each device resolves the
 * array's address; issue it before any stream capture begins. */
int llmt_rmsnorm_bwd_noalloc_max_cols(void);
int llmt_rmsnorm_bwd_noalloc(const void* dy, const void* x, const void* w, const float* rstd, void* dx, void* dw,
                             int rows, int cols, int dw_is_fp32, int accumulate_dw, cudaStream_t stream);

/* ---- SiLU*mul, RoPE, embedding (+NEFTune) (ops/liger_kernel/swiglu_op.py:36-39, ops/rope_op.py:10-20,
 *      lms/clm/clm.py:45-78) ---------------------------------------------------------------------------- */
int llmt_silu_mul_fwd(const void* gate, const void* up, void* out, int rows, int cols, int64_t in_stride,
                      int64_t out_stride, cudaStream_t stream);
int llmt_silu_mul_bwd(const void* dout, const void* gate, const void* up, void* dgate, void* dup, int rows, int cols,
                      int64_t in_stride, int64_t dout_stride, int64_t dgu_stride, cudaStream_t stream);
/* Same pass, plus act = silu(gate) * up written to a contiguous [rows, cols] buffer (bit-identical to
 * llmt_silu_mul_fwd): the recomputed wgrad input of the down projection without a second read of gate | up. */
int llmt_silu_mul_bwd_act(const void* dout, const void* gate, const void* up, void* dgate, void* dup, void* act,
                          int rows, int cols, int64_t in_stride, int64_t dout_stride, int64_t dgu_stride,
                          cudaStream_t stream);
int llmt_rope_inplace(void* x, const int* pos, const float* cos_t, const float* sin_t, int tokens, int heads,
                      int head_dim, int64_t row_stride, int backward, cudaStream_t stream);
int llmt_embedding_fwd(const int64_t* ids, const void* weight, void* out, const float* noise_scale, int tokens,
                       int dim, int vocab, uint64_t seed, cudaStream_t stream);
int llmt_embedding_bwd(const int64_t* ids, const void* dout, void* dweight, int tokens, int dim, int vocab,
                       int64_t padding_idx, int grad_is_fp32, cudaStream_t stream);
/* Scheduling-independent variant (no atomics): ids sorted (stable) + the permutation; dweight bf16, fully overwritten. */
int llmt_embedding_bwd_sorted(const int64_t* sorted_ids, const int64_t* perm, const void* dout, void* dweight,
                              int tokens, int dim, int vocab, int64_t padding_idx, cudaStream_t stream);

/* ---- cross-entropy / token log-prob on bf16 logits (ops/liger_kernel/cross_entropy_op.py:29-33,
 *      lms/dpo/dpo.py:110-114) ---------------------------------------------------------------------------- */
int llmt_ce_fwd(const void* logits, const int64_t* labels, float* lse, float* tgt, int rows, int vocab,
                int64_t row_stride, int64_t ignore_index, float softcap, cudaStream_t stream);
int llmt_ce_bwd_inplace(void* logits, const int64_t* labels, const float* lse, const float* row_scale, int rows,
                        int vocab, int64_t row_stride, int64_t ignore_index, float softcap, cudaStream_t stream);
int llmt_ce_reduce(const float* lse, const float* tgt, const int64_t* labels, float* out2, int rows,
                   int64_t ignore_index, cudaStream_t stream);

/* ---- optimizer (optim/master_weight_wrapper.py:63-80, fsdp2_precision.py:165-169) ------------------------
 * sumsq scratch: partial fp32 [llmt_sumsq_num_partials()]. */
int llmt_sumsq_num_partials(void);
int llmt_sumsq(const void* x, int64_t n, int is_fp32, float* partial, float* out, int accumulate, cudaStream_t stream);
int llmt_adamw(float* master, float* m, float* v, const void* grad, void* param_bf16, int64_t n, int grad_is_fp32,
               float lr, float beta1, float beta2, float eps, float weight_decay, int step, float grad_scale,
               float max_norm, const float* norm_sq, cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* LLMT_KERNELS_H_ */
