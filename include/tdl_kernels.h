/* C ABI of libtdl_kernels.so — the hand-written sm_100a kernels of trustworthy_dl.
 *
 * Conventions: every function returns 0 on success, non-zero on failure (tdl_last_error() holds the
 * reason); pointers are device pointers unless stated; `stream` is a cudaStream_t; dtype codes are
 * 0 = fp32, 1 = bf16, 2 = fp16.  No function allocates device memory or synchronises the device:
 * scratch is caller-provided, so every call can be captured into a CUDA graph.
 *
 * The reference (Tanmoy058/Trustworthy-Distributed-Deep-Learning) contains no native code
 * (SURVEY.md 2.6); the comment on each declaration names the Python site whose work it absorbs.
 */
#pragma once
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* tdl_last_error(void);
int tdl_version(void);
int tdl_device_sm_count(void);

/* ---- K5: tensor statistics (reference attack_detector.py:78,185-200) ------------------------- */
size_t tdl_stats_scratch_bytes(int64_t n);
/* out12 (float64): mean,std,min,max,median,skewness,kurtosis,p25,p75,L1,L2,Linf */
int tdl_tensor_stats(const void* x, int dtype, int64_t n, double* out12, void* scratch, size_t scratch_bytes, void* stream);
/* out10 (float64): n,mean,m2,m3,m4,min,max,L1,sum(x^2),Linf */
int tdl_tensor_moments(const void* x, int dtype, int64_t n, double* out10, void* scratch, size_t scratch_bytes, void* stream);

/* Same 12 statistics without any device allocation: the scratch is a fixed array inside the library
 * (one per GPU, ~44 KB, kept until exit), so callers that cannot obtain device memory mid-run can use
 * it.  No counterpart in the reference.  n must not exceed TDL_STATS_NOALLOC_MAX_ELEMS; the first call
 * on a device performs a one-time address lookup and should happen before any stream capture; calls
 * share the array and must not overlap on the GPU (see csrc/stats.cu). */
#define TDL_STATS_NOALLOC_MAX_ELEMS (1ll << 30)
int tdl_tensor_stats_noalloc(const void* x, int dtype, int64_t n, double* out12, void* stream);

/* ---- K6: tcgen05 GEMM with fused epilogues (cuBLAS calls under distributed_trainer.py:160-161) - */
/* C[M,N] (+)= A * B^T.  a/b_mn_major: 0 = operand stored [rows, K] (K contiguous), 1 = stored [K, rows].
 * epilogue: 0 none, 1 GELU (pre-activation saved to aux_out), 2 +residual(aux_in), 3 *GELU'(aux_in),
 *           4 GELU with GELU'(pre-activation) saved to aux_out, 5 *aux_in (the backward partner of 4).
 * out_f32_atomic: accumulate into an fp32 C with red.global.add (split_k > 1 allowed). */
int tdl_gemm_bf16(const void* A, const void* B, void* C, int64_t M, int64_t N, int64_t K, int64_t lda, int64_t ldb,
                  int64_t ldc, int a_mn_major, int b_mn_major, const void* bias, int bias_f32, int epilogue,
                  const void* aux_in, void* aux_out, int64_t ld_aux, int out_f32_atomic, int split_k, int block_n,
                  void* stream);
/* Same GEMM with a bf16 output, plus colsum[n] += sum_m C[m, n] taken from the epilogue's staged panels: the
 * bias gradient of the layer whose output gradient this GEMM produces, without another pass over C. */
int tdl_gemm_bf16_colsum(const void* A, const void* B, void* C, int64_t M, int64_t N, int64_t K, int64_t lda, int64_t ldb,
                         int64_t ldc, int a_mn_major, int b_mn_major, const void* bias, int bias_f32, int epilogue,
                         const void* aux_in, void* aux_out, int64_t ld_aux, int block_n, float* colsum, void* stream);

/* ---- K8 / K7: GEMM fused with the tensor-parallel collectives over peer-mapped memory ---------- */
/* all-gather -> GEMM: A is W row shards [shard_rows, K] in peer buffers, read by TMA as the tiles are needed */
int tdl_allgather_gemm_bf16(const void* const* a_shard_ptrs, int W, int my_rank, int64_t shard_rows, const void* B, void* C,
                            int64_t N, int64_t K, int64_t lda, int64_t ldb, int64_t ldc, const void* bias, int bias_f32,
                            int epilogue, const void* aux_in, void* aux_out, int64_t ld_aux, int block_n, void* stream);
/* GEMM -> reduce-scatter: partial tiles are stored from the epilogue into the owner rank's staging slots */
int tdl_gemm_scatter_bf16(const void* A, const void* B, void* const* c_ptrs, int W, int my_rank, int64_t shard_rows,
                          int64_t N, int64_t K, int64_t lda, int64_t ldb, int64_t ldc, int b_mn_major, int block_n,
                          void* stream);
int tdl_reduce_slots(const void* staging, int W, int64_t slot_elems, const void* residual, void* out, void* stream);
/* as above for a [rows, cols] slot, adding the row-parallel layer's replicated fp32 bias[cols] */
int tdl_reduce_slots_bias(const void* staging, int W, int64_t rows, int64_t cols, const float* bias, const void* residual,
                          void* out, void* stream);
/* K8 with backward-pass layouts: B may be MN-major ([K, N]); colsum (fp32 [N], may be NULL) += column sums of the output */
int tdl_allgather_gemm_bf16_ex(const void* const* a_shard_ptrs, int W, int my_rank, int64_t shard_rows, const void* B, void* C,
                               int64_t N, int64_t K, int64_t lda, int64_t ldb, int64_t ldc, int b_mn_major, const void* bias,
                               int bias_f32, int epilogue, const void* aux_in, void* aux_out, int64_t ld_aux, int block_n,
                               float* colsum, void* stream);

/* ---- K12: causal attention, head_dim 64 (ATen/HF attention under :160-161) -------------------- */
int tdl_attention_fwd(const void* qkv, void* out, float* lse, int64_t B, int64_t T, int64_t H, int64_t Dh, void* stream);
int tdl_attention_bwd(const void* qkv, const void* out, const void* d_out, const float* lse, float* delta, void* dqkv,
                      int64_t B, int64_t T, int64_t H, int64_t Dh, void* stream);
/* as above; dbias (fp32 [3C], may be NULL) += column sums of dqkv = the qkv projection's bias gradient */
int tdl_attention_bwd_dbias(const void* qkv, const void* out, const void* d_out, const float* lse, float* delta, void* dqkv,
                            float* dbias, int64_t B, int64_t T, int64_t H, int64_t Dh, void* stream);

/* ---- K10/K11/K14/K15 and friends ---------------------------------------------------------------- */
int tdl_layernorm_fwd(const void* x, const float* w, const float* b, void* y, float* mean, float* rstd, int64_t M,
                      int64_t C, float eps, void* stream);
int tdl_layernorm_bwd(const void* dy, const void* x, const float* w, const float* mean, const float* rstd,
                      const void* dres, void* dx, float* dw, float* db, int64_t M, int64_t C, void* stream);
/* as above; dxsum (fp32 [C], may be NULL) += column sums of the bf16 dx = bias gradient of the layer below */
int tdl_layernorm_bwd_colsum(const void* dy, const void* x, const float* w, const float* mean, const float* rstd,
                             const void* dres, void* dx, float* dw, float* db, float* dxsum, int64_t M, int64_t C, void* stream);
int tdl_bias_grad(const void* dy, float* db, int64_t M, int64_t N, int64_t ld, void* stream);
int tdl_embedding_fwd(const int64_t* idx, const void* wte, const void* wpe, void* out, int64_t BT, int64_t T, int64_t C, void* stream);
int tdl_embedding_bwd(const int64_t* idx, const void* dx, float* dwte, float* dwpe, int64_t BT, int64_t T, int64_t C, void* stream);
/* optimizer_step (distributed_trainer.py:441-446): fp32 master + bf16 working copy; hyper_dev (lr,bc1,bc2,t) optional */
int tdl_adamw(float* p32, void* p16, const void* grad, int grad_dtype, float* m, float* v, int64_t n, int64_t n_decay,
              float lr, float beta1, float beta2, float eps, float wd, int64_t step, const float* grad_scale,
              const float* hyper_dev, void* stream);
int tdl_adam_tick(float* hyper_dev, float lr, float beta1, float beta2, void* stream);
/* calculate_loss (:435-439): logits bf16 [M,Vp] are overwritten with dlogits */
int tdl_xent_fwd_bwd(void* logits, const int64_t* targets, float* row_loss, float* loss_sum, int64_t M, int64_t V,
                     int64_t Vp, float grad_scale, void* stream);
int tdl_cast_f32_bf16(const float* src, void* dst, int64_t n, void* stream);
int tdl_cast_bf16_f32(const void* src, float* dst, int64_t n, void* stream);

/* ---- K1/K2/K3: trust-weighted all-reduce with fused statistics (README.md:31; attack_detector.py:109-141) */
/* scale_dev: optional device float multiplied into the bucket (fault injection: scale / sign flip) */
int tdl_pack_norm(const float* src, void* dst_bf16, int64_t n, double* sumsq, const float* scale_dev, void* stream);
size_t tdl_exchange_pad_bytes(void);
int tdl_trust_reduce(void* const* in_ptrs, void* const* out_ptrs, int W, int dtype, int64_t vec_begin, int64_t vec_end,
                     const float* coeff, double* stats_partial, void* stream);
/* as tdl_trust_reduce; store_rank >= 0: only out_ptrs[store_rank] receives the result (ZeRO-1 owner slice); mc_out: NVLS
 * multicast address of the bucket (one multimem.st instead of W peer stores) or NULL; threads: 256; max_blocks <= 0: default */
int tdl_trust_reduce_ex(void* const* in_ptrs, void* const* out_ptrs, int W, int dtype, int64_t vec_begin, int64_t vec_end,
                        const float* coeff, double* stats_partial, int store_rank, void* mc_out, int threads, int max_blocks,
                        void* stream);
/* K4: plain (sum x scale) all-reduce of a slice of a symmetric bucket: multimem.ld_reduce + multimem.st through the
 * switch when mc_ptr != NULL, P2P loads/stores otherwise (the no-security arm the detection overhead is measured against) */
int tdl_sum_allreduce(void* const* buf_ptrs, void* mc_ptr, int W, int dtype, int64_t vec_begin, int64_t vec_end, float scale,
                      int max_blocks, void* stream);
int tdl_sumsq(const void* x, int dtype, int64_t n, double* sumsq, void* stream);
int tdl_exchange(void* const* pad_ptrs, int rank, int W, int which, const double* vals, int count,
                 unsigned int* epoch_counter, int zero_vals, void* stream);
int tdl_trust_coefficients(const void* my_pad, int W, const float* trust, const int* status, const float* params12,
                           float* coeff, float* norms_out, void* stream);
/* bucketed steps: acc_sq[W] (double) / acc_bad[W] (int) accumulate squared bucket norms and per-bucket rejections over
 * the coefficient calls of a step; finalize consumes and clears them (NULL: single-bucket behaviour) */
int tdl_trust_coefficients_acc(const void* my_pad, int W, const float* trust, const int* status, const float* params12,
                               float* coeff, float* norms_out, double* acc_sq, int* acc_bad, void* stream);
int tdl_trust_finalize_acc(const void* my_pad, int W, const float* norms, const float* out_dev, const float* params12,
                           float* stats, float* trust, int* status, int* flags_out, int update_trust, double* acc_sq,
                           int* acc_bad, void* stream);
int tdl_trust_finalize(const void* my_pad, int W, const float* norms, const float* out_dev, const float* params12,
                       float* stats, float* trust, int* status, int* flags_out, int update_trust, void* stream);

/* ---- K9: parameter all-gather with fused checksum + drift ("Param Monitor", README.md:22) -------- */
int tdl_param_gather(const void* const* shard_ptrs, int W, int dtype, void* dst, void* prev, int64_t shard_elems,
                     void* sums_u64, double* drift, void* stream);

/* device-side verdict of the gather: what the owners published (read back from the exchange pad) against the
 * checksums folded while pulling; flags[i] |= 1 when shard i does not match */
int tdl_exchange_read(const void* my_pad, int W, int which, int count, void* out_u64, void* stream);
int tdl_integrity_compare(const void* seen_u64, const void* published_u64, int W, int* flags, void* stream);
/* graph-capturable parameter-poisoning hook: adds *level_dev * N(0,1) to a bf16 shard while *level_dev != 0 */
int tdl_poison_if(void* x_bf16, int64_t n, const float* level_dev, int64_t seed, void* stream);

/* ---- K13: conv epilogue (ResNet/VGG stages left unimplemented at distributed_trainer.py:137-144) */
int tdl_conv_epilogue_fwd(const void* x, const float* scale, const float* shift, const void* residual, void* y, int64_t N,
                          int64_t C, int64_t HW, int dtype, int channels_last, int relu, void* stream);
int tdl_conv_epilogue_bwd(const void* dy, const void* y, const void* x, const float* scale, void* dx, void* dres,
                          float* dscale, float* dshift, int64_t N, int64_t C, int64_t HW, int dtype, int channels_last,
                          int relu, void* stream);

/* ---- misc security primitives ------------------------------------------------------------------- */
int tdl_multi_sqnorm(const int64_t* ptrs_dev, const int64_t* numels_dev, int num, int dtype, float* out, void* stream);
int tdl_checksum64(const void* words, int64_t nwords, void* out2_u64, void* stream);
int tdl_drift_update(const void* cur, void* prev, int64_t n, int dtype, float* out, void* stream);
int tdl_poison_inplace(void* x, int64_t n, int dtype, float a, float b, int64_t seed, void* stream);

#ifdef __cplusplus
}
#endif
