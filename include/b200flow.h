/* b200flow — C ABI of the sm_100a kernel library (libb200flow_cuda.so).
 *
 * Every entry point takes raw device pointers plus the CUDA stream to launch on
 * (`void* stream` == cudaStream_t) and returns 0 on success; on failure the
 * message is available from b2f_last_error() (thread-local).  bf16 tensors are
 * passed as `void*`.  Nothing here allocates device memory: scratch buffers are
 * supplied by the caller and their sizes are queried with the *_num_partials /
 * *_scratch_bytes helpers.
 *
 * The reference repository contains no native code; each group
 * below cites the upstream native component it stands in for.
 */
#ifndef B200FLOW_H_
#define B200FLOW_H_

#ifdef __cplusplus
extern "C" {
#endif

const char* b2f_last_error(void);

/* ---- K0: tcgen05 GEMM ------------------------------------------------------
 * C[b] = epilogue(alpha * op(A[b]) * op(B[b])), bf16 operands, fp32 accumulate.
 *   a_mn_major = 0: A stored [M, K] (K contiguous);  1: A stored [K, M].
 *   b_mn_major = 0: B stored [N, K] (K contiguous);  1: B stored [K, N].
 * flags: 1 gelu, 2 C is fp32, 4 C += (fp32 only), 8 relu, 16 bias per row.
 * Stands in for cuBLAS + apex fused_dense / fused_weight_gradient_mlp
 * (containers/megatron-deepspeed/Dockerfile:6-11). */
int b2f_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, long long lda,
                  long long ldb, long long ldc, int a_mn_major, int b_mn_major, int batch,
                  long long strideA, long long strideB, long long strideC, const void* bias,
                  const void* residual, long long ldr, long long strideR, void* aux,
                  long long ldaux, long long strideAux, float alpha, int flags, void* stream);

/* Same contract on CTA pairs (cta_group::2, 256x256 tile per pair, 6 stages). */
int b2f_gemm_bf16_2cta(const void* A, const void* B, void* C, int M, int N, int K, long long lda,
                       long long ldb, long long ldc, int a_mn_major, int b_mn_major, int batch,
                       long long strideA, long long strideB, long long strideC, const void* bias,
                       const void* residual, long long ldr, long long strideR, void* aux,
                       long long ldaux, long long strideAux, float alpha, int flags, void* stream);

/* fp8 (e4m3) GEMM for the MoE experts: C = (Aq Bq^T) * row_scale[m] * col_scale[n] with the usual
 * epilogue; operands quantised row-wise by b2f_quantize_rows_e4m3 (scale = amax / 448). */
int b2f_gemm_fp8(const void* Aq, const void* Bq, void* C, int M, int N, int K, long long lda,
                 long long ldb, long long ldc, const float* row_scale, const float* col_scale,
                 const void* bias, const void* residual, long long ldr, void* aux, long long ldaux,
                 int flags, void* stream);
int b2f_quantize_rows_e4m3(const void* x, void* q, float* scale, long long rows, int K, long long ldx,
                           void* stream);

/* ---- K8: LayerNorm / RMSNorm (apex fused_layer_norm_cuda) ------------------ */
int b2f_norm_fwd(const void* x, const void* residual, const void* gamma, const void* beta, void* y,
                 void* h_out, float* mean, float* rstd, int rows, int H, float eps, int rms,
                 void* stream);
int b2f_norm_bwd_num_partials(int rows);
/* scratch: fp32 [b2f_norm_bwd_num_partials(rows) * 2 * H] */
int b2f_norm_bwd(const void* dy, const void* x, const void* gamma, const float* mean,
                 const float* rstd, const void* dres, void* dx, float* dgamma, float* dbeta,
                 float* scratch, int rows, int H, int rms, int accumulate, void* stream);

/* Allocation-free norm backward.  No counterpart in the reference: it exists for callers
 * that must not (or cannot) allocate device memory while running — a host that owns the
 * whole GPU through its own pool and passes libraries raw pointers only, a process held to
 * a fixed memory partition, a loop that cannot take the device synchronisation an
 * allocation may cause.  Same arguments as b2f_norm_bwd without `scratch`: the partial
 * sums go to a 9.25 MiB file-scope device array that every GPU loading this library holds
 * until exit.  H <= b2f_norm_bwd_noalloc_max_h() (8192), otherwise an error is returned and
 * nothing runs.  The first call per device resolves the array's device address; issue it
 * before any stream capture.  Calls share the array: calls that can overlap on the GPU race. */
int b2f_norm_bwd_noalloc_max_h(void);
int b2f_norm_bwd_noalloc(const void* dy, const void* x, const void* gamma, const float* mean,
                         const float* rstd, const void* dres, void* dx, float* dgamma,
                         float* dbeta, int rows, int H, int rms, int accumulate, void* stream);

/* ---- K6 / K11: optimizer family (apex amp_C, DeepSpeed FusedAdam) ---------- */
int b2f_adamw_flat(float* master, float* m, float* v, const void* grad, int grad_bf16,
                   void* param_out, long long n, float lr, float beta1, float beta2, float eps,
                   float weight_decay, int step, const float* grad_scale, const int* skip,
                   void* stream);
int b2f_adamw_multi(const void* segs, int nsegs, long long total_chunks, int grad_bf16, float lr,
                    float beta1, float beta2, float eps, float weight_decay, int step,
                    const float* grad_scale, const int* skip, void* stream);
int b2f_sgd_momentum_flat(float* master, float* buf, const void* grad, int grad_bf16, void* param_out,
                          long long n, float lr, float momentum, float weight_decay,
                          const float* grad_scale, void* stream);
int b2f_multi_tensor_chunk(void);
int b2f_sumsq_num_partials(long long n);
/* scratch: fp32 [b2f_sumsq_num_partials(n)] */
int b2f_sumsq(const void* x, long long n, int dtype_fp32, float* scratch, float* out, int* overflow,
              int accumulate, void* stream);
int b2f_clip_coef(const float* sumsq, float max_norm, float extra_scale, float* coef_out,
                  float* norm_out, void* stream);
int b2f_scale(void* x, long long n, int dtype_fp32, const float* scale_ptr, float scale,
              void* stream);

/* ---- K10 + RoPE + cross-entropy (Megatron fused_bias_gelu, bias_dropout_add,
 *      apex fused_rotary_positional_embedding, scaled_masked_softmax) --------- */
int b2f_swiglu_fwd(const void* gu, void* out, long long rows, int F, void* stream);
int b2f_swiglu_bwd(const void* gu, const void* dout, void* dgu, long long rows, int F, void* stream);
int b2f_gelu_fwd(const void* x, const void* bias, void* y, long long rows, int H, void* stream);
int b2f_gelu_bwd(const void* x, const void* bias, const void* dy, void* dx, long long rows, int H,
                 void* stream);
int b2f_bias_dropout_add(const void* x, const void* bias, const void* residual, void* out,
                         void* mask, long long rows, int H, float p, unsigned long long seed,
                         void* stream);
int b2f_dropout_bwd(const void* dy, const void* mask, void* dx, long long n, float p, void* stream);
/* position of token t = (t / pos_div) % seq + pos_offset (pos_div = batch for [s, b] order) */
int b2f_rope(void* x, const float* cos_t, const float* sin_t, long long tokens, int seq, int nh,
             int d, long long token_stride, int pos_offset, int pos_div, int backward,
             void* stream);
int b2f_cross_entropy(const void* logits, const long long* targets, float* loss, void* dlogits,
                      long long rows, int V, long long ld, float grad_scale,
                      long long ignore_index, void* stream);
/* K9: cross-entropy over vocab-sharded logits (Megatron vocab_parallel_cross_entropy): per-row
 * statistics {max, sum-exp, target logit} stored into slot `rank` of a [world][T][3] fp32 buffer on
 * this GPU and (through mapped symmetric memory) on every peer, then loss + in-place gradient. */
int b2f_vocab_ce_stats(const void* logits, const long long* targets, long long T, int Vl, long long ld,
                       long long lo, int rank, int world, const long long* stats_ptrs, void* stream);
int b2f_vocab_ce_finish(void* logits, const long long* targets, long long T, int Vl, long long ld,
                        long long lo, int world, const float* stats, long long ignore_index,
                        float grad_scale, float* loss, void* stream);
int b2f_scaled_masked_softmax(const void* x, const void* mask, void* y, long long batch, int heads,
                              int sq, int sk, float scale, int causal, void* stream);
int b2f_cast_f32_bf16(const float* x, void* y, long long n, float scale, void* stream);

/* ---- K7: flash attention fwd/bwd on tcgen05 (flash-attn v2 / Megatron fused softmax).
 * q/k/v/o: [B, S, heads, D] views, heads packed, D in {64,128}; *_ts token stride,
 * *_bs batch stride (elements).  lse: fp32 [B, H, S].  delta: fp32 scratch of
 * 2 * B * H * S_pad floats (S_pad = S rounded up to 128), 16-byte aligned. */
int b2f_attn_fwd(const void* q, const void* k, const void* v, void* o, float* lse, int B, int S,
                 int H, int HKV, int D, long long q_ts, long long q_bs, long long k_ts,
                 long long k_bs, long long v_ts, long long v_bs, long long o_ts, long long o_bs,
                 float scale, int causal, const int* kv_lens, void* stream);
int b2f_attn_bwd(const void* q, const void* k, const void* v, const void* o, const void* d_o,
                 const float* lse, float* delta, void* dq, void* dk, void* dv, int B, int S, int H,
                 int HKV, int D, long long q_ts, long long q_bs, long long k_ts, long long k_bs,
                 long long v_ts, long long v_bs, long long o_ts, long long o_bs, long long do_ts,
                 long long do_bs, long long dq_ts, long long dq_bs, long long dk_ts, long long dk_bs,
                 long long dv_ts, long long dv_bs, float scale, int causal, const int* kv_lens,
                 void* stream);

/* The same kernels with attention dropout (Megatron `--attention-dropout`, GPT-2 345M values files:
 * examples/megatron-deepspeed/gpt2_345m/pretrain-ddp-zero1.yaml).  The keep mask is a counter-based hash of
 * (seed, batch*head, query, key): no mask tensor is stored, forward and both backward kernels regenerate the same
 * bits; kept probabilities are scaled by 1 / (1 - p_drop). */
int b2f_attn_fwd_drop(const void* q, const void* k, const void* v, void* o, float* lse, int B, int S,
                      int H, int HKV, int D, long long q_ts, long long q_bs, long long k_ts,
                      long long k_bs, long long v_ts, long long v_bs, long long o_ts, long long o_bs,
                      float scale, int causal, const int* kv_lens, float p_drop, unsigned seed,
                      void* stream);
int b2f_attn_bwd_drop(const void* q, const void* k, const void* v, const void* o, const void* d_o,
                      const float* lse, float* delta, void* dq, void* dk, void* dv, int B, int S,
                      int H, int HKV, int D, long long q_ts, long long q_bs, long long k_ts,
                      long long k_bs, long long v_ts, long long v_bs, long long o_ts, long long o_bs,
                      long long do_ts, long long do_bs, long long dq_ts, long long dq_bs,
                      long long dk_ts, long long dk_bs, long long dv_ts, long long dv_bs, float scale,
                      int causal, const int* kv_lens, float p_drop, unsigned seed, void* stream);

/* ---- K1 / K2: gradient collectives over NVLink peer memory, fused with scale / Adam
 * (torch DDP Reducer + NCCL, Horovod fusion buffer, DeepSpeed ZeRO-1 phases).
 * *_ptrs: host arrays of `world` device addresses of the same symmetric buffer on every
 * rank; flags: symmetric u32 region, zeroed once; epoch: monotonically increasing. */
int b2f_comm_max_blocks(void);
int b2f_allreduce_f32(const long long* grads_ptrs, long long multicast_ptr,
                      const long long* flags_ptrs, int rank, int world, long long off, long long n,
                      float scale, void* bf16_out, int blocks, unsigned epoch, void* stream);
int b2f_zero1_reduce(const long long* grads_ptrs, const long long* flags_ptrs, int rank, int world,
                     long long lo, long long n, float* reduced, float* partial, float* sumsq_out,
                     int blocks, unsigned epoch, void* stream);
int b2f_zero1_adam_gather(const float* reduced, float* master, float* m, float* v,
                          const long long* params_ptrs, const long long* flags_ptrs, int rank,
                          int world, long long lo, long long n, float lr, float beta1, float beta2,
                          float eps, float weight_decay, int step, const float* grad_scale,
                          const int* skip, int blocks, unsigned epoch, void* stream);
/* K1 with half-precision on the wire: the fp32 bucket is cast into a symmetric staging buffer, reduced there
 * (multimem.ld_reduce when stage_multicast_ptr != 0, peer loads otherwise) and written back as fp32 * scale
 * (torch DDP `bf16_compress_hook`; NCCL moves half the bytes). */
int b2f_allreduce_bf16wire(void* grads, const long long* stage_ptrs, long long stage_multicast_ptr,
                           const long long* flags_ptrs, int rank, int world, long long off, long long n,
                           float scale, int blocks, unsigned epoch, void* stream);
/* K2 through the switch: *_multicast_ptr != 0 reduces with multimem.ld_reduce and publishes the updated
 * half-precision shard with multimem.st; 0 falls back to the peer-pointer loops of the entry points above. */
int b2f_zero1_reduce_mc(const long long* grads_ptrs, long long grads_multicast_ptr,
                        const long long* flags_ptrs, int rank, int world, long long lo, long long n,
                        float* reduced, float* partial, float* sumsq_out, int blocks, unsigned epoch,
                        void* stream);
int b2f_zero1_adam_gather_mc(const float* reduced, float* master, float* m, float* v,
                             const long long* params_ptrs, long long params_multicast_ptr,
                             const long long* flags_ptrs, int rank, int world, long long lo, long long n,
                             float lr, float beta1, float beta2, float eps, float weight_decay, int step,
                             const float* grad_scale, const int* skip, int blocks, unsigned epoch,
                             void* stream);
int b2f_peer_barrier(const long long* flags_ptrs, int rank, int world, unsigned epoch, void* stream);
int b2f_peer_sum_test(const long long* buf_ptrs, const long long* flags_ptrs, int rank, int world,
                      long long n, float* out, int blocks, unsigned epoch, void* stream);

/* ---- K3 / K4: tensor-parallel GEMMs fused with the sequence-parallel collective
 * (Megatron Column/RowParallelLinear: NCCL all-gather + cuBLAS, cuBLAS + NCCL reduce-scatter). */
int b2f_ag_gemm_bf16(const long long* x_shard_ptrs, void* xg, const void* B, void* Y, int T, int N,
                     int K, long long ldb, long long ldy, int b_mn_major, const void* bias,
                     const long long* ready_ptrs, void* tile_flags, int rank, int world,
                     int comm_ctas, unsigned epoch, void* stream);
int b2f_gemm_rs_bf16(const void* X, const void* B, void* Y_shard, int T, int N, int K,
                     long long ldx, long long ldb, long long ldy, int b_mn_major, const void* bias,
                     const void* residual, long long ldr, const long long* recv_ptrs,
                     const long long* flag_ptrs, int rank, int world, unsigned epoch, void* stream);
/* K4': GEMM -> all-reduce, Y[T, N] complete on every rank (Megatron RowParallelLinear without
 * sequence parallelism and the dgrad of ColumnParallelLinear: cuBLAS + NCCL all-reduce).
 * out_ptrs: every rank's [T, N] bf16 output in symmetric memory; residual: this rank's own rows
 * [T/world, N]; done_ptrs: symmetric u32[world]; cta_counter: local u32 that is zero between calls. */
int b2f_gemm_ar_bf16(const void* X, const void* B, const long long* out_ptrs, int T, int N, int K,
                     long long ldx, long long ldb, long long ldy, int b_mn_major, const void* bias,
                     const void* residual, long long ldr, const long long* recv_ptrs,
                     const long long* flag_ptrs, const long long* done_ptrs, void* cta_counter,
                     int rank, int world, unsigned epoch, void* stream);

/* ---- K12-K15: Mask R-CNN detection kernels (tf.image.crop_and_resize / non_max_suppression,
 * tensorpack model_box / model_frcnn / model_mrcnn).  Features are NHWC bf16, one pointer per
 * FPN level (P2..P5); rois: fp32 [R, 5] = (image, x1, y1, x2, y2). */
int b2f_roi_align_fwd(const void* const* feats, const int* H, const int* W, const float* scales,
                      int levels, int C, const float* rois, int R, int P, int sampling, int aligned,
                      void* out, void* stream);
int b2f_roi_align_bwd(const void* const* feats, float* const* grads, const int* H, const int* W,
                      const float* scales, int levels, int C, const float* rois, int R, int P,
                      int sampling, int aligned, const void* dout, void* stream);
/* y = act(y + bias[c] (+ residual)) in place, NHWC half precision, C % 8 == 0: the epilogue of a convolution whose
 * frozen batch norm was folded into the weights (tensorpack BACKBONE.NORM=FreezeBN). */
int b2f_bias_act_nhwc(void* y, const void* bias, const void* residual, long long numel, int C, int relu,
                      void* stream);
/* Column sums of a half-precision [rows, C] matrix (bias gradient of a linear layer, or of a convolution over an NHWC
 * gradient) as fp32 partial rows partial[b2f_colsum_num_partials(rows, C)][C]; the caller adds the partial rows. */
int b2f_colsum_num_partials(long long rows, int C);
int b2f_colsum_partial(const void* g, float* partial, long long rows, int C, void* stream);
int b2f_box_decode(const float* anchors, const float* deltas, float* out, long long n, float wx,
                   float wy, float ww, float wh, float img_w, float img_h, void* stream);
/* mask_scratch: u64 [total_boxes * ceil(max_seg / 64)] */
int b2f_nms_batched(const float* boxes, const int* seg_off, int num_segs, int max_seg, float iou_thr,
                    unsigned long long* mask_scratch, unsigned char* keep, void* stream);
int b2f_match_proposals(const float* props, const int* prop_batch, long long n, const float* gts,
                        const int* gt_off, float* best_iou, int* best_idx, void* stream);
int b2f_pairwise_iou(const float* a, long long na, const float* b, int nb, float* out, void* stream);
int b2f_mask_crop(const unsigned char* masks, int H, int W, const float* rois, const int* gt_index,
                  int R, int M, float* out, void* stream);
int b2f_mask_paste(const float* probs, int M, const float* boxes, int R, int H, int W, float thr,
                   unsigned char* out, void* stream);

/* K16: polygons -> binary masks (pycocotools frPoly): even-odd test at pixel centres, parts OR-ed. */
int b2f_poly_to_mask(const float* xy, const int* poly_off, const int* obj_poly, int N, int H, int W,
                     unsigned char* masks, void* stream);

/* ---- K5a + MoE token movement (DeepSpeed top1gating + einsum dispatch/combine + all_to_all).
 * With peer_ptrs the dispatch stores rows straight into the expert owner's buffer
 * [src_rank, local_expert, capacity, H] and the combine pulls them back: the all-to-alls are
 * done by these kernels over NVLink. */
long long b2f_moe_router_scratch_bytes(int T, int E);
int b2f_moe_router_top1(const float* logits, int T, int E, int capacity, int* expert, int* slot,
                        float* gate, float* probs, int* counts, float* me, float* ce, void* scratch,
                        void* stream);
int b2f_moe_dispatch(const void* x, const int* expert, const int* slot, int T, int H, int capacity,
                     void* dst_local, const long long* peer_ptrs, int world, int experts_per_rank,
                     int src_rank, void* stream);
int b2f_moe_combine(const void* src_local, const long long* peer_ptrs, int world,
                    int experts_per_rank, int src_rank, const int* expert, const int* slot,
                    const float* gate, int T, int H, int capacity, void* y, void* stream);
int b2f_moe_gate_grad(const void* rows, const int* expert, const int* slot, const void* dy, int T,
                      int H, int capacity, float* dgate, void* stream);

/* ---- Block-scaled fp8 (MXFP8): e4m3 values with one UE8M0 scale per 32 elements along K, multiplied by
 * tcgen05.mma kind::mxf8f6f4.block_scale with the scale factors in TMEM (Transformer Engine MXFP8 recipe;
 * DeepSpeed-MoE expert GEMMs of examples/megatron-deepspeed/gpt2_moe/).
 * Scale-factor layout ("atoms"): sf[k/128][row/128][512 B], byte (row%32)*16 + ((row%128)/32)*4 + (k%128)/32;
 * rows are padded to 128 and K to 128 (b2f_mx_sf_bytes). */
int b2f_mx_sf_atoms(int rows);
long long b2f_mx_sf_bytes(int rows, int K);
int b2f_quantize_mx(const void* x, void* q, void* sf, int rows, int K, long long ldx, int batch,
                    long long x_batch, void* stream);
/* x [N, K] -> MX copy of x^T (rows K, contracting over N; N % 128 == 0) without materialising the transpose. */
int b2f_quantize_mx_t(const void* x, void* q, void* sf, int N, int K, long long ldx, int batch,
                      long long x_batch, void* stream);
int b2f_dequantize_mx(const void* q, const void* sf, void* y, int rows, int K, long long ldy, int batch,
                      long long y_batch, void* stream);
/* C[b] = epilogue(A[b] B[b]^T): A [M, K], B [N, K] quantised as above; flags as b2f_gemm_bf16. */
int b2f_gemm_mx(const void* Aq, const void* sfa, const void* Bq, const void* sfb, void* C, int M, int N,
                int K, int batch, long long ldc, long long strideC, const void* bias, long long bias_batch,
                const void* residual, long long ldr, long long strideR, void* aux, long long ldaux,
                long long strideAux, int flags, void* stream);

/* ---- K5: the expert FFN of an expert-parallel MoE layer as two fused kernels per direction
 * (DeepSpeed MOELayer: all_to_all -> experts -> all_to_all, three NCCL / cuBLAS phases upstream).
 *   up:   tokens leave as e4m3 + scales into the owners' symmetric buffers (bulk stores over NVLink) and the owner's
 *         GEMM 1 starts on a 128-row unit as soon as its flag shows the epoch; epilogue + bias, keep, GELU
 *         (backward: x gelu'(pre)), requantise for GEMM 2.
 *   down: GEMM 2, epilogue + bias, rows stored straight into the source rank's home buffer; a release counter per
 *         source tells it when all of its rows have landed (b2f_moe_wait_rows).
 * C % 128 == 0, H % 512 == 0.  *_ptrs: `world` addresses of the same symmetric buffer on every rank. */
int b2f_moe_token_of(const int* expert, const int* slot, int T, int C, int* token_of, void* stream);
int b2f_moe_up(const void* x, const int* token_of, const int* counts, int T, int H, int E, int C,
               const long long* xq_ptrs, const long long* xsf_ptrs, const long long* flag_ptrs,
               const long long* cnt_ptrs, const void* Wq, const void* wsf, int N, const void* bias,
               void* keep, const void* pre, void* hq, void* hsf, int backward, int rank, int world,
               unsigned epoch, void* stream);
int b2f_moe_down(const void* hq, const void* hsf, const void* Wq, const void* wsf, int K, int N,
                 const void* bias, const long long* rows_ptrs, const long long* done_ptrs,
                 const void* my_cnt, int E, int C, int rank, int world, unsigned epoch, void* stream);
int b2f_moe_wait_rows(const void* done, const int* counts, int E, int N, void* stream);
int b2f_moe_zero_inactive(void* buf, const void* my_cnt, int El, int world, int C, int N, void* stream);

#ifdef __cplusplus
}
#endif

/* ======================= host runtime: libb200flow_host.so ========================================
 * C++ (no CUDA) pieces of the runtime, loaded by b200flow/ops/_lib.py the same way. */
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* Gradient bucket planning and the ready-order bookkeeping of DDP (torch Reducer, Horovod fusion buffer).
 * plan rows: {first_slot, last_slot, elem_lo, elem_hi}.  `mark` returns the next bucket that may launch
 * (buckets launch strictly in index order on every rank) or -1; `next_ready` continues after a launch. */
int b2f_plan_buckets(const int64_t* offsets, const int64_t* numels, int nslots, int64_t total_elems,
                     int64_t elem_bytes, int64_t cap_bytes, int64_t first_cap_bytes, int64_t* out,
                     int max_buckets);
void* b2f_reducer_create(const int64_t* plan, int nb, int nslots, const int32_t* uses);
void b2f_reducer_destroy(void* h);
void b2f_reducer_reset(void* h);
int b2f_reducer_num_buckets(void* h);
int b2f_reducer_mark(void* h, int slot);
int b2f_reducer_next_ready(void* h);
int b2f_reducer_next_unlaunched(void* h);

/* Megatron dataset index builders (megatron/data/helpers.cpp, compiled by `make` in the reference image:
 * containers/megatron-deepspeed/Dockerfile). */
void b2f_build_doc_idx(int32_t* out, const int32_t* documents, long long num_docs, int32_t num_epochs,
                       uint64_t seed);
long long b2f_build_sample_idx(const int32_t* sizes, const int32_t* doc_idx, long long doc_idx_len,
                               int32_t seq_length, long long num_samples, int64_t* sample_idx);
void b2f_build_shuffle_idx(int64_t* out, long long n, uint64_t seed);
void b2f_build_blending_indices(uint8_t* dataset_index, int64_t* dataset_sample_index,
                                const double* weights, int32_t num_datasets, long long size);
long long b2f_build_bert_mapping(const int64_t* docs, long long num_docs, const int32_t* sizes,
                                 int32_t num_epochs, long long max_num_samples, int32_t max_seq_length,
                                 double short_seq_prob, uint64_t seed, int32_t min_num_sent,
                                 int32_t long_sentence_len, int64_t* maps);

/* Shared-memory message channel between data-loader workers and the trainer (torch DataLoader's
 * shared-memory queues / tensorpack's ZMQ pipes). */
void* b2f_chan_create(const char* name, uint64_t capacity);
void* b2f_chan_open(const char* name);
int b2f_chan_send(void* ch, const void* buf, uint32_t len, long timeout_ms);
long long b2f_chan_recv(void* ch, void* buf, uint32_t buf_len, long timeout_ms, uint32_t* need_out);
void b2f_chan_close(void* ch);
void b2f_chan_free(void* ch);

#ifdef __cplusplus
}
#endif
#endif /* B200FLOW_H_ */
