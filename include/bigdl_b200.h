/* C ABI of libbigdl_b200.so — the sm_100a kernel library behind bigdl_b200.
 *
 * Conventions: every pointer is a device pointer unless its name starts with h_; `stream` is a cudaStream_t
 * passed as void*; every function returns BDL_OK (0) or a negative error code and records a message readable
 * through bdl_last_error(). Nothing here synchronises the device; all work is enqueued on `stream`.
 * This library takes the place of the reference's JNI natives (core/mkl2017-xeon-blas, core/native-dnn,
 * core/bigquant): same role, different machine.
 */
#ifndef BIGDL_B200_H_
#define BIGDL_B200_H_

#ifdef __cplusplus
extern "C" {
#endif

#define BDL_OK 0
#define BDL_ERR_INVALID (-1)
#define BDL_ERR_CUDA (-2)

const char* bdl_last_error(void);
int bdl_version(void);
int bdl_sm_count(void);
/* 0 none, 1 capturing, 2 invalidated; diagnostic for CUDA-graph capture of training steps. */
int bdl_stream_capture_status(void* stream);

/* ---- tcgen05 GEMM (csrc/gemm_tcgen05.cu). Reference: MKL sgemm JNI, core/mkl2017-xeon-blas/src/main/c/mkl.c;
 *      S/nn/Linear.scala:78-167. D[M,N] = opA(A) * opB(B)^T.
 *      a_mn = 0: A is [M,K] row-major (lda >= K);  a_mn = 1: A is [K,M] row-major (lda >= M).
 *      b_mn = 0: B is [N,K] row-major (ldb >= K);  b_mn = 1: B is [K,N] row-major (ldb >= N).
 *      out_fp32 selects fp32 (else bf16) output with row pitch ldd. splits <= 0 picks split-K automatically
 *      (only with BDL_GEMM_ATOMIC_ADD, which accumulates into an fp32 D with red.global.add). */
#define BDL_GEMM_BIAS 1
#define BDL_GEMM_RELU 2
#define BDL_GEMM_ATOMIC_ADD 4
/* Residual operand of the NEXT bdl_gemm_bf16 / bdl_conv_igemm_bf16 (fprop) / bdl_gemm_fp8 launch issued by this host
 * thread: D = act(A*B + bias + residual), residual bf16 [M,N] with row pitch ld (elements). Cleared by that launch.
 * Reference: the conv + sum post-op of S/nn/mkldnn/Fusion.scala (fusionCAddTable). */
/* Per-column (sum, sum of squares) of the NEXT bf16 GEMM / conv fprop output on this thread, added into sums [2*N] fp32 by
 * the epilogue (batch-norm statistics without a pass over the output). bdl_gemm_stats_done(): 1 if that launch did it. */
int bdl_gemm_set_residual_mask(int on);   /* the armed residual masks instead of adding: out = residual > 0 ? out : 0 */
int bdl_gemm_set_stats(float* sums);
int bdl_gemm_stats_done(void);
int bdl_gemm_set_residual(const void* residual_bf16, long long ld);
/* CTA-pair tiles (tcgen05.mma.cta_group::2 on 2-CTA clusters: 256 rows per MMA, half of the B tile per SM) for GEMM,
 * conv fprop and dgrad: -1 = where the wave heuristic wants them (default; env BDL_GEMM_TWO), 0 = never, 1 = wherever
 * legal. Process-wide; returns the previous mode. Reference: the same S/nn/Linear.scala:78-167 /
 * S/nn/SpatialConvolution.scala:253-521 GEMMs as bdl_gemm_bf16. */
int bdl_gemm_set_pair_mode(int mode);
/* Block-scaled fp8: e4m3 operands with one UE8M0 (power-of-two) scale per (row, 32 consecutive K elements).
 * bdl_quantize_mxfp8 writes the e4m3 bytes (q, row pitch ldq, padding pre-zeroed) and the scale bytes (sf:
 * ceil(rows/128) * ceil(cols/128) atoms of 512 bytes, offset (r%32)*16 + (r%128)/32*4 + kb%4 inside an atom — the layout
 * tcgen05.cp 32x128b.warpx4 moves to TMEM). bdl_gemm_mxfp8 multiplies two such operands with
 * tcgen05.mma.kind::mxf8f6f4.block_scale (scales applied by the tensor core, per block) and adds bias / ReLU.
 * Reference: BigQuant's per-window scales (core/bigquant/native/ops/shuffle/shuffle_im2col.h),
 * S/nn/quantized/Quantization.scala:35-62, S/nn/quantized/Linear.scala. */
int bdl_quantize_mxfp8(const void* x, void* q, void* sf, long long rows, long long cols, long long ldx, long long ldq,
                       int is_bf16, void* stream);
int bdl_gemm_mxfp8(const void* A, const void* B, const void* sfa, const void* sfb, void* D, const float* bias, int M,
                   int N, int K, long long lda, long long ldb, long long ldd, int out_fp32, int flags, void* stream);
int bdl_gemm_bf16(const void* A, const void* B, void* D, const float* bias, int M, int N, int K, long long lda,
                  long long ldb, long long ldd, int a_mn, int b_mn, int out_fp32, int flags, int splits, void* stream);
/* Implicit-GEMM convolution on the same kernel: the gathered operand is fetched from the NHWC activation tensor by
 * TMA im2col loads (no column matrix). mode 0 fprop, 1 dgrad (caller passes the transposed-conv geometry, stride 1),
 * 2 wgrad (C % 64 == 0, accumulates with red.add). Reference: S/nn/SpatialConvolution.scala:253-521,
 * S/nn/NNPrimitive.scala im2col/col2im, S/nn/mkldnn/SpatialConvolution.scala:435-633. */
/* Linear forward with the all-gather of its row-sharded weight fused in: n tiles [t0[r], t1[r]) (tiles of block_n rows)
 * are loaded from peer_B[r] (rank r's copy, a peer-mapped pointer; plain non-multicast TMA loads) and written through to
 * B_local, all other tiles come from B_local. splits / ATOMIC_ADD as bdl_gemm_bf16.
 * Reference: AllReduceParameter.getWeights (S/parameters/AllReduceParameter.scala:201-228) fused into
 * S/nn/Linear.scala:78-104. */
int bdl_gemm_bf16_gathered(const void* A, void* B_local, void* D, const float* bias, int M, int N, int K, long long lda,
                           long long ldb, long long ldd, int out_fp32, int flags, int splits, int block_n,
                           const void* const* peer_B, const int* t0, const int* t1, int nranks, int rank, void* stream);
int bdl_conv_igemm_bf16(int mode, const void* act, int N, int H, int W, int C, long long ld_act, const void* other,
                        long long ld_other, void* D, long long ldd, const float* bias, int O, int Ho, int Wo, int kh,
                        int kw, int sh, int sw, int ph, int pw, int dh, int dw, int out_fp32, int flags, int splits,
                        void* stream);

/* ---- elementwise / reductions on bf16 activations (csrc/elementwise.cu).
 *      Reference: S/nn/ReLU.scala, S/nn/Dropout.scala:40-75, S/nn/CAddTable.scala,
 *      S/parameters/FP16CompressedTensor.scala:80-98,152-196,271-278. */
#include <stddef.h>
int bdl_cast_f32_to_bf16(const float* src, void* dst, size_t n, int truncate, void* stream);
int bdl_cast_bf16_to_f32(const void* src, float* dst, size_t n, void* stream);
int bdl_bf16_trunc_add(const void* a, const void* b, void* out, size_t n, void* stream);
/* op codes: 0 relu, 1 relu_bwd(a=dy,b=y), 2 add, 3 add_relu, 4 mul, 5 tanh, 6 sigmoid, 7 tanh_bwd, 8 sigmoid_bwd,
 *           9 scale(alpha), 10 axpy(alpha*a+b), 11 sub, 12 max, 13 relu6 */
int bdl_eltwise_bf16(int op, const void* a, const void* b, void* out, size_t n, float alpha, void* stream);
/* out[c] += sum_m x[m,c]; with y != NULL: dz = (y>0 ? x : 0) is written and summed (fused ReLU-bwd + bias grad). */
/* out = sum of `count` (2..8) dense bf16 tensors of n elements (n % 8 == 0), fp32 accumulate */
int bdl_addn_bf16(const void* const* ptrs, int count, void* out, size_t n, void* stream);
/* dz = (y > 0) ? sum of `count` (2..4) dense [M,C] bf16 tensors : 0, and out[c] += column sums of dz */
int bdl_colsum_sumn_relu_bf16(const void* const* ptrs, int count, const void* y, void* dz, float* out, long long M,
                              int C, void* stream);
int bdl_colsum_bf16(const void* x, const void* y, void* dz, float* out, long long M, int C, long long ldx,
                    long long ldy, long long ldz, void* stream);
int bdl_dropout_fwd_bf16(const void* x, void* y, void* mask, size_t n, float p, unsigned long long seed,
                         unsigned long long offset, void* stream);
/* Graph-replay friendly variant: Philox position read from *counter_dev (u64) and advanced by the call. */
int bdl_dropout_fwd_bf16_ctr(const void* x, void* y, void* mask, size_t n, float p, unsigned long long seed,
                             void* counter_dev, void* stream);
int bdl_counter_add(void* counter_dev, unsigned long long v, void* stream);
/* dst[k][0:len[k]) += src[off[k] : off[k]+len[k]) for up to 8 segments, then src is cleared (persistent scratch): the
 * bias gradients of the branch tails of a Concat out of one fused column-sum vector.
 * Reference: accGradParameters of S/nn/SpatialConvolution.scala:429-521 (gradBias). */
int bdl_scatter_acc_clear_f32(float* src, void* const* dst, const int* off, const int* len, int count, void* stream);
/* y = act(acc + bias) in bf16 from the fp32 accumulator of a split-K forward GEMM; re-zeroes acc for the next launch.
 * Reference: the bias add of S/nn/Linear.scala:93-103 (+ the fused ReLU of S/nn/mkldnn/Fusion.scala). */
int bdl_bias_act_finish(float* acc, const float* bias, void* y, long long M, int N, long long ld_acc, long long ldy,
                        int relu, void* stream);
int bdl_dropout_bwd_bf16(const void* dy, const void* mask, void* dx, size_t n, float p, void* stream);
int bdl_copy2d_bf16(const void* src, void* dst, long long rows, int C, long long lds, long long ldd, int accumulate,
                    void* stream);

/* ---- input pipeline on the GPU (csrc/data_ops.cu): uint8 HWC frames -> crop + flip + normalise -> bf16/fp32 batch.
 *      crop_xy: [N,2] int32 device (x0,y0) or NULL = centre crop; flip: [N] uint8 device or NULL. */
int bdl_image_batch_u8(const void* src_u8, void* dst, const int* crop_xy, const void* flip, int N, int H, int W, int C,
                       int Hc, int Wc, const float* mean, const float* std_, int out_fp32, int out_nchw, void* stream);

/* ---- NHWC im2col / col2im (csrc/conv_im2col.cu). Reference: S/nn/NNPrimitive.scala:24-1719. */
int bdl_im2col_nhwc_bf16(const void* x, void* col, int N, int H, int W, int C, int Ho, int Wo, int kh, int kw, int sh,
                         int sw, int ph, int pw, int dh, int dw, long long ldx, int Kp, void* stream);
/* space-to-depth of a stride-2 RGB stem input, horizontal taps folded into channels: [N,R,Wo,tw*16] bf16 (see
   conv_im2col.cu) */
int bdl_s2d_stem_bf16(const void* x, int x_is_fp32, void* y, int N, int C, int H, int W, long long sn, long long sc,
                      long long sh, long long sw, int pt, int pl, int R, int Wo, int tw, void* stream);
int bdl_col2im_nhwc_bf16(const void* col, void* dx, int N, int H, int W, int C, int Ho, int Wo, int kh, int kw, int sh,
                         int sw, int ph, int pw, int dh, int dw, long long ldx, int Kp, void* stream);

/* ---- pooling / LRN (csrc/pool_lrn.cu). Reference: S/nn/SpatialMaxPooling.scala, SpatialAveragePooling.scala,
 *      SpatialCrossMapLRN.scala:204-237, S/nn/mkldnn/{MaxPooling,AvgPooling,LRN}.scala. */
int bdl_maxpool_fwd_nhwc_bf16(const void* x, void* y, void* idx, int N, int H, int W, int C, int Ho, int Wo, int kh,
                              int kw, int sh, int sw, int ph, int pw, long long ldx, long long ldy, void* stream);
int bdl_maxpool_bwd_nhwc_bf16(const void* dy, const void* idx, void* dx, int N, int H, int W, int C, int Ho, int Wo,
                              int kh, int kw, int sh, int sw, int ph, int pw, long long ldx, long long ldy,
                              void* stream);
int bdl_avgpool_fwd_nhwc_bf16(const void* x, void* y, int N, int H, int W, int C, int Ho, int Wo, int kh, int kw,
                              int sh, int sw, int ph, int pw, long long ldx, long long ldy, int count_include_pad,
                              int divide, void* stream);
int bdl_avgpool_bwd_nhwc_bf16(const void* dy, void* dx, int N, int H, int W, int C, int Ho, int Wo, int kh, int kw,
                              int sh, int sw, int ph, int pw, long long ldx, long long ldy, int count_include_pad,
                              int divide, void* stream);
int bdl_lrn_fwd_nhwc_bf16(const void* x, void* y, long long pixels, int C, long long ldx, long long ldy, int size,
                          float alpha, float beta, float k, void* stream);
int bdl_lrn_bwd_nhwc_bf16(const void* x, const void* dy, void* dx, long long pixels, int C, long long ldx,
                          long long ldy, long long lddx, int size, float alpha, float beta, float k, void* stream);

/* SoftMax / LogSoftMax over the last dimension of a [rows, C] matrix (one warp per row, fp32 math; dtype codes 0 = fp32,
 * 1 = bf16; log_form selects LogSoftMax) and their backward from the OUTPUT y.
 * Reference: S/nn/SoftMax.scala, S/nn/LogSoftMax.scala:37-120, S/nn/mkldnn/SoftMax.scala. */
int bdl_softmax_fwd(const void* x, void* y, long long rows, int C, long long ldx, long long ldy, int in_bf16, int out_bf16,
                    int log_form, void* stream);
int bdl_softmax_bwd(const void* y, const void* dy, void* dx, long long rows, int C, long long ldy, long long ldd,
                    long long ldx, int y_bf16, int d_bf16, int log_form, void* stream);
/* ---- fused LogSoftMax + ClassNLL (csrc/loss_softmax.cu). Reference: S/nn/ClassNLLCriterion.scala:70-242,
 *      S/nn/CrossEntropyCriterion.scala:35-53, S/optim/ValidationMethod.scala:173,827.
 *      scratch: 3*B floats kept between fwd and bwd; out: {loss, total weight}. Targets are 1-based. */
int bdl_xent_fwd(const void* logits, const void* target, const float* class_w, float* scratch, float* out, long long B,
                 int C, long long ld, int logits_bf16, int target_i64, int padding_value, int size_average,
                 void* stream);
int bdl_xent_bwd(const void* logits, const void* target, const float* class_w, const float* scratch, const float* out,
                 void* dlogits, long long B, int C, long long ld, long long ldd, int logits_bf16, int target_i64,
                 int padding_value, int size_average, float gscale, void* stream);
/* Allocation-free variant of bdl_xent_fwd: no counterpart in the reference. For hosts that may not allocate device
 * memory at call time (pre-reserved arenas, real-time loops, plug-ins without access to the owner's allocator): the
 * 3*B-float scratch lives in a fixed array inside the library, so B is capped at 65536 rows (larger B is rejected) and
 * there is no backward. The first call on each device resolves the array's device address — make it before any stream
 * capture. All calls on a device share that array: calls that can overlap on the GPU race. out: {loss, total weight}. */
int bdl_xent_fwd_noalloc(const void* logits, const void* target, const float* class_w, float* out, long long B, int C,
                         long long ld, int logits_bf16, int target_i64, int padding_value, int size_average,
                         void* stream);
int bdl_topk_correct(const void* logits, const void* target, void* counts, long long B, int C, long long ld,
                     int logits_bf16, int target_i64, void* stream);

/* ---- flat OptimMethod update (csrc/optim.cu, csrc/optim_update.cuh). Reference: S/optim/{SGD,Adam,Adagrad,RMSprop,
 *      Adadelta,Adamax,Ftrl,LarsSGD}.scala; S/parameters/Util.scala:30-50.
 *      kind: 0 SGD, 1 Adam, 2 Adagrad, 3 RMSprop, 4 Adadelta, 5 Adamax, 6 Ftrl, 7 LARS.
 *      hp_host/hp_dev point at an OptimHP (bdl_optim_hp_size() bytes); hp_dev (device memory) wins when non-NULL. */
int bdl_optim_hp_size(void);
int bdl_optim_update(int kind, float* w, float* g, float* s0, float* s1, void* w16, size_t n, const void* hp_host,
                     const void* hp_dev, int truncate, int zero_grad, void* stream);
int bdl_sumsq_f32(const float* x, size_t n, float* out, void* stream);
int bdl_absmax(const void* x, size_t n, int is_bf16, float* out, void* stream);
int bdl_scale_f32(float* x, size_t n, float alpha, const float* norm_sq, float max_norm, void* stream);

/* ---- fused NVLink reduce-scatter + OptimMethod update + all-gather (csrc/comm_fused.cu).
 *      Reference: S/parameters/AllReduceParameter.scala:201-328, S/parameters/FP16CompressedTensor.scala:152-196,
 *      S/optim/DistriOptimizer.scala:358-396.
 *      mode bits: 1 reduce, 2 write fp32 grad slice, 4 update, 8 publish bf16 weights to all peers,
 *                 16 begin barrier, 32 end barrier, 64 reference-exact (truncating) arithmetic, 128 sum of squares,
 *                 256 divide by hp.gscale instead of multiplying.
 *      epoch_dev (optional, u64 in device memory) is added to `epoch` so a captured graph can advance the barrier
 *      generation between replays with bdl_counter_add. */
int bdl_compress_grads(float* g, void* out16, size_t n, int truncate, int zero_src, void* stream);
int bdl_fused_rs_update(const void* const* grad16_ptrs, void* const* w16_ptrs, void* const* flag_ptrs, int nranks,
                        int rank, long long start, long long length, float* master, float* grad_part, float* s0,
                        float* s1, int kind, const void* hp_host, const void* hp_dev, int mode, unsigned int epoch,
                        void* local_sync, float* sumsq, int max_blocks, const void* epoch_dev, void* stream);
/* cudaMalloc + cudaIpc handle exchange for the symmetric buffers (one process per GPU). handle64: 64 bytes. */
/* Bucketed form of the above, one launch per gradient bucket from inside backward (communication stream):
 * [compress my fp32 bucket -> bf16] -> begin barrier -> P2P reduce of my sub-slice -> average (-> device-side clip
 * scale) -> OptimMethod -> bf16 weights stored into every peer's compute buffer (CM_PUBLISH; [hole_start, +hole_len) and
 * CM_PUBLISH_LOCAL stay local for a consumer that gathers them itself) -> "ready" generation into every peer's flag
 * area (no end barrier; consumers acquire it per bucket, bdl_pull_gather). mode = CM_* bits (csrc/comm_fused.cu); generation = *epoch_dev * gen_mul + gen_add; flag_off = element offset
 * of this bucket's 3*nranks flag words; local_sync = 3 words shared by the stream-ordered launches of one plane.
 * Reference: S/optim/ParallelOptimizer.scala:481-513 (per-bucket exchange inside backward) applied to the sharded
 * update of S/optim/DistriOptimizer.scala:358-396 and S/parameters/AllReduceParameter.scala:236-328. */
int bdl_bucket_rs_update(const void* const* grad16_ptrs, void* const* w16_ptrs, void* const* flag_ptrs, int nranks,
                         int rank, long long b_start, long long b_len, long long start, long long length,
                         float* grads32, float* master, float* grad_part, float* s0, float* s1, int kind,
                         const void* hp_host, const void* hp_dev, int mode, unsigned int gen_mul, unsigned int gen_add,
                         const void* epoch_dev, long long flag_off, void* local_sync, float* sumsq,
                         const float* clip_scale, long long hole_start, long long hole_len, int max_blocks,
                         void* stream);
/* Pull side of the all-gather: copy remote sub-slices (segs = nsegs x {peer, start, len}) from their owners' compute
 * buffers into mine, each after acquiring the owner's ready generation of (*epoch_dev - prev_steps).
 * Reference: AllReduceParameter.getWeights, S/parameters/AllReduceParameter.scala:201-228. */
int bdl_pull_gather(void* const* w16_ptrs, const void* my_flags, int nranks, int rank, const long long* segs, int nsegs,
                    unsigned int gen_mul, unsigned int gen_add, int prev_steps, const void* epoch_dev,
                    long long flag_off, int max_blocks, void* stream);
/* Global-norm clipping on the device: exchange per-rank partial sums of squares through peer memory and write
 * scale = min(1, max_norm / ||g||) where the update kernels read it (CM_CLIP_DEV). Resets *sumsq.
 * Reference: L2NormClippingProcessor, S/parameters/ParameterOperations.scala:94-130. */
int bdl_peer_clip_scale(void* const* slot_ptrs, void* const* flag_ptrs, int nranks, int rank, unsigned int gen_mul,
                        unsigned int gen_add, const void* epoch_dev, long long flag_off, float* sumsq, float max_norm,
                        float* scale_out, float* norm_out, void* stream);
int bdl_ipc_alloc(size_t bytes, void** ptr, void* handle64);
int bdl_ipc_open(const void* handle64, void** ptr);
int bdl_ipc_close(void* ptr);
int bdl_ipc_free(void* ptr);

/* ---- batch normalisation, NHWC bf16 (csrc/batchnorm.cu). Reference: S/nn/BatchNormalization.scala:40-110,
 *      SpatialBatchNormalization.scala:485-494, S/nn/mkldnn/SpatialBatchNormalization.scala:44-83.
 *      sums buffers are [2*C] fp32 and must be zero on entry; between bdl_bn_stats and bdl_bn_finalize they may be
 *      all-reduced across ranks (sync-BN). */
int bdl_bn_stats(const void* x, float* sums, long long M, int C, long long ldx, void* stream);
int bdl_bn_finalize(const float* sums, float count, int C, float eps, float momentum, float* save_mean,
                    float* save_invstd, float* running_mean, float* running_var, void* stream);
int bdl_bn_apply(const void* x, void* y, const float* mean, const float* invstd_or_var, const float* gamma,
                 const float* beta, long long M, int C, long long ldx, long long ldy, float eps, int use_var, int relu,
                 void* stream);
int bdl_bn_bwd_reduce(const void* dy, const void* x, const void* y, const float* mean, const float* invstd,
                      float* sums, long long M, int C, long long lddy, long long ldx, long long ldy, void* stream);
int bdl_bn_bwd_apply(const void* dy, const void* x, const void* y, const float* mean, const float* invstd,
                     const float* gamma, const float* sums, void* dx, long long M, float count, int C, long long lddy,
                     long long ldx, long long ldy, long long lddx, void* stream);
/* backward of BN + fused ReLU with the mask recomputed from x (y > 0 <=> fma(x, invstd*gamma, beta - mean*invstd*gamma) > 0,
 * the forward's own arithmetic) instead of read back from the output: one tensor read less per pass */
int bdl_bn_bwd_reduce_remask(const void* dy, const void* x, const float* mean, const float* invstd, const float* gamma,
                             const float* beta, float* sums, long long M, int C, long long lddy, long long ldx,
                             void* stream);
int bdl_bn_bwd_apply_remask(const void* dy, const void* x, const float* mean, const float* invstd, const float* gamma,
                            const float* beta, const float* sums, void* dx, long long M, float count, int C,
                            long long lddy, long long ldx, long long lddx, void* stream);
/* residual-block fusions: y = relu?(bn(x) + res) in one pass; bwd_apply that also emits dy masked by (y > 0) */
int bdl_bn_apply_res(const void* x, const void* res, void* y, const float* mean, const float* invstd_or_var,
                     const float* gamma, const float* beta, long long M, int C, long long ldx, long long ldr,
                     long long ldy, float eps, int use_var, int relu, void* stream);
int bdl_bn_bwd_apply_mask(const void* dy, const void* x, const void* y, const float* mean, const float* invstd,
                          const float* gamma, const float* sums, void* dx, void* dmask, long long M, float count, int C,
                          long long lddy, long long ldx, long long ldy, long long lddx, long long lddm, void* stream);
int bdl_bn_param_grad(const float* sums, float* gw, float* gb, int C, float scale_w, float scale_b, void* stream);

/* ---- fused LSTM gate math (csrc/recurrent.cu). Reference: S/nn/LSTM.scala:150-175, S/nn/mkldnn/RNN.scala.
 *      Gate order along 4H: input, candidate, forget, output. c is kept in fp32. */
int bdl_lstm_gates_fwd(const void* g, const float* c_prev, void* h, float* c, void* gates, long long B, int H,
                       long long ldg, void* stream);
int bdl_lstm_gates_bwd(const void* dh, const float* dc_in, const void* gates, const float* c, const float* c_prev,
                       void* dg, float* dc_prev, long long B, int H, void* stream);
/* Persistent LSTM over a whole sequence (csrc/lstm_seq.cu): ONE cooperative launch per layer and direction runs all T
 * steps with W_hh resident in shared memory, split across the SMs by hidden unit; the input projection xw = x W_ih^T + b
 * [B,T,4H] is computed beforehand by one GEMM. Forward keeps hs [B,T,H] (hs_hread >= H: readable zero-padded row width,
 * for 16-byte staging loads), cs [T,B,H] and the activated gates [T,B,4H];
 * backward (BPTT, 2-D split of W_hh) produces dG [T,B,4,dg_hp] (dg_hp >= H: zero-padded, 16-byte aligned gate blocks),
 * the gradient w.r.t. the gate pre-activations, from which
 * dW_ih, dW_hh, db and dx are single GEMMs. counter: one zeroed uint32; dh_rec [2,B,H] / dc [B,H]: zeroed fp32 scratch
 * (or the upstream dh_T in slot 0 / dc_T); at return dh_rec slot (T & 1) = dh0, dc = dc0. Gate order [i, g, f, o].
 * Reference: S/nn/Recurrent.scala:243-527, S/nn/LSTM.scala:44-175, S/nn/mkldnn/RNN.scala:31-100. */
int bdl_lstm_seq_fwd(const void* xw, long long xw_sb, long long xw_st, const void* whh, const void* h0, const float* c0,
                     void* hs, long long hs_sb, long long hs_st, int hs_hread, float* cs, void* gates, float* c_last,
                     int B, int T, int H, int reverse, void* counter, void* stream);
int bdl_lstm_seq_bwd(const void* dhs, long long dhs_sb, long long dhs_st, const void* whh, const void* gates,
                     const float* cs, const float* c0, void* dG, int dg_hp, float* dh_rec, float* dc, int B, int T,
                     int H, int reverse, void* counter, void* stream);
/* GRU step, two pointwise stages around the recurrent GEMMs (S/nn/GRU.scala): all tensors bf16; ld* = row pitches */
int bdl_gru_rz_fwd(const void* xrz, const void* hg, const void* h, void* rz, void* hr, long long B, int H,
                   long long ldx, long long ldh, void* stream);
int bdl_gru_rz_bwd(const void* d_hr, const void* d_z, const void* rz, const void* h, void* d_pre, void* d_h,
                   long long B, int H, long long ldh, void* stream);
int bdl_gru_out_fwd(const void* xc, const void* hc, const void* z, const void* h, void* hh, void* out, long long B,
                    int H, long long ldx, long long ldz, long long ldh, void* stream);
int bdl_gru_out_bwd(const void* d_out, const void* hh, const void* z, const void* h, void* d_pre_c, void* d_z,
                    void* d_h, long long B, int H, long long ldz, long long ldh, void* stream);

/* ---- fp8 (e4m3) quantised inference (csrc/gemm_tcgen05.cu kFp8, csrc/quantize.cu). Reference: core/bigquant/native
 *      (MixPrecisionGEMM, quantize), S/nn/quantized/*, S/nn/MklInt8Convertible.scala, S/nn/mkldnn/SpatialConvolution.scala
 *      :237-247,325-372. D = (A_q * B_q^T) * col_scale[n] (+bias)(ReLU); conv_geom = NULL for a plain GEMM or 14 ints
 *      {N,H,W,C,Ho,Wo,kh,kw,sh,sw,ph,pw,dh,dw} to gather A from an NHWC e4m3 tensor with TMA im2col. out_fp32: 0 bf16, 1 fp32,
 *      2 e4m3 requantised with the device scalar *out_qscale (= 448 / absmax of the consumer's input). */
int bdl_gemm_fp8(const void* A, const void* B, void* D, const float* bias, const float* col_scale,
                 const float* out_qscale, const void* conv_geom, int M, int N, int K, long long lda, long long ldb, long long ldd, int out_fp32, int flags,
                 void* stream);
int bdl_row_absmax(const void* x, float* out, long long rows, long long cols, long long ld, int is_bf16, void* stream);
int bdl_quantize_e4m3(const void* x, void* q, const float* absmax, int per_row, long long rows, long long cols,
                      long long ldx, long long ldq, int is_bf16, void* stream);
int bdl_fp8_colscale(const float* act_absmax, const float* w_absmax, float* out, int n, void* stream);
int bdl_quantize_int8_rows(const float* w, void* q, const float* absmax, long long rows, long long cols, void* stream);

/* ---- fused multi-head attention forward (csrc/attention.cu): O = softmax(scale * Q K^T + bias) V per (batch, head), the
 *      logits stay in TMEM / shared memory. Reference: S/nn/Attention.scala:95-114. Q [B,Lq,heads*D], K/V [B,Lk,heads*D]
 *      bf16 (pitches in elements, multiples of 8), D = 64 or 128; bias fp32 with element strides over (batch, head,
 *      query), key stride 1, 0 = broadcast, NULL = none; causal masks key > query + (Lk - Lq); lse [B,heads,Lq] fp32. */
int bdl_attention_fwd(const void* Q, long long q_sb, long long q_sl, const void* K, long long k_sb, long long k_sl,
                      const void* V, long long v_sb, long long v_sl, const float* bias, long long bias_sb,
                      long long bias_sh, long long bias_sq, void* O, long long o_sb, long long o_sl, float* lse, int B,
                      int heads, int Lq, int Lk, int D, float scale, int causal, void* stream);
/* forward variant: -1 = heuristic (two 128-query tiles per CTA when the grid still fills the SMs), 0 = one tile per CTA,
 * 1 = two tiles wherever Lq > 128. Returns the previous mode. */
int bdl_attention_set_two_mode(int mode);
/* gradients w.r.t. Q, K, V (csrc/attention_bwd.cu): probabilities recomputed from lse; delta = [B,heads,Lq] fp32 scratch */
int bdl_attention_bwd(const void* Q, long long q_sb, long long q_sl, const void* K, long long k_sb, long long k_sl,
                      const void* V, long long v_sb, long long v_sl, const void* O, long long o_sb, long long o_sl,
                      const void* dO, long long do_sb, long long do_sl, const float* lse, float* delta,
                      const float* bias, long long bias_sb, long long bias_sh, long long bias_sq, void* dQ,
                      long long dq_sb, long long dq_sl, void* dK, long long dk_sb, long long dk_sl, void* dV,
                      long long dv_sb, long long dv_sl, int B, int heads, int Lq, int Lk, int D, float scale,
                      int causal, void* stream);

/* ---- NVLS (in-switch) all-reduce over a MULTICAST address (csrc/nvls.cu): multimem.ld_reduce / multimem.st. `mc` is the
 *      multicast mapping of a symmetric buffer (same offset on every rank); callers bracket each call with a group barrier.
 *      Reference: the aggregate step of S/utils/DistriParameterSynchronizer.scala:234-356. */
int bdl_nvls_reduce_bf16_to_f32(const void* mc, float* out, long long n, float scale, void* stream);
int bdl_nvls_allreduce_bf16_inplace(void* mc, long long n, int rank, int world, void* stream);
int bdl_nvls_allreduce_f32(const void* mc, float* out, long long n, float scale, void* stream);

/* ---- fp32 / fp64 BLAS + vector math for Tensor[Float] / Tensor[Double] (csrc/blas_vml.cu; host twins bdlh_* in
 *      csrc/host/blas_vml.cpp). Reference: core/mkl2017-xeon-blas/src/main/c/mkl.c (vsAdd ... vsgemm, vd*), called from
 *      S/tensor/TensorNumeric.scala:253-420. Column-major like BLAS; offsets are folded into the pointers.
 *      vml op: 0 Add 1 Sub 2 Mul 3 Div 4 Powx(a, scalar) 5 Ln 6 Exp 7 Sqrt 8 Tanh 9 Abs 10 Log1p; b is NULL for unary ops.
 *      dot writes its result to device memory (*result), deterministic summation order. */
#define BDL_DECLARE_BLAS(SUF, T)                                                                                       \
  int bdl_vml_##SUF(int op, long long n, const T* a, const T* b, T scalar, T* y, void* stream);                         \
  int bdl_blas_axpy_##SUF(long long n, T alpha, const T* x, long long incx, T* y, long long incy, void* stream);        \
  int bdl_blas_scal_##SUF(long long n, T alpha, T* x, long long incx, void* stream);                                    \
  int bdl_blas_dot_##SUF(long long n, const T* x, long long incx, const T* y, long long incy, T* result, void* stream); \
  int bdl_blas_ger_##SUF(int m, int n, T alpha, const T* x, long long incx, const T* y, long long incy, T* A,           \
                         long long lda, void* stream);                                                                 \
  int bdl_blas_gemv_##SUF(int trans, int m, int n, T alpha, const T* A, long long lda, const T* x, long long incx,      \
                          T beta, T* y, long long incy, void* stream);                                                 \
  int bdl_blas_gemm_##SUF(int transa, int transb, int m, int n, int k, T alpha, const T* A, long long lda, const T* B,  \
                          long long ldb, T beta, T* C, long long ldc, void* stream);
BDL_DECLARE_BLAS(f32, float)
BDL_DECLARE_BLAS(f64, double)

#ifdef __cplusplus
}
#endif
#endif /* BIGDL_B200_H_ */
