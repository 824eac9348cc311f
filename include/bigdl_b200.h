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

/* ---- tcgen05 GEMM (csrc/gemm_tcgen05.cu). Reference: MKL sgemm JNI, core/mkl2017-xeon-blas/src/main/c/mkl.c;
 *      S/nn/Linear.scala:78-167. D[M,N] = opA(A) * opB(B)^T.
 *      a_mn = 0: A is [M,K] row-major (lda >= K);  a_mn = 1: A is [K,M] row-major (lda >= M).
 *      b_mn = 0: B is [N,K] row-major (ldb >= K);  b_mn = 1: B is [K,N] row-major (ldb >= N).
 *      out_fp32 selects fp32 (else bf16) output with row pitch ldd. splits <= 0 picks split-K automatically
 *      (only with BDL_GEMM_ATOMIC_ADD, which accumulates into an fp32 D with red.global.add). */
#define BDL_GEMM_BIAS 1
#define BDL_GEMM_RELU 2
#define BDL_GEMM_ATOMIC_ADD 4
int bdl_gemm_bf16(const void* A, const void* B, void* D, const float* bias, int M, int N, int K, long long lda,
                  long long ldb, long long ldd, int a_mn, int b_mn, int out_fp32, int flags, int splits, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* BIGDL_B200_H_ */
