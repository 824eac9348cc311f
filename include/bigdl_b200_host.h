/* C ABI of the host runtime library (libbigdl_b200_host.so, csrc/host/*.cpp). No CUDA dependency.
 *
 * Counterparts in the reference: the JNI natives under core/ (bigquant int8 kernels, MKL VML/BLAS glue) and the JVM
 * utilities S/utils/{ThreadPool,RandomGenerator,Crc32}.scala, S/parameters/FP16CompressedTensor.scala,
 * S/dataset/image/MTLabeledBGRImgToBatch.scala, netty's Crc32c used by S/visualization/tensorboard/RecordWriter.scala. */
#ifndef BIGDL_B200_HOST_H_
#define BIGDL_B200_HOST_H_
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* ---- thread pool (S/utils/ThreadPool.scala, Engine.default/Engine.model) */
/* CPU affinity / topology (csrc/host/affinity.cpp). Reference: C/native-dnn/src/main/c/affinity.c, native_utils.c */
int bdlh_set_affinity(const int* cpus, int n);
int bdlh_get_affinity(int* cpus, int cap);
int bdlh_logical_cores(void);
int bdlh_physical_cores(void);
int bdlh_pool_size(void);
void bdlh_pool_resize(int threads);

/* ---- checksums */
uint32_t bdlh_crc32c(const void* data, size_t n, uint32_t seed);   /* Castagnoli; seed 0 for a fresh checksum */
uint32_t bdlh_crc32c_masked(const void* data, size_t n);           /* TFRecord mask: rot15(crc) + 0xa282ead8 */
uint32_t bdlh_crc32(const void* data, size_t n, uint32_t seed);    /* zlib polynomial (S/utils/Crc32.scala) */

/* ---- TFRecord framing (S/visualization/tensorboard/RecordWriter.scala, S/utils/tf/TFRecordIterator.scala) */
size_t bdlh_tfrecord_frame(const void* payload, size_t n, void* out /* n + 16 bytes */);
/* Returns payload length and sets *offset to the payload start, or -1 at a clean EOF, -2 on a corrupt record. */
long long bdlh_tfrecord_next(const void* buf, size_t buf_len, size_t pos, size_t* offset, int verify);

/* ---- truncated-bf16 ("fp16") gradient codec (S/parameters/FP16CompressedTensor.scala:150-270), multi-threaded */
void bdlh_bf16_compress(const float* src, uint16_t* dst, size_t n);          /* keep the upper 16 bits (truncate) */
void bdlh_bf16_decompress(const uint16_t* src, float* dst, size_t n);
void bdlh_bf16_add(uint16_t* acc, const uint16_t* other, size_t n);          /* acc = trunc(acc + other), parAdd */

/* ---- Mersenne-Twister stream with Torch7 semantics (S/utils/RandomGenerator.scala:60-260) */
void* bdlh_rng_new(uint64_t seed);
void bdlh_rng_free(void* rng);
void bdlh_rng_seed(void* rng, uint64_t seed);
uint32_t bdlh_rng_u32(void* rng);
double bdlh_rng_uniform(void* rng, double a, double b);
double bdlh_rng_normal(void* rng, double mean, double stdv);
void bdlh_rng_fill_uniform(void* rng, float* dst, size_t n, double a, double b);
void bdlh_rng_fill_normal(void* rng, float* dst, size_t n, double mean, double stdv);
void bdlh_rng_permutation(void* rng, int64_t* dst, size_t n);               /* Fisher-Yates, 0-based */

/* ---- mini-batch assembly (S/dataset/image/MTLabeledBGRImgToBatch.scala:60-130, BGRImgCropper, HFlip, Normalizer)
 * images: n pointers to uint8 HWC images (h[i] x w[i] x 3). Each is cropped to crop_h x crop_w (random offset from
 * the per-sample 64-bit seeds when random_crop, centre otherwise), optionally mirrored, normalised
 * (x/scale - mean[c]) / std[c] and written CHW (or HWC when nhwc) as fp32 or bf16 into out. Threads split the batch. */
int bdlh_batch_images(const uint8_t* const* images, const int* h, const int* w, int n, int crop_h, int crop_w,
                      int random_crop, int random_flip, const uint64_t* seeds, const float* mean, const float* stdv,
                      float scale, int nhwc, int out_bf16, void* out);
/* Gather n fixed-size byte records into one contiguous (pinned) buffer with the pool. */
void bdlh_gather_rows(const void* const* rows, size_t row_bytes, int n, void* out);

/* ---- BigQuant reference arithmetic on the CPU (core/bigquant: MixPrecisionGEMM; S/nn/quantized/Quantization.scala)
 * C[m,n] = sum_k A_q[m,k]*B_q[n,k] * (a_scale[m]/127) * (b_scale[n]/127) + bias[n]; int32 accumulation. */
void bdlh_quantize_rows_int8(const float* src, int8_t* dst, float* absmax, int rows, int cols);
void bdlh_int8_gemm(const int8_t* A, const float* a_absmax, const int8_t* B, const float* b_absmax, const float* bias,
                    float* C, int M, int N, int K);

const char* bdlh_version(void);

#ifdef __cplusplus
}
#endif
#endif /* BIGDL_B200_HOST_H_ */
