// Fused LSTM / GRU gate kernels: the recurrent GEMMs run on the tcgen05 GEMM (gemm_tcgen05.cu); these kernels do all of
// the per-element gate math of one time step in a single pass (forward and backward), so a step is
// "1 GEMM + 1 pointwise launch" instead of the reference's ~15-module cell graph
// (S/nn/LSTM.scala:150-175 buildLSTM: CAddTable, Reshape, 4x Select, 3x Sigmoid, 2x Tanh, 3x CMulTable, CAddTable)
// and the analogue of MKL-DNN's fused rnn cell (S/nn/mkldnn/RNN.scala).
// Gate order in the 4H axis follows the reference: [input, candidate, forget, output].
#include "bigdl_b200.h"
#include "host_common.h"
#include "ptx.cuh"

namespace bdl {

__device__ __forceinline__ float sigm(float v) { return 1.f / (1.f + __expf(-v)); }

// g: [B, 4H] bf16 pre-activations (x-projection + h-projection), c_prev: [B, H] fp32
// outputs: h [B, H] bf16, c [B, H] fp32, gates [B, 4H] bf16 (activated, kept for backward)
__global__ void lstm_fwd_kernel(const __nv_bfloat16* __restrict__ g, const float* __restrict__ c_prev,
                                __nv_bfloat16* __restrict__ h, float* __restrict__ c,
                                __nv_bfloat16* __restrict__ gates, long long B, int H, long long ldg) {
  const long long total = B * H;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const long long b = i / H;
    const int j = static_cast<int>(i - b * H);
    const __nv_bfloat16* row = g + b * ldg;
    const float ig = sigm(__bfloat162float(row[j]));
    const float zg = tanhf(__bfloat162float(row[H + j]));
    const float fg = sigm(__bfloat162float(row[2 * H + j]));
    const float og = sigm(__bfloat162float(row[3 * H + j]));
    const float cn = ig * zg + fg * c_prev[i];
    c[i] = cn;
    h[i] = __float2bfloat16_rn(tanhf(cn) * og);
    __nv_bfloat16* gr = gates + b * 4 * H;
    gr[j] = __float2bfloat16_rn(ig);
    gr[H + j] = __float2bfloat16_rn(zg);
    gr[2 * H + j] = __float2bfloat16_rn(fg);
    gr[3 * H + j] = __float2bfloat16_rn(og);
  }
}

// dh [B,H] bf16 (grad of h), dc_in [B,H] fp32 (grad flowing into c from the next step), gates, c, c_prev
// outputs: dg [B,4H] bf16 (grad of pre-activations), dc_prev [B,H] fp32
__global__ void lstm_bwd_kernel(const __nv_bfloat16* __restrict__ dh, const float* __restrict__ dc_in,
                                const __nv_bfloat16* __restrict__ gates, const float* __restrict__ c,
                                const float* __restrict__ c_prev, __nv_bfloat16* __restrict__ dg,
                                float* __restrict__ dc_prev, long long B, int H) {
  const long long total = B * H;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const long long b = i / H;
    const int j = static_cast<int>(i - b * H);
    const __nv_bfloat16* gr = gates + b * 4 * H;
    const float ig = __bfloat162float(gr[j]), zg = __bfloat162float(gr[H + j]);
    const float fg = __bfloat162float(gr[2 * H + j]), og = __bfloat162float(gr[3 * H + j]);
    const float tc = tanhf(c[i]);
    const float dhv = dh ? __bfloat162float(dh[i]) : 0.f;
    const float dcv = (dc_in ? dc_in[i] : 0.f) + dhv * og * (1.f - tc * tc);
    __nv_bfloat16* dr = dg + b * 4 * H;
    dr[j] = __float2bfloat16_rn(dcv * zg * ig * (1.f - ig));
    dr[H + j] = __float2bfloat16_rn(dcv * ig * (1.f - zg * zg));
    dr[2 * H + j] = __float2bfloat16_rn(dcv * c_prev[i] * fg * (1.f - fg));
    dr[3 * H + j] = __float2bfloat16_rn(dhv * tc * og * (1.f - og));
    dc_prev[i] = dcv * fg;
  }
}

static int rgrid(long long total) {
  const long long cap = static_cast<long long>(bdl_num_sms()) * 8;
  long long b = (total + 255) / 256;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return static_cast<int>(b);
}

}  // namespace bdl

using namespace bdl;

extern "C" int bdl_lstm_gates_fwd(const void* g, const float* c_prev, void* h, float* c, void* gates, long long B,
                                  int H, long long ldg, void* stream) {
  if (B <= 0 || H <= 0) return BDL_OK;
  lstm_fwd_kernel<<<rgrid(B * H), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(g), c_prev, static_cast<__nv_bfloat16*>(h), c,
      static_cast<__nv_bfloat16*>(gates), B, H, ldg);
  return bdl_check_launch("lstm_fwd_kernel");
}

extern "C" int bdl_lstm_gates_bwd(const void* dh, const float* dc_in, const void* gates, const float* c,
                                  const float* c_prev, void* dg, float* dc_prev, long long B, int H, void* stream) {
  if (B <= 0 || H <= 0) return BDL_OK;
  lstm_bwd_kernel<<<rgrid(B * H), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(dh), dc_in, static_cast<const __nv_bfloat16*>(gates), c, c_prev,
      static_cast<__nv_bfloat16*>(dg), dc_prev, B, H);
  return bdl_check_launch("lstm_bwd_kernel");
}
