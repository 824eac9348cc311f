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

// ---- GRU (S/nn/GRU.scala:72-160): two pointwise stages around the two recurrent GEMMs of a step.
// stage 1: r, z = sigmoid(x_rz + h.U_rz); hr = h * r          (hr feeds the candidate GEMM)
// stage 2: hh = tanh(x_c + hr.U_c);  h' = (1 - z) * hh + z * h
__global__ void gru_rz_fwd_kernel(const __nv_bfloat16* __restrict__ xrz, const __nv_bfloat16* __restrict__ hg,
                                  const __nv_bfloat16* __restrict__ h, __nv_bfloat16* __restrict__ rz,
                                  __nv_bfloat16* __restrict__ hr, long long B, int H, long long ldx, long long ldh) {
  const long long total = B * H;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const long long b = i / H;
    const int j = static_cast<int>(i - b * H);
    const float r = sigm(__bfloat162float(xrz[b * ldx + j]) + __bfloat162float(hg[b * 2 * H + j]));
    const float z = sigm(__bfloat162float(xrz[b * ldx + H + j]) + __bfloat162float(hg[b * 2 * H + H + j]));
    rz[b * 2 * H + j] = __float2bfloat16_rn(r);
    rz[b * 2 * H + H + j] = __float2bfloat16_rn(z);
    hr[i] = __float2bfloat16_rn(__bfloat162float(h[b * ldh + j]) * r);
  }
}

// d_hr: grad of hr; d_z: grad of z (from stage 2). Outputs d_pre [B,2H] (grad of both x_rz and h.U_rz), d_h [B,H].
__global__ void gru_rz_bwd_kernel(const __nv_bfloat16* __restrict__ d_hr, const __nv_bfloat16* __restrict__ d_z,
                                  const __nv_bfloat16* __restrict__ rz, const __nv_bfloat16* __restrict__ h,
                                  __nv_bfloat16* __restrict__ d_pre, __nv_bfloat16* __restrict__ d_h, long long B,
                                  int H, long long ldh) {
  const long long total = B * H;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const long long b = i / H;
    const int j = static_cast<int>(i - b * H);
    const float r = __bfloat162float(rz[b * 2 * H + j]), z = __bfloat162float(rz[b * 2 * H + H + j]);
    const float ghr = d_hr ? __bfloat162float(d_hr[i]) : 0.f;
    const float gz = d_z ? __bfloat162float(d_z[i]) : 0.f;
    const float hv = __bfloat162float(h[b * ldh + j]);
    d_pre[b * 2 * H + j] = __float2bfloat16_rn(ghr * hv * r * (1.f - r));
    d_pre[b * 2 * H + H + j] = __float2bfloat16_rn(gz * z * (1.f - z));
    d_h[i] = __float2bfloat16_rn(ghr * r);
  }
}

__global__ void gru_out_fwd_kernel(const __nv_bfloat16* __restrict__ xc, const __nv_bfloat16* __restrict__ hc,
                                   const __nv_bfloat16* __restrict__ z, const __nv_bfloat16* __restrict__ h,
                                   __nv_bfloat16* __restrict__ hh, __nv_bfloat16* __restrict__ out, long long B, int H,
                                   long long ldx, long long ldz, long long ldh) {
  const long long total = B * H;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const long long b = i / H;
    const int j = static_cast<int>(i - b * H);
    const float c = tanhf(__bfloat162float(xc[b * ldx + j]) + __bfloat162float(hc[i]));
    const float zv = __bfloat162float(z[b * ldz + j]);
    hh[i] = __float2bfloat16_rn(c);
    out[i] = __float2bfloat16_rn((1.f - zv) * c + zv * __bfloat162float(h[b * ldh + j]));
  }
}

// outputs: d_pre_c [B,H] (grad of x_c and of hr.U_c), d_z [B,H], d_h [B,H] (direct path through z * h)
__global__ void gru_out_bwd_kernel(const __nv_bfloat16* __restrict__ d_out, const __nv_bfloat16* __restrict__ hh,
                                   const __nv_bfloat16* __restrict__ z, const __nv_bfloat16* __restrict__ h,
                                   __nv_bfloat16* __restrict__ d_pre_c, __nv_bfloat16* __restrict__ d_z,
                                   __nv_bfloat16* __restrict__ d_h, long long B, int H, long long ldz, long long ldh) {
  const long long total = B * H;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const long long b = i / H;
    const int j = static_cast<int>(i - b * H);
    const float g = __bfloat162float(d_out[i]);
    const float c = __bfloat162float(hh[i]), zv = __bfloat162float(z[b * ldz + j]);
    const float hv = __bfloat162float(h[b * ldh + j]);
    d_pre_c[i] = __float2bfloat16_rn(g * (1.f - zv) * (1.f - c * c));
    d_z[i] = __float2bfloat16_rn(g * (hv - c));
    d_h[i] = __float2bfloat16_rn(g * zv);
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

extern "C" int bdl_gru_rz_fwd(const void* xrz, const void* hg, const void* h, void* rz, void* hr, long long B, int H,
                              long long ldx, long long ldh, void* stream) {
  if (B <= 0 || H <= 0) return BDL_OK;
  gru_rz_fwd_kernel<<<rgrid(B * H), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(xrz), static_cast<const __nv_bfloat16*>(hg),
      static_cast<const __nv_bfloat16*>(h), static_cast<__nv_bfloat16*>(rz), static_cast<__nv_bfloat16*>(hr), B, H, ldx,
      ldh);
  return bdl_check_launch("gru_rz_fwd_kernel");
}

extern "C" int bdl_gru_rz_bwd(const void* d_hr, const void* d_z, const void* rz, const void* h, void* d_pre, void* d_h,
                              long long B, int H, long long ldh, void* stream) {
  if (B <= 0 || H <= 0) return BDL_OK;
  gru_rz_bwd_kernel<<<rgrid(B * H), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(d_hr), static_cast<const __nv_bfloat16*>(d_z),
      static_cast<const __nv_bfloat16*>(rz), static_cast<const __nv_bfloat16*>(h), static_cast<__nv_bfloat16*>(d_pre),
      static_cast<__nv_bfloat16*>(d_h), B, H, ldh);
  return bdl_check_launch("gru_rz_bwd_kernel");
}

extern "C" int bdl_gru_out_fwd(const void* xc, const void* hc, const void* z, const void* h, void* hh, void* out,
                               long long B, int H, long long ldx, long long ldz, long long ldh, void* stream) {
  if (B <= 0 || H <= 0) return BDL_OK;
  gru_out_fwd_kernel<<<rgrid(B * H), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(xc), static_cast<const __nv_bfloat16*>(hc),
      static_cast<const __nv_bfloat16*>(z), static_cast<const __nv_bfloat16*>(h), static_cast<__nv_bfloat16*>(hh),
      static_cast<__nv_bfloat16*>(out), B, H, ldx, ldz, ldh);
  return bdl_check_launch("gru_out_fwd_kernel");
}

extern "C" int bdl_gru_out_bwd(const void* d_out, const void* hh, const void* z, const void* h, void* d_pre_c,
                               void* d_z, void* d_h, long long B, int H, long long ldz, long long ldh, void* stream) {
  if (B <= 0 || H <= 0) return BDL_OK;
  gru_out_bwd_kernel<<<rgrid(B * H), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(d_out), static_cast<const __nv_bfloat16*>(hh),
      static_cast<const __nv_bfloat16*>(z), static_cast<const __nv_bfloat16*>(h), static_cast<__nv_bfloat16*>(d_pre_c),
      static_cast<__nv_bfloat16*>(d_z), static_cast<__nv_bfloat16*>(d_h), B, H, ldz, ldh);
  return bdl_check_launch("gru_out_bwd_kernel");
}
