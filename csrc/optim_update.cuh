// Per-element OptimMethod updates as device functions, shared by the stand-alone flat update kernel (optim.cu) and
// the fused NVLink reduce-scatter + update kernel (comm_fused.cu).
//
// Formulas follow the reference exactly: S/optim/SGD.scala:61-125, Adam.scala:54-108, Adagrad.scala:47-79,
// RMSprop.scala:50-75, Adadelta.scala:48-77, Adamax.scala:50-85, Ftrl.scala:80-170, LarsSGD.scala:73-100.
// Host code resolves everything that is scalar per step (clr from the schedule, Adam's bias-corrected step size,
// LARS' per-layer rate) and passes it in OptimHP; the kernel only does the per-element math.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace bdl {

enum OptimKind : int {
  OPT_SGD = 0,
  OPT_ADAM = 1,
  OPT_ADAGRAD = 2,
  OPT_RMSPROP = 3,
  OPT_ADADELTA = 4,
  OPT_ADAMAX = 5,
  OPT_FTRL = 6,
  OPT_LARS = 7,
};

// Plain-old-data so it can be passed by value or read from device memory (CUDA-graph friendly).
struct OptimHP {
  float lr;         // positive step size for this step (already schedule-resolved; Adam: bias-corrected step)
  float wd;         // weight decay
  float momentum;   // SGD/LARS momentum, Adam beta1, Adamax beta1, RMSprop/Adadelta decay
  float dampening;  // SGD dampening, Adam beta2, Adamax beta2
  float eps;
  float gscale;     // multiplies the incoming gradient (1/numReplicas, loss-scale^-1, clip factor)
  float clip_min;   // constant clipping range, applied when clip_min < clip_max
  float clip_max;
  int nesterov;
  int first_step;   // SGD: momentum buffer is initialised to g on the first step
  // Ftrl
  float lr_power, l1, l2, l2_shrink;
};

template <int kKind>
__device__ __forceinline__ float optim_apply(float w, float g, float* s0, float* s1, const OptimHP& hp) {
  g *= hp.gscale;
  if (hp.clip_min < hp.clip_max) g = fminf(fmaxf(g, hp.clip_min), hp.clip_max);
  if (kKind == OPT_SGD) {
    if (hp.wd != 0.f) g += hp.wd * w;
    if (hp.momentum != 0.f) {
      float v = hp.first_step ? g : hp.momentum * (*s0) + (1.f - hp.dampening) * g;
      *s0 = v;
      g = hp.nesterov ? g + hp.momentum * v : v;
    }
    return w - hp.lr * g;
  } else if (kKind == OPT_ADAM) {
    const float m = hp.momentum * (*s0) + (1.f - hp.momentum) * g;
    const float v = hp.dampening * (*s1) + (1.f - hp.dampening) * g * g;
    *s0 = m;
    *s1 = v;
    return w - hp.lr * m / (sqrtf(v) + hp.eps);
  } else if (kKind == OPT_ADAGRAD) {
    if (hp.wd != 0.f) g += hp.wd * w;
    const float G = *s0 + g * g;
    *s0 = G;
    return w - hp.lr * g / (sqrtf(G) + 1e-10f);
  } else if (kKind == OPT_RMSPROP) {
    const float ss = hp.momentum * (*s0) + (1.f - hp.momentum) * g * g;
    *s0 = ss;
    return w - hp.lr * g / (sqrtf(ss) + hp.eps);
  } else if (kKind == OPT_ADADELTA) {
    const float var = hp.momentum * (*s0) + (1.f - hp.momentum) * g * g;
    const float delta = sqrtf(*s1 + hp.eps) / sqrtf(var + hp.eps) * g;
    *s0 = var;
    *s1 = hp.momentum * (*s1) + (1.f - hp.momentum) * delta * delta;
    return w - delta;
  } else if (kKind == OPT_ADAMAX) {
    const float m = hp.momentum * (*s0) + (1.f - hp.momentum) * g;
    const float u = fmaxf(hp.dampening * (*s1), fabsf(g) + hp.eps);
    *s0 = m;
    *s1 = u;
    return w - hp.lr * m / u;
  } else if (kKind == OPT_FTRL) {
    // s0 = accum, s1 = linear
    const float gs = g + 2.f * hp.l2_shrink * w;
    const float accum = *s0;
    const float accum_new = accum + gs * gs;
    const float p_old = __powf(accum, -hp.lr_power);
    const float p_new = __powf(accum_new, -hp.lr_power);
    const float linear = *s1 + gs + (p_old - p_new) / hp.lr * w;
    const float quadratic = p_new / hp.lr + 2.f * hp.l2;
    *s0 = accum_new;
    *s1 = linear;
    if (fabsf(linear) > hp.l1) {
      const float sgn = linear > 0.f ? hp.l1 : (linear < 0.f ? -hp.l1 : 0.f);
      return (sgn - linear) / quadratic;
    }
    return 0.f;
  } else {  // OPT_LARS: lr already holds globalLr * trust / scale for this layer
    const float v = hp.momentum * (*s0) + hp.lr * (g + hp.wd * w);
    *s0 = v;
    return w - v;
  }
}

__device__ __forceinline__ float optim_dispatch(int kind, float w, float g, float* s0, float* s1, const OptimHP& hp) {
  switch (kind) {
    case OPT_SGD: return optim_apply<OPT_SGD>(w, g, s0, s1, hp);
    case OPT_ADAM: return optim_apply<OPT_ADAM>(w, g, s0, s1, hp);
    case OPT_ADAGRAD: return optim_apply<OPT_ADAGRAD>(w, g, s0, s1, hp);
    case OPT_RMSPROP: return optim_apply<OPT_RMSPROP>(w, g, s0, s1, hp);
    case OPT_ADADELTA: return optim_apply<OPT_ADADELTA>(w, g, s0, s1, hp);
    case OPT_ADAMAX: return optim_apply<OPT_ADAMAX>(w, g, s0, s1, hp);
    case OPT_FTRL: return optim_apply<OPT_FTRL>(w, g, s0, s1, hp);
    default: return optim_apply<OPT_LARS>(w, g, s0, s1, hp);
  }
}

__device__ __forceinline__ uint16_t f32_to_bf16_bits(float v, int truncate) {
  const uint32_t u = __float_as_uint(v);
  if (truncate) return static_cast<uint16_t>(u >> 16);
  // round-to-nearest-even, NaN preserved
  if ((u & 0x7FFFFFFFu) > 0x7F800000u) return static_cast<uint16_t>((u >> 16) | 0x40u);
  return static_cast<uint16_t>((u + 0x7FFFu + ((u >> 16) & 1u)) >> 16);
}

}  // namespace bdl
