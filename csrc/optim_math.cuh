// optim_math.cuh — the step scalars (gradient scale, clip coefficient) and the Adam element update shared by the
// flat-arena optimizer kernel (optim_kernels.cu) and the sharded update of the peer-to-peer path (p2p_allreduce.cu).
// Reference math: hetseq/optim.py:206-229 (Adam), hetseq/controller.py:338-347 (scale, clip).
#pragma once
#include <cmath>
#include "vec_utils.cuh"
#include "../include/b200seq.h"

namespace b200 {

// ---- derived step scalars -------------------------------------------------------------------------
struct StepScalars {
  float scale;   // multiply every gradient by this (grad scale x clip coefficient)
  float gnorm;   // norm of the scaled, unclipped gradient (what the reference logs as gnorm)
  float clipped; // 1 if the clip coefficient was applied
};

__device__ __forceinline__ StepScalars derive_scalars(float mult, const double* sample_size, const float* sample_size_f32,
                                                      const float* norm_sq, float max_norm) {
  StepScalars s;
  float c = mult;
  if (sample_size != nullptr) {
    const double ss = *sample_size;
    c = ss > 0.0 ? (float)((double)mult / ss) : 0.f;
  } else if (sample_size_f32 != nullptr) {
    const float ss = *sample_size_f32;
    c = ss > 0.f ? mult / ss : 0.f;
  }
  s.scale = c;
  s.gnorm = 0.f;
  s.clipped = 0.f;
  if (norm_sq != nullptr) {
    s.gnorm = sqrtf(fmaxf(*norm_sq, 0.f)) * fabsf(c);
    if (max_norm > 0.f && s.gnorm > max_norm) {
      // torch.nn.utils.clip_grad_norm_: coef = max_norm / (norm + 1e-6), applied when norm > max_norm
      s.scale = c * (max_norm / (s.gnorm + 1e-6f));
      s.clipped = 1.f;
    }
  }
  return s;
}

struct AdamArgsDev {
  float lr, beta1, beta2, eps, weight_decay;
  float step_size;  // lr * sqrt(1-beta2^t) / (1-beta1^t), computed on the host in double
  float mult, max_norm;
  const float* norm_sq;
  const double* sample_size;
  float* stats_out;
  int zero_grad;
  const float* sample_size_f32;
  const float* hyper_dev;   // optional device float[4]: lr, step_size, (1 - wd*lr), unused — overrides the by-value copies
};

// one element of the reference Adam rule (decoupled weight decay applied to the weight before the moment step)
__device__ __forceinline__ void adam_element(float& w, float g, float& m, float& v, float scale, float beta1, float beta2,
                                             float one_m_b1, float one_m_b2, float eps, float step_size, bool decay_on,
                                             float decay) {
  const float gs = g * scale;
  m = beta1 * m + one_m_b1 * gs;
  v = beta2 * v + one_m_b2 * gs * gs;
  const float denom = sqrtf(v) + eps;
  if (decay_on) w *= decay;
  w = w - step_size * (m / denom);
}

inline void fill_adam_args(const b200_adam_args* a, AdamArgsDev& d) {
  d.lr = a->lr; d.beta1 = a->beta1; d.beta2 = a->beta2; d.eps = a->eps; d.weight_decay = a->weight_decay;
  const double bc1 = 1.0 - pow((double)a->beta1, (double)a->step);
  const double bc2 = 1.0 - pow((double)a->beta2, (double)a->step);
  d.step_size = (float)((double)a->lr * sqrt(bc2) / bc1);
  d.mult = a->grad_scale_mult; d.max_norm = a->max_norm;
  d.norm_sq = a->norm_sq; d.sample_size = a->sample_size; d.stats_out = a->stats_out; d.zero_grad = a->zero_grad;
  d.sample_size_f32 = a->sample_size_f32; d.hyper_dev = a->hyper_dev;
}

}  // namespace b200
