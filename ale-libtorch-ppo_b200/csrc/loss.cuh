// loss.cuh — per-sample PPO loss forward + analytic backward (device function).
//
// Follows ai::ppo::losses::{normalize_logits, compute, clipped_surrogate_objectives,
// compute_entropies} (src/ai/ppo/losses.cc:4-47) and the gradient autograd derives for
// mini_batch_update (src/ai/ppo/train.h:118-126):
//   logp = z - logsumexp(z)                                              losses.cc:45-47
//   rho = exp(logp[a] - old[a]); surr = min(rho*A, clamp(rho,1-e,1+e)*A) losses.cc:33-38
//   vl = 0.5 (v-R)^2 ; H = -sum exp(logp) logp                           losses.cc:15, 41-43
//   total = -surr + cv*vl - ce*H ; loss = sum(mask*total)/sum(mask)      losses.cc:17-19
// Backward rule for min/clamp (torch semantics): clamp passes gradient on the CLOSED range
// and min splits ties 1/2+1/2, so d surr/d rho = A when 1-e <= rho <= 1+e or the unclipped
// branch is the strict minimum, else 0.
#pragma once
#include "common.cuh"

namespace alegpu {

struct PpoSample {
  float surr, value_loss, entropy, total, ratio;
};

// z: raw logits (A values, any address space), old_lp: normalised old log-probs.
// scale = mask ? 1/count : 0.  dz (A values) and *dv receive d loss / d(z, v).
template <int kMaxA>
__device__ __forceinline__ PpoSample ppo_sample(const float *z, const float *old_lp, int A,
                                                int action, float adv, float v, float ret,
                                                float clip, float vcoef, float ecoef,
                                                float scale, float *dz, float *dv) {
  float mx = z[0];
  for (int j = 1; j < A; ++j)
    mx = fmaxf(mx, z[j]);
  float se = 0.0f;
  for (int j = 0; j < A; ++j)
    se += expf(z[j] - mx);
  const float lse = mx + logf(se);
  float lp[kMaxA], p[kMaxA];
  float ent = 0.0f;
#pragma unroll
  for (int j = 0; j < kMaxA; ++j) {
    if (j < A) {
      lp[j] = z[j] - lse;
      p[j] = expf(lp[j]);
      ent += p[j] * lp[j];
    }
  }
  ent = -ent;
  float lpa = 0.0f, olda = 0.0f;
#pragma unroll
  for (int j = 0; j < kMaxA; ++j)
    if (j == action) {
      lpa = lp[j];
      olda = old_lp[j];
    }
  PpoSample o;
  const float rho = expf(lpa - olda);
  const float lo = 1.0f - clip, hi = 1.0f + clip;
  const float crho = fminf(fmaxf(rho, lo), hi);
  const float un = rho * adv, cl = crho * adv;
  o.surr = fminf(un, cl);
  const float d = v - ret;
  o.value_loss = 0.5f * (d * d);
  o.entropy = ent;
  o.total = -o.surr + vcoef * o.value_loss - ecoef * ent;
  o.ratio = rho;
  const bool inside = rho >= lo && rho <= hi;
  const float g_rho = (inside || un < cl) ? adv : 0.0f;
  if (dz) {
    float g[kMaxA];
    float gsum = 0.0f;
#pragma unroll
    for (int j = 0; j < kMaxA; ++j)
      if (j < A) {
        g[j] = ecoef * p[j] * (lp[j] + 1.0f); // d(-ce*H)/dlogp_j
        if (j == action)
          g[j] += -g_rho * rho;               // d(-surr)/dlogp_a
        gsum += g[j];
      }
#pragma unroll
    for (int j = 0; j < kMaxA; ++j)
      if (j < A)
        dz[j] = (g[j] - p[j] * gsum) * scale; // log-softmax backward
  }
  if (dv)
    *dv = vcoef * d * scale;
  return o;
}

} // namespace alegpu
