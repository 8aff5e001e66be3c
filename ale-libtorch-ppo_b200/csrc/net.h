// net.h — Nature-CNN policy (src/bin/train.cc:230-265) data structures shared by the kernels.
#pragma once
#include "common.cuh"

namespace alegpu {

// Views into one flat fp32 buffer, tensor ORDER = reference `parameters()` order
// (c1w c1b c2w c2b c3w c3b fcw fcb piw pib vw vb); c2w/c3w/fcw are stored with the
// reduction index permuted to (ky,kx,ci) / ((y,x),c) — see net_f32.cu.
struct NetParams {
  int64_t hidden, actions;
  float *c1w, *c1b, *c2w, *c2b, *c3w, *c3b, *fcw, *fcb, *piw, *pib, *vw, *vb;
};

inline int64_t net_param_count(int64_t hidden, int64_t actions) {
  return 32 * 256 + 32 + 64 * 512 + 64 + 64 * 576 + 64 + hidden * 3136 + hidden +
         actions * hidden + actions + hidden + 1;
}

inline NetParams net_params_view(float *base, int64_t hidden, int64_t actions) {
  NetParams p;
  p.hidden = hidden;
  p.actions = actions;
  float *o = base;
  p.c1w = o; o += 32 * 256;
  p.c1b = o; o += 32;
  p.c2w = o; o += 64 * 512;
  p.c2b = o; o += 64;
  p.c3w = o; o += 64 * 576;
  p.c3b = o; o += 64;
  p.fcw = o; o += hidden * 3136;
  p.fcb = o; o += hidden;
  p.piw = o; o += actions * hidden;
  p.pib = o; o += actions;
  p.vw = o;  o += hidden;
  p.vb = o;
  return p;
}

// fp32 activation / gradient workspaces for one minibatch of up to `cap` rows (NHWC)
struct NetActs {
  int64_t cap = 0;
  float *act1 = nullptr, *act2 = nullptr, *act3 = nullptr, *fc = nullptr;
  float *dact1 = nullptr, *dact2 = nullptr, *dact3 = nullptr, *dfc = nullptr;
  float *logits = nullptr, *values = nullptr; // [cap,A], [cap]
  float *dout = nullptr;                      // [cap, A+1] = (dlogits, dvalue)
  void *dfc16 = nullptr;                      // optional bf16 copy of dfc (tensor-core path)
  bool dfc_round_tf32 = false;                // tf32 path: dfc is stored rounded to tf32 (cvt.rna)
  float *fc_ws = nullptr;                     // split-K partials of the FC forward (tensor-core path)
  size_t fc_ws_floats = 0;
};

int net_f32_forward(const NetParams &P, const uint8_t *obs, int64_t n, NetActs &A,
                    cudaStream_t s);
int net_f32_backward(const NetParams &P, const NetParams &G, const uint8_t *obs, int64_t n,
                     NetActs &A, float *ws, size_t ws_floats, cudaStream_t s);

// heads.cu --------------------------------------------------------------------------------------
struct LossArgs {
  const int64_t *actions;
  const float *old_log_probs, *advantages, *returns;
  const uint8_t *masks;
  float clip, vcoef, ecoef;
  const float *count; // device scalar: (global) number of unmasked rows in the minibatch
  // optional advantage normalisation: (count, sum, sum of squares) of the minibatch's unmasked
  // advantages in double; adv' = (adv - mean) / (std + 1e-8), unbiased std.  nullptr = off.
  const double *adv_stats;
  // optional clipped value loss (loss.cuh): the values stored at collection time; nullptr = off
  const float *old_values;
  // the LAST block of the heads kernel sums the per-block loss partials in index order into
  // *loss_sum (un-normalised: divide by *count); done_counter is a zeroed device counter it resets
  float *loss_sum;
  unsigned int *done_counter;
  int *error_flag; // optional: set to 1 when an action index lies outside [0, A)
  // per-sample metric outputs at their final offsets (any may be null) — Metrics::set, train.h:93-108
  float *m_clipped, *m_value, *m_entropy, *m_total, *m_ratio, *m_adv, *m_ret;
  uint8_t *m_masks;
};
// logits/value heads only (inference; src/bin/train.cc:262-264)
int heads_forward(const NetParams &P, const float *fc, int64_t n, float *logits, float *values,
                  cudaStream_t s);
// softmax + categorical draw per row (src/bin/train.cc:373-375) and strided writes of
// logits/values/actions (e.g. straight into column t of the [E,T,...] rollout storage)
int launch_act_sample(const float *logits, const float *values, int64_t n, int A,
                      const float *uniforms, uint64_t seed, uint64_t step, float *logits_out,
                      int64_t logits_stride, float *values_out, int64_t values_stride,
                      int64_t *actions_out, int64_t actions_stride, int64_t *actions_dense,
                      cudaStream_t s);
// heads forward + PPO loss + d loss/d(logits,value) + dfc, fused; loss_partials[grid] out
int heads_loss_backward(const NetParams &P, NetActs &A, int64_t n, const LossArgs &L,
                        float *loss_partials, int *nparts, cudaStream_t s);
// gradients of piw/pib/vw/vb from A.dout and A.fc (split over rows, fixed-order reduce)
int heads_wgrad(const NetParams &G, const NetActs &A, int64_t n, float *ws, size_t ws_floats,
                cudaStream_t s);
// Metrics::set tail: loss (broadcast over mb) = *loss_sum / *count; loss_out[0] if non-null
int finalize_metrics(const float *loss_sum, const float *count, float *m_loss, int64_t mb,
                     float *loss_out, cudaStream_t s);

// loss.cu
int launch_mask_count(const uint8_t *masks, int64_t n, int64_t groups, float *count,
                      cudaStream_t s);
int launch_normalize_logits(const float *logits, int64_t n, int64_t A, float *out, cudaStream_t s);

} // namespace alegpu
