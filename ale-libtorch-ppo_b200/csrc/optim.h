// optim.h — clip + Adam launchers (optim.cu).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace alegpu {

struct OptState {
  double lr;      // src/bin/train.cc:424-428 writes it once per rollout
  long long step; // torch Adam's per-parameter step (identical for all parameters)
  // written by the last block of grad_sumsq for the Adam kernel of the same step
  unsigned int done_count; // blocks of the running grad_sumsq that have stored their partial
  float coef;              // min(1, max_norm / (total + 1e-6))            train.cc:39-41
  float step_size;         // lr / (1 - beta1^t)
  float bc2_sqrt;          // sqrt(1 - beta2^t)
};

constexpr int kNormBlocks = 296; // 2 CTAs per SM, fixed => bit-reproducible partial order

// partial sums of g^2; the last block to finish sums them in fixed order, advances the step counter
// and leaves the clip coefficient / bias corrections in *st; total_norm_out <- the PRE-clip norm
// (over the first n_valid of the n_padded floats)
int launch_grad_sumsq(const float *grads, int64_t n_padded, int64_t n_valid, float *partials,
                      OptState *st, float max_norm, float *total_norm_out, cudaStream_t s);
// shadow (optional): operand-type copy of the parameters for the tensor-core path, bf16
// (shadow_elem_bytes 2) or tf32-rounded fp32 (4); c1w_perm: [32][256] conv1 weights of the same type
// in the space-to-depth k' order
int launch_adam_clip(float *params, float *grads, float *m, float *v, void *shadow, void *c1w_perm,
                     int shadow_elem_bytes, int64_t n_padded, const OptState *st, cudaStream_t s);
// ONE launch for norm + clip + Adam (+ the Metrics.loss broadcast when m_loss != nullptr): the gradient
// stays in registers across a grid barrier (optim.cu).  Only when adam_fused_fits(n_padded).
bool adam_fused_fits(int64_t n_padded);
int launch_adam_fused(float *params, float *grads, float *m, float *v, void *shadow, void *c1w_perm,
                      int shadow_elem_bytes, int64_t n_padded, int64_t n_valid, float *partials,
                      OptState *st, float max_norm, float *total_norm_out, const float *loss_sum,
                      const float *count, float *m_loss, int64_t mb, cudaStream_t s);
// ---- multi-GPU: the gradient exchange fused into the same kernel (no separate collective launch) ----
// Every rank maps its peers' gradient buffers, exchange buffers and signal blocks (CUDA IPC, set up
// once by the learner).  Rank r owns the r-th contiguous slice of the flat buffer.
constexpr int kArMaxRanks = 16;
struct ArSignals { // lives in every rank's exchange allocation; peers write their slot over NVLink
  unsigned long long ready[kArMaxRanks];   // [p]: rank p's local gradient of step e is complete
  unsigned long long reduced[kArMaxRanks]; // [p]: rank p's slice of step e is summed and published
  double partial[kArMaxRanks];             // [p]: sum of squares of rank p's reduced slice
};
struct ArParams {
  int world = 1, rank = 0;
  int64_t slice4 = 0;               // float4 elements per rank slice (the last slice may be shorter)
  float *grads[kArMaxRanks] = {};   // every rank's raw local gradient (flat, n_padded floats)
  float *xbuf[kArMaxRanks] = {};    // every rank's reduced slice (slice4 float4)
  ArSignals *sig[kArMaxRanks] = {}; // every rank's signal block
  int *error_flag = nullptr;        // mapped host int: 2 = a peer did not show up in time
};
int launch_adam_fused_allreduce(float *params, float *m, float *v, void *shadow, void *c1w_perm,
                                int shadow_elem_bytes, int64_t n_padded, int64_t n_valid,
                                float *partials, OptState *st, float max_norm, float *total_norm_out,
                                const float *count, float *m_loss, int64_t mb, const ArParams &ar,
                                cudaStream_t s);
int launch_f32_to_bf16(const float *src, void *dst, int64_t n, cudaStream_t s);

} // namespace alegpu
