// rollout.cu — the rollout buffer behind the C ABI: mirror of ai::buffer::Buffer
// (src/ai/buffer.h:27-58, src/ai/buffer.cc:5-77) plus prepare_batch (src/bin/train.cc:272-283).
//
// Storage is the reference's: env-major [E,T,...] device arrays (obs u8 [E,T,4,84,84], actions
// i64, f32 scalars, 1-byte flags, logits [E,T,A]).  `add` is ONE scatter kernel + one strided
// 2-D copy instead of 8 strided copy_ kernels (buffer.cc:47-54); producers that can write in
// place (alegpu_preprocess_step -> slot t+1) get the raw pointers through alegpu_rollout_view.
// `get` = fused clamp + GAE + returns + masks (gae.cu) + log-softmax of the stored logits.
#include <algorithm>
#include <cstdlib>
#include <vector>

#include "net.h"

using namespace alegpu;

struct alegpu_rollout {
  int64_t E = 0, T = 0, A = 0;
  int device = 0;
  int64_t write_index = 0; // Buffer::indices_
  uint8_t *obs = nullptr, *carry = nullptr;
  int64_t *actions = nullptr;
  float *rewards = nullptr, *values = nullptr, *advantages = nullptr, *returns = nullptr;
  float *logits = nullptr, *log_probs = nullptr;
  uint8_t *terminals = nullptr, *truncations = nullptr, *starts = nullptr, *masks = nullptr;
  int *error_flag = nullptr;
  // hand-off state (alegpu_rollout_ingest)
  // device ring of per-step frame slabs: as many slots as steps when that fits kRingBytes (then a
  // whole rollout's copies can run ahead of the compute stream, e.g. under the previous train()),
  // never fewer than kMinRing
  static constexpr int kMinRing = 4;
  static constexpr size_t kRingBytes = (size_t)6 << 30;
  int ring_slots = 0;
  cudaStream_t copy_stream = nullptr;
  uint8_t *ring = nullptr;
  std::vector<cudaEvent_t> copied, consumed;
  std::vector<char> used; // slot has been read by a kernel before (consumed[slot] is meaningful)
  cudaEvent_t start = nullptr;
};

namespace {

__global__ void rollout_add_kernel(int64_t E, int64_t T, int64_t A, int64_t t,
                                   const int64_t *__restrict__ actions,
                                   const float *__restrict__ rewards,
                                   const uint8_t *__restrict__ terminals,
                                   const uint8_t *__restrict__ truncations,
                                   const uint8_t *__restrict__ starts,
                                   const float *__restrict__ logits,
                                   const float *__restrict__ values, int64_t *__restrict__ o_actions,
                                   float *__restrict__ o_rewards, uint8_t *__restrict__ o_terminals,
                                   uint8_t *__restrict__ o_truncations, uint8_t *__restrict__ o_starts,
                                   float *__restrict__ o_logits, float *__restrict__ o_values) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E)
    return;
  const int64_t o = e * T + t; // buffer.cc:47-54: column t of every [E,T,...] array
  o_actions[o] = actions[e];
  o_rewards[o] = rewards[e];
  o_terminals[o] = terminals[e] ? 1 : 0;
  o_truncations[o] = truncations[e] ? 1 : 0;
  o_starts[o] = starts[e] ? 1 : 0;
  o_values[o] = values[e];
  for (int64_t a = 0; a < A; ++a)
    o_logits[o * A + a] = logits[e * A + a];
}

template <typename T> int dmalloc(T **p, size_t n) {
  ALEGPU_CUDA_CHECK(cudaMalloc(p, sizeof(T) * (n ? n : 1)));
  ALEGPU_CUDA_CHECK(cudaMemset(*p, 0, sizeof(T) * (n ? n : 1)));
  return ALEGPU_OK;
}

} // namespace

extern "C" int alegpu_rollout_create(int64_t E, int64_t T, int64_t A, int device,
                                     alegpu_rollout **out) {
  ALEGPU_REQUIRE(out && E >= 1 && T >= 1 && A >= 1, ALEGPU_INVALID_ARGUMENT,
                 "rollout_create: bad arguments");
  ALEGPU_CUDA_CHECK(cudaSetDevice(device));
  auto *r = new alegpu_rollout();
  r->E = E; r->T = T; r->A = A; r->device = device;
  const size_t n = (size_t)E * T;
  int rc = 0;
  if ((rc = dmalloc(&r->obs, n * 28224)) || (rc = dmalloc(&r->carry, (size_t)2 * E * 28224)) ||
      (rc = dmalloc(&r->actions, n)) || (rc = dmalloc(&r->rewards, n)) ||
      (rc = dmalloc(&r->values, n)) || (rc = dmalloc(&r->advantages, n)) ||
      (rc = dmalloc(&r->returns, n)) || (rc = dmalloc(&r->logits, n * A)) ||
      (rc = dmalloc(&r->log_probs, n * A)) || (rc = dmalloc(&r->terminals, n)) ||
      (rc = dmalloc(&r->truncations, n)) || (rc = dmalloc(&r->starts, n)) ||
      (rc = dmalloc(&r->masks, n)) || (rc = dmalloc(&r->error_flag, 1))) {
    alegpu_rollout_destroy(r);
    return rc;
  }
  *out = r;
  return ALEGPU_OK;
}

extern "C" int alegpu_rollout_destroy(alegpu_rollout *r) {
  if (!r)
    return ALEGPU_OK;
  cudaSetDevice(r->device);
  for (void *p : {(void *)r->obs, (void *)r->carry, (void *)r->actions, (void *)r->rewards,
                  (void *)r->values, (void *)r->advantages, (void *)r->returns,
                  (void *)r->logits, (void *)r->log_probs, (void *)r->terminals,
                  (void *)r->truncations, (void *)r->starts, (void *)r->masks,
                  (void *)r->error_flag})
    if (p)
      cudaFree(p);
  if (r->copy_stream) {
    cudaStreamSynchronize(r->copy_stream);
    cudaStreamDestroy(r->copy_stream);
    cudaFree(r->ring);
    for (int i = 0; i < r->ring_slots; ++i) {
      cudaEventDestroy(r->copied[i]);
      cudaEventDestroy(r->consumed[i]);
    }
    cudaEventDestroy(r->start);
  }
  delete r;
  return ALEGPU_OK;
}

extern "C" int alegpu_rollout_get_view(alegpu_rollout *r, alegpu_rollout_view *v) {
  ALEGPU_REQUIRE(r && v, ALEGPU_INVALID_ARGUMENT, "rollout_get_view: null argument");
  v->E = r->E; v->T = r->T; v->A = r->A;
  v->observations = r->obs; v->carry = r->carry; v->actions = r->actions;
  v->rewards = r->rewards; v->values = r->values; v->advantages = r->advantages;
  v->returns = r->returns; v->logits = r->logits; v->log_probabilities = r->log_probs;
  v->terminals = r->terminals; v->truncations = r->truncations;
  v->episode_starts = r->starts; v->masks = r->masks;
  return ALEGPU_OK;
}

extern "C" int alegpu_rollout_add(alegpu_rollout *r, const uint8_t *observations,
                                  const int64_t *actions, const float *rewards,
                                  const uint8_t *terminals, const uint8_t *truncations,
                                  const uint8_t *episode_starts, const float *logits,
                                  const float *values, void *stream) {
  ALEGPU_REQUIRE(r && actions && rewards && terminals && truncations && episode_starts && logits &&
                     values,
                 ALEGPU_INVALID_ARGUMENT, "rollout_add: null argument");
  ALEGPU_CUDA_CHECK(cudaSetDevice(r->device));
  cudaStream_t s = as_stream(stream);
  const int64_t t = r->write_index;
  if (observations) // null = the producer already wrote slot t in place
    ALEGPU_CUDA_CHECK(cudaMemcpy2DAsync(r->obs + t * 28224, (size_t)r->T * 28224, observations,
                                        28224, 28224, (size_t)r->E, cudaMemcpyDeviceToDevice, s));
  rollout_add_kernel<<<(unsigned)ceil_div(r->E, 128), 128, 0, s>>>(
      r->E, r->T, r->A, t, actions, rewards, terminals, truncations, episode_starts, logits,
      values, r->actions, r->rewards, r->terminals, r->truncations, r->starts, r->logits,
      r->values);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  r->write_index = (r->write_index + 1) % r->T; // buffer.cc:55
  return ALEGPU_OK;
}

// ---- the hand-off: pinned host ring -> side-stream cudaMemcpyAsync -> fused preprocess --------
// Replaces screen_buffers_ + torch::stack + per-env scalar device writes
// (src/ai/rollout.cc:83-93, 184-196, 216-218).  frames_host is PINNED memory laid out
// [T,E,2,210,160,3] (what the emulator threads fill, one [E,...] slab per env step).  Step t's
// slab is copied on the copy stream into one of kRing device slots while the compute stream
// is still preprocessing step t-1; events chain copy -> preprocess -> slot reuse.
// carry_index selects which of the two carry stacks holds the stack before step 0; the stack
// after step T-1 lands in the other one.
extern "C" int alegpu_rollout_ingest(alegpu_rollout *r, const uint8_t *frames_host,
                                     int carry_index, void *stream) {
  ALEGPU_REQUIRE(r && frames_host && (carry_index == 0 || carry_index == 1),
                 ALEGPU_INVALID_ARGUMENT, "rollout_ingest: bad argument");
  ALEGPU_CUDA_CHECK(cudaSetDevice(r->device));
  cudaStream_t s = as_stream(stream);
  const int64_t E = r->E, T = r->T;
  const size_t slab = (size_t)E * 2 * 100800;
  if (!r->copy_stream) {
    const int64_t fit = (int64_t)(alegpu_rollout::kRingBytes / slab);
    r->ring_slots = (int)std::max<int64_t>(alegpu_rollout::kMinRing, std::min<int64_t>(T, fit));
    if (const char *e = getenv("ALEGPU_RING_SLOTS")) // test knob: force the per-step ring branch
      if (atoi(e) >= 2)
        r->ring_slots = atoi(e);
    ALEGPU_CUDA_CHECK(cudaMalloc(&r->ring, slab * (size_t)r->ring_slots));
    ALEGPU_CUDA_CHECK(cudaStreamCreateWithFlags(&r->copy_stream, cudaStreamNonBlocking));
    r->copied.assign((size_t)r->ring_slots, nullptr);
    r->consumed.assign((size_t)r->ring_slots, nullptr);
    r->used.assign((size_t)r->ring_slots, 0);
    for (int i = 0; i < r->ring_slots; ++i) {
      ALEGPU_CUDA_CHECK(cudaEventCreateWithFlags(&r->copied[i], cudaEventDisableTiming));
      ALEGPU_CUDA_CHECK(cudaEventCreateWithFlags(&r->consumed[i], cudaEventDisableTiming));
    }
    ALEGPU_CUDA_CHECK(cudaEventCreateWithFlags(&r->start, cudaEventDisableTiming));
  }
  uint8_t *carry_in = r->carry + (size_t)carry_index * E * 28224;
  uint8_t *carry_out = r->carry + (size_t)(1 - carry_index) * E * 28224;
  const int slots = r->ring_slots;
  // The copies depend only on their ring slot being free (the kernel that last read it has
  // finished), NOT on what the compute stream is doing now: the frames of the next rollout stream
  // in under the current train().
  auto copy_slab = [&](int64_t t) -> int {
    const int slot = (int)(t % slots);
    if (r->used[slot])
      ALEGPU_CUDA_CHECK(cudaStreamWaitEvent(r->copy_stream, r->consumed[slot], 0));
    ALEGPU_CUDA_CHECK(cudaMemcpyAsync(r->ring + slab * slot, frames_host + slab * t, slab,
                                      cudaMemcpyHostToDevice, r->copy_stream));
    ALEGPU_CUDA_CHECK(cudaEventRecord(r->copied[slot], r->copy_stream));
    return ALEGPU_OK;
  };
  if (slots >= T) {
    // the ring holds the whole rollout as [T,E,2,210,160,3]: one fused kernel over all steps
    // (the T slabs are contiguous on both sides and share one reader, so they move as ONE copy)
    if (r->used[0])
      ALEGPU_CUDA_CHECK(cudaStreamWaitEvent(r->copy_stream, r->consumed[0], 0));
    ALEGPU_CUDA_CHECK(cudaMemcpyAsync(r->ring, frames_host, slab * (size_t)T, cudaMemcpyHostToDevice,
                                      r->copy_stream));
    ALEGPU_CUDA_CHECK(cudaEventRecord(r->copied[(size_t)(T - 1)], r->copy_stream));
    ALEGPU_CUDA_CHECK(cudaStreamWaitEvent(s, r->copied[(size_t)(T - 1)], 0));
    if (int rc = alegpu_preprocess_rollout(r->ring, r->starts, carry_in, carry_out, r->obs, E, T,
                                           stream))
      return rc;
    ALEGPU_CUDA_CHECK(cudaEventRecord(r->consumed[0], s)); // every slot was read by that kernel
    r->used[0] = 1;
    return ALEGPU_OK;
  }
  // slot 0 <- the carried stack (Rollout::observations_ persists across rollouts)
  ALEGPU_CUDA_CHECK(cudaMemcpy2DAsync(r->obs, (size_t)T * 28224, carry_in, 28224, 28224,
                                      (size_t)E, cudaMemcpyDeviceToDevice, s));
  for (int64_t t = 0; t < T; ++t) {
    const int slot = (int)(t % slots);
    uint8_t *dev = r->ring + slab * slot;
    if (int rc = copy_slab(t))
      return rc;
    ALEGPU_CUDA_CHECK(cudaStreamWaitEvent(s, r->copied[slot], 0));
    const bool last = t + 1 == T;
    // ring slab layout is [E,2,210,160,3]: frame a of env e at e*201600, frame b 100800 later,
    // so both frame arrays are the same base with a 2-frame env stride
    if (int rc = alegpu_preprocess_step_strided(
            dev, dev + 100800, 2 * 100800, r->starts + t, T, r->obs + t * 28224, T * 28224,
            last ? carry_out : r->obs + (t + 1) * 28224, last ? 28224 : T * 28224, E, stream))
      return rc;
    ALEGPU_CUDA_CHECK(cudaEventRecord(r->consumed[slot], s));
    r->used[slot] = 1;
  }
  return ALEGPU_OK;
}

extern "C" int alegpu_rollout_set_write_index(alegpu_rollout *r, int64_t index) {
  ALEGPU_REQUIRE(r && index >= 0 && index < r->T, ALEGPU_INVALID_ARGUMENT,
                 "rollout_set_write_index: out of range");
  r->write_index = index;
  return ALEGPU_OK;
}

extern "C" int alegpu_rollout_get(alegpu_rollout *r, const float *next_values, float gamma,
                                  float lambda, alegpu_batch *flat, int flags, void *stream) {
  ALEGPU_REQUIRE(r && next_values, ALEGPU_INVALID_ARGUMENT, "rollout_get: null argument");
  ALEGPU_CUDA_CHECK(cudaSetDevice(r->device));
  cudaStream_t s = as_stream(stream);
  ALEGPU_CUDA_CHECK(cudaMemsetAsync(r->error_flag, 0, sizeof(int), s));
  // the flag check (if requested) happens after BOTH kernels are enqueued
  if (int rc = alegpu_buffer_get(r->rewards, r->values, next_values, r->terminals, r->truncations,
                                 r->starts, r->E, r->T, gamma, lambda, r->write_index,
                                 r->advantages, r->returns, r->masks, 1, r->error_flag, stream))
    return rc;
  // prepare_batch: old log-probs = log_softmax(stored logits) (src/bin/train.cc:279)
  if (int rc = launch_normalize_logits(r->logits, r->E * r->T, r->A, r->log_probs, s))
    return rc;
  if (flat) { // flatten(0,1) / ravel views (src/bin/train.cc:274-282)
    flat->observations = r->obs;
    flat->actions = r->actions;
    flat->log_probabilities = r->log_probs;
    flat->advantages = r->advantages;
    flat->returns = r->returns;
    flat->masks = r->masks;
    flat->size = r->E * r->T;
    flat->values = r->values; // ai::buffer::Batch.values (buffer.h:11), for the opt-in clipped value loss
  }
  if (!(flags & 1)) {
    int h = 0;
    ALEGPU_CUDA_CHECK(cudaMemcpyAsync(&h, r->error_flag, sizeof(int), cudaMemcpyDeviceToHost, s));
    ALEGPU_CUDA_CHECK(cudaStreamSynchronize(s));
    if (h)
      return fail(ALEGPU_INVALID_ARGUMENT,
                  "Episode starts, terminals, and truncations must be mutually exclusive.");
  }
  return ALEGPU_OK;
}
