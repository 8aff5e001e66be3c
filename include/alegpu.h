/* alegpu.h — C ABI of the B200-native PPO learner hot path (libalegpu.so).
 *
 * Drop-in boundary for cemlyn007/ale-libtorch-ppo's learner path.  The reference has no FFI:
 * its "operator API" is the C++ signatures in src/ai/{gae,buffer,vision}.h and
 * src/ai/ppo/{losses,train}.h, all over torch::Tensor.  Each entry point below replaces one
 * of them (cited file:line) with raw device pointers + explicit sizes + a cudaStream_t
 * (passed as void*; NULL = the library's own stream for ctx calls / the legacy default
 * stream for stateless calls).  No torch types cross this boundary; the library links only
 * cudart (+ NCCL when built with ALEGPU_WITH_NCCL).
 *
 * Conventions
 *   - every function returns an alegpu_status; alegpu_last_error() gives the message
 *     (thread-local).  The status codes map 1:1 onto the C++ exception types the reference
 *     throws for the same condition (SURVEY.md §8b "Error convention"); the C++ shim in
 *     include/ai/ re-throws them.
 *   - all pointers are DEVICE pointers unless the name ends in _host.
 *   - layouts are the reference's: rollout storage env-major [E,T,...] (row = e*T + t), bools
 *     are 1 byte, actions int64, observations [.,4,84,84] u8 with plane 0 = newest,
 *     parameters flat in `parameters()` order, each tensor OIHW / [out,in].
 *   - there is NO CPU fallback: without a CUDA device every compute entry point fails with
 *     ALEGPU_CUDA_ERROR.
 */
#ifndef ALEGPU_H_
#define ALEGPU_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  ALEGPU_OK = 0,
  ALEGPU_INVALID_ARGUMENT = 1, /* std::invalid_argument (gae.cc:11,19,24-46,52; rollout.cc:46-63) */
  ALEGPU_RUNTIME_ERROR = 2,    /* std::runtime_error    (buffer.cc:65; train.h:141)             */
  ALEGPU_OUT_OF_RANGE = 3,     /* std::out_of_range     (rollout.cc:315)                        */
  ALEGPU_CUDA_ERROR = 4        /* a CUDA runtime / NCCL failure (no reference counterpart)       */
} alegpu_status;

const char *alegpu_last_error(void);
/* library / build identification, e.g. "alegpu 0.1 sm_100a nccl=1" */
const char *alegpu_version(void);

/* ---- device memory helpers (so a host without torch can drive the library) ------------ */
int alegpu_device_count(int *count);
int alegpu_set_device(int device);
int alegpu_malloc(void **dptr, size_t bytes);
int alegpu_free(void *dptr);
int alegpu_host_alloc(void **hptr, size_t bytes); /* pinned */
int alegpu_host_free(void *hptr);
int alegpu_memcpy_h2d(void *dst, const void *src_host, size_t bytes, void *stream);
int alegpu_memcpy_d2h(void *dst_host, const void *src, size_t bytes, void *stream);
int alegpu_memcpy_d2d(void *dst, const void *src, size_t bytes, void *stream);
int alegpu_memset(void *dst, int value, size_t bytes, void *stream);
int alegpu_stream_synchronize(void *stream);

/* ---- stateless operators ---------------------------------------------------------------- */

/* ai::gae::gae — src/ai/gae.h:4-7, src/ai/gae.cc:4-80.
 * advantages[E,T] (out), rewards/values [E,T] f32, next_values [E] f32, flags [E,T] 1-byte
 * bools.  Returns ALEGPU_INVALID_ARGUMENT if two flags co-occur anywhere (gae.cc:49-53);
 * like the reference this check synchronises the stream. */
int alegpu_gae(float *advantages, const float *rewards, const float *values,
               const float *next_values, const uint8_t *terminals, const uint8_t *truncations,
               const uint8_t *episode_starts, int64_t E, int64_t T, float gamma, float lambda,
               void *stream);

/* ai::buffer::Buffer::get — src/ai/buffer.cc:58-77, in pointer form: clamps rewards to
 * [-1,1] IN PLACE (:67), runs the GAE scan (:68), returns = advantages + values (:70-71),
 * masks = !episode_starts (:74).  write_index != 0 -> ALEGPU_RUNTIME_ERROR (:64-65).
 * One fused kernel.  flags: bit0 = skip the flag-exclusivity host sync (error then surfaces
 * through *device_error_flag if non-NULL). */
int alegpu_buffer_get(float *rewards_inout, const float *values, const float *next_values,
                      const uint8_t *terminals, const uint8_t *truncations,
                      const uint8_t *episode_starts, int64_t E, int64_t T, float gamma,
                      float lambda, int64_t write_index, float *advantages, float *returns,
                      uint8_t *masks, int flags, int *device_error_flag, void *stream);

/* Observation preprocessing for ONE env step of E environments.
 * Replaces MaxAndSkipEnvironment's 2-frame max (src/ai/environment/max_and_skip.cc:33-42),
 * the libtorch vision path resize_frame_stacked_rgb_images + rgb_to_grayscale_frame_stacked_images
 * (src/ai/vision.cc:38-48, 71-84; weights :51; area mode :8-10) and
 * Rollout::update_observations (src/ai/rollout.cc:184-196) with one fused kernel:
 *   frames_a/frames_b [E,210,160,3] u8 HWC (the last two frames of the skip window)
 *   -> max -> area-resize to 84x84 per channel -> 0.2125R+0.7154G+0.0721B -> floor(x+.5) u8
 *   -> dst stack[e] = starts[e] ? {new,new,new,new} : {new, src[e][0], src[e][1], src[e][2]}.
 * src/dst are [.,4,84,84] u8 stacks addressed as base + e*stride (bytes); they must not
 * alias.  starts[e*starts_stride] is env e's flag (stride T reads the buffer's [E,T] array).  gray_f32 (optional, [E,84,84]) receives the un-quantised grayscale for parity. */
int alegpu_preprocess_step(const uint8_t *frames_a, const uint8_t *frames_b,
                           const uint8_t *starts, int64_t starts_stride, const uint8_t *src_stack,
                           int64_t src_env_stride, uint8_t *dst_stack, int64_t dst_env_stride,
                           float *gray_f32, int64_t E, void *stream);

/* as above with frame a/b of env e at frames_x + e*frame_env_stride (interleaved [E,2,...] slabs) */
int alegpu_preprocess_step_strided(const uint8_t *frames_a, const uint8_t *frames_b,
                                   int64_t frame_env_stride, const uint8_t *starts,
                                   int64_t starts_stride, const uint8_t *src_stack,
                                   int64_t src_env_stride, uint8_t *dst_stack,
                                   int64_t dst_env_stride, int64_t E, void *stream);

/* The same operator over a whole rollout whose raw frames are resident:
 *   frames [T,E,2,210,160,3] u8 (time-major, as the ring delivers them),
 *   episode_starts [E,T] (env-major, the buffer's own array: starts[e,t] != 0 means step t
 *   of env e was a reset, so the update after step t fills all four planes),
 *   obs [E,T,4,84,84] u8 rollout storage; carry_in [E,4,84,84] is the stack before step 0
 *   (it becomes slot 0), carry_out receives the stack after step T-1.  carry_in != carry_out.
 * Equivalent to T calls of alegpu_preprocess_step with dst = slot t+1. */
int alegpu_preprocess_rollout(const uint8_t *frames, const uint8_t *episode_starts,
                              const uint8_t *carry_in, uint8_t *carry_out, uint8_t *obs,
                              int64_t E, int64_t T, void *stream);

/* ai::vision::resize_grayscale_image(const std::vector<unsigned char>&, width, height, new_width,
 * new_height) — src/ai/vision.h:15-17, src/ai/vision.cc:86-95 — the LIVE path's per-frame resize
 * (ResizeEnvironment::resize, src/ai/environment/resize.cc:34-41), batched over n images:
 * images [n,height,width] u8 -> out [n,new_height,new_width] u8, device pointers.
 * Arithmetic: stbir_resize_uint8_linear(..., STBIR_1CHANNEL) with stb's defaults (Mitchell when an
 * axis shrinks, Catmull-Rom when it grows, clamp edges, linear light, round to nearest), restated
 * from the published algorithm — stb_image_resize2 (stb@0.0.0-20241109-5c20573, MODULE.bazel:38)
 * is not under /root/reference and the reference has no test of this overload: PARITY UNPINNED.
 * out_f32 (optional, [n,new_height,new_width]) receives the value before rounding.
 * new_width/new_height <= 0 -> ALEGPU_INVALID_ARGUMENT (std::invalid_argument, resize.cc:15-16). */
int alegpu_resize_grayscale(const uint8_t *images, int64_t n, int width, int height,
                            int new_width, int new_height, uint8_t *out, float *out_f32,
                            void *stream);

/* One env step of the LIVE wrapper stack (src/ai/rollout.cc:146-161) for E environments whose
 * emulator delivers grayscale screens: frames_a/frames_b = the last two 210x160 u8 screens of the
 * skip window at frames_x + e*frame_env_stride; each is resized to 84x84 as above
 * (ResizeEnvironment), the two results are max'ed (MaxAndSkipEnvironment, max_and_skip.cc:33-42;
 * pass frames_a == frames_b when the window ended after one frame, :40-42) and the frame stack is
 * updated exactly as in alegpu_preprocess_step (Rollout::update_observations, rollout.cc:184-196).
 * One fused kernel; stacks 8-byte aligned, src != dst. */
int alegpu_preprocess_gray_step(const uint8_t *frames_a, const uint8_t *frames_b,
                                int64_t frame_env_stride, const uint8_t *starts,
                                int64_t starts_stride, const uint8_t *src_stack,
                                int64_t src_env_stride, uint8_t *dst_stack,
                                int64_t dst_env_stride, int64_t E, void *stream);

/* The stand-alone libtorch vision operators of src/ai/vision.h:6-13 (the preprocess oracle's own
 * interface; the learner path itself uses the fused alegpu_preprocess_step):
 * resize_grayscale_image(Tensor [210,160]) (vision.cc:12-20, planes = 1),
 * resize_frame_stacked_grayscale_images ([B,F,210,160], :22-32, planes = B*F) and
 * resize_frame_stacked_rgb_images ([B,F,3,210,160], :38-48, planes = B*F*3) are all
 * interpolate(size 84x84, mode area) over float planes: images [planes,210,160] f32 -> out [planes,84,84]. */
int alegpu_vision_resize_area(const float *images, int64_t planes, float *out, void *stream);
/* rgb_to_grayscale_frame_stacked_images — vision.cc:71-84: [n = B*F, 3, 84, 84] f32 -> [n, 84, 84] f32,
 * matmul with (0.2125, 0.7154, 0.0721) over the channel axis (:51). */
int alegpu_vision_rgb_to_grayscale(const float *images, int64_t n, float *out, void *stream);

/* ai::ppo::losses::normalize_logits — src/ai/ppo/losses.cc:45-47. [n,A] -> [n,A]. */
int alegpu_normalize_logits(const float *logits, int64_t n, int64_t A, float *out,
                            void *stream);

/* ai::ppo::losses::compute — src/ai/ppo/losses.h:6-37, losses.cc:4-43 — plus the gradient
 * autograd produces for mini_batch_update (train.h:118-126) w.r.t. the RAW logits and the
 * values.  logits [n,A] raw network outputs, old_log_probs [n,A] normalised, actions [n]
 * int64, masks [n] 1-byte.  Outputs (any may be NULL): loss [1], per-sample clipped_losses,
 * value_losses, entropies, total_losses, ratio [n], dlogits [n,A], dvalues [n]. A <= 32.
 * Uses a per-device scratch buffer: not re-entrant on one device (call it from one thread / stream at a
 * time) and it may allocate, so not inside a stream capture — alegpu_learner_train has neither limit. */
int alegpu_loss(const float *logits, const float *old_log_probs, const int64_t *actions,
                const float *advantages, const float *values, const float *returns,
                const uint8_t *masks, int64_t n, int64_t A, float clip_param,
                float value_loss_coef, float entropy_coef, float *loss, float *clipped_losses,
                float *value_losses, float *entropies, float *total_losses, float *ratio,
                float *dlogits, float *dvalues, void *stream);

/* alegpu_loss for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid allocation).  It has NO counterpart in the reference.
 * Same arguments, kernels and results as alegpu_loss, but the
 * scratch (mask count, loss slot, per-block partials) is a 16.25 KB file-scope `static __device__`
 * array inside the library, so the call performs no cudaMalloc / cudaFree / allocating library call.
 * Limits: n <= ALEGPU_LOSS_STATIC_MAX_ROWS (larger n -> ALEGPU_INVALID_ARGUMENT; use alegpu_loss).
 * Every GPU that loads the library holds the array until exit.  All calls on one device share it and
 * it keeps the last call's contents (the slots that must start at zero are cleared on `stream` by each
 * call): calls that overlap on the GPU — two streams, or two threads — race; serialise them.  The
 * first call on a device looks the array's device address up (cudaGetSymbolAddress, once per device
 * with that device current); make that first call outside a stream capture. */
#define ALEGPU_LOSS_STATIC_MAX_ROWS (1 << 20)
int alegpu_loss_static(const float *logits, const float *old_log_probs, const int64_t *actions,
                       const float *advantages, const float *values, const float *returns,
                       const uint8_t *masks, int64_t n, int64_t A, float clip_param,
                       float value_loss_coef, float entropy_coef, float *loss,
                       float *clipped_losses, float *value_losses, float *entropies,
                       float *total_losses, float *ratio, float *dlogits, float *dvalues,
                       void *stream);

/* streams / events for hosts that overlap the frame hand-off with compute */
int alegpu_stream_create(void **stream);
int alegpu_stream_destroy(void *stream);
int alegpu_event_create(void **event);
int alegpu_event_destroy(void *event);
int alegpu_event_record(void *event, void *stream);
int alegpu_stream_wait_event(void *stream, void *event);
int alegpu_event_synchronize(void *event);
int alegpu_event_elapsed_ms(void *start, void *stop, float *ms);
/* kernels launched by this thread through the library so far (bench.py's gpu_launches) */
int64_t alegpu_launch_count(void);
/* launch tracer (profiling aid): between _begin and _end a CUDA event follows every kernel launch
 * this thread makes on `stream`; _end writes "site count total_us" lines, sorted, into out. */
int alegpu_trace_begin(void *stream);
int alegpu_trace_end(char *out, size_t out_bytes);
/* a fresh 128-byte ncclUniqueId (rank 0 creates it, the host broadcasts it) */
int alegpu_nccl_unique_id(void *out128);

/* ---- configs/<name>.yaml: Config + load_config — src/bin/train.cc:33-60, 108-136 -------------------
 * The reference's flat YAML schema, key for key, with its inline defaults (:111-134).  size_t keys
 * are int64 here, bools are int.  action_size is in the shipped files but never read by the
 * reference (`const size_t action_size = 4`, :36): it is accepted and forced to 4.  The last four
 * members are this library's opt-in extras (absent keys = reference semantics). */
typedef struct {
  int64_t total_environments; /* 512    */
  int64_t hidden_size;        /* 512    */
  int64_t horizon;            /* 128    */
  int64_t max_steps;          /* 108000 */
  int64_t frame_stack;        /* 4      */
  double learning_rate;       /* 2.5e-4 */
  float clip_param;           /* 0.1    */
  float value_loss_coef;      /* 0.5    */
  float entropy_coef;         /* 0.01   */
  int64_t num_epochs;         /* 1      */
  int64_t mini_batch_size;    /* 2048   */
  int64_t num_mini_batches;   /* 32     */
  float gae_discount;         /* 0.99   */
  float gae_lambda;           /* 0.95   */
  float max_gradient_norm;    /* 0.5    */
  int64_t num_rollouts;       /* 7000   */
  int64_t num_workers;        /* 16     */
  int64_t worker_batch_size;  /* 32     */
  int64_t frame_skip;         /* 4      */
  float max_return;           /* -1     */
  int record_observation, record_video, cuda_graph, deterministic; /* false */
  int64_t action_size;        /* always 4 */
  int normalize_advantages;   /* extra (alegpu_train_options), default 0 */
  int clip_value_loss;        /* extra (alegpu_train_options), default 0 */
  int shuffle;                /* extra: the host passes a randperm per epoch as options.indices */
  int64_t num_gpus;           /* extra: ranks the envs are sharded over, default 1 */
} alegpu_config;
int alegpu_config_defaults(alegpu_config *config);
/* load_config(path): a missing file or a scalar that does not convert -> ALEGPU_RUNTIME_ERROR
 * (YAML::BadFile / YAML::BadConversion are std::runtime_error); unknown keys are ignored. */
int alegpu_config_load(const char *path, alegpu_config *config);
int alegpu_config_parse(const char *yaml_text, alegpu_config *config);
/* What the reference leaves unchecked: `mini_batch_size` only sizes its Metrics tensors
 * (src/bin/train.cc:383-387) while train() slices E*T / num_mini_batches rows (train.h:139-151).
 * mini_batch_size * num_mini_batches != total_environments * horizon -> ALEGPU_INVALID_ARGUMENT;
 * E*T not divisible by num_mini_batches -> ALEGPU_RUNTIME_ERROR (train.h:140-143). */
int alegpu_config_validate(const alegpu_config *config);

/* ---- the learner: policy + optimiser + trainer ----------------------------------------- */
typedef struct alegpu_learner alegpu_learner;
typedef struct alegpu_rollout alegpu_rollout;

typedef enum {
  ALEGPU_MATH_FP32 = 0, /* CUDA-core fp32 implicit GEMM: the fp32-faithful parity mode      */
  ALEGPU_MATH_BF16 = 1, /* tcgen05/TMEM tensor-core implicit GEMM, bf16 operands, fp32 accum */
  ALEGPU_MATH_TF32 = 2  /* the same engine with tcgen05.mma kind::tf32: fp32 activations / weights /
                           gradients rounded to tf32 where they are produced, fp32 accumulation —
                           the precision cuDNN gives the reference's convolutions on a GPU          */
} alegpu_math;

typedef struct {
  int64_t hidden_size;    /* configs/<name>.yaml hidden_size (src/bin/train.cc:113)               */
  int64_t action_size;    /* fixed 4 in the reference (src/bin/train.cc:36)                   */
  int64_t max_minibatch;  /* largest minibatch (rows) any call will use; sizes workspaces    */
  int math;               /* alegpu_math                                                      */
  double learning_rate;   /* torch::optim::Adam(lr, eps=1e-5), src/bin/train.cc:360-362       */
  int device;             /* CUDA device ordinal                                              */
  int world_size;         /* >1: data-parallel; gradients are summed over ranks (NCCL)        */
  int rank;
  const void *nccl_unique_id; /* 128-byte ncclUniqueId shared by all ranks (world_size>1)     */
} alegpu_learner_config;

/* hyper-parameters — ai::ppo::train::Hyperparameters, src/ai/ppo/train.h:27-32 */
typedef struct {
  float clip_param;
  float value_loss_coef;
  float entropy_coef;
  float max_gradient_norm;
} alegpu_hyperparameters;

/* ai::ppo::train::Batch — src/ai/ppo/train.h:34-57 (device pointers, N rows) */
typedef struct {
  const uint8_t *observations;    /* [N,4,84,84] u8 */
  const int64_t *actions;         /* [N] */
  const float *log_probabilities; /* [N,A] normalised (prepare_batch, src/bin/train.cc:279) */
  const float *advantages;        /* [N] */
  const float *returns;           /* [N] */
  const uint8_t *masks;           /* [N] */
  int64_t size;                   /* N */
  const float *values;            /* [N] values stored at collection time (ai::buffer::Batch.values,
                                     src/ai/buffer.h:11; prepare_batch drops them, src/bin/train.cc:272-283):
                                     read only when alegpu_train_options.clip_value_loss; may be NULL */
} alegpu_batch;

/* ai::ppo::train::Metrics — src/ai/ppo/train.h:64-109: [epochs,nmb,mb] f32 (masks 1-byte),
 * clipped_gradients [epochs,nmb].  Any pointer may be NULL (that metric is skipped). */
typedef struct {
  float *loss;
  float *clipped_losses;
  float *value_losses;
  float *entropies;
  float *ratio;
  float *total_losses;
  float *advantages;
  float *returns;
  uint8_t *masks;
  float *clipped_gradients;
} alegpu_metrics;

/* The scalars log_data (src/bin/train.cc:163-210) extracts from Metrics with one `.item()` sync
 * each: clipped_gradients.mean() (:180-181), loss.mean() (:182), and the means of
 * clipped_losses / value_losses / entropies / ratio over the UNMASKED samples (mean(tensor, mask),
 * :144-147, :183-190); plus the unmasked mean of total_losses and the unmasked count. */
typedef struct {
  float mean_clipped_gradient;
  float mean_loss;
  float mean_clipped_loss;
  float mean_value_loss;
  float mean_entropy;
  float mean_ratio;
  float mean_total_loss;
  float unmasked;
} alegpu_metrics_summary;
/* One kernel fills *summary (DEVICE); summary_host (optional, pinned for a truly async copy)
 * receives it with one cudaMemcpyAsync on `stream`.  metrics holds [epochs,nmb,mb] arrays (NULL
 * members are skipped and reported as 0). */
int alegpu_metrics_summarize(const alegpu_metrics *metrics, int64_t num_epochs,
                             int64_t num_mini_batches, int64_t mini_batch_size,
                             alegpu_metrics_summary *summary, alegpu_metrics_summary *summary_host,
                             void *stream);
/* tensor.masked_select(mask) — the histogram feeds gather(tensor, mask) of src/bin/train.cc:149-155,
 * 194-207: out receives the values[i] with masks[i] != 0 in order (stable), *count (DEVICE int64)
 * how many.  out must hold n floats.  Not re-entrant (shares one scratch buffer). */
int alegpu_masked_select(const float *values, const uint8_t *masks, int64_t n, float *out,
                         int64_t *count, void *stream);

/* ---- rollout buffer: ai::buffer::Buffer — src/ai/buffer.h:27-58, src/ai/buffer.cc:5-77 ---- */
/* raw device pointers of the [E,T,...] storage (ai::buffer::Batch aliases storage too,
 * buffer.cc:73-76), so producers can write slots in place.  carry = 2 x [E,4,84,84] ping-pong
 * stacks for the observation carried across rollouts (Rollout::observations_). */
typedef struct {
  int64_t E, T, A;
  uint8_t *observations;      /* [E,T,4,84,84] */
  uint8_t *carry;             /* [2,E,4,84,84] */
  int64_t *actions;           /* [E,T] */
  float *rewards, *values, *advantages, *returns; /* [E,T] */
  float *logits, *log_probabilities;              /* [E,T,A] */
  uint8_t *terminals, *truncations, *episode_starts, *masks; /* [E,T] */
} alegpu_rollout_view;

/* Buffer::Buffer(total_environments, capacity, {4,84,84}, action_size, device) — buffer.cc:5-39 */
int alegpu_rollout_create(int64_t E, int64_t T, int64_t A, int device, alegpu_rollout **out);
int alegpu_rollout_destroy(alegpu_rollout *rollout);
int alegpu_rollout_get_view(alegpu_rollout *rollout, alegpu_rollout_view *view);
/* Buffer::add — buffer.cc:41-56: per-step device arrays ([E,...]) go to column `write index`,
 * which then advances mod T.  observations may be NULL when the producer already wrote the
 * slot in place (alegpu_preprocess_step with dst = slot t). */
int alegpu_rollout_add(alegpu_rollout *rollout, const uint8_t *observations,
                       const int64_t *actions, const float *rewards, const uint8_t *terminals,
                       const uint8_t *truncations, const uint8_t *episode_starts,
                       const float *logits, const float *values, void *stream);
int alegpu_rollout_set_write_index(alegpu_rollout *rollout, int64_t index);
/* The hand-off (replaces screen_buffers_/torch::stack/H2D, src/ai/rollout.cc:83-93,184-196):
 * frames_host = PINNED [T,E,2,210,160,3]; each step's slab is cudaMemcpyAsync'ed on a side
 * stream into a 4-slot device ring and preprocessed into slot t+1 on `stream` as it lands.
 * episode_starts[E,T] must already be in the buffer.  carry_index in {0,1} names the carry
 * stack holding the observation before step 0; the other one receives the stack after T-1. */
int alegpu_rollout_ingest(alegpu_rollout *rollout, const uint8_t *frames_host, int carry_index,
                          void *stream);
/* Buffer::get (buffer.cc:58-77) + prepare_batch (src/bin/train.cc:272-283): fills *flat with
 * the [N = E*T] views train() consumes.  ALEGPU_RUNTIME_ERROR if the buffer is not full
 * (:64-65); ALEGPU_INVALID_ARGUMENT if flags co-occur (gae.cc:49-53) unless flags bit0. */
int alegpu_rollout_get(alegpu_rollout *rollout, const float *next_values, float gamma,
                       float lambda, alegpu_batch *flat, int flags, void *stream);

int alegpu_learner_create(const alegpu_learner_config *config, alegpu_learner **out);
int alegpu_learner_destroy(alegpu_learner *learner);
int64_t alegpu_learner_param_count(const alegpu_learner *learner);
void *alegpu_learner_stream(alegpu_learner *learner);
/* Stream contract.  Every learner entry point (train, act, forward, grads, set_lr) enqueues on the
 * learner's OWN non-blocking stream (alegpu_learner_stream) and returns without waiting; the stateless
 * operators and the rollout buffer enqueue on the stream they are given.  A caller that mixes the two on
 * different streams (the reference does everything on the current stream, SURVEY §8b "Threading") orders
 * them with these two calls, which record / wait an event and never block the host:
 *   alegpu_learner_wait_stream(L, s): work the learner enqueues from now on starts after everything
 *                                     already enqueued on s (e.g. Buffer::add / get of this rollout);
 *   alegpu_stream_wait_learner(L, s): work enqueued on s from now on starts after everything the
 *                                     learner has enqueued so far (e.g. a Metrics read after train()).
 * include/ai/alegpu_shim.hpp brackets every Learner call with them on the shim's current stream. */
int alegpu_learner_wait_stream(alegpu_learner *learner, void *stream);
int alegpu_stream_wait_learner(alegpu_learner *learner, void *stream);
/* how gradients are exchanged (no reference counterpart: the reference is single-device):
 * "single" (world_size 1), "fused-nvlink" (every rank maps its peers' gradient buffers through CUDA
 * IPC and the optimiser kernel pulls, sums and applies them over NVLink in one launch per minibatch), or
 * "nccl" (ncclAllReduce per minibatch: ALEGPU_NO_FUSED_AR=1, or when peer mapping is not possible) */
const char *alegpu_learner_comm_mode(const alegpu_learner *learner);

/* Weights in the reference's `parameters()` order and native tensor layouts (host arrays of
 * param_count floats) — needed to start from the oracle's orthogonal init
 * (src/bin/train.cc:212-248) and to compare after training. */
int alegpu_learner_set_params(alegpu_learner *learner, const float *params_host);
int alegpu_learner_get_params(alegpu_learner *learner, float *params_host);
/* Adam moments in the same order/layout, and the optimiser step counter. */
int alegpu_learner_get_adam(alegpu_learner *learner, float *exp_avg_host, float *exp_avg_sq_host,
                            int64_t *step);
/* Checkpoint / resume (the reference has none): together with set_params this restores
 * a learner exactly — parameters, both Adam moments and the step counter that drives the bias
 * correction — so training continues bit-identically. */
int alegpu_learner_set_adam(alegpu_learner *learner, const float *exp_avg_host,
                            const float *exp_avg_sq_host, int64_t step);
/* the per-rollout lr anneal of src/bin/train.cc:424-428 (lr lives in a device scalar so a
 * captured graph stays correct) */
int alegpu_learner_set_lr(alegpu_learner *learner, double lr);

/* NetworkImpl::forward — src/bin/train.cc:255-265 — inference on n <= max_minibatch rows. */
int alegpu_learner_forward(alegpu_learner *learner, const uint8_t *observations, int64_t n,
                           float *logits, float *values);
/* The rollout-side policy call — the action-selector lambda of src/bin/train.cc:367-379, called at
 * src/ai/rollout.cc:207,268: forward of E observation stacks ([E,4,84,84] u8, contiguous; E may
 * exceed max_minibatch, it is processed in chunks), probabilities = softmax(logits), one
 * categorical draw per env (torch::multinomial(p, 1)).  The draw is the inverse CDF of u in [0,1):
 * action = first a with cumsum(p)[a] > u.  uniforms [E] supplies u (parity runs); NULL = a
 * counter-based generator keyed by (seed, step, env) — libtorch's Philox stream is not
 * reproducible, so parity with the reference is distributional.
 * Outputs (any may be NULL) are written at base + e*stride (strides in ELEMENTS) so they can land
 * in column t of the [E,T,...] rollout storage (Buffer::add, src/ai/buffer.cc:41-56):
 * logits[e*logits_env_stride + a], values[e*values_env_stride], actions[e*actions_env_stride].
 * actions_dense [E] (device) also receives the actions; actions_host (pinned, needs actions_dense)
 * gets them in ONE async copy on the learner's stream — replaces the per-env `.item()` syncs of
 * src/ai/rollout.cc:313.  Asynchronous on the learner's stream. */
int alegpu_learner_act(alegpu_learner *learner, const uint8_t *observations, int64_t E,
                       const float *uniforms, uint64_t seed, uint64_t step, float *logits,
                       int64_t logits_env_stride, float *values, int64_t values_env_stride,
                       int64_t *actions, int64_t actions_env_stride, int64_t *actions_dense,
                       int64_t *actions_host);
/* Post-ReLU conv activations and the FC output of the last forward, converted to the
 * reference's NCHW f32 layout (host arrays; parity/debug only). */
int alegpu_learner_get_activations(alegpu_learner *learner, int64_t n, float *conv1_host,
                                   float *conv2_host, float *conv3_host, float *fc_host);

/* forward + loss + backward of one minibatch exactly as mini_batch_update does up to (not
 * including) clip/step — src/ai/ppo/train.h:118-126.  grads_host receives the un-clipped
 * gradient in reference order/layout; loss_host/logits/values optional. */
int alegpu_learner_grads(alegpu_learner *learner, const alegpu_batch *minibatch,
                         const alegpu_hyperparameters *hp, float *grads_host, float *loss_host,
                         float *logits, float *values);
/* clip_grad_norm_ (src/ai/ppo/train.cc:12-46) + Adam step on caller-provided gradients
 * (reference order/layout).  total_norm_host = the PRE-clip norm. */
int alegpu_learner_apply_grads(alegpu_learner *learner, const float *grads_host,
                               double max_norm, float *total_norm_host);

/* ai::ppo::train::train — src/ai/ppo/train.h:133-157: epochs x minibatches of
 * {forward, loss, backward, clip, Adam}, minibatch k = rows [k*mb,(k+1)*mb) (the reference's
 * randperm at :146 is never consumed).  batch.size % num_mini_batches != 0 ->
 * ALEGPU_RUNTIME_ERROR (:140-143).  Asynchronous on the learner's stream.
 * flags: bit0 = replay through a cached CUDA graph of this exact call shape. */
int alegpu_learner_train(alegpu_learner *learner, const alegpu_batch *batch,
                         const alegpu_metrics *metrics, int64_t num_epochs,
                         int64_t num_mini_batches, const alegpu_hyperparameters *hp, int flags);
/* Opt-in minibatch preparation that BASELINE.json's north_star lists on the path but the reference
 * does not perform (src/ai/ppo/train.h:146-151 slices contiguously — its randperm is generated
 * and never consumed — and nothing normalises advantages).  Zero-initialised = reference semantics.
 *   indices: DEVICE [num_epochs * N] int64; minibatch k of epoch e trains on rows
 *            indices[e*N + k*mb .. e*N + (k+1)*mb) (one randperm(N) per epoch gives the usual PPO
 *            shuffle), gathered by one HBM-bound kernel (2 x 28 224 + 45 B per row).  The per-sample
 *            metrics at [e,k,:] then describe the gathered rows.  An index outside [0,N) makes THIS
 *            call fail with ALEGPU_OUT_OF_RANGE before anything is trained (one validation pass over
 *            the index vector and a 4-byte read-back: shuffled calls synchronise once).
 *   normalize_advantages: A' = (A - mean) / (std + 1e-8) per minibatch, mean and unbiased std over
 *            the minibatch's UNMASKED rows (over all ranks when world_size > 1); Metrics.advantages
 *            stays the raw value.
 *   clip_value_loss: value loss = 0.5 max((v-R)^2, (v_old + clamp(v-v_old, -clip_param, clip_param) - R)^2)
 *            with v_old = batch.values (north_star's "clipped value" loss; the reference has
 *            the plain 0.5 (v-R)^2, src/ai/ppo/losses.cc:15).  Metrics.value_losses holds the clipped form. */
typedef struct {
  const int64_t *indices;
  int normalize_advantages;
  int clip_value_loss;
} alegpu_train_options;
int alegpu_learner_train_ex(alegpu_learner *learner, const alegpu_batch *batch,
                            const alegpu_metrics *metrics, int64_t num_epochs,
                            int64_t num_mini_batches, const alegpu_hyperparameters *hp,
                            const alegpu_train_options *options, int flags);
/* number of kernels the last alegpu_learner_train call launched (or replayed) */
int64_t alegpu_learner_last_launch_count(const alegpu_learner *learner);

/* ---- rollout collection: ai::rollout::Rollout::rollout() — src/ai/rollout.cc:198-278 -------------
 * The emulator stays on CPU threads behind a step callback (where ALE's step and the last two raw
 * frames of the MaxAndSkip window plug in, src/ai/environment/max_and_skip.cc:21-45,
 * src/ai/rollout.cc:289-328); the collector owns the worker threads, the pinned frame ring and the
 * per-step hand-off (one cudaMemcpyAsync of the step's [E,2,210,160,3] slab on a copy stream, one
 * fused preprocess kernel into slot t+1, rewards/flags as one small copy, the E actions back in one
 * pinned D2H instead of E .item() syncs).
 * step(user, env, action, reset, frame_a, frame_b, &reward, &terminated, &truncated, &game_over):
 * reset != 0 -> environment->reset() (rollout.cc:299-305: reward 0, flags false, frames = the
 * reset observation twice); else step with `action` and write the window's last two RGB frames
 * [210,160,3].  Called concurrently for different envs.  NULL = the built-in Breakout-shaped
 * frame source (ALE and its ROMs are not part of this library). */
typedef void (*alegpu_env_step_fn)(void *user, int64_t env, int64_t action, int reset,
                                   uint8_t *frame_a, uint8_t *frame_b, float *reward,
                                   uint8_t *terminated, uint8_t *truncated, uint8_t *game_over);
typedef struct alegpu_collector alegpu_collector;
/* ai::rollout::Log — src/ai/rollout.h (steps, episodes, the finished episodes of this rollout) */
typedef struct {
  int64_t steps, episodes, finished_episodes, finished_games;
  float mean_episode_return, mean_episode_length, mean_game_return, mean_game_length;
} alegpu_rollout_log;
int alegpu_collector_create(int64_t E, int64_t T, int64_t A, int num_workers,
                            alegpu_env_step_fn step, void *user, uint64_t seed, int device,
                            alegpu_collector **out);
int alegpu_collector_destroy(alegpu_collector *collector);
/* T steps of {act, step all envs, buffer column t, next observation stack, episode bookkeeping},
 * one more policy call for the bootstrap values, then Buffer::get: on return the rollout storage is
 * full and *flat holds the [N] views alegpu_learner_train consumes.  std::out_of_range
 * (ALEGPU_OUT_OF_RANGE) for an action index outside [0, A) (rollout.cc:314-316). */
int alegpu_collector_rollout(alegpu_collector *collector, alegpu_rollout *rollout,
                             alegpu_learner *learner, float gamma, float lambda,
                             alegpu_batch *flat, alegpu_rollout_log *log);
/* the pinned host frame ring [T,E,2,210,160,3] of the last rollout (parity tests replay it) */
const uint8_t *alegpu_collector_frames(const alegpu_collector *collector);

/* Self-test of the tcgen05/TMEM implicit-GEMM engine behind ALEGPU_MATH_BF16 (no reference
 * counterpart: cuBLAS plays this role under libtorch).  C[M][N] fp32 = A . B^T with fp32 inputs
 * rounded to bf16 on load.  a_mn == 0: A is [M][R] row-major; a_mn == 1: A is stored [R][M];
 * likewise b_mn for B and N.  bn = tile width (32, 64, 128, 256; an MN-major B needs bn >= 64).
 * splits > 1: the reduction is cut into ceil(R / rps) parts, rps = R/splits rounded up to 64,
 * and C receives that many [M][N] partial products back to back (not summed). */
int alegpu_selftest_tc_gemm(const float *A, const float *B, float *C, int64_t M, int64_t N,
                            int64_t R, int a_mn, int b_mn, int bn, int splits, void *stream);
/* the same product with A and B first rounded to bf16 device buffers and read by the engine's
 * cp.async (global -> swizzled shared memory, no register staging) producers — the path every bf16
 * operand of the network takes; the fp32 entry point above exercises the register producers. */
int alegpu_selftest_tc_gemm_async(const float *A, const float *B, float *C, int64_t M, int64_t N,
                                  int64_t R, int a_mn, int b_mn, int bn, int splits, void *stream);

/* the same product on the kind::tf32 path behind ALEGPU_MATH_TF32: fp32 inputs copied as they are by
 * the cp.async producers, the tensor core multiplies their upper 19 bits (tf32), fp32 accumulation.
 * Contiguous dimensions must be multiples of 4. */
int alegpu_selftest_tc_gemm_tf32(const float *A, const float *B, float *C, int64_t M, int64_t N,
                                 int64_t R, int a_mn, int b_mn, int bn, int splits, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* ALEGPU_H_ */
