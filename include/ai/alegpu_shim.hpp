// alegpu_shim.hpp — header-only C++ mirror of the reference's `ai::` operator API over the C ABI.
//
// cemlyn007/ale-libtorch-ppo has no FFI: its learner path is the C++ signatures of
// src/ai/gae.h:4-7, src/ai/buffer.h:5-58, src/ai/ppo/losses.h:6-37 and src/ai/ppo/train.h:27-157,
// all over torch::Tensor.  This header re-creates those names, argument orders and exception
// types on top of include/alegpu.h (libalegpu.so), over a minimal device-array view, so the
// reference's call sites (src/bin/train.cc:364-458, src/ai/rollout.cc:231,269-270) keep their
// shape.  With -DALEGPU_WITH_TORCH it adds torch::Tensor overloads that only unwrap
// data_ptr()/sizes() — no libtorch call sits on the hot path.
//
// Error convention: alegpu_status -> the exception the reference throws for the
// same condition: INVALID_ARGUMENT -> std::invalid_argument (gae.cc:11,19,24-46,52),
// RUNTIME_ERROR -> std::runtime_error (buffer.cc:65, train.h:141), OUT_OF_RANGE ->
// std::out_of_range (rollout.cc:315); a CUDA failure -> std::runtime_error.
#ifndef AI_ALEGPU_SHIM_HPP_
#define AI_ALEGPU_SHIM_HPP_

#include <cstddef>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "../alegpu.h"

#ifdef ALEGPU_WITH_TORCH
#include <torch/torch.h>
#endif

namespace ai {
namespace alegpu_shim {

inline void check(int rc) {
  if (rc == ALEGPU_OK)
    return;
  const std::string m = alegpu_last_error();
  if (rc == ALEGPU_INVALID_ARGUMENT)
    throw std::invalid_argument(m);
  if (rc == ALEGPU_OUT_OF_RANGE)
    throw std::out_of_range(m);
  throw std::runtime_error(m);
}

// A typed view of device memory with a shape: what the reference passes as torch::Tensor.
template <class T> struct View {
  T *data = nullptr;
  std::vector<int64_t> sizes;
  View() = default;
  View(T *p, std::vector<int64_t> s) : data(p), sizes(std::move(s)) {}
  int64_t size(size_t d) const { return sizes.at(d); }
  int64_t numel() const {
    int64_t n = 1;
    for (int64_t s : sizes)
      n *= s;
    return n;
  }
#ifdef ALEGPU_WITH_TORCH
  View(const torch::Tensor &t) // NOLINT: implicit on purpose, mirrors passing a tensor
      : data(static_cast<T *>(t.data_ptr())), sizes(t.sizes().begin(), t.sizes().end()) {}
#endif
};
using F32 = View<float>;
using U8 = View<uint8_t>;
using I64 = View<int64_t>;

// the CUDA stream every shim call is enqueued on (the reference uses the current stream)
inline void *&current_stream() {
  static thread_local void *s = nullptr;
  return s;
}

} // namespace alegpu_shim

// ---- src/ai/vision.h:15-17 — the live path's per-frame grayscale resize (vision.cc:86-95) ------------
namespace vision {
// images [n,height,width] u8 -> out [n,new_height,new_width] u8 (device views); n = 1 with a
// [height*width] view is the reference's single-image std::vector form.
inline void resize_grayscale_image(const alegpu_shim::U8 &image, int width, int height,
                                   int new_width, int new_height, alegpu_shim::U8 &out) {
  using namespace alegpu_shim;
  if (width <= 0 || height <= 0 || image.numel() % ((int64_t)width * height) != 0)
    throw std::invalid_argument("image.size() must be a multiple of width*height"); // vision.cc:89
  const int64_t n = image.numel() / ((int64_t)width * height);
  if (new_width > 0 && new_height > 0 && out.numel() != n * new_width * new_height)
    throw std::invalid_argument("out must hold n*new_width*new_height bytes");
  check(alegpu_resize_grayscale(image.data, n, width, height, new_width, new_height, out.data,
                                nullptr, current_stream()));
}
// The libtorch overloads of src/ai/vision.h:6-13 over device views of FLOAT planes (the header's own
// note: "when using the GPU, the dtype must be a float"); `out` is caller-owned, shapes as the reference
// returns them.  All are interpolate(size 84x84, mode area) / a matmul with the grayscale weights.
inline void resize_grayscale_image(const alegpu_shim::F32 &image, alegpu_shim::F32 &out) { // vision.cc:12-20
  using namespace alegpu_shim;
  if (image.sizes != std::vector<int64_t>{210, 160} || out.numel() != 84 * 84)
    throw std::invalid_argument("image must be [210,160], out [84,84]");
  check(alegpu_vision_resize_area(image.data, 1, out.data, current_stream()));
}
inline void resize_frame_stacked_grayscale_images(const alegpu_shim::F32 &images, // vision.cc:22-32
                                                  alegpu_shim::F32 &out) {
  using namespace alegpu_shim;
  if (images.sizes.size() != 4 || images.size(2) != 210 || images.size(3) != 160 ||
      out.numel() != images.size(0) * images.size(1) * 84 * 84)
    throw std::invalid_argument("images must be [B,F,210,160], out [B,F,84,84]");
  check(alegpu_vision_resize_area(images.data, images.size(0) * images.size(1), out.data, current_stream()));
}
inline void resize_frame_stacked_rgb_images(const alegpu_shim::F32 &images, // vision.cc:38-48
                                            alegpu_shim::F32 &out) {
  using namespace alegpu_shim;
  if (images.sizes.size() != 5 || images.size(2) != 3 || images.size(3) != 210 || images.size(4) != 160 ||
      out.numel() != images.size(0) * images.size(1) * 3 * 84 * 84)
    throw std::invalid_argument("images must be [B,F,3,210,160], out [B,F,3,84,84]");
  check(alegpu_vision_resize_area(images.data, images.size(0) * images.size(1) * 3, out.data,
                                  current_stream()));
}
inline void rgb_to_grayscale_frame_stacked_images(const alegpu_shim::F32 &images, // vision.cc:71-84
                                                  alegpu_shim::F32 &out) {
  using namespace alegpu_shim;
  if (images.sizes.size() != 5 || images.size(2) != 3 || images.size(3) != 84 || images.size(4) != 84 ||
      out.numel() != images.size(0) * images.size(1) * 84 * 84)
    throw std::invalid_argument("images must be [B,F,3,84,84], out [B,F,84,84]");
  check(alegpu_vision_rgb_to_grayscale(images.data, images.size(0) * images.size(1), out.data,
                                       current_stream()));
}
#ifdef ALEGPU_WITH_TORCH
// the reference's exact signatures (src/ai/vision.h:6-13): CUDA float tensors in, new tensors out;
// torch only allocates the result
inline torch::Tensor resize_grayscale_image(const torch::Tensor &image) {
  auto out = torch::empty({84, 84}, image.options());
  alegpu_shim::F32 o(out);
  resize_grayscale_image(alegpu_shim::F32(image.contiguous()), o);
  return out;
}
inline torch::Tensor resize_frame_stacked_grayscale_images(const torch::Tensor &images) {
  auto out = torch::empty({images.size(0), images.size(1), 84, 84}, images.options());
  alegpu_shim::F32 o(out);
  resize_frame_stacked_grayscale_images(alegpu_shim::F32(images.contiguous()), o);
  return out;
}
inline torch::Tensor resize_frame_stacked_rgb_images(const torch::Tensor &images) {
  auto out = torch::empty({images.size(0), images.size(1), 3, 84, 84}, images.options());
  alegpu_shim::F32 o(out);
  resize_frame_stacked_rgb_images(alegpu_shim::F32(images.contiguous()), o);
  return out;
}
inline torch::Tensor rgb_to_grayscale_frame_stacked_images(const torch::Tensor &images) {
  auto out = torch::empty({images.size(0), images.size(1), 84, 84}, images.options());
  alegpu_shim::F32 o(out);
  rgb_to_grayscale_frame_stacked_images(alegpu_shim::F32(images.contiguous()), o);
  return out;
}
#endif
} // namespace vision

// ---- src/ai/gae.h:4-7 -----------------------------------------------------------------------------
namespace gae {
inline void gae(alegpu_shim::F32 &advantages, const alegpu_shim::F32 &rewards,
                const alegpu_shim::F32 &values, const alegpu_shim::F32 &next_values,
                const alegpu_shim::U8 &terminals, const alegpu_shim::U8 &truncations,
                const alegpu_shim::U8 &episode_starts, float gamma, float lambda) {
  using namespace alegpu_shim;
  if (rewards.sizes.size() != 2) // gae.cc:9-12
    throw std::invalid_argument("rewards must be 2 dimensional [environments, steps]");
  const std::vector<int64_t> *shapes[5] = {&values.sizes, &advantages.sizes, &terminals.sizes,
                                           &truncations.sizes, &episode_starts.sizes};
  for (const std::vector<int64_t> *s : shapes)
    if (*s != rewards.sizes) // gae.cc:24-46
      throw std::invalid_argument("all [E,T] arguments must have the shape of rewards");
  if (next_values.sizes.size() != 1 || next_values.size(0) != rewards.size(0)) // gae.cc:17-20
    throw std::invalid_argument("next_values must be [environments]");
  check(alegpu_gae(advantages.data, rewards.data, values.data, next_values.data, terminals.data,
                   truncations.data, episode_starts.data, rewards.size(0), rewards.size(1), gamma,
                   lambda, current_stream()));
}
#ifdef ALEGPU_WITH_TORCH
// the reference's exact signature (src/ai/gae.h:4-7); bool tensors are 1-byte, as the C ABI expects
inline void gae(torch::Tensor &advantages, const torch::Tensor &rewards, const torch::Tensor &values,
                const torch::Tensor &next_values, const torch::Tensor &terminals,
                const torch::Tensor &truncations, const torch::Tensor &episode_starts, float gamma,
                float lambda) {
  alegpu_shim::F32 a(advantages);
  gae(a, alegpu_shim::F32(rewards), alegpu_shim::F32(values), alegpu_shim::F32(next_values),
      alegpu_shim::U8(terminals), alegpu_shim::U8(truncations), alegpu_shim::U8(episode_starts), gamma,
      lambda);
}
#endif
} // namespace gae

// ---- src/ai/buffer.h:5-58 ---------------------------------------------------------------------------
namespace buffer {
struct Batch { // buffer.h:5-25: members alias the buffer's storage (valid until the next add)
  alegpu_shim::U8 observations;
  alegpu_shim::I64 actions;
  alegpu_shim::F32 rewards;
  alegpu_shim::U8 masks;
  alegpu_shim::F32 logits, values, advantages, returns;
  alegpu_shim::F32 log_probabilities; // prepare_batch's normalize_logits, src/bin/train.cc:279
  alegpu_batch flat{};                // the [N = E*T] views train() consumes
};

class Buffer {
public:
  Buffer(size_t total_environments, size_t capacity, std::vector<size_t> observation_shape,
         size_t action_size, int device = 0)
      : E_((int64_t)total_environments), T_((int64_t)capacity), A_((int64_t)action_size) {
    if (observation_shape != std::vector<size_t>{4, 84, 84})
      throw std::invalid_argument("observation_shape must be {4, 84, 84}");
    alegpu_shim::check(alegpu_rollout_create(E_, T_, A_, device, &h_));
    alegpu_shim::check(alegpu_rollout_get_view(h_, &view_));
  }
  ~Buffer() {
    if (h_)
      alegpu_rollout_destroy(h_);
  }
  Buffer(const Buffer &) = delete;
  Buffer &operator=(const Buffer &) = delete;

  // buffer.cc:41-56; observations may be empty when alegpu_preprocess_step wrote the slot in place
  void add(const alegpu_shim::U8 &observations, const alegpu_shim::I64 &actions,
           const alegpu_shim::F32 &rewards, const alegpu_shim::U8 &terminals,
           const alegpu_shim::U8 &truncations, const alegpu_shim::U8 &episode_starts,
           const alegpu_shim::F32 &logits, const alegpu_shim::F32 &values) {
    alegpu_shim::check(alegpu_rollout_add(h_, observations.data, actions.data, rewards.data,
                                          terminals.data, truncations.data, episode_starts.data,
                                          logits.data, values.data, alegpu_shim::current_stream()));
  }
  // buffer.cc:58-77 (+ prepare_batch): clamp, GAE, returns, masks, old log-probs
  Batch get(const alegpu_shim::F32 &next_values, float discount, float lambda) {
    using namespace alegpu_shim;
    Batch b;
    check(alegpu_rollout_get(h_, next_values.data, discount, lambda, &b.flat, 0, current_stream()));
    b.observations = U8(view_.observations, {E_, T_, 4, 84, 84});
    b.actions = I64(view_.actions, {E_, T_});
    b.rewards = F32(view_.rewards, {E_, T_});
    b.masks = U8(view_.masks, {E_, T_});
    b.logits = F32(view_.logits, {E_, T_, A_});
    b.values = F32(view_.values, {E_, T_});
    b.advantages = F32(view_.advantages, {E_, T_});
    b.returns = F32(view_.returns, {E_, T_});
    b.log_probabilities = F32(view_.log_probabilities, {E_, T_, A_});
    return b;
  }
  const alegpu_rollout_view &view() const { return view_; }
  alegpu_rollout *handle() { return h_; }

private:
  int64_t E_, T_, A_;
  alegpu_rollout *h_ = nullptr;
  alegpu_rollout_view view_{};
};
} // namespace buffer

namespace ppo {
// ---- src/ai/ppo/losses.h:6-37 ----------------------------------------------------------------------
namespace losses {
struct Metrics { // losses.h:6-13; every member is a caller-owned device array, loss is [1]
  alegpu_shim::F32 loss, clipped_losses, value_losses, entropies, ratio, total_losses;
};
inline void normalize_logits(const alegpu_shim::F32 &logits, alegpu_shim::F32 &out) {
  using namespace alegpu_shim;
  check(alegpu_normalize_logits(logits.data, logits.numel() / logits.sizes.back(),
                                logits.sizes.back(), out.data, current_stream()));
}
// compute(logits, old log-probs, actions, advantages, values, returns, masks, clip, c_v, c_e):
// `logits` are the RAW network outputs (the reference normalises them first, train.h:118-119);
// dlogits/dvalues (optional) receive what autograd would give mini_batch_update.
inline void compute(Metrics &out, const alegpu_shim::F32 &logits,
                    const alegpu_shim::F32 &old_log_probabilities,
                    const alegpu_shim::I64 &actions, const alegpu_shim::F32 &advantages,
                    const alegpu_shim::F32 &values, const alegpu_shim::F32 &returns,
                    const alegpu_shim::U8 &masks, float clip_param, float value_loss_coef,
                    float entropy_coef, float *dlogits = nullptr, float *dvalues = nullptr) {
  using namespace alegpu_shim;
  check(alegpu_loss(logits.data, old_log_probabilities.data, actions.data, advantages.data,
                    values.data, returns.data, masks.data, actions.numel(), logits.sizes.back(),
                    clip_param, value_loss_coef, entropy_coef, out.loss.data,
                    out.clipped_losses.data, out.value_losses.data, out.entropies.data,
                    out.total_losses.data, out.ratio.data, dlogits, dvalues, current_stream()));
}
#ifdef ALEGPU_WITH_TORCH
// the reference's signature and return type (losses.h:6-37): tensors in, a Metrics of NEW tensors out.
// (log-softmax is idempotent, so passing the already normalised log-probabilities, as the reference's
// caller does at train.h:118-124, gives the same numbers as passing raw logits.)
struct TensorMetrics { // losses.h:6-13
  torch::Tensor loss, clipped_losses, value_losses, entropies, ratio, total_losses;
};
inline TensorMetrics compute(const torch::Tensor &log_probabilities,
                             const torch::Tensor &old_log_probabilities, const torch::Tensor &actions,
                             const torch::Tensor &advantages, const torch::Tensor &values,
                             const torch::Tensor &returns, const torch::Tensor &masks, float clip_param,
                             float value_loss_coef, float entropy_coef) {
  const auto f = log_probabilities.options();
  const int64_t n = actions.numel();
  TensorMetrics t{torch::empty({}, f),  torch::empty({n}, f), torch::empty({n}, f),
                  torch::empty({n}, f), torch::empty({n}, f), torch::empty({n}, f)};
  Metrics m{alegpu_shim::F32(t.loss),      alegpu_shim::F32(t.clipped_losses),
            alegpu_shim::F32(t.value_losses), alegpu_shim::F32(t.entropies),
            alegpu_shim::F32(t.ratio),     alegpu_shim::F32(t.total_losses)};
  const auto mk = masks.to(torch::kUInt8); // bool tensors are 1 byte; make the dtype explicit
  compute(m, alegpu_shim::F32(log_probabilities.contiguous()),
          alegpu_shim::F32(old_log_probabilities.contiguous()), alegpu_shim::I64(actions.contiguous()),
          alegpu_shim::F32(advantages.contiguous()), alegpu_shim::F32(values.contiguous()),
          alegpu_shim::F32(returns.contiguous()), alegpu_shim::U8(mk), clip_param, value_loss_coef,
          entropy_coef);
  return t;
}
#endif
} // namespace losses

// ---- src/ai/ppo/train.h:27-157 ----------------------------------------------------------------------
namespace train {
struct Hyperparameters { // train.h:27-32
  float clip_param, value_loss_coef, entropy_coef, max_gradient_norm;
};
struct Batch { // train.h:34-57, [N,...] rows
  alegpu_shim::U8 observations;
  alegpu_shim::I64 actions;
  alegpu_shim::F32 log_probabilities, advantages, returns;
  alegpu_shim::U8 masks;
};
struct Metrics { // train.h:64-109: [epochs,nmb,mb] (+ clipped_gradients [epochs,nmb])
  alegpu_shim::F32 loss, clipped_losses, value_losses, entropies, ratio, total_losses, advantages,
      returns;
  alegpu_shim::U8 masks;
  alegpu_shim::F32 clipped_gradients;
};

// Network (src/bin/train.cc:230-269) + torch::optim::Adam (:360-362) live in one alegpu_learner.
class Learner {
public:
  Learner(int64_t hidden_size, int64_t action_size, int64_t max_minibatch, double learning_rate,
          int math = ALEGPU_MATH_BF16, int device = 0) {
    alegpu_learner_config c{hidden_size, action_size, max_minibatch, math, learning_rate,
                            device,      1,           0,             nullptr};
    alegpu_shim::check(alegpu_learner_create(&c, &h_));
  }
  ~Learner() {
    if (h_)
      alegpu_learner_destroy(h_);
  }
  Learner(const Learner &) = delete;
  Learner &operator=(const Learner &) = delete;
  alegpu_learner *handle() { return h_; }
  int64_t parameter_count() const { return alegpu_learner_param_count(h_); }
  // parameters() order, OIHW / [out,in] (the reference's orthogonal init, train.cc:212-248)
  void load_parameters(const std::vector<float> &flat) {
    if ((int64_t)flat.size() != parameter_count())
      throw std::invalid_argument("load_parameters: wrong parameter count");
    alegpu_shim::check(alegpu_learner_set_params(h_, flat.data()));
  }
  std::vector<float> parameters() {
    std::vector<float> flat((size_t)parameter_count());
    alegpu_shim::check(alegpu_learner_get_params(h_, flat.data()));
    return flat;
  }
  void set_learning_rate(double lr) { alegpu_shim::check(alegpu_learner_set_lr(h_, lr)); } // :424-428
  // Stream contract (include/alegpu.h): the learner works on its own stream; every call made through
  // this shim is ordered after what the shim's current stream holds (Buffer::add / get, gae, ...) and
  // the current stream is ordered after it, as if everything ran on the reference's single current stream.
  void enter() { alegpu_shim::check(alegpu_learner_wait_stream(h_, alegpu_shim::current_stream())); }
  void leave() { alegpu_shim::check(alegpu_stream_wait_learner(h_, alegpu_shim::current_stream())); }
  // NetworkImpl::forward (src/bin/train.cc:255-265): logits [n,A], values [n] (caller-owned views)
  void forward(const alegpu_shim::U8 &observations, alegpu_shim::F32 &logits, alegpu_shim::F32 &values) {
    enter();
    alegpu_shim::check(alegpu_learner_forward(h_, observations.data, observations.size(0), logits.data,
                                              values.data));
    leave();
  }

private:
  alegpu_learner *h_ = nullptr;
};
// The two objects the reference's train() takes (train.h:133-136): a NetworkModel and an optimizer.
// Both are views of the one Learner that owns the flat parameter / Adam buffers.
struct Network { // satisfies the NetworkModel shape of train.h:12-25 where the trainer needs it
  Learner &learner;
  Network *operator->() { return this; }
  void train() {} // train.h:138: train mode (no dropout / batch-norm in this network)
};
struct Adam { // torch::optim::Adam(lr, eps = 1e-5), src/bin/train.cc:360-362
  Learner &learner;
  void set_lr(double lr) { learner.set_learning_rate(lr); } // :424-428
};

// train(network, optimizer, metrics, indices, batch, num_epochs, num_mini_batches, hp) —
// train.h:133-157.  `indices` is accepted and, exactly like the reference (:146), not consumed.
inline void train(Learner &learner, Metrics &metrics, alegpu_shim::I64 & /*indices*/, Batch &batch,
                  size_t num_epochs, size_t num_mini_batches, Hyperparameters &hp,
                  bool cuda_graph = true) {
  using namespace alegpu_shim;
  alegpu_batch b{batch.observations.data, batch.actions.data,  batch.log_probabilities.data,
                 batch.advantages.data,   batch.returns.data,  batch.masks.data,
                 batch.actions.numel(),   nullptr};
  alegpu_metrics m{metrics.loss.data,       metrics.clipped_losses.data, metrics.value_losses.data,
                   metrics.entropies.data,  metrics.ratio.data,          metrics.total_losses.data,
                   metrics.advantages.data, metrics.returns.data,        metrics.masks.data,
                   metrics.clipped_gradients.data};
  alegpu_hyperparameters h{hp.clip_param, hp.value_loss_coef, hp.entropy_coef, hp.max_gradient_norm};
  if (num_mini_batches == 0 || b.size % (int64_t)num_mini_batches != 0) // train.h:140-143
    throw std::runtime_error("Batch size must be divisible by num_mini_batches");
  learner.enter(); // after the Buffer::get / prepare_batch work on the caller's stream
  check(alegpu_learner_train(learner.handle(), &b, &m, (int64_t)num_epochs,
                             (int64_t)num_mini_batches, &h, cuda_graph ? 1 : 0));
  learner.leave(); // a Metrics read on the caller's stream sees the finished update
}
// the reference's argument list (train.h:133-136):
//   train(network, optimizer, metrics, indices, batch, num_epochs, num_mini_batches, hyperparameters)
inline void train(Network &network, Adam &optimizer, Metrics &metrics, alegpu_shim::I64 &indices,
                  Batch &batch, size_t num_epochs, size_t num_mini_batches, Hyperparameters &hp) {
  if (&network.learner != &optimizer.learner)
    throw std::invalid_argument("network and optimizer must belong to the same Learner");
  train(network.learner, metrics, indices, batch, num_epochs, num_mini_batches, hp);
}
} // namespace train
} // namespace ppo
} // namespace ai

#endif // AI_ALEGPU_SHIM_HPP_
