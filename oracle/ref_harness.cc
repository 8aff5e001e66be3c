// oracle/ref_harness.cc — TEST INFRASTRUCTURE, NOT PRODUCT.
//
// Thin extern "C" harness around the UNMODIFIED reference sources compiled where they
// lie under /root/reference/src (gae.cc, buffer.cc, vision.cc, ppo/losses.cc,
// ppo/train.cc + train.h), against the libtorch that ships in the pip wheel, CPU only.
// Built by oracle/Makefile into oracle/_ref/libref_oracle.so (git-ignored; travels to
// the GPU box). Used by tests/ (as the checker), by tools/make_golden.py (fixture
// generation) and by `bench.py --impl reference` / the cpu_baseline leg (timed baseline).
//
// Pieces of the reference that live in translation units which cannot be compiled here
// (they need ALE / yaml-cpp / tensorboard_logger) are restated below, each citing the
// lines it follows:
//   * Nature-CNN policy        — src/bin/train.cc:212-270
//   * prepare_batch            — src/bin/train.cc:272-283
//   * 2-frame max              — src/ai/environment/max_and_skip.cc:33-42
//   * frame-stack update       — src/ai/rollout.cc:184-196
#include "ai/buffer.h"
#include "ai/gae.h"
#include "ai/ppo/losses.h"
#include "ai/ppo/train.h"
#include "ai/vision.h"
#include "stb_image_resize2.h"

#include <ATen/Parallel.h>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <torch/torch.h>

namespace {

std::string g_error;

template <typename F> int guarded(F &&f) {
  try {
    f();
    return 0;
  } catch (const std::invalid_argument &e) {
    g_error = e.what();
    return 1;
  } catch (const std::runtime_error &e) {
    g_error = e.what();
    return 2;
  } catch (const std::exception &e) {
    g_error = e.what();
    return 3;
  }
}

torch::Tensor f32(const float *p, std::vector<int64_t> shape) {
  return torch::from_blob(const_cast<float *>(p), shape, torch::kFloat32);
}
torch::Tensor u8(const uint8_t *p, std::vector<int64_t> shape) {
  return torch::from_blob(const_cast<uint8_t *>(p), shape, torch::kUInt8);
}
torch::Tensor b8(const uint8_t *p, std::vector<int64_t> shape) {
  return torch::from_blob(const_cast<uint8_t *>(p), shape, torch::kUInt8)
      .to(torch::kBool);
}
torch::Tensor i64(const int64_t *p, std::vector<int64_t> shape) {
  return torch::from_blob(const_cast<int64_t *>(p), shape, torch::kInt64);
}
void out_f32(float *dst, const torch::Tensor &t) {
  if (!dst)
    return;
  auto c = t.detach().to(torch::kFloat32).contiguous();
  std::memcpy(dst, c.data_ptr<float>(), sizeof(float) * c.numel());
}
void out_u8(uint8_t *dst, const torch::Tensor &t) {
  if (!dst)
    return;
  auto c = t.detach().to(torch::kUInt8).contiguous();
  std::memcpy(dst, c.data_ptr<uint8_t>(), c.numel());
}

// ---- policy network: restated from src/bin/train.cc:212-270 -------------------------
template <typename Layer> Layer ortho(Layer layer, double gain = std::sqrt(2.0)) {
  torch::nn::init::orthogonal_(layer->weight, gain);
  torch::nn::init::constant_(layer->bias, 0.0);
  return layer;
}

struct PolicyImpl : torch::nn::Module {
  struct Out {
    torch::Tensor logits;
    torch::Tensor value;
  };
  PolicyImpl(int64_t hidden, int64_t actions)
      : trunk(ortho(torch::nn::Conv2d(torch::nn::Conv2dOptions(4, 32, 8).stride(4))),
              torch::nn::ReLU(),
              ortho(torch::nn::Conv2d(torch::nn::Conv2dOptions(32, 64, 4).stride(2))),
              torch::nn::ReLU(),
              ortho(torch::nn::Conv2d(torch::nn::Conv2dOptions(64, 64, 3).stride(1))),
              torch::nn::ReLU(), torch::nn::Flatten(),
              ortho(torch::nn::Linear(64 * 7 * 7, hidden))),
        pi(ortho(torch::nn::Linear(hidden, actions), 0.01)),
        vf(ortho(torch::nn::Linear(hidden, 1), 1.0)) {
    register_module("sequential", trunk);
    register_module("action_head", pi);
    register_module("value_head", vf);
  }
  Out forward(torch::Tensor x) {
    {
      torch::NoGradGuard guard;
      x = x.to(torch::kFloat32);
      x.divide_(255.0);
    }
    x = trunk->forward(x);
    return {pi->forward(x), vf->forward(x).squeeze(-1)};
  }
  torch::nn::Sequential trunk;
  torch::nn::Linear pi, vf;
};
TORCH_MODULE(Policy);

// prepare_batch: src/bin/train.cc:272-283
ai::ppo::train::Batch flatten_batch(ai::buffer::Batch &b) {
  auto logits = b.logits.view({-1, b.logits.size(2)});
  return {b.observations.flatten(0, 1),
          b.actions.ravel(),
          ai::ppo::losses::normalize_logits(logits),
          b.advantages.ravel(),
          b.returns.ravel(),
          b.masks.ravel()};
}

// 2-frame max (max_and_skip.cc:33-42) then the libtorch vision path (vision.cc:38-84).
// frames_*: [B,210,160,3] u8 HWC. Returns gray f32 [B,84,84] (no quantisation).
torch::Tensor preprocess_gray(const torch::Tensor &fa, const torch::Tensor &fb) {
  auto pooled = torch::max(fa, fb);                              // u8, exact
  auto chw = pooled.permute({0, 3, 1, 2}).to(torch::kFloat32);   // [B,3,210,160]
  auto in5 = chw.unsqueeze(1).contiguous();                      // [B,1,3,210,160]
  auto small = ai::vision::resize_frame_stacked_rgb_images(in5); // [B,1,3,84,84]
  auto gray = ai::vision::rgb_to_grayscale_frame_stacked_images(small);
  return gray.squeeze(1);
}
torch::Tensor quantise(const torch::Tensor &gray) {
  return torch::floor(gray + 0.5f).clamp(0, 255).to(torch::kUInt8);
}

// Frame-stack update: src/ai/rollout.cc:184-196. stack [E,4,84,84] u8, plane 0 newest.
void stack_update(torch::Tensor &stack, const torch::Tensor &frames,
                  const torch::Tensor &starts) {
  using torch::indexing::Slice;
  for (int64_t p = stack.size(1) - 1; p > 0; --p)
    stack.index_put_({Slice(), p}, stack.index({Slice(), p - 1}).clone());
  auto acc = starts.accessor<bool, 1>();
  for (int64_t e = 0; e < stack.size(0); ++e)
    if (acc[e])
      stack.select(0, e).copy_(frames.select(0, e));
  stack.index_put_({Slice(), 0}, frames);
}

struct Learner {
  Policy net{nullptr};
  std::unique_ptr<torch::optim::Adam> opt;
  std::vector<int64_t> sizes;
  int64_t total = 0;
  Learner(int64_t hidden, int64_t actions, uint64_t seed, double lr) {
    torch::manual_seed(seed);
    net = Policy(hidden, actions);
    opt = std::make_unique<torch::optim::Adam>(
        net->parameters(), torch::optim::AdamOptions(lr).eps(1e-5)); // train.cc:360-362
    for (auto &p : net->parameters()) {
      sizes.push_back(p.numel());
      total += p.numel();
    }
  }
  void set_lr(double lr) {
    static_cast<torch::optim::AdamOptions &>(opt->param_groups()[0].options()).lr(lr);
  }
};

} // namespace

// vision.cc:86-95 calls into the un-vendored stb_image_resize2 (absent offline). The symbol is
// defined here only so the library links; it is NOT an oracle and aborts if ever reached.
extern "C" unsigned char *stbir_resize_uint8_linear(const unsigned char *, int, int, int,
                                                    unsigned char *, int, int, int,
                                                    stbir_pixel_layout) {
  std::abort();
}

extern "C" {

const char *ref_last_error() { return g_error.c_str(); }
void ref_set_num_threads(int n) { at::set_num_threads(n); }
int ref_get_num_threads() { return at::get_num_threads(); }

// ai::gae::gae — src/ai/gae.cc:4-80. Flags are 1-byte bools.
int ref_gae(float *advantages, const float *rewards, const float *values,
            const float *next_values, const uint8_t *terminals,
            const uint8_t *truncations, const uint8_t *episode_starts, int64_t E,
            int64_t T, float gamma, float lambda) {
  return guarded([&] {
    auto adv = torch::zeros({E, T});
    ai::gae::gae(adv, f32(rewards, {E, T}), f32(values, {E, T}), f32(next_values, {E}),
                 b8(terminals, {E, T}), b8(truncations, {E, T}),
                 b8(episode_starts, {E, T}), gamma, lambda);
    out_f32(advantages, adv);
  });
}

// ai::buffer::Buffer add×T + get — src/ai/buffer.cc:41-77. All arrays env-major [E,T,..].
// obs may be null (then a 1-byte observation is used).
int ref_buffer_roundtrip(const uint8_t *obs, const int64_t *actions, const float *rewards,
                         const uint8_t *terminals, const uint8_t *truncations,
                         const uint8_t *episode_starts, const float *logits,
                         const float *values, const float *next_values, int64_t E,
                         int64_t T, int64_t A, float gamma, float lambda, int64_t steps_added,
                         uint8_t *obs_out, float *rewards_out, uint8_t *masks_out,
                         float *advantages_out, float *returns_out) {
  return guarded([&] {
    std::vector<size_t> shape = obs ? std::vector<size_t>{4, 84, 84} : std::vector<size_t>{1};
    std::vector<int64_t> oshape = obs ? std::vector<int64_t>{E, T, 4, 84, 84}
                                      : std::vector<int64_t>{E, T, 1};
    ai::buffer::Buffer buffer(E, T, shape, A, torch::kCPU);
    auto o = obs ? u8(obs, oshape) : torch::zeros(oshape, torch::kUInt8);
    auto a = i64(actions, {E, T});
    auto r = f32(rewards, {E, T});
    auto te = b8(terminals, {E, T}), tr = b8(truncations, {E, T}),
         es = b8(episode_starts, {E, T});
    auto lg = f32(logits, {E, T, A});
    auto v = f32(values, {E, T});
    for (int64_t t = 0; t < steps_added; ++t)
      buffer.add(o.select(1, t), a.select(1, t), r.select(1, t), te.select(1, t),
                 tr.select(1, t), es.select(1, t), lg.select(1, t), v.select(1, t));
    auto batch = buffer.get(f32(next_values, {E}), gamma, lambda);
    out_u8(obs_out, batch.observations);
    out_f32(rewards_out, batch.rewards);
    out_u8(masks_out, batch.masks);
    out_f32(advantages_out, batch.advantages);
    out_f32(returns_out, batch.returns);
  });
}

// 2-frame max + area resize + grayscale (reference libtorch vision path).
int ref_preprocess(const uint8_t *frames_a, const uint8_t *frames_b, int64_t B,
                   float *gray_f32, uint8_t *gray_u8) {
  return guarded([&] {
    auto g = preprocess_gray(u8(frames_a, {B, 210, 160, 3}), u8(frames_b, {B, 210, 160, 3}));
    out_f32(gray_f32, g);
    out_u8(gray_u8, quantise(g));
  });
}

// vision.cc:12-32 grayscale-only resize, [B,F,210,160] f32 -> [B,F,84,84] f32.
int ref_resize_gray_stacked(const float *images, int64_t B, int64_t F, float *out) {
  return guarded([&] {
    out_f32(out, ai::vision::resize_frame_stacked_grayscale_images(f32(images, {B, F, 210, 160})));
  });
}

int ref_stack_update(uint8_t *stack, const uint8_t *frames, const uint8_t *starts, int64_t E) {
  return guarded([&] {
    auto s = u8(stack, {E, 4, 84, 84});
    auto st = b8(starts, {E});
    stack_update(s, u8(frames, {E, 84, 84}), st);
  });
}

// losses::normalize_logits + losses::compute (losses.cc:4-47) with autograd gradients of
// the scalar loss w.r.t. the RAW logits and the values (what mini_batch_update back-props).
int ref_loss(const float *logits, const float *old_log_probs, const int64_t *actions,
             const float *advantages, const float *values, const float *returns,
             const uint8_t *masks, int64_t n, int64_t A, float clip, float vcoef, float ecoef,
             float *loss, float *clipped_losses, float *value_losses, float *entropies,
             float *total_losses, float *ratio, float *dlogits, float *dvalues) {
  return guarded([&] {
    auto lg = f32(logits, {n, A}).clone().requires_grad_(true);
    auto v = f32(values, {n}).clone().requires_grad_(true);
    auto logp = ai::ppo::losses::normalize_logits(lg);
    auto m = ai::ppo::losses::compute(logp, f32(old_log_probs, {n, A}), i64(actions, {n}),
                                      f32(advantages, {n}), v, f32(returns, {n}),
                                      b8(masks, {n}), clip, vcoef, ecoef);
    m.loss.backward();
    out_f32(loss, m.loss.reshape({1}));
    out_f32(clipped_losses, m.clipped_losses);
    out_f32(value_losses, m.value_losses);
    out_f32(entropies, m.entropies);
    out_f32(total_losses, m.total_losses);
    out_f32(ratio, m.ratio);
    out_f32(dlogits, lg.grad());
    out_f32(dvalues, v.grad());
  });
}

int ref_normalize_logits(const float *logits, int64_t n, int64_t A, float *out) {
  return guarded([&] { out_f32(out, ai::ppo::losses::normalize_logits(f32(logits, {n, A}))); });
}

// ---- learner handle --------------------------------------------------------------
void *ref_learner_create(int64_t hidden, int64_t actions, uint64_t seed, double lr) {
  Learner *l = nullptr;
  if (guarded([&] { l = new Learner(hidden, actions, seed, lr); }))
    return nullptr;
  return l;
}
void ref_learner_destroy(void *h) { delete static_cast<Learner *>(h); }
int64_t ref_learner_param_count(void *h) { return static_cast<Learner *>(h)->total; }
int64_t ref_learner_tensor_count(void *h) {
  return static_cast<int64_t>(static_cast<Learner *>(h)->sizes.size());
}
void ref_learner_tensor_sizes(void *h, int64_t *out) {
  auto *l = static_cast<Learner *>(h);
  for (size_t i = 0; i < l->sizes.size(); ++i)
    out[i] = l->sizes[i];
}
// flat parameter vector in `parameters()` order, each tensor in its native (OIHW / [out,in]) order
int ref_learner_get_params(void *h, float *out) {
  return guarded([&] {
    auto *l = static_cast<Learner *>(h);
    int64_t off = 0;
    for (auto &p : l->net->parameters()) {
      out_f32(out + off, p);
      off += p.numel();
    }
  });
}
int ref_learner_set_params(void *h, const float *in) {
  return guarded([&] {
    auto *l = static_cast<Learner *>(h);
    torch::NoGradGuard guard;
    int64_t off = 0;
    for (auto &p : l->net->parameters()) {
      p.copy_(f32(in + off, p.sizes().vec()));
      off += p.numel();
    }
  });
}
int ref_learner_set_lr(void *h, double lr) {
  static_cast<Learner *>(h)->set_lr(lr);
  return 0;
}
// Adam moments (zeros before the first step) and the step counter.
int ref_learner_get_adam(void *h, float *exp_avg, float *exp_avg_sq, int64_t *step) {
  return guarded([&] {
    auto *l = static_cast<Learner *>(h);
    int64_t off = 0;
    int64_t st = 0;
    for (auto &p : l->net->parameters()) {
      auto it = l->opt->state().find(p.unsafeGetTensorImpl());
      if (it == l->opt->state().end()) {
        std::memset(exp_avg + off, 0, sizeof(float) * p.numel());
        std::memset(exp_avg_sq + off, 0, sizeof(float) * p.numel());
      } else {
        auto &s = static_cast<torch::optim::AdamParamState &>(*it->second);
        out_f32(exp_avg + off, s.exp_avg());
        out_f32(exp_avg_sq + off, s.exp_avg_sq());
        st = s.step();
      }
      off += p.numel();
    }
    if (step)
      *step = st;
  });
}
int ref_learner_forward(void *h, const uint8_t *obs, int64_t n, float *logits, float *values) {
  return guarded([&] {
    auto *l = static_cast<Learner *>(h);
    torch::NoGradGuard guard;
    auto o = l->net->forward(u8(obs, {n, 4, 84, 84}));
    out_f32(logits, o.logits);
    out_f32(values, o.value);
  });
}
// Intermediate activations of the forward pass (post-ReLU conv outputs, NCHW; fc output).
int ref_learner_activations(void *h, const uint8_t *obs, int64_t n, float *c1, float *c2,
                            float *c3, float *fc) {
  return guarded([&] {
    auto *l = static_cast<Learner *>(h);
    torch::NoGradGuard guard;
    auto x = u8(obs, {n, 4, 84, 84}).to(torch::kFloat32) / 255.0;
    auto it = l->net->trunk->begin();
    x = (it++)->forward(x); x = (it++)->forward(x); out_f32(c1, x);
    x = (it++)->forward(x); x = (it++)->forward(x); out_f32(c2, x);
    x = (it++)->forward(x); x = (it++)->forward(x); out_f32(c3, x);
    x = (it++)->forward(x); x = (it++)->forward(x); out_f32(fc, x);
  });
}
// forward + loss + backward of ONE minibatch exactly as mini_batch_update does up to
// (not including) clip/step — train.h:118-126. Returns the un-clipped flat gradient.
int ref_learner_grads(void *h, const uint8_t *obs, const int64_t *actions,
                      const float *old_log_probs, const float *advantages,
                      const float *returns, const uint8_t *masks, int64_t n, int64_t A,
                      float clip, float vcoef, float ecoef, float *grads, float *loss,
                      float *logits_out, float *values_out) {
  return guarded([&] {
    auto *l = static_cast<Learner *>(h);
    l->net->train();
    auto o = l->net->forward(u8(obs, {n, 4, 84, 84}));
    auto logp = ai::ppo::losses::normalize_logits(o.logits);
    auto m = ai::ppo::losses::compute(logp, f32(old_log_probs, {n, A}), i64(actions, {n}),
                                      f32(advantages, {n}), o.value, f32(returns, {n}),
                                      b8(masks, {n}), clip, vcoef, ecoef);
    l->opt->zero_grad();
    m.loss.backward();
    int64_t off = 0;
    for (auto &p : l->net->parameters()) {
      out_f32(grads + off, p.grad());
      off += p.numel();
    }
    out_f32(loss, m.loss.reshape({1}));
    out_f32(logits_out, o.logits);
    out_f32(values_out, o.value);
  });
}
// clip_grad_norm_ (src/ai/ppo/train.cc:12-46) + Adam step on caller-provided gradients.
int ref_learner_apply_grads(void *h, const float *grads, double max_norm, float *total_norm) {
  return guarded([&] {
    auto *l = static_cast<Learner *>(h);
    int64_t off = 0;
    for (auto &p : l->net->parameters()) {
      p.mutable_grad() = f32(grads + off, p.sizes().vec()).clone();
      off += p.numel();
    }
    auto tn = ai::ppo::train::clip_grad_norm_(l->net->parameters(), max_norm);
    l->opt->step();
    out_f32(total_norm, tn.reshape({1}));
  });
}
// ai::ppo::train::train — train.h:133-157 — on a flat batch of N rows.
// metrics_* are [epochs,nmb,mb] (clipped_gradients [epochs,nmb]); any may be null.
int ref_learner_train(void *h, const uint8_t *obs, const int64_t *actions,
                      const float *log_probs, const float *advantages, const float *returns,
                      const uint8_t *masks, int64_t N, int64_t A, int64_t epochs, int64_t nmb,
                      float clip, float vcoef, float ecoef, float max_norm, float *m_loss,
                      float *m_clipped, float *m_value, float *m_entropy, float *m_ratio,
                      float *m_total, float *m_adv, float *m_ret, uint8_t *m_masks,
                      float *m_gradnorm) {
  return guarded([&] {
    auto *l = static_cast<Learner *>(h);
    ai::ppo::train::Batch batch{u8(obs, {N, 4, 84, 84}), i64(actions, {N}),
                                f32(log_probs, {N, A}),  f32(advantages, {N}),
                                f32(returns, {N}),       b8(masks, {N})};
    ai::ppo::train::Hyperparameters hp{clip, vcoef, ecoef, max_norm};
    int64_t mb = nmb > 0 && N % nmb == 0 ? N / nmb : 1;
    ai::ppo::train::Metrics metrics(epochs, nmb, mb, torch::kCPU);
    auto indices = torch::empty({N}, torch::kInt64);
    ai::ppo::train::train(l->net, *l->opt, metrics, indices, batch, epochs, nmb, hp);
    out_f32(m_loss, metrics.loss);
    out_f32(m_clipped, metrics.clipped_losses);
    out_f32(m_value, metrics.value_losses);
    out_f32(m_entropy, metrics.entropies);
    out_f32(m_ratio, metrics.ratio);
    out_f32(m_total, metrics.total_losses);
    out_f32(m_adv, metrics.advantages);
    out_f32(m_ret, metrics.returns);
    out_u8(m_masks, metrics.masks);
    out_f32(m_gradnorm, metrics.clipped_gradients);
  });
}

// ---- timed CPU baseline: one full learner-path iteration ----------------------------
// for t<T { 2-frame max -> f32 -> area resize -> gray -> u8 -> stack update -> Buffer::add }
// -> Buffer::get -> prepare_batch -> train().  Frames come from a caller-provided pool
// [pool,2,210,160,3] u8 (step t uses E consecutive pool entries, wrapping). Scalars are
// [E,T] env-major arrays. Returns wall seconds per stage in out_seconds[3] =
// {preprocess+add, Buffer::get, train}. `preprocess_steps` <= T bounds the sampled
// preprocess work (the remaining steps re-use the last stack; its time is scaled by the
// caller) so that a bounded sample of a large workload finishes in seconds.
int ref_iteration(void *h, const uint8_t *frame_pool, int64_t pool, const int64_t *actions,
                  const float *rewards, const uint8_t *terminals, const uint8_t *truncations,
                  const uint8_t *episode_starts, const float *logits, const float *values,
                  const float *next_values, int64_t E, int64_t T, int64_t A, int64_t epochs,
                  int64_t nmb, float gamma, float lambda, float clip, float vcoef, float ecoef,
                  float max_norm, int64_t preprocess_steps, double *out_seconds,
                  float *out_mean_loss) {
  return guarded([&] {
    using clock = std::chrono::steady_clock;
    auto *l = static_cast<Learner *>(h);
    auto secs = [](clock::time_point a, clock::time_point b) {
      return std::chrono::duration<double>(b - a).count();
    };
    ai::buffer::Buffer buffer(E, T, {4, 84, 84}, A, torch::kCPU);
    auto frames = u8(frame_pool, {pool, 2, 210, 160, 3});
    auto a = i64(actions, {E, T});
    auto r = f32(rewards, {E, T});
    auto te = b8(terminals, {E, T}), tr = b8(truncations, {E, T}),
         es = b8(episode_starts, {E, T});
    auto lg = f32(logits, {E, T, A});
    auto v = f32(values, {E, T});
    auto stack = torch::zeros({E, 4, 84, 84}, torch::kUInt8);
    auto idx = torch::arange(E, torch::kInt64);
    auto t0 = clock::now();
    {
      torch::NoGradGuard guard;
      for (int64_t t = 0; t < T; ++t) {
        buffer.add(stack, a.select(1, t), r.select(1, t), te.select(1, t), tr.select(1, t),
                   es.select(1, t), lg.select(1, t), v.select(1, t));
        if (t < preprocess_steps) {
          auto sel = frames.index_select(0, (idx + t * E) % pool);
          auto gray = quantise(preprocess_gray(sel.select(1, 0), sel.select(1, 1)));
          auto st = es.select(1, t).contiguous();
          stack_update(stack, gray, st);
        }
      }
    }
    auto t1 = clock::now();
    ai::buffer::Batch raw;
    {
      torch::NoGradGuard guard;
      raw = buffer.get(f32(next_values, {E}), gamma, lambda);
    }
    auto t2 = clock::now();
    auto batch = flatten_batch(raw);
    ai::ppo::train::Hyperparameters hp{clip, vcoef, ecoef, max_norm};
    ai::ppo::train::Metrics metrics(epochs, nmb, (E * T) / nmb, torch::kCPU);
    auto indices = torch::empty({E * T}, torch::kInt64);
    ai::ppo::train::train(l->net, *l->opt, metrics, indices, batch, epochs, nmb, hp);
    auto t3 = clock::now();
    out_seconds[0] = secs(t0, t1);
    out_seconds[1] = secs(t1, t2);
    out_seconds[2] = secs(t2, t3);
    if (out_mean_loss)
      *out_mean_loss = metrics.loss.mean().item<float>();
  });
}

} // extern "C"
