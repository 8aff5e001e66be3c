// include/ai/alegpu_shim.hpp built with -DALEGPU_WITH_TORCH: the reference's call sites, with real CUDA
// torch::Tensor arguments and the reference's argument lists (src/ai/gae.h:4-7, src/ai/vision.h:6-13,
// src/ai/ppo/losses.h:6-37, src/ai/ppo/train.h:133-136), run through libalegpu.so.  libtorch only
// allocates tensors here; every operator is the CUDA library.  Checked against libtorch's own CPU ops.
#include <cmath>
#include <cstdio>

#include "ai/alegpu_shim.hpp"

#define EXPECT(cond)                                                                     \
  do {                                                                                   \
    if (!(cond)) {                                                                       \
      std::printf("FAILED %s:%d %s\n", __FILE__, __LINE__, #cond);                       \
      return 1;                                                                          \
    }                                                                                    \
  } while (0)

int main() {
  if (!torch::cuda::is_available()) {
    std::printf("no cuda\n");
    return 2;
  }
  torch::manual_seed(7);
  const auto dev = torch::Device(torch::kCUDA, 0);
  const auto f32 = torch::TensorOptions().dtype(torch::kFloat32).device(dev);
  // ---- ai::gae::gae — test/ai/gae-test.cc:6-38 (SimpleNoTerminalsNoTruncations)
  {
    auto rewards = torch::ones({1, 3}, f32), values = torch::full({1, 3}, 0.5f, f32);
    auto next_values = torch::full({1}, 0.5f, f32), advantages = torch::zeros({1, 3}, f32);
    auto flags = torch::zeros({1, 3}, torch::TensorOptions().dtype(torch::kBool).device(dev));
    ai::gae::gae(advantages, rewards, values, next_values, flags, flags, flags, 0.99f, 0.95f);
    torch::cuda::synchronize();
    auto a = advantages.cpu();
    EXPECT(std::fabs(a[0][2].item<float>() - 0.995f) < 1e-5f);
    EXPECT(std::fabs(a[0][0].item<float>() - 2.8109f) < 1e-3f);
  }
  // ---- ai::vision (src/ai/vision.h:6-13) vs libtorch's own interpolate(area) / matmul on the CPU
  {
    namespace F = torch::nn::functional;
    const auto opt = F::InterpolateFuncOptions().size(std::vector<int64_t>({84, 84})).mode(torch::kArea);
    auto rgb = torch::rand({2, 4, 3, 210, 160}, f32) * 255.0f;
    auto small = ai::vision::resize_frame_stacked_rgb_images(rgb);          // [2,4,3,84,84]
    auto gray = ai::vision::rgb_to_grayscale_frame_stacked_images(small);   // [2,4,84,84]
    auto g2 = ai::vision::resize_frame_stacked_grayscale_images(rgb.select(2, 0).contiguous());
    auto one = ai::vision::resize_grayscale_image(rgb[0][0][0].contiguous());
    torch::cuda::synchronize();
    auto want_small = F::interpolate(rgb.cpu().flatten(0, 1), opt).unflatten(0, {2, 4});
    auto w = torch::tensor({0.2125f, 0.7154f, 0.0721f});
    auto want_gray = torch::matmul(want_small.permute({0, 1, 3, 4, 2}), w);
    EXPECT(small.sizes() == want_small.sizes() && gray.sizes() == want_gray.sizes());
    EXPECT((small.cpu() - want_small).abs().max().item<float>() < 1e-3f);
    EXPECT((gray.cpu() - want_gray).abs().max().item<float>() < 1e-3f);
    EXPECT((g2.cpu() - want_small.select(2, 0)).abs().max().item<float>() < 1e-3f);
    EXPECT((one.cpu() - want_small[0][0][0]).abs().max().item<float>() < 1e-3f);
    // test/ai/vision-test.cc:5-30: a constant image stays that constant
    auto c = ai::vision::resize_frame_stacked_grayscale_images(torch::full({1, 4, 210, 160}, 42.0f, f32));
    torch::cuda::synchronize();
    EXPECT((c.cpu() - 42.0f).abs().max().item<float>() == 0.0f);
  }
  // ---- ai::ppo::losses::compute returning Metrics (losses.h:6-37) vs the same expression in libtorch
  {
    const int64_t n = 257, A = 4;
    auto logits = torch::randn({n, A}, f32), old = torch::log_softmax(torch::randn({n, A}, f32), -1);
    auto actions = torch::randint(0, A, {n}, torch::TensorOptions().dtype(torch::kInt64).device(dev));
    auto adv = torch::randn({n}, f32), values = torch::randn({n}, f32), returns = torch::randn({n}, f32);
    auto masks = torch::rand({n}, f32) > 0.1f;
    auto logp = torch::log_softmax(logits, -1);
    auto m = ai::ppo::losses::compute(logp, old, actions, adv, values, returns, masks, 0.1f, 0.5f, 0.01f);
    torch::cuda::synchronize();
    auto lp = logp.cpu(), ol = old.cpu(), ac = actions.cpu().unsqueeze(-1);
    auto ratio = torch::exp(lp.gather(-1, ac).squeeze(-1) - ol.gather(-1, ac).squeeze(-1));
    auto surr = torch::min(ratio * adv.cpu(), torch::clamp(ratio, 0.9f, 1.1f) * adv.cpu());
    auto vl = 0.5f * (values.cpu() - returns.cpu()).pow(2);
    auto ent = -(lp.exp() * lp).sum(-1);
    auto total = -surr + 0.5f * vl - 0.01f * ent;
    auto mk = masks.cpu().to(torch::kFloat32);
    const float loss = ((total * mk).sum() / mk.sum()).item<float>();
    EXPECT(std::fabs(m.loss.cpu().item<float>() - loss) < 1e-5f * std::fabs(loss) + 1e-6f);
    EXPECT((m.ratio.cpu() - ratio).abs().max().item<float>() < 1e-4f);
    EXPECT((m.entropies.cpu() - ent).abs().max().item<float>() < 1e-5f);
    EXPECT((m.total_losses.cpu() - total).abs().max().item<float>() < 1e-4f);
  }
  // ---- ai::ppo::train::train(network, optimizer, metrics, indices, batch, epochs, nmb, hp) — train.h:133-136
  {
    const int64_t N = 64, A = 4, nmb = 2, mb = N / nmb;
    const auto u8 = torch::TensorOptions().dtype(torch::kUInt8).device(dev);
    ai::ppo::train::Learner learner(32, A, mb, 2.5e-4, ALEGPU_MATH_TF32);
    std::vector<float> w((size_t)learner.parameter_count());
    auto wt = torch::randn({(int64_t)w.size()}) * 0.05f;
    std::memcpy(w.data(), wt.data_ptr<float>(), w.size() * sizeof(float));
    learner.load_parameters(w);
    ai::ppo::train::Network network{learner};
    ai::ppo::train::Adam optimizer{learner};
    ai::ppo::train::Batch batch;
    auto obs = torch::randint(0, 256, {N, 4, 84, 84}, u8);
    auto actions = torch::randint(0, A, {N}, torch::TensorOptions().dtype(torch::kInt64).device(dev));
    auto logp = torch::log_softmax(torch::randn({N, A}, f32), -1);
    auto adv = torch::randn({N}, f32), ret = torch::randn({N}, f32);
    auto masks = (torch::rand({N}, f32) > 0.1f).to(torch::kUInt8);
    batch.observations = obs; batch.actions = actions; batch.log_probabilities = logp;
    batch.advantages = adv; batch.returns = ret; batch.masks = masks;
    ai::ppo::train::Metrics metrics;
    auto loss = torch::zeros({1, nmb, mb}, f32), gn = torch::zeros({1, nmb}, f32);
    metrics.loss = loss; metrics.clipped_gradients = gn;
    ai::ppo::train::Hyperparameters hp{0.1f, 0.5f, 0.01f, 0.5f};
    auto indices = torch::zeros({N}, torch::TensorOptions().dtype(torch::kInt64).device(dev));
    ai::alegpu_shim::I64 idx(indices);
    ai::ppo::train::train(network, optimizer, metrics, idx, batch, 1, nmb, hp);
    auto l = loss.cpu(), g = gn.cpu(); // torch's current (legacy default) stream: ordered by the shim's bridge
    EXPECT(std::isfinite(l[0][0][0].item<float>()) && std::isfinite(l[0][1][0].item<float>()));
    EXPECT(g[0][0].item<float>() > 0.0f && g[0][1].item<float>() > 0.0f);
    auto after = learner.parameters();
    double moved = 0.0;
    for (size_t i = 0; i < w.size(); ++i)
      moved += std::fabs((double)after[i] - w[i]);
    EXPECT(moved > 0.0);
  }
  std::printf("shim torch ok\n");
  return 0;
}
