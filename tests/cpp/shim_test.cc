// Host-side test of include/ai/alegpu_shim.hpp: the reference-named C++ API compiles, links against
// libalegpu.so and throws the exception types the reference throws for the same conditions
// (test/ai/gae-test.cc style checks that need no GPU).  With a GPU (argv[1] == "gpu") it also runs
// the SimpleNoTerminals known-answer case of test/ai/gae-test.cc:6-38 through ai::gae::gae.
#include <cmath>
#include <cstdio>
#include <cstring>

#include "ai/alegpu_shim.hpp"

using namespace ai::alegpu_shim;

#define EXPECT(cond)                                                                     \
  do {                                                                                   \
    if (!(cond)) {                                                                       \
      std::printf("FAILED %s:%d %s\n", __FILE__, __LINE__, #cond);                       \
      return 1;                                                                          \
    }                                                                                    \
  } while (0)

template <class E, class F> bool throws(F f) {
  try {
    f();
  } catch (const E &) {
    return true;
  } catch (...) {
    return false;
  }
  return false;
}

int main(int argc, char **argv) {
  // shape validation mirrors gae.cc:9-46 -> std::invalid_argument, before any device work
  F32 adv(nullptr, {2, 3}), rew(nullptr, {2, 3}), val(nullptr, {2, 3}), next(nullptr, {2});
  U8 term(nullptr, {2, 3}), trunc(nullptr, {2, 3}), start(nullptr, {2, 3});
  F32 rew1d(nullptr, {6}), next_bad(nullptr, {3}), val_bad(nullptr, {2, 4});
  EXPECT(throws<std::invalid_argument>(
      [&] { ai::gae::gae(adv, rew1d, val, next, term, trunc, start, 0.99f, 0.95f); }));
  EXPECT(throws<std::invalid_argument>(
      [&] { ai::gae::gae(adv, rew, val_bad, next, term, trunc, start, 0.99f, 0.95f); }));
  EXPECT(throws<std::invalid_argument>(
      [&] { ai::gae::gae(adv, rew, val, next_bad, term, trunc, start, 0.99f, 0.95f); }));
  EXPECT(throws<std::invalid_argument>([] { ai::buffer::Buffer b(2, 4, {1, 84, 84}, 4); }));
  // train.h:140-143 -> std::runtime_error, checked before the learner is touched
  {
    ai::ppo::train::Batch batch;
    batch.actions = I64(nullptr, {10});
    ai::ppo::train::Metrics met;
    ai::ppo::train::Hyperparameters hp{0.1f, 0.5f, 0.01f, 0.5f};
    I64 indices;
    ai::ppo::train::Learner *none = nullptr;
    EXPECT(throws<std::runtime_error>(
        [&] { ai::ppo::train::train(*none, met, indices, batch, 1, 3, hp); }));
  }
  // status -> exception mapping
  EXPECT(throws<std::invalid_argument>([] { check(ALEGPU_INVALID_ARGUMENT); }));
  EXPECT(throws<std::runtime_error>([] { check(ALEGPU_RUNTIME_ERROR); }));
  EXPECT(throws<std::out_of_range>([] { check(ALEGPU_OUT_OF_RANGE); }));
  EXPECT(throws<std::runtime_error>([] { check(ALEGPU_CUDA_ERROR); }));
  EXPECT(std::strstr(alegpu_version(), "sm_100a") != nullptr);

  if (argc > 1 && std::strcmp(argv[1], "gpu") == 0) {
    // test/ai/gae-test.cc:6-38 SimpleNoTerminalsNoTruncations: r = 1, v = .5, next = .5
    const int E = 1, T = 3;
    float h_r[T] = {1, 1, 1}, h_v[T] = {0.5f, 0.5f, 0.5f}, h_n[E] = {0.5f}, h_a[T] = {0, 0, 0};
    uint8_t h_z[T] = {0, 0, 0};
    float *d_r, *d_v, *d_n, *d_a;
    uint8_t *d_z;
    check(alegpu_malloc((void **)&d_r, sizeof h_r));
    check(alegpu_malloc((void **)&d_v, sizeof h_v));
    check(alegpu_malloc((void **)&d_n, sizeof h_n));
    check(alegpu_malloc((void **)&d_a, sizeof h_a));
    check(alegpu_malloc((void **)&d_z, sizeof h_z));
    check(alegpu_memcpy_h2d(d_r, h_r, sizeof h_r, nullptr));
    check(alegpu_memcpy_h2d(d_v, h_v, sizeof h_v, nullptr));
    check(alegpu_memcpy_h2d(d_n, h_n, sizeof h_n, nullptr));
    check(alegpu_memcpy_h2d(d_z, h_z, sizeof h_z, nullptr));
    F32 a(d_a, {E, T}), r(d_r, {E, T}), v(d_v, {E, T}), n(d_n, {E});
    U8 z(d_z, {E, T});
    ai::gae::gae(a, r, v, n, z, z, z, 0.99f, 0.95f);
    check(alegpu_memcpy_d2h(h_a, d_a, sizeof h_a, nullptr));
    check(alegpu_stream_synchronize(nullptr));
    float want[T], last = 0.0f, nvv = h_n[0]; // the expectation exactly as the reference test forms it
    for (int i = T - 1; i >= 0; --i) {
      const float delta = h_r[i] + 0.99f * nvv - h_v[i];
      want[i] = delta + 0.99f * 0.95f * last;
      last = want[i];
      nvv = h_v[i];
    }
    for (int t = 0; t < T; ++t)
      EXPECT(std::fabs(h_a[t] - want[t]) < 1e-5f);
    std::printf("gpu gae: %.4f %.4f %.4f\n", h_a[0], h_a[1], h_a[2]);
    // ai::vision::resize_grayscale_image (vision.cc:86-95): constant in -> constant out, bad size throws
    {
      std::vector<uint8_t> h_img(210 * 160, 93), h_out(84 * 84, 0);
      uint8_t *d_img, *d_out;
      check(alegpu_malloc((void **)&d_img, h_img.size()));
      check(alegpu_malloc((void **)&d_out, h_out.size()));
      check(alegpu_memcpy_h2d(d_img, h_img.data(), h_img.size(), nullptr));
      U8 img(d_img, {210 * 160}), out(d_out, {84 * 84});
      ai::vision::resize_grayscale_image(img, 160, 210, 84, 84, out);
      check(alegpu_memcpy_d2h(h_out.data(), d_out, h_out.size(), nullptr));
      check(alegpu_stream_synchronize(nullptr));
      bool all = true;
      for (uint8_t px : h_out)
        all = all && px == 93;
      EXPECT(all);
      EXPECT(throws<std::invalid_argument>(
          [&] { ai::vision::resize_grayscale_image(img, 160, 210, 0, 84, out); })); // resize.cc:15-16
      EXPECT(throws<std::invalid_argument>(
          [&] { ai::vision::resize_grayscale_image(img, 161, 210, 84, 84, out); }));
      std::printf("gpu vision: resize ok\n");
    }
  }
  if (argc > 1 && std::strcmp(argv[1], "gpu") == 0) {
    // The reference's call sequence with NO manual synchronisation between the caller's stream (legacy
    // default here, as in the reference) and the learner's own stream: Buffer::add x T, Buffer::get,
    // train(network, optimizer, ...), then a Metrics read on the caller's stream.  Run twice from the
    // same start: identical losses, finite, and the Adam step counter moved — the stream bridge of
    // ai::ppo::train::Learner orders the two streams.
    const int64_t E = 4, T = 8, A = 4, N = E * T, nmb = 2, mb = N / nmb;
    float loss_runs[2][2];
    for (int run = 0; run < 2; ++run) {
      ai::buffer::Buffer buf(E, T, {4, 84, 84}, A);
      std::vector<int64_t> act(E);
      std::vector<float> rew(E), lg(E * A), val(E), nv(E, 0.25f);
      std::vector<uint8_t> z(E, 0), st(E, 0);
      int64_t *d_act; float *d_rew, *d_lg, *d_val, *d_nv; uint8_t *d_z, *d_st;
      check(alegpu_malloc((void **)&d_act, E * 8)); check(alegpu_malloc((void **)&d_rew, E * 4));
      check(alegpu_malloc((void **)&d_lg, E * A * 4)); check(alegpu_malloc((void **)&d_val, E * 4));
      check(alegpu_malloc((void **)&d_nv, E * 4)); check(alegpu_malloc((void **)&d_z, E));
      check(alegpu_malloc((void **)&d_st, E));
      check(alegpu_memcpy_h2d(d_nv, nv.data(), E * 4, nullptr));
      check(alegpu_memcpy_h2d(d_z, z.data(), E, nullptr));
      for (int64_t t = 0; t < T; ++t) {
        for (int64_t e = 0; e < E; ++e) {
          act[e] = (e + t) % A; rew[e] = (float)((e * 7 + t) % 3); val[e] = 0.1f * (float)(e - t);
          st[e] = t == 0; for (int64_t a = 0; a < A; ++a) lg[e * A + a] = 0.01f * (float)(a * (t + 1) - e);
        }
        check(alegpu_memcpy_h2d(d_act, act.data(), E * 8, nullptr));
        check(alegpu_memcpy_h2d(d_rew, rew.data(), E * 4, nullptr));
        check(alegpu_memcpy_h2d(d_lg, lg.data(), E * A * 4, nullptr));
        check(alegpu_memcpy_h2d(d_val, val.data(), E * 4, nullptr));
        check(alegpu_memcpy_h2d(d_st, st.data(), E, nullptr));
        check(alegpu_stream_synchronize(nullptr)); // the host arrays above are reused next step
        buf.add(U8(nullptr, {}), I64(d_act, {E}), F32(d_rew, {E}), U8(d_z, {E}), U8(d_z, {E}),
                U8(d_st, {E}), F32(d_lg, {E, A}), F32(d_val, {E}));
      }
      ai::buffer::Batch rb = buf.get(F32(d_nv, {E}), 0.99f, 0.95f);
      ai::ppo::train::Learner learner(32, A, mb, 2.5e-4, ALEGPU_MATH_FP32);
      std::vector<float> w((size_t)learner.parameter_count());
      for (size_t i = 0; i < w.size(); ++i)
        w[i] = 0.02f * (float)((int)(i * 2654435761u % 201) - 100) / 100.0f;
      learner.load_parameters(w);
      ai::ppo::train::Network network{learner};
      ai::ppo::train::Adam optimizer{learner};
      ai::ppo::train::Batch tb;
      tb.observations = U8(rb.observations.data, {N, 4, 84, 84});
      tb.actions = I64(rb.actions.data, {N});
      tb.log_probabilities = F32(rb.log_probabilities.data, {N, A});
      tb.advantages = F32(rb.advantages.data, {N});
      tb.returns = F32(rb.returns.data, {N});
      tb.masks = U8(rb.masks.data, {N});
      float *d_loss, *d_gn;
      check(alegpu_malloc((void **)&d_loss, nmb * mb * 4)); check(alegpu_malloc((void **)&d_gn, nmb * 4));
      ai::ppo::train::Metrics met;
      met.loss = F32(d_loss, {1, nmb, mb});
      met.clipped_gradients = F32(d_gn, {1, nmb});
      ai::ppo::train::Hyperparameters hp{0.1f, 0.5f, 0.01f, 0.5f};
      I64 indices;
      ai::ppo::train::train(network, optimizer, met, indices, tb, 1, nmb, hp); // train.h:133-136
      std::vector<float> h_loss(nmb * mb);
      check(alegpu_memcpy_d2h(h_loss.data(), d_loss, nmb * mb * 4, nullptr)); // caller's stream, no manual sync
      check(alegpu_stream_synchronize(nullptr));
      loss_runs[run][0] = h_loss[0];
      loss_runs[run][1] = h_loss[mb];
      EXPECT(std::isfinite(h_loss[0]) && std::isfinite(h_loss[mb]) && h_loss[0] != 0.0f);
      int64_t step = 0;
      check(alegpu_learner_get_adam(learner.handle(), nullptr, nullptr, &step));
      EXPECT(step == nmb);
    }
    EXPECT(loss_runs[0][0] == loss_runs[1][0] && loss_runs[0][1] == loss_runs[1][1]);
    std::printf("gpu train via shim: loss %.6f %.6f\n", loss_runs[0][0], loss_runs[0][1]);
  }
  std::printf("shim ok\n");
  return 0;
}
