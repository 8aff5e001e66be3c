// collector.cu — the rollout collection loop around the learner: a mirror of
// ai::rollout::Rollout::rollout() (src/ai/rollout.cc:198-278) over the C ABI of this library.
//
//   per step t:  action selection on the current observation stacks        (rollout.cc:207)
//                -> emulator threads step every env with its action          (rollout.cc:211, 280-328)
//                -> rewards / flags of non-reset envs                        (rollout.cc:212-228)
//                -> buffer column t                                          (rollout.cc:231-233)
//                -> next observation stack (frame-stack update)              (rollout.cc:237)
//                -> episode bookkeeping                                      (rollout.cc:239-266)
//   then one more action selection for the bootstrap values and Buffer::get  (rollout.cc:268-270)
//
// The emulator stays on CPU threads behind `alegpu_env_step_fn` (where ALE's step + the last two
// raw frames of the MaxAndSkip window plug in, src/ai/environment/max_and_skip.cc:21-45); only the
// hand-off changes: workers write the two raw RGB frames of each env straight into a PINNED slab
// [E,2,210,160,3] of a [T,...] host ring, the slab goes to the device with one cudaMemcpyAsync on a
// copy stream and is preprocessed (max, resize, gray, stack) into rollout slot t+1 by one kernel;
// rewards and flags travel as four small pinned arrays; the E actions come back in one pinned
// D2H copy instead of E `.item()` syncs (rollout.cc:313).
#include <atomic>
#include <condition_variable>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>

#include "common.cuh"

namespace alegpu {
namespace {

constexpr size_t kFrame = 100800; // 210*160*3

// ---- a Breakout-shaped frame source -------------------------------------------------------------
// Not ALE (the emulator and ROMs are not in this image): a small deterministic brick-breaker with
// ALE's frame geometry (210x160 RGB, walls, 6 brick rows, paddle, ball), Breakout's minimal action
// set {NOOP, FIRE, RIGHT, LEFT}, brick rewards 1/1/4/4/7/7, 5 lives with episodic-life termination
// and a max-steps truncation, so rewards, flags and frames have the reference's statistics shape.
struct BrickEnv {
  uint32_t rng;
  int paddle_x, ball_x, ball_y, vx, vy, lives, steps;
  bool in_play, game_over;
  uint8_t bricks[6][18];
};
struct BrickWorld {
  std::vector<BrickEnv> envs;
  int max_steps, frame_skip;
};
inline uint32_t xorshift(uint32_t &s) {
  s ^= s << 13;
  s ^= s >> 17;
  s ^= s << 5;
  return s;
}
void brick_reset(BrickEnv &e, bool full) {
  if (full) {
    std::memset(e.bricks, 1, sizeof(e.bricks));
    e.lives = 5;
    e.game_over = false;
  }
  e.paddle_x = 72;
  e.ball_x = 20 + (int)(xorshift(e.rng) % 120);
  e.ball_y = 120;
  e.vx = (xorshift(e.rng) & 1) ? 2 : -2;
  e.vy = -3;
  e.in_play = false;
  e.steps = 0;
}
inline void fill_rect(uint8_t *f, int x0, int y0, int w, int h, uint8_t r, uint8_t g, uint8_t b) {
  for (int y = y0; y < y0 + h; ++y) {
    if (y < 0 || y >= 210)
      continue;
    uint8_t *row = f + ((size_t)y * 160) * 3;
    for (int x = x0; x < x0 + w; ++x)
      if (x >= 0 && x < 160) {
        row[x * 3] = r;
        row[x * 3 + 1] = g;
        row[x * 3 + 2] = b;
      }
  }
}
void brick_render(const BrickEnv &e, uint8_t *f) {
  static const uint8_t colors[6][3] = {{200, 72, 72},  {198, 108, 58}, {180, 122, 48},
                                       {162, 162, 42}, {72, 160, 72},  {66, 72, 200}};
  std::memset(f, 0, kFrame);
  fill_rect(f, 0, 17, 160, 15, 142, 142, 142); // top wall
  fill_rect(f, 0, 32, 8, 164, 142, 142, 142);  // side walls
  fill_rect(f, 152, 32, 8, 164, 142, 142, 142);
  for (int r = 0; r < 6; ++r)
    for (int c = 0; c < 18; ++c)
      if (e.bricks[r][c])
        fill_rect(f, 8 + c * 8, 57 + r * 6, 8, 6, colors[r][0], colors[r][1], colors[r][2]);
  fill_rect(f, e.paddle_x, 189, 16, 4, 200, 72, 72);
  fill_rect(f, e.ball_x, e.ball_y, 2, 4, 200, 72, 72);
  for (int l = 0; l < e.lives; ++l) // lives indicator
    fill_rect(f, 100 + l * 6, 5, 4, 10, 142, 142, 142);
}
// one emulator frame; returns the reward collected
float brick_tick(BrickEnv &e, int action, bool *life_lost) {
  static const float row_reward[6] = {7, 7, 4, 4, 1, 1};
  if (action == 2)
    e.paddle_x = e.paddle_x + 4 > 136 ? 136 : e.paddle_x + 4;
  if (action == 3)
    e.paddle_x = e.paddle_x - 4 < 8 ? 8 : e.paddle_x - 4;
  if (action == 1)
    e.in_play = true;
  float reward = 0.0f;
  if (!e.in_play)
    return reward;
  e.ball_x += e.vx;
  e.ball_y += e.vy;
  if (e.ball_x <= 8) { e.ball_x = 8; e.vx = -e.vx; }
  if (e.ball_x >= 150) { e.ball_x = 150; e.vx = -e.vx; }
  if (e.ball_y <= 32) { e.ball_y = 32; e.vy = -e.vy; }
  if (e.ball_y >= 57 && e.ball_y < 93) { // brick field
    const int r = (e.ball_y - 57) / 6, c = (e.ball_x - 8) / 8;
    if (c >= 0 && c < 18 && e.bricks[r][c]) {
      e.bricks[r][c] = 0;
      e.vy = -e.vy;
      reward += row_reward[r];
    }
  }
  if (e.ball_y >= 185 && e.ball_y <= 189 && e.vy > 0 && e.ball_x + 2 >= e.paddle_x &&
      e.ball_x <= e.paddle_x + 16) {
    e.vy = -e.vy;
    e.vx = (e.ball_x - e.paddle_x < 8) ? -2 : 2;
  }
  if (e.ball_y > 200) {
    *life_lost = true;
    e.in_play = false;
    e.ball_y = 120;
    e.vy = -3;
  }
  return reward;
}
void brick_step(void *user, int64_t env, int64_t action, int reset, uint8_t *frame_a,
                uint8_t *frame_b, float *reward, uint8_t *terminated, uint8_t *truncated,
                uint8_t *game_over) {
  BrickWorld &w = *static_cast<BrickWorld *>(user);
  BrickEnv &e = w.envs[(size_t)env];
  *reward = 0.0f;
  *terminated = *truncated = *game_over = 0;
  if (reset) { // environment->reset(): a new game after a game over, else the next life
    brick_reset(e, e.game_over || e.lives <= 0);
    brick_render(e, frame_a);
    std::memcpy(frame_b, frame_a, kFrame);
    return;
  }
  bool life_lost = false;
  for (int k = 0; k < w.frame_skip; ++k) { // MaxAndSkip: act frame_skip times, keep the last two frames
    *reward += brick_tick(e, (int)action, &life_lost);
    if (k == w.frame_skip - 2)
      brick_render(e, frame_a);
    if (k == w.frame_skip - 1)
      brick_render(e, frame_b);
    if (life_lost)
      break;
  }
  if (life_lost) { // the window ended early: both frames are the last one (max_and_skip.cc:40-42)
    e.lives -= 1;
    brick_render(e, frame_b);
    std::memcpy(frame_a, frame_b, kFrame);
    *terminated = 1; // episodic life
    if (e.lives <= 0) {
      e.game_over = true;
      *game_over = 1;
    }
  }
  bool cleared = true;
  for (int r = 0; r < 6 && cleared; ++r)
    for (int c = 0; c < 18; ++c)
      if (e.bricks[r][c]) {
        cleared = false;
        break;
      }
  if (cleared && !*terminated) {
    *terminated = 1;
    e.game_over = true;
    *game_over = 1;
  }
  e.steps += 1;
  if (!*terminated && e.steps >= w.max_steps)
    *truncated = 1; // never together with terminated (gae.cc:49-53)
}

// column t of the [E,T] reward / flag arrays <- this step's per-env values (Buffer::add's part
// that the action-selection kernel has not already written)
__global__ void put_step_scalars_kernel(int64_t E, int64_t T, int64_t t,
                                        const float *__restrict__ rewards,
                                        const uint8_t *__restrict__ terminals,
                                        const uint8_t *__restrict__ truncations,
                                        const uint8_t *__restrict__ starts,
                                        float *__restrict__ o_rewards,
                                        uint8_t *__restrict__ o_terminals,
                                        uint8_t *__restrict__ o_truncations,
                                        uint8_t *__restrict__ o_starts) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E)
    return;
  const int64_t o = e * T + t;
  o_rewards[o] = rewards[e];
  o_terminals[o] = terminals[e];
  o_truncations[o] = truncations[e];
  o_starts[o] = starts[e];
}

} // namespace
} // namespace alegpu

using namespace alegpu;

// ---- the collector --------------------------------------------------------------------------------
struct alegpu_collector {
  int64_t E = 0, T = 0, A = 0;
  int device = 0, workers = 1;
  alegpu_env_step_fn step_fn = nullptr;
  void *user = nullptr;
  BrickWorld *owned_world = nullptr;
  uint64_t seed = 0, act_calls = 0;
  // pinned host: frames [T,E,2,210,160,3], per-step scalars, actions
  uint8_t *frames = nullptr;
  float *h_rewards = nullptr;
  uint8_t *h_term = nullptr, *h_trunc = nullptr, *h_start = nullptr;
  int64_t *h_actions = nullptr;
  // per-step snapshots of the scalar arrays (the async H2D reads host memory when it EXECUTES, the
  // bookkeeping below edits the live arrays right after enqueueing it)
  float *h_stage_r[2] = {};
  uint8_t *h_stage_f[2] = {}; // 3*E: terminals | truncations | starts
  // device: per-step scalar staging (double buffered), dense actions, frame slab ring
  float *d_rewards[2] = {};
  uint8_t *d_flags[2] = {}; // 3*E: terminals | truncations | starts
  int64_t *d_actions = nullptr;
  float *d_next_values = nullptr;
  uint8_t *ring = nullptr;
  static constexpr int kRing = 3;
  cudaEvent_t copied[kRing] = {}, consumed[kRing] = {};
  cudaStream_t copy_stream = nullptr;
  // CPU-side episode state (Rollout's *_cpu_ members, rollout.cc:95-108)
  std::vector<uint8_t> start_cpu, game_over;
  std::vector<float> ep_return, game_return;
  std::vector<int64_t> ep_length, game_length;
  int carry_index = 0;
  int64_t total_steps = 0, episodes = 0;
  // worker pool
  std::vector<std::thread> pool;
  std::mutex mu;
  std::condition_variable cv_go, cv_done;
  int64_t generation = 0, pending = 0;
  int64_t cur_t = 0;
  bool stop = false;
};

namespace {
void step_range(alegpu_collector *c, int64_t t, int64_t e0, int64_t e1) {
  uint8_t *slab = c->frames + (size_t)t * c->E * 2 * kFrame;
  for (int64_t e = e0; e < e1; ++e) {
    float r = 0.0f;
    uint8_t term = 0, trunc = 0, over = 0;
    const int reset = c->start_cpu[(size_t)e] ? 1 : 0; // rollout.cc:299-305
    c->step_fn(c->user, e, c->h_actions[e], reset, slab + (size_t)e * 2 * kFrame,
               slab + (size_t)e * 2 * kFrame + kFrame, &r, &term, &trunc, &over);
    if (!reset) { // rollout.cc:213-227: a reset step leaves reward / flags / counters untouched
      c->h_rewards[e] = r;
      c->h_term[e] = term;
      c->h_trunc[e] = trunc;
      c->game_over[(size_t)e] = over;
      c->ep_return[(size_t)e] += r;
      c->ep_length[(size_t)e] += 1;
      c->game_return[(size_t)e] += r;
      c->game_length[(size_t)e] += 1;
    }
  }
}
void worker_main(alegpu_collector *c, int w) {
  int64_t seen = 0;
  for (;;) {
    int64_t t;
    {
      std::unique_lock<std::mutex> lk(c->mu);
      c->cv_go.wait(lk, [&] { return c->stop || c->generation != seen; });
      if (c->stop)
        return;
      seen = c->generation;
      t = c->cur_t;
    }
    const int64_t per = (c->E + c->workers - 1) / c->workers;
    const int64_t e0 = std::min<int64_t>(c->E, (int64_t)w * per), e1 = std::min<int64_t>(c->E, e0 + per);
    step_range(c, t, e0, e1);
    {
      std::lock_guard<std::mutex> lk(c->mu);
      if (--c->pending == 0)
        c->cv_done.notify_one();
    }
  }
}
void step_all(alegpu_collector *c, int64_t t) { // Rollout::step_all, rollout.cc:280-287
  if (c->workers <= 1) {
    step_range(c, t, 0, c->E);
    return;
  }
  {
    std::lock_guard<std::mutex> lk(c->mu);
    c->cur_t = t;
    c->pending = c->workers;
    c->generation += 1;
  }
  c->cv_go.notify_all();
  std::unique_lock<std::mutex> lk(c->mu);
  c->cv_done.wait(lk, [&] { return c->pending == 0; });
}
template <class T> int pinned(T **p, size_t n) {
  ALEGPU_CUDA_CHECK(cudaHostAlloc((void **)p, sizeof(T) * (n ? n : 1), cudaHostAllocDefault));
  std::memset(*p, 0, sizeof(T) * (n ? n : 1));
  return ALEGPU_OK;
}
} // namespace

extern "C" int alegpu_collector_destroy(alegpu_collector *c) {
  if (!c)
    return ALEGPU_OK;
  {
    std::lock_guard<std::mutex> lk(c->mu);
    c->stop = true;
  }
  c->cv_go.notify_all();
  for (auto &th : c->pool)
    th.join();
  cudaSetDevice(c->device);
  if (c->copy_stream) {
    cudaStreamSynchronize(c->copy_stream);
    cudaStreamDestroy(c->copy_stream);
  }
  for (int i = 0; i < alegpu_collector::kRing; ++i) {
    if (c->copied[i]) cudaEventDestroy(c->copied[i]);
    if (c->consumed[i]) cudaEventDestroy(c->consumed[i]);
  }
  for (void *p : {(void *)c->frames, (void *)c->h_rewards, (void *)c->h_term, (void *)c->h_trunc,
                  (void *)c->h_start, (void *)c->h_actions, (void *)c->h_stage_r[0],
                  (void *)c->h_stage_r[1], (void *)c->h_stage_f[0], (void *)c->h_stage_f[1]})
    if (p)
      cudaFreeHost(p);
  for (void *p : {(void *)c->d_rewards[0], (void *)c->d_rewards[1], (void *)c->d_flags[0],
                  (void *)c->d_flags[1], (void *)c->d_actions, (void *)c->d_next_values,
                  (void *)c->ring})
    if (p)
      cudaFree(p);
  delete c->owned_world;
  delete c;
  return ALEGPU_OK;
}

extern "C" int alegpu_collector_create(int64_t E, int64_t T, int64_t A, int num_workers,
                                       alegpu_env_step_fn step_fn, void *user, uint64_t seed,
                                       int device, alegpu_collector **out) {
  ALEGPU_REQUIRE(out && E >= 1 && T >= 1 && A >= 1 && num_workers >= 1, ALEGPU_INVALID_ARGUMENT,
                 "collector_create: bad arguments"); // rollout.cc:46-63
  const size_t slab = (size_t)E * 2 * kFrame;
  ALEGPU_REQUIRE(slab * (size_t)T <= ((size_t)16 << 30), ALEGPU_INVALID_ARGUMENT,
                 "collector_create: the pinned frame ring [T,E,2,210,160,3] would exceed 16 GiB");
  ALEGPU_CUDA_CHECK(cudaSetDevice(device));
  auto *c = new alegpu_collector();
  c->E = E; c->T = T; c->A = A; c->device = device; c->seed = seed;
  c->workers = (int)std::min<int64_t>(num_workers, E);
  if (step_fn) {
    c->step_fn = step_fn;
    c->user = user;
  } else { // the built-in Breakout-shaped frame source (A must cover its 4 actions)
    c->owned_world = new BrickWorld();
    c->owned_world->max_steps = 27000; // configs' max_steps order of magnitude
    c->owned_world->frame_skip = 4;    // configs/*.yaml frame_skip
    c->owned_world->envs.resize((size_t)E);
    for (int64_t e = 0; e < E; ++e) {
      BrickEnv &b = c->owned_world->envs[(size_t)e];
      b.rng = (uint32_t)(seed * 2654435761u + (uint64_t)e * 40503u + 1u) | 1u;
      b.game_over = true;
      b.lives = 0;
    }
    c->step_fn = brick_step;
    c->user = c->owned_world;
  }
  int rc = 0;
  if ((rc = pinned(&c->frames, slab * (size_t)T)) || (rc = pinned(&c->h_rewards, (size_t)E)) ||
      (rc = pinned(&c->h_term, (size_t)E)) || (rc = pinned(&c->h_trunc, (size_t)E)) ||
      (rc = pinned(&c->h_start, (size_t)E)) || (rc = pinned(&c->h_actions, (size_t)E)) ||
      (rc = pinned(&c->h_stage_r[0], (size_t)E)) || (rc = pinned(&c->h_stage_r[1], (size_t)E)) ||
      (rc = pinned(&c->h_stage_f[0], 3 * (size_t)E)) || (rc = pinned(&c->h_stage_f[1], 3 * (size_t)E))) {
    alegpu_collector_destroy(c);
    return rc;
  }
  cudaError_t ce = cudaSuccess;
  for (int i = 0; i < 2 && ce == cudaSuccess; ++i) {
    ce = cudaMalloc(&c->d_rewards[i], sizeof(float) * (size_t)E);
    if (ce == cudaSuccess)
      ce = cudaMalloc(&c->d_flags[i], 3 * (size_t)E);
  }
  if (ce == cudaSuccess) ce = cudaMalloc(&c->d_actions, sizeof(int64_t) * (size_t)E);
  if (ce == cudaSuccess) ce = cudaMalloc(&c->d_next_values, sizeof(float) * (size_t)E);
  if (ce == cudaSuccess) ce = cudaMalloc(&c->ring, slab * alegpu_collector::kRing);
  if (ce == cudaSuccess) ce = cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking);
  for (int i = 0; i < alegpu_collector::kRing && ce == cudaSuccess; ++i) {
    ce = cudaEventCreateWithFlags(&c->copied[i], cudaEventDisableTiming);
    if (ce == cudaSuccess)
      ce = cudaEventCreateWithFlags(&c->consumed[i], cudaEventDisableTiming);
  }
  if (ce != cudaSuccess) {
    alegpu_collector_destroy(c);
    return fail(ALEGPU_CUDA_ERROR, std::string("collector_create: ") + cudaGetErrorString(ce));
  }
  c->start_cpu.assign((size_t)E, 1); // rollout.cc:101: every env starts with a reset
  c->game_over.assign((size_t)E, 0);
  c->ep_return.assign((size_t)E, 0.0f);
  c->game_return.assign((size_t)E, 0.0f);
  c->ep_length.assign((size_t)E, 0);
  c->game_length.assign((size_t)E, 0);
  for (int64_t e = 0; e < E; ++e)
    c->h_start[e] = 1;
  if (c->workers > 1)
    for (int w = 0; w < c->workers; ++w)
      c->pool.emplace_back(worker_main, c, w);
  *out = c;
  return ALEGPU_OK;
}

// Rollout::rollout(), src/ai/rollout.cc:198-278.  On return (asynchronous on the learner's stream
// except for the per-step action read-back) the rollout storage is full, *flat holds the [N] views
// train() consumes and *log the episode statistics of this rollout.
extern "C" int alegpu_collector_rollout(alegpu_collector *c, alegpu_rollout *rollout,
                                        alegpu_learner *learner, float gamma, float lambda,
                                        alegpu_batch *flat, alegpu_rollout_log *log) {
  ALEGPU_REQUIRE(c && rollout && learner, ALEGPU_INVALID_ARGUMENT, "collector_rollout: null argument");
  alegpu_rollout_view v;
  if (int rc = alegpu_rollout_get_view(rollout, &v))
    return rc;
  ALEGPU_REQUIRE(v.E == c->E && v.T == c->T && v.A == c->A, ALEGPU_INVALID_ARGUMENT,
                 "collector_rollout: the rollout storage has a different shape");
  ALEGPU_CUDA_CHECK(cudaSetDevice(c->device));
  void *stream = alegpu_learner_stream(learner);
  cudaStream_t s = as_stream(stream);
  const int64_t E = c->E, T = c->T, A = c->A;
  const size_t slab = (size_t)E * 2 * kFrame;
  double ret_sum = 0.0, len_sum = 0.0, gret_sum = 0.0, glen_sum = 0.0;
  int64_t finished = 0, finished_games = 0, steps_inc = 0;
  uint8_t *carry[2] = {v.carry, v.carry + (size_t)E * 28224};
  int cur = c->carry_index; // which carry stack holds the current observation (Rollout::observations_)
  if (int rc = alegpu_rollout_set_write_index(rollout, 0))
    return rc;
  for (int64_t t = 0; t < T; ++t) {
    // observation at column t = the current stack (buffer_.add's observations_, rollout.cc:231)
    if (t == 0) // slot 0 <- the carried stack; later slots were written by the previous step
      ALEGPU_CUDA_CHECK(cudaMemcpy2DAsync(v.observations, (size_t)T * 28224, carry[cur], 28224, 28224,
                                          (size_t)E, cudaMemcpyDeviceToDevice, s));
    // ---- action selection (rollout.cc:207) on the current stacks.  For t > 0 they live in slot t
    // of the storage with an env stride of T*28224; the policy call wants them contiguous, so the
    // preprocess of step t-1 ALSO wrote them to the other carry stack (see below).
    if (int rc = alegpu_learner_act(learner, carry[cur], E, nullptr, c->seed, c->act_calls++,
                                    v.logits + t * A, T * A, v.values + t, T, v.actions + t, T,
                                    c->d_actions, c->h_actions))
      return rc;
    ALEGPU_CUDA_CHECK(cudaStreamSynchronize(s)); // the emulator threads need the actions
    for (int64_t e = 0; e < E; ++e)
      if (!c->start_cpu[(size_t)e] && (c->h_actions[e] < 0 || c->h_actions[e] >= A))
        return fail(ALEGPU_OUT_OF_RANGE, "Action index out of range for environment " +
                                             std::to_string(e)); // rollout.cc:314-316
    // ---- step all environments on the worker threads (rollout.cc:211)
    for (int64_t e = 0; e < E; ++e)
      c->h_start[e] = c->start_cpu[(size_t)e];
    step_all(c, t);
    for (int64_t e = 0; e < E; ++e)
      if (!c->start_cpu[(size_t)e])
        steps_inc += 1;
    // ---- hand-off: frames slab -> ring slot (copy stream), scalars -> column t
    const int slot = (int)(t % alegpu_collector::kRing), sb = (int)(t & 1);
    uint8_t *dev = c->ring + slab * slot;
    if (t >= alegpu_collector::kRing)
      ALEGPU_CUDA_CHECK(cudaStreamWaitEvent(c->copy_stream, c->consumed[slot], 0));
    ALEGPU_CUDA_CHECK(cudaMemcpyAsync(dev, c->frames + slab * (size_t)t, slab,
                                      cudaMemcpyHostToDevice, c->copy_stream));
    ALEGPU_CUDA_CHECK(cudaEventRecord(c->copied[slot], c->copy_stream));
    std::memcpy(c->h_stage_r[sb], c->h_rewards, sizeof(float) * (size_t)E);
    std::memcpy(c->h_stage_f[sb], c->h_term, (size_t)E);
    std::memcpy(c->h_stage_f[sb] + E, c->h_trunc, (size_t)E);
    std::memcpy(c->h_stage_f[sb] + 2 * E, c->h_start, (size_t)E);
    ALEGPU_CUDA_CHECK(cudaMemcpyAsync(c->d_rewards[sb], c->h_stage_r[sb], sizeof(float) * (size_t)E,
                                      cudaMemcpyHostToDevice, s));
    ALEGPU_CUDA_CHECK(cudaMemcpyAsync(c->d_flags[sb], c->h_stage_f[sb], 3 * (size_t)E,
                                      cudaMemcpyHostToDevice, s));
    put_step_scalars_kernel<<<(unsigned)ceil_div(E, 128), 128, 0, s>>>(
        E, T, t, c->d_rewards[sb], c->d_flags[sb], c->d_flags[sb] + E, c->d_flags[sb] + 2 * E,
        v.rewards, v.terminals, v.truncations, v.episode_starts);
    ALEGPU_COUNT_LAUNCH();
    ALEGPU_LAUNCH_CHECK();
    // ---- update_observations (rollout.cc:184-196, 237): new stack -> the other carry (contiguous,
    // what the next policy call reads) and, for t+1 < T, slot t+1 of the storage
    ALEGPU_CUDA_CHECK(cudaStreamWaitEvent(s, c->copied[slot], 0));
    if (int rc = alegpu_preprocess_step_strided(dev, dev + kFrame, 2 * kFrame, v.episode_starts + t, T,
                                                carry[cur], 28224, carry[1 - cur], 28224, E, stream))
      return rc;
    ALEGPU_CUDA_CHECK(cudaEventRecord(c->consumed[slot], s));
    if (t + 1 < T)
      ALEGPU_CUDA_CHECK(cudaMemcpy2DAsync(v.observations + (size_t)(t + 1) * 28224, (size_t)T * 28224,
                                          carry[1 - cur], 28224, 28224, (size_t)E,
                                          cudaMemcpyDeviceToDevice, s));
    cur = 1 - cur;
    // ---- episode bookkeeping (rollout.cc:239-266)
    for (int64_t e = 0; e < E; ++e) {
      const size_t i = (size_t)e;
      if (!c->h_start[e] && (c->h_term[e] || c->h_trunc[e])) {
        c->start_cpu[i] = 1;
        c->h_term[e] = 0; // is_terminated_/is_truncated_ are cleared for the coming reset step
        c->h_trunc[e] = 0;
        c->episodes += 1;
        ret_sum += c->ep_return[i];
        len_sum += (double)c->ep_length[i];
        finished += 1;
        c->ep_return[i] = 0.0f;
        c->ep_length[i] = 0;
        if (c->game_over[i]) {
          gret_sum += c->game_return[i];
          glen_sum += (double)c->game_length[i];
          finished_games += 1;
          c->game_return[i] = 0.0f;
          c->game_length[i] = 0;
        }
      } else if (c->start_cpu[i]) {
        c->start_cpu[i] = 0;
      }
    }
  }
  c->total_steps += steps_inc;
  c->carry_index = cur;
  // ---- bootstrap values of the final observation, then Buffer::get (rollout.cc:268-270)
  if (int rc = alegpu_learner_act(learner, carry[cur], E, nullptr, c->seed, c->act_calls++, nullptr,
                                  0, c->d_next_values, 1, nullptr, 0, nullptr, nullptr))
    return rc;
  if (int rc = alegpu_rollout_set_write_index(rollout, 0))
    return rc;
  if (int rc = alegpu_rollout_get(rollout, c->d_next_values, gamma, lambda, flat, 0, stream))
    return rc;
  if (log) {
    log->steps = c->total_steps;
    log->episodes = c->episodes;
    log->finished_episodes = finished;
    log->finished_games = finished_games;
    log->mean_episode_return = finished ? (float)(ret_sum / (double)finished) : 0.0f;
    log->mean_episode_length = finished ? (float)(len_sum / (double)finished) : 0.0f;
    log->mean_game_return = finished_games ? (float)(gret_sum / (double)finished_games) : 0.0f;
    log->mean_game_length = finished_games ? (float)(glen_sum / (double)finished_games) : 0.0f;
  }
  return ALEGPU_OK;
}

// the pinned frame ring [T,E,2,210,160,3] and this step's scalar arrays (parity tests replay them)
extern "C" const uint8_t *alegpu_collector_frames(const alegpu_collector *c) {
  return c ? c->frames : nullptr;
}
