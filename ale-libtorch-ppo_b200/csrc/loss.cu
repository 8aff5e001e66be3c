// loss.cu — stateless PPO loss operators (mirror of src/ai/ppo/losses.{h,cc}).
//
// Replaces ~25 ATen elementwise/reduction kernels forward + their autograd backward with one
// fused kernel: log-softmax, action gather, ratio, clamp, min, value loss, entropy, masked
// sum, per-sample metrics, dlogits, dvalues.  102 algorithmic bytes per sample.
// Reductions are deterministic: warp shuffles, a fixed tree per block, block partials summed
// in index order by one warp.
#include <mutex>

#include "loss.cuh"

namespace alegpu {
namespace {

constexpr int kThreads = 256;

__global__ void normalize_logits_kernel(const float *__restrict__ z, float *__restrict__ out,
                                        int64_t n, int A) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n)
    return;
  const float *zi = z + i * A;
  float mx = zi[0];
  for (int j = 1; j < A; ++j)
    mx = fmaxf(mx, zi[j]);
  float se = 0.0f;
  for (int j = 0; j < A; ++j)
    se += expf(zi[j] - mx);
  const float lse = mx + logf(se);
  for (int j = 0; j < A; ++j)
    out[i * A + j] = zi[j] - lse;
}

// count[b] = sum(masks[b*n .. (b+1)*n)) as float (exact up to 2^24 rows per minibatch)
__global__ void __launch_bounds__(kThreads)
    mask_count_kernel(const uint8_t *__restrict__ masks, int64_t n, float *__restrict__ count) {
  __shared__ float red[kThreads / 32];
  const uint8_t *mk = masks + (int64_t)blockIdx.x * n;
  float c = 0.0f;
  for (int64_t i = threadIdx.x; i < n; i += kThreads)
    c += mk[i] ? 1.0f : 0.0f;
  c = block_sum<float, kThreads>(c, red);
  if (threadIdx.x == 0)
    count[blockIdx.x] = c;
}

template <int kMaxA>
__global__ void __launch_bounds__(kThreads)
    loss_kernel(const float *__restrict__ logits, const float *__restrict__ old_lp,
                const int64_t *__restrict__ actions, const float *__restrict__ adv,
                const float *__restrict__ values, const float *__restrict__ returns,
                const uint8_t *__restrict__ masks, int64_t n, int A, float clip, float vcoef,
                float ecoef, const float *__restrict__ count, float *__restrict__ partials,
                float *__restrict__ o_clipped, float *__restrict__ o_value,
                float *__restrict__ o_entropy, float *__restrict__ o_total,
                float *__restrict__ o_ratio, float *__restrict__ dlogits,
                float *__restrict__ dvalues) {
  __shared__ float red[kThreads / 32];
  const int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x;
  float contrib = 0.0f;
  if (i < n) {
    const float inv = 1.0f / count[0];
    float z[kMaxA], ol[kMaxA], dz[kMaxA];
#pragma unroll
    for (int j = 0; j < kMaxA; ++j)
      if (j < A) {
        z[j] = logits[i * A + j];
        ol[j] = old_lp[i * A + j];
      }
    const bool m = masks[i] != 0;
    float dv;
    const PpoSample s =
        ppo_sample<kMaxA>(z, ol, A, (int)actions[i], adv[i], values[i], returns[i], clip, vcoef,
                          ecoef, m ? inv : 0.0f, dz, &dv);
    if (o_clipped) o_clipped[i] = s.surr;
    if (o_value) o_value[i] = s.value_loss;
    if (o_entropy) o_entropy[i] = s.entropy;
    if (o_total) o_total[i] = s.total;
    if (o_ratio) o_ratio[i] = s.ratio;
    if (dlogits) {
#pragma unroll
      for (int j = 0; j < kMaxA; ++j)
        if (j < A)
          dlogits[i * A + j] = dz[j];
    }
    if (dvalues) dvalues[i] = dv;
    contrib = m ? s.total : 0.0f;
  }
  contrib = block_sum<float, kThreads>(contrib, red);
  if (threadIdx.x == 0)
    partials[blockIdx.x] = contrib;
}

// loss = (sum of block partials, index order) / count   — one warp, fixed order
__global__ void loss_finalize_kernel(const float *__restrict__ partials, int nparts,
                                     const float *__restrict__ count, float *__restrict__ loss) {
  float s = 0.0f;
  for (int i = threadIdx.x; i < nparts; i += 32)
    s += partials[i];
  s = warp_sum(s);
  if (threadIdx.x == 0)
    loss[0] = s / count[0];
}

// scratch of the stateless alegpu_loss (count, loss, per-block partials): one per DEVICE, grown under a
// lock.  alegpu_loss is not re-entrant per device (two streams would share count / partials) and may
// allocate, so it must not be called inside a stream capture — documented in alegpu.h; the learner's own
// loss path (heads.cu) uses per-learner buffers and has neither restriction.
struct Scratch {
  float *buf = nullptr;
  size_t cap = 0;
};
std::mutex g_scratch_mutex;
Scratch g_scratches[64];
thread_local Scratch g_scratch; // the current device's scratch after ensure_scratch()

int ensure_scratch(size_t floats) {
  int dev = 0;
  ALEGPU_CUDA_CHECK(cudaGetDevice(&dev));
  ALEGPU_REQUIRE(dev >= 0 && dev < 64, ALEGPU_RUNTIME_ERROR, "loss: device ordinal out of range");
  std::lock_guard<std::mutex> lock(g_scratch_mutex);
  Scratch &sc = g_scratches[dev];
  if (sc.cap < floats) {
    if (sc.buf)
      cudaFree(sc.buf);
    sc.buf = nullptr;
    sc.cap = 0;
    ALEGPU_CUDA_CHECK(cudaMalloc(&sc.buf, floats * sizeof(float)));
    sc.cap = floats;
  }
  g_scratch = sc;
  return ALEGPU_OK;
}

} // namespace

// shared with learner.cu ------------------------------------------------------------------
int launch_mask_count(const uint8_t *masks, int64_t n, int64_t groups, float *count,
                      cudaStream_t s) {
  mask_count_kernel<<<(unsigned)groups, kThreads, 0, s>>>(masks, n, count);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}

int launch_normalize_logits(const float *logits, int64_t n, int64_t A, float *out,
                            cudaStream_t s) {
  if (n == 0)
    return ALEGPU_OK;
  normalize_logits_kernel<<<(unsigned)ceil_div(n, 256), 256, 0, s>>>(logits, out, n, (int)A);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}

} // namespace alegpu

using namespace alegpu;

extern "C" int alegpu_normalize_logits(const float *logits, int64_t n, int64_t A, float *out,
                                       void *stream) {
  ALEGPU_REQUIRE(n >= 0 && A >= 1, ALEGPU_INVALID_ARGUMENT, "normalize_logits: bad shape");
  ALEGPU_REQUIRE(n == 0 || (logits && out), ALEGPU_INVALID_ARGUMENT,
                 "normalize_logits: null pointer");
  return launch_normalize_logits(logits, n, A, out, as_stream(stream));
}

extern "C" int alegpu_loss(const float *logits, const float *old_log_probs,
                           const int64_t *actions, const float *advantages, const float *values,
                           const float *returns, const uint8_t *masks, int64_t n, int64_t A,
                           float clip_param, float value_loss_coef, float entropy_coef,
                           float *loss, float *clipped_losses, float *value_losses,
                           float *entropies, float *total_losses, float *ratio, float *dlogits,
                           float *dvalues, void *stream) {
  ALEGPU_REQUIRE(n >= 1 && A >= 1 && A <= 32, ALEGPU_INVALID_ARGUMENT,
                 "loss: need n >= 1 and 1 <= A <= 32");
  ALEGPU_REQUIRE(logits && old_log_probs && actions && advantages && values && returns && masks,
                 ALEGPU_INVALID_ARGUMENT, "loss: null input pointer");
  cudaStream_t s = as_stream(stream);
  const int blocks = (int)ceil_div(n, kThreads);
  if (int rc = ensure_scratch((size_t)blocks + 2))
    return rc;
  float *count = g_scratch.buf, *lossbuf = g_scratch.buf + 1, *partials = g_scratch.buf + 2;
  if (int rc = launch_mask_count(masks, n, 1, count, s))
    return rc;
  if (A <= 4)
    loss_kernel<4><<<blocks, kThreads, 0, s>>>(logits, old_log_probs, actions, advantages,
                                               values, returns, masks, n, (int)A, clip_param,
                                               value_loss_coef, entropy_coef, count, partials,
                                               clipped_losses, value_losses, entropies,
                                               total_losses, ratio, dlogits, dvalues);
  else
    loss_kernel<32><<<blocks, kThreads, 0, s>>>(logits, old_log_probs, actions, advantages,
                                                values, returns, masks, n, (int)A, clip_param,
                                                value_loss_coef, entropy_coef, count, partials,
                                                clipped_losses, value_losses, entropies,
                                                total_losses, ratio, dlogits, dvalues);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  loss_finalize_kernel<<<1, 32, 0, s>>>(partials, blocks, count, loss ? loss : lossbuf);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}

// ---- alegpu_loss_static: alegpu_loss for callers that cannot allocate ------------------------------
// No counterpart in the reference.  Same kernels and
// launch sequence as alegpu_loss, but the scratch (mask count, loss slot, per-block partials) lives in
// a file-scope device array instead of a cudaMalloc'ed buffer, so the entry point never calls
// cudaMalloc / cudaFree (or anything that allocates) and can be used under stream capture or from
// harnesses that forbid allocation.  Every GPU that loads the library holds the array
// (kLossStaticFloats * 4 bytes = 16.25 KB) until exit.  All calls on one device share it: calls that
// overlap on the GPU race, and the array keeps the previous call's contents — the header slots are
// cleared on the stream at the start of every call; the partials need no clearing (every slot that is
// read is written first by the same call).
static constexpr int64_t kLossStaticMaxRows = ALEGPU_LOSS_STATIC_MAX_ROWS;        // 1 << 20
static constexpr int kLossStaticHeader = 64;                                      // 256-byte header
static constexpr int kLossStaticBlocks = (int)(kLossStaticMaxRows / kThreads);    // 4096 partials
static constexpr int kLossStaticFloats = kLossStaticHeader + kLossStaticBlocks;
static_assert(kLossStaticMaxRows % kThreads == 0, "max rows must be whole blocks");
static __device__ __align__(256) float g_loss_static_scratch[kLossStaticFloats];

namespace {
// In host code g_loss_static_scratch is only a stand-in for the symbol, not a device address: the
// address each device gives it is looked up once per device, with that device current.
std::mutex g_loss_static_mutex;
float *g_loss_static_ptr[64] = {};

int loss_static_scratch(float **out) {
  int dev = 0;
  ALEGPU_CUDA_CHECK(cudaGetDevice(&dev));
  ALEGPU_REQUIRE(dev >= 0 && dev < 64, ALEGPU_RUNTIME_ERROR, "loss_static: device ordinal out of range");
  std::lock_guard<std::mutex> lock(g_loss_static_mutex);
  if (!g_loss_static_ptr[dev]) {
    float *ptr = nullptr;
    ALEGPU_CUDA_CHECK(cudaGetSymbolAddress((void **)&ptr, g_loss_static_scratch));
    ALEGPU_REQUIRE(ptr != nullptr, ALEGPU_RUNTIME_ERROR, "loss_static: null symbol address");
    g_loss_static_ptr[dev] = ptr;
  }
  *out = g_loss_static_ptr[dev];
  return ALEGPU_OK;
}
} // namespace

extern "C" int alegpu_loss_static(const float *logits, const float *old_log_probs,
                                  const int64_t *actions, const float *advantages,
                                  const float *values, const float *returns, const uint8_t *masks,
                                  int64_t n, int64_t A, float clip_param, float value_loss_coef,
                                  float entropy_coef, float *loss, float *clipped_losses,
                                  float *value_losses, float *entropies, float *total_losses,
                                  float *ratio, float *dlogits, float *dvalues, void *stream) {
  ALEGPU_REQUIRE(n >= 1 && A >= 1 && A <= 32, ALEGPU_INVALID_ARGUMENT,
                 "loss_static: need n >= 1 and 1 <= A <= 32");
  ALEGPU_REQUIRE(n <= kLossStaticMaxRows, ALEGPU_INVALID_ARGUMENT,
                 "loss_static: n = " + std::to_string(n) + " exceeds the static scratch (at most " +
                     std::to_string(kLossStaticMaxRows) + " rows per call; use alegpu_loss)");
  ALEGPU_REQUIRE(logits && old_log_probs && actions && advantages && values && returns && masks,
                 ALEGPU_INVALID_ARGUMENT, "loss_static: null input pointer");
  cudaStream_t s = as_stream(stream);
  float *scratch = nullptr;
  if (int rc = loss_static_scratch(&scratch))
    return rc;
  const int blocks = (int)ceil_div(n, kThreads);
  float *count = scratch, *lossbuf = scratch + 1, *partials = scratch + kLossStaticHeader;
  ALEGPU_CUDA_CHECK(cudaMemsetAsync(scratch, 0, kLossStaticHeader * sizeof(float), s));
  if (int rc = launch_mask_count(masks, n, 1, count, s))
    return rc;
  if (A <= 4)
    loss_kernel<4><<<blocks, kThreads, 0, s>>>(logits, old_log_probs, actions, advantages,
                                               values, returns, masks, n, (int)A, clip_param,
                                               value_loss_coef, entropy_coef, count, partials,
                                               clipped_losses, value_losses, entropies,
                                               total_losses, ratio, dlogits, dvalues);
  else
    loss_kernel<32><<<blocks, kThreads, 0, s>>>(logits, old_log_probs, actions, advantages,
                                                values, returns, masks, n, (int)A, clip_param,
                                                value_loss_coef, entropy_coef, count, partials,
                                                clipped_losses, value_losses, entropies,
                                                total_losses, ratio, dlogits, dvalues);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  loss_finalize_kernel<<<1, 32, 0, s>>>(partials, blocks, count, loss ? loss : lossbuf);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}
