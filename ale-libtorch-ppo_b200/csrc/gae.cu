// gae.cu — reverse-scan GAE + Buffer::get epilogue, one kernel.
//
// Replaces ai::gae::gae (src/ai/gae.cc:55-79: ~12 ATen kernels per timestep, T launches) and
// Buffer::get's clamp_/copy_/add_/logical_not (src/ai/buffer.cc:67-74) and the
// `.any().item()` exclusivity check (gae.cc:49-53, folded in as a device error flag).
//
// Storage is env-major [E,T] (time contiguous), the scan is sequential in time per env.
// A CTA owns 32 envs: all threads stage [32 envs x 64 steps] tiles of rewards/values/flags
// through shared memory with coalesced loads along T (rows padded to 65 words so the
// thread-per-env column walk is bank-conflict free), one warp runs the recurrence with one
// lane per env in EXACTLY the reference's operation order (no fma contraction, so results are
// bit-identical to a sequential fp32 evaluation), and all threads write advantages / returns
// / clamped rewards / masks back coalesced.  23 algorithmic bytes per (env, step).
#include <mutex>

#include "common.cuh"

namespace alegpu {
namespace {

constexpr int kEnvs = 32;
constexpr int kChunk = 64;
constexpr int kThreads = 128;

template <bool kBufferGet>
__global__ void __launch_bounds__(kThreads)
    gae_scan_kernel(float *__restrict__ adv_out, float *__restrict__ rewards,
                    const float *__restrict__ values, const float *__restrict__ next_values,
                    const uint8_t *__restrict__ terminals, const uint8_t *__restrict__ truncations,
                    const uint8_t *__restrict__ starts, float *__restrict__ returns,
                    uint8_t *__restrict__ masks, int64_t E, int64_t T, float gamma, float gl,
                    int *__restrict__ error_flag) {
  __shared__ float s_r[kEnvs][kChunk + 1];
  __shared__ float s_v[kEnvs][kChunk + 1];
  __shared__ float s_a[kEnvs][kChunk + 1];
  __shared__ uint8_t s_f[kEnvs][kChunk + 4];
  const int tid = threadIdx.x;
  const int64_t e0 = (int64_t)blockIdx.x * kEnvs;
  const int nenv = (int)min((int64_t)kEnvs, E - e0);
  float last = 0.0f, nv = 0.0f;
  if (tid < nenv)
    nv = next_values[e0 + tid];
  for (int64_t t_hi = T; t_hi > 0; t_hi -= kChunk) {
    const int64_t t_lo = t_hi > kChunk ? t_hi - kChunk : 0;
    const int len = (int)(t_hi - t_lo);
    for (int idx = tid; idx < nenv * len; idx += kThreads) {
      const int e = idx / len, tt = idx - e * len;
      const int64_t gi = (e0 + e) * T + t_lo + tt;
      float r = rewards[gi];
      if (kBufferGet)
        r = fminf(fmaxf(r, -1.0f), 1.0f); // buffer.cc:67
      s_r[e][tt] = r;
      s_v[e][tt] = values[gi];
      const int st = starts[gi] != 0, te = terminals[gi] != 0, tr = truncations[gi] != 0;
      if (st + te + tr > 1)
        *error_flag = 1; // gae.cc:49-53
      s_f[e][tt] = (uint8_t)(st | (te << 1) | (tr << 2));
    }
    __syncthreads();
    if (tid < nenv) {
      for (int tt = len - 1; tt >= 0; --tt) {
        const float r = s_r[tid][tt], v = s_v[tid][tt];
        const int f = s_f[tid][tt];
        const float gnv = __fmul_rn(gamma, nv);
        const float delta = __fsub_rn(__fadd_rn(r, gnv), v);          // gae.cc:65-66
        const float running = __fadd_rn(delta, __fmul_rn(gl, last));  // gae.cc:61-63
        float a = (f & 1) ? 0.0f : running;                           // gae.cc:68-69
        a = (f & 2) ? __fsub_rn(r, v) : a;                            // gae.cc:64,70-71
        a = (f & 4) ? delta : a;                                      // gae.cc:72-73
        s_a[tid][tt] = a;
        last = a; // gae.cc:77
        nv = v;   // gae.cc:78
      }
    }
    __syncthreads();
    for (int idx = tid; idx < nenv * len; idx += kThreads) {
      const int e = idx / len, tt = idx - e * len;
      const int64_t gi = (e0 + e) * T + t_lo + tt;
      const float a = s_a[e][tt];
      adv_out[gi] = a;
      if (kBufferGet) {
        rewards[gi] = s_r[e][tt];
        returns[gi] = __fadd_rn(a, s_v[e][tt]); // buffer.cc:70-71
        masks[gi] = (s_f[e][tt] & 1) ? 0 : 1;   // buffer.cc:74
      }
    }
    __syncthreads();
  }
}

// The stateless entry points need a device int for the flag-exclusivity check.  One per DEVICE (a
// process may drive several GPUs), created under a lock; the calls that use it end with a stream
// synchronise (the reference's `.any().item()`), so two host threads on the same device must not run
// them concurrently on different streams — documented in alegpu.h (the reference is single-threaded here).
std::mutex g_flag_mutex;
int *g_error_flags[64] = {};
thread_local int *g_error_flag = nullptr; // the current device's flag, set by ensure_flag()

int ensure_flag() {
  int dev = 0;
  ALEGPU_CUDA_CHECK(cudaGetDevice(&dev));
  ALEGPU_REQUIRE(dev >= 0 && dev < 64, ALEGPU_RUNTIME_ERROR, "gae: device ordinal out of range");
  std::lock_guard<std::mutex> lock(g_flag_mutex);
  if (!g_error_flags[dev])
    ALEGPU_CUDA_CHECK(cudaMalloc(&g_error_flags[dev], sizeof(int)));
  g_error_flag = g_error_flags[dev];
  return ALEGPU_OK;
}

int check_flag(int *flag, cudaStream_t s) {
  int h = 0;
  ALEGPU_CUDA_CHECK(cudaMemcpyAsync(&h, flag, sizeof(int), cudaMemcpyDeviceToHost, s));
  ALEGPU_CUDA_CHECK(cudaStreamSynchronize(s));
  if (h)
    return fail(ALEGPU_INVALID_ARGUMENT,
                "Episode starts, terminals, and truncations must be mutually exclusive.");
  return ALEGPU_OK;
}

} // namespace
} // namespace alegpu

using namespace alegpu;

extern "C" int alegpu_gae(float *advantages, const float *rewards, const float *values,
                          const float *next_values, const uint8_t *terminals,
                          const uint8_t *truncations, const uint8_t *episode_starts, int64_t E,
                          int64_t T, float gamma, float lambda, void *stream) {
  ALEGPU_REQUIRE(E >= 0 && T >= 0, ALEGPU_INVALID_ARGUMENT, "gae: negative size");
  if (E == 0 || T == 0)
    return ALEGPU_OK;
  ALEGPU_REQUIRE(advantages && rewards && values && next_values && terminals && truncations &&
                     episode_starts,
                 ALEGPU_INVALID_ARGUMENT, "gae: null pointer");
  if (int rc = ensure_flag())
    return rc;
  cudaStream_t s = as_stream(stream);
  ALEGPU_CUDA_CHECK(cudaMemsetAsync(g_error_flag, 0, sizeof(int), s));
  gae_scan_kernel<false><<<(unsigned)ceil_div(E, kEnvs), kThreads, 0, s>>>(
      advantages, const_cast<float *>(rewards), values, next_values, terminals, truncations,
      episode_starts, nullptr, nullptr, E, T, gamma, gamma * lambda, g_error_flag);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  return check_flag(g_error_flag, s);
}

extern "C" int alegpu_buffer_get(float *rewards_inout, const float *values,
                                 const float *next_values, const uint8_t *terminals,
                                 const uint8_t *truncations, const uint8_t *episode_starts,
                                 int64_t E, int64_t T, float gamma, float lambda,
                                 int64_t write_index, float *advantages, float *returns,
                                 uint8_t *masks, int flags, int *device_error_flag,
                                 void *stream) {
  ALEGPU_REQUIRE(write_index == 0, ALEGPU_RUNTIME_ERROR,
                 "Buffer is not full, cannot compute GAE."); // buffer.cc:64-65
  ALEGPU_REQUIRE(E >= 0 && T >= 0, ALEGPU_INVALID_ARGUMENT, "buffer_get: negative size");
  if (E == 0 || T == 0)
    return ALEGPU_OK;
  ALEGPU_REQUIRE(rewards_inout && values && next_values && terminals && truncations &&
                     episode_starts && advantages && returns && masks,
                 ALEGPU_INVALID_ARGUMENT, "buffer_get: null pointer");
  cudaStream_t s = as_stream(stream);
  int *flag = device_error_flag;
  if (!flag) {
    if (int rc = ensure_flag())
      return rc;
    flag = g_error_flag;
    ALEGPU_CUDA_CHECK(cudaMemsetAsync(flag, 0, sizeof(int), s));
  }
  gae_scan_kernel<true><<<(unsigned)ceil_div(E, kEnvs), kThreads, 0, s>>>(
      advantages, rewards_inout, values, next_values, terminals, truncations, episode_starts,
      returns, masks, E, T, gamma, gamma * lambda, flag);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  if (flags & 1)
    return ALEGPU_OK;
  return check_flag(flag, s);
}
