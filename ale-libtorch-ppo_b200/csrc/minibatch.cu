// minibatch.cu — the opt-in minibatch preparation steps of the PPO update that BASELINE.json's
// north_star names next to the reference path: shuffled minibatch gather and per-minibatch
// advantage normalisation.  The reference slices contiguously and never normalises
// (src/ai/ppo/train.h:146-151 — its randperm is generated but not consumed), so both are OFF
// unless alegpu_train_options asks for them (include/alegpu.h).
#include <algorithm>

#include "minibatch.h"

namespace alegpu {
namespace {

// One CTA per gathered row: 28 224 observation bytes = 1764 x 16 B, read and written with 16-byte
// accesses (the row start is 16-byte aligned: 28224 = 1764 * 16), scalars by the first threads.
// HBM-bound: 2 x 28 224 + 37 + 8 (index) bytes per row.
__global__ void __launch_bounds__(256)
    gather_rows_kernel(const int64_t *__restrict__ idx, int64_t n_rows_src, int A,
                       const uint8_t *__restrict__ obs, const int64_t *__restrict__ actions,
                       const float *__restrict__ logp, const float *__restrict__ adv,
                       const float *__restrict__ ret, const float *__restrict__ values,
                       const uint8_t *__restrict__ masks, uint8_t *__restrict__ o_obs, int64_t *__restrict__ o_actions,
                       float *__restrict__ o_logp, float *__restrict__ o_adv,
                       float *__restrict__ o_ret, float *__restrict__ o_values,
                       uint8_t *__restrict__ o_masks, int *__restrict__ error_flag) {
  const int64_t i = blockIdx.x;
  int64_t src = idx[i];
  if (src < 0 || src >= n_rows_src) { // std::out_of_range in a torch index_select; reported by the
    if (threadIdx.x == 0)             // SAME train() call (checked before it returns); the staged row
      atomicExch(error_flag, 1);      // is zero-filled and masked out, never stale
    uint4 *d = reinterpret_cast<uint4 *>(o_obs + i * 28224);
    for (int c = threadIdx.x; c < 1764; c += 256)
      d[c] = make_uint4(0u, 0u, 0u, 0u);
    if (threadIdx.x < A)
      o_logp[i * A + threadIdx.x] = 0.0f;
    if (threadIdx.x == 32) o_actions[i] = 0;
    if (threadIdx.x == 33) o_adv[i] = 0.0f;
    if (threadIdx.x == 34) o_ret[i] = 0.0f;
    if (threadIdx.x == 35) o_masks[i] = 0;
    if (threadIdx.x == 36 && o_values) o_values[i] = 0.0f;
    return;
  }
  const uint4 *s = reinterpret_cast<const uint4 *>(obs + src * 28224);
  uint4 *d = reinterpret_cast<uint4 *>(o_obs + i * 28224);
  for (int c = threadIdx.x; c < 1764; c += 256) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(s + c));
    d[c] = v;
  }
  if (threadIdx.x < A)
    o_logp[i * A + threadIdx.x] = logp[src * A + threadIdx.x];
  if (threadIdx.x == 32) o_actions[i] = actions[src];
  if (threadIdx.x == 33) o_adv[i] = adv[src];
  if (threadIdx.x == 34) o_ret[i] = ret[src];
  if (threadIdx.x == 35) o_masks[i] = masks[src];
  if (threadIdx.x == 36 && values) o_values[i] = values[src];
}

__global__ void check_indices_kernel(const int64_t *__restrict__ idx, int64_t n, int64_t n_rows_src,
                                     int *__restrict__ error_flag) {
  bool bad = false;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    bad |= idx[i] < 0 || idx[i] >= n_rows_src;
  if (bad)
    atomicExch(error_flag, 1);
}

// (count, sum, sum of squares) of the unmasked advantages of one minibatch, in double, one CTA,
// fixed reduction order (bit-reproducible).  stats[3].
__global__ void __launch_bounds__(1024)
    adv_stats_kernel(const float *__restrict__ adv, const uint8_t *__restrict__ masks, int64_t n,
                     double *__restrict__ stats) {
  __shared__ double sh[3][32];
  double c = 0.0, s = 0.0, q = 0.0;
  for (int64_t i = threadIdx.x; i < n; i += 1024)
    if (masks[i]) {
      const double a = (double)adv[i];
      c += 1.0;
      s += a;
      q += a * a;
    }
  for (int o = 16; o > 0; o >>= 1) {
    c += __shfl_down_sync(0xFFFFFFFFu, c, o);
    s += __shfl_down_sync(0xFFFFFFFFu, s, o);
    q += __shfl_down_sync(0xFFFFFFFFu, q, o);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) {
    sh[0][warp] = c;
    sh[1][warp] = s;
    sh[2][warp] = q;
  }
  __syncthreads();
  if (warp == 0) {
    c = sh[0][lane];
    s = sh[1][lane];
    q = sh[2][lane];
    for (int o = 16; o > 0; o >>= 1) {
      c += __shfl_down_sync(0xFFFFFFFFu, c, o);
      s += __shfl_down_sync(0xFFFFFFFFu, s, o);
      q += __shfl_down_sync(0xFFFFFFFFu, q, o);
    }
    if (lane == 0) {
      stats[0] = c;
      stats[1] = s;
      stats[2] = q;
    }
  }
}

} // namespace

int launch_gather_rows(const int64_t *idx, int64_t n, int64_t n_rows_src, int A,
                       const uint8_t *obs, const int64_t *actions, const float *logp,
                       const float *adv, const float *ret, const float *values,
                       const uint8_t *masks, const StagedBatch &out, int *error_flag, cudaStream_t s) {
  if (n == 0)
    return ALEGPU_OK;
  ALEGPU_REQUIRE(A <= 32, ALEGPU_INVALID_ARGUMENT, "gather: action_size > 32");
  gather_rows_kernel<<<(unsigned)n, 256, 0, s>>>(idx, n_rows_src, A, obs, actions, logp, adv, ret,
                                                 values, masks, out.obs, out.actions, out.logp,
                                                 out.adv, out.ret, out.values, out.masks, error_flag);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}

int launch_check_indices(const int64_t *idx, int64_t n, int64_t n_rows_src, int *error_flag,
                         cudaStream_t s) {
  if (n == 0)
    return ALEGPU_OK;
  check_indices_kernel<<<(unsigned)std::min<int64_t>(ceil_div(n, 256), 2 * kNumSMs), 256, 0, s>>>(
      idx, n, n_rows_src, error_flag);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}

int launch_adv_stats(const float *adv, const uint8_t *masks, int64_t n, double *stats,
                     cudaStream_t s) {
  adv_stats_kernel<<<1, 1024, 0, s>>>(adv, masks, n, stats);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}

} // namespace alegpu
