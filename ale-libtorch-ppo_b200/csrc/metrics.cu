// metrics.cu — on-device reduction of ai::ppo::train::Metrics for the logging feed.
//
// The reference's log_data (src/bin/train.cc:163-210) pulls ~20 scalars / vectors out of the
// [epochs,nmb,mb] metric tensors with `.mean().item()` and `masked_select(...).to(kCPU)` — one
// device sync each.  Here one kernel forms every scalar of that function (masked means taken over
// the UNMASKED samples, train.cc:144-147) into a small struct that goes to the host in one copy,
// and masked_select is a stable three-kernel stream compaction (histogram feeds, train.cc:149-155).
#include "common.cuh"

namespace alegpu {
namespace {

constexpr int kSumThreads = 1024;

__device__ __forceinline__ double block_sum(double v, double *sh) {
  for (int o = 16; o > 0; o >>= 1)
    v += __shfl_down_sync(0xFFFFFFFFu, v, o);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads(); // sh reuse
  if (lane == 0)
    sh[warp] = v;
  __syncthreads();
  v = threadIdx.x < (blockDim.x >> 5) ? sh[threadIdx.x] : 0.0;
  if (warp == 0)
    for (int o = 16; o > 0; o >>= 1)
      v += __shfl_down_sync(0xFFFFFFFFu, v, o);
  return v; // valid in thread 0
}

// one CTA, fixed order, double accumulation: bit-reproducible and well inside fp32 rounding of
// torch's own mean()
__global__ void __launch_bounds__(kSumThreads)
    metrics_summary_kernel(alegpu_metrics m, int64_t steps, int64_t n,
                           alegpu_metrics_summary *__restrict__ out) {
  __shared__ double sh[32];
  const float *series[5] = {m.clipped_losses, m.value_losses, m.entropies, m.ratio, m.total_losses};
  double acc[5] = {0, 0, 0, 0, 0}, cnt = 0.0, loss = 0.0, gn = 0.0;
  for (int64_t i = threadIdx.x; i < n; i += kSumThreads) {
    const bool on = m.masks ? m.masks[i] != 0 : true;
    if (on) {
      cnt += 1.0;
#pragma unroll
      for (int k = 0; k < 5; ++k)
        if (series[k])
          acc[k] += (double)series[k][i];
    }
    if (m.loss)
      loss += (double)m.loss[i]; // metrics.loss.mean(): unmasked mean of the broadcast scalar
  }
  if (m.clipped_gradients)
    for (int64_t i = threadIdx.x; i < steps; i += kSumThreads)
      gn += (double)m.clipped_gradients[i];
  cnt = block_sum(cnt, sh);
  loss = block_sum(loss, sh);
  gn = block_sum(gn, sh);
#pragma unroll
  for (int k = 0; k < 5; ++k)
    acc[k] = block_sum(acc[k], sh);
  if (threadIdx.x == 0) {
    const double inv = cnt > 0.0 ? 1.0 / cnt : 0.0;
    out->mean_clipped_gradient = steps > 0 ? (float)(gn / (double)steps) : 0.0f;
    out->mean_loss = n > 0 ? (float)(loss / (double)n) : 0.0f;
    out->mean_clipped_loss = (float)(acc[0] * inv);
    out->mean_value_loss = (float)(acc[1] * inv);
    out->mean_entropy = (float)(acc[2] * inv);
    out->mean_ratio = (float)(acc[3] * inv);
    out->mean_total_loss = (float)(acc[4] * inv);
    out->unmasked = (float)cnt;
  }
}

// ---- masked_select: stable compaction ---------------------------------------------------------
constexpr int kSelThreads = 256, kSelPer = 8, kSelTile = kSelThreads * kSelPer;

__global__ void __launch_bounds__(kSelThreads)
    select_count_kernel(const uint8_t *__restrict__ mask, int64_t n, int *__restrict__ tile_count) {
  __shared__ int sh[kSelThreads / 32];
  const int64_t base = (int64_t)blockIdx.x * kSelTile;
  int c = 0;
  for (int j = 0; j < kSelPer; ++j) {
    const int64_t i = base + (int64_t)threadIdx.x * kSelPer + j;
    c += (i < n && mask[i]) ? 1 : 0;
  }
  for (int o = 16; o > 0; o >>= 1)
    c += __shfl_down_sync(0xFFFFFFFFu, c, o);
  if ((threadIdx.x & 31) == 0)
    sh[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int w = 0; w < kSelThreads / 32; ++w)
      t += sh[w];
    tile_count[blockIdx.x] = t;
  }
}
// exclusive scan of the tile counts in place (one CTA walks them in chunks), total -> *count
__global__ void __launch_bounds__(1024)
    select_scan_kernel(int *__restrict__ tile_count, int tiles, int64_t *__restrict__ count) {
  __shared__ int sh[1024];
  __shared__ int carry;
  if (threadIdx.x == 0)
    carry = 0;
  __syncthreads();
  for (int b = 0; b < tiles; b += 1024) {
    const int i = b + threadIdx.x;
    const int v = i < tiles ? tile_count[i] : 0;
    sh[threadIdx.x] = v;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) { // Hillis-Steele inclusive scan
      const int t = threadIdx.x >= o ? sh[threadIdx.x - o] : 0;
      __syncthreads();
      sh[threadIdx.x] += t;
      __syncthreads();
    }
    if (i < tiles)
      tile_count[i] = carry + sh[threadIdx.x] - v;
    __syncthreads();
    if (threadIdx.x == 1023)
      carry += sh[1023];
    __syncthreads();
  }
  if (threadIdx.x == 0)
    *count = carry;
}
__global__ void __launch_bounds__(kSelThreads)
    select_scatter_kernel(const float *__restrict__ x, const uint8_t *__restrict__ mask, int64_t n,
                          const int *__restrict__ tile_offset, float *__restrict__ out) {
  __shared__ int sh[kSelThreads];
  const int64_t base = (int64_t)blockIdx.x * kSelTile + (int64_t)threadIdx.x * kSelPer;
  int c = 0;
  for (int j = 0; j < kSelPer; ++j)
    c += (base + j < n && mask[base + j]) ? 1 : 0;
  sh[threadIdx.x] = c;
  __syncthreads();
  for (int o = 1; o < kSelThreads; o <<= 1) {
    const int t = threadIdx.x >= o ? sh[threadIdx.x - o] : 0;
    __syncthreads();
    sh[threadIdx.x] += t;
    __syncthreads();
  }
  int pos = tile_offset[blockIdx.x] + sh[threadIdx.x] - c;
  for (int j = 0; j < kSelPer; ++j)
    if (base + j < n && mask[base + j])
      out[pos++] = x[base + j];
}

struct SelScratch {
  int *tiles = nullptr;
  size_t cap = 0;
} g_sel;

} // namespace
} // namespace alegpu

using namespace alegpu;

extern "C" int alegpu_metrics_summarize(const alegpu_metrics *metrics, int64_t num_epochs,
                                        int64_t num_mini_batches, int64_t mini_batch_size,
                                        alegpu_metrics_summary *summary,
                                        alegpu_metrics_summary *summary_host, void *stream) {
  ALEGPU_REQUIRE(metrics && summary, ALEGPU_INVALID_ARGUMENT, "metrics_summarize: null argument");
  ALEGPU_REQUIRE(num_epochs >= 0 && num_mini_batches >= 0 && mini_batch_size >= 0,
                 ALEGPU_INVALID_ARGUMENT, "metrics_summarize: negative shape");
  cudaStream_t s = as_stream(stream);
  const int64_t steps = num_epochs * num_mini_batches;
  metrics_summary_kernel<<<1, kSumThreads, 0, s>>>(*metrics, steps, steps * mini_batch_size, summary);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  if (summary_host)
    ALEGPU_CUDA_CHECK(cudaMemcpyAsync(summary_host, summary, sizeof(alegpu_metrics_summary),
                                      cudaMemcpyDeviceToHost, s));
  return ALEGPU_OK;
}

extern "C" int alegpu_masked_select(const float *values, const uint8_t *masks, int64_t n,
                                    float *out, int64_t *count, void *stream) {
  ALEGPU_REQUIRE(n >= 0 && count, ALEGPU_INVALID_ARGUMENT, "masked_select: bad argument");
  ALEGPU_REQUIRE(n == 0 || (values && masks && out), ALEGPU_INVALID_ARGUMENT,
                 "masked_select: null pointer");
  cudaStream_t s = as_stream(stream);
  if (n == 0) {
    ALEGPU_CUDA_CHECK(cudaMemsetAsync(count, 0, sizeof(int64_t), s));
    return ALEGPU_OK;
  }
  const int64_t tiles = ceil_div(n, (int64_t)kSelTile);
  ALEGPU_REQUIRE(tiles < (1 << 30), ALEGPU_INVALID_ARGUMENT, "masked_select: too many elements");
  if (g_sel.cap < (size_t)tiles) {
    if (g_sel.tiles)
      cudaFree(g_sel.tiles);
    g_sel.tiles = nullptr;
    g_sel.cap = 0;
    ALEGPU_CUDA_CHECK(cudaMalloc(&g_sel.tiles, sizeof(int) * (size_t)tiles));
    g_sel.cap = (size_t)tiles;
  }
  select_count_kernel<<<(unsigned)tiles, kSelThreads, 0, s>>>(masks, n, g_sel.tiles);
  ALEGPU_COUNT_LAUNCH();
  select_scan_kernel<<<1, 1024, 0, s>>>(g_sel.tiles, (int)tiles, count);
  ALEGPU_COUNT_LAUNCH();
  select_scatter_kernel<<<(unsigned)tiles, kSelThreads, 0, s>>>(values, masks, n, g_sel.tiles, out);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}
