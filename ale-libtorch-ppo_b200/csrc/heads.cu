// heads.cu — policy/value heads fused with the PPO loss.
//
// One warp per sample: the (A+1) x H head weights sit in shared memory, the warp forms the
// A logits and the value with shuffle reductions (src/bin/train.cc:262-264), evaluates the
// loss and its analytic gradient (loss.cuh == src/ai/ppo/losses.cc:4-47 + autograd), writes
// the per-sample metrics straight to their final [epoch,k,:] offsets (Metrics::set,
// src/ai/ppo/train.h:93-108) and back-propagates through the heads into dfc — so logits,
// values, dlogits and dvalues never round-trip through HBM between separate kernels.
#include <cstdlib>
#include <cuda_bf16.h>

#include "loss.cuh"
#include "net.h"

namespace alegpu {
namespace {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;

template <int kMaxA, bool kLoss>
__global__ void __launch_bounds__(kThreads)
    heads_kernel(const float *__restrict__ piw, const float *__restrict__ pib,
                 const float *__restrict__ vw, const float *__restrict__ vb,
                 const float *__restrict__ fc, int64_t n, int H, int A,
                 float *__restrict__ logits, float *__restrict__ values,
                 float *__restrict__ dfc, __nv_bfloat16 *__restrict__ dfc16, bool round_tf32,
                 float *__restrict__ dout, LossArgs L, float *__restrict__ partials) {
  extern __shared__ float sw[]; // [(A+1)][H] head weights, row A = value head
  __shared__ float red[kWarps];
  for (int i = threadIdx.x; i < (A + 1) * H; i += kThreads)
    sw[i] = i < A * H ? piw[i] : vw[i - A * H];
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float inv = 0.0f;
  if (kLoss)
    inv = 1.0f / L.count[0];
  float contrib = 0.0f;
  for (int64_t i = (int64_t)blockIdx.x * kWarps + warp; i < n; i += (int64_t)gridDim.x * kWarps) {
    const float *h = fc + i * H;
    float acc[kMaxA + 1];
#pragma unroll
    for (int a = 0; a <= kMaxA; ++a)
      acc[a] = 0.0f;
    for (int j = lane; j < H; j += 32) {
      const float x = h[j];
#pragma unroll
      for (int a = 0; a <= kMaxA; ++a)
        if (a <= A)
          acc[a] = fmaf(x, sw[a * H + j], acc[a]);
    }
    float z[kMaxA];
    float v = 0.0f;
#pragma unroll
    for (int a = 0; a <= kMaxA; ++a)
      if (a <= A) {
        const float t = warp_sum(acc[a]);
        if (a < A) {
          if (a < kMaxA)
            z[a] = t + pib[a];
        } else {
          v = t + vb[0];
        }
      }
    if (lane == 0) {
      if (logits)
        for (int a = 0; a < A; ++a)
          logits[i * A + a] = z[a];
      if (values)
        values[i] = v;
    }
    if (kLoss) {
      float ol[kMaxA], dz[kMaxA], dv;
#pragma unroll
      for (int a = 0; a < kMaxA; ++a)
        if (a < A)
          ol[a] = L.old_log_probs[i * A + a];
      const bool m = L.masks[i] != 0;
      if (lane == 0 && L.error_flag && (L.actions[i] < 0 || L.actions[i] >= A))
        atomicExch(L.error_flag, 1); // the reference's gather(-1, actions) would throw (losses.cc:12)
      const float adv_raw = L.advantages[i];
      float adv = adv_raw;
      const float ret = L.returns[i];
      if (L.adv_stats) { // every thread derives the same (mean, 1/(std+eps)) from the 3 doubles
        const double c = L.adv_stats[0], mean = L.adv_stats[1] / c;
        const double var = (L.adv_stats[2] - c * mean * mean) / (c - 1.0);
        adv = (float)(((double)adv - mean) / (sqrt(fmax(var, 0.0)) + 1e-8));
      }
      const PpoSample s = ppo_sample<kMaxA>(z, ol, A, (int)L.actions[i], adv, v, ret, L.clip,
                                            L.vcoef, L.ecoef, m ? inv : 0.0f, dz, &dv,
                                            L.old_values != nullptr,
                                            L.old_values ? L.old_values[i] : 0.0f);
      if (lane == 0) {
        if (L.m_clipped) L.m_clipped[i] = s.surr;
        if (L.m_value) L.m_value[i] = s.value_loss;
        if (L.m_entropy) L.m_entropy[i] = s.entropy;
        if (L.m_total) L.m_total[i] = s.total;
        if (L.m_ratio) L.m_ratio[i] = s.ratio;
        if (L.m_adv) L.m_adv[i] = adv_raw;
        if (L.m_ret) L.m_ret[i] = ret;
        if (L.m_masks) L.m_masks[i] = m ? 1 : 0;
        for (int a = 0; a < A; ++a)
          dout[i * (A + 1) + a] = dz[a];
        dout[i * (A + 1) + A] = dv;
        contrib += m ? s.total : 0.0f;
      }
      // dfc = dlogits . piw + dvalue * vw
      for (int j = lane; j < H; j += 32) {
        float g = dv * sw[A * H + j];
#pragma unroll
        for (int a = 0; a < kMaxA; ++a)
          if (a < A)
            g = fmaf(dz[a], sw[a * H + j], g);
        if (round_tf32) { // the tf32 tensor-core backward multiplies the upper 19 bits: round, not truncate
          uint32_t r;
          asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(g));
          g = __uint_as_float(r);
        }
        dfc[i * H + j] = g;
        if (dfc16)
          dfc16[i * H + j] = __float2bfloat16_rn(g);
      }
    }
  }
  if (kLoss) {
    if (lane == 0)
      red[warp] = contrib;
    __syncthreads();
    __shared__ bool s_last;
    if (threadIdx.x == 0) {
      float t = 0.0f;
      for (int w = 0; w < kWarps; ++w)
        t += red[w];
      partials[blockIdx.x] = t;
      s_last = false;
      if (L.loss_sum) {
        __threadfence();
        s_last = atomicAdd(L.done_counter, 1u) == gridDim.x - 1;
      }
    }
    __syncthreads();
    if (s_last && threadIdx.x < 32) { // every partial is visible: fixed-order sum (bit-reproducible)
      __threadfence();
      float t = 0.0f;
      for (int i = threadIdx.x; i < (int)gridDim.x; i += 32)
        t += __ldcg(partials + i);
      t = warp_sum(t);
      if (threadIdx.x == 0) {
        *L.loss_sum = t;
        *L.done_counter = 0u;
      }
    }
  }
}

// partial[z][a][j] = sum_{m in split z} dout[m][a] * (j < H ? fc[m][j] : 1)
template <int kMaxA1>
__global__ void __launch_bounds__(128)
    heads_wgrad_kernel(const float *__restrict__ dout, const float *__restrict__ fc, int64_t n,
                       int H, int A1, int rows_per_split, float *__restrict__ ws) {
  __shared__ float sd[64 * kMaxA1];
  const int j = blockIdx.x * 128 + threadIdx.x;
  const int64_t m0 = (int64_t)blockIdx.y * rows_per_split;
  const int64_t m1 = min(n, m0 + rows_per_split);
  float acc[kMaxA1];
#pragma unroll
  for (int a = 0; a < kMaxA1; ++a)
    acc[a] = 0.0f;
  for (int64_t mb = m0; mb < m1; mb += 64) {
    const int cnt = (int)min((int64_t)64, m1 - mb);
    __syncthreads();
    for (int i = threadIdx.x; i < cnt * A1; i += 128)
      sd[i] = dout[mb * A1 + i];
    __syncthreads();
    if (j <= H)
      for (int r = 0; r < cnt; ++r) {
        const float x = j < H ? fc[(mb + r) * H + j] : 1.0f;
#pragma unroll
        for (int a = 0; a < kMaxA1; ++a)
          if (a < A1)
            acc[a] = fmaf(sd[r * A1 + a], x, acc[a]);
      }
  }
  if (j <= H)
#pragma unroll
    for (int a = 0; a < kMaxA1; ++a)
      if (a < A1)
        ws[((size_t)blockIdx.y * A1 + a) * (H + 1) + j] = acc[a];
}

__global__ void heads_reduce_kernel(const float *__restrict__ ws, int splits, int H, int A,
                                    float *__restrict__ gpiw, float *__restrict__ gpib,
                                    float *__restrict__ gvw, float *__restrict__ gvb) {
  const int total = (A + 1) * (H + 1);
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += gridDim.x * blockDim.x) {
    float s = 0.0f;
    for (int z = 0; z < splits; ++z)
      s += ws[(size_t)z * total + idx];
    const int a = idx / (H + 1), j = idx - a * (H + 1);
    if (a < A) {
      if (j < H) gpiw[a * H + j] = s; else gpib[a] = s;
    } else {
      if (j < H) gvw[j] = s; else gvb[0] = s;
    }
  }
}

__global__ void finalize_metrics_kernel(const float *__restrict__ loss_sum,
                                        const float *__restrict__ count, float *__restrict__ m_loss,
                                        int64_t mb, float *__restrict__ loss_out) {
  const float loss = loss_sum[0] / count[0]; // losses.cc:19
  if (blockIdx.x == 0 && threadIdx.x == 0 && loss_out)
    loss_out[0] = loss;
  if (m_loss) // Metrics.loss[epoch,k,:] is the scalar broadcast over mb (train.h:98)
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < mb;
         i += (int64_t)gridDim.x * blockDim.x)
      m_loss[i] = loss;
}

int heads_grid(int64_t n) {
  return (int)std::max<int64_t>(1, std::min<int64_t>(ceil_div(n, kWarps), kNumSMs * 4));
}

template <bool kLoss>
int launch_heads(const NetParams &P, const float *fc, int64_t n, float *logits, float *values,
                 float *dfc, void *dfc16, bool round_tf32, float *dout, const LossArgs &L,
                 float *partials, int grid, cudaStream_t s) {
  const int H = (int)P.hidden, A = (int)P.actions;
  const size_t smem = (size_t)(A + 1) * H * sizeof(float);
  ALEGPU_REQUIRE(smem <= 200 * 1024, ALEGPU_INVALID_ARGUMENT, "heads: (A+1)*H too large");
  if (A <= 4) {
    auto k = heads_kernel<4, kLoss>;
    if (smem > 48 * 1024)
      ALEGPU_CUDA_CHECK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k<<<grid, kThreads, smem, s>>>(P.piw, P.pib, P.vw, P.vb, fc, n, H, A, logits, values, dfc,
                                   (__nv_bfloat16 *)dfc16, round_tf32, dout, L, partials);
  } else {
    auto k = heads_kernel<18, kLoss>;
    if (smem > 48 * 1024)
      ALEGPU_CUDA_CHECK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k<<<grid, kThreads, smem, s>>>(P.piw, P.pib, P.vw, P.vb, fc, n, H, A, logits, values, dfc,
                                   (__nv_bfloat16 *)dfc16, round_tf32, dout, L, partials);
  }
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}

// ---- rollout-side action selection ------------------------------------------------------------
// The reference's action selector (src/bin/train.cc:367-379): probabilities = softmax(logits),
// actions = multinomial(probabilities, 1).  One thread per env draws by inverse CDF: the action is
// the first a with cumsum(p)[a] > u, u in [0,1) either supplied per env (parity runs) or generated
// from a counter-based hash of (seed, step, env) — libtorch's Philox stream cannot be reproduced,
// so parity with the reference is distributional (tests: chi-square against softmax).
__device__ __forceinline__ float act_uniform(uint64_t seed, uint64_t step, uint64_t env) {
  uint64_t x = seed ^ (step * 0x9E3779B97F4A7C15ull) ^ (env * 0xD1B54A32D192ED03ull);
  x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull; // splitmix64 finaliser
  x ^= x >> 27; x *= 0x94D049BB133111EBull;
  x ^= x >> 31;
  return (float)(x >> 40) * (1.0f / 16777216.0f); // 24 random bits -> [0,1)
}
__global__ void act_sample_kernel(const float *__restrict__ logits_in,
                                  const float *__restrict__ values_in, int64_t n, int A,
                                  const float *__restrict__ uniforms, uint64_t seed, uint64_t step,
                                  float *__restrict__ logits_out, int64_t logits_stride,
                                  float *__restrict__ values_out, int64_t values_stride,
                                  int64_t *__restrict__ actions_out, int64_t actions_stride,
                                  int64_t *__restrict__ actions_dense) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n)
    return;
  const float *z = logits_in + i * A;
  float mx = z[0];
  for (int a = 1; a < A; ++a)
    mx = fmaxf(mx, z[a]);
  float sum = 0.0f;
  for (int a = 0; a < A; ++a)
    sum += expf(z[a] - mx);
  const float u = uniforms ? uniforms[i] : act_uniform(seed, step, (uint64_t)i);
  float cum = 0.0f;
  int act = A - 1;
  for (int a = 0; a < A; ++a) {
    cum += expf(z[a] - mx) / sum;
    if (cum > u) {
      act = a;
      break;
    }
  }
  if (logits_out)
    for (int a = 0; a < A; ++a)
      logits_out[i * logits_stride + a] = z[a];
  if (values_out)
    values_out[i * values_stride] = values_in[i];
  if (actions_out)
    actions_out[i * actions_stride] = act;
  if (actions_dense)
    actions_dense[i] = act;
}

} // namespace

int launch_act_sample(const float *logits, const float *values, int64_t n, int A,
                      const float *uniforms, uint64_t seed, uint64_t step, float *logits_out,
                      int64_t logits_stride, float *values_out, int64_t values_stride,
                      int64_t *actions_out, int64_t actions_stride, int64_t *actions_dense,
                      cudaStream_t s) {
  if (n == 0)
    return ALEGPU_OK;
  act_sample_kernel<<<(unsigned)ceil_div(n, 128), 128, 0, s>>>(
      logits, values, n, A, uniforms, seed, step, logits_out, logits_stride, values_out,
      values_stride, actions_out, actions_stride, actions_dense);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}

int heads_forward(const NetParams &P, const float *fc, int64_t n, float *logits, float *values,
                  cudaStream_t s) {
  if (n == 0)
    return ALEGPU_OK;
  LossArgs L{};
  return launch_heads<false>(P, fc, n, logits, values, nullptr, nullptr, false, nullptr, L, nullptr,
                             heads_grid(n), s);
}

int heads_loss_backward(const NetParams &P, NetActs &A, int64_t n, const LossArgs &L,
                        float *loss_partials, int *nparts, cudaStream_t s) {
  const int grid = heads_grid(n);
  *nparts = grid;
  return launch_heads<true>(P, A.fc, n, A.logits, A.values, A.dfc, A.dfc16, A.dfc_round_tf32, A.dout,
                            L, loss_partials, grid, s);
}

int heads_wgrad(const NetParams &G, const NetActs &A, int64_t n, float *ws, size_t ws_floats,
                cudaStream_t s) {
  const int H = (int)G.hidden, Ac = (int)G.actions, A1 = Ac + 1;
  const size_t per = (size_t)A1 * (H + 1);
  // rows per split: 64 by default (4 splits x 5 column blocks = 20 CTAs at v0's 256-row minibatch, 14 us in
  // profiles/r01e_launches_v0_default.txt); ALEGPU_HEADS_WGRAD_ROWS is the A/B switch for finer splits
  static const int64_t rows = [] {
    const char *e = getenv("ALEGPU_HEADS_WGRAD_ROWS");
    const int64_t v = e ? atoll(e) : 64;
    return v >= 1 ? v : 64;
  }();
  int splits = (int)std::max<int64_t>(1, std::min<int64_t>(n / rows, 2 * kNumSMs));
  splits = (int)std::min<size_t>(splits, ws_floats / per);
  ALEGPU_REQUIRE(splits >= 1, ALEGPU_RUNTIME_ERROR, "heads_wgrad: workspace too small");
  const int rps = (int)ceil_div(n, splits);
  splits = (int)ceil_div(n, rps);
  dim3 grid((unsigned)ceil_div(H + 1, 128), (unsigned)splits);
  if (A1 <= 5)
    heads_wgrad_kernel<5><<<grid, 128, 0, s>>>(A.dout, A.fc, n, H, A1, rps, ws);
  else
    heads_wgrad_kernel<19><<<grid, 128, 0, s>>>(A.dout, A.fc, n, H, A1, rps, ws);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  heads_reduce_kernel<<<(unsigned)ceil_div((int64_t)per, 256), 256, 0, s>>>(
      ws, splits, H, Ac, G.piw, G.pib, G.vw, G.vb);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}

int finalize_metrics(const float *loss_sum, const float *count, float *m_loss, int64_t mb,
                     float *loss_out, cudaStream_t s) {
  const int blocks = m_loss ? (int)std::min<int64_t>(ceil_div(mb, 256), 64) : 1;
  finalize_metrics_kernel<<<blocks, 256, 0, s>>>(loss_sum, count, m_loss, mb, loss_out);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}

} // namespace alegpu
