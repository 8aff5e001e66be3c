// optim.cu — global-norm gradient clip + Adam over one flat parameter buffer.
//
// Replaces clip_grad_norm_ (src/ai/ppo/train.cc:12-46: 10 norm kernels + stack + norm +
// clamp + 10 mul_) and torch::optim::Adam(lr, eps=1e-5) as configured at
// src/bin/train.cc:360-362 (betas 0.9/0.999, no weight decay, no amsgrad) with two launches:
//   grad_sumsq : deterministic per-block partial sums of g^2 (sqrt(sum_i ||g_i||^2) over the
//                per-tensor norms == the L2 norm of the flat gradient)
//                per-tensor norms == the L2 norm of the flat gradient); the LAST block to finish
//                sums the partials in index order and leaves coef = min(1, max_norm/(total+1e-6))
//                (train.cc:39-41), lr/(1-b1^t) and sqrt(1-b2^t) in the device-side OptState
//   adam_clip  : applies
//                g*=coef; m=b1 m+(1-b1)g; v=b2 v+(1-b2)g^2;
//                p -= (lr/(1-b1^t)) m / (sqrt(v)/sqrt(1-b2^t) + eps)      (torch's form)
// with 16-byte vector loads/stores: 28 B per parameter per step (+2 B for the bf16 shadow the
// tensor-core path consumes, + 4 B for the norm pass).  lr and the step counter live in a
// device struct so a captured CUDA graph stays correct across lr anneals.
#include <cuda_bf16.h>

#include "common.cuh"
#include "optim.h"

namespace alegpu {
namespace {

constexpr int kThreads = 256;

__global__ void __launch_bounds__(kThreads)
    grad_sumsq_kernel(const float *__restrict__ g, int64_t n4, int64_t n_valid,
                      float *__restrict__ partials,
                      OptState *__restrict__ st, float max_norm, float *__restrict__ total_norm_out) {
  __shared__ float red[kThreads / 32];
  __shared__ bool s_last;
  float acc = 0.0f;
  const float4 *g4 = reinterpret_cast<const float4 *>(g);
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n4;
       i += (int64_t)gridDim.x * kThreads) {
    float4 x = g4[i];
    if (4 * i + 3 >= n_valid) { // the padding behind the last parameter is not part of the norm
      x.y = 4 * i + 1 < n_valid ? x.y : 0.0f;
      x.z = 4 * i + 2 < n_valid ? x.z : 0.0f;
      x.w = 0.0f;
      x.x = 4 * i < n_valid ? x.x : 0.0f;
    }
    acc += x.x * x.x + x.y * x.y + x.z * x.z + x.w * x.w;
  }
  acc = block_sum<float, kThreads>(acc, red);
  if (threadIdx.x == 0) {
    partials[blockIdx.x] = acc;
    __threadfence();
    s_last = atomicAdd(&st->done_count, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!s_last || threadIdx.x >= 32)
    return;
  // the last block: every partial is visible; sum them in index order (bit-reproducible), then
  // everything the Adam kernel needs as three floats, so it starts straight on its loads
  __threadfence();
  double s = 0.0;
  for (int i = threadIdx.x; i < (int)gridDim.x; i += 32)
    s += (double)__ldcg(partials + i);
  s = warp_sum(s);
  if (threadIdx.x == 0) {
    const float total = (float)sqrt(s);
    const float coef = (float)((double)max_norm / ((double)total + 1e-6)); // train.cc:39
    st->coef = fminf(coef, 1.0f);                                          // train.cc:40-41
    const long long step = st->step + 1; // Adam's counter counts minibatch updates (train.h:129)
    st->step = step;
    const double t = (double)step;
    const double bc1 = 1.0 - pow(0.9, t), bc2 = 1.0 - pow(0.999, t);
    st->step_size = (float)(st->lr / bc1);
    st->bc2_sqrt = (float)sqrt(bc2);
    st->done_count = 0u;
    if (total_norm_out)
      *total_norm_out = total; // the PRE-clip norm (train.h:104)
  }
}

__device__ __forceinline__ float round_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}
// ST = element type of the operand shadow the tensor-core path reads: bf16, or float (tf32-rounded)
template <class ST>
__global__ void __launch_bounds__(kThreads)
    adam_clip_kernel(float *__restrict__ p, const float *__restrict__ g, float *__restrict__ m,
                     float *__restrict__ v, ST *__restrict__ shadow, ST *__restrict__ c1w_perm,
                     int64_t n4, const OptState *__restrict__ st) {
  float4 *p4 = reinterpret_cast<float4 *>(p);
  const float4 *g4 = reinterpret_cast<const float4 *>(g);
  float4 *m4 = reinterpret_cast<float4 *>(m);
  float4 *v4 = reinterpret_cast<float4 *>(v);
  const int64_t stride = (int64_t)gridDim.x * kThreads;
  int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x;
  const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
  float4 pp = zero4, gg = zero4, mm = zero4, vv = zero4;
  if (i < n4) {
    pp = p4[i];
    gg = g4[i];
    mm = m4[i];
    vv = v4[i];
  }
  const float coef = st->coef, step_size = st->step_size, bc2_sqrt = st->bc2_sqrt;
  const float b1 = 0.9f, b2 = 0.999f, omb1 = (float)(1.0 - 0.9), omb2 = (float)(1.0 - 0.999);
  const float eps = 1e-5f;
  while (i < n4) {
    const int64_t nxt = i + stride;
    float4 pn = zero4, gn = zero4, mn = zero4, vn = zero4;
    if (nxt < n4) { // next element's loads before this one's math and stores
      pn = p4[nxt];
      gn = g4[nxt];
      mn = m4[nxt];
      vn = v4[nxt];
    }
    float *pa = &pp.x, *ga = &gg.x, *ma = &mm.x, *va = &vv.x;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float gc = ga[k] * coef; // train.cc:42-44 (the scaled gradient is consumed here only)
      ma[k] = ma[k] * b1 + gc * omb1;
      va[k] = va[k] * b2 + (gc * gc) * omb2;
      const float denom = sqrtf(va[k]) / bc2_sqrt + eps;
      pa[k] = pa[k] - step_size * (ma[k] / denom);
    }
    p4[i] = pp;
    m4[i] = mm;
    v4[i] = vv;
    if (shadow) {
      if constexpr (sizeof(ST) == 2) {
        __nv_bfloat162 lo = __floats2bfloat162_rn(pp.x, pp.y), hi = __floats2bfloat162_rn(pp.z, pp.w);
        uint2 pk;
        pk.x = *reinterpret_cast<uint32_t *>(&lo);
        pk.y = *reinterpret_cast<uint32_t *>(&hi);
        reinterpret_cast<uint2 *>(shadow)[i] = pk;
      } else {
        reinterpret_cast<float4 *>(shadow)[i] =
            make_float4(round_tf32(pp.x), round_tf32(pp.y), round_tf32(pp.z), round_tf32(pp.w));
      }
      // conv1's weights (the first 32*256 parameters, OIHW: k = c*64 + ky*8 + kx) are also kept in
      // the (tap, y&3, x&3, c) order in which the tensor-core conv1 walks the space-to-depth observations
      if (c1w_perm && i < 32 * 256 / 4) {
        const int idx = (int)i * 4, co = idx >> 8, k = idx & 255;
        const int c = k >> 6, ky = (k >> 3) & 7, kx = k & 7;
        ST *o = c1w_perm + co * 256 + ((ky >> 2) * 2 + (kx >> 2)) * 64 + (ky & 3) * 16 +
                (kx & 3) * 4 + c; // net_tc.cu: conv1_k_of_kprime
        if constexpr (sizeof(ST) == 2) {
          o[0] = __float2bfloat16_rn(pp.x);
          o[4] = __float2bfloat16_rn(pp.y);
          o[8] = __float2bfloat16_rn(pp.z);
          o[12] = __float2bfloat16_rn(pp.w);
        } else {
          o[0] = round_tf32(pp.x);
          o[4] = round_tf32(pp.y);
          o[8] = round_tf32(pp.z);
          o[12] = round_tf32(pp.w);
        }
      }
    }
    pp = pn; gg = gn; mm = mn; vv = vn;
    i = nxt;
  }
}

// ---- one launch: ||g||^2 -> grid barrier -> clip + Adam (+ Metrics.loss broadcast) ---------------------
// kNormBlocks CTAs (2 per SM, all co-resident) each own a fixed slice of the flat buffer; a thread keeps
// its <= kCache float4 of the gradient IN REGISTERS across the barrier, so the gradient is read once
// (28 B per parameter in total, no 4 B/param norm pass), and its p / m / v loads are issued BEFORE it
// waits, so they are in flight while the last CTA forms the clip coefficient.  The barrier: every CTA
// reads the step counter, publishes its partial and arrives on done_count; the last one sums the
// partials in index order (bit-reproducible), writes coef / step_size / bc2_sqrt and then advances
// st->step, which is what the others spin on.  Used when the buffer fits the register cache
// (P <= kCache * 4 * 296 * 256 = 1.8 M floats: every shipped config); larger nets take the two kernels.
constexpr int kCache = 6;
template <class ST>
__global__ void __launch_bounds__(kThreads, 2)
    adam_fused_kernel(float *__restrict__ p, const float *__restrict__ g, float *__restrict__ m,
                      float *__restrict__ v, ST *__restrict__ shadow, ST *__restrict__ c1w_perm,
                      int64_t n4, int64_t n_valid, float *__restrict__ partials, OptState *st,
                      float max_norm, float *__restrict__ total_norm_out,
                      const float *__restrict__ loss_sum, const float *__restrict__ count,
                      float *__restrict__ m_loss, int64_t mb) {
  __shared__ float red[kThreads / 32];
  __shared__ bool s_last;
  float4 *p4 = reinterpret_cast<float4 *>(p);
  const float4 *g4 = reinterpret_cast<const float4 *>(g);
  float4 *m4 = reinterpret_cast<float4 *>(m);
  float4 *v4 = reinterpret_cast<float4 *>(v);
  const int64_t stride = (int64_t)gridDim.x * kThreads;
  const int64_t i0 = (int64_t)blockIdx.x * kThreads + threadIdx.x;
  const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
  float4 gg[kCache], pp[kCache], mm[kCache], vv[kCache];
  float acc = 0.0f;
#pragma unroll
  for (int c = 0; c < kCache; ++c) {
    const int64_t i = i0 + c * stride;
    gg[c] = i < n4 ? __ldcg(g4 + i) : zero4;
    float4 x = gg[c];
    if (4 * i + 3 >= n_valid) { // the padding behind the last parameter is not part of the norm
      x.x = 4 * i < n_valid ? x.x : 0.0f;
      x.y = 4 * i + 1 < n_valid ? x.y : 0.0f;
      x.z = 4 * i + 2 < n_valid ? x.z : 0.0f;
      x.w = 0.0f;
    }
    acc += x.x * x.x + x.y * x.y + x.z * x.z + x.w * x.w;
  }
  const long long step0 = *reinterpret_cast<volatile long long *>(&st->step);
  acc = block_sum<float, kThreads>(acc, red);
  if (threadIdx.x == 0) {
    partials[blockIdx.x] = acc;
    __threadfence();
    s_last = atomicAdd(&st->done_count, 1u) == gridDim.x - 1;
  }
#pragma unroll
  for (int c = 0; c < kCache; ++c) { // in flight while this CTA waits at the barrier
    const int64_t i = i0 + c * stride;
    pp[c] = i < n4 ? p4[i] : zero4;
    mm[c] = i < n4 ? m4[i] : zero4;
    vv[c] = i < n4 ? v4[i] : zero4;
  }
  __syncthreads();
  if (s_last) {
    if (threadIdx.x < 32) {
      __threadfence();
      double s = 0.0;
      for (int i = threadIdx.x; i < (int)gridDim.x; i += 32)
        s += (double)__ldcg(partials + i);
      s = warp_sum(s);
      if (threadIdx.x == 0) {
        const float total = (float)sqrt(s);
        const float coef = (float)((double)max_norm / ((double)total + 1e-6)); // train.cc:39
        st->coef = fminf(coef, 1.0f);                                          // train.cc:40-41
        const double t = (double)(step0 + 1); // Adam's counter counts minibatch updates (train.h:129)
        const double bc1 = 1.0 - pow(0.9, t), bc2 = 1.0 - pow(0.999, t);
        st->step_size = (float)(st->lr / bc1);
        st->bc2_sqrt = (float)sqrt(bc2);
        st->done_count = 0u;
        if (total_norm_out)
          *total_norm_out = total; // the PRE-clip norm (train.h:104)
        __threadfence();
        *reinterpret_cast<volatile long long *>(&st->step) = step0 + 1; // releases the other CTAs
      }
    }
  } else if (threadIdx.x == 0) {
    while (*reinterpret_cast<volatile long long *>(&st->step) == step0)
      __nanosleep(40);
    __threadfence();
  }
  __syncthreads();
  const float coef = __ldcg(&st->coef), step_size = __ldcg(&st->step_size), bc2_sqrt = __ldcg(&st->bc2_sqrt);
  const float b1 = 0.9f, b2 = 0.999f, omb1 = (float)(1.0 - 0.9), omb2 = (float)(1.0 - 0.999);
  const float eps = 1e-5f;
#pragma unroll
  for (int c = 0; c < kCache; ++c) {
    const int64_t i = i0 + c * stride;
    if (i >= n4)
      continue;
    float *pa = &pp[c].x, *ga = &gg[c].x, *ma = &mm[c].x, *va = &vv[c].x;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float gc = ga[k] * coef; // train.cc:42-44 (the scaled gradient is consumed here only)
      ma[k] = ma[k] * b1 + gc * omb1;
      va[k] = va[k] * b2 + (gc * gc) * omb2;
      const float denom = sqrtf(va[k]) / bc2_sqrt + eps;
      pa[k] = pa[k] - step_size * (ma[k] / denom);
    }
    p4[i] = pp[c];
    m4[i] = mm[c];
    v4[i] = vv[c];
    if (shadow) {
      if constexpr (sizeof(ST) == 2) {
        __nv_bfloat162 lo = __floats2bfloat162_rn(pp[c].x, pp[c].y), hi = __floats2bfloat162_rn(pp[c].z, pp[c].w);
        uint2 pk;
        pk.x = *reinterpret_cast<uint32_t *>(&lo);
        pk.y = *reinterpret_cast<uint32_t *>(&hi);
        reinterpret_cast<uint2 *>(shadow)[i] = pk;
      } else {
        reinterpret_cast<float4 *>(shadow)[i] = make_float4(round_tf32(pp[c].x), round_tf32(pp[c].y),
                                                            round_tf32(pp[c].z), round_tf32(pp[c].w));
      }
      if (c1w_perm && i < 32 * 256 / 4) { // conv1 weights also in the space-to-depth k' order (see adam_clip_kernel)
        const int idx = (int)i * 4, co = idx >> 8, k = idx & 255;
        const int ch = k >> 6, ky = (k >> 3) & 7, kx = k & 7;
        ST *o = c1w_perm + co * 256 + ((ky >> 2) * 2 + (kx >> 2)) * 64 + (ky & 3) * 16 + (kx & 3) * 4 + ch;
        if constexpr (sizeof(ST) == 2) {
          o[0] = __float2bfloat16_rn(pp[c].x);
          o[4] = __float2bfloat16_rn(pp[c].y);
          o[8] = __float2bfloat16_rn(pp[c].z);
          o[12] = __float2bfloat16_rn(pp[c].w);
        } else {
          o[0] = round_tf32(pp[c].x);
          o[4] = round_tf32(pp[c].y);
          o[8] = round_tf32(pp[c].z);
          o[12] = round_tf32(pp[c].w);
        }
      }
    }
  }
  if (m_loss) { // Metrics.loss[epoch,k,:] = the scalar loss broadcast over mb (train.h:98); losses.cc:19
    const float loss = loss_sum[0] / count[0];
    for (int64_t i = i0; i < mb; i += stride)
      m_loss[i] = loss;
  }
}

// ---- multi-GPU: gradient exchange + norm + clip + Adam in ONE kernel over NVLink peer memory ----------
// Two-shot all-reduce by PULL, fused into the optimiser step (replaces ncclAllReduce x2 + grad_sumsq +
// adam_clip + finalize_metrics per minibatch):
//   0. announce "my local gradient of step e is complete" in every peer's signal block; wait for all.
//   1. reduce-scatter: this rank sums ITS slice over all ranks' raw gradients (16-byte volatile loads
//      over NVLink, ranks added in index order -> bit-reproducible, identical on every rank), stores it
//      in its exchange buffer and accumulates the slice's sum of squares; grid barrier; the last CTA
//      publishes the slice's sum of squares and "slice e reduced" to every peer.
//   2. wait for all slices; total norm = sqrt(sum of the per-rank partials, rank order) — the same
//      bits on every rank; clip coefficient and bias corrections.
//   3. all-gather fused into Adam's gradient load: element i's gradient comes from its owner's
//      exchange buffer (local for the own slice, a volatile NVLink load otherwise); p / m / v were
//      prefetched before step 2's wait.  Parameters stay replicated and bit-identical on all ranks.
// Reuse hazards are closed by the protocol itself: a peer overwrites its raw gradient (next backward) only
// after it passed step 2, i.e. after this rank published its slice, i.e. after this rank finished reading;
// an exchange buffer is rewritten in step 1 of e+1, after every peer announced e+1, i.e. finished e.
__device__ __forceinline__ float4 ld_peer_f4(const float *p) {
  float4 r;
  asm volatile("ld.volatile.global.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p)
               : "memory");
  return r;
}
__device__ __forceinline__ unsigned long long ld_flag(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_flag(unsigned long long *p, unsigned long long v) {
  asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// spin until *p >= e (bounded: ~seconds; a missing peer raises the error flag instead of hanging the GPU)
__device__ __forceinline__ void wait_flag(const unsigned long long *p, unsigned long long e, int *err) {
  for (long long spins = 0; ld_flag(p) < e; ++spins) {
    __nanosleep(64);
    if (spins > (1ll << 25)) {
      if (err)
        atomicExch(err, 2);
      break;
    }
  }
}

template <class ST>
__global__ void __launch_bounds__(kThreads, 2)
    adam_fused_allreduce_kernel(float *__restrict__ p, float *__restrict__ m, float *__restrict__ v,
                                ST *__restrict__ shadow, ST *__restrict__ c1w_perm, int64_t n4,
                                int64_t n_valid, float *__restrict__ partials, OptState *st,
                                float max_norm, float *__restrict__ total_norm_out,
                                const float *__restrict__ count, float *__restrict__ m_loss, int64_t mb,
                                const ArParams ar) {
  __shared__ float red[kThreads / 32];
  __shared__ bool s_last;
  __shared__ float s_coef[3];
  float4 *p4 = reinterpret_cast<float4 *>(p);
  float4 *m4 = reinterpret_cast<float4 *>(m);
  float4 *v4 = reinterpret_cast<float4 *>(v);
  const int64_t stride = (int64_t)gridDim.x * kThreads;
  const int64_t i0 = (int64_t)blockIdx.x * kThreads + threadIdx.x;
  const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
  const long long step0 = *reinterpret_cast<volatile long long *>(&st->step);
  const unsigned long long e = (unsigned long long)(step0 + 1);
  ArSignals *mine = ar.sig[ar.rank];
  // ---- 0: announce, wait for every rank's gradient
  if (blockIdx.x == 0 && threadIdx.x < ar.world) {
    __threadfence_system();
    st_flag(&ar.sig[threadIdx.x]->ready[ar.rank], e);
  }
  if (threadIdx.x < ar.world)
    wait_flag(&mine->ready[threadIdx.x], e, ar.error_flag);
  __syncthreads();
  __threadfence_system();
  // ---- 1: reduce-scatter of this rank's slice
  const int64_t lo4 = ar.slice4 * ar.rank;
  const int64_t hi4 = lo4 + ar.slice4 < n4 ? lo4 + ar.slice4 : n4;
  float4 *xmine = reinterpret_cast<float4 *>(ar.xbuf[ar.rank]);
  float acc = 0.0f;
  for (int64_t i = lo4 + i0; i < hi4; i += stride) {
    // all ranks' loads are issued before the first one is consumed (NVLink round trips overlap instead
    // of adding up), then summed in fixed rank order: the same bits whoever computes the slice
    float4 xr[kArMaxRanks];
#pragma unroll
    for (int r = 0; r < kArMaxRanks; ++r)
      if (r < ar.world)
        xr[r] = ld_peer_f4(ar.grads[r] + 4 * i);
    float4 s = zero4;
#pragma unroll
    for (int r = 0; r < kArMaxRanks; ++r)
      if (r < ar.world) {
        s.x += xr[r].x; s.y += xr[r].y; s.z += xr[r].z; s.w += xr[r].w;
      }
    xmine[i - lo4] = s;
    float4 x = s;
    if (4 * i + 3 >= n_valid) { // the padding behind the last parameter (the loss sum) is not part of the norm
      x.x = 4 * i < n_valid ? x.x : 0.0f;
      x.y = 4 * i + 1 < n_valid ? x.y : 0.0f;
      x.z = 4 * i + 2 < n_valid ? x.z : 0.0f;
      x.w = 0.0f;
    }
    acc += x.x * x.x + x.y * x.y + x.z * x.z + x.w * x.w;
  }
  acc = block_sum<float, kThreads>(acc, red);
  if (threadIdx.x == 0) {
    partials[blockIdx.x] = acc;
    __threadfence_system(); // this CTA's slice stores before its arrival
    s_last = atomicAdd(&st->done_count, 1u) == gridDim.x - 1;
  }
  // p / m / v of this thread's Adam elements: in flight across the waits below
  float4 pp[kCache], mm[kCache], vv[kCache];
#pragma unroll
  for (int c = 0; c < kCache; ++c) {
    const int64_t i = i0 + c * stride;
    pp[c] = i < n4 ? p4[i] : zero4;
    mm[c] = i < n4 ? m4[i] : zero4;
    vv[c] = i < n4 ? v4[i] : zero4;
  }
  __syncthreads();
  if (s_last && threadIdx.x < 32) { // every CTA of this rank has stored its part of the slice
    __threadfence();
    double s = 0.0;
    for (int i = threadIdx.x; i < (int)gridDim.x; i += 32)
      s += (double)__ldcg(partials + i);
    s = warp_sum(s);
    if (threadIdx.x == 0)
      st->done_count = 0u;
    if (threadIdx.x < ar.world) {
      ar.sig[threadIdx.x]->partial[ar.rank] = s;
      __threadfence_system();
      st_flag(&ar.sig[threadIdx.x]->reduced[ar.rank], e);
    }
  }
  // ---- 2: all slices reduced; the global norm from the per-rank partials (rank order)
  if (threadIdx.x < ar.world)
    wait_flag(&mine->reduced[threadIdx.x], e, ar.error_flag);
  __syncthreads();
  __threadfence_system();
  if (threadIdx.x == 0) {
    double s = 0.0;
    for (int r = 0; r < ar.world; ++r)
      s += *reinterpret_cast<volatile double *>(&mine->partial[r]);
    const float total = (float)sqrt(s);
    const float coef = fminf((float)((double)max_norm / ((double)total + 1e-6)), 1.0f); // train.cc:39-41
    const double t = (double)(step0 + 1);
    const double bc1 = 1.0 - pow(0.9, t), bc2 = 1.0 - pow(0.999, t);
    s_coef[0] = coef;
    s_coef[1] = (float)(*reinterpret_cast<volatile double *>(&st->lr) / bc1);
    s_coef[2] = (float)sqrt(bc2);
    if (blockIdx.x == 0) { // every CTA derives the same three numbers; one records them
      st->coef = s_coef[0];
      st->step_size = s_coef[1];
      st->bc2_sqrt = s_coef[2];
      if (total_norm_out)
        *total_norm_out = total; // the PRE-clip norm of the GLOBAL gradient (train.h:104)
    }
  }
  __syncthreads();
  const float coef = s_coef[0], step_size = s_coef[1], bc2_sqrt = s_coef[2];
  const float b1 = 0.9f, b2 = 0.999f, omb1 = (float)(1.0 - 0.9), omb2 = (float)(1.0 - 0.999);
  const float eps = 1e-5f;
  // ---- 3: all-gather fused into Adam's gradient load (every pull issued before the first is used)
  float4 gq[kCache];
#pragma unroll
  for (int c = 0; c < kCache; ++c) {
    const int64_t i = i0 + c * stride;
    gq[c] = zero4;
    if (i < n4) {
      const int owner = (int)(i / ar.slice4);
      const float *src = ar.xbuf[owner] + 4 * (i - ar.slice4 * owner);
      gq[c] = owner == ar.rank ? *reinterpret_cast<const float4 *>(src) : ld_peer_f4(src);
    }
  }
#pragma unroll
  for (int c = 0; c < kCache; ++c) {
    const int64_t i = i0 + c * stride;
    if (i >= n4)
      continue;
    float *pa = &pp[c].x, *ga = &gq[c].x, *ma = &mm[c].x, *va = &vv[c].x;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float gc = ga[k] * coef;
      ma[k] = ma[k] * b1 + gc * omb1;
      va[k] = va[k] * b2 + (gc * gc) * omb2;
      const float denom = sqrtf(va[k]) / bc2_sqrt + eps;
      pa[k] = pa[k] - step_size * (ma[k] / denom);
    }
    p4[i] = pp[c];
    m4[i] = mm[c];
    v4[i] = vv[c];
    if (shadow) {
      if constexpr (sizeof(ST) == 2) {
        __nv_bfloat162 lo = __floats2bfloat162_rn(pp[c].x, pp[c].y), hi = __floats2bfloat162_rn(pp[c].z, pp[c].w);
        uint2 pk;
        pk.x = *reinterpret_cast<uint32_t *>(&lo);
        pk.y = *reinterpret_cast<uint32_t *>(&hi);
        reinterpret_cast<uint2 *>(shadow)[i] = pk;
      } else {
        reinterpret_cast<float4 *>(shadow)[i] = make_float4(round_tf32(pp[c].x), round_tf32(pp[c].y),
                                                            round_tf32(pp[c].z), round_tf32(pp[c].w));
      }
      if (c1w_perm && i < 32 * 256 / 4) {
        const int idx = (int)i * 4, co = idx >> 8, k = idx & 255;
        const int ch = k >> 6, ky = (k >> 3) & 7, kx = k & 7;
        ST *o = c1w_perm + co * 256 + ((ky >> 2) * 2 + (kx >> 2)) * 64 + (ky & 3) * 16 + (kx & 3) * 4 + ch;
        if constexpr (sizeof(ST) == 2) {
          o[0] = __float2bfloat16_rn(pp[c].x);
          o[4] = __float2bfloat16_rn(pp[c].y);
          o[8] = __float2bfloat16_rn(pp[c].z);
          o[12] = __float2bfloat16_rn(pp[c].w);
        } else {
          o[0] = round_tf32(pp[c].x);
          o[4] = round_tf32(pp[c].y);
          o[8] = round_tf32(pp[c].z);
          o[12] = round_tf32(pp[c].w);
        }
      }
    }
  }
  if (m_loss) { // Metrics.loss: the GLOBAL loss sum rode in the first padding slot of the gradient
    const int64_t li = n_valid, l4 = li / 4;
    const int owner = (int)(l4 / ar.slice4);
    const float *src = ar.xbuf[owner] + (li - 4 * ar.slice4 * owner);
    const float loss = *reinterpret_cast<const volatile float *>(src) / count[0];
    for (int64_t i = i0; i < mb; i += stride)
      m_loss[i] = loss;
  }
  // the step counter advances once every CTA has read it (all CTAs arrived before `reduced` was published)
  if (blockIdx.x == 0 && threadIdx.x == 0)
    *reinterpret_cast<volatile long long *>(&st->step) = step0 + 1;
}

__global__ void f32_to_bf16_kernel(const float *__restrict__ src, __nv_bfloat16 *__restrict__ dst,
                                   int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    dst[i] = __float2bfloat16_rn(src[i]);
}

} // namespace

int launch_grad_sumsq(const float *grads, int64_t n_padded, int64_t n_valid, float *partials,
                      OptState *st, float max_norm, float *total_norm_out, cudaStream_t s) {
  grad_sumsq_kernel<<<kNormBlocks, kThreads, 0, s>>>(grads, n_padded / 4, n_valid, partials, st,
                                                     max_norm, total_norm_out);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}

int launch_adam_clip(float *params, float *grads, float *m, float *v, void *shadow,
                     void *c1w_perm, int shadow_elem_bytes, int64_t n_padded, const OptState *st,
                     cudaStream_t s) {
  const int blocks = (int)std::min<int64_t>(ceil_div(n_padded / 4, kThreads), kNumSMs * 8);
  if (shadow && shadow_elem_bytes == 4)
    adam_clip_kernel<float><<<blocks, kThreads, 0, s>>>(params, grads, m, v, static_cast<float *>(shadow),
                                                        static_cast<float *>(c1w_perm), n_padded / 4, st);
  else
    adam_clip_kernel<__nv_bfloat16><<<blocks, kThreads, 0, s>>>(
        params, grads, m, v, static_cast<__nv_bfloat16 *>(shadow),
        static_cast<__nv_bfloat16 *>(c1w_perm), n_padded / 4, st);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}

bool adam_fused_fits(int64_t n_padded) { return n_padded / 4 <= (int64_t)kCache * kNormBlocks * kThreads; }

int launch_adam_fused(float *params, float *grads, float *m, float *v, void *shadow, void *c1w_perm,
                      int shadow_elem_bytes, int64_t n_padded, int64_t n_valid, float *partials,
                      OptState *st, float max_norm, float *total_norm_out, const float *loss_sum,
                      const float *count, float *m_loss, int64_t mb, cudaStream_t s) {
  ALEGPU_REQUIRE(adam_fused_fits(n_padded), ALEGPU_RUNTIME_ERROR, "adam_fused: buffer exceeds the register cache");
  if (shadow && shadow_elem_bytes == 4)
    adam_fused_kernel<float><<<kNormBlocks, kThreads, 0, s>>>(
        params, grads, m, v, static_cast<float *>(shadow), static_cast<float *>(c1w_perm), n_padded / 4,
        n_valid, partials, st, max_norm, total_norm_out, loss_sum, count, m_loss, mb);
  else
    adam_fused_kernel<__nv_bfloat16><<<kNormBlocks, kThreads, 0, s>>>(
        params, grads, m, v, static_cast<__nv_bfloat16 *>(shadow), static_cast<__nv_bfloat16 *>(c1w_perm),
        n_padded / 4, n_valid, partials, st, max_norm, total_norm_out, loss_sum, count, m_loss, mb);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}

int launch_adam_fused_allreduce(float *params, float *m, float *v, void *shadow, void *c1w_perm,
                                int shadow_elem_bytes, int64_t n_padded, int64_t n_valid,
                                float *partials, OptState *st, float max_norm, float *total_norm_out,
                                const float *count, float *m_loss, int64_t mb, const ArParams &ar,
                                cudaStream_t s) {
  ALEGPU_REQUIRE(adam_fused_fits(n_padded), ALEGPU_RUNTIME_ERROR,
                 "adam_fused_allreduce: buffer exceeds the register cache");
  if (shadow && shadow_elem_bytes == 4)
    adam_fused_allreduce_kernel<float><<<kNormBlocks, kThreads, 0, s>>>(
        params, m, v, static_cast<float *>(shadow), static_cast<float *>(c1w_perm), n_padded / 4, n_valid,
        partials, st, max_norm, total_norm_out, count, m_loss, mb, ar);
  else
    adam_fused_allreduce_kernel<__nv_bfloat16><<<kNormBlocks, kThreads, 0, s>>>(
        params, m, v, static_cast<__nv_bfloat16 *>(shadow), static_cast<__nv_bfloat16 *>(c1w_perm),
        n_padded / 4, n_valid, partials, st, max_norm, total_norm_out, count, m_loss, mb, ar);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}

int launch_f32_to_bf16(const float *src, void *dst, int64_t n, cudaStream_t s) {
  f32_to_bf16_kernel<<<kNumSMs * 2, 256, 0, s>>>(src, static_cast<__nv_bfloat16 *>(dst), n);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}

} // namespace alegpu
