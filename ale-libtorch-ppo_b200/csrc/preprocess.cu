// preprocess.cu — fused Atari observation preprocessing for sm_100a.
//
// One kernel does what the reference does in four places:
//   2-frame max            src/ai/environment/max_and_skip.cc:33-42
//   area resize 210x160->84x84 per channel   src/ai/vision.cc:8-10, 38-48 (interpolate(mode=area)
//                          == adaptive average pooling: bin [floor(o*in/out), ceil((o+1)*in/out)) )
//   RGB -> gray            src/ai/vision.cc:51 (0.2125, 0.7154, 0.0721), :71-84
//   frame-stack shift / reset-fill / plane-0 insert   src/ai/rollout.cc:184-196
// plus the Buffer::add observation copy (src/ai/buffer.cc:47) which disappears because the
// stack is written straight into rollout slot t+1.
//
// HBM-bound: 2 x 100 800 B read per env step, 28 224 B written (+21 168 B of carried planes
// read in step mode).  Geometry: 210/84 = 5/2, so output rows (2j, 2j+1) depend exactly on
// input rows 5j..5j+4 (row bins {5j..5j+2}, {5j+2..5j+4}); a CTA owns 7 such bands =
// 35 input rows = one contiguous 16 800 B span of each frame, read with 16-byte streaming
// loads (10 in flight per thread), max'ed with vmaxu4, column-summed with packed-u16
// arithmetic in registers and staged in shared memory for the horizontal pass.
#include "common.cuh"

namespace alegpu {

namespace {

constexpr int kFrameBytes = 210 * 160 * 3; // 100 800
constexpr int kRowBytes = 160 * 3;         // 480
constexpr int kPlane = 84 * 84;            // 7 056
constexpr int kStack = 4 * kPlane;         // 28 224
constexpr int kGroups = 6;                 // band groups per frame
constexpr int kBandsPerGroup = 7;
constexpr int kOutRowsPerGroup = 14;
constexpr int kGroupOut = kOutRowsPerGroup * 84; // 1 176 output bytes per plane per group
constexpr int kThreads = 256;

struct GroupSmem {
  // vertical sums of max(a,b): [14 output rows][480 byte-columns] u16
  alignas(16) uint16_t sums[kOutRowsPerGroup][kRowBytes];
  alignas(16) uint8_t out[kGroupOut];
};

// Phase 1+2: compute the 14x84 u8 tile of band-group g of one frame pair into sm.out.
__device__ __forceinline__ void gray_tile(const uint8_t *__restrict__ fa,
                                          const uint8_t *__restrict__ fb, int g, GroupSmem &sm,
                                          float *__restrict__ gray_f32_plane) {
  const int tid = threadIdx.x;
  if (tid < kBandsPerGroup * 30) {
    const int band = tid / 30, chunk = tid % 30;
    const size_t off = (size_t)(35 * g + 5 * band) * kRowBytes + 16 * chunk;
    uint4 a[5], b[5];
#pragma unroll
    for (int r = 0; r < 5; ++r) {
      a[r] = ld_stream_u4(fa + off + r * kRowBytes);
      b[r] = ld_stream_u4(fb + off + r * kRowBytes);
    }
    // packed u16 column sums: lo = bytes 0,2 ; hi = bytes 1,3 of each 32-bit word
    uint32_t top_lo[4], top_hi[4], bot_lo[4], bot_hi[4];
#pragma unroll
    for (int w = 0; w < 4; ++w)
      top_lo[w] = top_hi[w] = bot_lo[w] = bot_hi[w] = 0u;
#pragma unroll
    for (int r = 0; r < 5; ++r) {
      const uint32_t wa[4] = {a[r].x, a[r].y, a[r].z, a[r].w};
      const uint32_t wb[4] = {b[r].x, b[r].y, b[r].z, b[r].w};
#pragma unroll
      for (int w = 0; w < 4; ++w) {
        const uint32_t m = __vmaxu4(wa[w], wb[w]); // max_and_skip.cc:36-38, exact
        const uint32_t lo = m & 0x00FF00FFu, hi = (m >> 8) & 0x00FF00FFu;
        if (r <= 2) {
          top_lo[w] += lo;
          top_hi[w] += hi;
        }
        if (r >= 2) {
          bot_lo[w] += lo;
          bot_hi[w] += hi;
        }
      }
    }
    uint32_t t[8], bt[8];
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      t[2 * w] = __byte_perm(top_lo[w], top_hi[w], 0x5410);
      t[2 * w + 1] = __byte_perm(top_lo[w], top_hi[w], 0x7632);
      bt[2 * w] = __byte_perm(bot_lo[w], bot_hi[w], 0x5410);
      bt[2 * w + 1] = __byte_perm(bot_lo[w], bot_hi[w], 0x7632);
    }
    uint4 *dt = reinterpret_cast<uint4 *>(&sm.sums[2 * band][16 * chunk]);
    uint4 *db = reinterpret_cast<uint4 *>(&sm.sums[2 * band + 1][16 * chunk]);
    dt[0] = make_uint4(t[0], t[1], t[2], t[3]);
    dt[1] = make_uint4(t[4], t[5], t[6], t[7]);
    db[0] = make_uint4(bt[0], bt[1], bt[2], bt[3]);
    db[1] = make_uint4(bt[4], bt[5], bt[6], bt[7]);
  }
  __syncthreads();
  for (int idx = tid; idx < kGroupOut; idx += kThreads) {
    const int orow = idx / 84, ox = idx - orow * 84;
    const int x0 = (ox * 160) / 84, x1 = ((ox + 1) * 160 + 83) / 84; // floor / ceil bins
    uint32_t s0 = 0, s1 = 0, s2 = 0;
    for (int x = x0; x < x1; ++x) {
      s0 += sm.sums[orow][3 * x];
      s1 += sm.sums[orow][3 * x + 1];
      s2 += sm.sums[orow][3 * x + 2];
    }
    const float nx = (float)(x1 - x0);
    // adaptive average pool: sum / rows / cols (sums of <= 9 bytes are exact in fp32)
    const float m0 = __fdiv_rn(__fdiv_rn((float)s0, 3.0f), nx);
    const float m1 = __fdiv_rn(__fdiv_rn((float)s1, 3.0f), nx);
    const float m2 = __fdiv_rn(__fdiv_rn((float)s2, 3.0f), nx);
    float gr = __fmul_rn(0.2125f, m0);
    gr = __fadd_rn(gr, __fmul_rn(0.7154f, m1));
    gr = __fadd_rn(gr, __fmul_rn(0.0721f, m2)); // vision.cc:80, no fma contraction
    if (gray_f32_plane)
      gray_f32_plane[(14 * g + orow) * 84 + ox] = gr;
    float q = floorf(__fadd_rn(gr, 0.5f));
    q = fminf(fmaxf(q, 0.0f), 255.0f);
    sm.out[idx] = (uint8_t)q;
  }
  __syncthreads();
}

// write sm.out (1 176 B) to the band-group region of one destination plane, 8-byte stores
__device__ __forceinline__ void store_tile(uint8_t *__restrict__ plane, int g,
                                           const GroupSmem &sm, int lane147) {
  if (lane147 < kGroupOut / 8) {
    const uint2 v = reinterpret_cast<const uint2 *>(sm.out)[lane147];
    st_stream_u2(plane + kGroupOut * g + 8 * lane147, v);
  }
}

// ---- one env step -----------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads)
    preprocess_step_kernel(const uint8_t *__restrict__ frames_a,
                           const uint8_t *__restrict__ frames_b, int64_t frame_stride,
                           const uint8_t *__restrict__ starts, int64_t starts_stride,
                           const uint8_t *__restrict__ src, int64_t src_stride,
                           uint8_t *__restrict__ dst, int64_t dst_stride,
                           float *__restrict__ gray_f32) {
  __shared__ GroupSmem sm;
  const int g = blockIdx.x;
  const int64_t e = blockIdx.y;
  gray_tile(frames_a + e * frame_stride, frames_b + e * frame_stride, g, sm,
            gray_f32 ? gray_f32 + e * kPlane : nullptr);
  uint8_t *d = dst + e * dst_stride;
  const int tid = threadIdx.x;
  if (starts[e * starts_stride]) { // rollout.cc:189-193: all four planes <- new frame
    for (int i = tid; i < 4 * 147; i += kThreads)
      store_tile(d + (i / 147) * kPlane, g, sm, i % 147);
  } else {
    store_tile(d, g, sm, tid); // rollout.cc:194-195
    const uint8_t *s = src + e * src_stride;
    for (int i = tid; i < 3 * 147; i += kThreads) { // rollout.cc:185-188: planes 1..3 <- 0..2
      const int p = i / 147, l = i % 147;
      const uint2 v = ld_stream_u2(s + p * kPlane + kGroupOut * g + 8 * l);
      st_stream_u2(d + (p + 1) * kPlane + kGroupOut * g + 8 * l, v);
    }
  }
}

// ---- whole rollout, frames resident ---------------------------------------------------------
// plane p of slot s holds the frame of step s-1-p unless a reset-fill happened in between;
// so the tile of step t goes to slot t+1+d plane d (d = 0..3) until a later reset overwrites
// the stack, and if step t itself was a reset it also fills the older planes d..3.
__global__ void __launch_bounds__(kThreads)
    preprocess_rollout_kernel(const uint8_t *__restrict__ frames,
                              const uint8_t *__restrict__ starts, uint8_t *__restrict__ carry_out,
                              uint8_t *__restrict__ obs, int64_t E, int64_t T) {
  __shared__ GroupSmem sm;
  const int64_t lin = blockIdx.x;
  const int g = (int)(lin % kGroups);
  const int64_t et = lin / kGroups; // time-major frame index t*E + e
  const int64_t t = et / E, e = et - t * E;
  const uint8_t *fa = frames + (size_t)et * 2 * kFrameBytes;
  gray_tile(fa, fa + kFrameBytes, g, sm, nullptr);
  const uint8_t *st = starts + e * T;
  const bool reset_t = st[t] != 0;
  const int tid = threadIdx.x;
  for (int d = 0; d < 4; ++d) {
    const int64_t s = t + 1 + d;
    if (s > T)
      break;
    if (d >= 1 && st[t + d])
      break;
    uint8_t *base = (s == T) ? carry_out + e * kStack : obs + (size_t)(e * T + s) * kStack;
    if (reset_t) {
      for (int i = tid; i < (4 - d) * 147; i += kThreads)
        store_tile(base + (d + i / 147) * kPlane, g, sm, i % 147);
    } else {
      store_tile(base + d * kPlane, g, sm, tid);
    }
  }
}

// slots 0..3 inherit planes from the stack carried over from the previous rollout
__global__ void __launch_bounds__(kThreads)
    preprocess_carry_in_kernel(const uint8_t *__restrict__ carry_in,
                               const uint8_t *__restrict__ starts, uint8_t *__restrict__ carry_out,
                               uint8_t *__restrict__ obs, int64_t T) {
  const int g = blockIdx.x;
  const int64_t e = blockIdx.y;
  const uint8_t *src = carry_in + e * kStack + kGroupOut * g;
  const uint8_t *st = starts + e * T;
  const int tid = threadIdx.x;
  for (int s = 0; s < 4; ++s) {
    if (s > T)
      break;
    if (s >= 1 && st[s - 1])
      break;
    uint8_t *base =
        ((s == T) ? carry_out + e * kStack : obs + (size_t)(e * T + s) * kStack) + kGroupOut * g;
    for (int i = tid; i < (4 - s) * 147; i += kThreads) {
      const int p = s + i / 147, l = i % 147;
      const uint2 v = ld_stream_u2(src + (p - s) * kPlane + 8 * l);
      st_stream_u2(base + p * kPlane + 8 * l, v);
    }
  }
}

} // namespace
} // namespace alegpu

using namespace alegpu;

static int preprocess_step_impl(const uint8_t *frames_a, const uint8_t *frames_b,
                                int64_t frame_stride, const uint8_t *starts,
                                int64_t starts_stride, const uint8_t *src_stack,
                                int64_t src_env_stride, uint8_t *dst_stack,
                                int64_t dst_env_stride, float *gray_f32, int64_t E,
                                void *stream) {
  ALEGPU_REQUIRE(E >= 0 && E <= 65535, ALEGPU_INVALID_ARGUMENT,
                 "preprocess_step: E must be in [0, 65535]");
  if (E == 0)
    return ALEGPU_OK;
  ALEGPU_REQUIRE(frames_a && frames_b && starts && src_stack && dst_stack,
                 ALEGPU_INVALID_ARGUMENT, "preprocess_step: null pointer");
  ALEGPU_REQUIRE(src_stack != dst_stack, ALEGPU_INVALID_ARGUMENT,
                 "preprocess_step: src and dst stacks must not alias");
  ALEGPU_REQUIRE(((uintptr_t)frames_a % 16 == 0) && ((uintptr_t)frames_b % 16 == 0) &&
                     (frame_stride % 16 == 0) && ((uintptr_t)src_stack % 8 == 0) &&
                     ((uintptr_t)dst_stack % 8 == 0) &&
                     (src_env_stride % 8 == 0) && (dst_env_stride % 8 == 0),
                 ALEGPU_INVALID_ARGUMENT,
                 "preprocess_step: frames need 16-byte, stacks/strides 8-byte alignment");
  dim3 grid(kGroups, (unsigned)E);
  preprocess_step_kernel<<<grid, kThreads, 0, as_stream(stream)>>>(
      frames_a, frames_b, frame_stride, starts, starts_stride, src_stack, src_env_stride,
      dst_stack, dst_env_stride, gray_f32);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}

extern "C" int alegpu_preprocess_step(const uint8_t *frames_a, const uint8_t *frames_b,
                                      const uint8_t *starts, int64_t starts_stride,
                                      const uint8_t *src_stack, int64_t src_env_stride,
                                      uint8_t *dst_stack, int64_t dst_env_stride,
                                      float *gray_f32, int64_t E, void *stream) {
  return preprocess_step_impl(frames_a, frames_b, kFrameBytes, starts, starts_stride, src_stack,
                              src_env_stride, dst_stack, dst_env_stride, gray_f32, E, stream);
}

extern "C" int alegpu_preprocess_step_strided(const uint8_t *frames_a, const uint8_t *frames_b,
                                              int64_t frame_env_stride, const uint8_t *starts,
                                              int64_t starts_stride, const uint8_t *src_stack,
                                              int64_t src_env_stride, uint8_t *dst_stack,
                                              int64_t dst_env_stride, int64_t E, void *stream) {
  return preprocess_step_impl(frames_a, frames_b, frame_env_stride, starts, starts_stride,
                              src_stack, src_env_stride, dst_stack, dst_env_stride, nullptr, E,
                              stream);
}

extern "C" int alegpu_preprocess_rollout(const uint8_t *frames, const uint8_t *episode_starts,
                                         const uint8_t *carry_in, uint8_t *carry_out,
                                         uint8_t *obs, int64_t E, int64_t T, void *stream) {
  ALEGPU_REQUIRE(E >= 0 && T >= 0 && E <= 65535, ALEGPU_INVALID_ARGUMENT,
                 "preprocess_rollout: bad E/T");
  if (E == 0 || T == 0)
    return ALEGPU_OK;
  ALEGPU_REQUIRE(frames && episode_starts && carry_in && carry_out && obs,
                 ALEGPU_INVALID_ARGUMENT, "preprocess_rollout: null pointer");
  ALEGPU_REQUIRE(carry_in != carry_out, ALEGPU_INVALID_ARGUMENT,
                 "preprocess_rollout: carry_in and carry_out must not alias");
  ALEGPU_REQUIRE(((uintptr_t)frames % 16 == 0) && ((uintptr_t)obs % 8 == 0) &&
                     ((uintptr_t)carry_in % 8 == 0) && ((uintptr_t)carry_out % 8 == 0),
                 ALEGPU_INVALID_ARGUMENT, "preprocess_rollout: misaligned pointer");
  ALEGPU_REQUIRE(E * T * kGroups < (int64_t)2147483647, ALEGPU_INVALID_ARGUMENT,
                 "preprocess_rollout: E*T too large for one launch");
  preprocess_carry_in_kernel<<<dim3(kGroups, (unsigned)E), kThreads, 0, as_stream(stream)>>>(
      carry_in, episode_starts, carry_out, obs, T);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  preprocess_rollout_kernel<<<(unsigned)(E * T * kGroups), kThreads, 0, as_stream(stream)>>>(
      frames, episode_starts, carry_out, obs, E, T);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}
