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
#include <algorithm>

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

// ---- whole rollout, persistent + bulk-async pipeline ----------------------------------------------
// plane p of slot s holds the frame of step s-1-p unless a reset-fill happened in between; so the
// tile of step t goes to slot t+1+d plane d (d = 0..3) until a later reset overwrites the stack,
// and if step t itself was a reset it also fills the older planes d..3.
// Work items = (frame pair, band group), as in the step kernel, but each CTA stays
// resident and walks items blockIdx.x, +gridDim.x, ...: one thread streams the two contiguous
// 16 800-byte input spans of the NEXT item into shared memory with cp.async.bulk (TMA bulk copy,
// completion on an mbarrier) as soon as the current one has been reduced to column sums, so the
// copy overlaps the horizontal pass and the stores and costs no registers.  4 CTAs per SM keep up to
// 134 KB of loads in flight per SM and 32 warps to hide the shared-memory latencies.
constexpr int kSpan = 35 * kRowBytes; // 16 800 bytes of one frame per band group

struct PipeSmem {
  alignas(128) uint8_t in[2][kSpan]; // [frame a|b]
  alignas(16) uint16_t sums[kOutRowsPerGroup][kRowBytes];
  alignas(16) uint8_t out[kGroupOut + 8];
  alignas(8) uint64_t full;
};

__device__ __forceinline__ uint32_t smem_addr(const void *p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void bulk_load(uint32_t dst, const void *src, uint32_t bytes,
                                          uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ uint4 lds16(uint32_t addr) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "r"(addr));
  return r;
}
// x / 3 correctly rounded for the finite operand set of this kernel (integer sums <= 2295 and their
// thirds): q0 = x*(1/3), one residual correction; verified exhaustively against IEEE division.
__device__ __forceinline__ float div3_exact(float x) {
  const float r3 = 0.333333343267440796f; // fl(1/3)
  const float q0 = __fmul_rn(x, r3);
  return __fmaf_rn(__fmaf_rn(-3.0f, q0, x), r3, q0);
}

// slots 0..3 inherit planes from the stack carried over from the previous rollout (task = (env, group))
__device__ __forceinline__ void carry_in_task(const uint8_t *__restrict__ carry_in,
                                              const uint8_t *__restrict__ starts,
                                              uint8_t *__restrict__ carry_out,
                                              uint8_t *__restrict__ obs, int64_t T, int64_t e, int g,
                                              int tid) {
  const uint8_t *src = carry_in + e * kStack + kGroupOut * g;
  const uint8_t *st = starts + e * T;
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

__global__ void __launch_bounds__(kThreads, 4)
    preprocess_rollout_pipe_kernel(const uint8_t *__restrict__ frames,
                                   const uint8_t *__restrict__ starts,
                                   const uint8_t *__restrict__ carry_in,
                                   uint8_t *__restrict__ carry_out, uint8_t *__restrict__ obs,
                                   int64_t E, int64_t T) {
  extern __shared__ __align__(128) uint8_t pipe_raw[];
  PipeSmem &sm = *reinterpret_cast<PipeSmem *>(pipe_raw);
  const int tid = threadIdx.x;
  const uint32_t nitems = (uint32_t)(E * T * kGroups), E32 = (uint32_t)E; // < 2^31 (checked by the host)
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(&sm.full)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  auto issue = [&](uint32_t item) {
    const uint32_t et = item / kGroups, g = item - et * kGroups;
    const uint8_t *fa = frames + (size_t)et * 2 * kFrameBytes + (size_t)g * kSpan;
    const uint32_t bar = smem_addr(&sm.full);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar),
                 "r"(2u * kSpan)
                 : "memory");
    bulk_load(smem_addr(&sm.in[0][0]), fa, kSpan, bar);
    bulk_load(smem_addr(&sm.in[1][0]), fa + kFrameBytes, kSpan, bar);
  };
  uint32_t item = blockIdx.x;
  if (tid == 0 && item < nitems)
    issue(item);
  // while the first item streams in: the planes inherited from the previous rollout (they never
  // coincide with a plane written from this rollout's frames, so no ordering is needed)
  for (uint32_t task = blockIdx.x; task < E32 * kGroups; task += gridDim.x)
    carry_in_task(carry_in, starts, carry_out, obs, T, task / kGroups, (int)(task % kGroups), tid);
  // phase-2 geometry of this thread: one output column, rows rg, rg+3, ...
  const int ox = tid % 84, rg = tid / 84;
  const int x0 = (ox * 160) / 84, x1 = ((ox + 1) * 160 + 83) / 84;
  const bool three = (x1 - x0) == 3;
  for (int k = 0; item < nitems; ++k, item += gridDim.x) {
    {
      const uint32_t bar = smem_addr(&sm.full), parity = (uint32_t)k & 1u;
      asm volatile("{\n\t.reg .pred p;\n\tW_%=:\n\t"
                   "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
                   "@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}" ::"r"(bar),
                   "r"(parity)
                   : "memory");
    }
    // ---- phase 1: 2-frame max + vertical 3-row sums, 16 bytes of one band per thread
    if (tid < kBandsPerGroup * 30) {
      const int band = tid / 30, chunk = tid % 30;
      const uint32_t off = (uint32_t)(5 * band * kRowBytes + 16 * chunk);
      const uint32_t pa = smem_addr(&sm.in[0][0]) + off, pb = smem_addr(&sm.in[1][0]) + off;
      // packed u16 column sums in natural byte order: lo = (b0,b1), hi = (b2,b3) of each word
      uint32_t top_lo[4], top_hi[4], bot_lo[4], bot_hi[4];
#pragma unroll
      for (int w = 0; w < 4; ++w)
        top_lo[w] = top_hi[w] = bot_lo[w] = bot_hi[w] = 0u;
#pragma unroll
      for (int r = 0; r < 5; ++r) {
        const uint4 a = lds16(pa + r * kRowBytes), b = lds16(pb + r * kRowBytes);
        const uint32_t wa[4] = {a.x, a.y, a.z, a.w};
        const uint32_t wb[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
        for (int w = 0; w < 4; ++w) {
          const uint32_t m = __vmaxu4(wa[w], wb[w]); // max_and_skip.cc:36-38, exact
          const uint32_t lo = __byte_perm(m, 0u, 0x4140), hi = __byte_perm(m, 0u, 0x4342);
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
      uint4 *dt = reinterpret_cast<uint4 *>(&sm.sums[2 * band][16 * chunk]);
      uint4 *db = reinterpret_cast<uint4 *>(&sm.sums[2 * band + 1][16 * chunk]);
      dt[0] = make_uint4(top_lo[0], top_hi[0], top_lo[1], top_hi[1]);
      dt[1] = make_uint4(top_lo[2], top_hi[2], top_lo[3], top_hi[3]);
      db[0] = make_uint4(bot_lo[0], bot_hi[0], bot_lo[1], bot_hi[1]);
      db[1] = make_uint4(bot_lo[2], bot_hi[2], bot_lo[3], bot_hi[3]);
    }
    __syncthreads();
    { // the input buffer is consumed: stream the next item in while phases 2-3 run
      const uint32_t next = item + gridDim.x;
      if (tid == 0 && next < nitems)
        issue(next);
    }
    // ---- phase 2: horizontal bins, mean, luminance, round to u8 (bit-identical to the oracle)
    if (tid < 252) {
      // branch-free in the bin width (2 or 3 columns) so that a warp never diverges
      const uint16_t *col = &sm.sums[rg][3 * x0];
      const int o3 = three ? 6 : 0; // a 2-wide bin re-reads column 0 and masks it out
      const uint32_t keep = three ? 0xFFFFFFFFu : 0u;
#pragma unroll
      for (int j = 0; j < 5; ++j) {
        const int orow = rg + 3 * j;
        if (orow < kOutRowsPerGroup) {
          const uint16_t *row = col + j * 3 * kRowBytes;
          const uint32_t s0 = (uint32_t)row[0] + row[3] + (row[o3] & keep);
          const uint32_t s1 = (uint32_t)row[1] + row[4] + (row[o3 + 1] & keep);
          const uint32_t s2 = (uint32_t)row[2] + row[5] + (row[o3 + 2] & keep);
          float m0 = div3_exact((float)s0), m1 = div3_exact((float)s1), m2 = div3_exact((float)s2);
          const float h0 = __fmul_rn(m0, 0.5f), h1 = __fmul_rn(m1, 0.5f), h2 = __fmul_rn(m2, 0.5f);
          const float t0 = div3_exact(m0), t1 = div3_exact(m1), t2 = div3_exact(m2);
          m0 = three ? t0 : h0;
          m1 = three ? t1 : h1;
          m2 = three ? t2 : h2;
          float gr = __fmul_rn(0.2125f, m0);
          gr = __fadd_rn(gr, __fmul_rn(0.7154f, m1));
          gr = __fadd_rn(gr, __fmul_rn(0.0721f, m2)); // vision.cc:80, no fma contraction
          // 0 <= gr <= 255 (the weights sum to 1), so floor(gr + .5) needs no clamp
          sm.out[orow * 84 + ox] = (uint8_t)__float2uint_rd(__fadd_rn(gr, 0.5f));
        }
      }
    }
    __syncthreads();
    // ---- phase 3: the tile of step t is plane d of slot t+1+d until a reset overwrites the stack
    {
      const uint32_t et = item / kGroups, g = item - et * kGroups;
      const uint32_t t32 = et / E32;
      const int64_t t = t32, e = et - t32 * E32;
      const uint8_t *st = starts + e * T;
      const bool reset_t = st[t] != 0;
      for (int d = 0; d < 4; ++d) {
        const int64_t slot = t + 1 + d;
        if (slot > T)
          break;
        if (d >= 1 && st[t + d])
          break;
        uint8_t *base = (slot == T) ? carry_out + e * kStack : obs + (size_t)(e * T + slot) * kStack;
        const int nplanes = reset_t ? 4 - d : 1;
        for (int i = tid; i < nplanes * 147; i += kThreads) {
          const int pl = d + i / 147, l = i % 147;
          const uint2 v = *reinterpret_cast<const uint2 *>(&sm.out[8 * l]);
          st_stream_u2(base + pl * kPlane + kGroupOut * g + 8 * l, v);
        }
      }
    }
    // the next iteration's phase 1 only writes sm.sums (last read before the barrier above) and its
    // phase 2 writes sm.out after another barrier, so no extra synchronisation is needed here
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
  {
    constexpr int kPipeSmem = (int)sizeof(PipeSmem) + 128;
    static bool configured = false;
    if (!configured) {
      ALEGPU_CUDA_CHECK(cudaFuncSetAttribute(preprocess_rollout_pipe_kernel,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, kPipeSmem));
      configured = true;
    }
    const int64_t items = E * T * kGroups;
    const unsigned grid = (unsigned)std::min<int64_t>(items, 4 * kNumSMs);
    preprocess_rollout_pipe_kernel<<<grid, kThreads, kPipeSmem, as_stream(stream)>>>(
        frames, episode_starts, carry_in, carry_out, obs, E, T);
  }
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}
