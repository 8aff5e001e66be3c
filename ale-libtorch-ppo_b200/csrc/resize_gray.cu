// resize_gray.cu — the LIVE observation path of the reference: per-frame grayscale resize, then the
// 2-frame max, then the frame stack.
//
//   ai::vision::resize_grayscale_image(vector<u8>, w, h, nw, nh)      src/ai/vision.cc:86-95
//     = stbir_resize_uint8_linear(..., STBIR_1CHANNEL)                 (third-party stb_image_resize2,
//       Bazel module stb@0.0.0-20241109-5c20573, MODULE.bazel:38 — header NOT in /root/reference)
//   called per emulator frame by ResizeEnvironment::resize             src/ai/environment/resize.cc:34-41
//   which sits BELOW MaxAndSkipEnvironment in the wrapper stack        src/ai/rollout.cc:146-161
//   so the live order is resize(a), resize(b) -> max                   max_and_skip.cc:33-42
//   -> Rollout::update_observations                                    src/ai/rollout.cc:184-196
//
// PARITY UNPINNED: the stb source is absent offline and the reference has no test of this overload
// (test/ai/vision-test.cc covers the libtorch overloads only).  What is built here is the published
// algorithm of stbir_resize_uint8_linear with default settings, restated:
//   * linear light (no sRGB curve), u8 -> float unscaled (0..255), edge mode CLAMP;
//   * per axis: scale = out/in; scale < 1 (downsample) -> Mitchell-Netravali cubic (B = C = 1/3)
//     evaluated in OUTPUT space, w(o,i) = M((o+.5) - (i+.5)*scale) * scale, support 2 output pixels
//     (= 2/scale input pixels); scale >= 1 -> Catmull-Rom evaluated in INPUT space,
//     w(o,i) = C((o+.5)/scale - (i+.5)), support 2 input pixels;
//   * taps that fall outside the image are folded onto the edge pixel; every output pixel's
//     weights are normalised to sum to 1;
//   * separable, fp32; result = (u8)clamp(x + 0.5, 0, 255).
// oracle/restate.py:resize_grayscale_stb is the independent numpy statement of the same rules; the
// tests also check the properties any such filter has (constant in -> constant out, same-size
// resize is the identity, output max'ed bit-exactly, stack update bit-exact).
//
// Kernel: one CTA per (band of <= 14 output rows, image).  The input rows the band needs are staged in
// shared memory as fp32 (16-byte loads, each pixel converted once), the vertical pass runs 4 columns
// per thread off one 16-byte shared load per tap into an [rows][pitch] intermediate, the horizontal
// pass reads that intermediate.  HBM-bound by construction: w*h (+ band halo, served by L2) bytes in,
// nw*nh out per image; the fused step adds the 21 168 carried + 28 224 written stack bytes.
#include <algorithm>
#include <cmath>
#include <mutex>
#include <vector>

#include "common.cuh"

namespace alegpu {
namespace {

constexpr int kBandRows = 14;
constexpr int kThreads = 256;

struct Plan { // one per (device, w, h, nw, nh); device tables live for the life of the process
  int device = 0;
  int w = 0, h = 0, nw = 0, nh = 0;
  int kv = 0, kh = 0;     // taps per output row / column (padded with zero weights)
  int pitch = 0;          // shared-memory row pitch in bytes / floats (multiple of 16)
  int max_in_rows = 0;    // most input rows any band needs
  int *vstart = nullptr;  // [nh] first tap row (may be < 0: clamped at load)
  float *vw = nullptr;    // [nh][kv]
  int *hstart = nullptr;  // [nw]
  float *hw = nullptr;    // [nw][kh]
  size_t smem = 0;        // dynamic shared memory of one CTA (two-frame form)
};

float mitchell(float x) {
  x = std::fabs(x);
  if (x < 1.0f)
    return (16.0f + x * x * (21.0f * x - 36.0f)) / 18.0f;
  if (x < 2.0f)
    return (32.0f + x * (-60.0f + x * (36.0f - 7.0f * x))) / 18.0f;
  return 0.0f;
}
float catmull_rom(float x) {
  x = std::fabs(x);
  if (x < 1.0f)
    return 1.0f - x * x * (2.5f - 1.5f * x);
  if (x < 2.0f)
    return 2.0f - x * (4.0f + x * (0.5f * x - 2.5f));
  return 0.0f;
}

// gather-form weights of one axis: start[o], taps[o][k] (normalised, zero padded), returns k
int axis_weights(int in, int out, std::vector<int> &start, std::vector<float> &taps) {
  const float scale = (float)out / (float)in;
  const bool down = scale < 1.0f;
  const float reach = down ? 2.0f / scale : 2.0f; // support in input pixels
  std::vector<std::vector<float>> rows((size_t)out);
  start.assign((size_t)out, 0);
  int kmax = 1;
  for (int o = 0; o < out; ++o) {
    const float centre = ((float)o + 0.5f) / scale; // in input space
    const int lo = (int)std::floor(centre - reach - 0.5f) - 1;
    const int hi = (int)std::ceil(centre + reach - 0.5f) + 1;
    int first = 0, last = -1;
    std::vector<float> wts;
    for (int i = lo; i <= hi; ++i) {
      float wgt;
      if (down) {
        const float x = ((float)o + 0.5f) - ((float)i + 0.5f) * scale;
        wgt = mitchell(x) * scale;
      } else {
        wgt = catmull_rom(centre - ((float)i + 0.5f));
      }
      if (wgt == 0.0f && wts.empty())
        continue;
      if (wts.empty())
        first = i;
      wts.push_back(wgt);
      if (wgt != 0.0f)
        last = (int)wts.size() - 1;
    }
    wts.resize((size_t)(last + 1));
    float total = 0.0f;
    for (float v : wts)
      total += v;
    const float inv = 1.0f / total;
    for (float &v : wts)
      v *= inv;
    start[(size_t)o] = first;
    kmax = std::max(kmax, (int)wts.size());
    rows[(size_t)o] = std::move(wts);
  }
  taps.assign((size_t)out * kmax, 0.0f);
  for (int o = 0; o < out; ++o)
    std::copy(rows[(size_t)o].begin(), rows[(size_t)o].end(), taps.begin() + (size_t)o * kmax);
  return kmax;
}

std::mutex g_plan_mu;
std::vector<Plan> g_plans;

size_t align16(size_t x) { return (x + 15) & ~(size_t)15; }

int get_plan(int w, int h, int nw, int nh, const Plan **out) {
  std::lock_guard<std::mutex> lock(g_plan_mu);
  int device = 0;
  ALEGPU_CUDA_CHECK(cudaGetDevice(&device));
  for (const Plan &p : g_plans)
    if (p.device == device && p.w == w && p.h == h && p.nw == nw && p.nh == nh) {
      *out = &p;
      return ALEGPU_OK;
    }
  Plan p;
  p.device = device;
  p.w = w, p.h = h, p.nw = nw, p.nh = nh;
  std::vector<int> vs, hs;
  std::vector<float> vt, ht;
  p.kv = axis_weights(h, nh, vs, vt);
  p.kh = axis_weights(w, nw, hs, ht);
  p.pitch = (int)align16((size_t)w);
  for (int r0 = 0; r0 < nh; r0 += kBandRows) {
    const int r1 = std::min(nh, r0 + kBandRows) - 1;
    const int lo = std::min(std::max(vs[(size_t)r0], 0), h - 1);
    const int hi = std::min(std::max(vs[(size_t)r1] + p.kv - 1, 0), h - 1);
    p.max_in_rows = std::max(p.max_in_rows, hi - lo + 1);
  }
  p.smem = align16(sizeof(float) * p.max_in_rows * p.pitch)    // staged input rows, fp32 (one frame)
           + align16(sizeof(float) * kBandRows * p.pitch)      // vertical-pass intermediate
           + 2 * align16((size_t)kBandRows * nw)               // output tiles (frame a / b)
           + align16(sizeof(float) * kBandRows * p.kv) + align16(sizeof(int) * kBandRows) +
           align16(sizeof(float) * (size_t)nw * p.kh) + align16(sizeof(int) * (size_t)nw);
  ALEGPU_REQUIRE(p.smem <= 227 * 1024, ALEGPU_INVALID_ARGUMENT,
                 "resize_grayscale: image too wide for the shared-memory band plan");
  ALEGPU_CUDA_CHECK(cudaMalloc(&p.vstart, sizeof(int) * vs.size()));
  ALEGPU_CUDA_CHECK(cudaMalloc(&p.vw, sizeof(float) * vt.size()));
  ALEGPU_CUDA_CHECK(cudaMalloc(&p.hstart, sizeof(int) * hs.size()));
  ALEGPU_CUDA_CHECK(cudaMalloc(&p.hw, sizeof(float) * ht.size()));
  ALEGPU_CUDA_CHECK(cudaMemcpy(p.vstart, vs.data(), sizeof(int) * vs.size(), cudaMemcpyHostToDevice));
  ALEGPU_CUDA_CHECK(cudaMemcpy(p.vw, vt.data(), sizeof(float) * vt.size(), cudaMemcpyHostToDevice));
  ALEGPU_CUDA_CHECK(cudaMemcpy(p.hstart, hs.data(), sizeof(int) * hs.size(), cudaMemcpyHostToDevice));
  ALEGPU_CUDA_CHECK(cudaMemcpy(p.hw, ht.data(), sizeof(float) * ht.size(), cudaMemcpyHostToDevice));
  g_plans.reserve(64); // pointers handed out stay valid
  ALEGPU_REQUIRE(g_plans.size() < 64, ALEGPU_RUNTIME_ERROR, "resize_grayscale: too many distinct shapes");
  g_plans.push_back(p);
  *out = &g_plans.back();
  return ALEGPU_OK;
}

struct DevPlan { // by-value kernel argument
  int w, h, nw, nh, kv, kh, pitch, max_in_rows;
  const int *vstart;
  const float *vw;
  const int *hstart;
  const float *hw;
};
DevPlan dev_plan(const Plan &p) {
  return DevPlan{p.w, p.h, p.nw, p.nh, p.kv, p.kh, p.pitch, p.max_in_rows,
                 p.vstart, p.vw, p.hstart, p.hw};
}

struct BandSmem { // carved out of dynamic shared memory in the order get_plan() sizes it
  float *in;
  float *inter;
  uint8_t *tile[2];
  float *vw;
  int *vstart;
  float *hw;
  int *hstart;
};
__device__ __forceinline__ size_t dalign16(size_t x) { return (x + 15) & ~(size_t)15; }
__device__ __forceinline__ BandSmem carve(uint8_t *raw, const DevPlan &p) {
  BandSmem s;
  s.in = reinterpret_cast<float *>(raw);
  raw += dalign16(sizeof(float) * p.max_in_rows * p.pitch);
  s.inter = reinterpret_cast<float *>(raw);
  raw += dalign16(sizeof(float) * kBandRows * p.pitch);
  s.tile[0] = raw;
  raw += dalign16((size_t)kBandRows * p.nw);
  s.tile[1] = raw;
  raw += dalign16((size_t)kBandRows * p.nw);
  s.vw = reinterpret_cast<float *>(raw);
  raw += dalign16(sizeof(float) * kBandRows * p.kv);
  s.vstart = reinterpret_cast<int *>(raw);
  raw += dalign16(sizeof(int) * kBandRows);
  s.hw = reinterpret_cast<float *>(raw);
  raw += dalign16(sizeof(float) * (size_t)p.nw * p.kh);
  s.hstart = reinterpret_cast<int *>(raw);
  return s;
}

// tables of band [r0, r0+rows) into shared memory (visible after the caller's next barrier)
__device__ __forceinline__ void load_tables(const DevPlan &p, const BandSmem &s, int r0, int rows) {
  for (int i = threadIdx.x; i < rows * p.kv; i += kThreads)
    s.vw[i] = p.vw[(size_t)r0 * p.kv + i];
  for (int i = threadIdx.x; i < rows; i += kThreads)
    s.vstart[i] = p.vstart[r0 + i];
  for (int i = threadIdx.x; i < p.nw * p.kh; i += kThreads) { // [nw][kh] -> [kh][nw]: lanes walk ox
    const int ox = i / p.kh, k = i - ox * p.kh;
    s.hw[k * p.nw + ox] = p.hw[i];
  }
  for (int i = threadIdx.x; i < p.nw; i += kThreads)
    s.hstart[i] = p.hstart[i];
}

// One image, one band: u8 tile [rows][nw] -> tile (shared), optional pre-rounding fp32 to global.
// Ends with a barrier; the tables must be visible on entry.
__device__ __forceinline__ void resize_band(const DevPlan &p, const BandSmem &s,
                                            const uint8_t *__restrict__ img, int r0, int rows,
                                            uint8_t *__restrict__ tile,
                                            float *__restrict__ f32_plane) {
  const int tid = threadIdx.x;
  const int lo = min(max(s.vstart[0], 0), p.h - 1);
  const int hi = min(max(s.vstart[rows - 1] + p.kv - 1, 0), p.h - 1);
  const int n_in = hi - lo + 1;
  // ---- stage input rows lo..hi as fp32 (each pixel is converted once, not once per tap)
  const uint8_t *src = img + (size_t)lo * p.w;
  if ((p.w & 15) == 0 && ((uintptr_t)src & 3) == 0) { // pitch == w: one contiguous span
    const int words = n_in * (p.w >> 2); // a 32-bit load and a conflict-free 16-byte store per thread
    for (int c = tid; c < words; c += kThreads) {
      const uint32_t wd = __ldg(reinterpret_cast<const uint32_t *>(src) + c);
      // u8 -> fp32 exactly: 2^23 + b as a bit pattern, minus 2^23
      reinterpret_cast<float4 *>(s.in)[c] =
          make_float4(__uint_as_float(__byte_perm(wd, 0x4B000000u, 0x7650)) - 8388608.0f,
                      __uint_as_float(__byte_perm(wd, 0x4B000000u, 0x7651)) - 8388608.0f,
                      __uint_as_float(__byte_perm(wd, 0x4B000000u, 0x7652)) - 8388608.0f,
                      __uint_as_float(__byte_perm(wd, 0x4B000000u, 0x7653)) - 8388608.0f);
    }
  } else {
    const float inv_w = 1.0f / (float)p.w;
    for (int i = tid; i < n_in * p.w; i += kThreads) {
      const int r = (int)(((float)i + 0.5f) * inv_w), x = i - r * p.w;
      s.in[r * p.pitch + x] = (float)src[i];
    }
  }
  __syncthreads();
  // ---- vertical pass: 4 columns per thread, one 16-byte shared load per tap
  const int quads = p.pitch >> 2;
  const float inv_quads = 1.0f / (float)quads;
  for (int idx = tid; idx < rows * quads; idx += kThreads) {
    const int r = (int)(((float)idx + 0.5f) * inv_quads), q = idx - r * quads;
    const int first = s.vstart[r];
    const float *wk = s.vw + r * p.kv;
    float4 a = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    if (first >= 0 && first + p.kv <= p.h) { // interior: walk the rows
      const float4 *px = reinterpret_cast<const float4 *>(s.in + (size_t)(first - lo) * p.pitch) + q;
      for (int k = 0; k < p.kv; ++k, px += quads) {
        const float4 v = *px;
        const float wv = wk[k];
        a.x = fmaf(wv, v.x, a.x), a.y = fmaf(wv, v.y, a.y);
        a.z = fmaf(wv, v.z, a.z), a.w = fmaf(wv, v.w, a.w);
      }
    } else {
      for (int k = 0; k < p.kv; ++k) {
        const int row = min(max(first + k, 0), p.h - 1) - lo; // clamp edge == fold onto the edge pixel
        const float4 v = reinterpret_cast<const float4 *>(s.in + (size_t)row * p.pitch)[q];
        const float wv = wk[k];
        a.x = fmaf(wv, v.x, a.x), a.y = fmaf(wv, v.y, a.y);
        a.z = fmaf(wv, v.z, a.z), a.w = fmaf(wv, v.w, a.w);
      }
    }
    reinterpret_cast<float4 *>(s.inter + (size_t)r * p.pitch)[q] = a;
  }
  __syncthreads();
  // ---- horizontal pass + encode
  const float inv_nw = 1.0f / (float)p.nw;
  for (int idx = tid; idx < rows * p.nw; idx += kThreads) {
    const int r = (int)(((float)idx + 0.5f) * inv_nw), ox = idx - r * p.nw;
    const int first = s.hstart[ox];
    const float *line = s.inter + (size_t)r * p.pitch;
    const float *wk = s.hw + ox;
    float acc = 0.0f;
    if (first >= 0 && first + p.kh <= p.w) {
      for (int k = 0; k < p.kh; ++k)
        acc = fmaf(wk[k * p.nw], line[first + k], acc);
    } else {
      for (int k = 0; k < p.kh; ++k)
        acc = fmaf(wk[k * p.nw], line[min(max(first + k, 0), p.w - 1)], acc);
    }
    if (f32_plane)
      f32_plane[(size_t)(r0 + r) * p.nw + ox] = acc;
    const float qv = fminf(fmaxf(acc + 0.5f, 0.0f), 255.0f);
    tile[idx] = (uint8_t)(int)qv;
  }
  __syncthreads();
}

// ---- stand-alone resize: images [n,h,w] u8 -> out [n,nh,nw] u8 ---------------------------------
__global__ void __launch_bounds__(kThreads)
    resize_gray_kernel(DevPlan p, const uint8_t *__restrict__ images, uint8_t *__restrict__ out,
                       float *__restrict__ out_f32) {
  extern __shared__ __align__(16) uint8_t resize_raw[];
  const BandSmem s = carve(resize_raw, p);
  const int r0 = blockIdx.x * kBandRows, rows = min(kBandRows, p.nh - r0);
  const size_t n = blockIdx.y;
  load_tables(p, s, r0, rows);
  __syncthreads();
  resize_band(p, s, images + n * (size_t)p.w * p.h, r0, rows, s.tile[0],
              out_f32 ? out_f32 + n * (size_t)p.nw * p.nh : nullptr);
  uint8_t *dst = out + n * (size_t)p.nw * p.nh + (size_t)r0 * p.nw;
  for (int i = threadIdx.x; i < rows * p.nw; i += kThreads)
    dst[i] = s.tile[0][i];
}

// ---- fused live step: resize(a), resize(b) -> max -> frame stack (210x160 -> 84x84 only) ----------
constexpr int kPlane = 84 * 84, kTile = kBandRows * 84; // 1 176 bytes = 147 x 8
__global__ void __launch_bounds__(kThreads)
    preprocess_gray_step_kernel(DevPlan p, const uint8_t *__restrict__ frames_a,
                                const uint8_t *__restrict__ frames_b, int64_t frame_stride,
                                const uint8_t *__restrict__ starts, int64_t starts_stride,
                                const uint8_t *__restrict__ src, int64_t src_stride,
                                uint8_t *__restrict__ dst, int64_t dst_stride) {
  extern __shared__ __align__(16) uint8_t resize_raw[];
  const BandSmem s = carve(resize_raw, p);
  const int g = blockIdx.x, r0 = g * kBandRows, tid = threadIdx.x;
  const int64_t e = blockIdx.y;
  load_tables(p, s, r0, kBandRows);
  __syncthreads();
  resize_band(p, s, frames_a + e * frame_stride, r0, kBandRows, s.tile[0], nullptr);
  resize_band(p, s, frames_b + e * frame_stride, r0, kBandRows, s.tile[1], nullptr);
  for (int i = tid; i < kTile / 4; i += kThreads) { // max_and_skip.cc:36-38 on the two resized frames
    uint32_t *a = reinterpret_cast<uint32_t *>(s.tile[0]);
    a[i] = __vmaxu4(a[i], reinterpret_cast<const uint32_t *>(s.tile[1])[i]);
  }
  __syncthreads();
  uint8_t *d = dst + e * dst_stride + kTile * g;
  const uint2 *t8 = reinterpret_cast<const uint2 *>(s.tile[0]);
  if (starts[e * starts_stride]) { // rollout.cc:189-193
    for (int i = tid; i < 4 * 147; i += kThreads)
      st_stream_u2(d + (i / 147) * kPlane + 8 * (i % 147), t8[i % 147]);
  } else {
    if (tid < 147)
      st_stream_u2(d + 8 * tid, t8[tid]); // rollout.cc:194-195
    const uint8_t *sp = src + e * src_stride + kTile * g;
    for (int i = tid; i < 3 * 147; i += kThreads) { // rollout.cc:185-188
      const int pl = i / 147, l = i % 147;
      st_stream_u2(d + (pl + 1) * kPlane + 8 * l, ld_stream_u2(sp + pl * kPlane + 8 * l));
    }
  }
}

} // namespace
} // namespace alegpu

using namespace alegpu;

extern "C" int alegpu_resize_grayscale(const uint8_t *images, int64_t n, int width, int height,
                                       int new_width, int new_height, uint8_t *out,
                                       float *out_f32, void *stream) {
  ALEGPU_REQUIRE(width > 0 && height > 0, ALEGPU_INVALID_ARGUMENT,
                 "resize_grayscale: width and height must be > 0");
  ALEGPU_REQUIRE(new_width > 0 && new_height > 0, ALEGPU_INVALID_ARGUMENT,
                 "new_width and new_height must be > 0"); // resize.cc:15-16
  ALEGPU_REQUIRE(n >= 0 && n <= 65535, ALEGPU_INVALID_ARGUMENT, "resize_grayscale: 0 <= n <= 65535");
  if (n == 0)
    return ALEGPU_OK;
  ALEGPU_REQUIRE(images && out, ALEGPU_INVALID_ARGUMENT, "resize_grayscale: null pointer");
  const Plan *plan = nullptr;
  if (int rc = get_plan(width, height, new_width, new_height, &plan))
    return rc;
  if (plan->smem > 48 * 1024)
    ALEGPU_CUDA_CHECK(cudaFuncSetAttribute(resize_gray_kernel,
                                           cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)plan->smem));
  const dim3 grid((unsigned)ceil_div(new_height, kBandRows), (unsigned)n);
  resize_gray_kernel<<<grid, kThreads, plan->smem, as_stream(stream)>>>(dev_plan(*plan), images, out,
                                                                        out_f32);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}

extern "C" int alegpu_preprocess_gray_step(const uint8_t *frames_a, const uint8_t *frames_b,
                                           int64_t frame_env_stride, const uint8_t *starts,
                                           int64_t starts_stride, const uint8_t *src_stack,
                                           int64_t src_env_stride, uint8_t *dst_stack,
                                           int64_t dst_env_stride, int64_t E, void *stream) {
  ALEGPU_REQUIRE(E >= 0 && E <= 65535, ALEGPU_INVALID_ARGUMENT, "preprocess_gray_step: 0 <= E <= 65535");
  if (E == 0)
    return ALEGPU_OK;
  ALEGPU_REQUIRE(frames_a && frames_b && starts && src_stack && dst_stack, ALEGPU_INVALID_ARGUMENT,
                 "preprocess_gray_step: null pointer");
  ALEGPU_REQUIRE(src_stack != dst_stack, ALEGPU_INVALID_ARGUMENT,
                 "preprocess_gray_step: src and dst stacks must not alias");
  ALEGPU_REQUIRE((src_env_stride % 8) == 0 && (dst_env_stride % 8) == 0 &&
                     ((uintptr_t)src_stack % 8) == 0 && ((uintptr_t)dst_stack % 8) == 0,
                 ALEGPU_INVALID_ARGUMENT, "preprocess_gray_step: stacks must be 8-byte aligned");
  const Plan *plan = nullptr;
  if (int rc = get_plan(160, 210, 84, 84, &plan))
    return rc;
  if (plan->smem > 48 * 1024)
    ALEGPU_CUDA_CHECK(cudaFuncSetAttribute(preprocess_gray_step_kernel,
                                           cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)plan->smem));
  const dim3 grid(84 / kBandRows, (unsigned)E);
  preprocess_gray_step_kernel<<<grid, kThreads, plan->smem, as_stream(stream)>>>(
      dev_plan(*plan), frames_a, frames_b, frame_env_stride, starts, starts_stride, src_stack,
      src_env_stride, dst_stack, dst_env_stride);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}
