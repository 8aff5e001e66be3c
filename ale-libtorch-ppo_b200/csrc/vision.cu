// vision.cu — the stand-alone libtorch vision operators of the reference behind the C ABI:
//   ai::vision::resize_grayscale_image(Tensor)               src/ai/vision.cc:12-20
//   ai::vision::resize_frame_stacked_grayscale_images        src/ai/vision.cc:22-32
//   ai::vision::resize_frame_stacked_rgb_images              src/ai/vision.cc:38-48
//   ai::vision::rgb_to_grayscale_frame_stacked_images        src/ai/vision.cc:71-84
// All four are `interpolate(size = 84x84, mode = area)` (= adaptive average pooling: output (oy, ox)
// is the mean of input rows [floor(oy*210/84), ceil((oy+1)*210/84)) x columns [floor(ox*160/84),
// ceil((ox+1)*160/84))) over float planes, and a matmul with (0.2125, 0.7154, 0.0721) over the channel
// axis.  The learner hot path does not call them (alegpu_preprocess_step fuses max + resize + gray +
// quantise + stack in one kernel, preprocess.cu); they exist so the reference's own vision interface and
// its unit tests (test/ai/vision-test.cc) have a drop-in, and they share the bin arithmetic.
#include "common.cuh"

namespace alegpu {
namespace {

constexpr int kInH = 210, kInW = 160, kOut = 84;

// one thread per output pixel of one plane; rows then columns, fp32 sum / count (ATen's order)
__global__ void __launch_bounds__(256)
    resize_area_kernel(const float *__restrict__ in, float *__restrict__ out, int64_t planes) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= planes * kOut * kOut)
    return;
  const int64_t p = i / (kOut * kOut);
  const int o = (int)(i - p * kOut * kOut), oy = o / kOut, ox = o - oy * kOut;
  const int y0 = (oy * kInH) / kOut, y1 = ((oy + 1) * kInH + kOut - 1) / kOut;
  const int x0 = (ox * kInW) / kOut, x1 = ((ox + 1) * kInW + kOut - 1) / kOut;
  const float *src = in + p * (kInH * kInW);
  float s = 0.0f;
  for (int y = y0; y < y1; ++y)
    for (int x = x0; x < x1; ++x)
      s = __fadd_rn(s, __ldg(src + y * kInW + x));
  out[i] = s / (float)((y1 - y0) * (x1 - x0));
}

__global__ void __launch_bounds__(256)
    rgb_to_gray_kernel(const float *__restrict__ in, float *__restrict__ out, int64_t images) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= images * kOut * kOut)
    return;
  const int64_t n = i / (kOut * kOut);
  const int o = (int)(i - n * kOut * kOut);
  const float *src = in + n * 3 * kOut * kOut + o;
  // matmul over the channel axis, weights of vision.cc:51
  // (no FMA contraction: the restatement's fp32 operation order, bit for bit)
  out[i] = __fadd_rn(__fadd_rn(__fmul_rn(src[0], 0.2125f), __fmul_rn(src[kOut * kOut], 0.7154f)),
                     __fmul_rn(src[2 * kOut * kOut], 0.0721f));
}

} // namespace
} // namespace alegpu

using namespace alegpu;

extern "C" int alegpu_vision_resize_area(const float *images, int64_t planes, float *out, void *stream) {
  ALEGPU_REQUIRE(planes >= 0, ALEGPU_INVALID_ARGUMENT, "vision_resize_area: negative plane count");
  if (planes == 0)
    return ALEGPU_OK;
  ALEGPU_REQUIRE(images && out, ALEGPU_INVALID_ARGUMENT, "vision_resize_area: null pointer");
  const int64_t total = planes * kOut * kOut;
  resize_area_kernel<<<(unsigned)ceil_div(total, 256), 256, 0, as_stream(stream)>>>(images, out, planes);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}

extern "C" int alegpu_vision_rgb_to_grayscale(const float *images, int64_t n, float *out, void *stream) {
  ALEGPU_REQUIRE(n >= 0, ALEGPU_INVALID_ARGUMENT, "vision_rgb_to_grayscale: negative image count");
  if (n == 0)
    return ALEGPU_OK;
  ALEGPU_REQUIRE(images && out, ALEGPU_INVALID_ARGUMENT, "vision_rgb_to_grayscale: null pointer");
  const int64_t total = n * kOut * kOut;
  rgb_to_gray_kernel<<<(unsigned)ceil_div(total, 256), 256, 0, as_stream(stream)>>>(images, out, n);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}
