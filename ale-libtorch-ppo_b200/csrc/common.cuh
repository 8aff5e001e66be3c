// common.cuh — shared helpers for the sm_100a learner kernels.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <string>

#include "alegpu.h"

namespace alegpu {

// thread-local last-error string + status plumbing ----------------------------------------
void set_error(const std::string &msg);
int fail(int status, const std::string &msg);

#define ALEGPU_CUDA_CHECK(expr)                                                              \
  do {                                                                                       \
    cudaError_t _e = (expr);                                                                 \
    if (_e != cudaSuccess)                                                                   \
      return ::alegpu::fail(ALEGPU_CUDA_ERROR, std::string(#expr) + ": " +                   \
                                                   cudaGetErrorString(_e));                  \
  } while (0)

#define ALEGPU_LAUNCH_CHECK()                                                                \
  do {                                                                                       \
    cudaError_t _e = cudaGetLastError();                                                     \
    if (_e != cudaSuccess)                                                                   \
      return ::alegpu::fail(ALEGPU_CUDA_ERROR, std::string("kernel launch: ") +              \
                                                   cudaGetErrorString(_e));                  \
  } while (0)

#define ALEGPU_REQUIRE(cond, status, msg)                                                    \
  do {                                                                                       \
    if (!(cond))                                                                             \
      return ::alegpu::fail((status), (msg));                                                \
  } while (0)

constexpr int kNumSMs = 148; // B200: 2 dies x 74

inline cudaStream_t as_stream(void *s) { return static_cast<cudaStream_t>(s); }
inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

// launch counter (bench.py's gpu_launches claim): every kernel launch goes through COUNT()
extern thread_local int64_t g_launches;
// optional launch tracer (alegpu_trace_begin/_end): a CUDA event after every launch site
void trace_launch(const char *file, int line);
extern thread_local bool g_trace_on;
#define ALEGPU_COUNT_LAUNCH_NAMED(name)                                                      \
  do {                                                                                       \
    ++::alegpu::g_launches;                                                                  \
    if (::alegpu::g_trace_on)                                                                \
      ::alegpu::trace_launch(name, 0);                                                       \
  } while (0)
#define ALEGPU_COUNT_LAUNCH()                                                                \
  do {                                                                                       \
    ++::alegpu::g_launches;                                                                  \
    if (::alegpu::g_trace_on)                                                                \
      ::alegpu::trace_launch(__FILE__, __LINE__);                                            \
  } while (0)

#ifdef __CUDACC__
// 16-byte streaming load/store: data touched once goes around L1 -----------------------------
__device__ __forceinline__ uint4 ld_stream_u4(const void *p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ uint2 ld_stream_u2(const void *p) {
  uint2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];"
               : "=r"(r.x), "=r"(r.y)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream_u4(void *p, const uint4 &v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x),
               "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}
__device__ __forceinline__ void st_stream_u2(void *p, const uint2 &v) {
  asm volatile("st.global.L1::no_allocate.v2.u32 [%0], {%1,%2};" ::"l"(p), "r"(v.x), "r"(v.y)
               : "memory");
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1)
    v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1)
    v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Deterministic block reduction (fixed tree): every thread gets the total.
template <typename T, int kThreads> __device__ __forceinline__ T block_sum(T v, T *smem) {
  constexpr int kWarps = kThreads / 32;
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0)
    smem[warp] = v;
  __syncthreads();
  T t = smem[0];
#pragma unroll
  for (int i = 1; i < kWarps; ++i)
    t += smem[i];
  return t;
}
#endif

} // namespace alegpu
