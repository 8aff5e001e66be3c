// net_tc.cu — tensor-core (tcgen05/TMEM) Nature-CNN forward/backward for sm_100a.
//
// Replaces NetworkImpl::forward (src/bin/train.cc:255-265) and its autograd backward
// (src/ai/ppo/train.h:126) — cuDNN/cuBLAS in the reference — with the implicit-GEMM engine of
// tc_gemm.cuh.  Every layer, its data gradient and its weight gradient is ONE launch of that
// engine with a different pair of operand views (16-byte chunk loaders) and a fused epilogue:
//
//   layer        M        N     R (reduction)   A view (major)            B view (major)
//   conv1 fwd    n*400    32    256             im2col(u8 obs)      (K)   c1w [32][256]       (K)
//   conv2 fwd    n*81     64    512             im2col(act1)        (K)   c2w [64][512]       (K)
//   conv3 fwd    n*49     64    576             im2col(act2)        (K)   c3w [64][576]       (K)
//   fc    fwd    n        H     3136            act3                (K)   fcw [H][3136]       (K)
//   fc    dgrad  n        3136  H               dfc (fp32->bf16)    (K)   fcw^T               (MN)
//   fc    wgrad  H        3137  n               dfc^T               (MN)  act3^T | ones       (MN)
//   conv3 dgrad  n*81     64    576             col2im(dact3)       (K)   c3w^T               (MN)
//   conv3 wgrad  577      64    n*49            im2col(act2)^T|ones (MN)  dact3^T             (MN)
//   conv2 dgrad  4 x n*100 32   256             col2im(dact2),parity(K)   c2w^T (parity taps) (MN)
//   conv2 wgrad  513      64    n*81            im2col(act1)^T|ones (MN)  dact2^T             (MN)
//   conv1 wgrad  257      32    n*400           im2col(obs)^T|ones  (MN)  dact1^T             (MN)
//
// Activations and activation gradients are bf16 NHWC; weights are read from the bf16 shadow of
// the flat parameter buffer that the fused Adam kernel maintains (optim.cu).  u8 pixels are exact
// in bf16, so the 1/255 of train.cc:258-259 is applied to the fp32 accumulator in the epilogue.
// Weight gradients are split over the reduction and summed in fixed order (deterministic).
#include <algorithm>
#include <cstdlib>
#include <vector>

#include <cuda.h>

#include "net_tc.h"
#include "optim.h"
#include "tc_gemm.cuh"
#include "tc_shift.cuh"

namespace alegpu {
namespace {

using bf16 = __nv_bfloat16;
using tc::pack_bf16;

__device__ __forceinline__ uint4 ldg16(const void *p) {
  return __ldg(reinterpret_cast<const uint4 *>(p));
}
__device__ __forceinline__ uint4 zero16() { return make_uint4(0u, 0u, 0u, 0u); }
// the chunk (1, 0, 0, 0, 0, 0, 0, 0) in bf16: source of the "ones" row/column that turns a weight
// gradient GEMM into weight + bias gradient, for the cp.async producers
__device__ const uint4 g_ones_chunk = {0x00003F80u, 0u, 0u, 0u};
__device__ const uint4 g_ones_chunk_f32 = {0x3F800000u, 0u, 0u, 0u}; // (1, 0, 0, 0) in fp32 / tf32
template <int kElemBytes> __device__ __forceinline__ const void *ones_chunk_ptr() {
  if constexpr (kElemBytes == 4)
    return &g_ones_chunk_f32;
  else
    return &g_ones_chunk;
}
__device__ __forceinline__ uint4 ones_first16() { return make_uint4(0x00003F80u, 0u, 0u, 0u); }
// four u8 -> four bf16 (exact): bf16(v) = high half of float(v)
// (byte k dropped into the mantissa of 2^23, minus 2^23: PRMT + FADD instead of the slow I2F pipe)
__device__ __forceinline__ void u8x4_to_bf16(uint32_t w, uint32_t &lo, uint32_t &hi) {
  constexpr uint32_t kMagic = 0x4B000000u; // 8388608.0f
  const uint32_t f0 = __float_as_uint(__uint_as_float(__byte_perm(w, kMagic, 0x7540)) - 8388608.0f);
  const uint32_t f1 = __float_as_uint(__uint_as_float(__byte_perm(w, kMagic, 0x7541)) - 8388608.0f);
  const uint32_t f2 = __float_as_uint(__uint_as_float(__byte_perm(w, kMagic, 0x7542)) - 8388608.0f);
  const uint32_t f3 = __float_as_uint(__uint_as_float(__byte_perm(w, kMagic, 0x7543)) - 8388608.0f);
  lo = __byte_perm(f0, f1, 0x7632);
  hi = __byte_perm(f2, f3, 0x7632);
}
struct F32x8 {
  float4 a, b;
};
__device__ __forceinline__ F32x8 ldg_f32x8(const float *p) {
  F32x8 r;
  r.a = __ldg(reinterpret_cast<const float4 *>(p));
  r.b = __ldg(reinterpret_cast<const float4 *>(p) + 1);
  return r;
}
__device__ __forceinline__ uint4 f32x8_to_bf16(const F32x8 &v) {
  return make_uint4(pack_bf16(v.a.x, v.a.y), pack_bf16(v.a.z, v.a.w), pack_bf16(v.b.x, v.b.y),
                    pack_bf16(v.b.z, v.b.w));
}
// loaders whose raw form already is 8 bf16 / is 8 fp32
#define ALEGPU_RAW_BF16                                                                          \
  using Raw = uint4;                                                                             \
  static __device__ __forceinline__ Raw zero() { return make_uint4(0u, 0u, 0u, 0u); }            \
  static __device__ __forceinline__ uint4 cvt(const Raw &v) { return v; }
#define ALEGPU_RAW_F32                                                                           \
  using Raw = F32x8;                                                                             \
  static __device__ __forceinline__ Raw zero() {                                                 \
    F32x8 r;                                                                                     \
    r.a = r.b = make_float4(0.f, 0.f, 0.f, 0.f);                                                 \
    return r;                                                                                    \
  }                                                                                              \
  static __device__ __forceinline__ uint4 cvt(const Raw &v) { return f32x8_to_bf16(v); }

// ---- im2col / col2im chunk views -----------------------------------------------------------------
// Each view gives the RAW form of one 16-byte operand chunk (8 reduction- or row-contiguous
// elements); conversion to bf16 happens when the chunk is stored to shared memory.
// conv1: row m = (n,oy,ox) of 20x20, k = c*64 + ky*8 + kx ; chunk = the 8 kx of one (c,ky), raw u8
struct Col1 {
  static constexpr bool kAsyncSrc = false; // u8 pixels need the register path (convert at store)
  static constexpr int kPrefetch = 4;      // 8-byte raw chunks: loads run 3 k-blocks ahead
  using Raw = uint2;
  static __device__ __forceinline__ Raw zero() { return make_uint2(0u, 0u); }
  static __device__ __forceinline__ uint4 cvt(const Raw &v) {
    uint4 r;
    u8x4_to_bf16(v.x, r.x, r.y);
    u8x4_to_bf16(v.y, r.z, r.w);
    return r;
  }
  const uint8_t *obs; // [n,4,84,84] u8
  using PosPtr = const uint8_t *;
  __device__ __forceinline__ const void *anchor() const { return obs; }
  __device__ __forceinline__ PosPtr pos_ptr(int) const { return nullptr; } // never async
  __device__ __forceinline__ int k_off(int) const { return 0; }
  __device__ __forceinline__ Raw at(int m, int k) const {
    const int n = m / 400, pix = m - n * 400, oy = pix / 20, ox = pix - oy * 20;
    const int c = k >> 6, ky = (k >> 3) & 7;
    const uint32_t *p = reinterpret_cast<const uint32_t *>(
        obs + (size_t)n * 28224 + c * 7056 + (oy * 4 + ky) * 84 + ox * 4);
    return make_uint2(__ldg(p), __ldg(p + 1));
  }
};
// conv1 over the bf16 NHWC copy of the observations ([n,84,84,4], made once per train() by
// obs_to_s2d_bf16_kernel): row m = (n,oy,ox) of 20x20; on the space-to-depth layout the 8x8 stride-4
// window is 2x2 taps of 64 contiguous channels, so every 16-byte chunk can be copied with cp.async.  The conv1 weights are read from a shadow permuted to the same k' order.
template <class T> struct Col1N {
  static constexpr int kElemBytes = (int)sizeof(T);
  ALEGPU_RAW_BF16
  static constexpr bool kAsyncSrc = true;
  static constexpr int kPrefetch = 2;
  const T *x; // obs16, space-to-depth [n,21,21,64]: channel = (y&3)*16 + (x&3)*4 + c
  using PosPtr = const T *;
  __device__ __forceinline__ const void *anchor() const { return x; }
  __device__ __forceinline__ const T *pos_ptr(int m) const {
    const int n = m / 400, pix = m - n * 400, oy = pix / 20, ox = pix - oy * 20;
    return x + ((size_t)(n * 21 + oy) * 21 + ox) * 64;
  }
  // k' = tap*64 + ch, tap = a*2 + b: the 2x2 stride-1 convolution over 64 channels that the 8x8
  // stride-4 convolution becomes on this layout; one tap = 128 contiguous bytes
  __device__ __forceinline__ int k_off(int k) const {
    return ((k >> 7) * 21 + ((k >> 6) & 1)) * 64 + (k & 63);
  }
  __device__ __forceinline__ uint4 at(int m, int k) const { return ldg16(pos_ptr(m) + k_off(k)); }
};
// k' = (a*2+b)*64 + r*16 + s*4 + c  ->  the reference's OIHW reduction index k = c*64 + ky*8 + kx
// with ky = 4a + r, kx = 4b + s
__host__ __device__ __forceinline__ int conv1_k_of_kprime(int kp) {
  const int c = kp & 3, sx = (kp >> 2) & 3, ry = (kp >> 4) & 3, b = (kp >> 6) & 1, a = kp >> 7;
  return c * 64 + (4 * a + ry) * 8 + 4 * b + sx;
}
// conv2: row m = (n,oy,ox) of 9x9, k = (ky*4+kx)*32 + ci
template <class T> struct Col2 {
  static constexpr int kElemBytes = (int)sizeof(T);
  ALEGPU_RAW_BF16
  static constexpr bool kAsyncSrc = true;
  static constexpr int kPrefetch = 2;
  const T *x; // act1 [n,20,20,32]
  // address of element (m, k) = pos_ptr(m) + k_off(k): the position part and the tap part
  using PosPtr = const T *;
  __device__ __forceinline__ const void *anchor() const { return x; }
  __device__ __forceinline__ const T *pos_ptr(int m) const {
    const int n = m / 81, pix = m - n * 81, oy = pix / 9, ox = pix - oy * 9;
    return x + ((size_t)(n * 20 + 2 * oy) * 20 + 2 * ox) * 32;
  }
  __device__ __forceinline__ int k_off(int k) const {
    const int ky = k >> 7, kx = (k >> 5) & 3, ci = k & 31;
    return (ky * 20 + kx) * 32 + ci;
  }
  __device__ __forceinline__ uint4 at(int m, int k) const { return ldg16(pos_ptr(m) + k_off(k)); }
};
// conv3: row m = (n,oy,ox) of 7x7, k = (ky*3+kx)*64 + ci
template <class T> struct Col3 {
  static constexpr int kElemBytes = (int)sizeof(T);
  ALEGPU_RAW_BF16
  static constexpr bool kAsyncSrc = true;
  static constexpr int kPrefetch = 2;
  const T *x; // act2 [n,9,9,64]
  using PosPtr = const T *;
  __device__ __forceinline__ const void *anchor() const { return x; }
  __device__ __forceinline__ const T *pos_ptr(int m) const {
    const int n = m / 49, pix = m - n * 49, oy = pix / 7, ox = pix - oy * 7;
    return x + ((size_t)(n * 9 + oy) * 9 + ox) * 64;
  }
  __device__ __forceinline__ int k_off(int k) const {
    const int t = k >> 6, ky = t / 3, kx = t - ky * 3, ci = k & 63;
    return (ky * 9 + kx) * 64 + ci;
  }
  __device__ __forceinline__ uint4 at(int m, int k) const { return ldg16(pos_ptr(m) + k_off(k)); }
};

// K-major A operand fetched by TMA im2col loads: NHWC bf16 tensor with 64 channels (one tap = one
// 128-byte swizzle row), rows = output pixels (n, oy, ox) of OH x OW, reduction k = tap*64 + c with
// tap = ky*KW + kx.  LOW = lower corner of the pixel box (= -padding); SH = stride in h; kFlip = the transposed
// (data-gradient) walk, where tap (ky,kx) reads pixel (oy-ky, ox-kx).  One instruction per
// (128-row tile, k-block) replaces 1024 16-byte cp.async copies and their address arithmetic;
// pixels outside the tensor (padding, rows past the last image) arrive as zeros.
template <class T, int OH, int OW, int KW, int LOW, bool kFlip, int SH = 1> struct ColTma {
  ALEGPU_RAW_BF16
  static constexpr int kElemBytes = (int)sizeof(T);
  static constexpr int kBK = 128 / kElemBytes; // channels per k-block (64 bf16 / 32 fp32 of the 64 per pixel)
  static constexpr bool kMN = false;
  static constexpr bool kAsync = true;
  static constexpr bool kTma = true;
  const void *map; // CUtensorMap in global memory
  int n_base;      // image index of row 0 inside the mapped buffer
  struct TileCtx {
    int w, h, n;
  };
  __device__ __forceinline__ TileCtx tile(int m0) const {
    const int n = m0 / (OH * OW), pix = m0 - n * (OH * OW), oy = pix / OW, ox = pix - oy * OW;
    return TileCtx{ox + LOW, oy * SH + LOW, n + n_base}; // SH = traversal stride in h
  }
  __device__ __forceinline__ void tma(const TileCtx &t, uint32_t dst, int r0, int, uint64_t *bar) const {
    const int tap = r0 >> 6, ky = tap / KW, kx = tap - ky * KW;
    const int ow = kFlip ? -LOW - kx : kx, oh = kFlip ? -LOW - ky : ky;
    tc::tma_im2col_4d(dst, map, r0 & 63 & ~(kBK - 1), t.w, t.h, t.n, (uint16_t)ow, (uint16_t)oh, bar);
  }
};
template <class T> using Col3Tma = ColTma<T, 7, 7, 3, 0, false>;   // conv3 forward over act2 [n,9,9,64]
template <class T> using DCol3Tma = ColTma<T, 9, 9, 3, -2, true>;  // conv3 data gradient over dact3 [n,7,7,64]
// conv2 forward over act1 [n,20,20,32] viewed as pixel PAIRS [n,20,10,64]: with stride 2 the taps
// kx = (0,1) and (2,3) are the pairs ox and ox+1, so it is a 4 (ky, stride 2) x 2 (pairs, stride 1)
// convolution over 64 channels and k = (ky*2 + pair)*64 + (kx&1)*32 + ci is the weights' own order
template <class T> using Col2Tma = ColTma<T, 9, 9, 2, 0, false, 2>;
template <class T> using Col1Tma = ColTma<T, 20, 20, 2, 0, false>; // conv1 forward over the space-to-depth obs16 [n,21,21,64]
// conv2 data gradient: every parity class of the stride-2 transposed convolution reads the same
// 2x2-tap, padding-1 walk over dact2 [n,9,9,64] (rows = (n,qy,qx) of 10x10); only the weights differ
template <class T> using DCol2Tma = ColTma<T, 10, 10, 2, -1, true>;

// K-major A operand of a forward conv: (row m, reduction k..k+7)
template <class Col, class = void> struct ColElemBytes {
  static constexpr int value = 2;
};
template <class Col> struct ColElemBytes<Col, decltype((void)Col::kElemBytes)> {
  static constexpr int value = Col::kElemBytes;
};
template <class Col> struct ColK {
  static constexpr bool kMN = false;
  static constexpr int kElemBytes = ColElemBytes<Col>::value;
  using Raw = typename Col::Raw;
  static __device__ __forceinline__ Raw zero() { return Col::zero(); }
  static __device__ __forceinline__ uint4 cvt(const Raw &v) { return Col::cvt(v); }
  static constexpr bool kAsync = Col::kAsyncSrc;
  static constexpr int kPrefetch = Col::kPrefetch;
  Col src;
  __device__ __forceinline__ Raw load(int m, int k, int) const { return src.at(m, k); }
  using RowCtx = typename Col::PosPtr; // position part of the address
  using ColCtx = int;                                           // tap part (element offset)
  __device__ __forceinline__ RowCtx row(int m) const { return src.pos_ptr(m); }
  __device__ __forceinline__ ColCtx col(int k, int) const { return src.k_off(k); }
  __device__ __forceinline__ const void *anchor() const { return src.anchor(); }
  __device__ __forceinline__ const void *at(RowCtx rc, ColCtx cc) const { return rc + cc; }
};
// MN-major A operand of a conv wgrad: rows = k (plus a ones row at k == K for the bias), r = m.
// The ones row is encoded in the raw form as an all-ones first word (never a valid chunk of
// this operand: raw u8 0xFFFFFFFF|.. pixels or a bf16 NaN pair) and decoded in cvt().
template <class Col, int K> struct ColMN {
  static constexpr bool kMN = true;
  static constexpr int kElemBytes = ColElemBytes<Col>::value;
  struct Raw {
    typename Col::Raw v;
    bool ones;
  };
  static __device__ __forceinline__ Raw zero() { return Raw{Col::zero(), false}; }
  static __device__ __forceinline__ uint4 cvt(const Raw &r) {
    return r.ones ? ones_first16() : Col::cvt(r.v);
  }
  static constexpr bool kAsync = Col::kAsyncSrc;
  static constexpr int kPrefetch = Col::kPrefetch;
  static constexpr bool kPosFastest = !Col::kAsyncSrc; // register path (conv1): see tc_gemm.cuh
  Col src;
  __device__ __forceinline__ Raw load(int k0, int m, int) const {
    if (k0 < K)
      return Raw{src.at(m, k0), false};
    return Raw{Col::zero(), true};
  }
  using RowCtx = int; // tap offset of k0, or -1 for the ones row
  using ColCtx = typename Col::PosPtr;
  __device__ __forceinline__ RowCtx row(int k0) const { return k0 < K ? src.k_off(k0) : -1; }
  __device__ __forceinline__ ColCtx col(int m, int) const { return src.pos_ptr(m); }
  __device__ __forceinline__ const void *anchor() const { return src.anchor(); }
  __device__ __forceinline__ const void *at(RowCtx rc, ColCtx cc) const {
    return rc >= 0 ? (const void *)(cc + rc) : ones_chunk_ptr<kElemBytes>();
  }
};
// plain bf16 matrix [rows][ld]: K-major view (row, r..r+7)
template <class T> struct RowsK {
  static constexpr int kElemBytes = (int)sizeof(T);
  ALEGPU_RAW_BF16
  static constexpr bool kMN = false;
  static constexpr bool kAsync = true;
  const T *p;
  int ld;
  __device__ __forceinline__ uint4 load(int i, int r, int) const {
    return ldg16(p + (size_t)i * ld + r);
  }
  using RowCtx = const T *;
  using ColCtx = int;
  __device__ __forceinline__ RowCtx row(int i) const { return p + (size_t)i * ld; }
  __device__ __forceinline__ ColCtx col(int r, int) const { return r; }
  __device__ __forceinline__ const void *anchor() const { return p; }
  __device__ __forceinline__ const void *at(RowCtx rc, ColCtx cc) const { return rc + cc; }
};
// fp32 matrix [rows][ld] converted on the fly: K-major view
struct RowsKF32 {
  ALEGPU_RAW_F32
  static constexpr bool kMN = false;
  const float *p;
  int ld;
  __device__ __forceinline__ F32x8 load(int i, int r, int) const {
    return ldg_f32x8(p + (size_t)i * ld + r);
  }
};
// transposed view of a bf16 matrix stored [r][ld]: element (i, r) = p[r*ld + i]; MN-major;
// rows i >= ld (only i == ld is ever requested) give the ones row/column used for bias grads
template <class T> struct ColsMN {
  static constexpr int kElemBytes = (int)sizeof(T);
  ALEGPU_RAW_BF16
  static constexpr bool kMN = true;
  const T *p;
  int ld;
  static constexpr bool kAsync = true;
  __device__ __forceinline__ uint4 load(int i0, int r, int) const {
    return i0 < ld ? ldg16(p + (size_t)r * ld + i0) : ones_first16();
  }
  using RowCtx = int; // i0, or -1 for the ones row
  using ColCtx = const T *;
  __device__ __forceinline__ RowCtx row(int i0) const { return i0 < ld ? i0 : -1; }
  __device__ __forceinline__ ColCtx col(int r, int) const { return p + (size_t)r * ld; }
  __device__ __forceinline__ const void *anchor() const { return p; }
  __device__ __forceinline__ const void *at(RowCtx rc, ColCtx cc) const {
    return rc >= 0 ? (const void *)(cc + rc) : ones_chunk_ptr<kElemBytes>();
  }
};
// The same transposed view of a plain [rows r][64 channels] matrix as ColsMN, fetched by TMA: an MN-major
// operand tile is, per 128-byte group of channels, just BK consecutive matrix rows of 128 bytes (bf16: one
// 64-channel x 64-row box, SWIZZLE_128B; tf32: two 32-channel x 32-row boxes 4096 B apart, and the tensor
// map's SWIZZLE_128B_ATOM_32B is the BASE32B pattern the tf32 MN-major descriptor expects).  One elected
// thread, one or two bulk loads per k-block instead of 512 cp.async copies.  Rows past the split's r_end
// arrive as data (or zeros past the tensor) — harmless: the OTHER operand zero-fills them.
template <class T, int ROWS = 64> struct ColsMNTma {
  ALEGPU_RAW_BF16
  static constexpr int kElemBytes = (int)sizeof(T);
  static constexpr bool kMN = true;
  static constexpr bool kAsync = true;
  static constexpr bool kTma = true;
  static constexpr int kAtomRows = 128 / kElemBytes;              // matrix columns per 128-byte group
  static constexpr uint32_t kLbo = kElemBytes == 2 ? 8192u : 4096u; // bytes between the groups of one tile
  const void *map; // CUtensorMap (tiled, 2-d) in global memory
  int row_base;    // matrix row of reduction index 0
  struct TileCtx {
    int c0;
  };
  __device__ __forceinline__ TileCtx tile(int n0) const { return TileCtx{n0}; }
  __device__ __forceinline__ void tma(const TileCtx &t, uint32_t dst, int r0, int, uint64_t *bar) const {
#pragma unroll
    for (int a = 0; a < ROWS / kAtomRows; ++a) // columns past the matrix arrive as zeros
      tc::tma_tile_2d(dst + (uint32_t)a * kLbo, map, t.c0 + a * kAtomRows, row_base + r0, bar);
  }
};
// K-major view of a plain [rows][ld] matrix (RowsK) fetched by TMA: ROWS/64 boxes of 64 rows x one
// 128-byte k-block, SWIZZLE_128B.  Rows past the matrix are zeros, rows past M are masked by the epilogue.
template <class T, int ROWS> struct RowsKTma {
  ALEGPU_RAW_BF16
  static constexpr int kElemBytes = (int)sizeof(T);
  static constexpr bool kMN = false;
  static constexpr bool kAsync = true;
  static constexpr bool kTma = true;
  const void *map;
  struct TileCtx {
    int row0;
  };
  __device__ __forceinline__ TileCtx tile(int i0) const { return TileCtx{i0}; }
  __device__ __forceinline__ void tma(const TileCtx &t, uint32_t dst, int r0, int, uint64_t *bar) const {
#pragma unroll
    for (int h = 0; h < ROWS / 64; ++h)
      tc::tma_tile_2d(dst + (uint32_t)h * 8192u, map, r0, t.row0 + h * 64, bar);
  }
};
// The transposed conv weights of the data gradients (W3T, W2TAll) through 3-d tiled TMA.  Both views read,
// for one filter tap and a run of BK output channels co (the k-block), 128-byte groups of consecutive
// input-channel elements — boxes {128 bytes, 1 tap, BK co} of the weight tensor seen as [co][tap'][128 B]:
//   conv3 (W3T):    c3w [64 co][9 taps][64 ci]       -> group g = ci / (128 B), tap' = tap
//   conv2 (W2TAll): c2w [64 co][16 taps][32 ci]; output row j = cls*32 + ci with cls = (py,px) and the
//                   tap of (cls, t = (ty,tx)) = (py+2ty)*4 + px+2tx: bf16 groups of 64 rows are the two
//                   px of one py = two NEIGHBOURING taps (tap' = tap pair), tf32 groups of 32 rows one tap
template <class T, int NGROUPS, bool kConv2> struct WTapTma {
  ALEGPU_RAW_BF16
  static constexpr int kElemBytes = (int)sizeof(T);
  static constexpr bool kMN = true;
  static constexpr bool kAsync = true;
  static constexpr bool kTma = true;
  static constexpr int kAtomRows = 128 / kElemBytes;
  static constexpr uint32_t kLbo = kElemBytes == 2 ? 8192u : 4096u;
  const void *map;
  struct TileCtx {
    int n0;
  };
  __device__ __forceinline__ TileCtx tile(int n0) const { return TileCtx{n0}; }
  __device__ __forceinline__ void tma(const TileCtx &, uint32_t dst, int r0, int, uint64_t *bar) const {
    const int t = r0 >> 6, co0 = r0 & 63;
#pragma unroll
    for (int g = 0; g < NGROUPS; ++g) {
      int c0 = 0, tap = t;
      if constexpr (!kConv2) {
        c0 = g * kAtomRows; // conv3: group g = input channels [g*atom, (g+1)*atom) of tap t
      } else {
        const int ty = t >> 1, tx = t & 1;
        if constexpr (sizeof(T) == 2)
          tap = (g + 2 * ty) * 2 + tx; // bf16: group g = py, tap pair index ((py+2ty)*4 + 2tx) / 2
        else
          tap = ((g >> 1) + 2 * ty) * 4 + (g & 1) + 2 * tx; // tf32: group g = cls = (py, px)
      }
      tc::tma_tile_3d(dst + (uint32_t)g * kLbo, map, c0, tap, co0, bar);
    }
  }
};
// conv1 wgrad on the space-to-depth layout, A operand (MN-major): row m = tap*32 + oc (tap = a*2+b),
// reduction index r = s2d block (n, Y, X) of 21x21.  Element = dact1[n, Y-a, X-b, oc] or 0 outside
// 20x20: the FOUR TAPS ARE FOUR SHIFTED VIEWS OF THE (small) OUTPUT GRADIENT, so the observation
// operand (ColsMN over obs16, 128 contiguous bytes per r) is read exactly once, un-expanded:
//   dW[(a,b,oc)][ch] = sum_{n,Y,X} dact1[n, Y-a, X-b, oc] * S[n, Y, X, ch]
template <class T> struct DYShift {
  static constexpr int kElemBytes = (int)sizeof(T);
  ALEGPU_RAW_BF16
  static constexpr bool kMN = true;
  static constexpr bool kAsync = true;
  const T *dy; // dact1 [n,20,20,32]
  struct RowCtx {
    int off, a, b; // off = oc0 - (a*20 + b)*32
  };
  struct ColCtx {
    const T *base; // dy element of (n, oy = Y, ox = X), channel 0 (may lie outside: only an offset)
    int Y, X;
  };
  __device__ __forceinline__ RowCtx row(int m0) const {
    const int tap = m0 >> 5, a = tap >> 1, b = tap & 1;
    return RowCtx{(m0 & 31) - (a * 20 + b) * 32, a, b};
  }
  __device__ __forceinline__ ColCtx col(int r, int) const {
    const int n = r / 441, pix = r - n * 441, Y = pix / 21, X = pix - Y * 21;
    return ColCtx{dy + ((ptrdiff_t)(n * 20 + Y) * 20 + X) * 32, Y, X};
  }
  __device__ __forceinline__ const void *anchor() const { return dy; }
  __device__ __forceinline__ const void *at(const RowCtx &rc, const ColCtx &cc) const {
    const int oy = cc.Y - rc.a, ox = cc.X - rc.b;
    return ((unsigned)oy < 20u && (unsigned)ox < 20u) ? (const void *)(cc.base + rc.off) : nullptr;
  }
  __device__ __forceinline__ uint4 load(int m0, int r, int z) const {
    const void *p = at(row(m0), col(r, z));
    return p ? ldg16(p) : zero16();
  }
};
// DYShift through TMA (tf32 only: a tap's 32 output channels are exactly one 128-byte MN group; in bf16 a
// group would span two taps).  Per k-block of 32 space-to-depth blocks r = (n, Y, X) and per tap (a, b) ONE
// im2col load of 32 pixels x 32 channels of dact1 [n,20,20,32]: base pixel (X-1, Y-1) on the map's 21-wide
// traversal grid (lower corner -1, upper corner 0), tap offsets (1-b, 1-a); pixels outside 20x20 are zeros.
struct DYShiftTma {
  ALEGPU_RAW_BF16
  static constexpr int kElemBytes = 4;
  static constexpr bool kMN = true;
  static constexpr bool kAsync = true;
  static constexpr bool kTma = true;
  const void *map;
  struct TileCtx {
    int m0;
  };
  __device__ __forceinline__ TileCtx tile(int m0) const { return TileCtx{m0}; }
  __device__ __forceinline__ void tma(const TileCtx &, uint32_t dst, int r0, int, uint64_t *bar) const {
    const int n = r0 / 441, pix = r0 - n * 441, Y = pix / 21, X = pix - Y * 21;
#pragma unroll
    for (int tap = 0; tap < 4; ++tap)
      tc::tma_im2col_4d(dst + (uint32_t)tap * 4096u, map, 0, X - 1, Y - 1, n, (uint16_t)(1 - (tap & 1)),
                        (uint16_t)(1 - (tap >> 1)), bar);
  }
};
struct ColsMNF32 {
  ALEGPU_RAW_F32
  static constexpr bool kMN = true;
  const float *p;
  int ld;
  __device__ __forceinline__ F32x8 load(int i0, int r, int) const {
    return ldg_f32x8(p + (size_t)r * ld + i0);
  }
};
// conv3 dgrad A: rows = act2 pixels (n,iy,ix) of 9x9, r = (ky*3+kx)*64 + co
template <class T> struct DCol3 {
  static constexpr int kElemBytes = (int)sizeof(T);
  ALEGPU_RAW_BF16
  static constexpr bool kMN = false;
  const T *dy; // dact3 [n,7,7,64]
  __device__ __forceinline__ uint4 load(int i, int r, int) const {
    const int n = i / 81, pix = i - n * 81, iy = pix / 9, ix = pix - iy * 9;
    const int t = r >> 6, ky = t / 3, kx = t - ky * 3, co = r & 63;
    const int oy = iy - ky, ox = ix - kx;
    if (oy < 0 || oy >= 7 || ox < 0 || ox >= 7)
      return zero16();
    return ldg16(dy + ((size_t)(n * 7 + oy) * 7 + ox) * 64 + co);
  }
  static constexpr bool kAsync = true;
  struct RowCtx {
    const T *base; // dy element of (n, oy = iy, ox = ix), channel 0 (may lie outside: only offset)
    int iy, ix;
  };
  struct ColCtx {
    int off, ky, kx; // off = co - (ky*7 + kx)*64
  };
  __device__ __forceinline__ RowCtx row(int i) const {
    const int n = i / 81, pix = i - n * 81, iy = pix / 9, ix = pix - iy * 9;
    return RowCtx{dy + ((ptrdiff_t)(n * 7 + iy) * 7 + ix) * 64, iy, ix};
  }
  __device__ __forceinline__ ColCtx col(int r, int) const {
    const int t = r >> 6, ky = t / 3, kx = t - ky * 3, co = r & 63;
    return ColCtx{co - (ky * 7 + kx) * 64, ky, kx};
  }
  __device__ __forceinline__ const void *anchor() const { return dy; }
  __device__ __forceinline__ const void *at(const RowCtx &rc, const ColCtx &cc) const {
    const unsigned oy = (unsigned)(rc.iy - cc.ky), ox = (unsigned)(rc.ix - cc.kx);
    return (oy < 7u && ox < 7u) ? (const void *)(rc.base + cc.off) : nullptr;
  }
};
// conv3 dgrad B: (ci0..ci0+7, r = (t,co)) = c3w[co][t*64 + ci]
template <class T> struct W3T {
  static constexpr int kElemBytes = (int)sizeof(T);
  ALEGPU_RAW_BF16
  static constexpr bool kMN = true;
  const T *w;
  __device__ __forceinline__ uint4 load(int ci0, int r, int) const {
    const int t = r >> 6, co = r & 63;
    return ldg16(w + (size_t)co * 576 + t * 64 + ci0);
  }
  static constexpr bool kAsync = true;
  using RowCtx = int;
  using ColCtx = const T *;
  __device__ __forceinline__ RowCtx row(int ci0) const { return ci0; }
  __device__ __forceinline__ ColCtx col(int r, int) const {
    const int t = r >> 6, co = r & 63;
    return w + (size_t)co * 576 + t * 64;
  }
  __device__ __forceinline__ const void *anchor() const { return w; }
  __device__ __forceinline__ const void *at(RowCtx rc, ColCtx cc) const { return cc + rc; }
};
// conv2 dgrad A, parity class z = py*2+px: rows (n,qy,qx) of 10x10, r = (ty*2+tx)*64 + co
template <class T> struct DCol2 {
  static constexpr int kElemBytes = (int)sizeof(T);
  ALEGPU_RAW_BF16
  static constexpr bool kMN = false;
  const T *dy; // dact2 [n,9,9,64]
  __device__ __forceinline__ uint4 load(int i, int r, int) const {
    const int n = i / 100, pix = i - n * 100, qy = pix / 10, qx = pix - qy * 10;
    const int t = r >> 6, ty = t >> 1, tx = t & 1, co = r & 63;
    const int oy = qy - ty, ox = qx - tx;
    if (oy < 0 || oy >= 9 || ox < 0 || ox >= 9)
      return zero16();
    return ldg16(dy + ((size_t)(n * 9 + oy) * 9 + ox) * 64 + co);
  }
  static constexpr bool kAsync = true;
  struct RowCtx {
    const T *base; // dy element of (n, oy = qy, ox = qx), channel 0 (only an offset base)
    int qy, qx;
  };
  struct ColCtx {
    int off, ty, tx; // off = co - (ty*9 + tx)*64
  };
  __device__ __forceinline__ RowCtx row(int i) const {
    const int n = i / 100, pix = i - n * 100, qy = pix / 10, qx = pix - qy * 10;
    return RowCtx{dy + ((ptrdiff_t)(n * 9 + qy) * 9 + qx) * 64, qy, qx};
  }
  __device__ __forceinline__ ColCtx col(int r, int) const {
    const int t = r >> 6, ty = t >> 1, tx = t & 1, co = r & 63;
    return ColCtx{co - (ty * 9 + tx) * 64, ty, tx};
  }
  __device__ __forceinline__ const void *anchor() const { return dy; }
  __device__ __forceinline__ const void *at(const RowCtx &rc, const ColCtx &cc) const {
    const unsigned oy = (unsigned)(rc.qy - cc.ty), ox = (unsigned)(rc.qx - cc.tx);
    return (oy < 9u && ox < 9u) ? (const void *)(rc.base + cc.off) : nullptr;
  }
};
// conv2 dgrad B: (ci0.., r=(t,co), z) = c2w[co][(ky*4+kx)*32 + ci], ky = py+2ty, kx = px+2tx
template <class T> struct W2T {
  static constexpr int kElemBytes = (int)sizeof(T);
  ALEGPU_RAW_BF16
  static constexpr bool kMN = true;
  const T *w;
  __device__ __forceinline__ uint4 load(int ci0, int r, int z) const {
    const int t = r >> 6, ty = t >> 1, tx = t & 1, co = r & 63;
    const int ky = (z >> 1) + 2 * ty, kx = (z & 1) + 2 * tx;
    return ldg16(w + (size_t)co * 512 + (ky * 4 + kx) * 32 + ci0);
  }
  static constexpr bool kAsync = true;
  using RowCtx = int;
  using ColCtx = const T *;
  __device__ __forceinline__ RowCtx row(int ci0) const { return ci0; }
  __device__ __forceinline__ ColCtx col(int r, int z) const {
    const int t = r >> 6, ty = t >> 1, tx = t & 1, co = r & 63;
    const int ky = (z >> 1) + 2 * ty, kx = (z & 1) + 2 * tx;
    return w + (size_t)co * 512 + (ky * 4 + kx) * 32;
  }
  __device__ __forceinline__ const void *anchor() const { return w; }
  __device__ __forceinline__ const void *at(RowCtx rc, ColCtx cc) const { return cc + rc; }
};
// conv2 dgrad B for all four parity classes at once: row j = cls*32 + ci (cls = py*2+px),
// r = (t,co): c2w[co][(ky*4+kx)*32 + ci], ky = py+2ty, kx = px+2tx.  One pass over the (shared) A
// operand then produces the whole 2x2 output block of every (qy,qx).
template <class T> struct W2TAll {
  static constexpr int kElemBytes = (int)sizeof(T);
  ALEGPU_RAW_BF16
  static constexpr bool kMN = true;
  static constexpr bool kAsync = true;
  const T *w;
  struct RowCtx {
    int off; // (py*4 + px)*32 + ci0
  };
  using ColCtx = const T *; // w + co*512 + (2ty*4 + 2tx)*32
  __device__ __forceinline__ RowCtx row(int j0) const {
    const int cls = j0 >> 5;
    return RowCtx{((cls >> 1) * 4 + (cls & 1)) * 32 + (j0 & 31)};
  }
  __device__ __forceinline__ ColCtx col(int r, int) const {
    const int t = r >> 6, ty = t >> 1, tx = t & 1, co = r & 63;
    return w + (size_t)co * 512 + (ty * 8 + tx * 2) * 32;
  }
  __device__ __forceinline__ const void *anchor() const { return w; }
  __device__ __forceinline__ const void *at(const RowCtx &rc, ColCtx cc) const { return cc + rc.off; }
  __device__ __forceinline__ uint4 load(int j0, int r, int z) const { return ldg16(at(row(j0), col(r, z))); }
};

// ---- epilogues (16 consecutive columns of one row per call) --------------------------------------
__device__ __forceinline__ void store_bf16x16(bf16 *dst, const float (&v)[16]) {
  uint4 a = make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]),
                       pack_bf16(v[6], v[7]));
  uint4 b = make_uint4(pack_bf16(v[8], v[9]), pack_bf16(v[10], v[11]), pack_bf16(v[12], v[13]),
                       pack_bf16(v[14], v[15]));
  reinterpret_cast<uint4 *>(dst)[0] = a;
  reinterpret_cast<uint4 *>(dst)[1] = b;
}
__device__ __forceinline__ void load_bf16x16(const bf16 *src, float (&y)[16]) {
  const uint4 a = reinterpret_cast<const uint4 *>(src)[0], b = reinterpret_cast<const uint4 *>(src)[1];
  const uint32_t w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
  for (int t = 0; t < 8; ++t) {
    y[2 * t] = __uint_as_float(w[t] << 16);
    y[2 * t + 1] = __uint_as_float(w[t] & 0xFFFF0000u);
  }
}
// fp32 -> the value the tensor core will actually multiply: round-to-nearest tf32 (the MMA reads the
// upper 19 bits of an fp32 operand, so rounding where the data is produced replaces its truncation)
__device__ __forceinline__ float round_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}
__device__ __forceinline__ void store_x16(bf16 *dst, const float (&v)[16]) { store_bf16x16(dst, v); }
__device__ __forceinline__ void store_x16(float *dst, const float (&v)[16]) {
#pragma unroll
  for (int t = 0; t < 4; ++t)
    reinterpret_cast<float4 *>(dst)[t] = make_float4(round_tf32(v[4 * t]), round_tf32(v[4 * t + 1]),
                                                     round_tf32(v[4 * t + 2]), round_tf32(v[4 * t + 3]));
}
__device__ __forceinline__ void load_x16(const bf16 *src, float (&y)[16]) { load_bf16x16(src, y); }
__device__ __forceinline__ void load_x16(const float *src, float (&y)[16]) {
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    const float4 q = reinterpret_cast<const float4 *>(src)[t];
    y[4 * t] = q.x; y[4 * t + 1] = q.y; y[4 * t + 2] = q.z; y[4 * t + 3] = q.w;
  }
}
template <class T> struct EpiBiasRelu { // out[i][j] = relu(acc*scale + bias[j]) -> T, ld % 16 == 0
  T *out;
  const float *bias;
  int ld;
  float scale;
  __device__ void operator()(int i, int j0, float (&v)[16], int) const {
#pragma unroll
    for (int t = 0; t < 16; ++t)
      v[t] = fmaxf(fmaf(v[t], scale, __ldg(bias + j0 + t)), 0.0f);
    store_x16(out + (size_t)i * ld + j0, v);
  }
};
struct EpiBiasF32 { // out[i][j] = acc + bias[j] -> fp32 (the FC layer has no activation)
  float *out;
  const float *bias;
  int ld;
  __device__ void operator()(int i, int j0, float (&v)[16], int) const {
    float4 *o = reinterpret_cast<float4 *>(out + (size_t)i * ld + j0);
#pragma unroll
    for (int t = 0; t < 4; ++t)
      o[t] = make_float4(v[4 * t] + __ldg(bias + j0 + 4 * t), v[4 * t + 1] + __ldg(bias + j0 + 4 * t + 1),
                         v[4 * t + 2] + __ldg(bias + j0 + 4 * t + 2),
                         v[4 * t + 3] + __ldg(bias + j0 + 4 * t + 3));
  }
};
template <class T> struct EpiReluMask { // out[i][j] = y[i][j] > 0 ? acc : 0 -> T
  T *out;
  const T *y;
  int ld;
  __device__ void operator()(int i, int j0, float (&v)[16], int) const {
    const size_t o = (size_t)i * ld + j0;
    float yy[16];
    load_x16(y + o, yy);
#pragma unroll
    for (int t = 0; t < 16; ++t)
      v[t] = yy[t] > 0.0f ? v[t] : 0.0f;
    store_x16(out + o, v);
  }
};
// FC data gradient: dact3 = relu'(act3) * acc into BOTH layouts — compact [n,7,7,64] (conv3's weight
// gradient reads it) and zero-bordered [n,10,10,64] (the shift-convolution form of conv3's data gradient:
// the untouched columns / rows 7..9 are the padding)
template <class T> struct EpiReluMaskFcPad {
  T *out, *out_pad;
  const T *y;
  __device__ void operator()(int i, int j0, float (&v)[16], int) const {
    const size_t o = (size_t)i * 3136 + j0;
    float yy[16];
    load_x16(y + o, yy);
#pragma unroll
    for (int t = 0; t < 16; ++t)
      v[t] = yy[t] > 0.0f ? v[t] : 0.0f;
    store_x16(out + o, v);
    const int p = j0 >> 6, py = p / 7, px = p - py * 7;
    store_x16(out_pad + ((size_t)i * 100 + py * 10 + px) * 64 + (j0 & 63), v);
  }
};
// conv3 data gradient as a shift convolution: row i of the compact [n,9,9,64] layout (conv2's weight
// gradient reads it), row r of the zero-bordered [n,10,10,64] grid it was computed on (conv2's data
// gradient reads that)
template <class T> struct EpiReluMaskPad {
  T *out, *out_pad;
  const T *y;
  __device__ void operator()(int i, int j0, float (&v)[16], int r) const {
    const size_t o = (size_t)i * 64 + j0;
    float yy[16];
    load_x16(y + o, yy);
#pragma unroll
    for (int t = 0; t < 16; ++t)
      v[t] = yy[t] > 0.0f ? v[t] : 0.0f;
    store_x16(out + o, v);
    store_x16(out_pad + (size_t)r * 64 + j0, v);
  }
};
template <class T> struct EpiReluMaskConv2 { // scatter parity class z back into NHWC [n,20,20,32]
  T *out;
  const T *y;
  __device__ void operator()(int i, int j0, float (&v)[16], int z) const {
    const int n = i / 100, pix = i - n * 100, qy = pix / 10, qx = pix - qy * 10;
    const size_t o = ((size_t)(n * 20 + 2 * qy + (z >> 1)) * 20 + 2 * qx + (z & 1)) * 32 + j0;
    float yy[16];
    load_x16(y + o, yy);
#pragma unroll
    for (int t = 0; t < 16; ++t)
      v[t] = yy[t] > 0.0f ? v[t] : 0.0f;
    store_x16(out + o, v);
  }
};
template <class T> struct EpiReluMaskConv2All { // columns j = cls*32 + ci: all four parity classes of one (n,qy,qx) row
  T *out;
  const T *y;
  __device__ void operator()(int i, int j0, float (&v)[16], int) const {
    const int n = i / 100, pix = i - n * 100, qy = pix / 10, qx = pix - qy * 10, cls = j0 >> 5;
    const size_t o =
        ((size_t)(n * 20 + 2 * qy + (cls >> 1)) * 20 + 2 * qx + (cls & 1)) * 32 + (j0 & 31);
    float yy[16];
    load_x16(y + o, yy);
#pragma unroll
    for (int t = 0; t < 16; ++t)
      v[t] = yy[t] > 0.0f ? v[t] : 0.0f;
    store_x16(out + o, v);
  }
};
struct EpiPartial { // ws[z][i][j] fp32, row length ldn (any), columns >= ldn skipped
  float *ws;
  int rows, ldn;
  __device__ void operator()(int i, int j0, float (&v)[16], int z) const {
    float *o = ws + ((size_t)z * rows + i) * ldn + j0;
    if ((ldn & 3) == 0 && j0 + 16 <= ldn) {
#pragma unroll
      for (int t = 0; t < 4; ++t)
        reinterpret_cast<float4 *>(o)[t] = make_float4(v[4 * t], v[4 * t + 1], v[4 * t + 2], v[4 * t + 3]);
    } else {
#pragma unroll
      for (int t = 0; t < 16; ++t)
        if (j0 + t < ldn)
          o[t] = v[t];
    }
  }
};
struct EpiWgradDirect { // gw[i*K + j] = acc (j < K), gb[i] = acc (j == K)
  float *gw, *gb;
  int K;
  __device__ void operator()(int i, int j0, float (&v)[16], int) const {
    if (j0 + 16 <= K) {
      float4 *o = reinterpret_cast<float4 *>(gw + (size_t)i * K + j0);
#pragma unroll
      for (int t = 0; t < 4; ++t)
        o[t] = make_float4(v[4 * t], v[4 * t + 1], v[4 * t + 2], v[4 * t + 3]);
    } else {
#pragma unroll
      for (int t = 0; t < 16; ++t) {
        if (j0 + t < K)
          gw[(size_t)i * K + j0 + t] = v[t];
        else if (j0 + t == K)
          gb[i] = v[t];
      }
    }
  }
};

// gw[co*K + k] = scale * sum_z ws[z][k][co] (k < K), gb[co] = sum_z ws[z][K][co].
// One block per (k, 32 channels); the 8 thread rows sum interleaved z subsets which are then
// combined in fixed order, so the result is deterministic.  block = (32, 8), grid = (K+1, C/32)
__global__ void __launch_bounds__(256)
    reduce_wgrad_t_kernel(const float *__restrict__ ws, int splits, int K, int C, float scale,
                          float *__restrict__ gw, float *__restrict__ gb, bool conv1_perm) {
  __shared__ float part[8][32];
  const int k = blockIdx.x, co = blockIdx.y * 32 + threadIdx.x, ty = threadIdx.y;
  const size_t per = (size_t)(K + 1) * C;
  const float *p = ws + (size_t)k * C + co;
  float s = 0.0f;
  for (int z = ty; z < splits; z += 8)
    s += p[(size_t)z * per];
  part[ty][threadIdx.x] = s;
  __syncthreads();
  if (ty == 0) {
    float t = part[0][threadIdx.x];
#pragma unroll
    for (int g = 1; g < 8; ++g)
      t += part[g][threadIdx.x];
    if (k < K) // conv1 runs in the k' order of Col1N; the gradient goes back to OIHW
      gw[(size_t)co * K + (conv1_perm ? conv1_k_of_kprime(k) : k)] = t * scale;
    else
      gb[co] = t;
  }
}
// fc[i][j] = bias[j] + sum_z ws[z][i][j]  (split-K forward of the FC layer at small minibatches)
__global__ void reduce_fc_kernel(const float4 *__restrict__ ws, int splits, int64_t total4, int H4,
                                 const float4 *__restrict__ bias, float4 *__restrict__ out) {
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total4;
       idx += (int64_t)gridDim.x * blockDim.x) {
    float4 s = bias[idx % H4];
    for (int z = 0; z < splits; ++z) {
      const float4 v = ws[(size_t)z * total4 + idx];
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
    out[idx] = s;
  }
}
// gw[i*K + j] (j < K) / gb[i] (j == K) = sum_z ws[z][i][j]
__global__ void reduce_wgrad_kernel(const float *__restrict__ ws, int splits, int rows, int K,
                                    float *__restrict__ gw, float *__restrict__ gb) {
  const int ldn = K + 1;
  const int64_t total = (int64_t)rows * ldn;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    float s = 0.0f;
    for (int z = 0; z < splits; ++z)
      s += ws[(size_t)z * total + idx];
    const int i = (int)(idx / ldn), j = (int)(idx - (int64_t)i * ldn);
    if (j < K)
      gw[(size_t)i * K + j] = s;
    else
      gb[i] = s;
  }
}

// conv1 wgrad (space-to-depth form): gw[oc*256 + k] = scale * sum_z ws[z][(tap,oc)][ch] with
// k = conv1_k_of_kprime(tap*64 + ch); gb[oc] = sum over the column-sum partials bp[g][oc].
// grid = 256 blocks of weights + 1 block of biases.
__global__ void __launch_bounds__(1024)
    reduce_c1_s2d_kernel(const float *__restrict__ ws, int splits, float scale,
                         const float *__restrict__ bp, int bias_parts, float *__restrict__ gw,
                         float *__restrict__ gb) {
  // block (32, 32): 32 outputs x 32 interleaved z subsets (up to 296 splits: <= 10 loads per thread, all
  // independent), combined in fixed order -> deterministic.  (8 subsets of 37 dependent loads took 19.7 us
  // at 296 splits; this form is latency-shallow.)
  __shared__ float part[32][33];
  const int ty = threadIdx.y, tx = threadIdx.x;
  const bool bias = blockIdx.x == 256;
  const int idx = blockIdx.x * 32 + tx;
  const float *p = bias ? bp + tx : ws + idx;
  const size_t stride = bias ? 32 : 128 * 64;
  const int count = bias ? bias_parts : splits;
  float s0 = 0.0f, s1 = 0.0f;
  int z = ty;
  for (; z + 32 < count; z += 64) { // two independent chains per thread
    s0 += p[(size_t)z * stride];
    s1 += p[(size_t)(z + 32) * stride];
  }
  if (z < count)
    s0 += p[(size_t)z * stride];
  part[ty][tx] = s0 + s1;
  __syncthreads();
  if (ty == 0) {
    float t = part[0][tx];
#pragma unroll
    for (int g = 1; g < 32; ++g)
      t += part[g][tx];
    if (bias) {
      gb[tx] = t;
    } else {
      const int m = idx >> 6, ch = idx & 63, tap = m >> 5, oc = m & 31;
      gw[oc * 256 + conv1_k_of_kprime(tap * 64 + ch)] = t * scale;
    }
  }
}
// column sums of dact1 [rows][32] (the conv1 bias gradient): block g sums its contiguous row
// range -> bp[g][32].  A thread owns one 16-byte chunk (8 bf16 / 4 fp32 channels) of every kLanes-th
// row; the row lanes are combined in fixed order.
constexpr int kBiasParts = 296;
template <class T>
__global__ void __launch_bounds__(256)
    c1_bias_partial_kernel(const T *__restrict__ dy, int64_t rows, float *__restrict__ bp) {
  constexpr int kEPC = 16 / (int)sizeof(T), kCPR = 32 / kEPC, kLanes = 256 / kCPR;
  __shared__ float part[kLanes][33];
  const int64_t per = (rows + gridDim.x - 1) / gridDim.x;
  const int64_t r0 = (int64_t)blockIdx.x * per, r1 = r0 + per < rows ? r0 + per : rows;
  const int chk = threadIdx.x % kCPR, rl = threadIdx.x / kCPR;
  float acc[kEPC];
#pragma unroll
  for (int j = 0; j < kEPC; ++j)
    acc[j] = 0.f;
  for (int64_t r = r0 + rl; r < r1; r += kLanes) {
    const uint4 v = ldg16(dy + r * 32 + chk * kEPC);
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    if constexpr (sizeof(T) == 2) {
#pragma unroll
      for (int j = 0; j < 4; ++j) { // bf16 -> fp32: the bits are the upper half
        acc[2 * j] += __uint_as_float(w[j] << 16);
        acc[2 * j + 1] += __uint_as_float(w[j] & 0xFFFF0000u);
      }
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j)
        acc[j] += __uint_as_float(w[j]);
    }
  }
#pragma unroll
  for (int j = 0; j < kEPC; ++j)
    part[rl][chk * kEPC + j] = acc[j];
  __syncthreads();
  if (threadIdx.x < 32) {
    float t = 0.0f;
    for (int g = 0; g < kLanes; ++g)
      t += part[g][threadIdx.x];
    bp[blockIdx.x * 32 + threadIdx.x] = t;
  }
}

// how to split a reduction of length R over CTAs when the output has `tiles` tiles
void plan_splits(int64_t R, int64_t tiles, size_t per_split_floats, size_t ws_floats, int *splits,
                 int *r_per_split) {
  int64_t s = std::max<int64_t>(1, (2 * kNumSMs) / tiles);
  s = std::min<int64_t>(s, std::max<int64_t>(1, R / (2 * tc::BK)));
  s = std::min<int64_t>(s, (int64_t)(ws_floats / per_split_floats));
  s = std::max<int64_t>(s, 1);
  int64_t rps = ceil_div(ceil_div(R, s), tc::BK) * tc::BK;
  *r_per_split = (int)rps;
  *splits = (int)ceil_div(R, rps);
}

// conv weight gradient: ws[z][k][co] partials of  sum_m col(m,k) * dY[m][co], then fixed-order sum
template <class T, class Col, int K, int C>
int conv_wgrad(const Col &col, const T *dy, int64_t Mtot, float scale, float *gw, float *gb,
               float *ws, size_t ws_floats, cudaStream_t s, bool conv1_perm = false,
               const void *dy_map = nullptr /* TMA map of dy's [rows x 64] view, or nullptr = cp.async */) {
  const int64_t mtiles = ceil_div(K + 1, tc::BM);
  const size_t per = (size_t)(K + 1) * C;
  ALEGPU_REQUIRE(ws_floats >= per, ALEGPU_RUNTIME_ERROR, "wgrad: workspace too small");
  int splits, rps;
  plan_splits(Mtot, mtiles, per, ws_floats, &splits, &rps);
  if (dy_map && C == 64) {
    if (int rc = tc::launch_gemm_tc<64, 4>(ColMN<Col, K>{col}, ColsMNTma<T>{dy_map, 0},
                                           EpiPartial{ws, K + 1, C}, K + 1, C, Mtot, splits, rps, s))
      return rc;
  } else if (int rc = tc::launch_gemm_tc<64, 4>(ColMN<Col, K>{col}, ColsMN<T>{dy, C},
                                                EpiPartial{ws, K + 1, C}, K + 1, C, Mtot, splits, rps, s))
    return rc;
  static_assert(C % 32 == 0, "channel count must be a multiple of 32");
  reduce_wgrad_t_kernel<<<dim3(K + 1, C / 32), dim3(32, 8), 0, s>>>(
      ws, splits, K, C, scale, gw, gb, conv1_perm);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}

// conv1 weight gradient from the space-to-depth observations (see DYShift): one 128 x 64 output
// tile, split over the reduction (n*441 blocks); partials + bias partials summed in fixed order.
// The bias column sums run on stream `w` (the wgrad side stream) next to the GEMM; `join` (may be
// null when w == s) is recorded after them and awaited before the final reduction.
template <class T>
int conv1_wgrad_s2d(const T *s2d, const T *dy, int64_t n, float scale, float *gw, float *gb,
                    float *ws, size_t ws_floats, cudaStream_t s, cudaStream_t w, cudaEvent_t join,
                    const void *s2d_map = nullptr, int s2d_row_base = 0,
                    const void *dy_shift_map = nullptr /* tf32: DYShiftTma's im2col map of dact1 */) {
  constexpr size_t per = 128 * 64;
  const size_t bias_floats = (size_t)kBiasParts * 32;
  ALEGPU_REQUIRE(ws_floats >= per + bias_floats, ALEGPU_RUNTIME_ERROR, "conv1 wgrad: workspace too small");
  float *bp = ws + (ws_floats - bias_floats);
  int splits, rps;
  // one output tile: bf16 up to 148 splits (one CTA per SM).  Measured per iteration, v0 / v1, bf16: 296
  // splits 2.41 / 6.84 ms, 148 splits 2.32 / 6.82 ms, 74 splits 2.36 / 6.90 ms, 37 splits 2.58 / 8.07 ms
  // tf32 (twice the operand bytes per reduction row, HBM-bound): 296 splits = two CTAs per SM measured
  // 8.76 -> 8.30 ms per v1 train(); bf16 keeps 148.  ALEGPU_C1_WGRAD_TILES=1|2 overrides (A/B).
  static const int tiles_ab = getenv("ALEGPU_C1_WGRAD_TILES") ? atoi(getenv("ALEGPU_C1_WGRAD_TILES")) : 0;
  plan_splits(n * 441, tiles_ab > 0 ? tiles_ab : (sizeof(T) == 4 ? 1 : 2), per, ws_floats - bias_floats,
              &splits, &rps);
  c1_bias_partial_kernel<T><<<kBiasParts, 256, 0, w>>>(dy, n * 400, bp);
  ALEGPU_COUNT_LAUNCH();
  if (join)
    ALEGPU_CUDA_CHECK(cudaEventRecord(join, w));
  bool launched = false;
  if constexpr (sizeof(T) == 4) {
    if (s2d_map && dy_shift_map) { // both operands by TMA: no producer warp issues a copy
      if (int rc = tc::launch_gemm_tc<64, 4>(DYShiftTma{dy_shift_map}, ColsMNTma<T>{s2d_map, s2d_row_base},
                                             EpiPartial{ws, 128, 64}, 128, 64, n * 441, splits, rps, s))
        return rc;
      launched = true;
    }
  }
  if (launched) {
  } else if (s2d_map) { // the observation operand (read once, un-expanded) through TMA
    if (int rc = tc::launch_gemm_tc<64, 4>(DYShift<T>{dy}, ColsMNTma<T>{s2d_map, s2d_row_base},
                                           EpiPartial{ws, 128, 64}, 128, 64, n * 441, splits, rps, s))
      return rc;
  } else if (int rc = tc::launch_gemm_tc<64, 4>(DYShift<T>{dy}, ColsMN<T>{s2d, 64}, EpiPartial{ws, 128, 64},
                                                128, 64, n * 441, splits, rps, s))
    return rc;
  if (join)
    ALEGPU_CUDA_CHECK(cudaStreamWaitEvent(s, join, 0));
  reduce_c1_s2d_kernel<<<257, dim3(32, 32), 0, s>>>(ws, splits, scale, bp, kBiasParts, gw, gb);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}

__global__ void bf16_to_f32_kernel(const bf16 *__restrict__ src, float *__restrict__ dst, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    dst[i] = __bfloat162float(src[i]);
}

// u8 NCHW [n,4,84,84] -> space-to-depth NHWC [n,21,21,64] in the operand type T (bf16, or fp32 for
// the tf32 path; u8 is exact in both).  One thread = 4 pixels of one row: four coalesced 4-byte loads
// (one per channel plane), one 32-byte (bf16) / 64-byte (fp32) store.  HBM-bound: 28 224 B read +
// 56 448 / 112 896 B written per observation, once per train() (not per epoch / minibatch).
template <class T>
__global__ void __launch_bounds__(256)
    obs_to_s2d_kernel(const uint8_t *__restrict__ obs, T *__restrict__ out, int64_t total) {
  // total = n * 21 * 84 work items (n, Y, x4, y & 3) with y & 3 fastest: four neighbouring threads write
  // the four 16-channel groups of ONE 64-channel block, so a warp writes 8 whole blocks = 1 KB (bf16) /
  // 2 KB (fp32) of contiguous output (the first version ran x4 fastest: 32- / 64-byte pieces with a
  // 128- / 256-byte stride, 0.58 of the copy roofline)
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t blk = i / 84; // n*21 + Y
    const int j = (int)(i - blk * 84);
    const int x4 = j >> 2;
    const int64_t n = blk / 21;
    const int y = (int)(blk - n * 21) * 4 + (j & 3);
    const uint8_t *src = obs + n * 28224 + (size_t)y * 84 + x4 * 4;
    uint32_t w[4];
#pragma unroll
    for (int c = 0; c < 4; ++c)
      w[c] = __ldg(reinterpret_cast<const uint32_t *>(src + c * 7056));
    // space-to-depth: block (Y, X) = (y/4, x4), channel = (y&3)*16 + s*4 + c, s = pixel j of the four
    T *dstp = out + (((size_t)n * 21 + (y >> 2)) * 21 + x4) * 64 + (y & 3) * 16;
    if constexpr (sizeof(T) == 2) {
      uint32_t o[8]; // pixel j: channels 0..3 -> two packed bf16 pairs
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const uint32_t sel = 0x0040u + (uint32_t)j * 0x0011u; // byte j of the 1st, byte j of the 2nd
        const uint32_t px = (__byte_perm(w[0], w[1], sel) & 0xFFFFu) |               // c0, c1
                            (__byte_perm(w[2], w[3], sel) << 16);                    // c2, c3
        // px bytes = (c0, c1, c2, c3) of pixel j
        uint32_t lo, hi;
        u8x4_to_bf16(px, lo, hi);
        o[2 * j] = lo;
        o[2 * j + 1] = hi;
      }
      uint4 *dst = reinterpret_cast<uint4 *>(dstp);
      dst[0] = make_uint4(o[0], o[1], o[2], o[3]);
      dst[1] = make_uint4(o[4], o[5], o[6], o[7]);
    } else {
      float4 *dst = reinterpret_cast<float4 *>(dstp);
#pragma unroll
      for (int j = 0; j < 4; ++j) // pixel j: channels 0..3 as fp32 (exact)
        dst[j] = make_float4((float)((w[0] >> (8 * j)) & 0xFFu), (float)((w[1] >> (8 * j)) & 0xFFu),
                             (float)((w[2] >> (8 * j)) & 0xFFu), (float)((w[3] >> (8 * j)) & 0xFFu));
    }
  }
}
// c1w [32][256] OIHW order -> [32][256] in Col1N's k' order
template <class T> __global__ void permute_c1w_kernel(const T *__restrict__ w, T *__restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < 32 * 256) {
    const int co = i >> 8, kp = i & 255;
    out[i] = w[co * 256 + conv1_k_of_kprime(kp)];
  }
}
// fp32 -> tf32-rounded fp32 (the weight shadow of the tf32 path)
__global__ void f32_round_tf32_kernel(const float *__restrict__ src, float *__restrict__ dst, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    dst[i] = round_tf32(src[i]);
}

template <class T> struct Shadow { // operand-type views of the flat parameter buffer (same offsets as NetParams)
  const T *c1w, *c2w, *c3w, *fcw;
};
template <class T> Shadow<T> shadow_view(const TcState &S, const NetParams &P) {
  const T *base = static_cast<const T *>(S.w_shadow);
  Shadow<T> s;
  s.c1w = base + (P.c1w - P.c1w);
  s.c2w = base + (P.c2w - P.c1w);
  s.c3w = base + (P.c3w - P.c1w);
  s.fcw = base + (P.fcw - P.c1w);
  return s;
}

// ---- TMA im2col tensor maps ---------------------------------------------------------------------------
// kMapObsAll.. : one slot per (re)allocation of obs16_all, used round-robin, so a descriptor that an
// earlier kernel may still have cached is never rewritten in place
enum TmaSlot { kMapAct1 = 0, kMapAct2, kMapDact3, kMapDact2, kMapObsMb, kMapObsAll, kMapObsAllSlots = 16,
               // plain [rows x 64] tiled maps of the same buffers for the shift convolutions (tc_shift.cuh)
               kMapAct2Rows = kMapObsAll + kMapObsAllSlots, kMapAct1Rows, kMapDact3PadRows, kMapDact2PadRows,
               kMapDact3MN, kMapDact2MN, kMapDfcK, kMapAct3K, kMapFcwK, kMapFcwMN, kMapC2wK, kMapC3wTap, kMapC2wTap, kMapDact1Shift, kMapObsMbMN, kMapObsAllMN, // [rows x 64] boxes of one k-block of rows (ColsMNTma)
               kMapObsAllMNEnd = kMapObsAllMN + kMapObsAllSlots - 1,
               kMapObsMbRows, kMapObsAllRows,
               kMapCount = kMapObsAllRows + kMapObsAllSlots };
using EncodeIm2colFn = CUresult (*)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *,
                                    const cuuint64_t *, const cuuint64_t *, const int *, const int *,
                                    cuuint32_t, cuuint32_t, const cuuint32_t *, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeIm2colFn encode_im2col_fn() {
  static EncodeIm2colFn fn = [] {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return (EncodeIm2colFn)p;
  }();
  return fn;
}
// map slot <- NHWC tensor [N,H,W,64] of eb-byte elements at `base`, pixel box [lower, dim + upper) in
// both axes, 128 pixels x one 128-byte row of channels (64 bf16 / 32 fp32) per load, 128-byte swizzle,
// zeros outside
int set_im2col_map(TcState &T, int slot, void *base, int H, int W, int64_t N, int lower, int upper,
                   int upper_h = 1 << 30, int stride_h = 1) {
  if (upper_h == (1 << 30))
    upper_h = upper;
  EncodeIm2colFn fn = encode_im2col_fn();
  ALEGPU_REQUIRE(fn != nullptr, ALEGPU_RUNTIME_ERROR, "cuTensorMapEncodeIm2col is not available");
  alignas(64) CUtensorMap m;
  const cuuint64_t eb = (cuuint64_t)T.elem_bytes, px = 64 * eb;
  const cuuint64_t dims[4] = {64, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  const cuuint64_t strides[3] = {px, (cuuint64_t)W * px, (cuuint64_t)H * W * px};
  const int lo[2] = {lower, lower}, hi[2] = {upper, upper_h};
  const cuuint32_t estr[4] = {1, 1, (cuuint32_t)stride_h, 1};
  const CUresult r = fn(&m, eb == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16,
                        4, base, dims, strides, lo, hi, (cuuint32_t)(128 / eb), 128, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  ALEGPU_REQUIRE(r == CUDA_SUCCESS, ALEGPU_RUNTIME_ERROR,
                 "cuTensorMapEncodeIm2col failed: " + std::to_string((int)r));
  ALEGPU_CUDA_CHECK(cudaMemcpy((CUtensorMap *)T.tmaps + slot, &m, sizeof(m), cudaMemcpyHostToDevice));
  return ALEGPU_OK;
}
using EncodeTiledFn = CUresult (*)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                   const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                                   CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                   CUtensorMapFloatOOBfill);
EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = [] {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return (EncodeTiledFn)p;
  }();
  return fn;
}
// map slot <- the [rows x 64] matrix view of an NHWC buffer with 64 channels per pixel: boxes of
// box_rows rows (the layer's stage height) x one 128-byte channel panel, 128-byte swizzle, zeros past the last row
int set_rows_map(TcState &T, int slot, void *base, int64_t rows, int box_rows, bool mn_operand = false,
                 int64_t cols = 64) {
  EncodeTiledFn fn = encode_tiled_fn();
  ALEGPU_REQUIRE(fn != nullptr, ALEGPU_RUNTIME_ERROR, "cuTensorMapEncodeTiled is not available");
  alignas(64) CUtensorMap m;
  const cuuint64_t eb = (cuuint64_t)T.elem_bytes;
  const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)cols * eb};
  const cuuint32_t box[2] = {(cuuint32_t)(128 / eb), (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = fn(&m, eb == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                        base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        // an MN-major tf32 operand tile is swizzled in 32-byte units (UMMA SWIZZLE_128B_BASE32B)
                        mn_operand && eb == 4 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  ALEGPU_REQUIRE(r == CUDA_SUCCESS, ALEGPU_RUNTIME_ERROR,
                 "cuTensorMapEncodeTiled failed: " + std::to_string((int)r));
  ALEGPU_CUDA_CHECK(cudaMemcpy((CUtensorMap *)T.tmaps + slot, &m, sizeof(m), cudaMemcpyHostToDevice));
  return ALEGPU_OK;
}
// map slot <- a weight tensor seen as [co][taps'][group elements]: boxes {128 bytes, 1, one k-block of co},
// swizzled as an MN-major operand tile
int set_wtap_map(TcState &T, int slot, void *base, int group_elems_total, int taps, int co) {
  EncodeTiledFn fn = encode_tiled_fn();
  ALEGPU_REQUIRE(fn != nullptr, ALEGPU_RUNTIME_ERROR, "cuTensorMapEncodeTiled is not available");
  alignas(64) CUtensorMap m;
  const cuuint64_t eb = (cuuint64_t)T.elem_bytes;
  const cuuint64_t dims[3] = {(cuuint64_t)group_elems_total, (cuuint64_t)taps, (cuuint64_t)co};
  const cuuint64_t strides[2] = {(cuuint64_t)group_elems_total * eb, (cuuint64_t)group_elems_total * taps * eb};
  const cuuint32_t box[3] = {(cuuint32_t)(128 / eb), 1, (cuuint32_t)(128 / eb)};
  const cuuint32_t estr[3] = {1, 1, 1};
  const CUresult r = fn(&m, eb == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3,
                        base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        eb == 4 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  ALEGPU_REQUIRE(r == CUDA_SUCCESS, ALEGPU_RUNTIME_ERROR,
                 "cuTensorMapEncodeTiled (weights) failed: " + std::to_string((int)r));
  ALEGPU_CUDA_CHECK(cudaMemcpy((CUtensorMap *)T.tmaps + slot, &m, sizeof(m), cudaMemcpyHostToDevice));
  return ALEGPU_OK;
}
const void *tmap(const TcState &T, int slot) { return (const CUtensorMap *)T.tmaps + slot; }

} // namespace

// ---- public (net_tc.h) ----------------------------------------------------------------------------
int net_tc_init(TcState &T, int64_t hidden, int64_t actions, int64_t cap, int64_t Ppad,
                int elem_bytes, float *dfc_f32, cudaStream_t) {
  ALEGPU_REQUIRE(hidden % 16 == 0, ALEGPU_INVALID_ARGUMENT,
                 "the tensor-core path needs hidden_size % 16 == 0");
  ALEGPU_REQUIRE(elem_bytes == 2 || elem_bytes == 4, ALEGPU_INVALID_ARGUMENT,
                 "tensor-core operands are bf16 (2 bytes) or tf32 (4 bytes)");
  T.hidden = hidden;
  T.actions = actions;
  T.cap = cap;
  T.Ppad = Ppad;
  T.elem_bytes = elem_bytes;
  const size_t eb = (size_t)elem_bytes;
  auto alloc = [](void **p, size_t bytes) -> int {
    ALEGPU_CUDA_CHECK(cudaMalloc(p, bytes ? bytes : 16));
    ALEGPU_CUDA_CHECK(cudaMemset(*p, 0, bytes ? bytes : 16));
    return ALEGPU_OK;
  };
  if (int rc = alloc(&T.w_shadow, eb * (size_t)Ppad)) return rc;
  if (int rc = alloc(&T.c1w_perm, eb * 32 * 256)) return rc;
  if (int rc = alloc(&T.obs16_mb, eb * (size_t)cap * 28224)) return rc;
  if (int rc = alloc(&T.act1, eb * (size_t)cap * 12800)) return rc;
  if (int rc = alloc(&T.act2, eb * (size_t)cap * 5184)) return rc;
  if (int rc = alloc(&T.act3, eb * (size_t)cap * 3136)) return rc;
  if (int rc = alloc(&T.dact1, eb * (size_t)cap * 12800)) return rc;
  if (int rc = alloc(&T.dact2, eb * (size_t)cap * 5184)) return rc;
  if (int rc = alloc(&T.dact3, eb * (size_t)cap * 3136)) return rc;
  // zero-bordered [n,10,10,64] copies of dact3 / dact2 for the shift-convolution data gradients: only
  // the valid 7x7 / 9x9 positions are ever written, the rest stays the zero it was allocated as
  if (int rc = alloc(&T.dact3p, eb * (size_t)cap * 6400)) return rc;
  if (int rc = alloc(&T.dact2p, eb * (size_t)cap * 6400)) return rc;
  if (elem_bytes == 2) // tf32: the tensor core reads the (tf32-rounded) fp32 dfc itself
    if (int rc = alloc(&T.dfc16, eb * (size_t)cap * hidden)) return rc;
  T.ws_heads_n = (size_t)(actions + 1) * (hidden + 1) * 2 * kNumSMs;
  T.ws_fc_n = (size_t)hidden * 3137 * 2;
  T.ws_c3_n = (size_t)577 * 64 * 64;
  T.ws_c2_n = (size_t)513 * 64 * 64;
  T.ws_c1_n = (size_t)128 * 64 * 296 + (size_t)296 * 32; // conv1 wgrad: <= 296 split partials + bias partials
  if (int rc = alloc((void **)&T.ws_heads, sizeof(float) * T.ws_heads_n)) return rc;
  if (int rc = alloc((void **)&T.ws_fc, sizeof(float) * T.ws_fc_n)) return rc;
  if (int rc = alloc((void **)&T.ws_c3, sizeof(float) * T.ws_c3_n)) return rc;
  if (int rc = alloc((void **)&T.ws_c2, sizeof(float) * T.ws_c2_n)) return rc;
  if (int rc = alloc((void **)&T.ws_c1, sizeof(float) * T.ws_c1_n)) return rc;
  ALEGPU_CUDA_CHECK(cudaStreamCreateWithFlags(&T.side, cudaStreamNonBlocking));
  for (cudaEvent_t *e : {&T.ev_fork, &T.ev_d3, &T.ev_d2, &T.ev_d1, &T.ev_join, &T.ev_fcw})
    ALEGPU_CUDA_CHECK(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
  const char *env = getenv("ALEGPU_TC_SERIAL");
  T.serial = env && env[0] == '1';
  env = getenv("ALEGPU_CONV1_U8"); // the u8 register producers convert to bf16: bf16 mode only
  T.conv1_u8 = env && env[0] == '1' && elem_bytes == 2;
  env = getenv("ALEGPU_CONV1_WGRAD_U8");
  T.conv1_wgrad_u8 = env && env[0] == '1' && elem_bytes == 2;
  env = getenv("ALEGPU_NO_TMA");
  T.use_tma = !(env && env[0] == '1');
  if (T.use_tma) {
    ALEGPU_CUDA_CHECK(cudaMalloc(&T.tmaps, sizeof(CUtensorMap) * kMapCount));
    if (int rc = set_im2col_map(T, kMapAct1, T.act1, 20, 10, cap, 0, -1, -3, 2)) return rc; // pairs
    if (int rc = set_im2col_map(T, kMapAct2, T.act2, 9, 9, cap, 0, -2)) return rc;   // 3x3, no padding
    if (int rc = set_im2col_map(T, kMapDact3, T.dact3, 7, 7, cap, -2, 0)) return rc; // 3x3, padding 2
    if (int rc = set_im2col_map(T, kMapDact2, T.dact2, 9, 9, cap, -1, 0)) return rc; // 2x2, padding 1
    if (int rc = set_im2col_map(T, kMapObsMb, T.obs16_mb, 21, 21, cap, 0, -1)) return rc; // 2x2 taps
    if (int rc = set_rows_map(T, kMapAct2Rows, T.act2, cap * 81, tc::shift_stage_rows(3, 3, 9))) return rc;
    if (int rc = set_rows_map(T, kMapAct1Rows, T.act1, cap * 200, tc::shift_stage_rows(4, 2, 10))) return rc; // pixel pairs [n,20,10,64]
    if (int rc = set_rows_map(T, kMapDact3PadRows, T.dact3p, cap * 100, tc::shift_stage_rows(3, 3, 10))) return rc;
    if (int rc = set_rows_map(T, kMapDact2PadRows, T.dact2p, cap * 100, tc::shift_stage_rows(2, 2, 10))) return rc;
    const int bk = 128 / elem_bytes; // rows of one k-block
    if (int rc = set_rows_map(T, kMapDact3MN, T.dact3, cap * 49, bk, true)) return rc;
    if (int rc = set_rows_map(T, kMapDact2MN, T.dact2, cap * 81, bk, true)) return rc;
    if (int rc = set_rows_map(T, kMapObsMbMN, T.obs16_mb, cap * 441, bk, true)) return rc;
    // the FC layers' plain matrices: dfc [cap][H], act3 [cap][3136], the fc weight shadow [H][3136] (K-major
    // 64-row boxes, and the transposed MN-major view the data gradient reads)
    // conv weights of the shadow: c2w [64][512] K-major rows (conv2 forward), and the per-tap views the data
    // gradients read (conv3: [64 co][9][64 ci]; conv2: [64 co][8 tap pairs][64] in bf16, [64 co][16 taps][32] in tf32)
    char *c2w = (char *)T.w_shadow + eb * (size_t)(32 * 256 + 32);
    char *c3w = c2w + eb * (size_t)(64 * 512 + 64);
    if (int rc = set_rows_map(T, kMapC2wK, c2w, 64, 64, false, 512)) return rc;
    if (int rc = set_wtap_map(T, kMapC3wTap, c3w, 64, 9, 64)) return rc;
    if (int rc = set_wtap_map(T, kMapC2wTap, c2w, elem_bytes == 2 ? 64 : 32, elem_bytes == 2 ? 8 : 16, 64)) return rc;
    if (elem_bytes == 4) { // dact1 [n,20,20,32] fp32 for DYShiftTma: 32-pixel x 32-channel boxes, MN-major swizzle
      EncodeIm2colFn fn = encode_im2col_fn();
      ALEGPU_REQUIRE(fn != nullptr, ALEGPU_RUNTIME_ERROR, "cuTensorMapEncodeIm2col is not available");
      alignas(64) CUtensorMap m;
      const cuuint64_t dims[4] = {32, 20, 20, (cuuint64_t)cap};
      const cuuint64_t strides[3] = {128, 20 * 128, 400 * 128};
      const int lo[2] = {-1, -1}, hi[2] = {0, 0};
      const cuuint32_t estr[4] = {1, 1, 1, 1};
      const CUresult r = fn(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, T.dact1, dims, strides, lo, hi, 32, 32, estr,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B,
                            CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      ALEGPU_REQUIRE(r == CUDA_SUCCESS, ALEGPU_RUNTIME_ERROR,
                     "cuTensorMapEncodeIm2col (dact1) failed: " + std::to_string((int)r));
      ALEGPU_CUDA_CHECK(cudaMemcpy((CUtensorMap *)T.tmaps + kMapDact1Shift, &m, sizeof(m), cudaMemcpyHostToDevice));
    }
    void *dfc = elem_bytes == 2 ? T.dfc16 : (void *)dfc_f32;
    char *fcw = (char *)T.w_shadow + eb * (size_t)(32 * 256 + 32 + 64 * 512 + 64 + 64 * 576 + 64);
    T.fc_tma = dfc != nullptr;
    if (T.fc_tma) {
      if (int rc = set_rows_map(T, kMapDfcK, dfc, cap, 64, false, hidden)) return rc;
      if (int rc = set_rows_map(T, kMapAct3K, T.act3, cap, 64, false, 3136)) return rc;
      if (int rc = set_rows_map(T, kMapFcwK, fcw, hidden, 64, false, 3136)) return rc;
      if (int rc = set_rows_map(T, kMapFcwMN, fcw, hidden, bk, true, 3136)) return rc;
    }
    if (int rc = set_rows_map(T, kMapObsMbRows, T.obs16_mb, cap * 441, tc::shift_stage_rows(2, 2, 21))) return rc;
  }
  env = getenv("ALEGPU_NO_TMA_WEIGHTS"); // A/B: conv2 forward / conv3, conv2 data-gradient weights via cp.async
  T.use_tma_weights = T.use_tma && !(env && env[0] == '1');
  env = getenv("ALEGPU_NO_TMA_FC"); // A/B: the FC forward / data-gradient operands through cp.async
  T.fc_tma = T.fc_tma && T.use_tma && !(env && env[0] == '1');
  env = getenv("ALEGPU_NO_TMA_WGRAD"); // A/B: the weight gradients' plain-matrix operand through cp.async
  T.use_tma_wgrad = T.use_tma && !(env && env[0] == '1');
  env = getenv("ALEGPU_NO_SHIFT"); // A/B: conv1 / conv3 forward through the im2col engine instead
  T.use_shift = T.use_tma && !(env && env[0] == '1');
  // conv3 / conv2 data gradients as shift convolutions over zero-bordered gradient copies: measured at
  // 1280 rows, bf16 28.4 -> 26.5 us and 40.8 -> 38.9 us (train() 5.44 -> 5.24 ms); tf32 47.4 -> 49.1 us
  // and 67.5 -> 63.9 us with the extra bordered stores costing more than that (8.30 -> 8.51 ms): in tf32
  // these kernels are bound by the tensor core's operand fetch from shared memory (a K = 8 UMMA of
  // 128 x 64 moves 6 KB for 65 k MACs, half the bf16 ratio), not by L2 traffic.  So: on for bf16, off
  // for tf32; ALEGPU_SHIFT_DGRAD=0|1 overrides (A/B).
  env = getenv("ALEGPU_SHIFT_DGRAD");
  T.use_shift_dgrad = T.use_shift && (env ? env[0] == '1' : elem_bytes == 2);
  // the matrix-base-offset field of the row-shifted descriptors stays 0: measured on B200, the 128-byte
  // swizzle is a function of the absolute shared-memory address, so a descriptor may start at any
  // 128-byte row of a 1024-byte aligned panel as it is (setting (start >> 7) & 7 there gives wrong
  // products; ALEGPU_SHIFT_BASE_OFFSET=1 reproduces that)
  env = getenv("ALEGPU_SHIFT_BASE_OFFSET");
  T.shift_base_offset = env && env[0] == '1';
  return ALEGPU_OK;
}

void net_tc_destroy(TcState &T) {
  for (void *p : {T.c1w_perm, T.obs16_mb, T.obs16_all, T.w_shadow, T.act1, T.act2, T.act3, T.dact1, T.dact2, T.dact3, T.dact3p, T.dact2p, T.fc16, T.dfc16,
                  T.wt, T.tmaps, (void *)T.stage, (void *)T.ws_heads, (void *)T.ws_fc, (void *)T.ws_c3,
                  (void *)T.ws_c2, (void *)T.ws_c1})
    if (p)
      cudaFree(p);
  for (cudaEvent_t e : {T.ev_fork, T.ev_d3, T.ev_d2, T.ev_d1, T.ev_join, T.ev_fcw})
    if (e)
      cudaEventDestroy(e);
  if (T.side)
    cudaStreamDestroy(T.side);
  T = TcState{};
}

int net_tc_refresh_weights(TcState &T, const float *params, int64_t Ppad, cudaStream_t s) {
  if (T.elem_bytes == 4) {
    f32_round_tf32_kernel<<<kNumSMs * 2, 256, 0, s>>>(params, (float *)T.w_shadow, Ppad);
    ALEGPU_COUNT_LAUNCH();
    permute_c1w_kernel<float><<<32, 256, 0, s>>>((const float *)T.w_shadow, (float *)T.c1w_perm);
  } else {
    if (int rc = launch_f32_to_bf16(params, T.w_shadow, Ppad, s))
      return rc;
    permute_c1w_kernel<bf16><<<32, 256, 0, s>>>((const bf16 *)T.w_shadow, (bf16 *)T.c1w_perm); // c1w is first
  }
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}

namespace {
int convert_obs(const TcState &T, const uint8_t *obs, int64_t n, void *out, cudaStream_t s) {
  const int64_t total = n * 84 * 21;
  if (total == 0)
    return ALEGPU_OK;
  // one work item per thread, no grid-stride cap: measured per v1 train() (tf32) 722 us at 148 x 16 blocks,
  // 665 at x64, 646 at x256, 636 uncapped (0.70 of the copy peak); ALEGPU_S2D_BLOCKS caps it for the A/B
  static const int64_t cap_ab = getenv("ALEGPU_S2D_BLOCKS") ? atoll(getenv("ALEGPU_S2D_BLOCKS")) : ((int64_t)1 << 30);
  const int blocks = (int)std::min<int64_t>(ceil_div(total, 256), cap_ab);
  if (T.elem_bytes == 4)
    obs_to_s2d_kernel<float><<<blocks, 256, 0, s>>>(obs, (float *)out, total);
  else
    obs_to_s2d_kernel<bf16><<<blocks, 256, 0, s>>>(obs, (bf16 *)out, total);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}
constexpr int64_t kObs16MaxRows = (int64_t)1 << 17; // 131 072 rows = 7.4 GB of bf16 (14.8 GB of fp32) observations
} // namespace

// allocation only (never inside a stream capture): room for n rows of operand-type observations
int net_tc_reserve_obs(TcState &T, int64_t n) {
  if (T.conv1_u8 || n > kObs16MaxRows || T.obs16_all_rows >= n)
    return ALEGPU_OK;
  if (T.obs16_all)
    ALEGPU_CUDA_CHECK(cudaFree(T.obs16_all));
  T.obs16_all = nullptr;
  T.obs16_all_rows = 0;
  ALEGPU_CUDA_CHECK(cudaMalloc(&T.obs16_all, (size_t)T.elem_bytes * (size_t)n * 28224));
  T.obs16_all_rows = n;
  if (T.use_tma) {
    T.obs_all_map = (T.obs_all_map + 1) % kMapObsAllSlots;
    if (int rc = set_rows_map(T, kMapObsAllRows + T.obs_all_map, T.obs16_all, n * 441, tc::shift_stage_rows(2, 2, 21)))
      return rc;
    if (int rc = set_rows_map(T, kMapObsAllMN + T.obs_all_map, T.obs16_all, n * 441, 128 / T.elem_bytes, true))
      return rc;
    return set_im2col_map(T, kMapObsAll + T.obs_all_map, T.obs16_all, 21, 21, n, 0, -1);
  }
  return ALEGPU_OK;
}

int net_tc_prepare_obs(TcState &T, const uint8_t *obs, int64_t n, const void **obs16,
                       cudaStream_t s) {
  *obs16 = nullptr;
  if (T.conv1_u8 || T.obs16_all_rows < n)
    return ALEGPU_OK; // not reserved (too big to keep): the forward converts per minibatch
  if (int rc = convert_obs(T, obs, n, T.obs16_all, s))
    return rc;
  *obs16 = T.obs16_all;
  return ALEGPU_OK;
}

namespace {
template <class T>
int forward_impl(const NetParams &P, TcState &S, const uint8_t *obs, int64_t n, NetActs &A,
                 cudaStream_t s, const void *obs16) {
  const int H = (int)P.hidden;
  const Shadow<T> W = shadow_view<T>(S, P);
  T *act1 = (T *)S.act1, *act2 = (T *)S.act2, *act3 = (T *)S.act3;
  if constexpr (sizeof(T) == 2) {
    if (S.conv1_u8) { // A/B switch: conv1 straight from the u8 observations (register producers)
      if (int rc = tc::launch_gemm_tc<32, 4>(ColK<Col1>{Col1{obs}}, RowsK<T>{W.c1w, 256},
                                             EpiBiasRelu<T>{act1, P.c1b, 32, 1.0f / 255.0f},
                                             n * 400, 32, 256, 1, 0, s))
        return rc;
      obs16 = nullptr;
    }
  }
  if (!S.conv1_u8) {
    if (!obs16) {
      if (int rc = convert_obs(S, obs, n, S.obs16_mb, s))
        return rc;
      obs16 = S.obs16_mb;
    }
    const T *x16 = (const T *)obs16, *all = (const T *)S.obs16_all;
    const bool in_all = all && x16 >= all && x16 < all + (size_t)S.obs16_all_rows * 28224;
    if (S.use_shift && (x16 == (const T *)S.obs16_mb || in_all)) {
      // shift convolution: the 2x2 taps are row-shifted views of one staged copy of the s2d rows
      if (int rc = tc::launch_conv_shift<T, 32, 2, 2, 21, 21, 20, 20, 1, 3, 1, false>(
              tmap(S, in_all ? kMapObsAllRows + S.obs_all_map : kMapObsMbRows),
              in_all ? (int64_t)((x16 - all) / 28224) * 441 : 0, tc::WKMajor<T>{(const T *)S.c1w_perm, 256},
              EpiBiasRelu<T>{act1, P.c1b, 32, 1.0f / 255.0f}, n, s, S.shift_base_offset))
        return rc;
    } else if (S.use_tma && (x16 == (const T *)S.obs16_mb || in_all)) {
      const Col1Tma<T> a{tmap(S, in_all ? kMapObsAll + S.obs_all_map : kMapObsMb),
                         in_all ? (int)((x16 - all) / 28224) : 0};
      if (int rc = tc::launch_gemm_tc<32, 4>(a, RowsK<T>{(const T *)S.c1w_perm, 256},
                                             EpiBiasRelu<T>{act1, P.c1b, 32, 1.0f / 255.0f},
                                             n * 400, 32, 256, 1, 0, s))
        return rc;
    } else if (int rc = tc::launch_gemm_tc<32, 4>(ColK<Col1N<T>>{Col1N<T>{x16}},
                                                  RowsK<T>{(const T *)S.c1w_perm, 256},
                                                  EpiBiasRelu<T>{act1, P.c1b, 32, 1.0f / 255.0f},
                                                  n * 400, 32, 256, 1, 0, s))
      return rc;
  }
  // conv2 (stride 2) as a shift convolution runs on 200 grid rows per 81 outputs; measured at 1280 rows
  // (tf32): 44.1 us against 40.4 us for the im2col form — the shift kernels are bound by the UMMA issue
  // rate of 128 x 64 x 8 instructions (operands stream from shared memory per instruction), so 2.5x
  // the tensor work costs more than the 8x less L2 traffic saves.  ALEGPU_SHIFT_CONV2=1 selects it.
  static const bool shift_conv2 = getenv("ALEGPU_SHIFT_CONV2") != nullptr;
  if (S.use_shift && shift_conv2) { // 4 (ky) x 2 (pixel pairs) taps over act1's pair grid, outputs on its even rows
    if (int rc = tc::launch_conv_shift<T, 64, 4, 2, 20, 10, 9, 9, 2, (sizeof(T) == 4 ? 2 : 3), 1, false>(
            tmap(S, kMapAct1Rows), 0, tc::WKMajor<T>{W.c2w, 512}, EpiBiasRelu<T>{act2, P.c2b, 64, 1.0f}, n, s,
            S.shift_base_offset))
      return rc;
  } else if (S.use_tma && S.use_tma_weights) { // both operands by TMA: im2col boxes and the weight rows
    if (int rc = tc::launch_gemm_tc<64, 4>(Col2Tma<T>{tmap(S, kMapAct1), 0}, RowsKTma<T, 64>{tmap(S, kMapC2wK)},
                                           EpiBiasRelu<T>{act2, P.c2b, 64, 1.0f}, n * 81, 64, 512,
                                           1, 0, s))
      return rc;
  } else if (S.use_tma) {
    if (int rc = tc::launch_gemm_tc<64, 4>(Col2Tma<T>{tmap(S, kMapAct1), 0}, RowsK<T>{W.c2w, 512},
                                           EpiBiasRelu<T>{act2, P.c2b, 64, 1.0f}, n * 81, 64, 512,
                                           1, 0, s))
      return rc;
  } else if (int rc = tc::launch_gemm_tc<64, 4>(ColK<Col2<T>>{Col2<T>{act1}}, RowsK<T>{W.c2w, 512},
                                                EpiBiasRelu<T>{act2, P.c2b, 64, 1.0f}, n * 81, 64,
                                                512, 1, 0, s))
    return rc;
  if (S.use_shift) { // 3x3 taps = 9 row-shifted views of one staged copy of act2's rows
    if (int rc = tc::launch_conv_shift<T, 64, 3, 3, 9, 9, 7, 7, 1, (sizeof(T) == 4 ? 2 : 3), 1, false>(
            tmap(S, kMapAct2Rows), 0, tc::WKMajor<T>{W.c3w, 576}, EpiBiasRelu<T>{act3, P.c3b, 64, 1.0f}, n, s,
            S.shift_base_offset))
      return rc;
  } else if (S.use_tma) {
    if (int rc = tc::launch_gemm_tc<64, 4>(Col3Tma<T>{tmap(S, kMapAct2), 0}, RowsK<T>{W.c3w, 576},
                                           EpiBiasRelu<T>{act3, P.c3b, 64, 1.0f}, n * 49, 64, 576,
                                           1, 0, s))
      return rc;
  } else if (int rc = tc::launch_gemm_tc<64, 4>(ColK<Col3<T>>{Col3<T>{act2}}, RowsK<T>{W.c3w, 576},
                                                EpiBiasRelu<T>{act3, P.c3b, 64, 1.0f}, n * 49, 64,
                                                576, 1, 0, s))
    return rc;
  // FC: with few output tiles (small minibatch) the 3136-long reduction is split over CTAs and
  // summed in fixed order; otherwise one pass with the bias fused in the epilogue
  const int64_t tiles64 = ceil_div(n, tc::BM) * ceil_div(H, 64);
  if (tiles64 <= kNumSMs && A.fc_ws && (size_t)n * H * 2 <= A.fc_ws_floats) {
    int splits = (int)std::min<int64_t>({(int64_t)(2 * kNumSMs) / tiles64, (int64_t)(3136 / (2 * tc::BK)),
                                         (int64_t)(A.fc_ws_floats / ((size_t)n * H))});
    splits = std::max(splits, 1);
    const int rps = (int)(ceil_div(ceil_div(3136, splits), tc::BK) * tc::BK);
    splits = (int)ceil_div(3136, rps);
    if (S.fc_tma) { // both plain matrices through TMA (K-major 64-row boxes)
      if (int rc = tc::launch_gemm_tc<64, 4>(RowsKTma<T, 128>{tmap(S, kMapAct3K)},
                                             RowsKTma<T, 64>{tmap(S, kMapFcwK)},
                                             EpiPartial{A.fc_ws, (int)n, H}, n, H, 3136, splits, rps, s))
        return rc;
    } else if (int rc = tc::launch_gemm_tc<64, 4>(RowsK<T>{act3, 3136}, RowsK<T>{W.fcw, 3136},
                                                  EpiPartial{A.fc_ws, (int)n, H}, n, H, 3136, splits, rps, s))
      return rc;
    const int64_t total4 = n * H / 4;
    reduce_fc_kernel<<<(unsigned)std::min<int64_t>(ceil_div(total4, 256), 2 * kNumSMs), 256, 0, s>>>(
        reinterpret_cast<const float4 *>(A.fc_ws), splits, total4, H / 4,
        reinterpret_cast<const float4 *>(P.fcb), reinterpret_cast<float4 *>(A.fc));
    ALEGPU_COUNT_LAUNCH();
    ALEGPU_LAUNCH_CHECK();
    return ALEGPU_OK;
  }
  if (S.fc_tma)
    return tc::launch_gemm_tc<128, 3>(RowsKTma<T, 128>{tmap(S, kMapAct3K)}, RowsKTma<T, 128>{tmap(S, kMapFcwK)},
                                      EpiBiasF32{A.fc, P.fcb, H}, n, H, 3136, 1, 0, s);
  return tc::launch_gemm_tc<128, 3>(RowsK<T>{act3, 3136}, RowsK<T>{W.fcw, 3136},
                                    EpiBiasF32{A.fc, P.fcb, H}, n, H, 3136, 1, 0, s);
}

// A.dfc [n,H] fp32 is d loss / d fc-output and A.dout the head gradients (heads.cu).  Fills every
// gradient.  Two branches: the data-gradient chain fc -> conv3 -> conv2 (+ conv1's weight gradient)
// on `s`, and the other weight gradients on S.side as soon as their inputs exist.
template <class T>
int backward_impl(const NetParams &P, const NetParams &G, TcState &S, const uint8_t *obs,
                  int64_t n, NetActs &A, cudaStream_t s, const void *obs16) {
  const int H = (int)P.hidden;
  const Shadow<T> W = shadow_view<T>(S, P);
  T *act1 = (T *)S.act1, *act2 = (T *)S.act2, *act3 = (T *)S.act3;
  T *dact1 = (T *)S.dact1, *dact2 = (T *)S.dact2, *dact3 = (T *)S.dact3;
  // bf16: the copy of A.dfc written by the heads kernel; tf32: A.dfc itself (tf32-rounded there)
  const T *dfcT = sizeof(T) == 2 ? (const T *)S.dfc16 : (const T *)A.dfc;
  cudaStream_t w = S.serial ? s : S.side;
  if (!S.serial) {
    ALEGPU_CUDA_CHECK(cudaEventRecord(S.ev_fork, s));
    ALEGPU_CUDA_CHECK(cudaStreamWaitEvent(w, S.ev_fork, 0));
  }
  // ---- side: heads + FC weight gradients (need only dout / dfc / act3)
  if (int rc = heads_wgrad(G, A, n, S.ws_heads, S.ws_heads_n, w))
    return rc;
  {
    // [H][3136 | bias] = dfc^T . [act3 | 1], stored directly (no split partials, no reduction pass).
    // Up to 1024 rows 128-wide tiles (100 CTAs) are the fastest; above, 64-wide tiles give 200 CTAs each
    // walking the whole reduction (measured at 1280 rows: 2 splits of 128-wide tiles + the 16.8 us
    // reduction of 2 x 6.4 MB partials = 58.7 us in bf16; ALEGPU_FC_WGRAD_SPLIT=1 keeps that form for A/B)
    static const bool split_ab = getenv("ALEGPU_FC_WGRAD_SPLIT") != nullptr;
    if (n <= 1024) {
      if (int rc = tc::launch_gemm_tc<128, 3>(ColsMN<T>{dfcT, H}, ColsMN<T>{act3, 3136},
                                              EpiWgradDirect{G.fcw, G.fcb, 3136}, H, 3137, n, 1, 0, w))
        return rc;
    } else if (!split_ab) {
      if (int rc = tc::launch_gemm_tc<64, 4>(ColsMN<T>{dfcT, H}, ColsMN<T>{act3, 3136},
                                             EpiWgradDirect{G.fcw, G.fcb, 3136}, H, 3137, n, 1, 0, w))
        return rc;
    } else {
      const int64_t tiles = ceil_div(H, tc::BM) * ceil_div(3137, 128);
      const size_t per = (size_t)H * 3137;
      int splits, rps;
      plan_splits(n, tiles, per, S.ws_fc_n, &splits, &rps);
      if (int rc = tc::launch_gemm_tc<128, 3>(ColsMN<T>{dfcT, H}, ColsMN<T>{act3, 3136},
                                              EpiPartial{S.ws_fc, H, 3137}, H, 3137, n, splits, rps, w))
        return rc;
      const int blocks = (int)std::min<int64_t>(ceil_div((int64_t)per, 256), 4 * kNumSMs);
      reduce_wgrad_kernel<<<blocks, 256, 0, w>>>(S.ws_fc, splits, H, 3136, G.fcw, G.fcb);
      ALEGPU_COUNT_LAUNCH();
      ALEGPU_LAUNCH_CHECK();
    }
  }
  if (!S.serial) // the FC + heads gradients (95 % of the bytes) can already be exchanged
    ALEGPU_CUDA_CHECK(cudaEventRecord(S.ev_fcw, w));
  // ---- main: FC dgrad: dact3 = (dfc . fcw) * relu'(act3)
  T *dact3p = (T *)S.dact3p, *dact2p = (T *)S.dact2p;
  {
    // operands: dfc [n][H] K-major, fcw^T MN-major — through TMA (S.fc_tma) or cp.async; epilogue: relu'
    // mask into dact3, plus the zero-bordered copy when conv3's data gradient runs as a shift convolution;
    // small minibatches take 64-wide tiles (twice the CTAs)
    const bool small = ceil_div(n, tc::BM) * 25 < kNumSMs;
    auto run = [&](auto epi) -> int {
      if (S.fc_tma) {
        if (small)
          return tc::launch_gemm_tc<64, 4>(RowsKTma<T, 128>{tmap(S, kMapDfcK)},
                                           ColsMNTma<T, 64>{tmap(S, kMapFcwMN), 0}, epi, n, 3136, H, 1, 0, s);
        return tc::launch_gemm_tc<128, 3>(RowsKTma<T, 128>{tmap(S, kMapDfcK)},
                                          ColsMNTma<T, 128>{tmap(S, kMapFcwMN), 0}, epi, n, 3136, H, 1, 0, s);
      }
      if (small)
        return tc::launch_gemm_tc<64, 4>(RowsK<T>{dfcT, H}, ColsMN<T>{W.fcw, 3136}, epi, n, 3136, H, 1, 0, s);
      return tc::launch_gemm_tc<128, 3>(RowsK<T>{dfcT, H}, ColsMN<T>{W.fcw, 3136}, epi, n, 3136, H, 1, 0, s);
    };
    if (S.use_shift_dgrad) {
      if (int rc = run(EpiReluMaskFcPad<T>{dact3, dact3p, act3}))
        return rc;
    } else if (int rc = run(EpiReluMask<T>{dact3, act3, 3136}))
      return rc;
  }
  if (!S.serial) {
    ALEGPU_CUDA_CHECK(cudaEventRecord(S.ev_d3, s));
    ALEGPU_CUDA_CHECK(cudaStreamWaitEvent(w, S.ev_d3, 0));
  }
  // ---- side: conv3 wgrad (needs dact3)
  if (int rc = conv_wgrad<T, Col3<T>, 576, 64>(Col3<T>{act2}, dact3, n * 49, 1.0f, G.c3w, G.c3b,
                                               S.ws_c3, S.ws_c3_n, w, false,
                                               S.use_tma_wgrad ? tmap(S, kMapDact3MN) : nullptr))
    return rc;
  // ---- main: conv3 dgrad
  if (S.use_shift_dgrad) { // 9 flipped taps = row-shifted views of one staged copy of the bordered dact3
    if (int rc = tc::launch_conv_shift<T, 64, 3, 3, 10, 10, 9, 9, 1, (sizeof(T) == 4 ? 2 : 3), 1, true>(
            tmap(S, kMapDact3PadRows), 0, tc::WMNView<W3T<T>>{W3T<T>{W.c3w}},
            EpiReluMaskPad<T>{dact2, dact2p, act2}, n, s, S.shift_base_offset))
      return rc;
  } else if (S.use_tma && S.use_tma_weights) {
    if (int rc = tc::launch_gemm_tc<64, 4>(DCol3Tma<T>{tmap(S, kMapDact3), 0},
                                           WTapTma<T, (sizeof(T) == 2 ? 1 : 2), false>{tmap(S, kMapC3wTap)},
                                           EpiReluMask<T>{dact2, act2, 64}, n * 81, 64, 576, 1, 0, s))
      return rc;
  } else if (S.use_tma) {
    if (int rc = tc::launch_gemm_tc<64, 4>(DCol3Tma<T>{tmap(S, kMapDact3), 0}, W3T<T>{W.c3w},
                                           EpiReluMask<T>{dact2, act2, 64}, n * 81, 64, 576, 1, 0, s))
      return rc;
  } else if (int rc = tc::launch_gemm_tc<64, 4>(DCol3<T>{dact3}, W3T<T>{W.c3w},
                                                EpiReluMask<T>{dact2, act2, 64}, n * 81, 64, 576, 1,
                                                0, s))
    return rc;
  if (!S.serial) {
    ALEGPU_CUDA_CHECK(cudaEventRecord(S.ev_d2, s));
    ALEGPU_CUDA_CHECK(cudaStreamWaitEvent(w, S.ev_d2, 0));
  }
  // ---- side: conv2 wgrad (needs dact2)
  if (int rc = conv_wgrad<T, Col2<T>, 512, 64>(Col2<T>{act1}, dact2, n * 81, 1.0f, G.c2w, G.c2b,
                                               S.ws_c2, S.ws_c2_n, w, false,
                                               S.use_tma_wgrad ? tmap(S, kMapDact2MN) : nullptr))
    return rc;
  if (!S.serial)
    ALEGPU_CUDA_CHECK(cudaEventRecord(S.ev_join, w));
  // ---- main: conv2 dgrad (the 4 parity classes of the stride-2 transposed convolution as 128 output
  // columns of ONE GEMM over the TMA-fetched 2x2-tap walk of dact2; cp.async fallback: 4 z-slices)
  if (S.use_shift_dgrad) { // 2x2 flipped taps over the bordered dact2, all four parity classes as 128 columns
    if (int rc = tc::launch_conv_shift<T, 128, 2, 2, 10, 10, 10, 10, 1, (sizeof(T) == 4 ? 2 : 3), 1, true>(
            tmap(S, kMapDact2PadRows), 0, tc::WMNView<W2TAll<T>>{W2TAll<T>{W.c2w}},
            EpiReluMaskConv2All<T>{dact1, act1}, n, s, S.shift_base_offset))
      return rc;
  } else if (S.use_tma && S.use_tma_weights) {
    if (int rc = tc::launch_gemm_tc<128, 3>(DCol2Tma<T>{tmap(S, kMapDact2), 0},
                                            WTapTma<T, (sizeof(T) == 2 ? 2 : 4), true>{tmap(S, kMapC2wTap)},
                                            EpiReluMaskConv2All<T>{dact1, act1}, n * 100, 128, 256, 1, 0,
                                            s))
      return rc;
  } else if (S.use_tma) {
    if (int rc = tc::launch_gemm_tc<128, 3>(DCol2Tma<T>{tmap(S, kMapDact2), 0}, W2TAll<T>{W.c2w},
                                            EpiReluMaskConv2All<T>{dact1, act1}, n * 100, 128, 256, 1, 0,
                                            s))
      return rc;
  } else if (int rc = tc::launch_gemm_tc<64, 4>(DCol2<T>{dact2}, W2T<T>{W.c2w},
                                                EpiReluMaskConv2<T>{dact1, act1}, n * 100, 32, 256, 4,
                                                0, s))
    return rc;
  // conv1 (input needs no gradient); the 1/255 of the input scale lands on the weight gradient.
  // On the space-to-depth copy the observation operand is read once, un-expanded, and the four
  // taps become shifted views of dact1 (DYShift); ALEGPU_CONV1_WGRAD_U8=1 keeps the im2col form
  // over the u8 observations (register producers, bf16 mode only) for A/B.
  const T *x16 = (const T *)(obs16 ? obs16 : S.obs16_mb);
  bool done = false;
  if constexpr (sizeof(T) == 2) {
    if (S.conv1_u8 || S.conv1_wgrad_u8 || !x16) {
      if (int rc = conv_wgrad<T, Col1, 256, 32>(Col1{obs}, dact1, n * 400, 1.0f / 255.0f, G.c1w,
                                                G.c1b, S.ws_c1, S.ws_c1_n, s))
        return rc;
      done = true;
    }
  }
  if (!done) { // the forward of these rows left their operand copy in obs16_mb unless the caller has one
    // the bias column sums run next to the GEMM on the side stream once dact1 exists — when they are
    // long enough to pay for the fork/join edges (measured: 7.72 -> 7.51 ms per v1 iteration at 1280
    // rows, but 2.51 -> 2.59 ms at v0's 256 rows, where they stay in front of the GEMM)
    const bool fork = !S.serial && n > 512;
    if (fork) {
      ALEGPU_CUDA_CHECK(cudaEventRecord(S.ev_d1, s));
      ALEGPU_CUDA_CHECK(cudaStreamWaitEvent(w, S.ev_d1, 0));
    }
    const T *all = (const T *)S.obs16_all;
    const bool in_all = all && x16 >= all && x16 < all + (size_t)S.obs16_all_rows * 28224;
    const void *xmap = nullptr;
    int xbase = 0;
    if (S.use_tma_wgrad && (in_all || x16 == (const T *)S.obs16_mb)) {
      xmap = tmap(S, in_all ? kMapObsAllMN + S.obs_all_map : kMapObsMbMN);
      xbase = in_all ? (int)((x16 - all) / 28224) * 441 : 0;
    }
    if (int rc = conv1_wgrad_s2d<T>(x16, dact1, n, 1.0f / 255.0f, G.c1w, G.c1b, S.ws_c1, S.ws_c1_n,
                                    s, fork ? w : s, fork ? S.ev_join : nullptr, xmap, xbase,
                                    S.use_tma_wgrad && S.elem_bytes == 4 && !getenv("ALEGPU_NO_TMA_DYSHIFT")
                                        ? tmap(S, kMapDact1Shift)
                                        : nullptr))
      return rc;
  }
  if (!S.serial)
    ALEGPU_CUDA_CHECK(cudaStreamWaitEvent(s, S.ev_join, 0));
  return ALEGPU_OK;
}
} // namespace

int net_tc_forward(const NetParams &P, TcState &S, const uint8_t *obs, int64_t n, NetActs &A,
                   cudaStream_t s, const void *obs16) {
  if (S.elem_bytes == 4)
    return forward_impl<float>(P, S, obs, n, A, s, obs16);
  return forward_impl<bf16>(P, S, obs, n, A, s, obs16);
}

int net_tc_backward(const NetParams &P, const NetParams &G, TcState &S, const uint8_t *obs,
                    int64_t n, NetActs &A, cudaStream_t s, const void *obs16) {
  if (S.elem_bytes == 4)
    return backward_impl<float>(P, G, S, obs, n, A, s, obs16);
  return backward_impl<bf16>(P, G, S, obs, n, A, s, obs16);
}

int net_tc_fetch_activation(TcState &T, int which, int64_t n, float *host, cudaStream_t s) {
  const int64_t per = which == 1 ? 12800 : which == 2 ? 5184 : 3136;
  const void *src = which == 1 ? T.act1 : which == 2 ? T.act2 : T.act3;
  const int64_t total = n * per;
  if (T.elem_bytes == 4) {
    ALEGPU_CUDA_CHECK(cudaMemcpyAsync(host, src, sizeof(float) * total, cudaMemcpyDeviceToHost, s));
    ALEGPU_CUDA_CHECK(cudaStreamSynchronize(s));
    return ALEGPU_OK;
  }
  if (!T.stage)
    ALEGPU_CUDA_CHECK(cudaMalloc(&T.stage, sizeof(float) * (size_t)T.cap * 12800));
  bf16_to_f32_kernel<<<kNumSMs * 2, 256, 0, s>>>((const bf16 *)src, T.stage, total);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  ALEGPU_CUDA_CHECK(cudaMemcpyAsync(host, T.stage, sizeof(float) * total, cudaMemcpyDeviceToHost, s));
  ALEGPU_CUDA_CHECK(cudaStreamSynchronize(s));
  return ALEGPU_OK;
}

} // namespace alegpu

// ---- engine self-test (C ABI, include/alegpu.h) ---------------------------------------------------
// C[M][N] (fp32) = A . B^T through the tcgen05 engine with fp32 inputs rounded to bf16.
// a_mn == 0: A is [M][R] row-major (K-major operand); a_mn == 1: A is stored [R][M] (MN-major).
// Same for B with N.  bn selects the tile width instantiation (32, 64, 128 or 256).
// async == 0: fp32 loaders (register path: load -> convert -> swizzled st.shared);
// async == 1: the inputs are first rounded to bf16 buffers and read by the cp.async producers.
namespace alegpu {
namespace {
template <class AOpK, class AOpMN, class BOpK, class BOpMN>
int selftest_dispatch(const AOpK &ak, const AOpMN &am, const BOpK &bk, const BOpMN &bm, float *C,
                      int64_t M, int64_t N, int64_t R, int a_mn, int b_mn, int bn, int z, int rps,
                      cudaStream_t s) {
  EpiPartial e{C, (int)M, (int)N}; // splits > 1: C must hold z*M*N floats (partials, not summed)
#define ALEGPU_ST(BN, ST)                                                                          \
  do {                                                                                             \
    if (!a_mn && !b_mn)                                                                            \
      return tc::launch_gemm_tc<BN, ST>(ak, bk, e, M, N, R, z, rps, s);                            \
    if (a_mn && !b_mn)                                                                             \
      return tc::launch_gemm_tc<BN, ST>(am, bk, e, M, N, R, z, rps, s);                            \
    if constexpr (BN >= 64) {                                                                      \
      if (!a_mn && b_mn)                                                                           \
        return tc::launch_gemm_tc<BN, ST>(ak, bm, e, M, N, R, z, rps, s);                          \
      return tc::launch_gemm_tc<BN, ST>(am, bm, e, M, N, R, z, rps, s);                            \
    }                                                                                              \
  } while (0)
  if (bn == 32)
    ALEGPU_ST(32, 4);
  else if (bn == 64)
    ALEGPU_ST(64, 4);
  else if (bn == 128)
    ALEGPU_ST(128, 3);
  else if (bn == 256)
    ALEGPU_ST(256, 3);
#undef ALEGPU_ST
  return fail(ALEGPU_INVALID_ARGUMENT, "selftest_tc_gemm: unsupported bn / major combination");
}
} // namespace
} // namespace alegpu

static int selftest_impl(const float *A, const float *B, float *C, int64_t M, int64_t N, int64_t R,
                         int a_mn, int b_mn, int bn, int splits, int async, void *stream) {
  using namespace alegpu;
  ALEGPU_REQUIRE(A && B && C && M > 0 && N > 0 && R > 0, ALEGPU_INVALID_ARGUMENT,
                 "selftest_tc_gemm: bad argument");
  ALEGPU_REQUIRE(splits >= 1, ALEGPU_INVALID_ARGUMENT, "selftest_tc_gemm: splits >= 1");
  ALEGPU_REQUIRE((a_mn ? M : R) % 8 == 0 && (b_mn ? N : R) % 8 == 0, ALEGPU_INVALID_ARGUMENT,
                 "selftest_tc_gemm: the contiguous dimension of each operand must be a multiple of 8");
  cudaStream_t s = as_stream(stream);
  int rps = 0;
  if (splits > 1)
    rps = (int)(ceil_div(ceil_div(R, splits), tc::BK) * tc::BK);
  const int z = splits > 1 ? (int)ceil_div(R, rps) : 1;
  if (!async)
    return selftest_dispatch(RowsKF32{A, (int)R}, ColsMNF32{A, (int)M}, RowsKF32{B, (int)R},
                             ColsMNF32{B, (int)N}, C, M, N, R, a_mn, b_mn, bn, z, rps, s);
  bf16 *a16 = nullptr, *b16 = nullptr;
  ALEGPU_CUDA_CHECK(cudaMalloc(&a16, sizeof(bf16) * (size_t)M * R));
  ALEGPU_CUDA_CHECK(cudaMalloc(&b16, sizeof(bf16) * (size_t)N * R));
  int rc = launch_f32_to_bf16(A, a16, M * R, s);
  if (!rc)
    rc = launch_f32_to_bf16(B, b16, N * R, s);
  if (!rc) // ColsMN's ones row (i0 >= ld) is never requested here: rows < M == ld
    rc = selftest_dispatch(RowsK<bf16>{a16, (int)R}, ColsMN<bf16>{a16, (int)M}, RowsK<bf16>{b16, (int)R},
                           ColsMN<bf16>{b16, (int)N}, C, M, N, R, a_mn, b_mn, bn, z, rps, s);
  cudaStreamSynchronize(s);
  cudaFree(a16);
  cudaFree(b16);
  return rc;
}

// kind::tf32: the fp32 inputs are read as they are by the cp.async producers (the tensor core uses the
// upper 19 bits of every element); contiguous dimensions must be multiples of 4
extern "C" int alegpu_selftest_tc_gemm_tf32(const float *A, const float *B, float *C, int64_t M,
                                            int64_t N, int64_t R, int a_mn, int b_mn, int bn,
                                            int splits, void *stream) {
  using namespace alegpu;
  ALEGPU_REQUIRE(A && B && C && M > 0 && N > 0 && R > 0, ALEGPU_INVALID_ARGUMENT,
                 "selftest_tc_gemm_tf32: bad argument");
  ALEGPU_REQUIRE(splits >= 1, ALEGPU_INVALID_ARGUMENT, "selftest_tc_gemm_tf32: splits >= 1");
  ALEGPU_REQUIRE((a_mn ? M : R) % 4 == 0 && (b_mn ? N : R) % 4 == 0, ALEGPU_INVALID_ARGUMENT,
                 "selftest_tc_gemm_tf32: the contiguous dimension of each operand must be a multiple of 4");
  cudaStream_t s = as_stream(stream);
  int rps = 0;
  if (splits > 1)
    rps = (int)(ceil_div(ceil_div(R, splits), tc::BK) * tc::BK);
  const int z = splits > 1 ? (int)ceil_div(R, rps) : 1;
  return selftest_dispatch(RowsK<float>{A, (int)R}, ColsMN<float>{A, (int)M}, RowsK<float>{B, (int)R},
                           ColsMN<float>{B, (int)N}, C, M, N, R, a_mn, b_mn, bn, z, rps, s);
}

extern "C" int alegpu_selftest_tc_gemm(const float *A, const float *B, float *C, int64_t M, int64_t N,
                                       int64_t R, int a_mn, int b_mn, int bn, int splits,
                                       void *stream) {
  return selftest_impl(A, B, C, M, N, R, a_mn, b_mn, bn, splits, 0, stream);
}
extern "C" int alegpu_selftest_tc_gemm_async(const float *A, const float *B, float *C, int64_t M,
                                             int64_t N, int64_t R, int a_mn, int b_mn, int bn,
                                             int splits, void *stream) {
  return selftest_impl(A, B, C, M, N, R, a_mn, b_mn, bn, splits, 1, stream);
}
