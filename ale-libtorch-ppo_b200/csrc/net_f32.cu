// net_f32.cu — fp32 CUDA-core implicit-GEMM Nature-CNN (the fp32-faithful parity mode).
//
// Replaces NetworkImpl::forward (src/bin/train.cc:255-265) and its autograd backward
// (src/ai/ppo/train.h:126) — cuDNN/cuBLAS calls in the reference — with one generic
// register-blocked shared-memory GEMM  C[i][j] = sum_r A(i,r) * B(j,r)  whose operands are
// gathered on the fly (im2col / transposed-conv / weight-gradient views are functors, never
// materialised) and whose epilogue fuses bias+ReLU (forward), the ReLU mask of the producing
// layer (dgrad) or the split-R partial store (wgrad; reduced in fixed order => deterministic).
// The u8 -> f32 /255 cast (train.cc:258-259) is fused into conv1's operand load.
//
// Internal layouts (an implementation detail behind alegpu_learner_{set,get}_params):
//   activations NHWC: act1 [n,20,20,32], act2 [n,9,9,64], act3 [n,7,7,64] (= fc input, k=(y,x,c))
//   weights  c1w [32][c,ky,kx]  c2w [64][ky,kx,ci]  c3w [64][ky,kx,ci]  fcw [H][(y,x),c]
#include "net.h"

namespace alegpu {
namespace {

constexpr int BK = 16;
constexpr int kThreads = 256;

template <int BM, int BN, typename AOp, typename BOp, typename Epi>
__global__ void __launch_bounds__(kThreads)
    gemm_f32_kernel(AOp aop, BOp bop, Epi epi, int M, int N, int R, int r_per_split) {
  constexpr int TM = BM / 16, TN = BN / 16;
  constexpr int LA = BM * BK / kThreads, LB = BN * BK / kThreads;
  __shared__ __align__(16) float As[BK][BM + 4];
  __shared__ __align__(16) float Bs[BK][BN + 4];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int i0 = blockIdx.y * BM, j0 = blockIdx.x * BN, z = blockIdx.z;
  // r_per_split > 0: z indexes a split of the reduction; == 0: z is an operand class
  const int r_begin = r_per_split > 0 ? z * r_per_split : 0;
  const int r_end = r_per_split > 0 ? min(R, r_begin + r_per_split) : R;
  float acc[TM][TN];
#pragma unroll
  for (int a = 0; a < TM; ++a)
#pragma unroll
    for (int b = 0; b < TN; ++b)
      acc[a][b] = 0.0f;
  float ra[LA], rb[LB];
  auto fetch = [&](int r0) {
#pragma unroll
    for (int q = 0; q < LA; ++q) {
      const int idx = tid + q * kThreads;
      const int il = AOp::kContigR ? idx / BK : idx % BM;
      const int rl = AOp::kContigR ? idx % BK : idx / BM;
      const int i = i0 + il, r = r0 + rl;
      ra[q] = (i < M && r < r_end) ? aop(i, r, z) : 0.0f;
    }
#pragma unroll
    for (int q = 0; q < LB; ++q) {
      const int idx = tid + q * kThreads;
      const int jl = BOp::kContigR ? idx / BK : idx % BN;
      const int rl = BOp::kContigR ? idx % BK : idx / BN;
      const int j = j0 + jl, r = r0 + rl;
      rb[q] = (j < N && r < r_end) ? bop(j, r, z) : 0.0f;
    }
  };
  fetch(r_begin);
  for (int r0 = r_begin; r0 < r_end; r0 += BK) {
#pragma unroll
    for (int q = 0; q < LA; ++q) {
      const int idx = tid + q * kThreads;
      const int il = AOp::kContigR ? idx / BK : idx % BM;
      const int rl = AOp::kContigR ? idx % BK : idx / BM;
      As[rl][il] = ra[q];
    }
#pragma unroll
    for (int q = 0; q < LB; ++q) {
      const int idx = tid + q * kThreads;
      const int jl = BOp::kContigR ? idx / BK : idx % BN;
      const int rl = BOp::kContigR ? idx % BK : idx / BN;
      Bs[rl][jl] = rb[q];
    }
    __syncthreads();
    if (r0 + BK < r_end)
      fetch(r0 + BK);
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      float av[TM], bv[TN];
#pragma unroll
      for (int a = 0; a < TM; ++a)
        av[a] = As[k][ty * TM + a];
#pragma unroll
      for (int b = 0; b < TN; ++b)
        bv[b] = Bs[k][tx * TN + b];
#pragma unroll
      for (int a = 0; a < TM; ++a)
#pragma unroll
        for (int b = 0; b < TN; ++b)
          acc[a][b] = fmaf(av[a], bv[b], acc[a][b]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int a = 0; a < TM; ++a)
#pragma unroll
    for (int b = 0; b < TN; ++b) {
      const int i = i0 + ty * TM + a, j = j0 + tx * TN + b;
      if (i < M && j < N)
        epi(i, j, acc[a][b], z);
    }
}

template <int BM, int BN, typename AOp, typename BOp, typename Epi>
int launch_gemm(AOp a, BOp b, Epi e, int64_t M, int64_t N, int64_t R, int zdim, int r_per_split,
                cudaStream_t s) {
  if (M == 0 || N == 0)
    return ALEGPU_OK;
  dim3 grid((unsigned)ceil_div(N, BN), (unsigned)ceil_div(M, BM), (unsigned)zdim);
  gemm_f32_kernel<BM, BN><<<grid, kThreads, 0, s>>>(a, b, e, (int)M, (int)N, (int)R, r_per_split);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}

// ---- operand views ------------------------------------------------------------------------
// im2col of the u8 NCHW observation for conv1 (k = c*64 + ky*8 + kx), scaled by 1/255
struct Col1 {
  static constexpr bool kContigR = true;
  const uint8_t *obs;
  __device__ float operator()(int m, int k, int) const {
    const int n = m / 400, pix = m - n * 400, oy = pix / 20, ox = pix - oy * 20;
    const int c = k >> 6, ky = (k >> 3) & 7, kx = k & 7;
    const uint8_t v = obs[(size_t)n * 28224 + c * 7056 + (oy * 4 + ky) * 84 + ox * 4 + kx];
    return __fdiv_rn((float)v, 255.0f); // train.cc:258-259
  }
};
// im2col of NHWC act1 for conv2 (k = (ky*4+kx)*32 + ci)
struct Col2 {
  static constexpr bool kContigR = true;
  const float *x;
  __device__ float operator()(int m, int k, int) const {
    const int n = m / 81, pix = m - n * 81, oy = pix / 9, ox = pix - oy * 9;
    const int ky = k >> 7, kx = (k >> 5) & 3, ci = k & 31;
    return x[((size_t)(n * 20 + 2 * oy + ky) * 20 + 2 * ox + kx) * 32 + ci];
  }
};
// im2col of NHWC act2 for conv3 (k = (ky*3+kx)*64 + ci)
struct Col3 {
  static constexpr bool kContigR = true;
  const float *x;
  __device__ float operator()(int m, int k, int) const {
    const int n = m / 49, pix = m - n * 49, oy = pix / 7, ox = pix - oy * 7;
    const int t = k >> 6, ky = t / 3, kx = t - ky * 3, ci = k & 63;
    return x[((size_t)(n * 9 + oy + ky) * 9 + ox + kx) * 64 + ci];
  }
};
// plain row-major matrix [rows][ld], element (row, r)
struct RowMajor {
  static constexpr bool kContigR = true;
  const float *p;
  int ld;
  __device__ float operator()(int i, int r, int) const { return p[(size_t)i * ld + r]; }
};
// transposed view: element (i, r) = p[r*ld + i]  (contiguous along i)
struct ColMajor {
  static constexpr bool kContigR = false;
  const float *p;
  int ld;
  __device__ float operator()(int i, int r, int) const { return p[(size_t)r * ld + i]; }
};
// wgrad "B" views: (k, m) -> col(m, k) with an extra all-ones column k == K for the bias grad
template <typename Col, int K> struct ColT {
  static constexpr bool kContigR = false;
  Col col;
  __device__ float operator()(int k, int m, int) const { return k < K ? col(m, k, 0) : 1.0f; }
};
struct RowMajorOnes { // (k, m) -> x[m*ld + k], k == ld -> 1
  static constexpr bool kContigR = false;
  const float *p;
  int ld;
  __device__ float operator()(int k, int m, int) const {
    return k < ld ? p[(size_t)m * ld + k] : 1.0f;
  }
};
// conv3 dgrad A view: rows = act2 pixels (n,iy,ix), r = (ky,kx,co)
struct DCol3 {
  static constexpr bool kContigR = true;
  const float *dy; // dact3 NHWC [n,7,7,64]
  __device__ float operator()(int i, int r, int) const {
    const int n = i / 81, pix = i - n * 81, iy = pix / 9, ix = pix - iy * 9;
    const int t = r >> 6, ky = t / 3, kx = t - ky * 3, co = r & 63;
    const int oy = iy - ky, ox = ix - kx;
    if (oy < 0 || oy >= 7 || ox < 0 || ox >= 7)
      return 0.0f;
    return dy[((size_t)(n * 7 + oy) * 7 + ox) * 64 + co];
  }
};
struct W3T { // (ci, r=(ky,kx,co)) -> c3w[co][(ky,kx),ci]
  static constexpr bool kContigR = false;
  const float *w;
  __device__ float operator()(int ci, int r, int) const {
    const int t = r >> 6, co = r & 63;
    return w[(size_t)co * 576 + t * 64 + ci];
  }
};
// conv2 dgrad, one parity class z = py*2+px of the 20x20 input: rows (n,qy,qx), r = (ty,tx,co)
struct DCol2 {
  static constexpr bool kContigR = true;
  const float *dy; // dact2 NHWC [n,9,9,64]
  __device__ float operator()(int i, int r, int) const {
    const int n = i / 100, pix = i - n * 100, qy = pix / 10, qx = pix - qy * 10;
    const int t = r >> 6, ty = t >> 1, tx = t & 1, co = r & 63;
    const int oy = qy - ty, ox = qx - tx;
    if (oy < 0 || oy >= 9 || ox < 0 || ox >= 9)
      return 0.0f;
    return dy[((size_t)(n * 9 + oy) * 9 + ox) * 64 + co];
  }
};
struct W2T { // (ci, r=(ty,tx,co), z) -> c2w[co][(ky,kx),ci], ky = py+2ty, kx = px+2tx
  static constexpr bool kContigR = false;
  const float *w;
  __device__ float operator()(int ci, int r, int z) const {
    const int t = r >> 6, ty = t >> 1, tx = t & 1, co = r & 63;
    const int ky = (z >> 1) + 2 * ty, kx = (z & 1) + 2 * tx;
    return w[(size_t)co * 512 + (ky * 4 + kx) * 32 + ci];
  }
};

// ---- epilogues ------------------------------------------------------------------------------
struct EpiBiasAct { // out[i*ld + j] = act(acc + bias[j])
  float *out;
  const float *bias;
  int ld;
  bool relu;
  __device__ void operator()(int i, int j, float acc, int) const {
    const float v = acc + bias[j];
    out[(size_t)i * ld + j] = (relu && v < 0.0f) ? 0.0f : v;
  }
};
struct EpiReluMask { // out[i*ld + j] = y[i*ld + j] > 0 ? acc : 0
  float *out;
  const float *y;
  int ld;
  __device__ void operator()(int i, int j, float acc, int) const {
    const size_t o = (size_t)i * ld + j;
    out[o] = y[o] > 0.0f ? acc : 0.0f;
  }
};
struct EpiReluMaskConv2 { // scatter parity class z back into NHWC [n,20,20,32]
  float *out;
  const float *y;
  __device__ void operator()(int i, int j, float acc, int z) const {
    const int n = i / 100, pix = i - n * 100, qy = pix / 10, qx = pix - qy * 10;
    const size_t o = ((size_t)(n * 20 + 2 * qy + (z >> 1)) * 20 + 2 * qx + (z & 1)) * 32 + j;
    out[o] = y[o] > 0.0f ? acc : 0.0f;
  }
};
struct EpiPartial { // ws[z][i][j], row length ldn
  float *ws;
  int rows, ldn;
  __device__ void operator()(int i, int j, float acc, int z) const {
    ws[((size_t)z * rows + i) * ldn + j] = acc;
  }
};

// gw[i*K + j] (j < K) / gb[i] (j == K) = sum_z ws[z][i][j], z ascending => deterministic
__global__ void reduce_splits_kernel(const float *__restrict__ ws, int splits, int rows, int K,
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

// weight gradient of one layer: [rows][K+1] = sum_m dY[m][rows] * colT(k, m), split over m
template <int BM, typename BOp>
int wgrad(const float *dy, int rows, BOp colT, int K, int64_t Mtot, float *gw, float *gb,
          float *ws, size_t ws_floats, cudaStream_t s) {
  const int64_t tiles = ceil_div(rows, BM) * ceil_div(K + 1, 64);
  int splits = (int)std::max<int64_t>(1, std::min<int64_t>((4 * kNumSMs) / tiles, Mtot / 128));
  const size_t per = (size_t)rows * (K + 1);
  splits = (int)std::min<size_t>(splits, ws_floats / per);
  if (splits < 1)
    return fail(ALEGPU_RUNTIME_ERROR, "wgrad: workspace too small");
  int rps = (int)ceil_div(Mtot, splits);
  rps = (int)(ceil_div(rps, BK) * BK);
  splits = (int)ceil_div(Mtot, rps);
  ColMajor a{dy, rows};
  EpiPartial e{ws, rows, K + 1};
  if (int rc = launch_gemm<BM, 64>(a, colT, e, rows, K + 1, Mtot, splits, rps, s))
    return rc;
  const int64_t total = (int64_t)per;
  reduce_splits_kernel<<<(unsigned)std::min<int64_t>(ceil_div(total, 256), 4 * kNumSMs), 256, 0,
                         s>>>(ws, splits, rows, K, gw, gb);
  ALEGPU_COUNT_LAUNCH();
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}

} // namespace

// ---- public (net.h) ---------------------------------------------------------------------------
int net_f32_forward(const NetParams &P, const uint8_t *obs, int64_t n, NetActs &A,
                    cudaStream_t s) {
  const int H = (int)P.hidden;
  if (int rc = launch_gemm<64, 32>(Col1{obs}, RowMajor{P.c1w, 256},
                                   EpiBiasAct{A.act1, P.c1b, 32, true}, n * 400, 32, 256, 1, 256, s))
    return rc;
  if (int rc = launch_gemm<64, 64>(Col2{A.act1}, RowMajor{P.c2w, 512},
                                   EpiBiasAct{A.act2, P.c2b, 64, true}, n * 81, 64, 512, 1, 512, s))
    return rc;
  if (int rc = launch_gemm<64, 64>(Col3{A.act2}, RowMajor{P.c3w, 576},
                                   EpiBiasAct{A.act3, P.c3b, 64, true}, n * 49, 64, 576, 1, 576, s))
    return rc;
  return launch_gemm<64, 64>(RowMajor{A.act3, 3136}, RowMajor{P.fcw, 3136},
                             EpiBiasAct{A.fc, P.fcb, H, false}, n, H, 3136, 1, 3136, s);
}

// dfc [n,H] is d loss / d fc-output (from the heads kernel). Fills every gradient except the
// heads' (heads.cu does those).
int net_f32_backward(const NetParams &P, const NetParams &G, const uint8_t *obs, int64_t n,
                     NetActs &A, float *ws, size_t ws_floats, cudaStream_t s) {
  const int H = (int)P.hidden;
  // FC: dact3 = (dfc . fcw) * relu'(act3) ; dfcw = dfc^T . act3 ; dfcb = colsum(dfc)
  if (int rc = launch_gemm<64, 64>(RowMajor{A.dfc, H}, ColMajor{P.fcw, 3136},
                                   EpiReluMask{A.dact3, A.act3, 3136}, n, 3136, H, 1, H, s))
    return rc;
  if (int rc = wgrad<64>(A.dfc, H, RowMajorOnes{A.act3, 3136}, 3136, n, G.fcw, G.fcb, ws,
                         ws_floats, s))
    return rc;
  // conv3
  if (int rc = launch_gemm<64, 64>(DCol3{A.dact3}, W3T{P.c3w}, EpiReluMask{A.dact2, A.act2, 64},
                                   n * 81, 64, 576, 1, 576, s))
    return rc;
  if (int rc = wgrad<64>(A.dact3, 64, ColT<Col3, 576>{Col3{A.act2}}, 576, n * 49, G.c3w, G.c3b,
                         ws, ws_floats, s))
    return rc;
  // conv2 (4 parity classes of the stride-2 transposed convolution in one launch)
  if (int rc = launch_gemm<64, 32>(DCol2{A.dact2}, W2T{P.c2w}, EpiReluMaskConv2{A.dact1, A.act1},
                                   n * 100, 32, 256, 4, 0, s))
    return rc;
  if (int rc = wgrad<64>(A.dact2, 64, ColT<Col2, 512>{Col2{A.act1}}, 512, n * 81, G.c2w, G.c2b,
                         ws, ws_floats, s))
    return rc;
  // conv1 (input needs no gradient)
  return wgrad<32>(A.dact1, 32, ColT<Col1, 256>{Col1{obs}}, 256, n * 400, G.c1w, G.c1b, ws,
                   ws_floats, s);
}

} // namespace alegpu
