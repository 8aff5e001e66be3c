// tc_shift.cuh — stride-1 convolutions as "shift GEMMs": the input is staged ONCE per output tile and
// every filter tap is a shifted view of that one staged copy.
//
// The im2col engine (tc_gemm.cuh) feeds the tensor core one [128 pixels x 64 channels] box per tap, so
// a KHxKW convolution pulls every input pixel KH*KW times L2 -> SM (conv3: 9x, conv1 on the
// space-to-depth observations: 4x) and re-fetches the weights for every tile; the conv layers of this
// network sit on that L2 -> SM path, not on the tensor core.  Here the output is
// computed on the INPUT's pixel grid: rows r = image*G + gy*GW + gx run over all G = GH*GW input
// pixels of every image, and
//       out(r)[co] = sum_{tap (ky,kx)} sum_c X[r + ky*GW + kx][c] * W[co][(ky*KW+kx)*64 + c]
// which equals the convolution wherever (gy, gx) is a valid output position (gy < OH, gx < OW); the
// other rows are computed and dropped (conv3: 81 grid rows per 49 outputs, conv1: 441 per 400; conv2,
// stride 2, runs on act1's pixel-PAIR grid [20 x 10] where a 4x4 stride-2 filter is a 4x2 stride-(2,1)
// one: outputs sit on the even grid rows, 200 grid rows per 81 outputs — tensor work the kernel has to
// spare).  So the A operand of EVERY tap is the same 128 + (KH-1)*GW + KW-1 rows
// of the plain [rows x 64] input matrix: one tiled TMA load per 128-byte channel panel stages them in the
// 128-byte-swizzled K-major layout, and the UMMA shared-memory descriptor of tap (ky,kx) simply starts
// (ky*GW + kx) rows = (ky*GW + kx)*128 bytes further in.  The weights (<= 147 KB) are loaded once per
// CTA and stay resident; a CTA walks its tiles with a double-buffered TMEM accumulator.  Per tile the SM
// ingests 19 KB (bf16) / 38 KB (tf32) instead of 147 / 295 KB (+ 74 / 147 KB of weights) for conv3.
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = MMA issuer + TMEM owner, warps 2-5 =
// epilogue (TMEM lane quarter = warp & 3).
#pragma once
#include "tc_gemm.cuh"

namespace alegpu {
namespace tc {

__device__ __forceinline__ void tma_tile_2d(uint32_t dst, const void *map, int c0, int c1,
                                            uint64_t *bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
               " [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
               "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void tma_tile_3d(uint32_t dst, const void *map, int c0, int c1, int c2,
                                            uint64_t *bar) {
  asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
               " [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(dst),
               "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}
// the producer's single arrival on a stage's full barrier, carrying the bytes its TMA loads will deliver
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
// K-major SWIZZLE_128B descriptor whose start address is any 128-byte row of a 1024-byte aligned panel.
// Measured on B200: the swizzle XOR is a function of the absolute shared-memory address (as TMA's is),
// so the shifted start needs NO matrix-base-offset (bits 49-51 stay 0); use_base_offset puts
// (start >> 7) & 7 there and gives wrong products — kept only as the experiment that showed it
__device__ __forceinline__ uint64_t make_smem_desc_row(uint32_t smem_addr, bool use_base_offset) {
  uint64_t d = make_smem_desc(smem_addr, 16u, 1024u);
  if (use_base_offset)
    d |= (uint64_t)((smem_addr >> 7) & 7u) << 49;
  return d;
}

constexpr int kShiftThreads = 192;
// rows one staged copy holds: 128 + the filter's largest tap shift, rounded up to 8 (the TMA box height)
__host__ __device__ constexpr int shift_stage_rows(int kh, int kw, int gw) {
  return ((128 + (kh - 1) * gw + kw - 1 + 7) / 8) * 8;
}

template <class T, int BN, int KH, int KW, int GW> struct ShiftGeom {
  static constexpr int kEB = (int)sizeof(T);
  static constexpr int kBKE = 128 / kEB;       // elements of one 128-byte panel row
  static constexpr int kCP = 64 / kBKE;        // panels per 64-channel pixel (bf16 1, tf32 2)
  static constexpr int kTaps = KH * KW;
  static constexpr int kK = kTaps * 64;        // reduction length
  static constexpr int kNKB = kTaps * kCP;     // 128-byte k-blocks
  static constexpr int kRows = shift_stage_rows(KH, KW, GW);
  static_assert(kRows <= 256, "TMA box height");
  static constexpr int kPanel = kRows * 128;
  static constexpr int kStage = kCP * kPanel;
  static constexpr int kBPanel = BN * 128;
  static constexpr int kBBytes = kNKB * kBPanel;
  static constexpr int smem(int stages) { return kBBytes + stages * kStage + 1024; }
};

// Where the resident weights come from.  WKMajor: a plain [BN][K] matrix, reduction index contiguous
// (forward layers).  WMNView<Op>: one of the transposed-weight views of net_tc.cu (row(j0) / col(k, 0) /
// at(rc, cc): 16-byte chunks contiguous along the OUTPUT index j), stored MN-major (data gradients).
template <class T> struct WKMajor {
  static constexpr bool kMN = false;
  const T *w;
  int K;
  __device__ __forceinline__ const void *chunk(int j, int kcg) const { // row j, 16-byte chunk kcg of its K
    return w + (size_t)j * K + kcg * (16 / (int)sizeof(T));
  }
  __device__ __forceinline__ const void *anchor() const { return w; }
};
template <class Op> struct WMNView {
  static constexpr bool kMN = true;
  Op op;
  __device__ __forceinline__ const void *chunk(int c, int k) const { // rows c*EPC.., reduction index k
    return op.at(op.row(c * (16 / Op::kElemBytes)), op.col(k, 0));
  }
  __device__ __forceinline__ const void *anchor() const { return op.anchor(); }
};

// Epi(i_out, j0, v, r) receives 16 consecutive output channels of OUTPUT pixel row i_out (image-major
// OHxOW), as the epilogues of tc_gemm.cuh do; r = the row on the computation grid (image*G + gy*GW + gx).
//
// kFlip = the data-gradient walk: out(r) = sum_taps X[r - (ky*GW + kx)] W_tap.  The input X then is a
// zero-bordered copy of the output gradient on the OUTPUT's grid pitch GW (trailing zero columns / rows
// double as the left / top padding of the next row / image), written by the producing layer's epilogue.
template <class T, int BN, int KH, int KW, int GH, int GW, int OH, int OW, int SY, int STAGES, int OCC,
          bool kFlip, class WSrc, class Epi>
__global__ void __launch_bounds__(kShiftThreads, OCC)
    conv_shift_kernel(const void *xmap, int row_base, const WSrc w, const Epi epi, int64_t grid_rows,
                      int use_base_offset) {
  using Geo = ShiftGeom<T, BN, KH, KW, GW>;
  constexpr int G = GH * GW;
  constexpr int kMaxShift = (KH - 1) * GW + KW - 1;
  constexpr uint32_t kIdesc = make_idesc(BN, false, WSrc::kMN, Geo::kEB);
  constexpr uint32_t kCols = tmem_cols(BN);
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[STAGES], empty_bar[STAGES], acc_full[2], acc_empty[2], b_ready;
  __shared__ uint32_t tmem_slot;
  uint8_t *base = reinterpret_cast<uint8_t *>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t b_s = smem_u32(base), st_s = b_s + Geo::kBBytes;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles = (int)((grid_rows + BM - 1) / BM);

  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
#pragma unroll
    for (int b = 0; b < 2; ++b) {
      mbar_init(&acc_full[b], 1);
      mbar_init(&acc_empty[b], 4);
    }
    mbar_init(&b_ready, 4);
    fence_barrier_init();
  }
  if (warp == 1)
    tmem_alloc(&tmem_slot, 2 * kCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  if (warp == 0) {
    if (lane == 0) { // ===== TMA producer: one staged copy of the input rows per tile
      int it = 0;
      for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++it) {
        const int s = it % STAGES;
        mbar_wait(&empty_bar[s], (((uint32_t)(it / STAGES)) & 1u) ^ 1u);
        mbar_arrive_expect_tx(&full_bar[s], (uint32_t)Geo::kStage);
#pragma unroll
        for (int cp = 0; cp < Geo::kCP; ++cp)
          tma_tile_2d(st_s + (uint32_t)(s * Geo::kStage + cp * Geo::kPanel), xmap, cp * Geo::kBKE,
                      row_base + tile * BM - (kFlip ? kMaxShift : 0), &full_bar[s]); // (rows < 0: zeros)
      }
    }
  } else if (warp == 1) {
    if (lane == 0) { // ===== MMA issuer: every tap is a row-shifted view of the staged rows
      mbar_wait(&b_ready, 0u); // the resident weights have landed
      int it = 0;
      for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++it) {
        const int s = it % STAGES, buf = it & 1;
        if (it >= 2)
          mbar_wait(&acc_empty[buf], (((uint32_t)(it >> 1)) & 1u) ^ 1u);
        mbar_wait(&full_bar[s], ((uint32_t)(it / STAGES)) & 1u);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)(buf * kCols);
        const uint32_t stage = st_s + (uint32_t)(s * Geo::kStage);
#pragma unroll
        for (int t = 0; t < Geo::kTaps; ++t) {
          const int tap_shift = (t / KW) * GW + (t % KW);
          const int shift = kFlip ? kMaxShift - tap_shift : tap_shift; // row of the stage the tap starts at
          // MN-major resident weights: the layouts / descriptor constants of tc_gemm.cuh (mn_chunk_offset)
          constexpr uint32_t kLboMN = Geo::kEB == 2 ? 8192u : 4096u, kSboMN = Geo::kEB == 2 ? 1024u : 512u;
          constexpr uint32_t kTypeMN = Geo::kEB == 2 ? 2u : 1u;
          constexpr int kAdvB = WSrc::kMN ? (Geo::kEB == 2 ? 2048 : 1024) / 16 : 2;
#pragma unroll
          for (int cp = 0; cp < Geo::kCP; ++cp) {
            const uint64_t adesc = make_smem_desc_row(stage + (uint32_t)(cp * Geo::kPanel + shift * 128),
                                                      use_base_offset != 0);
            const uint32_t b_addr = b_s + (uint32_t)((t * Geo::kCP + cp) * Geo::kBPanel);
            const uint64_t bdesc = WSrc::kMN ? make_smem_desc(b_addr, kLboMN, kSboMN, kTypeMN)
                                             : make_smem_desc(b_addr, 16u, 1024u);
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
              const uint32_t acc = (uint32_t)((t | cp | kk) != 0);
              if constexpr (Geo::kEB == 4)
                umma_tf32(tmem_d, adesc + (uint64_t)(kk * 2), bdesc + (uint64_t)(kk * kAdvB), kIdesc, acc);
              else
                umma_bf16(tmem_d, adesc + (uint64_t)(kk * 2), bdesc + (uint64_t)(kk * kAdvB), kIdesc, acc);
            }
          }
        }
        umma_commit(&empty_bar[s]); // the stage is free once these UMMAs have read it
        umma_commit(&acc_full[buf]);
      }
    }
  } else { // ===== epilogue: warps 2..5, TMEM lane quarter = warp & 3
    // first: the whole weight matrix [BN][K] (K-major) -> resident 128-byte-swizzled panels, one per
    // k-block, while the producer warp already streams the first input stages in
    {
      if constexpr (!WSrc::kMN) {
        constexpr int kRowChunks = Geo::kK * Geo::kEB / 16;
        for (int q = threadIdx.x - 64; q < BN * kRowChunks; q += 128) {
          const int j = q / kRowChunks, kcg = q - j * kRowChunks, kb = kcg >> 3, kc = kcg & 7;
          const uint32_t dst = b_s + (uint32_t)(kb * Geo::kBPanel + j * 128 + ((kc ^ (j & 7)) << 4));
          cp_async16(dst, w.chunk(j, kcg), w.anchor());
        }
      } else { // one [BN x BKE] MN-major panel per k-block: chunk c = 16 bytes of consecutive output indices
        constexpr int kEPC = 16 / Geo::kEB, kRowChunks = BN / kEPC;
        for (int q = threadIdx.x - 64; q < kRowChunks * Geo::kK; q += 128) {
          const int k = q / kRowChunks, c = q - k * kRowChunks, kb = k / Geo::kBKE, kl = k - kb * Geo::kBKE;
          const uint32_t dst = b_s + (uint32_t)(kb * Geo::kBPanel) + mn_chunk_offset<Geo::kEB>(c, kl);
          cp_async16(dst, w.chunk(c, k), w.anchor());
        }
      }
      cp_async_commit();
      cp_async_wait<0>();
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0)
        mbar_arrive(&b_ready);
    }
    const int quarter = warp & 3;
    int it = 0;
    for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++it) {
      const int buf = it & 1;
      mbar_wait(&acc_full[buf], ((uint32_t)(it >> 1)) & 1u);
      tc_fence_after();
      const int64_t r = (int64_t)tile * BM + quarter * 32 + lane;
      const int64_t img = r / G;
      const int q = (int)(r - img * G), gy = q / GW, gx = q - gy * GW;
      // SY = 2: a stride-2 layer on its input grid — only every second grid row is an output row
      const bool valid = r < grid_rows && (gy % SY) == 0 && gy / SY < OH && gx < OW;
      const int64_t i_out = (img * OH + gy / SY) * OW + gx;
#pragma unroll
      for (int c = 0; c < BN; c += 16) {
        float v[16];
        tmem_ld16(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(buf * kCols + c), v);
        if (valid)
          epi((int)i_out, c, v, (int)r);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0)
        mbar_arrive(&acc_empty[buf]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1)
    tmem_dealloc(tmem_base, 2 * kCols);
}

// OCC = CTAs per SM the shared-memory footprint allows
template <class T, int BN, int KH, int KW, int GH, int GW, int OH, int OW, int SY, int STAGES, int OCC,
          bool kFlip, class WSrc, class Epi>
int launch_conv_shift(const void *xmap, int64_t row_base, const WSrc &w, const Epi &e, int64_t images,
                      cudaStream_t s, bool use_base_offset) {
  if (images <= 0)
    return ALEGPU_OK;
  using Geo = ShiftGeom<T, BN, KH, KW, GW>;
  auto kern = conv_shift_kernel<T, BN, KH, KW, GH, GW, OH, OW, SY, STAGES, OCC, kFlip, WSrc, Epi>;
  constexpr int kSmem = Geo::smem(STAGES);
  static_assert(OCC * (kSmem + 1024) <= 228 * 1024, "shift conv: shared memory over budget");
  static bool configured = false; // per template instantiation
  if (!configured) {
    ALEGPU_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
    configured = true;
  }
  const int64_t grid_rows = images * GH * GW, tiles = ceil_div(grid_rows, BM);
  ALEGPU_REQUIRE(row_base + grid_rows + Geo::kRows < ((int64_t)1 << 31), ALEGPU_INVALID_ARGUMENT,
                 "shift conv: row coordinate exceeds the TMA range");
  const unsigned grid = (unsigned)std::min<int64_t>(tiles, (int64_t)kNumSMs * OCC);
  kern<<<grid, kShiftThreads, kSmem, s>>>(xmap, (int)row_base, w, e, grid_rows, use_base_offset ? 1 : 0);
  ALEGPU_COUNT_LAUNCH_NAMED(__PRETTY_FUNCTION__);
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}

} // namespace tc
} // namespace alegpu
