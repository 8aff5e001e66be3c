// tc_gemm.cuh — the tcgen05/TMEM implicit-GEMM engine of the tensor-core Nature-CNN path.
//
//   C[i][j] = sum_r A(i,r) * B(j,r)        bf16 (kind::f16) or fp32-as-tf32 (kind::tf32) operands,
//                                          fp32 accumulation in tensor memory
//
// One CTA computes one 128 x BN tile.  Operands are never materialised in global memory: eight
// producer warps gather 16-byte chunks (8 bf16) through loader functors (im2col, transposed-conv
// and weight-gradient views live in net_tc.cu) and store them into shared memory directly in the
// 128-byte-swizzled canonical layout the UMMA shared-memory descriptors describe.  A ninth warp
// owns the tensor memory allocation and has one elected thread issue tcgen05.mma (UMMA 128xBNx16,
// kind::f16, cta_group::1); full/empty mbarriers form a STAGES-deep ring between the producers and
// the tensor core, tcgen05.commit releases ring slots and finally signals the epilogue.  The same
// eight warps then read the accumulator with tcgen05.ld (32 lanes x 16 columns per instruction)
// and run the fused epilogue functor (bias/ReLU, ReLU-mask, scatter, split partials).
//
// Element size: an operand view declares kElemBytes = 4 for fp32 data consumed as TF32 (the tensor
// core reads the upper 19 bits; producers round with cvt.rna.tf32 where they create the data); the
// default is 2 (bf16).  A k-block is always 128 bytes of reduction per row: 64 bf16 or 32 tf32
// elements = 4 UMMAs (K = 16 / K = 8), a 16-byte chunk is 8 bf16 / 4 tf32 elements, and the
// K-major byte layout below is the same for both.  MN-major tf32 operands have ONE legal layout,
// SWIZZLE_128B_BASE32B (descriptor layout type 1): atoms of 32 rows x 4 reduction indices (512 B), the
// 128-byte row of reduction index kr swizzled in 32-byte units: 16-byte chunk c sits at
// kr*128 + (((c>>1) ^ kr) << 5) + ((c&1) << 4); reduction groups of 4 are 512 B apart (SBO), 32-row
// blocks 4096 B (LBO); one UMMA (K = 8) spans two groups = 1024 B.
//
// Shared-memory tile layouts (one k-block = 64 bf16 / 32 tf32 reduction elements):
//   K-major operand  (reduction index contiguous in the 16-byte chunk):
//       row i owns 128 contiguous bytes; chunk kc sits at i*128 + ((kc ^ (i&7)) << 4);
//       8-row groups are 1024 B apart (SBO); one UMMA (K=16) advances the start address by 32 B.
//   MN-major operand (row index contiguous in the chunk, e.g. transposed weights / wgrad views):
//       64 rows x 8 reduction indices form a 1024-B atom: chunk c of reduction row kr sits at
//       kr*128 + ((c ^ kr) << 4); atoms of successive reduction groups are 1024 B apart (SBO),
//       atoms of successive 64-row blocks 8192 B apart (LBO); one UMMA advances by 2048 B.
#pragma once
#include <cuda_bf16.h>

#include <algorithm>

#include "common.cuh"

namespace alegpu {
namespace tc {

constexpr int BM = 128;          // UMMA M (cta_group::1)
constexpr int BK = 64;           // reduction elements per pipeline stage (4 UMMAs of K=16)
constexpr int kProducerWarps = 8;
constexpr int kThreads = (kProducerWarps + 1) * 32;
// bytes per operand element: 2 (bf16, kind::f16) unless the view says kElemBytes = 4 (tf32)
template <class Op, class = void> struct ElemBytes {
  static constexpr int value = 2;
};
template <class Op> struct ElemBytes<Op, decltype((void)Op::kElemBytes)> {
  static constexpr int value = Op::kElemBytes;
};
__host__ __device__ constexpr int bk_of(int elem_bytes) { return 128 / elem_bytes; }

// ---- PTX wrappers -------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
// TMA im2col load of one [pixels x channels] box (4-d NHWC tensor map): base pixel (w, h, n), filter
// offsets (off_w, off_h); lands as 128-byte swizzled rows, completion counted on `bar` in bytes
__device__ __forceinline__ void tma_im2col_4d(uint32_t dst, const void *map, int c, int w, int h,
                                              int n, uint16_t off_w, uint16_t off_h, uint64_t *bar) {
  asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.im2col.mbarrier::complete_tx::bytes"
               " [%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8};" ::"r"(dst),
               "l"(map), "r"(smem_u32(bar)), "r"(c), "r"(w), "r"(h), "r"(n), "h"(off_w), "h"(off_h)
               : "memory");
}
constexpr uint32_t kSuspendHintNs = 20000u;
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  const uint32_t a = smem_u32(bar);
  asm volatile("{\n\t"
               ".reg .pred p;\n\t"
               "WAIT_%=:\n\t"
               "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t"
               "@p bra DONE_%=;\n\t"
               "bra WAIT_%=;\n\t"
               "DONE_%=:\n\t"
               "}" ::"r"(a),
               "r"(parity), "r"(kSuspendHintNs) // the wait sleeps in hardware instead of spinning
               : "memory");
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
// generic-proxy shared-memory writes -> visible to the async proxy (tcgen05.mma operand reads)
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tmem_alloc(uint32_t *slot, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                   smem_u32(slot)),
               "r"(cols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols)
               : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc]
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc,
                                          uint32_t idesc, uint32_t accumulate) {
  asm volatile("{\n\t"
               ".reg .pred p;\n\t"
               "setp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
               "}" ::"r"(tmem_d),
               "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
               : "memory");
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc,
                                          uint32_t idesc, uint32_t accumulate) {
  asm volatile("{\n\t"
               ".reg .pred p;\n\t"
               "setp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
               "}" ::"r"(tmem_d),
               "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
               : "memory");
}
// arrive on an mbarrier once every tcgen05.mma issued so far by this thread has completed
__device__ __forceinline__ void umma_commit(uint64_t *bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                   smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
  uint32_t r[16];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 "
               "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]),
                 "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]),
                 "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
               : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i)
    v[i] = __uint_as_float(r[i]);
}

// ---- descriptors ----------------------------------------------------------------------------------
// UMMA shared-memory descriptor (SWIZZLE_128B): start>>4 [0,14) | LBO>>4 [16,30) | SBO>>4 [32,46)
// | version=1 [46,48) | layout_type=2 [61,64)
// layout_type: 2 = SWIZZLE_128B, 1 = SWIZZLE_128B_BASE32B (the MN-major tf32 layout)
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr, uint32_t lbo_bytes,
                                                   uint32_t sbo_bytes, uint32_t layout_type = 2u) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFFu);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)layout_type << 61;
  return d;
}
// byte offset inside an MN-major tile of the 16-byte chunk c (8 bf16 / 4 tf32 rows) of reduction index k
template <int kEB> __device__ __forceinline__ uint32_t mn_chunk_offset(int c, int k) {
  if constexpr (kEB == 2) { // 64-row x 8-k atoms (1024 B), 16-byte swizzle; 64-row blocks 8192 B apart
    const int kr = k & 7;
    return (uint32_t)((c >> 3) * 8192 + (k >> 3) * 1024 + kr * 128 + (((c & 7) ^ kr) << 4));
  } else { // 32-row x 4-k atoms (512 B), 32-byte swizzle; 32-row blocks 4096 B apart
    const int kr = k & 3, cc = c & 7;
    return (uint32_t)((c >> 3) * 4096 + (k >> 2) * 512 + kr * 128 + (((cc >> 1) ^ kr) << 5) +
                      ((cc & 1) << 4));
  }
}
// UMMA instruction descriptor, kind::f16 / kind::tf32: D=f32, A=B=bf16 (format 1) or tf32 (format 2),
// M=128, N=BN, per-operand major bits
__host__ __device__ constexpr uint32_t make_idesc(int n, bool a_mn, bool b_mn, int elem_bytes = 2) {
  const uint32_t fmt = elem_bytes == 4 ? 2u : 1u;
  return (1u << 4) | (fmt << 7) | (fmt << 10) | ((a_mn ? 1u : 0u) << 15) |
         ((b_mn ? 1u : 0u) << 16) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
}

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t *>(&h);
}

// ---- tile producer ----------------------------------------------------------------------------------
// One thread's share of a [ROWS x 64] operand tile, held in registers between the global gather
// and the swizzled shared-memory store so that the loads of k-block kb+1 are in flight while
// k-block kb is stored and consumed.  Loader contract:
//   K-major  (Op::kMN == false): Op::load(row, r, z)  -> Raw of elements (row, r..r+7)
//   MN-major (Op::kMN == true) : Op::load(row0, r, z) -> Raw of elements (row0..row0+7, r)
//   Op::cvt(Raw) -> uint4 of 8 bf16; Op::zero() -> the Raw that converts to eight zeros
// The loader is only called for row < rows_total and r < r_end; everything else is zero-filled.
__device__ __forceinline__ void sts16(uint32_t addr, const uint4 &v) {
  asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w)
               : "memory");
}
// MN-major register-path tiles: which index runs fastest over the threads.  Default: the 8-row
// chunks of one reduction index (contiguous 16-byte pieces of a plain matrix row).  With
// Op::kPosFastest the 64 reduction indices (im2col positions) run fastest instead: a thread keeps
// ONE position per k-block (one div/mod chain) and a warp reads neighbouring, overlapping windows.
template <class Op, class = void> struct PosFastest {
  static constexpr bool value = false;
};
template <class Op> struct PosFastest<Op, decltype((void)Op::kPosFastest)> {
  static constexpr bool value = Op::kPosFastest;
};
template <int ROWS, class Op> struct TileRegs {
  static_assert(ElemBytes<Op>::value == 2, "the register path converts to bf16; tf32 operands are async");
  static constexpr int kChunks = ROWS * 8;
  static constexpr int kProducers = kProducerWarps * 32;
  static constexpr int kPer = (kChunks + kProducers - 1) / kProducers;
  typename Op::Raw v[kPer]; // as loaded; converted to 8 bf16 only when stored (keeps loads in flight)

  __device__ __forceinline__ void load(const Op &op, int row0, int rows_total, int r0, int r_end,
                                       int z, int ptid) {
#pragma unroll
    for (int it = 0; it < kPer; ++it) {
      const int q = ptid + it * kProducers;
      v[it] = Op::zero();
      if (kChunks % kProducers == 0 || q < kChunks) {
        if constexpr (!Op::kMN) {
          const int kc = q & 7, il = q >> 3;
          const int i = row0 + il, r = r0 + kc * 8;
          if (i < rows_total && r < r_end)
            v[it] = op.load(i, r, z);
        } else {
          constexpr int kRowChunks = ROWS / 8;
          const int c = PosFastest<Op>::value ? q / BK : q % kRowChunks;
          const int k = PosFastest<Op>::value ? q % BK : q / kRowChunks;
          const int i = row0 + c * 8, r = r0 + k;
          if (i < rows_total && r < r_end)
            v[it] = op.load(i, r, z);
        }
      }
    }
  }
  __device__ __forceinline__ void store(uint32_t tile, int ptid) const {
#pragma unroll
    for (int it = 0; it < kPer; ++it) {
      const int q = ptid + it * kProducers;
      if (kChunks % kProducers == 0 || q < kChunks) {
        uint32_t off;
        if constexpr (!Op::kMN) {
          const int kc = q & 7, il = q >> 3;
          off = (uint32_t)(il * 128 + ((kc ^ (il & 7)) << 4));
        } else {
          constexpr int kRowChunks = ROWS / 8;
          const int c = PosFastest<Op>::value ? q / BK : q % kRowChunks;
          const int k = PosFastest<Op>::value ? q % BK : q / kRowChunks;
          const int kr = k & 7;
          off = (uint32_t)((c >> 3) * 8192 + (k >> 3) * 1024 + kr * 128 + (((c & 7) ^ kr) << 4));
        }
        sts16(tile + off, Op::cvt(v[it]));
      }
    }
  }
};

// ---- asynchronous tile producer (cp.async / LDGSTS) -------------------------------------------------
// Operands that already are bf16 in global memory skip the register round trip: every thread copies
// its 16-byte chunks global -> swizzled shared memory with cp.async.cg (zero-fill for chunks outside
// the matrix), so STAGES-1 k-blocks of loads are in flight per thread at no register cost.
// Loader contract (Op::kAsync == true): see AsyncCtx below.
// `fallback` is any mapped global address: with src-size 0 nothing is read and the 16 bytes are
// zero-filled, so chunks outside the matrix need no branch
__device__ __forceinline__ void cp_async16(uint32_t dst, const void *src, const void *fallback) {
  const uint32_t bytes = src ? 16u : 0u;
  const void *p = src ? src : fallback;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(p), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() {
  asm volatile("cp.async.commit_group;" ::: "memory");
}
template <int N> __device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
template <class Op, class = void> struct IsAsync {
  static constexpr bool value = false;
};
template <class Op> struct IsAsync<Op, decltype((void)Op::kAsync)> {
  static constexpr bool value = Op::kAsync;
};
// register buffers a register-path operand wants (Op::kPrefetch, default 2): small raw forms
// (u8 pixels) can afford to run their loads several k-blocks ahead
template <class Op, class = void> struct PrefetchDepth {
  static constexpr int value = 2;
};
template <class Op> struct PrefetchDepth<Op, decltype((void)Op::kPrefetch)> {
  static constexpr int value = Op::kPrefetch;
};
// Per-thread state of an asynchronous operand, set up once before the k-loop.  The address of a
// chunk is split into a ROW part and a REDUCTION part so the div/mod of the im2col views is not
// repeated for every chunk:
//   Op::row(i)      -> RowCtx   (i = matrix row; for MN-major operands the first of 8 rows)
//   Op::col(r, z)   -> ColCtx   (r = reduction index; K-major: first of 8)
//   Op::at(rc, cc)  -> address of the 16-byte chunk, or nullptr for eight zeros
// K-major tiles: a thread always serves reduction chunk kc = ptid & 7 of rows il = ptid/8 + 32*it,
// so its kPer row contexts are fixed and ONE column context per k-block serves all of them.
// MN-major tiles: a thread always serves row chunk c = ptid % (ROWS/8); only r changes.
// Op::kTma: a K-major operand fetched by ONE elected thread with a TMA im2col load per k-block
// (Op::tile(row0) -> TileCtx, Op::tma(ctx, dst, r0, z, bar)); the bytes complete on the stage's
// full barrier next to the producer warps' arrivals.
template <class Op, class = void> struct IsTma {
  static constexpr bool value = false;
};
template <class Op> struct IsTma<Op, decltype((void)Op::kTma)> {
  static constexpr bool value = Op::kTma;
};
template <int ROWS, class Op, bool kAsyncOp = IsAsync<Op>::value, bool kTmaOp = IsTma<Op>::value>
struct AsyncCtx {
  __device__ __forceinline__ void init(const Op &, int, int, int) {}
};
template <int ROWS, class Op> struct AsyncCtx<ROWS, Op, true, true> {
  typename Op::TileCtx tc;
  __device__ __forceinline__ void init(const Op &op, int row0, int, int ptid) {
    if (ptid == 0)
      tc = op.tile(row0);
  }
  __device__ __forceinline__ void issue(const Op &op, uint32_t tile, int r0, int, int z, int ptid,
                                        uint64_t *bar) const {
    if (ptid == 0) {
      mbar_expect_tx(bar, (uint32_t)(ROWS * 128));
      op.tma(tc, tile, r0, z, bar);
    }
  }
};
template <int ROWS, class Op> struct AsyncCtx<ROWS, Op, true, false> {
  static constexpr int kChunks = ROWS * 8;
  static constexpr int kProducers = kProducerWarps * 32;
  static_assert(kChunks % kProducers == 0, "tile rows must be a multiple of 32");
  static constexpr int kPer = kChunks / kProducers;
  static constexpr int kEB = ElemBytes<Op>::value;
  static constexpr int kEPC = 16 / kEB;               // elements per 16-byte chunk
  static constexpr int kRowChunks = ROWS / kEPC;      // MN-major: chunks along the row index
  static_assert(Op::kMN ? kProducers % kRowChunks == 0 : true, "row chunks must divide 256");
  static constexpr int kCtx = Op::kMN ? 1 : kPer;
  typename Op::RowCtx rc[kCtx];
  uint32_t valid; // bit it: row of chunk `it` is inside the matrix
  uint32_t off0;  // swizzled offset of chunk 0 inside the tile

  __device__ __forceinline__ void init(const Op &op, int row0, int rows_total, int ptid) {
    valid = 0u;
    if constexpr (!Op::kMN) {
      const int kc = ptid & 7, il0 = ptid >> 3;
      off0 = (uint32_t)(il0 * 128 + ((kc ^ (il0 & 7)) << 4));
#pragma unroll
      for (int it = 0; it < kPer; ++it) {
        const int i = row0 + il0 + it * 32;
        rc[it] = op.row(i < rows_total ? i : row0);
        valid |= (i < rows_total ? 1u : 0u) << it;
      }
    } else {
      const int c = ptid % kRowChunks;
      const int i = row0 + c * kEPC;
      rc[0] = op.row(i < rows_total ? i : row0);
      valid = i < rows_total ? 1u : 0u;
      off0 = 0u;
    }
  }
  __device__ __forceinline__ void issue(const Op &op, uint32_t tile, int r0, int r_end, int z,
                                        int ptid, uint64_t *) const {
    if constexpr (!Op::kMN) {
      const int r = r0 + (ptid & 7) * kEPC;
      const bool rv = r < r_end;
      const typename Op::ColCtx cc = op.col(rv ? r : r0, z);
#pragma unroll
      for (int it = 0; it < kPer; ++it) {
        const void *src = (rv && ((valid >> it) & 1u)) ? op.at(rc[it], cc) : nullptr;
        cp_async16(tile + off0 + (uint32_t)(it * 32 * 128), src, op.anchor());
      }
    } else {
      const int c = ptid % kRowChunks, k0 = ptid / kRowChunks;
#pragma unroll
      for (int it = 0; it < kPer; ++it) {
        const int k = k0 + it * (kProducers / kRowChunks);
        const int r = r0 + k;
        const uint32_t off = off0 + mn_chunk_offset<kEB>(c, k);
        const void *src = nullptr;
        if (valid && r < r_end)
          src = op.at(rc[0], op.col(r, z));
        cp_async16(tile + off, src, op.anchor());
      }
    }
  }
};
// register-path tile that compiles to nothing for asynchronous operands
template <int ROWS, class Op, bool kAsyncOp = IsAsync<Op>::value> struct TileStage {
  TileRegs<ROWS, Op> t;
  __device__ __forceinline__ void load(const Op &op, int row0, int rows_total, int r0, int r_end,
                                       int z, int ptid) {
    t.load(op, row0, rows_total, r0, r_end, z, ptid);
  }
  __device__ __forceinline__ void put(const Op &, const AsyncCtx<ROWS, Op> &, uint32_t tile, int,
                                      int, int, int ptid, uint64_t *) const {
    t.store(tile, ptid);
  }
};
template <int ROWS, class Op> struct TileStage<ROWS, Op, true> {
  __device__ __forceinline__ void load(const Op &, int, int, int, int, int, int) {}
  __device__ __forceinline__ void put(const Op &op, const AsyncCtx<ROWS, Op> &ctx, uint32_t tile,
                                      int r0, int r_end, int z, int ptid, uint64_t *bar) const {
    ctx.issue(op, tile, r0, r_end, z, ptid, bar);
  }
};

__host__ __device__ constexpr int tile_bytes(int rows) { return rows * 128; }
__host__ __device__ constexpr int smem_bytes(int bn, int stages) {
  return stages * (tile_bytes(BM) + tile_bytes(bn < 64 ? 64 : bn)) + 1024;
}
__host__ __device__ constexpr uint32_t tmem_cols(int bn) { return bn <= 32 ? 32u : bn <= 64 ? 64u : bn <= 128 ? 128u : 256u; }

// Epilogue contract: epi(i, j0, v, z) receives 16 consecutive columns j0..j0+15 of row i
// (only for i < M and j0 < N; columns >= N inside the group must be skipped by the functor).
//
// A CTA owns one (n-tile, split z) and walks the m-tiles blockIdx.x, blockIdx.x + gridDim.x, ...
// (gridDim.x == number of m-tiles gives the plain one-tile-per-CTA kernel).  Tiles are software
// pipelined: the accumulator is double-buffered in tensor memory, the operand ring keeps running
// across tile boundaries, and the producer warps issue the loads of tile t+1 BEFORE they run the
// epilogue of tile t, so for the short-reduction layers (4 k-blocks per tile) load latency, MMA
// and epilogue of neighbouring tiles overlap instead of adding up per CTA.
template <int BN, int STAGES, class AOp, class BOp, class Epi>
__global__ void __launch_bounds__(kThreads, BN <= 128 ? 2 : 1)
    gemm_tc_kernel(const AOp aop, const BOp bop, const Epi epi, int M, int N, int R,
                   int r_per_split) {
  static_assert(BN % 32 == 0 && BN >= 32 && BN <= 256, "BN must be 32..256, multiple of 32");
  static_assert(!BOp::kMN || BN >= 64, "an MN-major B tile needs at least one 64-row atom");
  constexpr int kATile = tile_bytes(BM), kBTile = tile_bytes(BN < 64 ? 64 : BN);
  constexpr int kStage = kATile + kBTile;
  constexpr int kEB = ElemBytes<AOp>::value; // 2: bf16 / kind::f16, 4: tf32 / kind::tf32
  static_assert(kEB == ElemBytes<BOp>::value, "both operands must have the same element type");
  constexpr int BK = bk_of(kEB);             // reduction elements per k-block (128 bytes per row)
  constexpr uint32_t kIdesc = make_idesc(BN, AOp::kMN, BOp::kMN, kEB);
  constexpr uint32_t kCols = tmem_cols(BN); // per accumulator buffer; two are allocated

  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[STAGES], empty_bar[STAGES], acc_full[2], acc_empty[2];
  __shared__ uint32_t tmem_slot;
  // SWIZZLE_128B atoms need 1024-byte alignment
  uint8_t *tiles = reinterpret_cast<uint8_t *>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.y * BN, z = blockIdx.z;
  const int m_tiles = (M + BM - 1) / BM;
  const int r_begin = r_per_split > 0 ? z * r_per_split : 0;
  const int r_end = r_per_split > 0 ? min(R, r_begin + r_per_split) : R;
  const int nkb = (r_end - r_begin + BK - 1) / BK;

  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], kProducerWarps);
      mbar_init(&empty_bar[s], 1);
    }
#pragma unroll
    for (int b = 0; b < 2; ++b) {
      mbar_init(&acc_full[b], 1);
      mbar_init(&acc_empty[b], kProducerWarps);
    }
    fence_barrier_init();
  }
  if (warp == kProducerWarps)
    tmem_alloc(&tmem_slot, 2 * kCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  if (warp < kProducerWarps) {
    // ===== producers (and epilogue) ================================================================
    const int ptid = threadIdx.x;
    constexpr int kBRows = BN < 64 ? 64 : BN;
    const uint32_t tiles_s = smem_u32(tiles);
    // Register-path operands (u8 / fp32 sources) are loaded kPF-1 k-blocks ahead and converted when
    // stored; asynchronous operands are copied straight into the stage.  A stage is handed to the
    // tensor core kLag k-blocks after it was issued, when its cp.async group has landed, so up to
    // STAGES-1 k-blocks of loads are in flight per thread.
    constexpr int kPF = PrefetchDepth<AOp>::value > PrefetchDepth<BOp>::value
                            ? PrefetchDepth<AOp>::value
                            : PrefetchDepth<BOp>::value;
    constexpr int kLag = STAGES - 2; // k-blocks between issuing a stage and signalling it
    static_assert(kLag >= 0 && kLag <= 4, "drain below is unrolled for kLag <= 4");
    TileStage<BM, AOp> ra[kPF];
    TileStage<kBRows, BOp> rb[kPF];
    AsyncCtx<BM, AOp> ca;
    AsyncCtx<kBRows, BOp> cb;
    cb.init(bop, n0, N, ptid);
    auto signal = [&](int g) { // g = k-block number counted over all tiles of this CTA
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0)
        mbar_arrive(&full_bar[g % STAGES]);
    };
    // issue every k-block of m-tile mt (global k-block numbers gbase..); signals all but the last
    // kLag of them
    auto issue_tile = [&](int mt, int gbase) {
      const int m0 = mt * BM;
      ca.init(aop, m0, M, ptid);
#pragma unroll
      for (int u = 0; u < kPF - 1; ++u)
        if (u < nkb) {
          ra[u].load(aop, m0, M, r_begin + u * BK, r_end, z, ptid);
          rb[u].load(bop, n0, N, r_begin + u * BK, r_end, z, ptid);
        }
      for (int kb0 = 0; kb0 < nkb; kb0 += kPF) {
#pragma unroll
        for (int u = 0; u < kPF; ++u) {
          const int kb = kb0 + u;
          if (kb < nkb) {
            if (kb + kPF - 1 < nkb) {
              ra[(u + kPF - 1) % kPF].load(aop, m0, M, r_begin + (kb + kPF - 1) * BK, r_end, z, ptid);
              rb[(u + kPF - 1) % kPF].load(bop, n0, N, r_begin + (kb + kPF - 1) * BK, r_end, z, ptid);
            }
            const int g = gbase + kb, s = g % STAGES;
            const uint32_t ph = (uint32_t)(g / STAGES) & 1u;
            mbar_wait(&empty_bar[s], ph ^ 1u);
            const uint32_t st = tiles_s + (uint32_t)(s * kStage);
            ra[u].put(aop, ca, st, r_begin + kb * BK, r_end, z, ptid, &full_bar[s]);
            rb[u].put(bop, cb, st + kATile, r_begin + kb * BK, r_end, z, ptid, &full_bar[s]);
            cp_async_commit();
            if (kb >= kLag) {
              cp_async_wait<kLag>();
              signal(g - kLag);
            }
          }
        }
      }
    };
    auto drain_tile = [&](int gbase) { // the last kLag stages of the tile
#define ALEGPU_TC_DRAIN(T)                                                                         \
  if constexpr (kLag > (T)) {                                                                      \
    if (nkb - 1 - (T) >= 0) {                                                                      \
      cp_async_wait<(T)>();                                                                        \
      signal(gbase + nkb - 1 - (T));                                                               \
    }                                                                                              \
  }
      ALEGPU_TC_DRAIN(3)
      ALEGPU_TC_DRAIN(2)
      ALEGPU_TC_DRAIN(1)
      ALEGPU_TC_DRAIN(0)
#undef ALEGPU_TC_DRAIN
    };
    // TMEM -> registers -> fused functor, for the tile whose accumulator sits in buffer it & 1
    auto epilogue = [&](int mt, int it) {
      const int buf = it & 1;
      if (nkb > 0) {
        mbar_wait(&acc_full[buf], (uint32_t)(it >> 1) & 1u);
        tc_fence_after();
      }
      const int quarter = warp & 3, half = warp >> 2;
      const int i = mt * BM + quarter * 32 + lane;
      constexpr int kColsPerHalf = BN / 2;
#pragma unroll
      for (int c = 0; c < kColsPerHalf; c += 16) {
        const int col = half * kColsPerHalf + c;
        float v[16];
        if (nkb > 0) {
          tmem_ld16(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(buf * kCols + col), v);
        } else {
#pragma unroll
          for (int t = 0; t < 16; ++t)
            v[t] = 0.0f;
        }
        if (i < M && n0 + col < N)
          epi(i, n0 + col, v, z);
      }
      if (nkb > 0) { // this warp has read its part of the accumulator buffer: hand it back
        tc_fence_before();
        __syncwarp();
        if (lane == 0)
          mbar_arrive(&acc_empty[buf]);
      }
    };
    int mt = blockIdx.x, it = 0;
    if (mt < m_tiles) {
      issue_tile(mt, 0);
      drain_tile(0);
    }
    for (; mt < m_tiles; mt += gridDim.x, ++it) {
      const int nxt = mt + gridDim.x;
      if (nxt < m_tiles)
        issue_tile(nxt, (it + 1) * nkb);
      epilogue(mt, it);
      if (nxt < m_tiles)
        drain_tile((it + 1) * nkb);
    }
  } else if (lane == 0) {
    // ===== MMA issuer: one thread drives the tensor core =========================================
    int g = 0, it = 0;
    for (int mt = blockIdx.x; mt < m_tiles && nkb > 0; mt += gridDim.x, ++it) {
      const int buf = it & 1;
      if (it >= 2) // the epilogue of tile it-2 has drained this accumulator buffer
        mbar_wait(&acc_empty[buf], ((uint32_t)(it >> 1) & 1u) ^ 1u);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + (uint32_t)(buf * kCols);
      for (int kb = 0; kb < nkb; ++kb, ++g) {
        const int s = g % STAGES;
        const uint32_t ph = (uint32_t)(g / STAGES) & 1u;
        mbar_wait(&full_bar[s], ph);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(tiles + s * kStage), b_addr = a_addr + kATile;
        // MN-major (see mn_chunk_offset): bf16 LBO 8192 / SBO 1024 / SWIZZLE_128B, one UMMA (K = 16)
        // = two 8-index groups = 2048 B; tf32 LBO 4096 / SBO 512 / SWIZZLE_128B_BASE32B, one UMMA
        // (K = 8) = two 4-index groups = 1024 B.  K-major: SBO 1024, SWIZZLE_128B, 32 B per UMMA.
        constexpr uint32_t kLboMN = kEB == 2 ? 8192u : 4096u, kSboMN = kEB == 2 ? 1024u : 512u;
        constexpr uint32_t kTypeMN = kEB == 2 ? 2u : 1u;
        constexpr int kUmmaK = 32 / kEB, kAdvMN = (kEB == 2 ? 2048 : 1024) / 16;
        const uint64_t adesc = AOp::kMN ? make_smem_desc(a_addr, kLboMN, kSboMN, kTypeMN)
                                        : make_smem_desc(a_addr, 16u, 1024u);
        const uint64_t bdesc = BOp::kMN ? make_smem_desc(b_addr, kLboMN, kSboMN, kTypeMN)
                                        : make_smem_desc(b_addr, 16u, 1024u);
#pragma unroll
        for (int kk = 0; kk < BK / kUmmaK; ++kk) {
          const uint64_t ad = adesc + (uint64_t)(kk * (AOp::kMN ? kAdvMN : 2));
          const uint64_t bd = bdesc + (uint64_t)(kk * (BOp::kMN ? kAdvMN : 2));
          if constexpr (kEB == 4)
            umma_tf32(tmem_d, ad, bd, kIdesc, (uint32_t)((kb | kk) != 0));
          else
            umma_bf16(tmem_d, ad, bd, kIdesc, (uint32_t)((kb | kk) != 0));
        }
        umma_commit(&empty_bar[s]); // slot reusable once these UMMAs have read it
      }
      umma_commit(&acc_full[buf]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kProducerWarps)
    tmem_dealloc(tmem_base, 2 * kCols);
}

template <int BN, int STAGES, class AOp, class BOp, class Epi>
int launch_gemm_tc(const AOp &a, const BOp &b, const Epi &e, int64_t M, int64_t N, int64_t R,
                   int zdim, int r_per_split, cudaStream_t s, bool persistent = true) {
  if (M <= 0 || N <= 0)
    return ALEGPU_OK;
  auto kern = gemm_tc_kernel<BN, STAGES, AOp, BOp, Epi>;
  constexpr int kSmem = smem_bytes(BN, STAGES);
  static bool configured = false; // per template instantiation
  if (!configured) {
    ALEGPU_CUDA_CHECK(
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
    configured = true;
  }
  // more tiles than resident CTA slots: each CTA walks several m-tiles (pipelined across tiles)
  int64_t gx = ceil_div(M, BM);
  const int64_t gy = ceil_div(N, BN), slots = (int64_t)kNumSMs * (BN <= 128 ? 2 : 1);
  if (persistent && gx * gy * zdim > slots)
    gx = std::max<int64_t>(1, std::min<int64_t>(gx, slots / (gy * zdim)));
  dim3 grid((unsigned)gx, (unsigned)gy, (unsigned)zdim);
  kern<<<grid, kThreads, kSmem, s>>>(a, b, e, (int)M, (int)N, (int)R, r_per_split);
  ALEGPU_COUNT_LAUNCH_NAMED(__PRETTY_FUNCTION__);
  ALEGPU_LAUNCH_CHECK();
  return ALEGPU_OK;
}

} // namespace tc
} // namespace alegpu
