// gpflow_b200 -- tcgen05 / TMEM / TMA GEMM engine (sm_100a) for the wide GP-flow shapes (C5: d=64, H=256; C3 padded).
//
// What it replaces: the per-sample `bmm` tangent propagation J_l = diag(s') W_l J_{l-1} of VectorField.jacobian
// (src/gp_flows/vector_field.py:92-128), the nn.Linear layers of VectorField.eval (:75-90), the masked matmuls of
// divergence_free_vector_func (velocity_field.py:53-79) and everything autograd derives from them, restated as
// layer-wise GEMMs over a CHUNK of samples (rows = samples x tangent directions), see gpflow_tc.cuh.
//
// One persistent, warp-specialised CTA per SM computes 128 x N output tiles D = A B^T (fp32 in, fp32 out):
//   * operands live in shared memory as K-major SWIZZLE_128B tiles of 32 fp32 (= 128 B) per row, each fp32 value split
//     x = hi + lo into two TF32 tiles; D += Ahi Bhi + Alo Bhi + Ahi Blo (3xTF32, fp32 accumulate) -- one-pass TF32
//     (2^-11) would miss the 1e-5 parity bar, the dropped lo*lo term is 2^-22;
//   * weight operands (pre-split once per call) arrive by TMA (`cp.async.bulk.tensor.2d`, SASS UTMALDG) with the 128-byte
//     swizzle applied by the hardware; activation operands are produced by 8 producer warps (coalesced global loads or
//     generated on the fly, split, swizzled st.shared) so that any layout (K-major rows, transposed, per-sample
//     coefficient vectors) feeds the same K-major tile format;
//   * one elected thread issues `tcgen05.mma.cta_group::1.kind::tf32` (SASS UTCHMMA-family), accumulators live in TMEM
//     (two N-column buffers: the epilogue of tile t overlaps the MMAs of tile t+1);
//   * 4 epilogue warps read the accumulator back with `tcgen05.ld.32x32b` (SASS LDTM), apply the op's fused epilogue
//     (bias + tanh, per-sample scaling, transposed / accumulating stores) and write global memory;
//   * smem full/empty and TMEM full/empty pipelines are mbarriers; MMA completion is signalled by `tcgen05.commit`.
// Every wait is bounded (GPF_TC_SPIN_LIMIT polls, then trap) so that a protocol bug cannot hang the GPU.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "gpflow_common.cuh"

namespace gpflow {
namespace tc {

constexpr int KC = 32;                 // fp32 elements per K chunk = one 128-byte swizzle row
constexpr int MT = 128;                // rows of A per tile = TMEM lanes
constexpr int kProdWarps = 8;
constexpr int kProdThreads = kProdWarps * 32;
constexpr int kGroupWarps = kProdWarps / 2;      // two producer groups work on alternate chunks (stages)
constexpr int kGroupThreads = kGroupWarps * 32;   // threads that cooperate on one operand tile
constexpr int kEpiWarps = 4;
constexpr int kThreads = (2 + kEpiWarps + kProdWarps) * 32;   // warp 0: TMA, warp 1: MMA, 2..5: epilogue, 6..13: producers
constexpr uint32_t kATileBytes = MT * 128;                     // 16 KB per hi / lo tile

#ifndef GPF_TC_SPIN_LIMIT
#define GPF_TC_SPIN_LIMIT (1u << 27)
#endif

// ---------------------------------------------------------------------------------------------------------------
// PTX wrappers
// ---------------------------------------------------------------------------------------------------------------
GPF_DEV void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
GPF_DEV bool mbar_try(uint64_t* bar, uint32_t parity) {
  uint32_t done;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(done)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return done != 0;
}
#ifdef GPF_TC_PROF
#define GPF_TC_PROF_DECL long long prof_wait = 0, prof_t0 = clock64();
#define GPF_TC_WAIT(bar, par) do { const long long c0_ = clock64(); mbar_wait_bounded(bar, par); prof_wait += clock64() - c0_; } while (0)
#define GPF_TC_PROF_STORE(role) do { if (p.prof != nullptr) { p.prof[(blockIdx.x * 8 + (role)) * 2] = prof_wait; p.prof[(blockIdx.x * 8 + (role)) * 2 + 1] = clock64() - prof_t0; } } while (0)
#else
#define GPF_TC_PROF_DECL
#define GPF_TC_WAIT(bar, par) mbar_wait_bounded(bar, par)
#define GPF_TC_PROF_STORE(role)
#endif
GPF_DEV void mbar_wait_bounded(uint64_t* bar, uint32_t parity) {
  uint32_t spins = 0;
  while (!mbar_try(bar, parity)) {
    if (++spins > GPF_TC_SPIN_LIMIT) {
      printf("gpflow tc: mbarrier wait timed out (block %d thread %d)\n", (int)blockIdx.x, (int)threadIdx.x);
      __trap();
    }
  }
}
GPF_DEV void tma_load_2d(void* smem_dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
          smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(smem_u32(bar))
      : "memory");
}
GPF_DEV void prefetch_tensormap(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}
GPF_DEV void tcgen05_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
GPF_DEV void tcgen05_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
GPF_DEV void tmem_alloc(uint32_t* smem_holder, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_holder)), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
GPF_DEV void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T, kind::tf32, M = 128, one K step = 8 tf32 (32 bytes)
GPF_DEV void umma_tf32(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
GPF_DEV void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
GPF_DEV void tmem_ld32(uint32_t taddr, float (&v)[32]) {
  uint32_t r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
        "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
        "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}
GPF_DEV void tmem_ld16(uint32_t taddr, float (&v)[32]) {   // 16 columns into v[0..15]
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// UMMA shared-memory descriptor: K-major, SWIZZLE_128B, rows of 128 bytes, 8-row groups 1024 bytes apart
// (cute::UMMA::SmemDescriptor: start >> 4 [0,14), LBO >> 4 [16,30), SBO >> 4 [32,46), version = 1 [46,48), layout [61,64))
GPF_DEV uint64_t make_desc_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)(1024u >> 4) << 32;   // stride byte offset between 8-row core groups
  d |= (uint64_t)1 << 46;              // descriptor version (Blackwell)
  d |= (uint64_t)2 << 61;              // SWIZZLE_128B
  return d;
}
// Instruction descriptor (cute::UMMA::InstrDescriptor): D = F32 [4,6), A = B = TF32 [7,10) [10,13), K-major both,
// N >> 3 at [17,23), M >> 4 at [24,29)
__host__ __device__ constexpr uint32_t make_idesc_tf32(int m, int n) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

// tf32 split x = hi + lo: hi = round-to-nearest tf32, lo = exact remainder (the tensor core truncates lo to tf32: 2^-21)
GPF_DEV void split_tf32(float x, float& hi, float& lo) {
  uint32_t h;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(h) : "f"(x));
  hi = __uint_as_float(h);
  const float rem = x - hi;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(h) : "f"(rem));   // round (not truncate) the remainder: unbiased 2^-22
  lo = __uint_as_float(h);
}
// store 4 consecutive K elements (16-byte chunk `c` of row `r`) of a K-major SWIZZLE_128B tile, hi and lo parts
GPF_DEV void store_chunk(char* t_hi, char* t_lo, int r, int c, float4 x) {
  float4 h, l;
  split_tf32(x.x, h.x, l.x);
  split_tf32(x.y, h.y, l.y);
  split_tf32(x.z, h.z, l.z);
  split_tf32(x.w, h.w, l.w);
  const uint32_t off = (uint32_t)r * 128u + (uint32_t)((c ^ (r & 7)) << 4);
  *reinterpret_cast<float4*>(t_hi + off) = h;
  *reinterpret_cast<float4*>(t_lo + off) = l;
}

// ---- operand fragments (called by the 256 producer threads, ptid = 0..255) -------------------------------------------
// An operand tile of R rows x 32 k is produced in two steps so that the global loads of chunk i+1 are in flight while
// the producers wait for the shared-memory stage of chunk i+1 to be released: fetch_* fills a register fragment
// (R / 8 floats per thread), commit_* splits it into tf32 hi / lo and writes the swizzled K-major tiles.
//
// ptid below is the thread index inside its producer GROUP (0 .. kGroupThreads - 1).
// ROWS mapping (operand rows are K-contiguous in memory): q = ptid + 128 it, row r = q >> 3, 16-byte chunk c = q & 7;
// 8 consecutive threads read one 128-byte row segment.
template <int R>
GPF_DEV void fetch_rows(float* frag, int ptid, const float* __restrict__ X, long long ld, long long row0, long long rows_valid,
                        int k0, int k_valid) {
#pragma unroll
  for (int it = 0; it < R * 8 / kGroupThreads; ++it) {
    const int q = ptid + it * kGroupThreads;
    const int r = q >> 3, c = q & 7;
    float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
    if (row0 + r < rows_valid && c * 4 < k_valid) x = __ldg(reinterpret_cast<const float4*>(X + (row0 + r) * ld + k0 + c * 4));
    frag[4 * it] = x.x; frag[4 * it + 1] = x.y; frag[4 * it + 2] = x.z; frag[4 * it + 3] = x.w;
  }
}
template <int R>
GPF_DEV void commit_rows(const float* frag, int ptid, char* t_hi, char* t_lo) {
#pragma unroll
  for (int it = 0; it < R * 8 / kGroupThreads; ++it) {
    const int q = ptid + it * kGroupThreads;
    store_chunk(t_hi, t_lo, q >> 3, q & 7, make_float4(frag[4 * it], frag[4 * it + 1], frag[4 * it + 2], frag[4 * it + 3]));
  }
}
// UNITS mapping (consecutive operand rows are contiguous in memory: transposed operands, per-sample column vectors):
// unit u = warp + 4 it covers rows (u >> 2) * 32 + lane and k = 8 (u & 3) .. + 7; f(m, kq, v) fills v[0..7] = tile[m][8 kq + i].
template <int R, class F>
GPF_DEV void fetch_units(float* frag, int ptid, F f) {
  const int warp = ptid >> 5, lane = ptid & 31;
#pragma unroll
  for (int it = 0; it < R / 32; ++it) {
    const int u = warp + it * kGroupWarps;
    float v[8];
    f((u >> 2) * 32 + lane, u & 3, v);
#pragma unroll
    for (int i = 0; i < 8; ++i) frag[8 * it + i] = v[i];
  }
}
template <int R>
GPF_DEV void commit_units(const float* frag, int ptid, char* t_hi, char* t_lo) {
  const int warp = ptid >> 5, lane = ptid & 31;
#pragma unroll
  for (int it = 0; it < R / 32; ++it) {
    const int u = warp + it * kGroupWarps;
    const int m = (u >> 2) * 32 + lane, kq = u & 3;
    store_chunk(t_hi, t_lo, m, 2 * kq, make_float4(frag[8 * it], frag[8 * it + 1], frag[8 * it + 2], frag[8 * it + 3]));
    store_chunk(t_hi, t_lo, m, 2 * kq + 1, make_float4(frag[8 * it + 4], frag[8 * it + 5], frag[8 * it + 6], frag[8 * it + 7]));
  }
}
enum { MODE_ROWS = 0, MODE_UNITS = 1 };
template <int R, int MODE>
GPF_DEV void commit_frag(const float* frag, int ptid, char* t_hi, char* t_lo) {
  if constexpr (MODE == MODE_ROWS) commit_rows<R>(frag, ptid, t_hi, t_lo);
  else commit_units<R>(frag, ptid, t_hi, t_lo);
}

// ---- epilogue helper: one warp's 32 rows x 32 columns staged through shared memory for coalesced row stores --------
struct EpiCtx {
  float* stage;   // this warp's [32 rows][32 floats] staging buffer (swizzled 16-byte chunks)
  int lane;
  int row0;       // first tile row of this warp (32 * TMEM lane quarter)
  // v[i] = value of tile row (row0 + lane), column c0 + i.  rowptr(r) -> global address of (tile row r, tile column 0)
  // or nullptr to skip the row.  ncols = 32 or 16.  accumulate: red.global.add instead of a plain store.
  template <class RowPtr>
  GPF_DEV void store_rows(const float (&v)[32], int nc, int c0, RowPtr rowptr, bool accumulate) const {
    char* sb = reinterpret_cast<char*>(stage);
    __syncwarp();
#pragma unroll
    for (int c = 0; c < 8; ++c)
      *reinterpret_cast<float4*>(sb + lane * 128 + ((c ^ (lane & 7)) << 4)) = make_float4(v[4 * c], v[4 * c + 1], v[4 * c + 2], v[4 * c + 3]);
    __syncwarp();
    const int cc = lane & 7;
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const int rr = it * 4 + (lane >> 3);
      const float4 x = *reinterpret_cast<const float4*>(sb + rr * 128 + ((cc ^ (rr & 7)) << 4));
      float* g = rowptr(row0 + rr);
      if (g != nullptr && cc * 4 < nc) {
        g += c0 + cc * 4;
        if (accumulate)
          asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(g), "f"(x.x), "f"(x.y), "f"(x.z), "f"(x.w) : "memory");
        else
          *reinterpret_cast<float4*>(g) = x;
      }
    }
  }
};

// ---------------------------------------------------------------------------------------------------------------
// The kernel.  Op supplies
//   NT (N tile, multiple of 16, <= 256), STAGES, B_TMA, A_MODE, B_MODE (MODE_ROWS / MODE_UNITS of the fragments),
//   Params (POD, passed by value), Tile {int nk (K chunks, may be 0 = empty tile); int n (MMA N, multiple of 16); ...},
//   num_tiles(p), decode(p, t, tile),
//   fetch_a(p, tile, kc, ptid, float* frag)          frag: MT / 4 floats in A_MODE order; ptid = thread in its group
//   fetch_b(p, tile, kc, ptid, float* frag)          when !B_TMA (NT / 4 floats, B_MODE order)
//   tmap(p, tile), tma_coords(p, tile, kc, c0, r_hi, r_lo) when B_TMA
//   epilogue(p, tile, ctx, row, col0, ncols, v)      v[i] = D[row][col0 + i]
// ---------------------------------------------------------------------------------------------------------------
// Op::A_TMA (optional, default false): the A operand (128 rows, pre-split weights) also arrives by TMA; the op then supplies
// tmap_a(p, tile) and tma_coords_a(p, tile, kc, c0, r_hi, r_lo) and no fetch_a.
template <class Op, class = void> struct OpATma { static constexpr bool value = false; };
template <class Op> struct OpATma<Op, decltype((void)Op::A_TMA)> { static constexpr bool value = Op::A_TMA; };

// Op::N_CAT (optional, default false; needs NT <= 128): the hi and lo tiles of B lie back to back in a stage, so ONE
// instruction of N = NT + n multiplies A_hi by both: D[:, 0:NT) += A_hi B_hi, D[:, NT:NT+n) += A_hi B_lo, then A_lo B_hi
// goes into the first block -- 2 instructions per K step instead of 3, and the epilogue adds the two column blocks.
// A tcgen05.mma from shared-memory operands costs ~107 cycles at N = 64 and ~114 at N = 128 (measured: the A tile is
// re-read for every instruction), so a 64-wide layer is bound by the instruction count, not by the flops.
template <class Op, class = void> struct OpNCat { static constexpr bool value = false; };
template <class Op> struct OpNCat<Op, decltype((void)Op::N_CAT)> { static constexpr bool value = Op::N_CAT; };

template <class Op>
struct GemmSmem {
  static constexpr uint32_t kAccCols = OpNCat<Op>::value ? 2 * Op::NT : Op::NT;   // TMEM columns of one accumulator buffer
  static_assert(!OpNCat<Op>::value || Op::NT <= 128, "N_CAT: NT + n <= 256");
  static constexpr uint32_t kBTileBytes = Op::NT * 128;
  static constexpr uint32_t kStageBytes = 2 * kATileBytes + 2 * kBTileBytes;
  static constexpr uint32_t kEpiOff = Op::STAGES * kStageBytes;
  static constexpr uint32_t kBarOff = kEpiOff + kEpiWarps * 4096;
  static constexpr uint32_t kBytes = kBarOff + 256 + 1024;   // + barriers + slack for the 1024-byte alignment
  static constexpr uint32_t kTmemCols = 2 * kAccCols <= 32 ? 32 : 2 * kAccCols <= 64 ? 64 : 2 * kAccCols <= 128 ? 128 : 2 * kAccCols <= 256 ? 256 : 512;
  static_assert(Op::NT % 16 == 0 && Op::NT >= 16 && Op::NT <= 256, "N tile");
  static_assert(kBytes <= 227 * 1024, "shared memory");
};

template <class Op>
__global__ void __launch_bounds__(kThreads, 1) tc_gemm_kernel(const __grid_constant__ typename Op::Params p) {
  using SM = GemmSmem<Op>;
  extern __shared__ unsigned char tc_smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(tc_smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + SM::kBarOff);
  uint64_t* empty_bar = full_bar + Op::STAGES;
  uint64_t* accfull_bar = empty_bar + Op::STAGES;
  uint64_t* accempty_bar = accfull_bar + 2;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(accempty_bar + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ntiles = Op::num_tiles(p);

  if (threadIdx.x == 0) {
    for (int s = 0; s < Op::STAGES; ++s) {
      mbar_init(&full_bar[s], ((!OpATma<Op>::value || !Op::B_TMA) ? kGroupWarps : 0) + ((OpATma<Op>::value || Op::B_TMA) ? 1 : 0));
      mbar_init(&empty_bar[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&accfull_bar[b], 1);
      mbar_init(&accempty_bar[b], kEpiWarps);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_holder, SM::kTmemCols);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_holder;

  if (warp == 0) {
    // ===== TMA producer of the weight operand(s) =====
    if constexpr (Op::B_TMA || OpATma<Op>::value) {
      if (lane == 0) {
        GPF_TC_PROF_DECL
        uint32_t stage = 0, phase = 0;
        for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
          typename Op::Tile tile;
          Op::decode(p, t, tile);
          for (int kc = 0; kc < tile.nk; ++kc) {
            GPF_TC_WAIT(&empty_bar[stage], phase ^ 1);
            unsigned char* st = smem + stage * SM::kStageBytes;
#ifdef GPF_TC_PROF
            if (p.dbg & 2) {   // timing experiment: no weight loads
              mbar_arrive(&full_bar[stage]);
              if (++stage == Op::STAGES) { stage = 0; phase ^= 1; }
              continue;
            }
#endif
            mbar_expect_tx(&full_bar[stage], (Op::B_TMA ? 2 * SM::kBTileBytes : 0) + (OpATma<Op>::value ? 2 * kATileBytes : 0));
            int c0, r_hi, r_lo;
            if constexpr (OpATma<Op>::value) {
              Op::tma_coords_a(p, tile, kc, c0, r_hi, r_lo);
              tma_load_2d(st, Op::tmap_a(p, tile), c0, r_hi, &full_bar[stage]);
              tma_load_2d(st + kATileBytes, Op::tmap_a(p, tile), c0, r_lo, &full_bar[stage]);
            }
            if constexpr (Op::B_TMA) {
              Op::tma_coords(p, tile, kc, c0, r_hi, r_lo);
              tma_load_2d(st + 2 * kATileBytes, Op::tmap(p, tile), c0, r_hi, &full_bar[stage]);
              tma_load_2d(st + 2 * kATileBytes + SM::kBTileBytes, Op::tmap(p, tile), c0, r_lo, &full_bar[stage]);
            }
            if (++stage == Op::STAGES) { stage = 0; phase ^= 1; }
          }
        }
        GPF_TC_PROF_STORE(0);
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread) =====
    if (lane == 0) {
      GPF_TC_PROF_DECL
      long long prof_wait_acc = 0;
      uint32_t stage = 0, phase = 0, acc = 0, acc_phase = 0;
      for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
        typename Op::Tile tile;
        Op::decode(p, t, tile);
        if (tile.nk == 0) continue;
        { const long long c0_ = clock64(); mbar_wait_bounded(&accempty_bar[acc], acc_phase ^ 1); prof_wait_acc += clock64() - c0_; }
        tcgen05_fence_after();
        const uint32_t tmem_d = tmem_base + acc * SM::kAccCols;
        const uint32_t idesc = make_idesc_tf32(MT, tile.n);
        const uint32_t idesc_cat = make_idesc_tf32(MT, Op::NT + tile.n);
        (void)idesc_cat;
        for (int kc = 0; kc < tile.nk; ++kc) {
          GPF_TC_WAIT(&full_bar[stage], phase);
          tcgen05_fence_after();
          const uint32_t sa = smem_u32(smem + stage * SM::kStageBytes);
          const uint64_t da_hi = make_desc_sw128(sa), da_lo = make_desc_sw128(sa + kATileBytes);
          const uint64_t db_hi = make_desc_sw128(sa + 2 * kATileBytes), db_lo = make_desc_sw128(sa + 2 * kATileBytes + SM::kBTileBytes);
          if constexpr (OpNCat<Op>::value) {
#pragma unroll
            for (int k = 0; k < KC / 8; ++k) {
              const uint64_t adv = (uint64_t)(k * 2);
              umma_tf32(tmem_d, da_hi + adv, db_hi + adv, idesc_cat, (kc | k) != 0);   // B rows: NT of the hi tile, n of the lo tile
              umma_tf32(tmem_d, da_lo + adv, db_hi + adv, idesc, 1u);
            }
            (void)db_lo;
          } else {
          // the two correction products first (they are ~2^-11 of the main one), then hi x hi
#pragma unroll
          for (int k = 0; k < KC / 8; ++k) {
            const uint64_t adv = (uint64_t)(k * 2);   // 8 tf32 = 32 bytes = 2 x 16-byte units inside the swizzle row
            umma_tf32(tmem_d, da_lo + adv, db_hi + adv, idesc, (kc | k) != 0);
            umma_tf32(tmem_d, da_hi + adv, db_lo + adv, idesc, 1u);
          }
#pragma unroll
          for (int k = 0; k < KC / 8; ++k) {
            const uint64_t adv = (uint64_t)(k * 2);
            umma_tf32(tmem_d, da_hi + adv, db_hi + adv, idesc, 1u);
          }
          }
          umma_commit(&empty_bar[stage]);   // frees the smem stage when these MMAs have read it
          if (++stage == Op::STAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit(&accfull_bar[acc]);      // accumulator complete
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
      GPF_TC_PROF_STORE(1);
#ifdef GPF_TC_PROF
      if (p.prof != nullptr) p.prof[(blockIdx.x * 8 + 4) * 2] = prof_wait_acc;
#endif
      (void)prof_wait_acc;
    }
  } else if (warp < 2 + kEpiWarps) {
    // ===== epilogue: TMEM -> registers -> (shared-memory transpose) -> global =====
    const int quarter = warp & 3;             // a warp may only touch TMEM lanes 32 (warp % 4) .. + 31
    const int row = quarter * 32 + lane;
    EpiCtx ctx;
    ctx.stage = reinterpret_cast<float*>(smem + SM::kEpiOff + (warp - 2) * 4096);
    ctx.lane = lane;
    ctx.row0 = quarter * 32;
    GPF_TC_PROF_DECL
    uint32_t acc = 0, acc_phase = 0;
    for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
      typename Op::Tile tile;
      Op::decode(p, t, tile);
      if (tile.nk == 0) continue;
      GPF_TC_WAIT(&accfull_bar[acc], acc_phase);
      tcgen05_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + acc * SM::kAccCols;
      for (int c0 = 0; c0 < tile.n; c0 += 32) {
        float v[32];
        const int nc = tile.n - c0 < 32 ? tile.n - c0 : 32;   // 16 or 32
        if (nc == 32) tmem_ld32(taddr + c0, v);
        else tmem_ld16(taddr + c0, v);
        if constexpr (OpNCat<Op>::value) {   // + the A_hi B_lo block
          float u[32];
          if (nc == 32) tmem_ld32(taddr + Op::NT + c0, u);
          else tmem_ld16(taddr + Op::NT + c0, u);
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] += (i < nc) ? u[i] : 0.f;
        }
        Op::epilogue(p, tile, ctx, row, c0, nc, v);
      }
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&accempty_bar[acc]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if (threadIdx.x == 64) GPF_TC_PROF_STORE(2);
  } else {
    // ===== operand producers: two groups of 4 warps work on alternate chunks (group g: chunks g, g + 2, ...), so that
    // one group's fence / arrive / wait-for-empty overlaps the other group's loads and stores; inside a group the
    // global loads of its next chunk are in flight while the current one is split and written =====
    const int ptid = threadIdx.x - (2 + kEpiWarps) * 32;
    const int grp = ptid >> 7, gtid = ptid & (kGroupThreads - 1);
    GPF_TC_PROF_DECL
    constexpr int kFB = Op::B_TMA ? 1 : Op::NT / 4;
    float fa[OpATma<Op>::value ? 1 : MT / 4], fb[kFB];
    struct Pos { int t, kc; typename Op::Tile tile; bool ok; };
    auto seek = [&](Pos& q) {        // first chunk of the first non-empty tile at or after q.t
      while (q.t < ntiles) {
        Op::decode(p, q.t, q.tile);
        if (q.tile.nk > 0) { q.kc = 0; q.ok = true; return; }
        q.t += gridDim.x;
      }
      q.ok = false;
    };
    auto advance = [&](Pos& q) {
      if (!q.ok) return;
      if (++q.kc < q.tile.nk) return;
      q.t += gridDim.x;
      seek(q);
    };
    auto fetch = [&](const Pos& q) {
      if (!q.ok) return;
      if constexpr (!OpATma<Op>::value) Op::fetch_a(p, q.tile, q.kc, gtid, fa);
      if constexpr (!Op::B_TMA) Op::fetch_b(p, q.tile, q.kc, gtid, fb);
    };
    Pos q;
    q.t = blockIdx.x; q.kc = 0; q.ok = false;
    if constexpr (!(OpATma<Op>::value && Op::B_TMA)) seek(q);     // both operands by TMA: nothing to produce
    if (grp == 1) advance(q);
    fetch(q);
    uint32_t c = grp;                // sequence number of the chunk in this CTA's stream
    while (q.ok) {
      const uint32_t stage = c % Op::STAGES, phase = (c / Op::STAGES) & 1;
      GPF_TC_WAIT(&empty_bar[stage], phase ^ 1);
      unsigned char* st = smem + stage * SM::kStageBytes;
      if constexpr (!OpATma<Op>::value)
        commit_frag<MT, Op::A_MODE>(fa, gtid, reinterpret_cast<char*>(st), reinterpret_cast<char*>(st + kATileBytes));
      if constexpr (!Op::B_TMA)
        commit_frag<Op::NT, Op::B_MODE>(fb, gtid, reinterpret_cast<char*>(st + 2 * kATileBytes),
                                        reinterpret_cast<char*>(st + 2 * kATileBytes + SM::kBTileBytes));
      fence_proxy_async();     // generic-proxy stores -> visible to the tensor core's async proxy
      __syncwarp();
      if (lane == 0) mbar_arrive(&full_bar[stage]);
      c += 2;
      advance(q);
      advance(q);
      fetch(q);
    }
    if (ptid == 0) GPF_TC_PROF_STORE(3);
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc(tmem_base, SM::kTmemCols);
  }
}

// ---- host side ---------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
inline EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
      f = nullptr;
    cudaGetLastError();
    return reinterpret_cast<EncodeTiledFn>(f);
  }();
  return fn;
}
// 2-D fp32 tensor [rows][cols] (row stride ld elements, ld % 4 == 0, base 16-byte aligned); box = 32 columns x box_rows;
// out-of-bounds elements read as zero.
inline bool make_tmap_2d(CUtensorMap* m, const float* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (fn == nullptr) return false;
  const cuuint64_t gdim[2] = {cols, rows};
  const cuuint64_t gstride[1] = {ld * sizeof(float)};
  const cuuint32_t box[2] = {(cuuint32_t)KC, box_rows};
  const cuuint32_t estr[2] = {1, 1};
  return fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

template <class Op>
inline cudaError_t launch_gemm(const typename Op::Params& p, int ntiles, int sm_count, cudaStream_t st) {
  using SM = GemmSmem<Op>;
  static bool attr_done = false;   // idempotent; a benign race only repeats the call
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(tc_gemm_kernel<Op>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SM::kBytes);
    if (e != cudaSuccess) return e;
    attr_done = true;
  }
  if (ntiles <= 0) return cudaSuccess;
  const int grid = ntiles < sm_count ? ntiles : sm_count;
  tc_gemm_kernel<Op><<<grid, kThreads, SM::kBytes, st>>>(p);
  return cudaGetLastError();
}

}  // namespace tc
}  // namespace gpflow
