// gemm.cu — bf16 GEMM on the 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM, operands
// staged by TMA with 128-byte swizzle) for the paths that really are dense contractions: prefill and
// batched decode.  Y[t][r] = sum_k X[t][k] * W[r][k]   (X: tokens x K bf16, W: rows x K bf16, Y fp32).
//
// "Swap-AB" orientation: the weight tile is the M operand (128 rows = 128 TMEM lanes), the token tile is
// the N operand (BN = 128 or 256 TMEM columns).  Both operands are K-major, which is how the weights
// already sit in HBM for the GEMV path, so bf16 models share one copy between decode and prefill; small
// token counts (batched decode) only shrink N.  The epilogue reads the accumulator with tcgen05.ld
// (lane = weight row, column = token), so for each token 32 lanes store 32 consecutive floats.
//
// Warp roles: warp 0 = TMA producer (one elected lane), warp 1 = TMEM allocator + MMA issuer (one elected
// lane), warps 2.. = epilogue (TMEM lane quadrant = warp % 4).  A ring of STAGES {A (MT*128)x64, B BNx64}
// tiles is handed between them with full/empty mbarriers; tcgen05.commit releases a stage when the MMAs
// that read it have retired.
//   MT = 1 (128-row weight tile): two CTAs fit per SM (smem <= 113 KB, TMEM 2 x <=256 columns), so one
//           CTA's epilogue overlaps the other's main loop.
//   MT = 2 (256 x 256 output tile, two 128x256 accumulators = all 512 TMEM columns, one CTA per SM): both
//           MMAs of a K step share the token tile, so a CTA moves 64 KB of operands per 256x256x64 MACs
//           instead of 48 KB per 128x256x64.  These GEMMs run against the L2 -> SM throughput cap (ncu:
//           tensor pipe 49 % active at MT = 1), so a third less operand traffic is a third more math.
//
// Numerics (oracle mode 1): inputs bf16, products exact, fp32 accumulation in TMEM.
#include <cuda.h>

#include <cstdlib>
#include <mutex>
#include <unordered_map>

#include "common.cuh"
#include "kernels.h"

namespace b200 {
namespace {

constexpr int BM = 128, BK = 64;  // BK * 2 B = 128 B = one swizzle row
constexpr uint32_t A_BYTES = BM * BK * 2;

DEVINL void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
      ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(bar)
      : "memory");
}
DEVINL void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
DEVINL void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
DEVINL void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc]^T, kind::f16 (bf16 inputs, fp32 accumulate), issued by ONE thread
DEVINL void tc_mma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// K-major, SWIZZLE_128B shared-memory matrix descriptor (cute::UMMA::SmemDescriptor, sm_100 version 1):
// start>>4 [0,14) | LBO>>4 [16,30) (unused for swizzled K-major) | SBO>>4 [32,46) = 1024 B between 8-row
// groups | version=1 [46,48) | layout=SWIZZLE_128B(2) [61,64)
DEVINL uint64_t smem_desc_sw128(uint32_t smem_addr) {
  return (uint64_t)((smem_addr & 0x3FFFFu) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) |
         ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}
// Instruction descriptor (cute::UMMA::InstrDescriptor): c=F32 [4,6)=1, a=BF16 [7,10)=1, b=BF16 [10,13)=1,
// a/b K-major (bits 15,16 = 0), N>>3 [17,23), M>>4 [24,29)
__host__ __device__ constexpr uint32_t instr_desc(int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

template <int BN, int MT>
struct GemmCfg {
  // BN = 32 (<= 32 tokens, weight streaming): 20 KB stages, five of them: more weight bytes in flight per CTA
  static constexpr int STAGES = BN == 32 ? 5 : ((MT == 2 && BN == 128) ? 4 : ((MT == 2 || BN == 128) ? 3 : 2));
  static constexpr int THREADS = 64 + 128 * MT;
  static constexpr uint32_t AT_BYTES = MT * A_BYTES;
  static constexpr uint32_t B_BYTES = BN * BK * 2;
  static constexpr uint32_t STAGE = AT_BYTES + B_BYTES;
  static constexpr uint32_t TMEM_COLS = BN * MT;
  static constexpr size_t SMEM = (size_t)STAGES * STAGE + 1024 /*align slack*/ + 256 /*barriers*/;
};

// Epilogues.  GEPI_F32: Y[tok][row] fp32.  GEPI_SWIGLU: the weight rows are interleaved gate_i, up_i (the ffn
// gate/up matrix), so a row pair sits in neighbouring TMEM lanes = neighbouring threads of the epilogue warp:
// H[tok][row/2] = bf16(silu(gate) * up) leaves the kernel instead of two fp32 values that a separate kernel
// would read back (same arithmetic as swiglu_bf16_kernel in prefill.cu, which stays for split-K outputs).
enum { GEPI_F32 = 0, GEPI_SWIGLU = 1 };

// One accumulator chunk: v[j] = (row, token tok0 + j) for j = 0..31 of this thread's weight row.
template <int EPI>
DEVINL void epi_store_chunk(const uint32_t (&v)[32], int tok0, int n_tokens, int row, bool row_ok, int lane, float* y, size_t ldy,
                            uint16_t* h, size_t ldh) {
  if (EPI == GEPI_F32) {
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      const int tok = tok0 + j;
      if (tok < n_tokens && row_ok) y[(size_t)tok * ldy + row] = __uint_as_float(v[j]);
    }
  } else {
    // even lane = gate row, odd lane = up row.  The even lane finishes token j, the odd lane token j + 1: one
    // exchange per token pair hands each the half it lacks.
    const bool odd = lane & 1;
#pragma unroll
    for (int j = 0; j < 32; j += 2) {
      const float x = __shfl_xor_sync(0xffffffffu, __uint_as_float(odd ? v[j] : v[j + 1]), 1);
      const float gate = odd ? x : __uint_as_float(v[j]);
      const float up = odd ? __uint_as_float(v[j + 1]) : x;
      const int tok = tok0 + j + (odd ? 1 : 0);
      if (tok < n_tokens && row_ok) h[(size_t)tok * ldh + (row >> 1)] = f32_to_bf16_bits(silu_mul(gate, up));
    }
  }
}

struct GemmArgs {
  int n_fast;     // grid.x = token tiles, grid.y = weight-row tiles (else the other way round)
  float* y;
  uint16_t* h;   // GEPI_SWIGLU: bf16 output [tok][rows / 2], ldh elements between tokens
  int ldh;
  int ldy;       // floats between consecutive tokens of y
  int n_tokens;  // valid tokens (TMA zero-fills the rest of a token tile)
  int n_rows;    // valid weight rows (the last row tile may be ragged: vocab-parallel lm_head shards)
  int kb_total;  // K / 64 blocks; split-K: CTA z of gridDim.z takes blocks [kb_total*z/Z, kb_total*(z+1)/Z) (the runs need not be equal)
  long long split_stride;  // floats between the partial outputs of consecutive K splits
};

template <int BN, int MT, int EPI>
__global__ void __launch_bounds__(64 + 128 * MT) gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA,
                                                                   const __grid_constant__ CUtensorMap tmB, const GemmArgs g) {
  using C = GemmCfg<BN, MT>;
  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B tiles must start on 1024-byte boundaries
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t tiles = (raw + 1023u) & ~1023u;
  uint8_t* bar_mem = smem_raw + (tiles - raw) + C::STAGES * C::STAGE;
  uint64_t* full = reinterpret_cast<uint64_t*>(bar_mem);
  uint64_t* empty = full + C::STAGES;
  uint64_t* tmem_full = empty + C::STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);
  const uint32_t full_s = smem_u32(full), empty_s = smem_u32(empty), tfull_s = smem_u32(tmem_full);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // n_fast: the token tile is the fastest grid dimension, so the CTAs of a wave cover few weight-row tiles x all token
  // tiles and share the weight fetches in L2 (many token tiles: batched prefill); else the row tile is fastest
  const int m0 = (g.n_fast ? blockIdx.y : blockIdx.x) * (BM * MT), n0 = (g.n_fast ? blockIdx.x : blockIdx.y) * BN;
  const int kb0 = (int)((long long)g.kb_total * blockIdx.z / gridDim.z);  // split-K: this CTA's first K block ...
  const int k_blocks = (int)((long long)g.kb_total * (blockIdx.z + 1) / gridDim.z) - kb0;  // ... and how many

  if (threadIdx.x == 0) {
    for (int s = 0; s < C::STAGES; ++s) {
      mbar_init(full_s + 8 * s, 1);
      mbar_init(empty_s + 8 * s, 1);
    }
    mbar_init(tfull_s, 1);
    mbar_fence_init();
  }
  if (warp == 1) {  // whole warp: allocate the TMEM columns (power of two >= 32), then give up the permit
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(C::TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {  // ===== TMA producer =====
      int s = 0;
      uint32_t ph = 0;
      for (int kb = 0; kb < k_blocks; ++kb) {
        mbar_wait(empty_s + 8 * s, ph ^ 1);
        const uint32_t dst = tiles + s * C::STAGE;
        mbar_expect_tx(full_s + 8 * s, C::STAGE);
#pragma unroll
        for (int t = 0; t < MT; ++t) tma_load_2d(dst + t * A_BYTES, &tmA, (kb0 + kb) * BK, m0 + t * BM, full_s + 8 * s);
        tma_load_2d(dst + C::AT_BYTES, &tmB, (kb0 + kb) * BK, n0, full_s + 8 * s);
        if (++s == C::STAGES) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {  // ===== MMA issuer =====
      constexpr uint32_t idesc = instr_desc(BM, BN);
      int s = 0;
      uint32_t ph = 0;
      for (int kb = 0; kb < k_blocks; ++kb) {
        mbar_wait(full_s + 8 * s, ph);
        tc_fence_after();
        const uint64_t ad = smem_desc_sw128(tiles + s * C::STAGE);
        const uint64_t bd = smem_desc_sw128(tiles + s * C::STAGE + C::AT_BYTES);
#pragma unroll
        for (int k = 0; k < BK / 16; ++k) {  // UMMA_K = 16 bf16 = 32 B: advance the start address by 2 (x16 B)
#pragma unroll
          for (int t = 0; t < MT; ++t)  // accumulator t: TMEM columns [t*BN, (t+1)*BN), weight rows [128t, 128t+128)
            tc_mma(tmem_base + t * BN, ad + t * (A_BYTES >> 4) + 2 * k, bd + 2 * k, idesc, (uint32_t)(kb | k));
        }
        tc_commit(empty_s + 8 * s);  // stage reusable once these MMAs have read it
        if (++s == C::STAGES) { s = 0; ph ^= 1; }
      }
      tc_commit(tfull_s);  // accumulator complete
    }
  } else {  // ===== epilogue: TMEM -> registers -> global =====
    mbar_wait(tfull_s, 0);
    tc_fence_after();
    const int quad = warp & 3;  // a warp may only touch TMEM lanes [32*(warp%4), +32)
    const int mt = (warp - 2) >> 2;  // which accumulator (MT = 2: warps 2..5 -> 0, warps 6..9 -> 1)
    const int row = m0 + mt * BM + quad * 32 + lane;
#pragma unroll 1
    for (int c = 0; c < BN / 32; ++c) {
      uint32_t v[32];
      const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(mt * BN + c * 32);
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
          "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
          : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
            "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
            "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
            "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
          : "r"(taddr));
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      epi_store_chunk<EPI>(v, n0 + c * 32, g.n_tokens, row, row < g.n_rows, lane, g.y + (size_t)blockIdx.z * g.split_stride, (size_t)g.ldy, g.h,
                           (size_t)g.ldh);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(C::TMEM_COLS) : "memory");
}

// ---------------------------------------------------------------------------------------------------------
// Stream-K variant of the 256 x 256 kernel: one persistent CTA per SM, the (tile, K block) work items of the
// whole GEMM are cut into equal contiguous runs, one per CTA, so the GPU is evenly loaded whatever the tile
// count (gate/up at 512 tokens is 224 tiles = 1.51 waves of 148).  A run that starts inside a tile computes
// that tile's remaining K blocks first and leaves its partial accumulator in a per-CTA workspace slot (flag
// = ready); a run that ends inside a tile holds its first K blocks, computed last, and that CTA is the tile's
// owner: it adds the partials of the following CTAs in CTA order (deterministic) and stores the tile.  The
// partial producers run at the start of their timeline and the owner at the end of its own, so nobody waits.
struct StreamKArgs {
  float* y;
  uint16_t* h;        // GEPI_SWIGLU output
  int ldh;
  int ldy, n_tokens;
  int tiles_n, n_fast;  // token tiles; n_fast: tile index = m_tile * tiles_n + n_tile, else n_tile * tiles_m + m_tile
  int tiles_m;        // weight-row tiles; tile index = n_tile * tiles_m + m_tile (row tile fastest: a CTA's
                      // consecutive tiles share the token tile, and a row tile's token tiles run close in time)
  int kb_total;       // K blocks (of 64) per tile
  int units;          // work items per CTA
  int total;          // tiles * kb_total
  float* ws;          // [grid][256 tokens][256 rows] fp32 partials
  unsigned int* flags;  // [grid]; 1 = that CTA's partial is in ws (the owner clears it)
};

constexpr size_t kWsSlot = (size_t)256 * 2 * BM;  // floats per workspace slot: [256 tokens][256 rows]
constexpr uint32_t kFixBytes = 32 * 2 * BM * 4;  // one chunk of a partial: 32 tokens x 256 rows fp32 (the fix-up stages them in the operand ring)

// The work of CTA p as a sequence of segments (tile, K blocks [k0, k1)).  The producer, the MMA issuer and the
// epilogue warps each walk their own copy.
struct SegIter {
  int KB, u, u_end;
  DEVINL SegIter(const StreamKArgs& g, int p) : KB(g.kb_total) {
    u = p * g.units;
    u_end = min(u + g.units, g.total);
  }
  DEVINL bool next(int& tile, int& k0, int& k1) {
    if (u >= u_end) return false;
    tile = u / KB; k0 = u - tile * KB; k1 = min(KB, k0 + (u_end - u));
    u += k1 - k0;
    return true;
  }
};

DEVINL void bar_sync_epi(int n) { asm volatile("bar.sync 1, %0;" ::"r"(n) : "memory"); }

template <int BN, int EPI>
__global__ void __launch_bounds__(320) gemm_streamk_kernel(const __grid_constant__ CUtensorMap tmA,
                                                            const __grid_constant__ CUtensorMap tmB, const StreamKArgs g) {
  using C = GemmCfg<BN, 2>;
  constexpr int kFixBufs = (int)(C::STAGES * C::STAGE / kFixBytes);  // 5 (32-token tiles) or 6: the more partial chunks in flight, the shorter the fix-up (4 -> 5: 107 -> 105 us)
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t tiles = (raw + 1023u) & ~1023u;
  uint8_t* bar_mem = smem_raw + (tiles - raw) + C::STAGES * C::STAGE;
  uint64_t* full = reinterpret_cast<uint64_t*>(bar_mem);
  uint64_t* empty = full + C::STAGES;
  uint64_t* tmem_full = empty + C::STAGES;
  uint64_t* tmem_empty = tmem_full + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 1);
  const uint32_t full_s = smem_u32(full), empty_s = smem_u32(empty), tfull_s = smem_u32(tmem_full), tempty_s = smem_u32(tmem_empty);
  const uint32_t fix_s = smem_u32(bar_mem + 128);  // kFixBufs barriers of the owner's fix-up (below)
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int p = blockIdx.x, KB = g.kb_total;

  if (threadIdx.x == 0) {
    for (int s = 0; s < C::STAGES; ++s) {
      mbar_init(full_s + 8 * s, 1);
      mbar_init(empty_s + 8 * s, 1);
    }
    mbar_init(tfull_s, 1);
    mbar_init(tempty_s, 8);  // one arrival per epilogue warp
    for (int b = 0; b < kFixBufs; ++b) mbar_init(fix_s + 8 * b, 1);
    mbar_fence_init();
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(C::TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {  // ===== TMA producer: runs ahead of the MMAs across segment boundaries =====
      int s = 0;
      uint32_t ph = 0;
      SegIter it(g, p);
      for (int tile, k0, k1; it.next(tile, k0, k1);) {
        const int m0 = (g.n_fast ? tile / g.tiles_n : tile % g.tiles_m) * (2 * BM), n0 = (g.n_fast ? tile % g.tiles_n : tile / g.tiles_m) * BN;
        for (int kb = k0; kb < k1; ++kb) {
          mbar_wait(empty_s + 8 * s, ph ^ 1);
          const uint32_t dst = tiles + s * C::STAGE;
          mbar_expect_tx(full_s + 8 * s, C::STAGE);
          tma_load_2d(dst, &tmA, kb * BK, m0, full_s + 8 * s);
          tma_load_2d(dst + A_BYTES, &tmA, kb * BK, m0 + BM, full_s + 8 * s);
          tma_load_2d(dst + C::AT_BYTES, &tmB, kb * BK, n0, full_s + 8 * s);
          if (++s == C::STAGES) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {  // ===== MMA issuer =====
      constexpr uint32_t idesc = instr_desc(BM, BN);
      int s = 0, seg = 0;
      uint32_t ph = 0;
      SegIter it(g, p);
      for (int tile, k0, k1; it.next(tile, k0, k1); ++seg) {
        mbar_wait(tempty_s, (uint32_t)((seg & 1) ^ 1));  // the previous segment's accumulators have been read out
        tc_fence_after();
        for (int kb = k0; kb < k1; ++kb) {
          mbar_wait(full_s + 8 * s, ph);
          tc_fence_after();
          const uint64_t ad = smem_desc_sw128(tiles + s * C::STAGE);
          const uint64_t bd = smem_desc_sw128(tiles + s * C::STAGE + C::AT_BYTES);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            tc_mma(tmem_base, ad + 2 * k, bd + 2 * k, idesc, (uint32_t)((kb - k0) | k));
            tc_mma(tmem_base + BN, ad + (A_BYTES >> 4) + 2 * k, bd + 2 * k, idesc, (uint32_t)((kb - k0) | k));
          }
          tc_commit(empty_s + 8 * s);
          if (++s == C::STAGES) { s = 0; ph ^= 1; }
        }
        tc_commit(tfull_s);
      }
    }
  } else {  // ===== epilogue warps 2..9 =====
    const int quad = warp & 3, mt = (warp - 2) >> 2;
    const int rloc = mt * BM + quad * 32 + lane;  // row within the 256-row tile
    int seg = 0;
    SegIter it(g, p);
    for (int tile, k0, k1; it.next(tile, k0, k1); ++seg) {
      const int m0 = (g.n_fast ? tile / g.tiles_n : tile % g.tiles_m) * (2 * BM), n0 = (g.n_fast ? tile % g.tiles_n : tile / g.tiles_m) * BN;
      const bool partial = k0 > 0;            // somebody else owns this tile: leave the accumulator in the workspace
      const bool owner_split = k0 == 0 && k1 < KB;  // this CTA owns the tile and others hold its later K blocks
      int n_parts = 0;
      if (owner_split) {  // CTAs p+1 .. p+n_parts hold [k1, KB)
        n_parts = (KB - k1 + g.units - 1) / g.units;
        if (threadIdx.x == 64) {
          for (int j = 1; j <= n_parts; ++j) {
            const long long t0 = clock64();
            unsigned int v;
            do {
              asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(g.flags + p + j) : "memory");
            } while (v == 0u && clock64() - t0 < 4000000000ll);  // bounded: a lost CTA must not hang the GPU ...
            if (v == 0u) __trap();                                // ... and must not go unnoticed either: the launch fails
          }
        }
        bar_sync_epi(256);
      }
      mbar_wait(tfull_s, (uint32_t)(seg & 1));
      tc_fence_after();
      // The owner's fix-up: the other CTAs' partials come in through the TMA unit, 32 KB (32 tokens x 256 rows) at a
      // time, into the operand ring — idle now: the owned segment is the last of this CTA's run and its MMAs have
      // retired.  (Plain loads, 32 per thread in flight, took 1.6 us per 32 KB round under load: 27-50 us for one or
      // two partials of a 256 x 256 tile.)  Item i = chunk i / n_parts of partial i % n_parts + 1.
      const int n_chunks = min(BN / 32, (g.n_tokens - n0 + 31) / 32);
      const int n_items = n_parts * n_chunks;
      auto fix_issue = [&](int i) {
        const int c = i / n_parts, q = i % n_parts + 1, b = i % kFixBufs;
        mbar_expect_tx(fix_s + 8 * b, kFixBytes);
        bulk_g2s(tiles + (uint32_t)b * kFixBytes, g.ws + (size_t)(p + q) * kWsSlot + (size_t)(c * 32) * (2 * BM), kFixBytes, fix_s + 8 * b,
                 l2_policy_evict_first());
      };
      if (n_items > 0 && threadIdx.x == 64) {
        asm volatile("fence.proxy.async;" ::: "memory");  // the partials were written with ordinary stores (acquired above)
        for (int i = 0; i < min(kFixBufs, n_items); ++i) fix_issue(i);
      }
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        if (n0 + c * 32 >= g.n_tokens) break;  // token columns past the batch hold zeros: nothing to store or reduce
        uint32_t v[32];
        const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(mt * BN + c * 32);
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
            "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
            : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
              "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
              "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
              "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
            : "r"(taddr));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        if (partial) {
          float* w = g.ws + (size_t)p * kWsSlot + (size_t)(c * 32) * (2 * BM) + rloc;
#pragma unroll
          for (int j = 0; j < 32; ++j) w[(size_t)j * (2 * BM)] = __uint_as_float(v[j]);
        } else {
          for (int q = 1; q <= n_parts; ++q) {  // partials in CTA order
            const int i = c * n_parts + q - 1, b = i % kFixBufs;
            mbar_wait(fix_s + 8 * b, (uint32_t)((i / kFixBufs) & 1));
            const float* w = reinterpret_cast<const float*>(smem_raw + (tiles - raw) + (size_t)b * kFixBytes) + rloc;
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = __float_as_uint(__uint_as_float(v[j]) + w[j * (2 * BM)]);
            bar_sync_epi(256);  // everybody has read buffer b
            if (threadIdx.x == 64 && i + kFixBufs < n_items) {
              fence_proxy_async();
              fix_issue(i + kFixBufs);
            }
          }
          epi_store_chunk<EPI>(v, n0 + c * 32, g.n_tokens, m0 + rloc, true, lane, g.y, (size_t)g.ldy, g.h, (size_t)g.ldh);
        }
      }
      tc_fence_before();
      if (lane == 0) mbar_arrive(tempty_s);
      if (partial) {  // publish the partial
        __threadfence();
        bar_sync_epi(256);
        if (threadIdx.x == 64) asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(g.flags + p), "r"(1u) : "memory");
      }
      if (owner_split) {  // all reads of the partials are done: re-arm the flags for the next launch
        bar_sync_epi(256);
        if (threadIdx.x == 64)
          for (int j = 1; j <= n_parts; ++j) g.flags[p + j] = 0u;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(C::TMEM_COLS) : "memory");
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// bf16 [rows][K] row-major -> 2-D tensor map with a {64 (K), box_rows} box, 128-byte swizzle.
// Encoded once per (buffer, shape, box) and kept: the same few hundred operands (weight matrices, the engine's
// activation buffers) come back on every prefill and every batched step.
struct MapKey {
  const void* base;
  uint64_t rows, K;
  uint32_t box_rows;
  bool operator==(const MapKey& o) const { return base == o.base && rows == o.rows && K == o.K && box_rows == o.box_rows; }
};
struct MapKeyHash {
  size_t operator()(const MapKey& k) const {
    uint64_t h = (uint64_t)(uintptr_t)k.base * 0x9E3779B97F4A7C15ull;
    h ^= (k.rows + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2));
    h ^= (k.K * 0xC2B2AE3D27D4EB4Full + (h << 6) + (h >> 2));
    return (size_t)(h ^ k.box_rows);
  }
};
bool make_map(CUtensorMap* m, const void* base, uint64_t rows, uint64_t K, uint32_t box_rows) {
  static std::mutex mu;
  static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
  const MapKey key{base, rows, K, box_rows};
  {
    std::lock_guard<std::mutex> lk(mu);
    auto it = cache.find(key);
    if (it != cache.end()) { *m = it->second; return true; }
  }
  EncodeTiledFn fn = encode_fn();
  if (!fn) return false;
  const cuuint64_t dims[2] = {K, rows};
  const cuuint64_t strides[1] = {K * 2};
  const cuuint32_t box[2] = {(cuuint32_t)BK, box_rows};
  const cuuint32_t estr[2] = {1, 1};
  if (fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
         CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
    return false;
  std::lock_guard<std::mutex> lk(mu);
  if (cache.size() > 8192) cache.clear();  // hooks that upload a fresh matrix per call: bounded
  cache.emplace(key, *m);
  return true;
}

// multiprocessors of the current device (asked once per device): the split heuristics deal CTAs out over them
int sm_count_now() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 148;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n < 1) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

// Environment switches of the schedule heuristics.  They exist for experiments and for the tests, which flip them
// between calls inside one process, so they are looked up per launch (a handful of getenv calls, ~0.1 us each, on the
// host thread that is enqueueing ahead of the GPU anyway).
bool env_set(const char* name) { return getenv(name) != nullptr; }
#define kNoStreamK env_set("B200_NO_STREAMK")
#define kNoBn32 env_set("B200_NO_GEMM_BN32")
int env_int(const char* name, int dflt) {
  const char* e = getenv(name);
  return e ? atoi(e) : dflt;
}

// Which tile dimension runs fastest over the grid of the whole-tile kernel.  With the row tile fastest a wave of 148
// CTAs covers many weight-row tiles of one or two token tiles, and every further token tile re-reads the weights from
// HBM; with the token tile fastest a wave covers ~148 / tiles_n row tiles under ALL token tiles, which share the
// weight fetches in L2.  Measured at 4096 tokens (batched prefill): gate/up 827 -> 729 us, qkv 205 -> 170, o 127 ->
// 115, down 335 -> 324; at two token tiles (512 tokens) the orders are equivalent and the row tile stays fastest.
// (B200_GEMM_NFAST = 0 / 1 forces it, for experiments.)
int n_fast_order(int tiles_m, int tiles_n) {
  const int ov = env_int("B200_GEMM_NFAST", -1);
  if (tiles_n <= 1 || tiles_m >= 65535) return 0;
  if (ov >= 0) return ov;
  return tiles_n >= 4;
}

// Where a launch puts its result: fp32 Y (possibly split-K partials) or, for GEPI_SWIGLU, bf16 H.
struct GemmOut {
  float* y = nullptr;
  int64_t ldy = 0;
  uint16_t* h = nullptr;
  int64_t ldh = 0;
};

template <int BN, int MT, int EPI>
cudaError_t launch_bn(const uint16_t* W, int64_t rows, int64_t K, const uint16_t* X, int64_t n_tokens, const GemmOut& o, int splits,
                      int64_t split_stride, cudaStream_t s) {
  using C = GemmCfg<BN, MT>;
  CUtensorMap ma, mb;
  if (!make_map(&ma, W, (uint64_t)rows, (uint64_t)K, BM) || !make_map(&mb, X, (uint64_t)n_tokens, (uint64_t)K, BN))
    return cudaErrorInvalidValue;
  auto kern = gemm_bf16_kernel<BN, MT, EPI>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM);
  if (e != cudaSuccess) return e;
  GemmArgs g;
  g.y = o.y; g.ldy = (int)o.ldy; g.h = o.h; g.ldh = (int)o.ldh;
  g.n_tokens = (int)n_tokens; g.n_rows = (int)rows; g.kb_total = (int)(K / BK); g.split_stride = split_stride;
  const unsigned tm = (unsigned)((rows + BM * MT - 1) / (BM * MT)), tn = (unsigned)((n_tokens + BN - 1) / BN);
  g.n_fast = n_fast_order((int)tm, (int)tn);
  dim3 grid(g.n_fast ? tn : tm, g.n_fast ? tm : tn, (unsigned)splits);
  count_launch(); kern<<<grid, C::THREADS, C::SMEM, s>>>(ma, mb, g);
  return cudaGetLastError();
}

template <int BN, int EPI>
cudaError_t launch_streamk(const uint16_t* W, int64_t rows, int64_t K, const uint16_t* X, int64_t n_tokens, const GemmOut& o,
                           const GemmWorkspace& wk, cudaStream_t s) {
  using C = GemmCfg<BN, 2>;
  CUtensorMap ma, mb;
  if (!make_map(&ma, W, (uint64_t)rows, (uint64_t)K, BM) || !make_map(&mb, X, (uint64_t)n_tokens, (uint64_t)K, BN))
    return cudaErrorInvalidValue;
  cudaError_t e = cudaFuncSetAttribute(gemm_streamk_kernel<BN, EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM);
  if (e != cudaSuccess) return e;
  StreamKArgs g;
  g.y = o.y; g.ldy = (int)o.ldy; g.h = o.h; g.ldh = (int)o.ldh; g.n_tokens = (int)n_tokens;
  g.tiles_m = (int)(rows / (2 * BM));
  g.tiles_n = (int)((n_tokens + BN - 1) / BN);
  g.n_fast = 0;  // stream-K keeps the row tile fastest (measured: the token-tile order costs it 8-12 % at 512-4096 tokens)
  g.kb_total = (int)(K / BK);
  const int tiles = g.tiles_m * (int)((n_tokens + BN - 1) / BN);
  g.total = tiles * g.kb_total;
  g.units = (g.total + wk.sm_count - 1) / wk.sm_count;
  const int grid = (g.total + g.units - 1) / g.units;
  g.ws = wk.ws; g.flags = wk.flags;
  count_launch(); gemm_streamk_kernel<BN, EPI><<<grid, C::THREADS, C::SMEM, s>>>(ma, mb, g);
  return cudaGetLastError();
}

// One token tile (n_tokens <= 128): the GEMM streams the weights once and is HBM-bound; what matters is that
// every SM streams the same number of bytes with no per-tile set-up in between.  Whole tiles cannot do that
// (gate/up: 224 row tiles of 128 on 148 SMs -> 75 us where the batch-1 GEMV needs 42), the stream-K schedule
// can: 148 persistent CTAs, each an equal run of (256-row tile, K block) items, partial tiles (tokens x 256
// floats, tiny here) reduced by the tile's owner in CTA order.
bool streamk_small_n(int64_t rows, int64_t K, int64_t n_tokens, const GemmWorkspace* wk) {
  if (!wk || !wk->ws || !wk->flags || wk->sm_count <= 0 || kNoStreamK) return false;
  if (n_tokens > 128 || rows % (2 * BM) || (size_t)wk->sm_count * kWsSlot * 4 > wk->ws_bytes) return false;
  const int64_t total = (rows / (2 * BM)) * (K / BK);
  if (env_int("B200_GEMM_STREAMK", 0) == 1) return true;  // tests: takes the schedule for every shape it can express
  // Measured (8B, 8 tokens): gate/up 37 us vs 75 us in whole tiles, lm_head 171 vs 232; but matrices with few row
  // tiles (o, down, qkv: 16-24) are cut into ~10 CTAs per tile and the owner's serial fix-up outweighs the balance
  // (down 49 vs 23 us), and from 64 tokens up the partial traffic does.  So: at most 32 tokens, and at least half a
  // row tile per CTA.
  return n_tokens <= 32 && 2 * (rows / (2 * BM)) >= wk->sm_count && total >= 4 * (int64_t)wk->sm_count;
}

// B200_GEMM_M256 = 0 never / 1 whenever the shape allows / unset: the heuristic below (experiments only)
int m256_override() { return env_int("B200_GEMM_M256", -1); }

}  // namespace

bool gemm_shape_ok(int64_t rows, int64_t K, int64_t n_tokens) {
  // rows: any multiple of 8 (ragged last tile: TMA zero-fills, the epilogue masks); K: whole 64-column blocks
  return rows > 0 && rows % 8 == 0 && K > 0 && K % BK == 0 && n_tokens > 0 && n_tokens <= (1 << 20);
}

namespace {

// Prefill-sized GEMMs whose 256 x 256 tiles do not come to whole waves: stream-K over all tiles.  Measured on the
// gate/up matrix (28672 x 4096), whole tiles -> stream-K: 512 tokens (224 tiles = 1.51 waves) 119.7 -> 102 us, 1024
// tokens 235 -> 183, 2048 tokens 433 -> 392, 4096 tokens (12.1 waves) 827 -> 688 us = 1.4 PFLOP/s.  Besides the
// balance, CTAs at different K positions contend less for L2 than a wave in lock-step (0.63 vs 0.83 us per K block).
// The other side of that staggering: CTAs that work on the same weight rows are at different K and no longer share
// the fetch in L2.  So it is taken for matrices with many row tiles (gate/up, lm_head: the CTAs of a moment are on
// different rows anyway), and otherwise only when K is short and there are between one and one and a half waves of
// tiles (qkv at 2048 tokens: 113 -> 99 us); ffn_down at 4096 tokens (K = 14336, 16 row tiles x 16 token tiles) went
// from 324 to 574 us under it — every tile streamed its 7 MB weight slab from HBM on its own.
// Not for GEMMs of less than a wave (qkv, o, down at 512 tokens: 32-48 tiles): a CTA's extra epilogues — 5 us to
// leave a partial, 12 us for the owner to pull one in and add it — outweigh the balance there; split-K stays.
// (Also measured: whole tiles for the full waves and K runs only for the last, partial wave.  Every CTA then writes
// its partial and every owner reads at the same moment, the owners wait for the slowest producer: 120-150 us at 512
// tokens.  Not kept.)   B200_GEMM_SK256 = 0 never / 1 whenever the shape allows (experiments, tests).
bool streamk_large_n(int64_t rows, int64_t K, int64_t n_tokens, const GemmWorkspace* wk) {
  if (!wk || !wk->ws || !wk->flags || wk->sm_count <= 0 || kNoStreamK) return false;
  if (n_tokens <= 128 || rows % (2 * BM) || (size_t)wk->sm_count * kWsSlot * 4 > wk->ws_bytes) return false;
  const int ov = env_int("B200_GEMM_SK256", -1);
  if (ov == 0) return false;
  if (ov == 1) return true;
  const int64_t tiles_m = rows / (2 * BM), tiles = tiles_m * ((n_tokens + 255) / 256), n_sm = wk->sm_count, rem = tiles % n_sm;
  if (tiles < n_sm || rem == 0) return false;
  if (tiles_m >= 64) return rem * 20 <= n_sm * 17;
  return K <= 8192 && tiles < 2 * n_sm && rem * 2 <= n_sm;
}

// One schedule for both epilogues.  Returns cudaErrorNotSupported when EPI needs the whole K in one CTA and the
// schedule that fits the shape splits it (the caller then takes the fp32 path and the separate kernel).
template <int EPI>
cudaError_t dispatch(const uint16_t* W, int64_t rows, int64_t K, const uint16_t* X, int64_t n_tokens, const GemmOut& o, int64_t split_stride,
                     int* splits, cudaStream_t s, const GemmWorkspace* wk) {
  if (!gemm_shape_ok(rows, K, n_tokens)) return cudaErrorInvalidValue;
  if (streamk_small_n(rows, K, n_tokens, wk)) {
    *splits = 1;
    if (n_tokens <= 32) return launch_streamk<32, EPI>(W, rows, K, X, n_tokens, o, *wk, s);
    return launch_streamk<128, EPI>(W, rows, K, X, n_tokens, o, *wk, s);
  }
  if (streamk_large_n(rows, K, n_tokens, wk)) {
    *splits = 1;
    return launch_streamk<256, EPI>(W, rows, K, X, n_tokens, o, *wk, s);
  }
  const int max_splits = EPI != GEPI_F32 ? 1 : (*splits < 1 ? 1 : (*splits > 16 ? 16 : *splits));
  const int64_t kb = K / BK, n_sm = wk && wk->sm_count > 0 ? wk->sm_count : sm_count_now();
  const int ov = m256_override();
  // 256 x 256 tiles, one CTA per SM: the fewest operand bytes per MAC.  Taken when, with K split up to
  // max_splits ways (each split keeps >= 24 K blocks: shorter loops do not amortise the CTA set-up), the grid comes to between 65 % of the SMs and all of them, or has several waves.
  if (rows % (2 * BM) == 0 && n_tokens > 128 && ov != 0) {
    const int64_t tiles = (rows / (2 * BM)) * ((n_tokens + 255) / 256);
    int use = 1;
    for (int sp = 2; sp <= max_splits; sp *= 2)
      if (kb % sp == 0 && kb / sp >= 24 && tiles * sp <= n_sm) use = sp;
    if (tiles * use >= (n_sm * 13) / 20 || ov == 1) {  // 96 of 148: the qkv GEMM at 512 tokens (48 tiles x 2 splits) sits exactly there
      *splits = use;
      return launch_bn<256, 2, EPI>(W, rows, K, X, n_tokens, o, use, split_stride, s);
    }
  }
  // 128-row tiles, two CTAs per SM; token tile 256 when that still yields >= 2 CTAs per SM of work, else 128
  const int64_t tiles256 = ((rows + BM - 1) / BM) * ((n_tokens + 255) / 256);
  const int64_t tiles128 = ((rows + BM - 1) / BM) * ((n_tokens + 127) / 128);
  int use = 1;
  if (n_tokens <= 128 && max_splits >= 2) {
    // One token tile: the GEMM streams the weights once and is HBM-bound, so what matters is that every SM gets
    // the same share of the rows (224 row tiles on 148 SMs leave half the GPU waiting for the SMs that got two:
    // measured 75 us where the batch-1 GEMV takes 42).  Partials are tiny here (tokens x rows), so K is split until
    // the CTAs deal out evenly: the s (>= 8 blocks per CTA) with the best
    // CTAs / (rounds x SMs), the smallest such s on ties.
    double best = 0.0;
    for (int sp = 1; sp <= max_splits && sp <= 16; ++sp) {
      if (kb / sp < 8) continue;  // (the runs may be unequal: qkv, 48 row tiles x 64 K blocks, takes 3 x 21-22 = 144 CTAs)
      const int64_t ctas = tiles128 * sp, rounds = (ctas + n_sm - 1) / n_sm;
      const double eff = (double)ctas / (double)(rounds * n_sm);
      if (eff > best + 0.02) { best = eff; use = sp; }
    }
  } else if (max_splits >= 2 && tiles128 < 2 * n_sm && tiles128 >= 32 && kb % 2 == 0 && K >= 2048) use = 2;
  *splits = use;
  if (n_tokens > 128 && tiles256 >= 2 * n_sm) return launch_bn<256, 1, EPI>(W, rows, K, X, n_tokens, o, 1, 0, s);
  if (n_tokens <= 32 && !kNoBn32) return launch_bn<32, 1, EPI>(W, rows, K, X, n_tokens, o, use, split_stride, s);
  return launch_bn<128, 1, EPI>(W, rows, K, X, n_tokens, o, use, split_stride, s);
}

}  // namespace

// Y[t*ldy + r] = sum_k X[t][k] W[r][k];  W, X bf16 (K-major, 16-byte aligned rows), Y fp32.
cudaError_t launch_gemm_bf16(const uint16_t* W, int64_t rows, int64_t K, const uint16_t* X, int64_t n_tokens, float* Y,
                             int64_t ldy, cudaStream_t s, const GemmWorkspace* wk) {
  int splits = 1;
  return launch_gemm_bf16_splitk(W, rows, K, X, n_tokens, Y, ldy, 0, &splits, s, wk);
}

// Same, but a GEMM with too few output tiles to fill the GPU may split K: partial p is written to
// Y + p*split_stride and the caller adds the partials (deterministically, ascending p) where it consumes Y.
// *splits: in = maximum allowed (1..16), out = what was used.
cudaError_t launch_gemm_bf16_splitk(const uint16_t* W, int64_t rows, int64_t K, const uint16_t* X, int64_t n_tokens, float* Y,
                                    int64_t ldy, int64_t split_stride, int* splits, cudaStream_t s, const GemmWorkspace* wk) {
  GemmOut o;
  o.y = Y; o.ldy = ldy;
  return dispatch<GEPI_F32>(W, rows, K, X, n_tokens, o, split_stride, splits, s, wk);
}

// H[t*ldh + i] = bf16( silu(g) * u ),  g = X[t] . W[2i], u = X[t] . W[2i+1]  (ffn gate/up, rows interleaved).
// The epilogue does it when one CTA holds a tile's whole K (every schedule except the balance-by-split one at
// 33..128 tokens); otherwise fp32 partials go to `scratch` ([max_splits][n_tokens][rows]) and swiglu_bf16_kernel
// finishes.  H must not overlap X.
cudaError_t launch_gemm_bf16_swiglu(const uint16_t* W, int64_t rows, int64_t K, const uint16_t* X, int64_t n_tokens, uint16_t* H, int64_t ldh,
                                    float* scratch, int max_splits, cudaStream_t s, const GemmWorkspace* wk) {
  if (rows % 2) return cudaErrorInvalidValue;
  if (!env_set("B200_NO_GEMM_SWIGLU")) {
    // would the F32 schedule split K here?  Only the <= 128-token whole-tile path does, for balance.
    const bool balance_split = n_tokens <= 128 && max_splits >= 2 && !streamk_small_n(rows, K, n_tokens, wk);
    if (!balance_split) {
      GemmOut o;
      o.h = H; o.ldh = ldh;
      int one = 1;
      return dispatch<GEPI_SWIGLU>(W, rows, K, X, n_tokens, o, 0, &one, s, wk);
    }
  }
  if (ldh != rows / 2) return cudaErrorInvalidValue;  // the separate kernel writes H densely
  int splits = max_splits;
  const int64_t stride = n_tokens * rows;
  cudaError_t e = launch_gemm_bf16_splitk(W, rows, K, X, n_tokens, scratch, rows, stride, &splits, s, wk);
  if (e != cudaSuccess) return e;
  return launch_swiglu_bf16(scratch, H, n_tokens * (rows / 2), splits, stride, s);
}

}  // namespace b200
