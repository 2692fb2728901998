// Prefill-regime grouped GEMM on the 5th-generation tensor cores (tcgen05 + TMEM), NF4 weights dequantised on the fly.
//
//   D[n, m] = sum_k W_e[n, k] * X[m, k]      n: 128 weight rows of expert e (MMA M, one TMEM lane each)
//                                             m: up to 192 routed tokens of expert e (MMA N, one TMEM column each)
//
// The WEIGHTS are the A operand and live in TENSOR MEMORY: a dequant thread owns one weight row, so the fp32 absmax of
// a 64-wide NF4 block is a per-thread scalar, the 32 packed bytes of the block are two LDS.128, every byte is decoded
// through a bank-conflict-free shared-memory table (byte -> the two fp32 NF4 codes), scaled in fp32 and rounded ONCE
// to fp16 -- bit-identical to bitsandbytes' dequantize_4bit, so the tensor cores see exactly the fp16 matrix the
// reference multiplies with -- and the 32 packed half2 go to TMEM with one tcgen05.st.32x32b.x32.  The ACTIVATIONS are
// the B operand, loaded by TMA (128B swizzle, K-major) straight into the canonical UMMA layout.  Each dequantised
// weight therefore feeds up to 192 tokens, 1.5x fewer dequants than the usual "tokens = M = 128" arrangement, and no
// dequantised weight ever touches shared or global memory.
//
// Roles (512 threads, one persistent CTA per SM, registers re-balanced per warpgroup with setmaxnreg):
//   warp 0       TMA producer: activation tiles [<=192 tokens x 64] per k-block, as 32-row boxes matching the ragged tile
//   warp 1       TMEM allocator and the single MMA-issuing thread (tcgen05.mma kind::f16, A in TMEM, B in smem)
//   warp 2       TMA producer: the tile's absmax (one bulk copy per tile, double-buffered) and raw NF4 tiles [128 rows x 128 B]
//                per 4 k-blocks, on their own rings so they run ahead of the MMA
//   warps 4-11   dequant: two groups of four warps (one lane quarter each) on alternate k-blocks
//   warps 12-15  epilogue: tcgen05.ld of one of the TWO accumulators while the next tile's main loop fills the other;
//                gpt-oss clamped SwiGLU on adjacent lanes (rows 2j / 2j+1 = gate / up) for w13, scatter to (token, slot)
//                order for w2
// Replaces vLLM's dequantise-everything + Triton fused_experts (V/.../quantization/bitsandbytes.py:478-506, 588-603).
#pragma once
#include <cuda.h>

#include <cstdio>

#include "common.cuh"

namespace nf4moe {

constexpr int kGemmAutoThreshold = 64;  // tokens; above this nf4moe_moe_forward(AUTO) uses the tcgen05 path
constexpr int kGemmThreads = 512;
constexpr int kGemmTokTile = 192;       // max MMA N: two accumulators of 192 columns + four A stages fill the 512 TMEM columns
constexpr int kGemmRowTile = 128;       // MMA M
constexpr int kGemmBoxRows = 32;        // token rows per activation TMA box (a tile is ceil(rows/32) boxes)
constexpr int kAStages = 4;             // A (dequantised weights) stages in TMEM, 32 columns each
constexpr int kMaxBStages = 4;          // activation stages: exactly kAStages, stage = k-block & 3
constexpr int kGemmWPrefetch = 6;        // raw-weight stages requested into L2 ahead of the shared-memory ring
constexpr int kMaxWStages = 4;
constexpr uint32_t kWStageBytes = 128 * 128;   // raw NF4: 128 rows x 4 k-blocks x 32 B
constexpr uint32_t kGemmSmemBytes = 232448 - 1024;  // dynamic part; the kernel's static __shared__ control block takes the rest
constexpr uint32_t kGemmLutBytes = 32768;
constexpr uint32_t kTmemCols = 512;
constexpr uint32_t kTmemACol = 2 * kGemmTokTile;   // accumulators at columns 0 and 192, A stages at [384, 384 + 4*32)

inline bool gemm_available() { return true; }

// ---- tcgen05 / TMA wrappers -------------------------------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem desc]
__device__ __forceinline__ void tc_mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(acc)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
      ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(bar)
      : "memory");
}
// One lane of a CONVERGED warp.  The single-thread roles run their loops warp-uniformly and predicate only the
// tcgen05 / TMA instruction on this: operands then stay in uniform registers (a `lane == 0` branch makes ptxas emit an
// ELECT + R2UR waterfall loop around every UTCHMMA / UTMALDG, ~100 cycles each).
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ float2 lds_f32x2(uint32_t addr) {
  float2 v;
  asm("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(addr));
  return v;
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
      "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]),
      "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]),
      "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}

struct GemmArgs {
  QW w;
  void* out;                      // W13: h fp16 [rows, I] (or raw [rows, 2I]);  W2: y fp16 [T*k, rows]
  const int32_t* sorted_pairs;    // W2 scatter
  const int32_t* meta;            // meta[2] = number of token tiles
  const int32_t* tile_list;       // (expert, first sorted row, rows, -) per token tile
  int act;
  long long* dbg;                 // optional [64] cycle counters written by CTA 0 (profiling builds of the caller only)
  int tile_tokens;                // largest token tile (256, or 32 for small batches): sizes the activation stages
};

// instruction descriptor: D f32, A/B f16, both K-major, M = 128, N = n
__device__ __forceinline__ uint32_t make_idesc(int n) {
  return (1u << 4) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(kGemmRowTile >> 4) << 24);
}
// K-major 128B-swizzled operand tile: 8-row atoms of 1024 B (SBO), descriptor version 1, layout SWIZZLE_128B
__device__ __forceinline__ uint64_t make_bdesc(uint32_t smem_addr) {
  return (uint64_t)((smem_addr & 0x3FFFFu) >> 4) | ((uint64_t)(1024u >> 4) << 32) | (1ull << 46) | (2ull << 61);
}

#define GEMM_DBG_T0() long long dbg_t = dbg_on ? clock64() : 0
#define GEMM_DBG_ADD(slot)                                  \
  do {                                                      \
    if (dbg_on) {                                           \
      const long long now__ = clock64();                    \
      dbg_acc[slot] += now__ - dbg_t;                       \
      dbg_t = now__;                                        \
    }                                                       \
  } while (0)

// barrier slots inside s_bar[]
constexpr int kBarWFull = 0, kBarWEmpty = kBarWFull + kMaxWStages, kBarBFull = kBarWEmpty + kMaxWStages,
              kBarAFull = kBarBFull + kMaxBStages, kBarAEmpty = kBarAFull + kAStages,
              kBarDFull = kBarAEmpty + kAStages, kBarDEmpty = kBarDFull + 2, kBarMFull = kBarDEmpty + 2,
              kBarMEmpty = kBarMFull + 2, kNumGemmBars = kBarMEmpty + 2;

template <int N>
__device__ __forceinline__ void set_max_regs_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }
template <int N>
__device__ __forceinline__ void set_max_regs_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }

template <bool W13>
__global__ void __launch_bounds__(kGemmThreads, 1) gemm_kernel(const GemmArgs a, const __grid_constant__ CUtensorMap map_w,
                                                              const __grid_constant__ CUtensorMap map_x) {
  extern __shared__ __align__(1024) uint8_t dyn_smem[];
  __shared__ __align__(8) uint64_t s_bar[kNumGemmBars];
  __shared__ uint32_t s_tmem;
  __shared__ uint32_t s_wst[kMaxWStages], s_bst[kMaxBStages], s_mst[2];
  __shared__ int s_nw, s_nb;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nkb = a.w.nkb;
  const bool nested = a.w.absmax_f32 == nullptr;
  const uint32_t b_bytes = (uint32_t)a.tile_tokens * 128u;
  // absmax stage: the tile's [128 rows x nkb] absmax (fp32, or u8 + the expert's 1 KB code2 table when nested)
  const uint32_t m_tile = (uint32_t)(kGemmRowTile * nkb) * (nested ? 1u : 4u);
  const uint32_t m_code2 = (m_tile + 15u) & ~15u;            // offset of code2 inside the stage
  const uint32_t m_bytes = (m_code2 + (nested ? 1024u : 0u) + 1023u) & ~1023u;
  const uint32_t base = smem_u32(dyn_smem);
  // decode table: byte -> (fp32 code of the high nibble, fp32 code of the low nibble), one 8-byte copy per lane of a
  // half-warp (an LDS.64 is served 16 lanes per wavefront), 32 KB, 32 KB-aligned so the address is (byte << 7) | lane bits
  const uint32_t lut = pin((base + kGemmLutBytes - 1u) & ~(kGemmLutBytes - 1u));
  const uint32_t bar0 = smem_u32(s_bar);
  auto bar = [&](int i) -> uint32_t { return bar0 + 8u * (uint32_t)i; };

  // ---- one-time setup ----
  if (tid == 0) {
    // carve [base, lut) and [lut + 32K, end) into 1024-aligned blocks: 4 activation stages, 2 absmax stages, 2 raw-weight
    // stages, then more raw-weight stages while they fit
    uint32_t cur[2] = {(base + 1023u) & ~1023u, lut + kGemmLutBytes};
    const uint32_t hi[2] = {lut, base + kGemmSmemBytes};
    int ib = 0, iw = 0, im = 0;
    auto place = [&](uint32_t bytes) -> uint32_t {
      int best = -1;
      uint32_t best_room = 0;
      for (int r = 0; r < 2; ++r) {   // best fit: the region with the least room that still takes the block
        const uint32_t room = hi[r] > cur[r] ? hi[r] - cur[r] : 0u;
        if (room >= bytes && (best < 0 || room < best_room)) { best = r; best_room = room; }
      }
      if (best < 0) return 0u;
      const uint32_t at = cur[best];
      cur[best] += bytes;
      return at;
    };
    bool ok = true;
    for (int i = 0; i < kMaxBStages && ok; ++i) { const uint32_t at = place(b_bytes); if (at) s_bst[ib++] = at; else ok = false; }
    for (int i = 0; i < 2 && ok; ++i) { const uint32_t at = place(m_bytes); if (at) s_mst[im++] = at; else ok = false; }
    for (int i = 0; i < 2 && ok; ++i) { const uint32_t at = place(kWStageBytes); if (at) s_wst[iw++] = at; else ok = false; }
    while (ok && iw < kMaxWStages) { const uint32_t at = place(kWStageBytes); if (!at) break; s_wst[iw++] = at; }
    if (!ok) ib = 0;
    s_nb = ib; s_nw = iw;
    for (int i = 0; i < kMaxWStages; ++i) { mbar_init(bar(kBarWFull + i), 1); mbar_init(bar(kBarWEmpty + i), 8); }
    for (int i = 0; i < kMaxBStages; ++i) mbar_init(bar(kBarBFull + i), 1);
    for (int i = 0; i < kAStages; ++i) { mbar_init(bar(kBarAFull + i), 4); mbar_init(bar(kBarAEmpty + i), 1); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar(kBarDFull + i), 1); mbar_init(bar(kBarDEmpty + i), 4);
      mbar_init(bar(kBarMFull + i), 1); mbar_init(bar(kBarMEmpty + i), 8);
    }
    fence_barrier_init();
  }
  for (int i = tid; i < 256 * 16; i += kGemmThreads) {
    const int b = i >> 4, l = i & 15;
    const uint32_t p = lut + (uint32_t)b * 128u + (uint32_t)l * 8u;
    sts32(p, __float_as_uint(kNF4[b >> 4]));        // element 2j   (high nibble)
    sts32(p + 4u, __float_as_uint(kNF4[b & 15]));   // element 2j+1 (low nibble)
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(kTmemCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;
  const int n_bst = s_nb, n_wst = s_nw;

  const bool dbg_on = a.dbg != nullptr && blockIdx.x == 0 && lane == 0;
  long long dbg_acc[6] = {0, 0, 0, 0, 0, 0};
  const int n_row_tiles = (a.w.rows + kGemmRowTile - 1) / kGemmRowTile;
  const int n_items = n_bst > 0 ? a.meta[2] * n_row_tiles : 0;   // (n_bst == 0: K too large for the stage budget; host rejects it)

  if (warp < 4) {
    set_max_regs_dec<56>();
    if (warp == 0) {
      // =============================== TMA producer: activation tiles, one stage per k-block ===============================
      uint32_t phb = 0;   // one phase bit per stage; stage = k-block & 3, the same slot the A operand of that k-block uses
      for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        const int tile = item / n_row_tiles;
        const int row0 = a.tile_list[4 * tile + 1], n_box = (a.tile_list[4 * tile + 2] + kGemmBoxRows - 1) / kGemmBoxRows;
        for (int kb = 0; kb < nkb; ++kb) {
          const int sb = kb & 3;
          GEMM_DBG_T0();
          mbar_wait(bar(kBarAEmpty + sb), ((phb >> sb) & 1u) ^ 1u);
          GEMM_DBG_ADD(0);
          const uint32_t dst = s_bst[sb];
          if (elect_one()) {
            mbar_expect_tx(bar(kBarBFull + sb), (uint32_t)n_box * (kGemmBoxRows * 128u));
            for (int bx = 0; bx < n_box; ++bx)   // only the 32-row boxes the tile really has: L2 -> SM traffic follows the ragged tile
              tma_load_2d(dst + (uint32_t)bx * (kGemmBoxRows * 128u), &map_x, kb * kBlock, row0 + bx * kGemmBoxRows, bar(kBarBFull + sb));
          }
          __syncwarp();
          phb ^= 1u << sb;
          GEMM_DBG_ADD(1);
        }
      }
      if (dbg_on) for (int i = 0; i < 6; ++i) a.dbg[0 + i] = dbg_acc[i];
    } else if (warp == 2) {
      // ============ TMA producer: the tile's absmax (one bulk copy per item, double-buffered) and the raw NF4 bytes, one
      // stage per 4 k-blocks, on their own rings so they run ahead of the MMA ============
      int sw = 0;
      uint32_t phw = 0;
      const uint64_t pol = l2_policy_evict_last();
      const uint32_t am_elt = nested ? 1u : 4u;
      const uint8_t* am_base = nested ? a.w.absmax_u8 : reinterpret_cast<const uint8_t*>(a.w.absmax_f32);
      const int spi = (nkb + 3) >> 2;                       // raw-weight stages per item
      const int my_items = n_items > (int)blockIdx.x ? (n_items - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
      const int total_st = my_items * spi;
      auto stage_coords = [&](int st, int& c0, int& c1) {   // global stage number -> TMA coordinates (byte column, weight row)
        const int it_s = st / spi, item = (int)blockIdx.x + it_s * (int)gridDim.x;
        const int tile = item / n_row_tiles, rt = item - tile * n_row_tiles;
        c0 = (st - it_s * spi) * 128;
        c1 = a.tile_list[4 * tile] * a.w.rows + rt * kGemmRowTile;
      };
      int pf = 0;                                           // L2 prefetch cursor: kGemmWPrefetch stages ahead of the loads
      for (int st = 0; st < total_st; ++st) {
        for (; pf < total_st && pf <= st + kGemmWPrefetch; ++pf) {
          int c0, c1;
          stage_coords(pf, c0, c1);
          if (elect_one())
            asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global.tile [%0, {%1, %2}];" ::"l"(&map_w), "r"(c0), "r"(c1) : "memory");
        }
        if (st % spi == 0) {
          const int it = st / spi, item = (int)blockIdx.x + it * (int)gridDim.x;
          const int tile = item / n_row_tiles, rt = item - tile * n_row_tiles;
          const int e = a.tile_list[4 * tile];
          const int mb = it & 1;
          const uint32_t mph = (uint32_t)(it >> 1) & 1u;
          const int rows_here = min(kGemmRowTile, a.w.rows - rt * kGemmRowTile);
          const uint32_t bytes = (uint32_t)(rows_here * nkb) * am_elt;
          mbar_wait(bar(kBarMEmpty + mb), mph ^ 1u);
          if (elect_one()) {
            mbar_expect_tx(bar(kBarMFull + mb), bytes + (nested ? 1024u : 0u));
            bulk_g2s(s_mst[mb], am_base + ((long long)e * a.w.nb + (long long)rt * kGemmRowTile * nkb) * am_elt, bytes,
                     bar(kBarMFull + mb), pol);
            if (nested) bulk_g2s(s_mst[mb] + m_code2, a.w.code2 + (long long)e * a.w.code2_stride, 1024u, bar(kBarMFull + mb), pol);
          }
          __syncwarp();
        }
        int c0, c1;
        stage_coords(st, c0, c1);
        GEMM_DBG_T0();
        mbar_wait(bar(kBarWEmpty + sw), phw ^ 1u);
        GEMM_DBG_ADD(0);
        if (elect_one()) {
          mbar_expect_tx(bar(kBarWFull + sw), kWStageBytes);
          tma_load_2d(s_wst[sw], &map_w, c0, c1, bar(kBarWFull + sw));
        }
        __syncwarp();
        if (++sw == n_wst) { sw = 0; phw ^= 1u; }
      }
      if (dbg_on) for (int i = 0; i < 6; ++i) a.dbg[8 + i] = dbg_acc[i];
    } else if (warp == 1) {
      // =============================== MMA issuer (one thread) ===============================
      int it = 0;
      uint32_t pha = 0;   // one phase bit per stage (A in TMEM and B in smem of k-block kb both live in slot kb & 3)
      for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++it) {
        const int tile = item / n_row_tiles;
        const int n_tok = a.tile_list[4 * tile + 2];
        const uint32_t idesc = make_idesc(max(16, (n_tok + 15) & ~15));
        const int db = it & 1;
        const uint32_t d_t = tmem + (uint32_t)db * kGemmTokTile;
        GEMM_DBG_T0();
        mbar_wait(bar(kBarDEmpty + db), (((uint32_t)(it >> 1)) & 1u) ^ 1u);   // the epilogue has drained this accumulator
        GEMM_DBG_ADD(0);
        tc_fence_after();
        for (int kb = 0; kb < nkb; ++kb) {
          const int sa = kb & 3;
          mbar_wait(bar(kBarAFull + sa), (pha >> sa) & 1u);
          GEMM_DBG_ADD(1);
          mbar_wait(bar(kBarBFull + sa), (pha >> sa) & 1u);
          GEMM_DBG_ADD(2);
          tc_fence_after();
          const uint32_t a_t = tmem + kTmemACol + (uint32_t)sa * 32u;
          const uint32_t b_s = s_bst[sa];
          if (elect_one()) {
#pragma unroll
            for (int kk = 0; kk < 4; ++kk)
              tc_mma_ts(d_t, a_t + 8u * kk, make_bdesc(b_s + 32u * kk), idesc, (uint32_t)(kb | kk));
            tc_commit(bar(kBarAEmpty + sa));     // frees the A stage (dequant warps) and the B stage (TMA producer) together
            if (kb == nkb - 1) tc_commit(bar(kBarDFull + db));
          }
          __syncwarp();
          pha ^= 1u << sa;
          GEMM_DBG_ADD(3);
        }
      }
      if (dbg_on) for (int i = 0; i < 6; ++i) a.dbg[16 + i] = dbg_acc[i];
    }
  }
  else if (warp < 12) {
    // =============================== dequant: NF4 -> fp16 A operand in TMEM ===============================
    set_max_regs_inc<168>();
    const int q = warp & 3, g = (warp - 4) >> 2;
    const int r = q * 32 + lane;
    const uint32_t lut_lane = pin(lut | ((uint32_t)(lane & 15) << 3));
    const uint32_t t_lane = tmem + ((uint32_t)(q * 32) << 16);
    const uint32_t swz = (uint32_t)(r & 7);
    int sw = 0, it = 0;
    uint32_t phw = 0, use0 = 0, use1 = 0;   // use counts of this group's two A stages (g and g+2)
    for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++it) {
      const int tile = item / n_row_tiles, rt = item - tile * n_row_tiles;
      const int e = a.tile_list[4 * tile];
      const int n = rt * kGemmRowTile + r;
      const int mb = it & 1;
      // nested absmax: code2[u8] * absmax2[blk >> 8] + offset; this row's 45 blocks touch at most two absmax2 entries
      float a2_lo = 0.f, a2_hi = 0.f, off_e = 0.f;
      long long blk_hi = 0;   // first block index that uses a2_hi
      if (nested) {
        const long long blk0 = (long long)min(n, a.w.rows - 1) * nkb;
        const long long i2 = blk0 >> 8;
        a2_lo = __ldg(a.w.absmax2 + (long long)e * a.w.n2 + i2);
        a2_hi = __ldg(a.w.absmax2 + (long long)e * a.w.n2 + min(i2 + 1, a.w.n2 - 1));
        off_e = __ldg(a.w.offset + e);
        blk_hi = ((i2 + 1) << 8) - blk0;   // relative to this row's first block
      }
      GEMM_DBG_T0();
      mbar_wait(bar(kBarMFull + mb), (uint32_t)(it >> 1) & 1u);
      GEMM_DBG_ADD(0);
      const uint32_t m_st = pin(s_mst[mb]);
      const uint32_t m_row = m_st + (uint32_t)(r * nkb) * (nested ? 1u : 4u);
      for (int kb0 = 0; kb0 < nkb; kb0 += 4) {
        mbar_wait(bar(kBarWFull + sw), phw);
        GEMM_DBG_ADD(1);
        const uint32_t wrow = pin(s_wst[sw]) + (uint32_t)r * 128u;
#pragma unroll
        for (int jj = 0; jj < 2; ++jj) {
          const int j = g + 2 * jj;
          const int kb = kb0 + j;
          if (kb < nkb) {
            float am;
            if (nested) {
              const float c = lds_f32(m_st + m_code2 + 4u * lds_u8(m_row + (uint32_t)kb));
              am = __fadd_rn(__fmul_rn(c, kb < blk_hi ? a2_lo : a2_hi), off_e);
            } else {
              am = lds_f32(m_row + 4u * (uint32_t)kb);
            }
            const uint4 v0 = lds128(wrow + (((uint32_t)(2 * j) ^ swz) << 4));
            const uint4 v1 = lds128(wrow + (((uint32_t)(2 * j + 1) ^ swz) << 4));
            const uint32_t wd[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
            uint32_t regs[32];
#pragma unroll
            for (int wi = 0; wi < 8; ++wi)
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const uint32_t sh = i == 0 ? wd[wi] << 7 : wd[wi] >> (8 * i - 7);
                const float2 f = lds_f32x2((sh & 0x7F80u) | lut_lane);
                regs[wi * 4 + i] = pack_half2(__fmul_rn(f.x, am), __fmul_rn(f.y, am));
              }
            uint32_t& use = jj ? use1 : use0;
            GEMM_DBG_ADD(2);
            mbar_wait(bar(kBarAEmpty + j), (use & 1u) ^ 1u);   // the MMAs that read this stage last time are done
            GEMM_DBG_ADD(3);
            ++use;
            tc_fence_after();
            tmem_st32(t_lane + kTmemACol + (uint32_t)j * 32u, regs);
            asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar(kBarAFull + j));
            GEMM_DBG_ADD(4);
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(bar(kBarWEmpty + sw));
        if (++sw == n_wst) { sw = 0; phw ^= 1u; }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(kBarMEmpty + mb));
    }
    if (dbg_on && (warp == 4 || warp == 8)) for (int i = 0; i < 6; ++i) a.dbg[(warp == 4 ? 24 : 32) + i] = dbg_acc[i];
  } else {
    // =============================== epilogue: TMEM -> registers -> global, overlapped with the next tile's main loop ======
    set_max_regs_dec<112>();
    const int q = warp & 3;
    const int r = q * 32 + lane;
    const uint32_t t_lane = tmem + ((uint32_t)(q * 32) << 16);
    int it = 0;
    for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++it) {
      const int tile = item / n_row_tiles, rt = item - tile * n_row_tiles;
      const int row0 = a.tile_list[4 * tile + 1], n_tok = a.tile_list[4 * tile + 2];
      const int n = rt * kGemmRowTile + r;
      const int db = it & 1;
      GEMM_DBG_T0();
      mbar_wait(bar(kBarDFull + db), (uint32_t)(it >> 1) & 1u);
      GEMM_DBG_ADD(0);
      tc_fence_after();
      const bool row_ok = n < a.w.rows;
      for (int c0 = 0; c0 < n_tok; c0 += 32) {
        uint32_t v[32];
        tmem_ld32(t_lane + (uint32_t)(db * kGemmTokTile + c0), v);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        if (W13 && a.act == NF4MOE_ACT_SWIGLU_OAI) {
          // rows 2j / 2j+1 (gate / up) sit in adjacent lanes.  Per column pair (i, i+1) the two lanes swap one value, after
          // which the even lane owns (gate, up) of column i and the odd lane of column i+1: 16 shuffles, 16 activations
          // and 16 stores per lane and chunk, all independent so they pipeline.
          __half* h = reinterpret_cast<__half*>(a.out);
          const int I = a.w.rows >> 1;
          const bool odd = lane & 1;
          float gt[16], up[16];
#pragma unroll
          for (int pr = 0; pr < 16; ++pr) {
            const float v0 = __half2float(__float2half_rn(__uint_as_float(v[2 * pr])));      // fp16 storage of the product
            const float v1 = __half2float(__float2half_rn(__uint_as_float(v[2 * pr + 1])));
            const float got = __shfl_xor_sync(0xffffffffu, odd ? v0 : v1, 1);
            gt[pr] = odd ? got : v0;
            up[pr] = odd ? v1 : got;
          }
          __half* dst = h + (long long)(row0 + c0 + (odd ? 1 : 0)) * I + (n >> 1);
#pragma unroll
          for (int pr = 0; pr < 16; ++pr) {
            const __half res = __float2half_rn(swiglu_oai(gt[pr], up[pr]));
            if (row_ok && c0 + 2 * pr + (odd ? 1 : 0) < n_tok) dst[(long long)(2 * pr) * I] = res;
          }
        } else if (W13) {
          __half* raw = reinterpret_cast<__half*>(a.out);
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (row_ok && c0 + i < n_tok) raw[(long long)(row0 + c0 + i) * a.w.rows + n] = __float2half_rn(__uint_as_float(v[i]));
        } else {
          __half* y = reinterpret_cast<__half*>(a.out);
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (row_ok && c0 + i < n_tok)
              y[(long long)__ldg(a.sorted_pairs + row0 + c0 + i) * a.w.rows + n] = __float2half_rn(__uint_as_float(v[i]));
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(kBarDEmpty + db));
      GEMM_DBG_ADD(1);
    }
    if (dbg_on && warp == 12) for (int i = 0; i < 6; ++i) a.dbg[40 + i] = dbg_acc[i];
  }

  // ---- teardown ----
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(kTmemCols) : "memory");
  }
}

// ---- host side ------------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) == cudaSuccess &&
        qr == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// 2-D row-major tensor [rows, inner] of `elt`-byte elements, box [box_rows, box_inner], 128B swizzle, OOB reads as zero
inline bool make_map_2d(CUtensorMap* m, const void* ptr, CUtensorMapDataType dt, int elt, uint64_t inner, uint64_t rows,
                        uint32_t box_inner, uint32_t box_rows) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return false;
  const cuuint64_t dims[2] = {inner, rows};
  const cuuint64_t strides[1] = {inner * (uint64_t)elt};
  const cuuint32_t box[2] = {box_inner, box_rows};
  const cuuint32_t es[2] = {1, 1};
  return fn(m, dt, 2, const_cast<void*>(ptr), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
            CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

inline long long*& gemm_debug_buffer() {
  static long long* p = nullptr;
  return p;
}

inline int gemm_tile_tokens(int T, int k, int E) {
  // small batches: 32-token tiles keep the activation stages at 4 KB; otherwise the full 256-column accumulator
  return ((long long)T * k <= 32ll * E) ? 32 : kGemmTokTile;
}

template <bool W13>
int gemm_launch(const void* act_in, const nf4moe_qweight* w, const int32_t* sorted_pairs, const int32_t* meta,
                const int32_t* tile_list, int T, int k, int E, int act, void* out, void* stream, char* err, size_t err_len,
                int* launches) {
  const char* name = W13 ? "gemm_w13" : "gemm_w2";
  auto fail = [&](int code, const char* msg) { snprintf(err, err_len, "%s: %s", name, msg); return code; };
  if (!act_in || !w || !w->packed || !meta || !tile_list || !out || T < 0 || k <= 0 || E <= 0) return fail(NF4MOE_EINVAL, "bad arguments");
  if (!W13 && !sorted_pairs) return fail(NF4MOE_EINVAL, "sorted_pairs required");
  if (w->rows <= 0 || w->cols <= 0 || w->cols % kBlock) return fail(NF4MOE_EINVAL, "cols must be a multiple of 64");
  if (W13 && (w->rows & 1)) return fail(NF4MOE_EINVAL, "w13 rows must be even");
  if (!w->absmax_f32 && !(w->absmax_u8 && w->absmax2 && w->code2 && w->offset)) return fail(NF4MOE_EINVAL, "absmax missing");
  if (w->rows % 16 || w->cols > 4096) return fail(NF4MOE_EINVAL, "the tcgen05 regime needs rows % 16 == 0 and cols <= 4096");
  if ((reinterpret_cast<uintptr_t>(act_in) & 15u) || (reinterpret_cast<uintptr_t>(w->packed) & 15u) ||
      (reinterpret_cast<uintptr_t>(w->absmax_f32) & 15u) || (reinterpret_cast<uintptr_t>(w->absmax_u8) & 15u) ||
      (!w->absmax_f32 && ((reinterpret_cast<uintptr_t>(w->code2) & 15u) || (w->code2_stride & 3))))
    return fail(NF4MOE_EALIGN, "activations, packed weights, absmax and code2 must be 16-byte aligned");
  if (W13 && act != NF4MOE_ACT_SWIGLU_OAI && act != NF4MOE_ACT_CHUNK_SILU) return fail(NF4MOE_EINVAL, "bad activation");
  if (T == 0) return 0;
  int dev = 0, sms = 0, cc = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaDeviceGetAttribute(&cc, cudaDevAttrComputeCapabilityMajor, dev);
  if (cc != 10) return fail(NF4MOE_EARCH, "needs an sm_100 device");
  GemmArgs a;
  a.w = make_qw(*w);
  a.out = out;
  a.sorted_pairs = sorted_pairs;
  a.meta = meta;
  a.tile_list = tile_list;
  a.act = act;
  a.tile_tokens = gemm_tile_tokens(T, k, E);
  a.dbg = gemm_debug_buffer();
  CUtensorMap map_w, map_x;
  if (!make_map_2d(&map_w, w->packed, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, (uint64_t)w->cols / 2, (uint64_t)E * w->rows, 128, 128))
    return fail(NF4MOE_EINVAL, "cuTensorMapEncodeTiled failed for the packed weights");
  if (!make_map_2d(&map_x, act_in, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, (uint64_t)w->cols, (uint64_t)T * k, kBlock, kGemmBoxRows))
    return fail(NF4MOE_EINVAL, "cuTensorMapEncodeTiled failed for the activations");
  static thread_local bool attr_set = false;
  auto kern = gemm_kernel<W13>;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kGemmSmemBytes);
    if (e != cudaSuccess) { snprintf(err, err_len, "%s: cudaFuncSetAttribute: %s", name, cudaGetErrorString(e)); return (int)e; }
    attr_set = true;
  }
  kern<<<sms, kGemmThreads, kGemmSmemBytes, (cudaStream_t)stream>>>(a, map_w, map_x);
  ++*launches;
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { snprintf(err, err_len, "%s: %s", name, cudaGetErrorString(e)); return (int)e; }
  return 0;
}

inline int gemm_w13_entry(const void* x_sorted, const nf4moe_qweight* w13, const int32_t*, const int32_t* meta,
                          const int32_t* tile_list, int T, int k, int E, int act, void* h_sorted, void* stream, char* err,
                          size_t err_len, int* launches) {
  return gemm_launch<true>(x_sorted, w13, nullptr, meta, tile_list, T, k, E, act, h_sorted, stream, err, err_len, launches);
}
inline int gemm_w2_entry(const void* h_sorted, const nf4moe_qweight* w2, const int32_t*, const int32_t* sorted_pairs,
                         const int32_t* meta, const int32_t* tile_list, int T, int k, int E, void* y, void* stream, char* err,
                         size_t err_len, int* launches) {
  return gemm_launch<false>(h_sorted, w2, sorted_pairs, meta, tile_list, T, k, E, 0, y, stream, err, err_len, launches);
}

}  // namespace nf4moe
