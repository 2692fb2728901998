// Decode-regime grouped GEMM on tcgen05: the HBM-bound kernel for 8 < T <= ~300 tokens.
//
// Weights = A operand in TENSOR MEMORY (one dequant thread per weight row, so absmax is a per-thread scalar), activations =
// B operand in shared memory, accumulators in TMEM for the whole K.  The pipeline is built for streaming 12.4 MB of packed
// NF4 per expert at HBM speed rather than for tensor throughput:
//   * an item = 128 weight rows of one expert x all K, for the expert's <= 32 routed tokens (MMA N = 16 or 32);
//   * a STAGE = 4 k-blocks: one 16 KB TMA box of raw NF4 [128 rows x 128 B, 128B swizzle] + the matching activation
//     tiles [4 x N x 64] in the canonical K-major UMMA layout, all on one mbarrier; 3-6 stages in flight per SM;
//   * dequant warps in kDecGroups groups; stage cs of the CTA belongs to group cs % kDecGroups.  Inside a group a warp is
//     (TMEM lane quarter q, share hh of the stage's 4 k-blocks): with kDecSplit = 1 a warp decodes all four k-blocks of
//     its 32 rows (8 LDS.128, 4 absmax, 128 x (PRMT, LDS, HMUL2), 4 x tcgen05.st, one wait in, two arrivals out), so
//     the per-stage bookkeeping is paid once per 128 packed bytes per thread;
//   * operand modes (template FAST): FAST = byte -> half2 table look-up x fp16(absmax) (one HMUL2 per packed byte);
//     EXACT = byte -> two fp32 codes, fp32 multiply, ONE rounding to fp16: bit-identical to bnb dequantize_4bit;
//   * nested (double-quant) absmax in FAST mode: code2[u8] comes from a per-lane copy of the 256-entry table parked in the
//     unused upper halves of the decode table's 256-byte rows (bank-conflict-free; 1.7 M of the 13.5 M shared-memory
//     wavefronts per launch of the round-1 kernel were bank conflicts of this look-up);
//   * one MMA-issuing thread: per stage ONE wait, <= 16 tcgen05.mma (M 128, N 16/32, K 16), one commit (+1 per item);
//   * four epilogue warps read the finished accumulator (double-buffered) while the next item streams: gpt-oss
//     clamped SwiGLU on adjacent lanes for w13 (+ optional fused up-projection bias), scatter to (token, slot) order for w2.
// The shared-memory plan (two absmax stages, the activation ring, the raw ring around the 64 KB-aligned decode table) is
// computed by dec_plan() -- the same function on the host (dec_launch refuses a shape that does not fit) and on the device
// (a plan that does not fit there raises bit 1 of the status word instead of silently doing nothing).
// Replaces bnb Linear4bit on single rows, P/unsloth_to_vllm_converter.py:309-318, for all routed pairs at once.
#pragma once
#include "gemm_tc.cuh"

namespace nf4moe {

#ifndef NF4MOE_DEC_GROUPS
#define NF4MOE_DEC_GROUPS 3
#endif
#ifndef NF4MOE_DEC_SPLIT
#define NF4MOE_DEC_SPLIT 1
#endif
#ifndef NF4MOE_DEC_ARITH
#define NF4MOE_DEC_ARITH 0
#endif
#ifndef NF4MOE_DEC_PRE
#define NF4MOE_DEC_PRE 1
#endif
constexpr int kDecArith = NF4MOE_DEC_ARITH;       // packed words (of 8 per k-block and thread) decoded without the shared-memory table
constexpr int kDecGroups = NF4MOE_DEC_GROUPS;     // dequant groups; stage cs belongs to group cs % kDecGroups
constexpr int kDecSplit = NF4MOE_DEC_SPLIT;       // warps sharing one (stage, lane quarter): 1 or 2
constexpr int kDecKpw = 4 / kDecSplit;            // k-blocks a warp decodes per stage
constexpr int kDecStageWarps = 4 * kDecSplit;     // warps that consume one stage
constexpr int kDecDqWarps = kDecStageWarps * kDecGroups;
static_assert(kDecSplit == 1 || kDecSplit == 2, "NF4MOE_DEC_SPLIT must be 1 or 2");
static_assert(kDecDqWarps >= 8 && kDecDqWarps <= 24, "8..24 dequant warps");
constexpr int kDecThreads = (kDecDqWarps + 4 + 4) * 32;   // dequant, epilogue warpgroup, control warpgroup
constexpr int kDecWarpProd = kDecDqWarps + 4, kDecWarpMma = kDecDqWarps + 5, kDecWarpAct = kDecDqWarps + 6;
constexpr int kDecMaxRaw = 8;                              // raw-NF4 stages (16 KB each) in the shared-memory ring
constexpr int kDecMaxB = 6;                                // activation stages (their own ring: they live until the MMAs retire)
constexpr int kDecSlots = 3;                               // A-operand slots in TMEM: one stage = 4 k-blocks = 128 columns each
constexpr int kDecAccCols = 32;                            // per accumulator (two of them); 64 in a launch with 64-row token tiles
constexpr uint32_t kDecRawBytes = 128 * 128;               // raw NF4 of one stage: 128 rows x 4 k-blocks x 32 B
constexpr uint32_t kDecSmemBytes = 232448 - 1024;
// Registers per thread.  The CTA is launched with kDecRegsLaunch per thread (what __launch_bounds__(kDecThreads, 1) lets ptxas
// use, and with setmaxnreg in the kernel it takes all of it); setmaxnreg then moves registers from the control warpgroup to
// the dequant warps.  The pool is what the CTA was LAUNCHED with -- an increase beyond what the decreases released blocks
// forever -- so the dequant figure is derived from the budget, and dec_launch() checks the budget against the kernel's
// real register count (cudaFuncGetAttributes) before the first launch instead of trusting this arithmetic.
constexpr int kDecRegsLaunch = (65536 / kDecThreads) / 8 * 8 > 255 ? 248 : (65536 / kDecThreads) / 8 * 8;
#ifndef NF4MOE_DEC_REGS_CTL
#define NF4MOE_DEC_REGS_CTL 32
#endif
#ifndef NF4MOE_DEC_REGS_EPI
#define NF4MOE_DEC_REGS_EPI 64
#endif
constexpr int kDecRegsCtl = NF4MOE_DEC_REGS_CTL;
// The epilogue warpgroup gives registers away too: it runs once per item (12 stages), so what it spills costs nothing, and the
// dequant warps at 128 registers keep their addresses and ring state live instead of rebuilding them every stage.
constexpr int kDecRegsEpi = kDecRegsLaunch < NF4MOE_DEC_REGS_EPI ? kDecRegsLaunch : NF4MOE_DEC_REGS_EPI;
constexpr int kDecRegsDqBudget = (kDecRegsLaunch * kDecThreads - 128 * kDecRegsEpi - 128 * kDecRegsCtl) / (32 * kDecDqWarps) / 8 * 8;
constexpr int kDecRegsDq = kDecRegsDqBudget > 232 ? 232 : kDecRegsDqBudget;
static_assert(kDecRegsDq >= kDecRegsLaunch, "the dequant warps never give registers away");
static_assert(32 * kDecDqWarps * kDecRegsDq + 128 * kDecRegsEpi + 128 * kDecRegsCtl <= kDecRegsLaunch * kDecThreads,
              "setmaxnreg budget: increases must fit in what the decreases release");

__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, int c0, int c1, int c2, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
      ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(bar)
      : "memory");
}

// Activations [rows, cols] fp16 viewed as a 3-D tensor (64 inner, rows, cols / 64 k-blocks) so ONE box [64, box_rows, 4] lands the
// four K-major 128B-swizzled UMMA tiles of a stage back to back (k-blocks past the end read as zero).
inline bool make_map_x3d(CUtensorMap* m, const void* ptr, uint64_t cols, uint64_t rows, uint32_t box_rows) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return false;
  const cuuint64_t dims[3] = {(cuuint64_t)kBlock, rows, cols / kBlock};
  const cuuint64_t strides[2] = {cols * 2, (cuuint64_t)kBlock * 2};
  const cuuint32_t box[3] = {(cuuint32_t)kBlock, box_rows, 4};
  const cuuint32_t es[3] = {1, 1, 1};
  return fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<void*>(ptr), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}


// ---- shared-memory plan -------------------------------------------------------------------------------------------------
// Two regions around the 64 KB-aligned decode table: [align1024(base), lut) and [lut + 64 KB, base + kDecSmemBytes).  Wanted:
// two absmax stages (the item's [128 rows x nkb] absmax: fp32, or u8 (+ the expert's 1 KB code2 table unless it lives in the
// table's holes)), ONE contiguous activation ring of n_b stages (4 x N x 128 B each) and as many 16 KB raw stages as still fit.
struct DecPlan {
  uint32_t m_code2, m_bytes, b_stage;
  int n_b, n_raw;
  uint32_t mst[2], b0, raw[kDecMaxRaw];
};

__host__ __device__ inline uint32_t dec_absmax_stage_bytes(int nkb, bool nested, bool stage_code2, uint32_t* code2_off) {
  const uint32_t m_tile = (uint32_t)(kGemmRowTile * nkb) * (nested ? 1u : 4u);
  const uint32_t m_code2 = (m_tile + 15u) & ~15u;
  if (code2_off) *code2_off = m_code2;
  return (m_code2 + (nested && stage_code2 ? 1024u : 0u) + 1023u) & ~1023u;
}

__host__ __device__ inline bool dec_plan_try(uint32_t base, uint32_t lut, uint32_t m_bytes, uint32_t b_stage, int n_b, DecPlan* p) {
  uint32_t cur[2] = {(base + 1023u) & ~1023u, lut + kGemmLutBytes};
  const uint32_t hi[2] = {lut, base + kDecSmemBytes};
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
  for (int i = 0; i < 2; ++i) { p->mst[i] = place(m_bytes); if (!p->mst[i]) return false; }
  p->b0 = place((uint32_t)n_b * b_stage);
  if (!p->b0) return false;
  p->n_b = n_b;
  p->n_raw = 0;
  for (int i = 0; i < kDecMaxRaw; ++i) { const uint32_t at = place(kDecRawBytes); if (!at) break; p->raw[p->n_raw++] = at; }
  // A raw stage must always be consumed by the SAME dequant group (stage cs -> group cs % kDecGroups, raw slot cs % n_raw): an
  // mbarrier parity wait is only sound when the waiter has itself seen the slot's previous phase complete; a group that met
  // the slot for the first time at its second use took "nothing has landed yet" for "landed" (round 2: wrong results and
  // faults with 2 or 3 raw stages under 4 groups).  So: a multiple of the group count.
  p->n_raw = p->n_raw / kDecGroups * kDecGroups;
  return p->n_raw >= kDecGroups;
}

// bmax = largest MMA N of the launch (16, 32 or -- 64-row token tiles, nested absmax only: the flat form does not fit -- 64).  Takes the deepest activation ring that still leaves one raw stage per
// dequant group (measured, K = 2880 nested, 64 tokens: 3 raw stages 80.4 us vs 6 raw stages 79.7 us per w13 launch, but 3
// activation stages 91.6 us vs 6 stages 79.7 us: the activation ring is what has to be deep, its slots are only released by
// retiring MMAs).  The activation ring is never shorter
// than the group count either: a dequant warp waits on Done[stage - 3] knowing only that Done[stage - 3 - groups] completed,
// which covers the barrier's previous phase (stage - 3 - n_b) only if n_b >= groups.  Returns false when nothing fits (K too
// large for this pipeline).
__host__ __device__ inline bool dec_plan(uint32_t base, uint32_t lut, int nkb, bool nested, bool stage_code2, int bmax, DecPlan* p) {
  p->m_bytes = dec_absmax_stage_bytes(nkb, nested, stage_code2, &p->m_code2);
  p->b_stage = 4u * (uint32_t)bmax * 128u;
  const int want_b = bmax == 16 ? kDecMaxB : bmax == 32 ? 5 : 3;
  for (int n_b = want_b; n_b >= (kDecGroups > 3 ? kDecGroups : 3); --n_b) {
    DecPlan t = *p;
    if (!dec_plan_try(base, lut, p->m_bytes, p->b_stage, n_b, &t)) continue;
    *p = t;   // dec_plan_try only succeeds with n_raw >= kDecGroups
    return true;
  }
  p->n_raw = 0;
  return false;
}

// One launch serves the up projection (side 0: w13, SwiGLU epilogue, output h in sorted-row order), the down projection
// (side 1: w2, output scattered to (token, slot) rows), or BOTH (template SIDES = 2): the persistent CTAs walk the w13 items
// and carry straight on into the w2 items, whose activations are the h rows other CTAs have just written.  A w2 item of token
// tile t may start once all row tiles of w13 for tile t are stored: the w13 epilogue counts them in tile_list[4 t + 3] (zeroed
// by the bucket kernel), the activation producer of the w2 item spins on that word.  All w13 items precede all w2 items in
// every CTA's sequence and every CTA is resident (one per SM), so the wait cannot deadlock; the weights of the first w2 items
// stream in while the last w13 items are still being multiplied, which is the point: two separate launches pay the pipeline
// fill, the drain and the launch gap twice (7.6 us from CTA entry to the first MMA + ~1 us drain, measured, of a 79 + 46 us
// pair).
struct DecSide {
  QW w;
  void* out;
};
struct DecArgs {
  DecSide s[2];
  const int32_t* sorted_pairs;
  const int32_t* meta;
  int32_t* tile_list;   // [4 t + 3] is the tile's count of finished w13 row tiles when SIDES = 2
  int act;
  const __half* bias;   // w13 only, optional: fp16 [E, rows] up-projection bias, added (fp16) before the fused SwiGLU
  int32_t* status;      // optional device word: bit 1 is raised when the shared-memory plan does not fit (nothing is computed then)
  long long* dbg;       // NF4MOE_EXPERIMENTS builds only
  int b3d;              // the activation maps are the 3-D form (one TMA per stage) rather than 2-D (one per k-block)
  int hole_code2;       // FAST + nested + one code2 table for all experts: the table sits, per lane, in the decode table's holes
  // The shared-memory plans, computed on the host for the two places the dynamic shared memory can start (1 KB / 2 KB into the
  // CTA's window) x the two MMA widths (16 / 32): a lone device thread needs ~1.5 us for dec_plan(), on the critical path of
  // every launch.  The kernel falls back to computing it only if its window starts somewhere else.
  uint32_t plan_base[2];
  DecPlan plan[6];      // [base index * 3 + width class (16 / 32 / 64)]; n_raw == 0: that width does not fit
  int max_tile;         // token-tile size the tile list was built with (32 or 64)
};

// barrier slots
//   RFull[i]  <- TMA bytes of raw stage i: waited on by the dequant warps that own the stage
//   BFull[c]  <- TMA bytes of activation stage c (ring of n_b): waited on by the MMA thread.  (Round 2 first counted them on
//                RFull, so that the MMA thread had a single wait; but then a stage's dequant could not start before its
//                activations had landed, and those are only requested once the MMAs n_b stages back retire.)
//   REmpty[i] <- the stage's dequant warps, as soon as their LDS.128 are done (the stage refills while they decode)
//   AFull[s]  <- the stage's dequant warps, A operand of the stage written to TMEM slot s
//   Done[c]   <- ONE tcgen05.commit per stage (ring of n_b): frees A slot c % 3 for the dequant warps of stage + 3
//                and activation stage c for the producer of stage + n_b
constexpr int kDbRFull = 0, kDbREmpty = kDbRFull + kDecMaxRaw, kDbDone = kDbREmpty + kDecMaxRaw,
              kDbBFull = kDbDone + kDecMaxB, kDbAFull = kDbBFull + kDecMaxB, kDbDFull = kDbAFull + kDecSlots, kDbDEmpty = kDbDFull + 2,
              kDbMFull = kDbDEmpty + 2, kDbMEmpty = kDbMFull + 2, kNumDecBars = kDbMEmpty + 2;

__device__ __forceinline__ void ring_advance(int& idx, uint32_t& phase, int step, int n) {
  idx += step;
  while (idx >= n) { idx -= n; phase ^= 1u; }
}

// Warp roles (one persistent CTA per SM).  cs = the CTA's running stage counter (12 per item for K = 2880); the rings follow
// it: raw stage cs % n_raw, activation stage cs % n_b, A slot cs % 3.
//   warps 0..kDecDqWarps-1   dequant: q = warp & 3 (TMEM lane quarter = weight rows 32q..32q+31), hh = share of the stage's
//                            k-blocks, pp = group (stages cs = pp mod kDecGroups)
//   next 4 warps             epilogue
//   then                     TMA producer (absmax per item; raw NF4 per stage), MMA issuer (+ TMEM allocator), TMA producer
//                            of the activation stages, one idle warp
// SIDES: 0 = w13 only, 1 = w2 only, 2 = w13 then w2 (see DecSide).  The tensor maps of the side that is not served are unused.
template <int SIDES, bool FAST, bool DBG>
__global__ void __launch_bounds__(kDecThreads, 1) dec_kernel(const DecArgs a, const __grid_constant__ CUtensorMap map_w13,
                                                            const __grid_constant__ CUtensorMap map_x16,
                                                            const __grid_constant__ CUtensorMap map_x32,
                                                            const __grid_constant__ CUtensorMap map_w2,
                                                            const __grid_constant__ CUtensorMap map_h16,
                                                            const __grid_constant__ CUtensorMap map_h32,
                                                            const __grid_constant__ CUtensorMap map_x64,
                                                            const __grid_constant__ CUtensorMap map_h64) {
  extern __shared__ __align__(1024) uint8_t dyn_smem[];
  __shared__ __align__(8) uint64_t s_bar[kNumDecBars];
  __shared__ uint32_t s_tmem;
  __shared__ DecPlan s_plan;
  grid_dep_launch();
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // DBG builds: a timeline (globaltimer, ns) of CTA 0 (slots 32..) and of the last CTA (slots 48..) -- see scripts/dec_roles.py
  const bool tl_on = DBG && a.dbg != nullptr && lane == 0 && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1);
  long long* tl = DBG && a.dbg != nullptr ? a.dbg + (blockIdx.x == 0 ? 32 : 48) : nullptr;
  auto stamp = [&](int i) {
    if (DBG && tl_on) { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); tl[i] = (long long)t; }
  };
  if (tid == 0) stamp(0);
  constexpr int kSide0 = SIDES == 1 ? 1 : 0;            // the first (or only) side of this launch
  const bool nested = a.s[kSide0].w.absmax_f32 == nullptr;   // both sides of a launch use the same absmax form
  const bool hole = FAST && a.hole_code2 != 0;
  const uint32_t base = smem_u32(dyn_smem);
  const uint32_t lut = pin((base + kGemmLutBytes - 1u) & ~(kGemmLutBytes - 1u));
  const uint32_t bar0 = pin(smem_u32(s_bar));
  auto bar = [&](int i) -> uint32_t { return bar0 + 8u * (uint32_t)i; };

  // ---- prologue that depends on nothing the previous kernel wrote: decode table (weights are constants), TMEM allocation.
  // Under programmatic dependent launch it overlaps the previous kernel's tail; everything below the wait may read bucket /
  // GEMM outputs ----
  {
    constexpr int kFill = (256 * 32 + kDecThreads - 1) / kDecThreads;
    float c2[kFill];
    if (hole) {   // all of a thread's code2 loads in flight at once (one dependent round trip per row cost 2 us per launch)
#pragma unroll
      for (int j = 0; j < kFill; ++j) {
        const int i = tid + j * kDecThreads;
        c2[j] = i < 256 * 32 ? __ldg(a.s[kSide0].w.code2 + (i >> 5)) : 0.f;   // SIDES = 2: both sides share the table (host-checked)
      }
    }
#pragma unroll
    for (int j = 0; j < kFill; ++j) {
      const int i = tid + j * kDecThreads;
      if (i >= 256 * 32) break;
      const int b = i >> 5, l = i & 31;
      if (FAST) {
        sts32(lut + (uint32_t)b * 256u + (uint32_t)l * 4u, pack_half2(kNF4[b >> 4], kNF4[b & 15]));
        if (hole) sts32(lut + (uint32_t)b * 256u + 128u + (uint32_t)l * 4u, __float_as_uint(c2[j]));
      } else {
        const uint32_t p = lut + (uint32_t)b * 256u + (uint32_t)l * 8u;
        sts32(p, __float_as_uint(kNF4[b >> 4]));        // element 2j   (high nibble)
        sts32(p + 4u, __float_as_uint(kNF4[b & 15]));   // element 2j+1 (low nibble)
      }
    }
  }
  if (warp == kDecWarpMma) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(kTmemCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (lane == 0 && (warp == kDecWarpProd || warp == kDecWarpAct)) {   // descriptor fetches off the first TMA's critical path
    auto pf = [](const CUtensorMap* m) { asm volatile("prefetch.tensormap [%0];" ::"l"(m) : "memory"); };
    if (warp == kDecWarpProd) {
      if (SIDES != 1) pf(&map_w13);
      if (SIDES != 0) pf(&map_w2);
    } else {
      if (SIDES != 1) { pf(&map_x16); pf(&map_x32); pf(&map_x64); }
      if (SIDES != 0) { pf(&map_h16); pf(&map_h32); pf(&map_h64); }
    }
  }
  if (tid == kDecThreads - 1) {   // the barriers depend on nothing either (the idle control warp's lane: off everybody's path)
    for (int i = 0; i < kDecMaxRaw; ++i) { mbar_init(bar(kDbRFull + i), 1); mbar_init(bar(kDbREmpty + i), kDecStageWarps); }
    for (int i = 0; i < kDecMaxB; ++i) { mbar_init(bar(kDbDone + i), 1); mbar_init(bar(kDbBFull + i), 1); }
    for (int i = 0; i < kDecSlots; ++i) mbar_init(bar(kDbAFull + i), kDecStageWarps);
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar(kDbDFull + i), 1); mbar_init(bar(kDbDEmpty + i), 4);
      mbar_init(bar(kDbMFull + i), 1); mbar_init(bar(kDbMEmpty + i), kDecDqWarps);
    }
    fence_barrier_init();
  }
  if (tid == 0) stamp(1);
  grid_dep_wait();
  if (tid == 0) stamp(2);
  // largest MMA N of this launch (meta[3] = largest per-expert count; tiles never exceed max_tile rows)
  const int bmax = a.meta[3] <= 16 ? 16 : (a.meta[3] <= 32 || a.max_tile <= 32) ? 32 : 64;
  const uint32_t acc_cols = bmax == 64 ? 64u : (uint32_t)kDecAccCols;
  // ---- one-time setup ----
  if (tid == 0) {
    const int bi = base == a.plan_base[0] ? 0 : base == a.plan_base[1] ? 1 : -1;
    if (bi >= 0) s_plan = a.plan[bi * 3 + (bmax == 16 ? 0 : bmax == 32 ? 1 : 2)];
    else dec_plan(base, lut, SIDES == 2 ? max(a.s[0].w.nkb, a.s[1].w.nkb) : a.s[kSide0].w.nkb, nested, !hole, bmax, &s_plan);
    // the tile list was built (nf4moe_bucket records it in meta[5]) for larger token tiles than this launch was planned for: refuse
    // the whole launch -- an MMA N clamped to max_tile would silently drop the tiles' remaining rows
    if (a.meta[5] > a.max_tile) s_plan.n_raw = 0;
    if (s_plan.n_raw == 0 && a.meta[2] > 0 && a.status) atomicOr(a.status, 2);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (tid == 0) stamp(3);
  const uint32_t tmem = s_tmem;
  const int n_raw = s_plan.n_raw, n_b = s_plan.n_b;
  const uint32_t b_ring = s_plan.b0, b_stage = s_plan.b_stage;
  // Items: (token tile, 128-row weight tile) of the first side, tile-major; with SIDES = 2 the second side's items follow.
  const int nrt0 = (a.s[kSide0].w.rows + kGemmRowTile - 1) / kGemmRowTile;
  const int nrt1 = SIDES == 2 ? (a.s[1].w.rows + kGemmRowTile - 1) / kGemmRowTile : 0;
  const int n_items0 = n_raw > 0 ? a.meta[2] * nrt0 : 0;
  const int n_items = n_items0 + (n_raw > 0 ? a.meta[2] * nrt1 : 0);
  auto item_of = [&](int item, int& side, int& tile, int& rt) {
    if (SIDES == 2 && item >= n_items0) {
      side = 1;
      item -= n_items0;
      tile = item / nrt1;
      rt = item - tile * nrt1;
    } else {
      side = kSide0;
      tile = item / nrt0;
      rt = item - tile * nrt0;
    }
  };
  const uint32_t acol0 = 2u * acc_cols;   // TMEM: two accumulators (32 or 64 columns each), the three 128-column A slots after them
  const bool dbg_on = DBG && a.dbg != nullptr && blockIdx.x == 0 && lane == 0;   // DBG = false: every counter below folds away
  long long dbg_acc[6] = {0, 0, 0, 0, 0, 0};

  if (warp < kDecDqWarps) {
    // =============================== dequant: NF4 -> fp16 A operand in TMEM ===============================
    if constexpr (kDecRegsDq > kDecRegsLaunch) set_max_regs_inc<kDecRegsDq>();
    const int q = warp & 3, hh = (warp >> 2) % kDecSplit, pp = (warp >> 2) / kDecSplit;   // pp = dequant group
    const int r = q * 32 + lane;
    const uint32_t lut_lane = pin(lut | ((uint32_t)lane << (FAST ? 2 : 3)));
    const uint32_t hole_lane = pin(lut | 128u | ((uint32_t)lane << 2));
    const uint32_t t_lane = tmem + ((uint32_t)(q * 32) << 16) + acol0;
    const uint32_t rfull0 = bar(kDbRFull), rempty0 = bar(kDbREmpty), afull0 = bar(kDbAFull), done0 = bar(kDbDone);
    // this lane's 16-byte chunks of a raw stage row (128B swizzle: chunk c of row r sits at chunk c ^ (r & 7)).  Stages are
    // 1 KB-aligned and a row is 128 B, so chunk j of the stage at `wst` is at (wst + c_row) ^ (j << 4): one add per stage and one
    // LOP3 with an immediate per chunk (round 2: ptxas rebuilt all eight addresses from %tid every stage, 38 instructions)
    const uint32_t c_row = pin(((uint32_t)r * 128u) | ((uint32_t)(r & 7) << 4));
    // this warp owns stages cs = pp (mod kDecGroups).  Running ring positions of its NEXT stage: raw stage sw, A slot `slot`,
    // and the Done barrier of stage cs - 3 (index dn, parity pdn; "stage -3 .. -1" = the phase before the first one, which a
    // fresh barrier reports complete)
    int sw = 0, slot = pp % kDecSlots, dn = 0, st = pp, it = 0;
    uint32_t phw = 0, pdn = 1;
    ring_advance(sw, phw, pp, max(n_raw, 1));
    ring_advance(dn, pdn, pp + n_b - 3, max(n_b, 1));
    for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++it) {
      int side, tile, rt;
      item_of(item, side, tile, rt);
      const QW& w = a.s[side].w;
      const int nkb = w.nkb, nst = (nkb + 3) >> 2;          // stages of this item (the last one may be partly padding)
      const int e = a.tile_list[4 * tile];
      const int n = rt * kGemmRowTile + r;
      const int mb = it & 1;
      // nested absmax: code2[u8] * absmax2[blk >> 8] + offset; this row's blocks touch at most two absmax2 entries
      float a2_lo = 0.f, a2_hi = 0.f, off_e = 0.f;
      int blk_hi = 0;                                      // first k-block of this row that belongs to the next absmax2 entry (1..256)
      if (nested) {
        const long long blk0 = (long long)min(n, w.rows - 1) * nkb;
        const long long i2 = blk0 >> 8;
        a2_lo = __ldg(w.absmax2 + (long long)e * w.n2 + i2);
        a2_hi = __ldg(w.absmax2 + (long long)e * w.n2 + min(i2 + 1, w.n2 - 1));
        off_e = __ldg(w.offset + e);
        blk_hi = (int)(((i2 + 1) << 8) - blk0);
      }
      GEMM_DBG_T0();
      mbar_wait(bar(kDbMFull + mb), (uint32_t)(it >> 1) & 1u);
      GEMM_DBG_ADD(0);
      const uint32_t m_st = pin(s_plan.mst[mb]);
      const uint32_t m_row = m_st + (uint32_t)(r * nkb) * (nested ? 1u : 4u);
      const uint32_t m_code2 = ((uint32_t)(kGemmRowTile * nkb) * (nested ? 1u : 4u) + 15u) & ~15u;   // staged code2 table, if any
      for (; st < nst; st += kDecGroups) {
        mbar_wait(rfull0 + 8u * (uint32_t)sw, phw);
        if (DBG && warp == 0 && it == 0 && st == 0) stamp(4);
        GEMM_DBG_ADD(1);
        const uint32_t wst = pin(s_plan.raw[sw]) + c_row;
        const int kb0 = st * 4 + hh * kDecKpw;
        uint4 v[2 * kDecKpw];
#pragma unroll
        for (int j = 0; j < 2 * kDecKpw; ++j) v[j] = lds128(wst ^ ((uint32_t)(hh * 2 * kDecKpw + j) << 4));
        float am[kDecKpw];
#pragma unroll
        for (int u = 0; u < kDecKpw; ++u) {
          const int kb = min(kb0 + u, nkb - 1);
          if (nested) {
            const uint32_t c = lds_u8(m_row + (uint32_t)kb);
            const float cv = hole ? lds_f32(hole_lane + (c << 8)) : lds_f32(m_st + m_code2 + 4u * c);
            am[u] = __fadd_rn(__fmul_rn(cv, kb < blk_hi ? a2_lo : a2_hi), off_e);
          } else {
            am[u] = lds_f32(m_row + 4u * (uint32_t)kb);
          }
        }
        // the raw bytes are in registers: hand the stage back now (mbarrier.arrive has release semantics: the loads above
        // are performed before the arrival is visible), so its refill overlaps the decode
        __syncwarp();
        if (lane == 0) mbar_arrive(rempty0 + 8u * (uint32_t)sw);
        const uint32_t t_st = t_lane + (uint32_t)slot * 128u + (uint32_t)(hh * kDecKpw) * 32u;
        // kDecPre k-blocks are decoded BEFORE the wait for the A slot (the MMAs of stage cs - 3).  Decoding two ahead removes the
        // wait (ncu: 12.6 % of the dequant warps' samples spin on Done with one) but is 2 % SLOWER end to end (same-box A/B,
        // profiles/r2_notes.md): the time only moves to the raw-stage wait, the kernel is bound by issue + the shared-memory
        // pipe, not by this dependency.  What did help (2.6 %) is the register split that came with it: 128 per dequant thread.
        constexpr int kDecPre = (kDecKpw >= 2 && NF4MOE_DEC_PRE >= 2) ? 2 : 1;
        uint32_t ra[32], rb[32];
        const bool real0 = kb0 < nkb, real1 = kDecPre > 1 && kb0 + 1 < nkb;
        if (real0) decode_block<FAST, kDecArith>(v[0], v[1], lut_lane, am[0], ra);
        if constexpr (kDecPre > 1) {
          if (real1) decode_block<FAST, kDecArith>(v[2], v[3], lut_lane, am[1], rb);
        }
        GEMM_DBG_ADD(2);
        mbar_wait(done0 + 8u * (uint32_t)dn, pdn);          // the MMAs of stage cs - 3, which read this A slot, are done
        tc_fence_after();
        GEMM_DBG_ADD(3);
        if (real0) tmem_st32(t_st, ra);
        if constexpr (kDecPre > 1) {
          if (real1) tmem_st32(t_st + 32u, rb);
        }
#pragma unroll
        for (int u = kDecPre; u < kDecKpw; ++u) {
          if (kb0 + u < nkb) {
            if (u & 1) { decode_block<FAST, kDecArith>(v[2 * u], v[2 * u + 1], lut_lane, am[u], rb); tmem_st32(t_st + (uint32_t)u * 32u, rb); }
            else { decode_block<FAST, kDecArith>(v[2 * u], v[2 * u + 1], lut_lane, am[u], ra); tmem_st32(t_st + (uint32_t)u * 32u, ra); }
          }
        }
        GEMM_DBG_ADD(4);
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(afull0 + 8u * (uint32_t)slot);
        GEMM_DBG_ADD(5);
        ring_advance(sw, phw, kDecGroups, n_raw);
        slot = (slot + kDecGroups) % kDecSlots;
        ring_advance(dn, pdn, kDecGroups, n_b);
      }
      st -= nst;
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(kDbMEmpty + mb));
    }
    if (dbg_on && (warp == 0 || warp == kDecDqWarps - 3)) for (int i = 0; i < 6; ++i) a.dbg[(warp ? 8 : 0) + i] = dbg_acc[i];
  } else if (warp < kDecDqWarps + 4) {
    // =============================== epilogue: TMEM -> registers -> global, overlapped with the next item ===============
    if constexpr (kDecRegsEpi < kDecRegsLaunch) set_max_regs_dec<kDecRegsEpi>();
    const int q = warp & 3;
    const int r = q * 32 + lane;
    const uint32_t t_lane = tmem + ((uint32_t)(q * 32) << 16);
    int it = 0;
    for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++it) {
      int side, tile, rt;
      item_of(item, side, tile, rt);
      const bool up_side = side == 0;                       // w13: activation epilogue; w2: scatter epilogue
      const int w_rows = a.s[side].w.rows;
      void* const out = a.s[side].out;
      const int row0 = a.tile_list[4 * tile + 1], n_tok = a.tile_list[4 * tile + 2];
      const int n = rt * kGemmRowTile + r;
      const int db = it & 1;
      uint32_t bg2 = 0u, bu2 = 0u;   // this lane pair's gate / up bias, broadcast to both halves (zero = none)
      if (up_side && a.bias != nullptr && a.act == NF4MOE_ACT_SWIGLU_OAI && n < w_rows) {
        const __half* be = a.bias + (long long)a.tile_list[4 * tile] * w_rows + (n & ~1);
        const uint32_t g = __half_as_ushort(__ldg(be)), u = __half_as_ushort(__ldg(be + 1));
        bg2 = g | (g << 16);
        bu2 = u | (u << 16);
      }
      mbar_wait(bar(kDbDFull + db), (uint32_t)(it >> 1) & 1u);
      if (DBG && q == 0) stamp(it == 0 ? 6 : 7);          // first / last accumulator finished
      tc_fence_after();
      const bool row_ok = n < w_rows;
      const int n_chunks = n_tok > 32 ? 2 : 1;              // 64-row token tiles: the accumulator is read 32 columns at a time
#pragma unroll 1
      for (int ch = 0; ch < n_chunks; ++ch) {
      const int c0 = ch * 32;
      uint32_t v[32];
      tmem_ld32(t_lane + (uint32_t)db * acc_cols + (uint32_t)c0, v);
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      if (ch == n_chunks - 1) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar(kDbDEmpty + db));   // the accumulator is in registers: hand it back before the stores
      }
      if (up_side && a.act == NF4MOE_ACT_SWIGLU_OAI) {
        // rows 2j / 2j+1 (gate / up) sit in adjacent lanes; products rounded to fp16 first (the reference stores the GEMM
        // result as fp16 before the activation); even lane takes tokens 0..15, odd lane tokens 16..31 of the tile
        __half* h = reinterpret_cast<__half*>(out);
        const int I = w_rows >> 1;
        const bool odd = lane & 1;
        uint32_t p[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) p[i] = pack_half2(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1]));
        uint32_t gt[8], up[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const uint32_t got = __shfl_xor_sync(0xffffffffu, odd ? p[i] : p[8 + i], 1);
          gt[i] = odd ? got : p[i];
          up[i] = odd ? p[8 + i] : got;
        }
        if (a.bias != nullptr) {
#pragma unroll
          for (int i = 0; i < 8; ++i) { gt[i] = hadd2_u32(gt[i], bg2); up[i] = hadd2_u32(up[i], bu2); }
        }
        const int t0 = c0 + (odd ? 16 : 0);
        __half* dst = h + (long long)(row0 + t0) * I + (n >> 1);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          if (row_ok && t0 + 2 * i < n_tok) {
            const uint32_t r2 = swiglu_oai_h2(gt[i], up[i]);
            dst[(long long)(2 * i) * I] = __ushort_as_half((unsigned short)(r2 & 0xFFFFu));
            if (t0 + 2 * i + 1 < n_tok) dst[(long long)(2 * i + 1) * I] = __ushort_as_half((unsigned short)(r2 >> 16));
          }
        }
      } else if (up_side) {
        __half* raw = reinterpret_cast<__half*>(out);
#pragma unroll
        for (int i = 0; i < 32; ++i)
          if (row_ok && c0 + i < n_tok) raw[(long long)(row0 + c0 + i) * w_rows + n] = __float2half_rn(__uint_as_float(v[i]));
      } else {
        __half* y = reinterpret_cast<__half*>(out);
#pragma unroll
        for (int i = 0; i < 32; ++i)
          if (row_ok && c0 + i < n_tok)
            y[(long long)a.sorted_pairs[row0 + c0 + i] * w_rows + n] = __float2half_rn(__uint_as_float(v[i]));
      }
      }   // chunks
      if (SIDES == 2 && up_side) {
        // publish this row tile of h: every epilogue thread's stores are ordered before the count (fence, then the four warps
        // meet on named barrier 1, then ONE release-increment at GPU scope); the w2 items of this token tile wait for all nrt0
        __threadfence();
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (q == 0 && lane == 0) asm volatile("red.release.gpu.global.add.s32 [%0], 1;" ::"l"(a.tile_list + 4 * tile + 3) : "memory");
      }
    }
  } else {
    set_max_regs_dec<kDecRegsCtl>();
    if (warp == kDecWarpProd) {
      // =============================== TMA producer: absmax per item, raw NF4 per stage ===============================
      const uint64_t pol = l2_policy_evict_first();   // weights are read exactly once per step
      const uint32_t am_elt = nested ? 1u : 4u;
      const uint32_t rfull0 = bar(kDbRFull), rempty0 = bar(kDbREmpty);
      const bool stage_code2 = nested && !hole;
      int sw = 0, it = 0;
      uint32_t phw = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++it) {
        GEMM_DBG_T0();
        int side, tile, rt;
        item_of(item, side, tile, rt);
        const QW& w = a.s[side].w;
        const int nkb = w.nkb, nst = (nkb + 3) >> 2;
        const uint8_t* am_base = nested ? w.absmax_u8 : reinterpret_cast<const uint8_t*>(w.absmax_f32);
        const CUtensorMap* map_w = side == 0 ? &map_w13 : &map_w2;
        const uint32_t m_code2 = ((uint32_t)(kGemmRowTile * nkb) * am_elt + 15u) & ~15u;
        const int e = a.tile_list[4 * tile];
        const int c1 = e * w.rows + rt * kGemmRowTile;
        {
          const int mb = it & 1;
          const int rows_here = min(kGemmRowTile, w.rows - rt * kGemmRowTile);
          const uint32_t bytes = (uint32_t)(rows_here * nkb) * am_elt;
          mbar_wait(bar(kDbMEmpty + mb), ((uint32_t)(it >> 1) & 1u) ^ 1u);
          if (elect_one()) {
            mbar_expect_tx(bar(kDbMFull + mb), bytes + (stage_code2 ? 1024u : 0u));
            bulk_g2s(s_plan.mst[mb], am_base + ((long long)e * w.nb + (long long)rt * kGemmRowTile * nkb) * am_elt, bytes,
                     bar(kDbMFull + mb), pol);
            if (stage_code2) bulk_g2s(s_plan.mst[mb] + m_code2, w.code2 + (long long)e * w.code2_stride, 1024u, bar(kDbMFull + mb), pol);
          }
          __syncwarp();
        }
        GEMM_DBG_ADD(0);
        for (int st = 0; st < nst; ++st) {
          mbar_wait(rempty0 + 8u * (uint32_t)sw, phw ^ 1u);          // the dequant warps of the stage n_raw back have read it
          GEMM_DBG_ADD(1);
          if (elect_one()) {
            const uint32_t full = rfull0 + 8u * (uint32_t)sw;
            mbar_expect_tx(full, kDecRawBytes);
            // (a probe that read every stage as one contiguous 16 KB block instead of 128 row segments 1440 B apart changed
            // nothing -- 123.9 vs 123.5 us per layer -- so the weights keep bnb's row-major layout)
            tma_load_2d(s_plan.raw[sw], map_w, st * 128, c1, full);
          }
          __syncwarp();
          GEMM_DBG_ADD(2);
          if (++sw == n_raw) { sw = 0; phw ^= 1u; }
        }
      }
      if (dbg_on) for (int i = 0; i < 6; ++i) a.dbg[24 + i] = dbg_acc[i];
    } else if (warp == kDecWarpAct) {
      // =============================== TMA producer: activations per stage ===============================
      // Runs n_b stages ahead of the MMAs, independent of the raw-weight ring.
      const uint32_t bfull0 = bar(kDbBFull), done0 = bar(kDbDone);
      int sb = 0;
      uint32_t phb = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        int side, tile, rt;
        item_of(item, side, tile, rt);
        const int nkb = a.s[side].w.nkb, nst = (nkb + 3) >> 2;
        const int row0 = a.tile_list[4 * tile + 1], n_tok = a.tile_list[4 * tile + 2];
        const int nb = n_tok <= 16 ? 16 : n_tok <= 32 ? 32 : 64;
        const uint32_t b_tile = (uint32_t)nb * 128u;
        const CUtensorMap* map_x = side == 0 ? (nb == 16 ? &map_x16 : nb == 32 ? &map_x32 : &map_x64)
                                             : (nb == 16 ? &map_h16 : nb == 32 ? &map_h32 : &map_h64);
        if (SIDES == 2 && side == 1) {
          // the tile's h rows come from the w13 items of this very launch: wait until all nrt0 of them are counted, then order
          // the (generic-proxy) acquire before the TMA (async-proxy) reads that follow
          const int32_t* cnt = a.tile_list + 4 * tile + 3;
          int seen;
          unsigned int polls = 0;
          do {
            asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(seen) : "l"(cnt) : "memory");
            if (seen < nrt0) {
              __nanosleep(100);
              // every CTA is resident and w13 items wait for nothing, so this ends within microseconds; if the grid was NOT
              // fully resident (SMs held by something that itself waits for this stream) give up loudly rather than hang or
              // multiply garbage: set the bit and trap, so the host's next synchronisation fails
              if (++polls > (1u << 26)) {
                if (lane == 0 && a.status) atomicOr(a.status, 16);
                __threadfence();
                __trap();
              }
            }
          } while (seen < nrt0);
          asm volatile("fence.proxy.async;" ::: "memory");
        }
        for (int st = 0; st < nst; ++st) {
          mbar_wait(done0 + 8u * (uint32_t)sb, phb ^ 1u);            // the MMAs of the stage n_b back have read its activations
          const int nreal = min(4, nkb - st * 4);
          if (elect_one()) {
            const uint32_t full = bfull0 + 8u * (uint32_t)sb, dst = b_ring + (uint32_t)sb * b_stage;
            mbar_expect_tx(full, (uint32_t)(a.b3d ? 4 : nreal) * b_tile);   // a 3-D box always delivers 4 tiles (zero-filled)
            if (a.b3d) {
              tma_load_3d(dst, map_x, 0, row0, st * 4, full);
            } else {
              for (int u = 0; u < nreal; ++u) tma_load_2d(dst + (uint32_t)u * b_tile, map_x, (st * 4 + u) * kBlock, row0, full);
            }
          }
          __syncwarp();
          if (++sb == n_b) { sb = 0; phb ^= 1u; }
        }
      }
    } else if (warp == kDecWarpMma) {
      // =============================== MMA issuer (one thread) ===============================
      // A lone warp runs at one dependent instruction per ~6 cycles and every mbarrier / tcgen05 instruction costs it 60-100, so
      // this loop is kept to one wait, the MMAs (running addresses, constant accumulate flag except in an item's first stage)
      // and one commit per stage.
      const uint32_t afull0 = bar(kDbAFull), done0 = bar(kDbDone), bfull0 = bar(kDbBFull);
      const uint32_t desc_hi = (1024u >> 4) | (1u << 14) | (2u << 29);   // SBO 1024 B, descriptor version 1, SWIZZLE_128B
      const uint32_t blo0 = (b_ring & 0x3FFFFu) >> 4, bstep = b_stage >> 4;
      const uint32_t a_t0 = tmem + acol0;
      int sb = 0, slot = 0, it = 0;
      uint32_t pha = 0, phb = 0;
      auto issue = [&](const uint32_t d_t, const uint32_t a_t, const uint32_t b_lo, const uint32_t blo_step, const uint32_t idesc,
                       const int nreal, const bool first) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          if (u < nreal) {
            const uint64_t desc = ((uint64_t)desc_hi << 32) | (uint64_t)(b_lo + (uint32_t)u * blo_step);
#pragma unroll
            for (int kk = 0; kk < 4; ++kk)
              tc_mma_ts(d_t, a_t + (uint32_t)(u * 32 + 8 * kk), desc + (uint64_t)(2 * kk), idesc, first ? (uint32_t)(u | kk) : 1u);
          }
        }
      };
      for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++it) {
        int side, tile, rt;
        item_of(item, side, tile, rt);
        const int nkb = a.s[side].w.nkb, nst = (nkb + 3) >> 2;
        const int n_tok = a.tile_list[4 * tile + 2];
        const int nb = n_tok <= 16 ? 16 : n_tok <= 32 ? 32 : 64;   // rows of the activation box (the tile's stride in shared memory)
        const uint32_t idesc = make_idesc((n_tok + 15) & ~15);     // MMA N: the token count in 16s (48 of a 64-row box is fine)
        const uint32_t blo_step = ((uint32_t)nb * 128u) >> 4;
        const int db = it & 1;
        const uint32_t d_t = tmem + (uint32_t)db * acc_cols;
        GEMM_DBG_T0();
        mbar_wait(bar(kDbDEmpty + db), (((uint32_t)(it >> 1)) & 1u) ^ 1u);   // the epilogue has drained this accumulator
        GEMM_DBG_ADD(0);
        for (int st = 0; st < nst; ++st) {
          mbar_wait(bfull0 + 8u * (uint32_t)sb, phb);       // activations landed (long ago, normally: they run n_b stages ahead)
          mbar_wait(afull0 + 8u * (uint32_t)slot, pha);     // A operand written by the stage's dequant warps
          if (DBG && it == 0 && st == 0) stamp(5);
          tc_fence_after();
          GEMM_DBG_ADD(2);
          const int nreal = min(4, nkb - st * 4);
          if (elect_one()) {
            const uint32_t a_t = a_t0 + (uint32_t)slot * 128u, b_lo = blo0 + (uint32_t)sb * bstep;
            if (st == 0) issue(d_t, a_t, b_lo, blo_step, idesc, nreal, true);
            else issue(d_t, a_t, b_lo, blo_step, idesc, nreal, false);
            tc_commit(done0 + 8u * (uint32_t)sb);            // frees the A slot and the activation stage
            if (st == nst - 1) tc_commit(bar(kDbDFull + db));
          }
          __syncwarp();
          GEMM_DBG_ADD(3);
          if (++slot == kDecSlots) { slot = 0; pha ^= 1u; }
          if (++sb == n_b) { sb = 0; phb ^= 1u; }
        }
      }
      if (dbg_on) for (int i = 0; i < 6; ++i) a.dbg[16 + i] = dbg_acc[i];
    }
  }

  // ---- teardown ----
  if (DBG && warp == kDecDqWarps) stamp(8);               // epilogue warp 0 done with its stores
  tc_fence_before();
  __syncthreads();
  if (tid == 0) stamp(9);
  if (warp == kDecWarpMma) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(kTmemCols) : "memory");
  }
}

// ---- host side ------------------------------------------------------------------------------------------------------
// True when the decode-regime tcgen05 kernel serves this call: token tiles of at most 64 rows (either operand mode; 64-row tiles
// exist for nested absmax only, see gemm_tile_tokens).
constexpr int kDecMaxTile = 64;
inline bool dec_serves(int T, int k, int E, int mode, bool wide = false) {
  return (mode == NF4MOE_GEMM_FAST || mode == NF4MOE_GEMM_EXACT) && gemm_tile_tokens(T, k, E, wide && mode == NF4MOE_GEMM_FAST) <= kDecMaxTile;
}

struct DecHostSide {
  const nf4moe_qweight* w;
  const void* act_in;   // fp16 [T * k, w->cols] rows in sorted order
  void* out;
};

// SIDES as in dec_kernel: 0 = w13 (hs[0]), 1 = w2 (hs[1]), 2 = both in one launch (hs[1].act_in == hs[0].out: the h rows).
template <int SIDES>
int dec_launch_sides(const DecHostSide (&hs)[2], const int32_t* sorted_pairs, const int32_t* meta, const int32_t* tile_list, int T,
                     int k, int E, int act, int mode, void* stream, char* err, size_t err_len, int* launches, const void* bias,
                     int32_t* status, int tile_rows = 32) {
  const char* name = SIDES == 0 ? "gemm_w13(decode)" : SIDES == 1 ? "gemm_w2(decode)" : "gemm_w13_w2(decode)";
  auto fail = [&](int code, const char* msg) { snprintf(err, err_len, "%s: %s", name, msg); return code; };
  constexpr int s_lo = SIDES == 1 ? 1 : 0, s_hi = SIDES == 0 ? 0 : 1;
  if (!meta || !tile_list || T < 0 || k <= 0 || E <= 0) return fail(NF4MOE_EINVAL, "bad arguments");
  if (SIDES != 0 && !sorted_pairs) return fail(NF4MOE_EINVAL, "sorted_pairs required");
  if (mode != NF4MOE_GEMM_EXACT && mode != NF4MOE_GEMM_FAST) return fail(NF4MOE_EINVAL, "mode must be NF4MOE_GEMM_EXACT or NF4MOE_GEMM_FAST");
  for (int sd = s_lo; sd <= s_hi; ++sd) {
    const nf4moe_qweight* w = hs[sd].w;
    if (!hs[sd].act_in || !w || !w->packed || !hs[sd].out) return fail(NF4MOE_EINVAL, "bad arguments");
    if (w->rows <= 0 || w->cols <= 0 || w->cols % kBlock) return fail(NF4MOE_EINVAL, "cols must be a multiple of 64");
    if (sd == 0 && (w->rows & 1)) return fail(NF4MOE_EINVAL, "w13 rows must be even");
    if (!w->absmax_f32 && !(w->absmax_u8 && w->absmax2 && w->code2 && w->offset)) return fail(NF4MOE_EINVAL, "absmax missing");
    if (w->rows % 16 || w->cols > 4096) return fail(NF4MOE_EINVAL, "the tcgen05 regime needs rows % 16 == 0 and cols <= 4096");
    if ((reinterpret_cast<uintptr_t>(hs[sd].act_in) & 15u) || (reinterpret_cast<uintptr_t>(w->packed) & 15u) ||
        (reinterpret_cast<uintptr_t>(w->absmax_f32) & 15u) || (reinterpret_cast<uintptr_t>(w->absmax_u8) & 15u) ||
        (!w->absmax_f32 && ((reinterpret_cast<uintptr_t>(w->code2) & 15u) || (w->code2_stride & 3))))
      return fail(NF4MOE_EALIGN, "activations, packed weights, absmax and code2 must be 16-byte aligned");
  }
  if (SIDES != 1 && act != NF4MOE_ACT_SWIGLU_OAI && act != NF4MOE_ACT_CHUNK_SILU) return fail(NF4MOE_EINVAL, "bad activation");
  if (SIDES == 2) {
    if (act != NF4MOE_ACT_SWIGLU_OAI) return fail(NF4MOE_EINVAL, "the merged launch needs the fused activation");
    if ((hs[0].w->absmax_f32 == nullptr) != (hs[1].w->absmax_f32 == nullptr)) return fail(NF4MOE_EINVAL, "w13 and w2 must use the same absmax form");
    if (hs[1].act_in != hs[0].out || hs[1].w->cols * 2 != hs[0].w->rows) return fail(NF4MOE_EINVAL, "w2 must read the rows w13 writes");
  }
  if (T == 0) return 0;
  int dev = 0, sms = 0, cc = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaDeviceGetAttribute(&cc, cudaDevAttrComputeCapabilityMajor, dev);
  if (cc != 10) return fail(NF4MOE_EARCH, "needs an sm_100 device");
  const bool fast = mode == NF4MOE_GEMM_FAST;
  const bool nested = hs[s_lo].w->absmax_f32 == nullptr;
  DecArgs a;
  memset(&a, 0, sizeof(a));
  int nkb_max = 0;
  bool hole = fast && nested;
  for (int sd = s_lo; sd <= s_hi; ++sd) {
    a.s[sd].w = make_qw(*hs[sd].w);
    a.s[sd].out = hs[sd].out;
    nkb_max = std::max(nkb_max, a.s[sd].w.nkb);
    hole = hole && hs[sd].w->code2_stride == 0;
  }
  // one code2 table for every expert of every side served (same device pointer) lets it live in the decode table's holes
  if (SIDES == 2 && hs[0].w->code2 != hs[1].w->code2) hole = false;
  a.sorted_pairs = sorted_pairs;
  a.meta = meta;
  a.tile_list = const_cast<int32_t*>(tile_list);
  a.act = act;
  a.bias = SIDES != 1 ? (const __half*)bias : nullptr;
  a.status = status ? status : const_cast<int32_t*>(meta) + 4;
  a.dbg = gemm_debug_buffer();
  a.hole_code2 = hole;
  if (tile_rows != 16 && tile_rows != 32 && tile_rows != 64) return fail(NF4MOE_EINVAL, "the decode-regime kernel takes token tiles of 16, 32 or 64 rows");
  a.max_tile = tile_rows;
  for (int bi = 0; bi < 2; ++bi) {
    a.plan_base[bi] = bi ? 2048u : 1024u;
    for (int wide = 0; wide < 3; ++wide) {
      const bool fits = dec_plan(a.plan_base[bi], kGemmLutBytes, nkb_max, nested, nested && !hole, 16 << wide, &a.plan[bi * 3 + wide]);
      if (!fits && (16 << wide) <= std::max(tile_rows, 32))   // (a 64-wide plan is only needed when the tile list has 64-row tiles)
        return fail(NF4MOE_EINVAL, "K is too large for the decode pipeline's shared-memory plan with this absmax form and token-tile size");
    }
  }
#ifdef NF4MOE_EXPERIMENTS
  for (int i = 0; i < 6; ++i) {   // shallower rings for pipeline-depth measurements
    if (a.plan[i].n_raw == 0) continue;
    a.plan[i].n_raw = std::max(kDecGroups, std::min(a.plan[i].n_raw, experiment_env("NF4MOE_DEC_MAXRAW", kDecMaxRaw)) / kDecGroups * kDecGroups);
    a.plan[i].n_b = std::max(std::max(3, kDecGroups), std::min(a.plan[i].n_b, experiment_env("NF4MOE_DEC_MAXB", kDecMaxB)));
  }
#endif
  CUtensorMap map_w[2], map_a16[2], map_a32[2], map_a64[2];
  memset(map_w, 0, sizeof(map_w)); memset(map_a16, 0, sizeof(map_a16)); memset(map_a32, 0, sizeof(map_a32)); memset(map_a64, 0, sizeof(map_a64));
  a.b3d = 1;
  for (int pass = 0; pass < 2; ++pass) {   // pass 0 tries the 3-D activation maps for every side, pass 1 falls back to 2-D for all
    bool ok = true;
    for (int sd = s_lo; sd <= s_hi && ok; ++sd) {
      const nf4moe_qweight* w = hs[sd].w;
      if (pass == 0) ok = make_map_x3d(&map_a16[sd], hs[sd].act_in, (uint64_t)w->cols, (uint64_t)T * k, 16) &&
                          make_map_x3d(&map_a32[sd], hs[sd].act_in, (uint64_t)w->cols, (uint64_t)T * k, 32) &&
                          make_map_x3d(&map_a64[sd], hs[sd].act_in, (uint64_t)w->cols, (uint64_t)T * k, 64);
      else ok = make_map_2d(&map_a16[sd], hs[sd].act_in, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, (uint64_t)w->cols, (uint64_t)T * k, kBlock, 16) &&
                make_map_2d(&map_a32[sd], hs[sd].act_in, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, (uint64_t)w->cols, (uint64_t)T * k, kBlock, 32) &&
                make_map_2d(&map_a64[sd], hs[sd].act_in, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, (uint64_t)w->cols, (uint64_t)T * k, kBlock, 64);
    }
    if (ok) break;
    if (pass == 1) return fail(NF4MOE_EINVAL, "cuTensorMapEncodeTiled failed for the activations");
    a.b3d = 0;
  }
  for (int sd = s_lo; sd <= s_hi; ++sd) {
    const nf4moe_qweight* w = hs[sd].w;
    if (!make_map_2d(&map_w[sd], w->packed, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, (uint64_t)w->cols / 2, (uint64_t)E * w->rows, 128, 128))
      return fail(NF4MOE_EINVAL, "cuTensorMapEncodeTiled failed for the packed weights");
  }
  cudaError_t e;
  {   // once per kernel: the register pool the CTA is launched with must cover the setmaxnreg plan (else the kernel would hang)
    static std::atomic<int> regs_ok[2] = {{0}, {0}};
    int ok = regs_ok[fast].load(std::memory_order_acquire);
    if (ok == 0) {
      cudaFuncAttributes fa;
      e = fast ? cudaFuncGetAttributes(&fa, dec_kernel<SIDES, true, false>) : cudaFuncGetAttributes(&fa, dec_kernel<SIDES, false, false>);
      if (e != cudaSuccess) { snprintf(err, err_len, "%s: cudaFuncGetAttributes: %s", name, cudaGetErrorString(e)); return (int)e; }
      const int launched = (fa.numRegs + 7) / 8 * 8;
      ok = (32 * kDecDqWarps * std::max(kDecRegsDq, launched) + 128 * std::min(kDecRegsEpi, launched) + 128 * std::min(kDecRegsCtl, launched) <= launched * kDecThreads &&
            launched <= kDecRegsDq) ? 1 : -1;
      regs_ok[fast].store(ok, std::memory_order_release);
    }
    if (ok < 0) return fail(NF4MOE_EINVAL, "register plan of the decode kernel does not fit the registers it is launched with (build problem)");
  }
  const dim3 grid(sms), block(kDecThreads);
  const cudaStream_t st = (cudaStream_t)stream;
#ifdef NF4MOE_EXPERIMENTS
  if (SIDES == 0 && fast && a.dbg != nullptr) {   // the cycle-counter build exists for FAST w13 only (scripts/dec_roles.py)
    e = ensure_dyn_smem<dec_kernel<0, true, true>>((int)kDecSmemBytes);
    if (e == cudaSuccess) e = launch_pdl(dec_kernel<0, true, true>, grid, block, kDecSmemBytes, st, a, map_w[0], map_a16[0], map_a32[0], map_w[1], map_a16[1], map_a32[1], map_a64[0], map_a64[1]);
  } else
#endif
  if (fast) {
    e = ensure_dyn_smem<dec_kernel<SIDES, true, false>>((int)kDecSmemBytes);
    if (e == cudaSuccess) e = launch_pdl(dec_kernel<SIDES, true, false>, grid, block, kDecSmemBytes, st, a, map_w[0], map_a16[0], map_a32[0], map_w[1], map_a16[1], map_a32[1], map_a64[0], map_a64[1]);
  } else {
    e = ensure_dyn_smem<dec_kernel<SIDES, false, false>>((int)kDecSmemBytes);
    if (e == cudaSuccess) e = launch_pdl(dec_kernel<SIDES, false, false>, grid, block, kDecSmemBytes, st, a, map_w[0], map_a16[0], map_a32[0], map_w[1], map_a16[1], map_a32[1], map_a64[0], map_a64[1]);
  }
  ++*launches;
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e != cudaSuccess) { snprintf(err, err_len, "%s: %s", name, cudaGetErrorString(e)); return (int)e; }
  return 0;
}

template <bool W13>
int dec_launch(const void* act_in, const nf4moe_qweight* w, const int32_t* sorted_pairs, const int32_t* meta,
               const int32_t* tile_list, int T, int k, int E, int act, int mode, void* out, void* stream, char* err, size_t err_len,
               int* launches, const void* bias = nullptr, int32_t* status = nullptr, int tile_rows = 32) {
  DecHostSide hs[2] = {{nullptr, nullptr, nullptr}, {nullptr, nullptr, nullptr}};
  hs[W13 ? 0 : 1] = DecHostSide{w, act_in, out};
  return dec_launch_sides<W13 ? 0 : 1>(hs, sorted_pairs, meta, tile_list, T, k, E, act, mode, stream, err, err_len, launches, bias, status,
                                       tile_rows);
}

// w13 + SwiGLU and w2 in ONE launch: x_sorted [T k, H] -> h [T k, I] -> y [(token, slot) rows, H]
inline int dec_launch_both(const void* x_sorted, const nf4moe_qweight* w13, void* h, const nf4moe_qweight* w2, void* y,
                           const int32_t* sorted_pairs, const int32_t* meta, const int32_t* tile_list, int T, int k, int E, int act,
                           int mode, void* stream, char* err, size_t err_len, int* launches, const void* bias, int32_t* status,
                           int tile_rows = 32) {
  const DecHostSide hs[2] = {{w13, x_sorted, h}, {w2, h, y}};
  return dec_launch_sides<2>(hs, sorted_pairs, meta, tile_list, T, k, E, act, mode, stream, err, err_len, launches, bias, status, tile_rows);
}

}  // namespace nf4moe
