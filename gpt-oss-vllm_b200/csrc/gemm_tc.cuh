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
// Roles (640 threads, one persistent CTA per SM, registers re-balanced per warpgroup with setmaxnreg; the role index is the
// hardware warp id rotated by 4 so that the single-thread roles get the HIGHEST warp ids, which the sub-partition arbiter
// serves first):
//   role 0        TMA producer: activation tiles [<= tile_tokens x 64] per k-block, as whole-tile or 32-row boxes, own ring
//   role 1        TMEM allocator and the single MMA-issuing thread (tcgen05.mma kind::f16, A in TMEM, B in smem)
//   role 2        TMA producer: the item's absmax (one bulk copy per item, double-buffered) and raw NF4 stages [128 rows x 128 B]
//                 per 4 k-blocks, L2-prefetched kGemmWPrefetch stages ahead
//   roles 4-15    dequant: THREE warpgroups; k-block c of the CTA's run belongs to group c % 3
//   roles 16-19   epilogue: tcgen05.ld of one of the TWO accumulators while the next item's main loop fills the other;
//                 gpt-oss clamped SwiGLU on adjacent lanes (rows 2j / 2j+1 = gate / up) for w13, scatter to (token, slot)
//                 order for w2
// Replaces vLLM's dequantise-everything + Triton fused_experts (V/.../quantization/bitsandbytes.py:478-506, 588-603).
#pragma once
#include <cuda.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>

#include "common.cuh"

namespace nf4moe {

constexpr int kDec64MinMeanRows = 20;   // nested absmax: mean rows per expert from which the decode kernel uses 64-row token tiles ...
constexpr int kDec64MaxMeanRows = 80;   // ... and up to which it does (above: the prefill grouped GEMM, 160-row tiles)
constexpr int kDecMaxMeanRows = 40;     // mean routed rows per expert up to which token tiles are 32 rows (decode tcgen05 kernel)
constexpr int kDecAutoMinTokens = 8;    // tokens; up to this nf4moe_moe_forward(AUTO) uses the GEMV, above it a tcgen05 kernel
constexpr int kDqGroups = 3;            // dequant warpgroups; k-block c of a CTA belongs to group c % 3
#ifndef NF4MOE_GEMM_CLUSTER
#define NF4MOE_GEMM_CLUSTER 1
#endif
// CTAs per thread-block cluster of the prefill kernel.  The CTAs of a cluster work on neighbouring 128-row weight tiles of the
// SAME token tile, so they all need the same activation tile: each loads 1/C of its rows and TMA-multicasts them into every
// CTA's shared memory.  At M = 128 x N = 160 a lone CTA pulls 24 KB per k-block from L2 (20 KB of it activations): 9 TB/s over
// the chip, the L2 -> SM fabric's ceiling, with the tensor pipe idle half the time (round 1: 46 % active); sharing the tile
// across C CTAs divides the activation traffic by C.
constexpr int kGemmCluster = NF4MOE_GEMM_CLUSTER;
static_assert(kGemmCluster == 1 || kGemmCluster == 2 || kGemmCluster == 4, "NF4MOE_GEMM_CLUSTER must be 1, 2 or 4");
constexpr int kGemmThreads = (4 + 4 * kDqGroups + 4) * 32;   // 640: control WG, 3 dequant WGs, epilogue WG
constexpr int kGemmTokTile = 160;       // largest MMA N: two accumulators of 160 columns leave 192 TMEM columns = 3 slots of 2 k-blocks
constexpr int kGemmRowTile = 128;       // MMA M
constexpr int kGemmBoxRows = 32;        // token rows per activation TMA box (a tile is ceil(rows/32) boxes)
constexpr int kMaxAStages = 12;         // A (dequantised weights) stages in TMEM, 32 columns each: (512 - 2*tile) / 32
constexpr int kMaxBStages = 12;         // activation stages in shared memory (own ring, own empty barriers)
constexpr int kGemmWPrefetch = 6;       // raw-weight stages requested into L2 ahead of the shared-memory ring
constexpr int kMaxWStages = 6;
constexpr uint32_t kWStageBytes = 128 * 128;   // raw NF4: 128 rows x 4 k-blocks x 32 B
constexpr uint32_t kGemmSmemBytes = 232448 - 1024;  // dynamic part; the kernel's static __shared__ control block takes the rest
constexpr uint32_t kGemmLutBytes = 65536;
constexpr uint32_t kTmemCols = 512;

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
// The same load delivered to every CTA of `mask` (same shared-memory offset, same mbarrier offset in each of them).
__device__ __forceinline__ void tma_load_2d_mc(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar, uint16_t mask) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%2, %3}], [%4], %5;"
      ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(bar), "h"(mask)
      : "memory");
}
// tcgen05.commit that arrives on the mbarrier at this offset in EVERY CTA of `mask`
__device__ __forceinline__ void tc_commit_mc(uint32_t bar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar), "h"(mask)
               : "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
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
__device__ __forceinline__ uint32_t lds32_volatile(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr));
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
  int tile_tokens;                // largest token tile in tile_list (multiple of 32): sizes the activation stages and the accumulators
  int kps;                        // k-blocks per pipeline slot (2 or 4): one barrier wait and one commit serve kps k-blocks
  int n_a;                        // slots in TMEM: (512 - 2 * tile_tokens) / (32 * kps), each kps x 32 columns of A operand
  int want_b;                     // activation stages wanted (the kernel takes fewer if shared memory runs out, never below 2)
  int32_t* status;                // optional device word: bit 2 is raised when the shared-memory plan does not fit
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
// The main loop moves in SLOTS of kps k-blocks (kps = 2 for big token tiles, 4 for decode-sized ones).  With c = the CTA's
// running k-block count (each item padded to a multiple of kps), k-block c lives in slot (c / kps) % n_a, sub-block c % kps:
//   full[slot]  <- 4 dequant warps per sub-block (A operand written to TMEM) + the activation producer's expect_tx and
//                  TMA bytes: ONE wait per slot on the MMA thread.  Every mbarrier / tcgen05 / TMA instruction costs a
//                  single-thread role 100-300 cycles of latency here (measured), so they are rationed per slot, not per k-block
//   empty[slot] <- ONE tcgen05.commit per slot; waited on by the dequant warps (TMEM reuse, n_a slots later) and by the
//                  activation producer (shared-memory reuse, n_b / kps <= n_a slots later)
constexpr int kBarWFull = 0, kBarWEmpty = kBarWFull + kMaxWStages, kBarFull = kBarWEmpty + kMaxWStages,
              kBarEmpty = kBarFull + kMaxAStages, kBarDFull = kBarEmpty + kMaxAStages, kBarDEmpty = kBarDFull + 2,
              kBarMFull = kBarDEmpty + 2, kBarMEmpty = kBarMFull + 2, kNumGemmBars = kBarMEmpty + 2;

template <int N>
__device__ __forceinline__ void set_max_regs_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }
template <int N>
__device__ __forceinline__ void set_max_regs_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }

__device__ __forceinline__ uint32_t hadd2_u32(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("add.rn.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ uint32_t hmul2_u32(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("mul.rn.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}

// One weight row x one 64-wide NF4 block (32 packed bytes in v0, v1) -> 32 half2 registers in K order.
// Every byte is one table look-up; PRMT drops the byte into bits 8..15 of the 64 KB-aligned table address.
//   EXACT: entry = the two fp32 NF4 codes; each is scaled in fp32 and rounded ONCE to fp16: bit-identical to bitsandbytes'
//          dequantize_4bit (fp32 multiply, store as half), so the tensor cores see the very matrix the reference multiplies.
//   FAST : entry = the two codes as half2, scaled by fp16(absmax) with one HMUL2: half the shared-memory wavefronts and
//          3 instead of 5 instructions per byte; differs from EXACT by fp16 roundings of code and absmax (~2^-11 relative).
// NF4 -> fp16 for the 8 nibbles of one packed word WITHOUT shared memory (FAST operand only): the 16 fp16 codes as two byte
// planes (high bytes, low bytes), each a pair of 8-entry PRMT tables; PRMT takes four 3-bit indices at a time straight from
// the packed data (selector = data & 0x7777), bit 3 of every nibble picks the table half through a byte mask that a
// sign-replicating PRMT builds from the data and the data << 4, and two more PRMTs interleave the planes into half2 pairs in
// the order the look-up table uses (element 2j = high nibble in the low half).  21 ALU instructions per word against 4 PRMT +
// 4 LDS: the decode kernel is bound by the shared-memory data pipe (profiles/r2_notes.md), so a share of the words is moved
// here.  Same bits as the table: the tables ARE the fp16 roundings of kNF4 (tests/test_abi_host.py checks the constants).
constexpr uint32_t kNF4HiA = 0xb6b8b9bcu, kNF4HiB = 0x00adb1b4u, kNF4HiC = 0x3533312du, kNF4HiD = 0x3c393837u;
constexpr uint32_t kNF4LoA = 0x52339200u, kNF4LoB = 0x00d4ea8du, kNF4LoC = 0x68e02618u, kNF4LoD = 0x00c9800du;
__device__ __forceinline__ uint32_t sel_bytes(uint32_t a, uint32_t b, uint32_t m) { return (a & ~m) | (b & m); }   // one LOP3
__device__ __forceinline__ void decode_word_alu(const uint32_t w, const uint32_t am2, uint32_t* out) {
  const uint32_t s = w & 0x77777777u, w4 = w << 4;
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    const uint32_t sel = half ? (s >> 16) : s;
    const uint32_t m = prmt(w4, w, half ? 0xFBEAu : 0xD9C8u);
    const uint32_t rh = sel_bytes(prmt(kNF4HiA, kNF4HiB, sel), prmt(kNF4HiC, kNF4HiD, sel), m);
    const uint32_t rl = sel_bytes(prmt(kNF4LoA, kNF4LoB, sel), prmt(kNF4LoC, kNF4LoD, sel), m);
    out[2 * half] = hmul2_u32(prmt(rl, rh, 0x4051u), am2);
    out[2 * half + 1] = hmul2_u32(prmt(rl, rh, 0x6273u), am2);
  }
}

// ARITH = how many of the k-block's 8 packed words per thread take the shared-memory-free path (FAST only)
template <bool FAST, int ARITH = 0>
__device__ __forceinline__ void decode_block(const uint4 v0, const uint4 v1, const uint32_t lut_lane, const float am,
                                             uint32_t (&regs)[32]) {
  const uint32_t wd[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
  const uint32_t am2 = FAST ? pack_half2(am, am) : 0u;
#pragma unroll
  for (int wi = 0; wi < 8; ++wi) {
    if (FAST && wi >= 8 - ARITH) {
      decode_word_alu(wd[wi], am2, &regs[wi * 4]);
      continue;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const uint32_t addr = prmt(wd[wi], lut_lane, 0x7604u | ((uint32_t)i << 4));
      if (FAST) {
        regs[wi * 4 + i] = hmul2_u32(lds32(addr), am2);
      } else {
        const float2 f = lds_f32x2(addr);
        regs[wi * 4 + i] = pack_half2(__fmul_rn(f.x, am), __fmul_rn(f.y, am));
      }
    }
  }
}

// Shared-memory plan around the 64 KB-aligned decode table -- the same function on the host (gemm_launch refuses a shape that
// does not fit) and on the device (a plan that does not fit there raises NF4MOE_STATUS_GEMM_NOFIT instead of doing nothing).
// The activation ring is ONE contiguous block right above the table (stage addresses are then base + i * b_bytes: no table
// look-up on the single-thread roles); raw-weight stages fill the rest above; the two absmax stages and more raw-weight stages
// go below.
struct GemmPlan {
  int n_b, n_w;
  uint32_t wst[6], mst[2];
};
__host__ __device__ inline bool gemm_plan(uint32_t base, uint32_t lut, uint32_t smem_bytes, int nkb, bool nested, int tile_tokens,
                                          int kps, int n_a, int want_b, GemmPlan* p) {
  const uint32_t b_bytes = (uint32_t)tile_tokens * 128u;
  const uint32_t m_tile = (uint32_t)(128 * nkb) * (nested ? 1u : 4u);
  const uint32_t m_bytes = (((m_tile + 15u) & ~15u) + (nested ? 1024u : 0u) + 1023u) & ~1023u;
  const uint32_t above0 = lut + 65536u, above1 = base + smem_bytes;
  int nb = want_b < 12 ? want_b : 12;
  if (nb > n_a * kps) nb = n_a * kps;                        // n_b / kps <= n_a keeps the shared slots in phase
  if (above1 <= above0) return false;
  if (nb > (int)((above1 - above0) / b_bytes)) nb = (int)((above1 - above0) / b_bytes);
  nb -= nb % kps;                                            // whole slots only
  uint32_t cur[2] = {(base + 1023u) & ~1023u, above0 + (uint32_t)nb * b_bytes};
  const uint32_t hi[2] = {lut, above1};
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
  int iw = 0, im = 0;
  for (int i = 0; i < 2; ++i) { const uint32_t at = place(m_bytes); if (at) p->mst[im++] = at; }
  for (int i = 0; i < 6; ++i) { const uint32_t at = place(128u * 128u); if (at) p->wst[iw++] = at; else break; }
  p->n_w = iw;
  p->n_b = (im == 2 && iw >= 2 && nb >= kps) ? nb : 0;
  return p->n_b > 0;
}

// Warp roles (640 threads, one persistent CTA per SM, registers re-balanced per warpgroup with setmaxnreg):
//   warp 0        TMA producer: activation tiles [<= tile_tokens x 64] per k-block as 32-row boxes, own ring of n_b stages
//   warp 1        TMEM allocator and the single MMA-issuing thread (tcgen05.mma kind::f16, A in TMEM, B in smem)
//   warp 2        TMA producer: the tile's absmax (one bulk copy per item, double-buffered) and raw NF4 stages
//                 [128 rows x 128 B] = 4 k-blocks, L2-prefetched kGemmWPrefetch stages ahead
//   warps 4-15    dequant: three warpgroups; k-block c (counted over the CTA's whole run) belongs to group c % 3, A stage c % n_a.
//                 A thread owns one weight row: absmax is a per-thread scalar, the result goes to TMEM with one tcgen05.st.x32
//   warps 16-19   epilogue: tcgen05.ld of one accumulator while the next item's main loop fills the other
template <bool W13, bool FAST>
__global__ void __launch_bounds__(kGemmThreads, 1) gemm_kernel(const GemmArgs a, const __grid_constant__ CUtensorMap map_w,
                                                              const __grid_constant__ CUtensorMap map_x,
                                                              const __grid_constant__ CUtensorMap map_xw) {
  // map_x: 32-row activation boxes (ragged tiles); map_xw: boxes of tile_tokens / kGemmCluster rows (this CTA's share of a
  // whole tile)
  extern __shared__ __align__(1024) uint8_t dyn_smem[];
  __shared__ __align__(8) uint64_t s_bar[kNumGemmBars];
  __shared__ uint32_t s_tmem;
  __shared__ GemmPlan s_plan;
  // Role index: the control warps take the HIGHEST hardware warp ids (16-19).  The sub-partition arbiter serves the highest
  // warp id first, so as warps 0-3 the single-thread roles were starved of issue slots by the twelve dequant warps (every
  // mbarrier / tcgen05 / TMA instruction of theirs took 100-300 cycles); `warp` below is the ROLE index (0-3 control, 4-15
  // dequant, 16-19 epilogue), `warp & 3` is still the hardware TMEM lane quarter.
  grid_dep_launch();
  const int tid = threadIdx.x, lane = tid & 31;
  const int hw_warp = tid >> 5;
  const int warp = (hw_warp + 4) % (kGemmThreads / 32);
  const int nkb = a.w.nkb;
  const int n_a = a.n_a, kps = a.kps;
  const int nslot = (nkb + kps - 1) / kps;                  // slots per item (the last one may be partly padding)
  const bool nested = a.w.absmax_f32 == nullptr;
  const uint32_t b_bytes = (uint32_t)a.tile_tokens * 128u;
  const uint32_t acol0 = 2u * (uint32_t)a.tile_tokens;      // TMEM: accumulators at columns 0 and tile_tokens, A stages after them
  // absmax stage: the tile's [128 rows x nkb] absmax (fp32, or u8 + the expert's 1 KB code2 table when nested)
  const uint32_t m_tile = (uint32_t)(kGemmRowTile * nkb) * (nested ? 1u : 4u);
  const uint32_t m_code2 = (m_tile + 15u) & ~15u;            // offset of code2 inside the stage
  const uint32_t m_bytes = (m_code2 + (nested ? 1024u : 0u) + 1023u) & ~1023u;
  const uint32_t base = smem_u32(dyn_smem);
  // decode table: 256 rows of 256 B, row b = one entry per lane (EXACT: 8 B = fp32 codes of the high and low nibble,
  // FAST: 4 B = the same two codes as half2), 64 KB-aligned so a table address is [hi16 | byte | lane offset]
  const uint32_t lut = pin((base + kGemmLutBytes - 1u) & ~(kGemmLutBytes - 1u));
  const uint32_t bar0 = pin(smem_u32(s_bar));   // pinned: otherwise every use re-derives it from SR_CgaCtaId
  auto bar = [&](int i) -> uint32_t { return bar0 + 8u * (uint32_t)i; };
  const uint32_t b_ring = pin(lut + kGemmLutBytes);   // activation stage i lives at b_ring + i * b_bytes

  // ---- one-time setup ----
  if (tid == 0) {
    gemm_plan(base, lut, kGemmSmemBytes, nkb, nested, a.tile_tokens, a.kps, n_a, a.want_b, &s_plan);
    for (int i = 0; i < kMaxWStages; ++i) { mbar_init(bar(kBarWFull + i), 1); mbar_init(bar(kBarWEmpty + i), 4 * kDqGroups); }
    for (int i = 0; i < kMaxAStages; ++i) { mbar_init(bar(kBarFull + i), 4 * a.kps + 1); mbar_init(bar(kBarEmpty + i), kGemmCluster); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar(kBarDFull + i), 1); mbar_init(bar(kBarDEmpty + i), 4);
      mbar_init(bar(kBarMFull + i), 1); mbar_init(bar(kBarMEmpty + i), 4 * kDqGroups);
    }
    fence_barrier_init();
  }
  for (int i = tid; i < 256 * 32; i += kGemmThreads) {
    const int b = i >> 5, l = i & 31;
    if (FAST) {
      sts32(lut + (uint32_t)b * 256u + (uint32_t)l * 4u, pack_half2(kNF4[b >> 4], kNF4[b & 15]));
    } else {
      const uint32_t p = lut + (uint32_t)b * 256u + (uint32_t)l * 8u;
      sts32(p, __float_as_uint(kNF4[b >> 4]));        // element 2j   (high nibble)
      sts32(p + 4u, __float_as_uint(kNF4[b & 15]));   // element 2j+1 (low nibble)
    }
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(kTmemCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  // launched with launch_pdl(): everything above (decode table, barriers, TMEM allocation) overlaps the previous kernel's tail;
  // everything below reads what the bucket / permute (and, for w2, the w13) kernel wrote
  grid_dep_wait();
  tc_fence_before();
  __syncthreads();
  if (kGemmCluster > 1) cluster_sync_all();   // every CTA's barriers exist before a peer multicasts into them
  tc_fence_after();
  const uint32_t tmem = s_tmem;
  const int n_bst = s_plan.n_b, n_wst = s_plan.n_w;
  if (tid == 0 && n_bst == 0 && a.status != nullptr && a.meta[2] > 0) atomicOr(a.status, 4);

  const bool dbg_on = a.dbg != nullptr && blockIdx.x == 0 && lane == 0;
  long long dbg_acc[6] = {0, 0, 0, 0, 0, 0};
  // Work items.  A CLUSTER walks (token tile, group of kGemmCluster neighbouring row tiles); CTA `cr` of the cluster takes row
  // tile rtg * C + cr.  When the row tiles do not divide by C the last group's surplus CTAs redo the last real row tile (they
  // must keep loading their share of the activations and committing) and store nothing.
  const int n_row_tiles = (a.w.rows + kGemmRowTile - 1) / kGemmRowTile;
  const int n_rtg = (n_row_tiles + kGemmCluster - 1) / kGemmCluster;
  const int cr = kGemmCluster > 1 ? (int)cluster_ctarank() : 0;
  const int item0 = (int)blockIdx.x / kGemmCluster, item_step = (int)gridDim.x / kGemmCluster;
  const int n_items = n_bst > 0 ? a.meta[2] * n_rtg : 0;   // (n_bst == 0: K too large for the stage budget: status bit 2)
  auto item_tile = [&](int item, int& tile, int& rt) -> bool {   // returns false for a surplus (store-nothing) CTA
    tile = item / n_rtg;
    const int raw = (item - tile * n_rtg) * kGemmCluster + cr;
    rt = min(raw, n_row_tiles - 1);
    return raw < n_row_tiles;
  };
  constexpr uint16_t kMcMask = (uint16_t)((1u << kGemmCluster) - 1u);

  if (warp < 4) {
    set_max_regs_dec<48>();
    if (warp == 0) {
      // =============================== TMA producer: activation tiles, kps stages per slot ===============================
      // A lone warp runs dependent instructions at ~6 cycles each, so this loop is kept to a few dozen instructions per slot:
      // running addresses instead of index arithmetic, no shared-memory look-ups, one wait and one expect_tx per slot.
      const int lead = n_bst / kps;                       // slots the shared-memory ring holds
      const uint32_t full0 = bar(kBarFull), full_end = bar(kBarFull + n_a);
      const uint32_t empty0 = bar(kBarEmpty), empty_end = bar(kBarEmpty + n_a);
      const uint32_t slot_bytes = (uint32_t)kps * b_bytes, ring_end = b_ring + (uint32_t)n_bst * b_bytes;
      uint32_t full_a = full0, empty_a = empty0, dst = b_ring, pe = 0;
      int cs = 0;                                         // slots issued so far
      const int share = a.tile_tokens / kGemmCluster;      // rows of a whole tile this CTA loads (for every CTA of the cluster)
      for (int item = item0; item < n_items; item += item_step) {
        const int tile = item / n_rtg;
        const int row0 = a.tile_list[4 * tile + 1], n_tok = a.tile_list[4 * tile + 2];
        const bool whole = n_tok == a.tile_tokens;   // a full tile is one box per CTA; a ragged one only the 32-row boxes it really has
        const int n_box = (n_tok + kGemmBoxRows - 1) / kGemmBoxRows;
        // bytes that LAND in this CTA's stage, whoever loads them
        const uint32_t kb_bytes = whole ? b_bytes : (uint32_t)n_box * (kGemmBoxRows * 128u);
        for (int kb0 = 0; kb0 < nkb; kb0 += kps, ++cs) {
          if (cs >= lead) {                          // the MMAs that read these B stages (slot cs - lead) are done
            mbar_wait(empty_a, pe);
            empty_a += 8u;
            if (empty_a == empty_end) { empty_a = empty0; pe ^= 1u; }
          }
          const int nreal = min(kps, nkb - kb0);
          if (elect_one()) {
            mbar_expect_tx(full_a, (uint32_t)nreal * kb_bytes);
            for (int u = 0; u < nreal; ++u) {
              if (kGemmCluster == 1) {
                if (whole) {
                  tma_load_2d(dst + (uint32_t)u * b_bytes, &map_xw, (kb0 + u) * kBlock, row0, full_a);
                } else {
                  for (int bx = 0; bx < n_box; ++bx)
                    tma_load_2d(dst + (uint32_t)u * b_bytes + (uint32_t)bx * (kGemmBoxRows * 128u), &map_x, (kb0 + u) * kBlock,
                                row0 + bx * kGemmBoxRows, full_a);
                }
              } else if (whole) {   // this CTA's 1/C of the rows, delivered to the same stage of every CTA of the cluster
                tma_load_2d_mc(dst + (uint32_t)u * b_bytes + (uint32_t)(cr * share) * 128u, &map_xw, (kb0 + u) * kBlock,
                               row0 + cr * share, full_a, kMcMask);
              } else {              // ragged: the 32-row boxes are dealt round-robin
                for (int bx = cr; bx < n_box; bx += kGemmCluster)
                  tma_load_2d_mc(dst + (uint32_t)u * b_bytes + (uint32_t)bx * (kGemmBoxRows * 128u), &map_x, (kb0 + u) * kBlock,
                                 row0 + bx * kGemmBoxRows, full_a, kMcMask);
              }
            }
          }
          __syncwarp();
          dst += slot_bytes;
          if (dst == ring_end) dst = b_ring;
          full_a += 8u;
          if (full_a == full_end) full_a = full0;
        }
      }
      if (kGemmCluster > 1) {
        // A peer's MMA commits arrive on THIS CTA's `empty` barriers; none may still be in flight when the CTA exits.  Waiting
        // out the slots not yet waited for (in order, as above) sees every peer's last arrival.
        for (int i = min(cs, lead); i > 0; --i) {
          mbar_wait(empty_a, pe);
          empty_a += 8u;
          if (empty_a == empty_end) { empty_a = empty0; pe ^= 1u; }
        }
      }
    } else if (warp == 2) {
      // ============ TMA producer: the tile's absmax (one bulk copy per item, double-buffered) and the raw NF4 bytes, one
      // stage per 4 k-blocks, on their own ring so they run ahead of the MMA.  Lean on purpose (see the activation producer):
      // per-item coordinates are computed once, the per-stage work is a wait, an expect_tx, a TMA and an L2 prefetch. ======
      const uint64_t pol = l2_policy_evict_last();
      const uint32_t am_elt = nested ? 1u : 4u;
      const uint8_t* am_base = nested ? a.w.absmax_u8 : reinterpret_cast<const uint8_t*>(a.w.absmax_f32);
      const int spi = (nkb + 3) >> 2;                       // raw-weight stages per item
      const uint32_t wfull0 = bar(kBarWFull), wempty0 = bar(kBarWEmpty);
      int sw = 0, it = 0;
      uint32_t phw = 0;
      // L2 prefetch runs kGemmWPrefetch stages ahead: (pf_item, pf_st) is its cursor, pf_row the weight row of pf_item's tile
      int pf_item = item0, pf_st = 0, pf_ahead = 0, pf_row = 0;
      auto item_row = [&](int item) -> int {
        int tile, rt;
        item_tile(item, tile, rt);
        return a.tile_list[4 * tile] * a.w.rows + rt * kGemmRowTile;
      };
      if (pf_item < n_items) pf_row = item_row(pf_item);
      for (int item = item0; item < n_items; item += item_step, ++it) {
        int tile, rt;
        item_tile(item, tile, rt);
        const int e = a.tile_list[4 * tile];
        const int c1 = e * a.w.rows + rt * kGemmRowTile;
        {
          const int mb = it & 1;
          const int rows_here = min(kGemmRowTile, a.w.rows - rt * kGemmRowTile);
          const uint32_t bytes = (uint32_t)(rows_here * nkb) * am_elt;
          mbar_wait(bar(kBarMEmpty + mb), ((uint32_t)(it >> 1) & 1u) ^ 1u);
          if (elect_one()) {
            mbar_expect_tx(bar(kBarMFull + mb), bytes + (nested ? 1024u : 0u));
            bulk_g2s(s_plan.mst[mb], am_base + ((long long)e * a.w.nb + (long long)rt * kGemmRowTile * nkb) * am_elt, bytes,
                     bar(kBarMFull + mb), pol);
            if (nested) bulk_g2s(s_plan.mst[mb] + m_code2, a.w.code2 + (long long)e * a.w.code2_stride, 1024u, bar(kBarMFull + mb), pol);
          }
          __syncwarp();
        }
        for (int st = 0; st < spi; ++st) {
          while (pf_ahead <= kGemmWPrefetch && pf_item < n_items) {
            if (elect_one())
              asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global.tile [%0, {%1, %2}];" ::"l"(&map_w), "r"(pf_st * 128), "r"(pf_row) : "memory");
            ++pf_ahead;
            if (++pf_st == spi) {
              pf_st = 0;
              pf_item += item_step;
              if (pf_item < n_items) pf_row = item_row(pf_item);
            }
          }
          --pf_ahead;
          mbar_wait(wempty0 + 8u * (uint32_t)sw, phw ^ 1u);
          if (elect_one()) {
            mbar_expect_tx(wfull0 + 8u * (uint32_t)sw, kWStageBytes);
            tma_load_2d(s_plan.wst[sw], &map_w, st * 128, c1, wfull0 + 8u * (uint32_t)sw);
          }
          __syncwarp();
          if (++sw == n_wst) { sw = 0; phw ^= 1u; }
        }
      }
    } else if (warp == 1) {
      // =============================== MMA issuer (one thread) ===============================
      // Per slot: ONE barrier wait, 4 * kps MMAs, ONE commit; running addresses / descriptors, nothing looked up.
      const uint32_t full0 = bar(kBarFull), full_end = bar(kBarFull + n_a);
      const uint32_t d_empty0 = bar(kBarDEmpty), d_full0 = bar(kBarDFull);
      const uint32_t desc_hi = (1024u >> 4) | (1u << 14) | (2u << 29);            // SBO 1024 B, descriptor version 1, SWIZZLE_128B
      const uint32_t blo0 = (b_ring & 0x3FFFFu) >> 4, blo_step = b_bytes >> 4, blo_end = blo0 + (uint32_t)n_bst * blo_step;
      const uint32_t a_t0 = tmem + acol0, a_step = (uint32_t)kps * 32u, a_end = a_t0 + (uint32_t)n_a * a_step;
      uint32_t full_a = full0, pha = 0, b_lo = blo0, a_t = a_t0;
      int it = 0;
      for (int item = item0; item < n_items; item += item_step, ++it) {
        const int tile = item / n_rtg;
        const int n_tok = a.tile_list[4 * tile + 2];
        const uint32_t idesc = make_idesc(max(16, (n_tok + 15) & ~15));
        const int db = it & 1;
        const uint32_t d_t = tmem + (uint32_t)(db * a.tile_tokens);
        mbar_wait(d_empty0 + 8u * (uint32_t)db, (((uint32_t)(it >> 1)) & 1u) ^ 1u);   // the epilogue has drained this accumulator
        tc_fence_after();
        for (int kb0 = 0; kb0 < nkb; kb0 += kps) {
          mbar_wait(full_a, pha);
          tc_fence_after();
          const int nreal = min(kps, nkb - kb0);
          if (elect_one()) {
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              if (u < nreal) {
                const uint64_t desc = ((uint64_t)desc_hi << 32) | (uint64_t)(b_lo + (uint32_t)u * blo_step);
#pragma unroll
                for (int kk = 0; kk < 4; ++kk)
                  tc_mma_ts(d_t, a_t + (uint32_t)(u * 32 + 8 * kk), desc + (uint64_t)(2 * kk), idesc, (uint32_t)(kb0 | u | kk));
              }
            }
            // frees the slot's A operand and activation stages -- the stages in EVERY CTA of the cluster, whose producers
            // write into this CTA's copy too
            if (kGemmCluster == 1) tc_commit(full_a + 8u * (uint32_t)(kBarEmpty - kBarFull));
            else tc_commit_mc(full_a + 8u * (uint32_t)(kBarEmpty - kBarFull), kMcMask);
            if (kb0 + kps >= nkb) tc_commit(d_full0 + 8u * (uint32_t)db);
          }
          __syncwarp();
          full_a += 8u;
          a_t += a_step;
          if (full_a == full_end) { full_a = full0; a_t = a_t0; pha ^= 1u; }
          b_lo += (uint32_t)kps * blo_step;
          if (b_lo == blo_end) b_lo = blo0;
        }
      }
      (void)a_end;
    }
  }
  else if (warp < 4 + 4 * kDqGroups) {
    // =============================== dequant: NF4 -> fp16 A operand in TMEM ===============================
    set_max_regs_inc<112>();
    const int q = warp & 3, g = (warp - 4) >> 2;
    const int r = q * 32 + lane;
    const uint32_t lut_lane = pin(lut | ((uint32_t)lane << (FAST ? 2 : 3)));
    const uint32_t t_lane = tmem + ((uint32_t)(q * 32) << 16) + acol0;
    const uint32_t swz = (uint32_t)(r & 7);
    const uint32_t full0 = bar(kBarFull), empty0 = bar(kBarEmpty);
    const int vkb = nslot * kps;                 // k-blocks per item including the padding of the last slot
    const int spi = (nkb + 3) >> 2;              // raw-weight stages per item
    // This group owns the virtual k-blocks c = g, g + 3, g + 6, ... of the CTA's run; (sl, sub, pha) track slot, sub-block and
    // slot-use parity of the NEXT owned one, v its index inside the current item.
    int sw = 0, it = 0, sl = 0, sub = 0, v = g;
    uint32_t phw = 0, pha = 0;
    while (sub >= kps) { sub -= kps; ++sl; }
    sub = g % kps; sl = g / kps;                 // g < 3 <= n_a * kps
    for (int item = item0; item < n_items; item += item_step, ++it) {
      int tile, rt;
      item_tile(item, tile, rt);
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
      const uint32_t m_st = pin(s_plan.mst[mb]);
      const uint32_t m_row = m_st + (uint32_t)(r * nkb) * (nested ? 1u : 4u);
      int ws = 0;                                // raw-weight stage of the item this warp has reached (not yet released)
      mbar_wait(bar(kBarWFull + sw), phw);
      uint32_t wrow = pin(s_plan.wst[sw]) + (uint32_t)r * 128u;
      for (; v < vkb; v += kDqGroups) {
        const bool real = v < nkb;
        const int ws_need = real ? (v >> 2) : spi - 1;
        while (ws < ws_need) {                   // release the stages this warp is done with, wait for the next one
          __syncwarp();
          if (lane == 0) mbar_arrive(bar(kBarWEmpty + sw));
          if (++sw == n_wst) { sw = 0; phw ^= 1u; }
          ++ws;
          mbar_wait(bar(kBarWFull + sw), phw);
          wrow = pin(s_plan.wst[sw]) + (uint32_t)r * 128u;
        }
        GEMM_DBG_ADD(1);
        uint32_t regs[32];
        if (real) {
          float am;
          if (nested) {
            const float c = lds_f32(m_st + m_code2 + 4u * lds_u8(m_row + (uint32_t)v));
            am = __fadd_rn(__fmul_rn(c, v < blk_hi ? a2_lo : a2_hi), off_e);
          } else {
            am = lds_f32(m_row + 4u * (uint32_t)v);
          }
          const uint32_t j2 = (uint32_t)(v & 3) * 2u;
          const uint4 v0 = lds128(wrow + ((j2 ^ swz) << 4));
          const uint4 v1 = lds128(wrow + (((j2 + 1u) ^ swz) << 4));
          decode_block<FAST>(v0, v1, lut_lane, am, regs);
        }
        GEMM_DBG_ADD(2);
        mbar_wait(empty0 + 8u * (uint32_t)sl, pha ^ 1u);   // the MMAs that read this slot last time are done
        GEMM_DBG_ADD(3);
        if (real) {
          tc_fence_after();
          tmem_st32(t_lane + (uint32_t)(sl * kps + sub) * 32u, regs);
          asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
          tc_fence_before();
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(full0 + 8u * (uint32_t)sl);   // padding sub-blocks arrive too, so the slot completes
        GEMM_DBG_ADD(4);
        sub += kDqGroups;
        while (sub >= kps) { sub -= kps; if (++sl == n_a) { sl = 0; pha ^= 1u; } }
      }
      v -= vkb;                                  // position of the next owned k-block inside the next item
      while (ws < spi) {                         // release the item's remaining raw-weight stages
        __syncwarp();
        if (lane == 0) mbar_arrive(bar(kBarWEmpty + sw));
        if (++sw == n_wst) { sw = 0; phw ^= 1u; }
        if (++ws < spi) mbar_wait(bar(kBarWFull + sw), phw);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(kBarMEmpty + mb));
    }
    if (dbg_on && q == 0) for (int i = 0; i < 6; ++i) a.dbg[24 + 8 * g + i] = dbg_acc[i];
  } else {
    // =============================== epilogue: TMEM -> registers -> global, overlapped with the next tile's main loop ======
    const int q = warp & 3;
    const int r = q * 32 + lane;
    const uint32_t t_lane = tmem + ((uint32_t)(q * 32) << 16);
    int it = 0;
    for (int item = item0; item < n_items; item += item_step, ++it) {
      int tile, rt;
      const bool stores = item_tile(item, tile, rt);
      const int row0 = a.tile_list[4 * tile + 1], n_tok = a.tile_list[4 * tile + 2];
      const int n = rt * kGemmRowTile + r;
      const int db = it & 1;
      GEMM_DBG_T0();
      mbar_wait(bar(kBarDFull + db), (uint32_t)(it >> 1) & 1u);
      GEMM_DBG_ADD(0);
      tc_fence_after();
      const bool row_ok = stores && n < a.w.rows;
      for (int c0 = 0; c0 < n_tok; c0 += 32) {
        uint32_t v[32];
        tmem_ld32(t_lane + (uint32_t)(db * a.tile_tokens + c0), v);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        if (W13 && a.act == NF4MOE_ACT_SWIGLU_OAI) {
          // Rows 2j / 2j+1 (gate / up) sit in adjacent lanes.  The products are rounded to fp16 (the reference stores the GEMM
          // result as fp16 before the activation) and packed two tokens per register; the even lane then takes tokens
          // 0..15 of the chunk and the odd lane tokens 16..31, so ONE exchange of 8 packed registers gives each lane the
          // (gate, up) pairs of its 16 tokens.  Clamps run in half2 (exact), the sigmoid in fp32.
          __half* h = reinterpret_cast<__half*>(a.out);
          const int I = a.w.rows >> 1;
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
          const int t0 = c0 + (odd ? 16 : 0);                       // first token of this lane's 16
          __half* dst = h + (long long)(row0 + t0) * I + (n >> 1);
          const bool full = row_ok && t0 + 16 <= n_tok;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const uint32_t r2 = swiglu_oai_h2(gt[i], up[i]);
            const __half lo = __ushort_as_half((unsigned short)(r2 & 0xFFFFu)), hi = __ushort_as_half((unsigned short)(r2 >> 16));
            if (full) {
              dst[(long long)(2 * i) * I] = lo;
              dst[(long long)(2 * i + 1) * I] = hi;
            } else if (row_ok) {
              if (t0 + 2 * i < n_tok) dst[(long long)(2 * i) * I] = lo;
              if (t0 + 2 * i + 1 < n_tok) dst[(long long)(2 * i + 1) * I] = hi;
            }
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
              y[(long long)a.sorted_pairs[row0 + c0 + i] * a.w.rows + n] = __float2half_rn(__uint_as_float(v[i]));
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(kBarDEmpty + db));
      GEMM_DBG_ADD(1);
    }
    if (dbg_on && q == 0) for (int i = 0; i < 6; ++i) a.dbg[56 + i] = dbg_acc[i];
  }

  // ---- teardown ----
  tc_fence_before();
  __syncthreads();
  if (kGemmCluster > 1) cluster_sync_all();   // no CTA leaves while a peer may still write its shared memory or barriers
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

// Token-tile size (multiple of 32).  Small batches: 32-token tiles (4 KB activation stages, three 4-k-block slots in TMEM).
// Otherwise 160: two 160-column accumulators leave 192 TMEM columns = three 2-k-block slots, and four 20 KB activation
// stages fit beside the 64 KB decode table and two 23 KB flat-absmax stages.
// wide: 64-row token tiles may be used (the caller knows it runs the decode kernel with the FAST operand on nested absmax: what
// fits that kernel's shared memory).  Without it: round 1's 32 / 160 split, which suits either absmax form.
inline int gemm_tile_tokens(int T, int k, int E, bool wide = false) {
  const int tile_env = experiment_env("NF4MOE_TILE", 0);   // crossover measurements (NF4MOE_EXPERIMENTS builds only)
  if (tile_env == 32 || tile_env == 160 || (tile_env == 64 && wide)) return tile_env;
  // 32-row tiles (decode tcgen05 kernel; an expert with more rows gets several tiles and its weights are decoded once per tile)
  // while the mean routed rows per expert stay <= 40 (measured crossover with the 160-row prefill pipeline: 48 rows per expert at
  // E = 32), or <= 128 with at most 8 experts (an expert-parallel rank: few experts, so the extra tiles are what fills 148 SMs;
  // E = 4..8 measured equal or faster up to 128 rows per expert)
  const long long pairs = (long long)T * k;
  const bool small = pairs <= (long long)kDecMaxMeanRows * E || (E <= 8 && pairs <= 128ll * E);
  if (!wide) return small ? 32 : 160;
  // nested absmax (12 KB of absmax stages instead of 46 KB) leaves room for 64-row token tiles in the decode kernel: from a mean
  // of kDec64MinMeanRows rows per expert a good share of the experts would otherwise need two 32-row tiles, i.e. decode their
  // weights twice (round 2, T = 256 at E = 32: 183 us per layer with 32-row tiles)
  if (pairs <= (long long)kDec64MinMeanRows * E) return 32;
  if (pairs <= (long long)kDec64MaxMeanRows * E || (E <= 8 && pairs <= 160ll * E)) return 64;
  return 160;
}

template <bool W13>
int gemm_launch(const void* act_in, const nf4moe_qweight* w, const int32_t* sorted_pairs, const int32_t* meta,
                const int32_t* tile_list, int T, int k, int E, int act, int mode, void* out, void* stream, char* err,
                size_t err_len, int* launches, int32_t* status = nullptr, int tile_rows = 0) {
  const char* name = W13 ? "gemm_w13" : "gemm_w2";
  auto fail = [&](int code, const char* msg) { snprintf(err, err_len, "%s: %s", name, msg); return code; };
  if (!act_in || !w || !w->packed || !meta || !tile_list || !out || T < 0 || k <= 0 || E <= 0) return fail(NF4MOE_EINVAL, "bad arguments");
  if (!W13 && !sorted_pairs) return fail(NF4MOE_EINVAL, "sorted_pairs required");
  if (mode != NF4MOE_GEMM_EXACT && mode != NF4MOE_GEMM_FAST) return fail(NF4MOE_EINVAL, "mode must be NF4MOE_GEMM_EXACT or NF4MOE_GEMM_FAST");
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
  a.tile_tokens = tile_rows > 0 ? tile_rows : gemm_tile_tokens(T, k, E);   // stand-alone entry points: the size nf4moe_bucket(tile_rows = 0) used
  a.kps = a.tile_tokens <= 32 ? 4 : 2;
  a.n_a = std::min(kMaxAStages, (512 - 2 * a.tile_tokens) / (32 * a.kps));
  a.want_b = a.tile_tokens <= 32 ? 12 : 4;
  // a dequant group must meet every use of an A slot it waits on (mbarrier parity waits are only sound for a waiter that saw the
  // previous phase complete): the ring of n_a x kps k-blocks has to be a whole number of rounds over the groups
  if ((a.n_a * a.kps) % kDqGroups) return fail(NF4MOE_EINVAL, "token-tile size leaves an A-operand ring the dequant groups do not divide");
  a.dbg = gemm_debug_buffer();
  a.status = status ? status : const_cast<int32_t*>(meta) + 4;
  {   // the dynamic shared memory starts 1 KB (static control block) or 2 KB (+ the driver's reserved 1 KB) into the window
    GemmPlan plan;
    for (uint32_t base : {1024u, 2048u})
      if (!gemm_plan(base, kGemmLutBytes, kGemmSmemBytes, a.w.nkb, w->absmax_f32 == nullptr, a.tile_tokens, a.kps, a.n_a, a.want_b, &plan))
        return fail(NF4MOE_EINVAL, "K is too large for the grouped GEMM's shared-memory plan with this absmax form");
  }
  CUtensorMap map_w, map_x;
  if (!make_map_2d(&map_w, w->packed, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, (uint64_t)w->cols / 2, (uint64_t)E * w->rows, 128, 128))
    return fail(NF4MOE_EINVAL, "cuTensorMapEncodeTiled failed for the packed weights");
  if (!make_map_2d(&map_x, act_in, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, (uint64_t)w->cols, (uint64_t)T * k, kBlock, kGemmBoxRows))
    return fail(NF4MOE_EINVAL, "cuTensorMapEncodeTiled failed for the activations");
  CUtensorMap map_xw;   // one CTA's share of a full tile: boxes [tile_tokens / cluster x 64]
  if (a.tile_tokens % (8 * kGemmCluster)) return fail(NF4MOE_EINVAL, "tile rows must split into 1 KB-aligned shares across the cluster");
  if (!make_map_2d(&map_xw, act_in, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, (uint64_t)w->cols, (uint64_t)T * k, kBlock,
                   (uint32_t)(a.tile_tokens / kGemmCluster)))
    return fail(NF4MOE_EINVAL, "cuTensorMapEncodeTiled failed for the activation tiles");
  const bool fast = mode == NF4MOE_GEMM_FAST;
  cudaError_t e = fast ? ensure_dyn_smem<gemm_kernel<W13, true>>((int)kGemmSmemBytes) : ensure_dyn_smem<gemm_kernel<W13, false>>((int)kGemmSmemBytes);
  if (e != cudaSuccess) { snprintf(err, err_len, "%s: cudaFuncSetAttribute: %s", name, cudaGetErrorString(e)); return (int)e; }
  int n_clusters = sms / kGemmCluster;
  if (kGemmCluster > 1) {
    // clusters live inside one GPC: ask how many can be resident at once (a GPC with an SM count that C does not divide
    // strands SMs) and launch exactly that many, so that no cluster waits for another to finish
    static std::atomic<int> cached[64];
    int c = dev < 64 ? cached[dev].load(std::memory_order_acquire) : 0;
    if (c == 0) {
      cudaLaunchConfig_t q = {};
      q.gridDim = dim3((unsigned)(n_clusters * kGemmCluster));
      q.blockDim = dim3(kGemmThreads);
      q.dynamicSmemBytes = kGemmSmemBytes;
      cudaLaunchAttribute qa[1];
      qa[0].id = cudaLaunchAttributeClusterDimension;
      qa[0].val.clusterDim.x = kGemmCluster;
      qa[0].val.clusterDim.y = 1;
      qa[0].val.clusterDim.z = 1;
      q.attrs = qa;
      q.numAttrs = 1;
      int got = 0;
      cudaError_t qe = fast ? cudaOccupancyMaxActiveClusters(&got, gemm_kernel<W13, true>, &q) : cudaOccupancyMaxActiveClusters(&got, gemm_kernel<W13, false>, &q);
      if (qe != cudaSuccess || got <= 0) { cudaGetLastError(); got = n_clusters; }
      c = std::min(got, n_clusters);
      if (dev < 64) cached[dev].store(c, std::memory_order_release);
    }
    n_clusters = c;
  }
  const dim3 grid((unsigned)(n_clusters * kGemmCluster));
  if (fast) e = launch_pdl_cluster(gemm_kernel<W13, true>, grid, dim3(kGemmThreads), kGemmSmemBytes, (cudaStream_t)stream, kGemmCluster, a, map_w, map_x, map_xw);
  else e = launch_pdl_cluster(gemm_kernel<W13, false>, grid, dim3(kGemmThreads), kGemmSmemBytes, (cudaStream_t)stream, kGemmCluster, a, map_w, map_x, map_xw);
  ++*launches;
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e != cudaSuccess) { snprintf(err, err_len, "%s: %s", name, cudaGetErrorString(e)); return (int)e; }
  return 0;
}

}  // namespace nf4moe
