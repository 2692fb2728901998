// Decode-regime grouped GEMM on tcgen05: the HBM-bound kernel for 8 < T <= ~200 tokens.
//
// Same arithmetic as the FAST mode of gemm_tc.cuh (weights = A operand in TENSOR MEMORY, one dequant thread per weight
// row, byte -> half2 table look-up, one HMUL2 by fp16(absmax), activations = B operand in shared memory), but the
// pipeline is built for streaming 12.4 MB of packed NF4 per expert at HBM speed rather than for tensor throughput:
//   * an item = 128 weight rows of one expert x all K, for the expert's <= 32 routed tokens (MMA N = 16 or 32);
//   * a STAGE = 4 k-blocks: one 16 KB TMA box of raw NF4 [128 rows x 128 B, 128B swizzle] + the matching activation
//     tiles [4 x N x 64] in the canonical K-major UMMA layout, all on one mbarrier; 3-6 stages in flight per SM;
//   * SIXTEEN dequant warps: warp = (TMEM lane quarter q, k-block pair of the stage, stage parity).  Every warp does
//     the same thing per stage it owns: two k-blocks of its 32 rows = 4 LDS.128 + 64 x (PRMT, LDS, HMUL2) +
//     2 x tcgen05.st.x32, one barrier wait in, two arrivals out.  No k-split partial sums, no reductions: the fp32
//     accumulators live in TMEM for the whole K;
//   * one MMA-issuing thread: per stage ONE wait, <= 16 tcgen05.mma (M 128, N 16/32, K 16), TWO commits (stage + A slot);
//   * four epilogue warps read the finished accumulator (double-buffered) while the next item streams: gpt-oss
//     clamped SwiGLU on adjacent lanes for w13, scatter to (token, slot) order for w2.
// Instruction budget: ~19 weights per issued instruction (the mma.sync GEMV of gemv.cuh: 8.3, measured), shared-memory
// wavefronts ~27 per 512 packed bytes (16 of them the table look-ups, which bound this design at ~5.5 TB/s).
// Replaces bnb Linear4bit on single rows, P/unsloth_to_vllm_converter.py:309-318, for all routed pairs at once.
#pragma once
#include "gemm_tc.cuh"

namespace nf4moe {

constexpr int kDecDqWarps = 16;
constexpr int kDecThreads = (kDecDqWarps + 4 + 4) * 32;   // 768: dequant, epilogue warpgroup, control warpgroup
constexpr int kDecMaxStages = 6;
constexpr int kDecSlots = 3;                               // A-operand slots in TMEM (one stage = 4 k-blocks = 128 columns each)
constexpr int kDecAccCols = 32;                            // per accumulator (two of them)
constexpr uint32_t kDecRawBytes = 128 * 128;               // raw NF4 of one stage
constexpr uint32_t kDecSmemBytes = 232448 - 1024;

struct DecArgs {
  QW w;
  void* out;
  const int32_t* sorted_pairs;
  const int32_t* meta;
  const int32_t* tile_list;
  int act;
  long long* dbg;
};

// barrier slots
constexpr int kDbWFull = 0, kDbWEmpty = kDbWFull + kDecMaxStages, kDbAFull = kDbWEmpty + kDecMaxStages,
              kDbAEmpty = kDbAFull + kDecSlots, kDbDFull = kDbAEmpty + kDecSlots, kDbDEmpty = kDbDFull + 2,
              kDbMFull = kDbDEmpty + 2, kDbMEmpty = kDbMFull + 2, kNumDecBars = kDbMEmpty + 2;

// Warp roles (768 threads, one persistent CTA per SM):
//   warps 0-15   dequant: q = warp & 3 (TMEM lane quarter = weight rows 32q..32q+31), hh = (warp >> 2) & 1 (k-blocks 2hh, 2hh+1
//                of a stage), pp = warp >> 3 (stages with running index cs = pp mod 2)
//   warps 16-19  epilogue
//   warp 20      TMA producer (absmax per item, raw NF4 + activations per stage)
//   warp 21      TMEM allocator and the MMA-issuing thread
template <bool W13>
__global__ void __launch_bounds__(kDecThreads, 1) dec_kernel(const DecArgs a, const __grid_constant__ CUtensorMap map_w,
                                                            const __grid_constant__ CUtensorMap map_x16,
                                                            const __grid_constant__ CUtensorMap map_x32) {
  extern __shared__ __align__(1024) uint8_t dyn_smem[];
  __shared__ __align__(8) uint64_t s_bar[kNumDecBars];
  __shared__ uint32_t s_tmem;
  __shared__ uint32_t s_st[kDecMaxStages], s_mst[2];
  __shared__ int s_nst;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nkb = a.w.nkb;
  const int nst = (nkb + 3) >> 2;                       // stages per item (the last one may be partly padding)
  const bool nested = a.w.absmax_f32 == nullptr;
  const int bmax = a.meta[3] <= 16 ? 16 : 32;           // largest MMA N of this launch (meta[3] = largest per-expert count)
  const uint32_t stage_bytes = kDecRawBytes + 4u * (uint32_t)bmax * 128u;
  const uint32_t m_tile = (uint32_t)(kGemmRowTile * nkb) * (nested ? 1u : 4u);
  const uint32_t m_code2 = (m_tile + 15u) & ~15u;
  const uint32_t m_bytes = (m_code2 + (nested ? 1024u : 0u) + 1023u) & ~1023u;
  const uint32_t base = smem_u32(dyn_smem);
  const uint32_t lut = pin((base + kGemmLutBytes - 1u) & ~(kGemmLutBytes - 1u));
  const uint32_t bar0 = pin(smem_u32(s_bar));
  auto bar = [&](int i) -> uint32_t { return bar0 + 8u * (uint32_t)i; };

  // ---- one-time setup ----
  if (tid == 0) {
    uint32_t cur[2] = {(base + 1023u) & ~1023u, lut + kGemmLutBytes};
    const uint32_t hi[2] = {lut, base + kDecSmemBytes};
    auto place = [&](uint32_t bytes) -> uint32_t {
      int best = -1;
      uint32_t best_room = 0;
      for (int r = 0; r < 2; ++r) {
        const uint32_t room = hi[r] > cur[r] ? hi[r] - cur[r] : 0u;
        if (room >= bytes && (best < 0 || room < best_room)) { best = r; best_room = room; }
      }
      if (best < 0) return 0u;
      const uint32_t at = cur[best];
      cur[best] += bytes;
      return at;
    };
    int im = 0, is = 0;
    for (int i = 0; i < 2; ++i) { const uint32_t at = place(m_bytes); if (at) s_mst[im++] = at; }
    for (int i = 0; i < kDecMaxStages; ++i) { const uint32_t at = place(stage_bytes); if (at) s_st[is++] = at; else break; }
    s_nst = (im == 2 && is >= 2) ? is : 0;
    for (int i = 0; i < kDecMaxStages; ++i) { mbar_init(bar(kDbWFull + i), 1); mbar_init(bar(kDbWEmpty + i), kDecDqWarps / 2 + 1); }
    for (int i = 0; i < kDecSlots; ++i) { mbar_init(bar(kDbAFull + i), kDecDqWarps / 2); mbar_init(bar(kDbAEmpty + i), 1); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar(kDbDFull + i), 1); mbar_init(bar(kDbDEmpty + i), 4);
      mbar_init(bar(kDbMFull + i), 1); mbar_init(bar(kDbMEmpty + i), kDecDqWarps);
    }
    fence_barrier_init();
  }
  for (int i = tid; i < 256 * 32; i += kDecThreads) {
    const int b = i >> 5, l = i & 31;
    sts32(lut + (uint32_t)b * 256u + (uint32_t)l * 4u, pack_half2(kNF4[b >> 4], kNF4[b & 15]));
  }
  if (warp == 21) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(kTmemCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;
  const int n_st = s_nst;
  const int n_row_tiles = (a.w.rows + kGemmRowTile - 1) / kGemmRowTile;
  const int n_items = n_st > 0 ? a.meta[2] * n_row_tiles : 0;
  const uint32_t acol0 = 2u * kDecAccCols;   // TMEM: accumulators at columns 0 and 32, A slots after them

  if (warp < kDecDqWarps) {
    // =============================== dequant: NF4 -> fp16 A operand in TMEM ===============================
    const int q = warp & 3, hh = (warp >> 2) & 1, pp = warp >> 3;
    const int r = q * 32 + lane;
    const uint32_t lut_lane = pin(lut | ((uint32_t)lane << 2));
    const uint32_t t_lane = tmem + ((uint32_t)(q * 32) << 16) + acol0 + (uint32_t)hh * 64u;
    const uint32_t swz = (uint32_t)(r & 7);
    const uint32_t wfull0 = bar(kDbWFull), wempty0 = bar(kDbWEmpty), afull0 = bar(kDbAFull), aempty0 = bar(kDbAEmpty);
    // running stage counter cs = it * nst + st; this warp owns cs = pp (mod 2); (sw, phw) / (slot, pha) follow cs
    int sw = pp, slot = pp, st = pp, it = 0;
    uint32_t phw = 0, pha = 0;
    for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++it) {
      const int tile = item / n_row_tiles, rt = item - tile * n_row_tiles;
      const int e = a.tile_list[4 * tile];
      const int n = rt * kGemmRowTile + r;
      const int mb = it & 1;
      // nested absmax: code2[u8] * absmax2[blk >> 8] + offset; this row's blocks touch at most two absmax2 entries
      float a2_lo = 0.f, a2_hi = 0.f, off_e = 0.f;
      long long blk_hi = 0;
      if (nested) {
        const long long blk0 = (long long)min(n, a.w.rows - 1) * nkb;
        const long long i2 = blk0 >> 8;
        a2_lo = __ldg(a.w.absmax2 + (long long)e * a.w.n2 + i2);
        a2_hi = __ldg(a.w.absmax2 + (long long)e * a.w.n2 + min(i2 + 1, a.w.n2 - 1));
        off_e = __ldg(a.w.offset + e);
        blk_hi = ((i2 + 1) << 8) - blk0;
      }
      mbar_wait(bar(kDbMFull + mb), (uint32_t)(it >> 1) & 1u);
      const uint32_t m_st = pin(s_mst[mb]);
      const uint32_t m_row = m_st + (uint32_t)(r * nkb) * (nested ? 1u : 4u);
      for (; st < nst; st += 2) {
        mbar_wait(wfull0 + 8u * (uint32_t)sw, phw);
        const uint32_t wrow = pin(s_st[sw]) + (uint32_t)r * 128u;
        const uint32_t t_slot = t_lane + (uint32_t)slot * 128u;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int kb = st * 4 + hh * 2 + j;
          const bool real = kb < nkb;
          uint32_t regs[32];
          if (real) {
            float am;
            if (nested) {
              const float c = lds_f32(m_st + m_code2 + 4u * lds_u8(m_row + (uint32_t)kb));
              am = __fadd_rn(__fmul_rn(c, kb < blk_hi ? a2_lo : a2_hi), off_e);
            } else {
              am = lds_f32(m_row + 4u * (uint32_t)kb);
            }
            const uint32_t c2 = (uint32_t)(hh * 2 + j) * 2u;
            const uint4 v0 = lds128(wrow + ((c2 ^ swz) << 4));
            const uint4 v1 = lds128(wrow + (((c2 + 1u) ^ swz) << 4));
            decode_block<true>(v0, v1, lut_lane, am, regs);
          }
          if (j == 0) {
            mbar_wait(aempty0 + 8u * (uint32_t)slot, pha ^ 1u);   // the MMAs that read this slot three stages ago are done
            tc_fence_after();
          }
          if (real) tmem_st32(t_slot + (uint32_t)j * 32u, regs);
        }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(afull0 + 8u * (uint32_t)slot);
          mbar_arrive(wempty0 + 8u * (uint32_t)sw);
        }
        sw += 2;
        if (sw >= n_st) { sw -= n_st; phw ^= 1u; }
        slot += 2;
        if (slot >= kDecSlots) { slot -= kDecSlots; pha ^= 1u; }
      }
      st -= nst;
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(kDbMEmpty + mb));
    }
  } else if (warp < kDecDqWarps + 4) {
    // =============================== epilogue: TMEM -> registers -> global, overlapped with the next item ===============
    const int q = warp & 3;
    const int r = q * 32 + lane;
    const uint32_t t_lane = tmem + ((uint32_t)(q * 32) << 16);
    int it = 0;
    for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++it) {
      const int tile = item / n_row_tiles, rt = item - tile * n_row_tiles;
      const int row0 = a.tile_list[4 * tile + 1], n_tok = a.tile_list[4 * tile + 2];
      const int n = rt * kGemmRowTile + r;
      const int db = it & 1;
      mbar_wait(bar(kDbDFull + db), (uint32_t)(it >> 1) & 1u);
      tc_fence_after();
      const bool row_ok = n < a.w.rows;
      uint32_t v[32];
      tmem_ld32(t_lane + (uint32_t)(db * kDecAccCols), v);
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(kDbDEmpty + db));   // the accumulator is in registers: hand it back before the stores
      if (W13 && a.act == NF4MOE_ACT_SWIGLU_OAI) {
        // rows 2j / 2j+1 (gate / up) sit in adjacent lanes; products rounded to fp16 first (the reference stores the GEMM
        // result as fp16 before the activation); even lane takes tokens 0..15, odd lane tokens 16..31 of the tile
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
        const int t0 = odd ? 16 : 0;
        __half* dst = h + (long long)(row0 + t0) * I + (n >> 1);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          if (row_ok && t0 + 2 * i < n_tok) {
            const uint32_t r2 = swiglu_oai_h2(gt[i], up[i]);
            dst[(long long)(2 * i) * I] = __ushort_as_half((unsigned short)(r2 & 0xFFFFu));
            if (t0 + 2 * i + 1 < n_tok) dst[(long long)(2 * i + 1) * I] = __ushort_as_half((unsigned short)(r2 >> 16));
          }
        }
      } else if (W13) {
        __half* raw = reinterpret_cast<__half*>(a.out);
#pragma unroll
        for (int i = 0; i < 32; ++i)
          if (row_ok && i < n_tok) raw[(long long)(row0 + i) * a.w.rows + n] = __float2half_rn(__uint_as_float(v[i]));
      } else {
        __half* y = reinterpret_cast<__half*>(a.out);
#pragma unroll
        for (int i = 0; i < 32; ++i)
          if (row_ok && i < n_tok)
            y[(long long)__ldg(a.sorted_pairs + row0 + i) * a.w.rows + n] = __float2half_rn(__uint_as_float(v[i]));
      }
    }
  } else if (warp == 20) {
    // =============================== TMA producer ===============================
    const uint64_t pol = l2_policy_evict_first();   // weights are read exactly once per step
    const uint32_t am_elt = nested ? 1u : 4u;
    const uint8_t* am_base = nested ? a.w.absmax_u8 : reinterpret_cast<const uint8_t*>(a.w.absmax_f32);
    const uint32_t wfull0 = bar(kDbWFull), wempty0 = bar(kDbWEmpty);
    int sw = 0, it = 0;
    uint32_t phw = 0;
    for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++it) {
      const int tile = item / n_row_tiles, rt = item - tile * n_row_tiles;
      const int e = a.tile_list[4 * tile], row0 = a.tile_list[4 * tile + 1], n_tok = a.tile_list[4 * tile + 2];
      const int nb = n_tok <= 16 ? 16 : 32;
      const uint32_t b_tile = (uint32_t)nb * 128u;
      const CUtensorMap* map_x = nb == 16 ? &map_x16 : &map_x32;
      const int c1 = e * a.w.rows + rt * kGemmRowTile;
      {
        const int mb = it & 1;
        const int rows_here = min(kGemmRowTile, a.w.rows - rt * kGemmRowTile);
        const uint32_t bytes = (uint32_t)(rows_here * nkb) * am_elt;
        mbar_wait(bar(kDbMEmpty + mb), ((uint32_t)(it >> 1) & 1u) ^ 1u);
        if (elect_one()) {
          mbar_expect_tx(bar(kDbMFull + mb), bytes + (nested ? 1024u : 0u));
          bulk_g2s(s_mst[mb], am_base + ((long long)e * a.w.nb + (long long)rt * kGemmRowTile * nkb) * am_elt, bytes,
                   bar(kDbMFull + mb), pol);
          if (nested) bulk_g2s(s_mst[mb] + m_code2, a.w.code2 + (long long)e * a.w.code2_stride, 1024u, bar(kDbMFull + mb), pol);
        }
        __syncwarp();
      }
      for (int st = 0; st < nst; ++st) {
        mbar_wait(wempty0 + 8u * (uint32_t)sw, phw ^ 1u);
        const int nreal = min(4, nkb - st * 4);
        if (elect_one()) {
          const uint32_t full = wfull0 + 8u * (uint32_t)sw, dst = s_st[sw];
          mbar_expect_tx(full, kDecRawBytes + (uint32_t)nreal * b_tile);
          tma_load_2d(dst, &map_w, st * 128, c1, full);
          for (int u = 0; u < nreal; ++u)
            tma_load_2d(dst + kDecRawBytes + (uint32_t)u * b_tile, map_x, (st * 4 + u) * kBlock, row0, full);
        }
        __syncwarp();
        if (++sw == n_st) { sw = 0; phw ^= 1u; }
      }
    }
  } else if (warp == 21) {
    // =============================== MMA issuer (one thread) ===============================
    const uint32_t wfull0 = bar(kDbWFull), wempty0 = bar(kDbWEmpty), afull0 = bar(kDbAFull), aempty0 = bar(kDbAEmpty);
    const uint32_t desc_hi = (1024u >> 4) | (1u << 14) | (2u << 29);   // SBO 1024 B, descriptor version 1, SWIZZLE_128B
    int sw = 0, slot = 0, it = 0;
    uint32_t phw = 0, pha = 0;
    for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++it) {
      const int tile = item / n_row_tiles;
      const int n_tok = a.tile_list[4 * tile + 2];
      const int nb = n_tok <= 16 ? 16 : 32;
      const uint32_t idesc = make_idesc(nb);
      const uint32_t blo_step = ((uint32_t)nb * 128u) >> 4;
      const int db = it & 1;
      const uint32_t d_t = tmem + (uint32_t)(db * kDecAccCols);
      mbar_wait(bar(kDbDEmpty + db), (((uint32_t)(it >> 1)) & 1u) ^ 1u);   // the epilogue has drained this accumulator
      tc_fence_after();
      for (int st = 0; st < nst; ++st) {
        mbar_wait(wfull0 + 8u * (uint32_t)sw, phw);      // activations landed (TMA)
        mbar_wait(afull0 + 8u * (uint32_t)slot, pha);    // A operand written (8 dequant warps)
        tc_fence_after();
        const int nreal = min(4, nkb - st * 4);
        if (elect_one()) {
          const uint32_t b_lo = ((s_st[sw] + kDecRawBytes) & 0x3FFFFu) >> 4;
          const uint32_t a_t = tmem + acol0 + (uint32_t)slot * 128u;
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            if (u < nreal) {
              const uint64_t desc = ((uint64_t)desc_hi << 32) | (uint64_t)(b_lo + (uint32_t)u * blo_step);
#pragma unroll
              for (int kk = 0; kk < 4; ++kk)
                tc_mma_ts(d_t, a_t + (uint32_t)(u * 32 + 8 * kk), desc + (uint64_t)(2 * kk), idesc, (uint32_t)(st | u | kk));
            }
          }
          tc_commit(wempty0 + 8u * (uint32_t)sw);        // the stage's activations are consumed
          tc_commit(aempty0 + 8u * (uint32_t)slot);      // ... and so is the A slot
          if (st == nst - 1) tc_commit(bar(kDbDFull + db));
        }
        __syncwarp();
        if (++sw == n_st) { sw = 0; phw ^= 1u; }
        if (++slot == kDecSlots) { slot = 0; pha ^= 1u; }
      }
    }
  }

  // ---- teardown ----
  tc_fence_before();
  __syncthreads();
  if (warp == 21) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(kTmemCols) : "memory");
  }
}

// ---- host side ------------------------------------------------------------------------------------------------------
// True when the decode-regime tcgen05 kernel serves this call: FAST operand mode and token tiles of at most 32 rows.
inline bool dec_serves(int T, int k, int E, int mode) { return mode == NF4MOE_GEMM_FAST && gemm_tile_tokens(T, k, E) <= 32; }

template <bool W13>
int dec_launch(const void* act_in, const nf4moe_qweight* w, const int32_t* sorted_pairs, const int32_t* meta,
               const int32_t* tile_list, int T, int k, int E, int act, void* out, void* stream, char* err, size_t err_len,
               int* launches) {
  const char* name = W13 ? "gemm_w13(decode)" : "gemm_w2(decode)";
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
  DecArgs a;
  a.w = make_qw(*w);
  a.out = out;
  a.sorted_pairs = sorted_pairs;
  a.meta = meta;
  a.tile_list = tile_list;
  a.act = act;
  a.dbg = nullptr;
  CUtensorMap map_w, map_x16, map_x32;
  if (!make_map_2d(&map_w, w->packed, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, (uint64_t)w->cols / 2, (uint64_t)E * w->rows, 128, 128))
    return fail(NF4MOE_EINVAL, "cuTensorMapEncodeTiled failed for the packed weights");
  if (!make_map_2d(&map_x16, act_in, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, (uint64_t)w->cols, (uint64_t)T * k, kBlock, 16) ||
      !make_map_2d(&map_x32, act_in, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, (uint64_t)w->cols, (uint64_t)T * k, kBlock, 32))
    return fail(NF4MOE_EINVAL, "cuTensorMapEncodeTiled failed for the activations");
  auto kern = dec_kernel<W13>;
  static thread_local bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kDecSmemBytes);
    if (e != cudaSuccess) { snprintf(err, err_len, "%s: cudaFuncSetAttribute: %s", name, cudaGetErrorString(e)); return (int)e; }
    attr_set = true;
  }
  kern<<<sms, kDecThreads, kDecSmemBytes, (cudaStream_t)stream>>>(a, map_w, map_x16, map_x32);
  ++*launches;
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { snprintf(err, err_len, "%s: %s", name, cudaGetErrorString(e)); return (int)e; }
  return 0;
}

}  // namespace nf4moe
