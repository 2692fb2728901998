// Decode-regime grouped GEMV: the HBM-bound kernel of the path.
//
// One persistent CTA per SM walks a contiguous range of work items; an item is 16 consecutive weight rows of one
// (expert, token-chunk) = one contiguous span of packed NF4 bytes (16*K/2) plus its absmax (16*K/64 values).
//   * warp 15 (one lane) is the producer: TMA 1-D bulk copies (cp.async.bulk, SASS UBLKCP) of the raw packed bytes
//     and absmax into a shared-memory ring, completion on mbarriers, L2 evict-first (weights are read exactly once);
//   * warps 0..14 are consumers, grouped in TEAMS that own one item at a time.  Inside a team warp w owns KBW k-blocks
//     (team of 15 x 3 blocks or 5 x 9 blocks = 45 blocks = K 2880 = H = I of gpt-oss), keeps the activations of those
//     k for up to 8*NT routed tokens in REGISTERS as mma B fragments for the whole expert, reads its 8 packed bytes per
//     row per block with one LDS.64, decodes each byte through a 64 KB bank-conflict-free shared-memory LUT
//     (byte -> half2 of the two NF4 codes; one PRMT forms the address, one LDS fetches), feeds mma.sync.m16n8k16
//     (weights = A, tokens = B) and applies the fp32 absmax to the per-block fp32 partial sums;
//   * the team's k-slices are summed through shared memory (mbarrier hand-off, no CTA-wide barrier); the reducer warp
//     applies the fused epilogue: gpt-oss interleaved clamped SwiGLU for w13 (rows 2j/2j+1 = gate/up,
//     V/model_executor/layers/activation.py:389-397) or the scatter back to (token,slot) order for w2.
// Experts with more than 8*NT routed tokens are split into "virtual experts" of <= 8*NT tokens that each stream the
// weights once (rare in the decode regime; the prefill regime uses the tcgen05 kernel instead).
// Replaces bnb Linear4bit on single rows, P/unsloth_to_vllm_converter.py:309-318, for all routed pairs at once.
#pragma once
#include "common.cuh"

namespace nf4moe {

constexpr int kGemvConsumers = 15;
constexpr int kGemvThreads = (kGemvConsumers + 1) * 32;
constexpr int kGemvRows = 16;
constexpr int kGemvMaxK = 45 * kBlock;  // 2880
constexpr uint32_t kLutBytes = 65536;
constexpr int kMaxStages = 8;
constexpr int kNumPartBars = 32;   // >= teams * buffers * 2
constexpr int kMaxVirt = 1024;     // (expert, token-chunk) entries
constexpr uint32_t kGemvSmemBytes = 232448;  // 227 KB opt-in maximum

struct GemvArgs {
  QW w;
  const __half* x;         // activations, row stride = w.cols
  void* out;               // W13: h_sorted fp16 [n_pairs, I] (or raw [n_pairs, 2I]); W2: y fp16 [n_pairs, rows]
  const int32_t* expert_offsets;
  const int32_t* sorted_pairs;
  const int32_t* active_experts;
  const int32_t* meta;
  int k;
  int act;                 // W13 only: 0 fused swiglu_oai, 1 raw store (activation applied by a second kernel)
};

// One 64-wide k-block of a 16-row item for NTA token tiles: 16 table look-ups -> 4 x NTA HMMA -> fp32 absmax scaling.
template <int NTA, int NT, int KBW>
__device__ __forceinline__ void block_step(const uint2 w0, const uint2 w1, const uint32_t lut_lane,
                                           const uint32_t (&b)[NT][KBW][8], const int j, const float am0,
                                           const float am1, float (&acc)[NT][4]) {
  float d[NTA][2][4];  // two independent accumulator chains per tile (m even / odd) halve the HMMA dependency depth
#pragma unroll
  for (int nt = 0; nt < NTA; ++nt)
#pragma unroll
    for (int c = 0; c < 2; ++c) d[nt][c][0] = d[nt][c][1] = d[nt][c][2] = d[nt][c][3] = 0.f;
#pragma unroll
  for (int m = 0; m < 4; ++m) {
    const uint32_t s0w = m < 2 ? w0.x : w0.y, s1w = m < 2 ? w1.x : w1.y;
    const uint32_t selA = 0x7604u | ((uint32_t)((2 * m) & 3) << 4);
    const uint32_t selB = 0x7604u | ((uint32_t)((2 * m + 1) & 3) << 4);
    const uint32_t a0 = lds32(prmt(s0w, lut_lane, selA));  // row g  , k 16t+4m   , +1
    const uint32_t a1 = lds32(prmt(s1w, lut_lane, selA));  // row g+8, same k
    const uint32_t a2 = lds32(prmt(s0w, lut_lane, selB));  // row g  , k 16t+4m+2 , +3
    const uint32_t a3 = lds32(prmt(s1w, lut_lane, selB));  // row g+8
#pragma unroll
    for (int nt = 0; nt < NTA; ++nt) mma_16816(d[nt][m & 1], a0, a1, a2, a3, b[nt][j][2 * m], b[nt][j][2 * m + 1]);
  }
#pragma unroll
  for (int nt = 0; nt < NTA; ++nt) {
    acc[nt][0] = fmaf(am0, d[nt][0][0] + d[nt][1][0], acc[nt][0]);
    acc[nt][1] = fmaf(am0, d[nt][0][1] + d[nt][1][1], acc[nt][1]);
    acc[nt][2] = fmaf(am1, d[nt][0][2] + d[nt][1][2], acc[nt][2]);
    acc[nt][3] = fmaf(am1, d[nt][0][3] + d[nt][1][3], acc[nt][3]);
  }
}

// NT    : token tiles (8 tokens each) whose activations a warp keeps in registers
// KBW   : k-blocks per warp; a TEAM of 45/KBW warps covers K = 2880 and owns one item at a time; 15/(45/KBW) teams/CTA
//         (KBW=3: one 15-warp team, NT<=2;  KBW=9: three 5-warp teams, NT=1 -- 3x fewer, 3x longer k-slices per item)
// NKB_T : compile-time K/64 (45 for gpt-oss: every bound and offset becomes an immediate) or 0 for a runtime K.
template <int NT, int KBW, bool W13, bool NESTED, int NKB_T>
__global__ void __launch_bounds__(kGemvThreads, 1) gemv_kernel(const GemvArgs a) {
  extern __shared__ __align__(1024) uint8_t dyn_smem[];
  constexpr int TW = 45 / KBW;                 // warps per team
  constexpr int NTEAMS = kGemvConsumers / TW;  // teams per CTA
  constexpr int NRED = 4 * NT;                 // reducer warps per item (one table hole each)
  constexpr int NB = 240 / (NTEAMS * TW * NRED) >= 4 ? 4 : 240 / (NTEAMS * TW * NRED);  // partial buffers per team
  static_assert(NB >= 2 && NRED <= TW && NTEAMS * NB * 2 <= kNumPartBars, "bad team configuration");
  grid_dep_launch();
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nkb = NKB_T ? NKB_T : a.w.nkb;
  const uint32_t row_bytes = (uint32_t)nkb * 32u;
  const uint32_t w_bytes = kGemvRows * row_bytes;
  const uint32_t a_bytes = kGemvRows * (uint32_t)nkb * (NESTED ? 1u : 4u);
  const uint32_t stage_bytes = (w_bytes + a_bytes + 127u) & ~127u;

  // ---- carve shared memory around a 64 KB-ALIGNED lookup table (the PRMT address trick needs bits 0..15 free).
  // Table row b (256 B) = 32 lanes x half2 in its first 128 B; the second 128 B of each row is a "hole"; the holes
  // carry the k-split partial sums (teams x NB buffers x TW warps x NRED holes <= 240), so the table costs nothing extra.
  constexpr uint32_t kBarBytes = (2u * kMaxStages + kNumPartBars) * 8u;
  constexpr uint32_t kCtlBytes = kBarBytes + 128u + kMaxVirt * 4u;   // barriers, n_virt, virtual-expert list
  const uint32_t base = smem_u32(dyn_smem);
  const uint32_t lut = pin((base + kLutBytes - 1) & ~(kLutBytes - 1));
  uint32_t a_lo = base + kCtlBytes, a_hi = lut;          // region A: below the table (after the control block)
  const uint32_t b_lo = lut + kLutBytes;                 // region B: above it
  uint32_t b_hi = base + kGemvSmemBytes;
  uint32_t ctl = base;
  if (lut - base < kCtlBytes) {                          // table at the very start: control block goes to the top end
    b_hi -= kCtlBytes;
    ctl = b_hi;
    a_lo = a_hi;
  }
  a_lo = (a_lo + 127u) & ~127u;
  const int n_a = a_hi > a_lo ? (int)((a_hi - a_lo) / stage_bytes) : 0;
  const int n_b = (int)((b_hi - b_lo) / stage_bytes);
  const int n_stages = min(kMaxStages, n_a + n_b);
  const uint32_t bar_full = pin(ctl), bar_empty = bar_full + kMaxStages * 8u;
  const uint32_t bar_part = bar_empty + kMaxStages * 8u;  // per team: NB "full" then NB "empty"
  int* s_nvirt = reinterpret_cast<int*>(dyn_smem + (ctl - base) + kBarBytes);
  const uint32_t* s_virt = reinterpret_cast<const uint32_t*>(dyn_smem + (ctl - base) + kBarBytes + 128u);
  auto stage_addr = [&](int si) -> uint32_t {
    return si < n_a ? a_lo + (uint32_t)si * stage_bytes : b_lo + (uint32_t)(si - n_a) * stage_bytes;
  };

  for (int i = tid; i < 256 * 32; i += kGemvThreads) {
    const int b = i >> 5, l = i & 31;
    sts32(lut + (uint32_t)b * 256u + (uint32_t)l * 4u, pack_half2(kNF4[b >> 4], kNF4[b & 15]));
  }
  if (tid == 0) {
    for (int s = 0; s < kMaxStages; ++s) {
      mbar_init(bar_full + 8u * s, 1);
      mbar_init(bar_empty + 8u * s, TW);
    }
    for (int q = 0; q < NTEAMS; ++q)
      for (int b = 0; b < NB; ++b) {
        mbar_init(bar_part + 8u * (q * 2 * NB + b), TW);
        mbar_init(bar_part + 8u * (q * 2 * NB + NB + b), NRED);
      }
    fence_barrier_init();
  }
  // launched with launch_pdl(): table fill and barrier setup above overlap the previous kernel's tail; everything below reads
  // what the bucket (and, for w2, the w13) kernel wrote
  grid_dep_wait();
  if (warp == 1) {
    // virtual experts: (expert e, chunk c) for every active expert and every 8*NT-token chunk of its routed rows
    uint32_t* virt = const_cast<uint32_t*>(s_virt);
    const int n_act = a.meta[0];
    int total = 0;
    if (a.meta[3] <= 8 * NT) {
      // no expert has more rows than one chunk holds (always so under AUTO, which sends at most 8 tokens here): the virtual experts ARE
      // the active experts -- one parallel load instead of the dependent active_experts -> expert_offsets round trips below
      for (int ai = lane; ai < n_act && ai < kMaxVirt; ai += 32) virt[ai] = (uint32_t)a.active_experts[ai];
      total = n_act;
    } else
    for (int a0 = 0; a0 < n_act; a0 += 32) {
      const int ai = a0 + lane;
      const int e = ai < n_act ? a.active_experts[ai] : 0;
      const int m = ai < n_act ? a.expert_offsets[e + 1] - a.expert_offsets[e] : 0;
      const int nc = (m + 8 * NT - 1) / (8 * NT);
      int inc = nc;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int n = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += n;
      }
      const int pos = total + inc - nc;
      for (int c = 0; c < nc; ++c)
        if (pos + c < kMaxVirt) virt[pos + c] = (uint32_t)e | ((uint32_t)c << 16);
      total += __shfl_sync(0xffffffffu, inc, 31);
    }
    if (lane == 0) *s_nvirt = min(total, kMaxVirt);
  }
  __syncthreads();

  // ---- work range: items = (virtual expert, 16-row group), contiguous per CTA, dealt round-robin to the teams ----
  const int n_groups = a.w.rows / kGemvRows;
  const long long n_items = (long long)(*s_nvirt) * n_groups;
  const long long it_lo = n_items * blockIdx.x / gridDim.x;
  const int n_my = (int)(n_items * (blockIdx.x + 1) / gridDim.x - it_lo);
  if (n_my <= 0) return;

  constexpr uint32_t kBufStride = (uint32_t)(TW * NRED) * 256u;   // one partial buffer of one team
  if (warp == kGemvConsumers) {
    // =============================== producer: one lane issues TMA bulk copies ===============================
    if (lane == 0) {
      const uint64_t pol = l2_policy_evict_first();
      const uint32_t am_elt = NESTED ? 1u : 4u;
      const uint8_t* am_base = NESTED ? a.w.absmax_u8 : reinterpret_cast<const uint8_t*>(a.w.absmax_f32);
      int v = (int)(it_lo / n_groups);
      int grp = (int)(it_lo - (long long)v * n_groups);
      int s = 0;
      uint32_t ph = 0, st = stage_addr(0);
      int e = (int)(s_virt[v] & 0xffffu);
      const uint8_t* wsrc = a.w.packed + (long long)e * a.w.packed_stride + (long long)grp * w_bytes;
      const uint8_t* asrc = am_base + ((long long)e * a.w.nb + (long long)grp * kGemvRows * nkb) * am_elt;
      for (int it = 0; it < n_my; ++it) {
        mbar_wait(bar_empty + 8u * s, ph ^ 1u);
        mbar_expect_tx(bar_full + 8u * s, w_bytes + a_bytes);
        bulk_g2s(st, wsrc, w_bytes, bar_full + 8u * s, pol);
        bulk_g2s(st + w_bytes, asrc, a_bytes, bar_full + 8u * s, pol);
        if (++s == n_stages) { s = 0; ph ^= 1u; }
        st = stage_addr(s);
        if (++grp == n_groups) {
          grp = 0;
          ++v;
          if (it + 1 < n_my) {
            e = (int)(s_virt[v] & 0xffffu);
            wsrc = a.w.packed + (long long)e * a.w.packed_stride;
            asrc = am_base + (long long)e * a.w.nb * am_elt;
          }
        } else {
          wsrc += w_bytes;
          asrc += a_bytes;
        }
      }
    }
    return;
  }

  // =============================== consumers ===============================
  const int team = warp / TW, wt = warp - team * TW;
  const int g = lane >> 2, t = lane & 3;
  const uint32_t lut_lane = pin(lut | ((uint32_t)lane << 2));
  const int kb0 = wt * KBW;                                        // first k-block of this warp
  const int nkb_w = NKB_T ? KBW : min(KBW, max(0, nkb - kb0));     // how many of its blocks exist for this K
  const uint32_t w_off = (uint32_t)g * row_bytes + (uint32_t)kb0 * 32u + (uint32_t)t * 8u;
  const uint32_t am_off = w_bytes + (uint32_t)(g * nkb + kb0) * 4u;          // flat form, row g (row g+8: + 32*nkb)
  const uint32_t u8_off = w_bytes + (uint32_t)((lane & 15) * nkb + kb0 + (lane >> 4));  // nested form, conversion lane
  // partial sums: team q, buffer f, warp w, hole h (= nt*4 + t), word p -> lut + 128 + (((q*NB+f)*TW + w)*NRED + h)*256 + p*4
  const uint32_t hole_team = pin(lut + 128u + (uint32_t)(team * NB) * kBufStride);
  const uint32_t hole_w = (uint32_t)(wt * NRED + t) * 256u;
  const uint32_t p0 = 4u * ((uint32_t)(g + 8 * t) & 31u), p1 = 4u * ((uint32_t)(g + 8 + 8 * t) & 31u);
  const uint32_t p2 = 4u * ((uint32_t)(g + 16 + 8 * t) & 31u), p3 = 4u * ((uint32_t)(g + 24 + 8 * t) & 31u);
  const uint32_t bar_pfull = bar_part + 8u * (uint32_t)(team * 2 * NB), bar_pempty = bar_pfull + 8u * NB;

  int it = team;  // this team's first item
  if (it >= n_my) return;
  int v, grp;
  {
    const long long gi = it_lo + it;
    v = (int)(gi / n_groups);
    grp = (int)(gi - (long long)v * n_groups);
  }
  int s = it % n_stages;
  uint32_t ph = (uint32_t)(it / n_stages) & 1u;

  uint32_t breg[NT][KBW][8];
  int e = 0, s0 = 0, m_v = 0;   // expert, first sorted row and row count of the current virtual expert
  const float* code2_e = nullptr;
  const float* absmax2_e = nullptr;
  float off_e = 0.f;
  int pb = 0;                   // partial buffer in use
  uint32_t pph = 0;

  auto new_virtual_expert = [&]() {
    const uint32_t ve = s_virt[v];
    e = (int)(ve & 0xffffu);
    const int off = a.expert_offsets[e];
    s0 = off + (int)(ve >> 16) * 8 * NT;
    m_v = min(8 * NT, a.expert_offsets[e + 1] - s0);
    if (NESTED) {
      code2_e = a.w.code2 + (long long)e * a.w.code2_stride;
      absmax2_e = a.w.absmax2 + (long long)e * a.w.n2;
      off_e = __ldg(a.w.offset + e);
    }
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const int col = nt * 8 + g;
      const bool valid = col < m_v;
      long long src_row = 0;
      if (valid) src_row = W13 ? (long long)(a.sorted_pairs[s0 + col] / a.k) : (long long)(s0 + col);
#pragma unroll
      for (int j = 0; j < KBW; ++j) {
        uint4 v0 = make_uint4(0, 0, 0, 0), v1 = v0;
        if (valid && j < nkb_w) {
          const uint4* p = reinterpret_cast<const uint4*>(a.x + src_row * (nkb * kBlock) + (kb0 + j) * kBlock + 16 * t);
          v0 = p[0];
          v1 = p[1];
        }
        breg[nt][j][0] = v0.x; breg[nt][j][1] = v0.y; breg[nt][j][2] = v0.z; breg[nt][j][3] = v0.w;
        breg[nt][j][4] = v1.x; breg[nt][j][5] = v1.y; breg[nt][j][6] = v1.z; breg[nt][j][7] = v1.w;
      }
    }
  };
  new_virtual_expert();

  int rot = 0;                  // reducer rotation
  bool have_prev = false;
  int prev_pb = 0, prev_rot = 0, prev_grp = 0, prev_s0 = 0, prev_m = 0;
  uint32_t prev_pph = 0;
  auto reduce_item = [&](int rpb, uint32_t rpph, int rrot, int rgrp, int rs0, int rm) {
    int rw = wt - rrot;
    if (rw < 0) rw += TW;
    if (rw >= NRED) return;
    mbar_wait(bar_pfull + 8u * rpb, rpph);
    const uint32_t pr = pin(hole_team + (uint32_t)rpb * kBufStride + (uint32_t)rw * 256u + (uint32_t)lane * 4u);
    float v0 = 0.f, v1 = 0.f, v2 = 0.f;
#pragma unroll
    for (int w = 0; w < TW; ++w) {
      const float pv = lds_f32(pr + (uint32_t)(w * NRED) * 256u);
      if (w % 3 == 0) v0 += pv; else if (w % 3 == 1) v1 += pv; else v2 += pv;
    }
    float vs = (v0 + v1) + v2;
    asm volatile("" : "+f"(vs));  // the sums (hence the loads) are complete before the buffer is handed back
    __syncwarp();
    if (lane == 0) mbar_arrive(bar_pempty + 8u * rpb);
    const int tq = rw & 3, x = (lane - 8 * tq) & 31;
    const int row = x & 15, col = (rw >> 2) * 8 + 2 * tq + (x >> 4);
    const bool valid = col < rm;
    const int srow = rs0 + col;
    vs = __half2float(__float2half_rn(vs));  // fp16 storage of the product, as the oracle / reference does
    if (W13) {
      if (a.act == NF4MOE_ACT_SWIGLU_OAI) {
        const float up = __shfl_down_sync(0xffffffffu, vs, 1);
        if (valid && !(row & 1))
          reinterpret_cast<__half*>(a.out)[(long long)srow * (a.w.rows >> 1) + rgrp * 8 + (row >> 1)] =
              __float2half_rn(swiglu_oai(vs, up));
      } else if (valid) {
        reinterpret_cast<__half*>(a.out)[(long long)srow * a.w.rows + rgrp * kGemvRows + row] = __float2half_rn(vs);
      }
    } else if (valid) {
      reinterpret_cast<__half*>(a.out)[(long long)a.sorted_pairs[srow] * a.w.rows + rgrp * kGemvRows + row] =
          __float2half_rn(vs);
    }
  };

#pragma unroll 1
  for (;;) {
    mbar_wait(bar_full + 8u * s, ph);
    const uint32_t st = pin(stage_addr(s));   // every stage load below depends on this post-wait value

    const bool two = NT > 1 && m_v > 8;
    float acc[NT][4];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f;

    // nested absmax: bnb double-quant, two roundings: code2[u8] * absmax2[b >> 8] + offset.  16*KBW values per warp,
    // converted two k-blocks at a time by the 32 lanes (row = lane & 15, block = 2r + (lane >> 4)), handed out by shuffle.
    float cv[(KBW + 1) / 2];
    if (NESTED) {
      const int bidx = (grp * kGemvRows + (lane & 15)) * nkb + kb0 + (lane >> 4);
#pragma unroll
      for (int r = 0; r < (KBW + 1) / 2; ++r) {
        cv[r] = 0.f;
        if (2 * r + (lane >> 4) < nkb_w) {
          const uint32_t c = lds_u8(st + u8_off + 2u * r);
          cv[r] = __fadd_rn(__fmul_rn(__ldg(code2_e + c), __ldg(absmax2_e + ((bidx + 2 * r) >> 8))), off_e);
        }
      }
    }
#pragma unroll
    for (int j = 0; j < KBW; ++j) {
      if (j < nkb_w) {
        float am0, am1;
        if (NESTED) {
          am0 = __shfl_sync(0xffffffffu, cv[j >> 1], (j & 1) * 16 + g);
          am1 = __shfl_sync(0xffffffffu, cv[j >> 1], (j & 1) * 16 + g + 8);
        } else {
          am0 = lds_f32(st + am_off + 4u * j);
          am1 = lds_f32(st + am_off + 4u * j + 32u * (uint32_t)nkb);
        }
        const uint2 w0 = lds64(st + w_off + 32u * j);                   // row g   : k = (kb0+j)*64 + 16t .. +15
        const uint2 w1 = lds64(st + w_off + 32u * j + 8u * row_bytes);  // row g+8
        if (two) block_step<NT, NT, KBW>(w0, w1, lut_lane, breg, j, am0, am1, acc);
        else block_step<1, NT, KBW>(w0, w1, lut_lane, breg, j, am0, am1, acc);
      }
    }

    // ---- hand the k-slice partial sums to this item's reducer warps (mbarrier-decoupled, bank-conflict-free) ----
    mbar_wait(bar_pempty + 8u * pb, pph ^ 1u);
    const uint32_t hb = hole_team + (uint32_t)pb * kBufStride;
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const uint32_t hq = hb + hole_w + (uint32_t)nt * 1024u;
      sts32(hq + p0, __float_as_uint(acc[nt][0]));   // x = g      (row g  , col 2t)
      sts32(hq + p1, __float_as_uint(acc[nt][2]));   // x = g+8    (row g+8, col 2t)
      sts32(hq + p2, __float_as_uint(acc[nt][1]));   // x = 16+g   (row g  , col 2t+1)
      sts32(hq + p3, __float_as_uint(acc[nt][3]));   // x = 24+g   (row g+8, col 2t+1)
    }
    __syncwarp();
    if (lane == 0) {
      mbar_arrive(bar_pfull + 8u * pb);
      mbar_arrive(bar_empty + 8u * s);       // the stage's bytes are all consumed (they fed acc)
    }
    // DEFERRED reduction: sum and store the PREVIOUS item of this team now -- every warp of the team finished it at
    // least one item ago, so the wait below does not stall on stragglers.  Duty rotates over the team's warps.
    if (have_prev) reduce_item(prev_pb, prev_pph, prev_rot, prev_grp, prev_s0, prev_m);
    have_prev = true;
    prev_pb = pb; prev_pph = pph; prev_rot = rot; prev_grp = grp; prev_s0 = s0; prev_m = m_v;
    rot += NRED;
    if (rot >= TW) rot -= TW;
    if (++pb == NB) { pb = 0; pph ^= 1u; }

    // ---- next item of this team ----
    it += NTEAMS;
    if (it >= n_my) break;
    s += NTEAMS;
    while (s >= n_stages) { s -= n_stages; ph ^= 1u; }
    grp += NTEAMS;
    if (grp >= n_groups) {
      do { grp -= n_groups; ++v; } while (grp >= n_groups);
      new_virtual_expert();
    }
  }
  reduce_item(prev_pb, prev_pph, prev_rot, prev_grp, prev_s0, prev_m);
}

// reference-variant activation on a raw [n, 2I] buffer: silu(h[:, :I]) * h[:, I:]  (P/unsloth_to_vllm_converter.py:312-315)
__global__ void __launch_bounds__(256) chunk_silu_kernel(const __half* raw, const int32_t* meta,
                                                         int I, __half* __restrict__ h) {
  const long long total = (long long)ld_dep_s32(meta + 1) * I;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / I, c = i - r * I;
    h[i] = __float2half_rn(silu_mul(__half2float(raw[r * 2 * I + c]), __half2float(raw[r * 2 * I + I + c])));
  }
}

// Biased up-projection: raw [rows, 2I] (fp16 products of the w13 kernel) + w13_bias[e] -> activation -> h [rows, I].  The row's
// expert is ids[sorted_pairs[row]].  Replaces the bias add of gpt-oss's expert MLP (V/model_executor/models/gpt_oss.py:199),
// which neither the reference nor vLLM's bnb method has.  act: interleaved clamped SwiGLU or chunk + SiLU.
__global__ void __launch_bounds__(256) bias_act_kernel(const __half* raw, const __half* __restrict__ bias,
                                                       const int32_t* ids, const int32_t* sorted_pairs,
                                                       const int32_t* meta, int I, int act, __half* __restrict__ h) {
  grid_dep_launch();
  grid_dep_wait();
  const long long total = (long long)ld_dep_s32(meta + 1) * I;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / I, c = i - r * I;
    const int e = ids[sorted_pairs[r]];
    const __half* b = bias + (long long)e * 2 * I;
    const __half* x = raw + r * 2 * I;
    float g, u;
    if (act == NF4MOE_ACT_SWIGLU_OAI) {
      g = __fadd_rn(__half2float(x[2 * c]), __half2float(b[2 * c]));
      u = __fadd_rn(__half2float(x[2 * c + 1]), __half2float(b[2 * c + 1]));
      h[i] = __float2half_rn(swiglu_oai(g, u));
    } else {
      g = __fadd_rn(__half2float(x[c]), __half2float(b[c]));
      u = __fadd_rn(__half2float(x[I + c]), __half2float(b[I + c]));
      h[i] = __float2half_rn(silu_mul(g, u));
    }
  }
}

}  // namespace nf4moe
