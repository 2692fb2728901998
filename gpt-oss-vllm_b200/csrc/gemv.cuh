// Decode-regime grouped GEMV: the HBM-bound kernel of the path.
//
// One persistent CTA per SM walks a contiguous range of work items; an item is 16 consecutive weight rows of one
// active expert = one contiguous span of packed NF4 bytes (16*K/2) plus its absmax (16*K/64 values).
//   * warp 15 (one lane) is the producer: TMA 1-D bulk copies (cp.async.bulk, SASS UBLKCP) of the raw packed bytes
//     and absmax into a shared-memory ring, completion on mbarriers, L2 evict-first (weights are read exactly once);
//   * warps 0..14 are consumers: warp w owns k-blocks [3w, 3w+3) (15*3*64 = 2880 = H = I of gpt-oss), keeps the
//     activations of those k for up to 8*NT routed tokens in REGISTERS as mma B fragments for the whole expert, reads
//     its 8 packed bytes per row per block with one LDS.64, decodes each byte through a 64 KB bank-conflict-free
//     shared-memory LUT (byte -> half2 of the two NF4 codes; one PRMT forms the address, one LDS fetches), feeds
//     mma.sync.m16n8k16 (weights = A, tokens = B) and applies the fp32 absmax to the per-block fp32 partial sums;
//   * the 15 k-slices are reduced through shared memory; the reducer applies the fused epilogue: gpt-oss interleaved
//     clamped SwiGLU for w13 (rows 2j/2j+1 = gate/up, V/model_executor/layers/activation.py:389-397) or the scatter
//     back to (token,slot) order for w2.
// Replaces bnb Linear4bit on single rows, P/unsloth_to_vllm_converter.py:309-318, for all routed pairs at once.
#pragma once
#include "common.cuh"

namespace nf4moe {

constexpr int kGemvConsumers = 15;
constexpr int kGemvThreads = (kGemvConsumers + 1) * 32;
constexpr int kGemvRows = 16;
constexpr int kKBW = 3;  // k-blocks per consumer warp
constexpr int kGemvMaxK = kGemvConsumers * kKBW * kBlock;  // 2880
constexpr uint32_t kLutBytes = 65536;
constexpr int kMaxStages = 8;
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

template <int NT, bool W13, bool NESTED>
__global__ void __launch_bounds__(kGemvThreads, 1) gemv_kernel(const GemvArgs a) {
  extern __shared__ __align__(128) uint8_t dyn_smem[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int K = a.w.cols, nkb = a.w.nkb;
  const uint32_t row_bytes = (uint32_t)K / 2;
  const uint32_t w_bytes = kGemvRows * row_bytes;
  const uint32_t a_bytes = kGemvRows * (uint32_t)nkb * (NESTED ? 1u : 4u);
  const uint32_t stage_bytes = (w_bytes + a_bytes + 127u) & ~127u;

  // ---- carve shared memory around a 64 KB-ALIGNED lookup table (the PRMT address trick needs bits 0..15 free) ----
  const uint32_t base = smem_u32(dyn_smem);
  const uint32_t lut = (base + kLutBytes - 1) & ~(kLutBytes - 1);
  uint32_t a_lo = base, a_hi = lut;                        // region A: below the table
  uint32_t b_lo = lut + kLutBytes, b_hi = base + kGemvSmemBytes;  // region B: above it
  constexpr uint32_t kPartialBytes = 2u * kGemvConsumers * 128u * NT * 4u;
  constexpr uint32_t kSmallBytes = 2u * kMaxStages * 8u + kPartialBytes;
  uint32_t small;
  if (a_hi - a_lo >= kSmallBytes) { small = a_lo; a_lo += (kSmallBytes + 127u) & ~127u; }
  else { small = b_lo; b_lo += (kSmallBytes + 127u) & ~127u; }
  const uint32_t bar_full = small, bar_empty = small + kMaxStages * 8u;
  float* partial = reinterpret_cast<float*>(dyn_smem + (small + 2u * kMaxStages * 8u - base));
  const int n_a = (int)((a_hi - a_lo) / stage_bytes);
  const int n_b = (int)((b_hi - b_lo) / stage_bytes);
  const int n_stages = min(kMaxStages, n_a + n_b);
  auto stage_addr = [&](int s) -> uint32_t {
    return s < n_a ? a_lo + (uint32_t)s * stage_bytes : b_lo + (uint32_t)(s - n_a) * stage_bytes;
  };

  // table: entry (byte b, lane l) at lut + b*256 + l*4 = half2{ NF4[b>>4], NF4[b&15] } -> bank l for every b
  for (int i = tid; i < 256 * 32; i += kGemvThreads) {
    const int b = i >> 5, l = i & 31;
    sts32(lut + (uint32_t)b * 256u + (uint32_t)l * 4u, pack_half2(kNF4[b >> 4], kNF4[b & 15]));
  }
  if (tid == 0) {
    for (int s = 0; s < kMaxStages; ++s) {
      mbar_init(bar_full + 8u * s, 1);
      mbar_init(bar_empty + 8u * s, kGemvConsumers);
    }
    fence_barrier_init();
  }
  __syncthreads();

  // ---- work range: items = (active expert, 16-row group), contiguous per CTA ----
  const int n_groups = a.w.rows / kGemvRows;
  const long long n_items = (long long)a.meta[0] * n_groups;
  const long long it_lo = n_items * blockIdx.x / gridDim.x;
  const long long it_hi = n_items * (blockIdx.x + 1) / gridDim.x;

  if (warp == kGemvConsumers) {
    // =============================== producer ===============================
    if (lane == 0) {
      const uint64_t pol = l2_policy_evict_first();
      int s = 0;
      uint32_t ph = 0;
      for (long long it = it_lo; it < it_hi; ++it) {
        const int ae = (int)(it / n_groups), grp = (int)(it - (long long)ae * n_groups);
        const int e = a.active_experts[ae];
        mbar_wait(bar_empty + 8u * s, ph ^ 1u);
        mbar_expect_tx(bar_full + 8u * s, w_bytes + a_bytes);
        const uint32_t dst = stage_addr(s);
        bulk_g2s(dst, a.w.packed + (long long)e * a.w.packed_stride + (long long)grp * w_bytes, w_bytes,
                 bar_full + 8u * s, pol);
        const long long blk0 = (long long)e * a.w.nb + (long long)grp * kGemvRows * nkb;
        const void* am_src = NESTED ? (const void*)(a.w.absmax_u8 + blk0) : (const void*)(a.w.absmax_f32 + blk0);
        bulk_g2s(dst + w_bytes, am_src, a_bytes, bar_full + 8u * s, pol);
        if (++s == n_stages) { s = 0; ph ^= 1u; }
      }
    }
    return;
  }

  // =============================== consumers ===============================
  const int g = lane >> 2, t = lane & 3;
  const uint32_t lut_lane = lut | ((uint32_t)lane << 2);
  uint32_t breg[NT][kKBW][8];
  int cur_e = -1, s0 = 0, m_e = 0, n_chunks = 0;
  int s = 0;
  uint32_t ph = 0, flip = 0;

  auto load_b = [&](int chunk) {
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const int col = chunk * 8 * NT + nt * 8 + g;
      const bool valid = col < m_e;
      long long src_row = 0;
      if (valid) src_row = W13 ? (long long)(a.sorted_pairs[s0 + col] / a.k) : (long long)(s0 + col);
#pragma unroll
      for (int j = 0; j < kKBW; ++j) {
        const int kb = warp * kKBW + j;
        uint4 v0 = make_uint4(0, 0, 0, 0), v1 = v0;
        if (valid && kb < nkb) {
          const uint4* p = reinterpret_cast<const uint4*>(a.x + src_row * K + kb * kBlock + 16 * t);
          v0 = __ldg(p);
          v1 = __ldg(p + 1);
        }
        breg[nt][j][0] = v0.x; breg[nt][j][1] = v0.y; breg[nt][j][2] = v0.z; breg[nt][j][3] = v0.w;
        breg[nt][j][4] = v1.x; breg[nt][j][5] = v1.y; breg[nt][j][6] = v1.z; breg[nt][j][7] = v1.w;
      }
    }
  };

  for (long long it = it_lo; it < it_hi; ++it) {
    const int ae = (int)(it / n_groups), grp = (int)(it - (long long)ae * n_groups);
    const int e = a.active_experts[ae];
    if (e != cur_e) {
      cur_e = e;
      s0 = a.expert_offsets[e];
      m_e = a.expert_offsets[e + 1] - s0;
      n_chunks = (m_e + 8 * NT - 1) / (8 * NT);
      if (n_chunks == 1) load_b(0);
    }
    mbar_wait(bar_full + 8u * s, ph);
    const uint32_t st = stage_addr(s);
    const uint32_t am_base = st + w_bytes;

    for (int chunk = 0; chunk < n_chunks; ++chunk) {
      if (n_chunks > 1) load_b(chunk);
      const int m_chunk = min(8 * NT, m_e - chunk * 8 * NT);
      float acc[NT][4];
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f;

#pragma unroll
      for (int j = 0; j < kKBW; ++j) {
        const int kb = warp * kKBW + j;
        if (kb < nkb) {
          const uint32_t wa = st + (uint32_t)g * row_bytes + (uint32_t)kb * 32u + (uint32_t)t * 8u;
          const uint2 w0 = lds64(wa);                   // row g   : k = kb*64 + 16t .. +15
          const uint2 w1 = lds64(wa + 8u * row_bytes);  // row g+8
          float am0, am1;
          if (NESTED) {
            uint32_t c0, c1;
            asm volatile("ld.shared.u8 %0, [%1];" : "=r"(c0) : "r"(am_base + (uint32_t)(g * nkb + kb)));
            asm volatile("ld.shared.u8 %0, [%1];" : "=r"(c1) : "r"(am_base + (uint32_t)((g + 8) * nkb + kb)));
            const long long b0 = (long long)(grp * kGemvRows + g) * nkb + kb;
            const long long b1 = b0 + 8ll * nkb;
            const float* code2 = a.w.code2 + (long long)e * a.w.code2_stride;
            const float off = __ldg(a.w.offset + e);
            am0 = __fadd_rn(__fmul_rn(__ldg(code2 + c0), __ldg(a.w.absmax2 + (long long)e * a.w.n2 + (b0 >> 8))), off);
            am1 = __fadd_rn(__fmul_rn(__ldg(code2 + c1), __ldg(a.w.absmax2 + (long long)e * a.w.n2 + (b1 >> 8))), off);
          } else {
            am0 = lds_f32(am_base + 4u * (uint32_t)(g * nkb + kb));
            am1 = lds_f32(am_base + 4u * (uint32_t)((g + 8) * nkb + kb));
          }
          float d[NT][4];
#pragma unroll
          for (int nt = 0; nt < NT; ++nt) d[nt][0] = d[nt][1] = d[nt][2] = d[nt][3] = 0.f;
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
            for (int nt = 0; nt < NT; ++nt)
              if (nt == 0 || nt * 8 < m_chunk) mma_16816(d[nt], a0, a1, a2, a3, breg[nt][j][2 * m], breg[nt][j][2 * m + 1]);
          }
#pragma unroll
          for (int nt = 0; nt < NT; ++nt) {
            acc[nt][0] = fmaf(am0, d[nt][0], acc[nt][0]);
            acc[nt][1] = fmaf(am0, d[nt][1], acc[nt][1]);
            acc[nt][2] = fmaf(am1, d[nt][2], acc[nt][2]);
            acc[nt][3] = fmaf(am1, d[nt][3], acc[nt][3]);
          }
        }
      }

      // ---- cross-warp reduction of the 15 k-slices; layout [flip][warp][col 8*NT][row 16] ----
      float* pw = partial + ((size_t)flip * kGemvConsumers + warp) * (128 * NT);
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        const int c = nt * 8 + 2 * t;
        pw[c * 16 + g] = acc[nt][0];
        pw[(c + 1) * 16 + g] = acc[nt][1];
        pw[c * 16 + g + 8] = acc[nt][2];
        pw[(c + 1) * 16 + g + 8] = acc[nt][3];
      }
      named_bar_sync(1, kGemvConsumers * 32);
      if (tid < 128 * NT) {
        const float* pr = partial + (size_t)flip * kGemvConsumers * (128 * NT) + tid;
        float v = 0.f;
#pragma unroll
        for (int w = 0; w < kGemvConsumers; ++w) v += pr[w * 128 * NT];
        const int col = tid >> 4, row = tid & 15;
        const bool valid = col < m_chunk;
        const int srow = s0 + chunk * 8 * NT + col;
        v = __half2float(__float2half_rn(v));  // fp16 storage of the product, as the oracle / reference does
        if (W13) {
          if (a.act == NF4MOE_ACT_SWIGLU_OAI) {
            const float up = __shfl_down_sync(0xffffffffu, v, 1);
            if (valid && !(row & 1)) {
              const int I = a.w.rows / 2;
              reinterpret_cast<__half*>(a.out)[(long long)srow * I + grp * 8 + (row >> 1)] =
                  __float2half_rn(swiglu_oai(v, up));
            }
          } else if (valid) {
            reinterpret_cast<__half*>(a.out)[(long long)srow * a.w.rows + grp * kGemvRows + row] = __float2half_rn(v);
          }
        } else if (valid) {
          reinterpret_cast<__half*>(a.out)[(long long)a.sorted_pairs[srow] * a.w.rows + grp * kGemvRows + row] =
              __float2half_rn(v);
        }
      }
      flip ^= 1u;
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(bar_empty + 8u * s);
    if (++s == n_stages) { s = 0; ph ^= 1u; }
  }
}

// reference-variant activation on a raw [n, 2I] buffer: silu(h[:, :I]) * h[:, I:]  (P/unsloth_to_vllm_converter.py:312-315)
__global__ void __launch_bounds__(256) chunk_silu_kernel(const __half* __restrict__ raw, const int32_t* __restrict__ meta,
                                                         int I, __half* __restrict__ h) {
  const long long total = (long long)meta[1] * I;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / I, c = i - r * I;
    h[i] = __float2half_rn(silu_mul(__half2float(raw[r * 2 * I + c]), __half2float(raw[r * 2 * I + I + c])));
  }
}

}  // namespace nf4moe
