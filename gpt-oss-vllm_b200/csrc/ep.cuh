// Expert-parallel dispatch / combine helpers: experts are sharded contiguously over the ranks
// (owner(e) = e / (E / world)), tokens stay data-parallel, and the only data-path exchange is one fixed-capacity
// all-to-all of activation rows each way.  The kernels here write straight into the send buffer and read straight from
// the receive buffer, so there is no separate pack / unpack pass on either side of the collective.
#pragma once
#include "common.cuh"

namespace nf4moe {

// The bucket kernel sorts the (token, slot) pairs by GLOBAL expert id; owners are contiguous expert ranges, so that order
// is already grouped by owner rank and, inside a rank, by expert.  Slot j of peer p in the send buffer therefore is sorted
// position expert_offsets[p * E_local] + j.  One warp per send slot:
//   send_rows[p * cap + j, :] = x[pair / k, :]      send_ids[p * cap + j] = local expert id (or -1 for an unused slot)
//   pair_slot[pair]           = p * cap + j          (the row the combine reads after the return all-to-all)
// Pairs past the capacity are dropped and flagged in status[0] (cannot happen with cap = T * k, the default).
__global__ void __launch_bounds__(256) ep_pack_kernel(const __half* __restrict__ x, const int32_t* __restrict__ ids,
                                                      const int32_t* __restrict__ expert_offsets,
                                                      const int32_t* __restrict__ sorted_pairs, int k, int H, int world,
                                                      int e_local, int cap, __half* __restrict__ send_rows,
                                                      int32_t* __restrict__ send_ids, int32_t* __restrict__ pair_slot,
                                                      int32_t* __restrict__ status) {
  const int lane = threadIdx.x & 31;
  const int n_slots = world * cap;
  const int vec = H / 8;
  for (int slot = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; slot < n_slots; slot += (gridDim.x * blockDim.x) >> 5) {
    const int p = slot / cap, j = slot - p * cap;
    const int start = expert_offsets[p * e_local];
    const int cnt = expert_offsets[(p + 1) * e_local] - start;
    if (j == 0 && lane == 0 && cnt > cap) atomicOr(status, 1);
    if (j < cnt) {
      const int pair = sorted_pairs[start + j];
      const uint4* src = reinterpret_cast<const uint4*>(x + (long long)(pair / k) * H);
      uint4* dst = reinterpret_cast<uint4*>(send_rows + (long long)slot * H);
      for (int i = lane; i < vec; i += 32) dst[i] = __ldg(src + i);
      if (lane == 0) {
        send_ids[slot] = ids[pair] - p * e_local;
        pair_slot[pair] = slot;
      }
    } else if (lane == 0) {
      send_ids[slot] = -1;
    }
  }
}

// out[t, :] = fp16( sum_slot w[t, slot] * (y[pair_slot[t*k+slot], :] + bias[ids[t*k+slot], :]) ), fp32, slot ascending: the same
// arithmetic and order as combine_kernel, reading the rows where the return all-to-all left them.  pair_slot < 0: pair dropped.
// bias (optional): the FULL down-projection bias table fp16 [E, H] (replicated on every rank: 184 KB for gpt-oss-20b), indexed by
// the home rank's GLOBAL expert ids -- added at home, in fp32, so the expert-parallel sum equals the single-GPU one bit for bit.
__global__ void __launch_bounds__(256) combine_indexed_kernel(const __half* __restrict__ y, const float* __restrict__ w,
                                                              const int32_t* __restrict__ pair_slot, const int32_t* __restrict__ ids,
                                                              const __half* __restrict__ bias, int E, int T, int k, int H,
                                                              __half* __restrict__ out) {
  const int vec = H / 8;
  const long long total = (long long)T * vec;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int t = (int)(i / vec), c = (int)(i - (long long)t * vec);
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int s = 0; s < k; ++s) {
      const int row = pair_slot[(long long)t * k + s];
      if (row < 0) continue;
      const float ws = w[(long long)t * k + s];
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(y + (long long)row * H) + c);
      const __half2* h = reinterpret_cast<const __half2*>(&v);
      uint4 bv = make_uint4(0u, 0u, 0u, 0u);
      if (bias) {
        const int e = ids[(long long)t * k + s];
        if (e >= 0 && e < E) bv = __ldg(reinterpret_cast<const uint4*>(bias + (long long)e * H) + c);
      }
      const __half2* bh = reinterpret_cast<const __half2*>(&bv);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float2 f = __half22float2(h[j]);
        if (bias) { const float2 b = __half22float2(bh[j]); f.x = __fadd_rn(f.x, b.x); f.y = __fadd_rn(f.y, b.y); }
        acc[2 * j] = __fadd_rn(acc[2 * j], __fmul_rn(ws, f.x));
        acc[2 * j + 1] = __fadd_rn(acc[2 * j + 1], __fmul_rn(ws, f.y));
      }
    }
    uint4 o;
    uint32_t* ow = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
    for (int j = 0; j < 4; ++j) ow[j] = pack_half2(acc[2 * j], acc[2 * j + 1]);
    reinterpret_cast<uint4*>(out + (long long)t * H)[c] = o;
  }
}

}  // namespace nf4moe
