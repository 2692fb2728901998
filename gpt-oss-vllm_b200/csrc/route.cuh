// Router top-k, expert bucketing (stable counting sort), row permute and the weighted un-permute epilogue.
#pragma once
#include <cuda_bf16.h>

#include "common.cuh"

namespace nf4moe {

// ---------------------------------------------------------------------------------------------------------------
// (1) router: one warp per token, E/32 logits per lane, k rounds of warp arg-max with shuffles
// (ties -> lowest expert index, like torch.topk on the reference's P/unsloth_to_vllm_converter.py:291), then a
// softmax over the k selected logits in fp32 (:292) == softmax -> top-k -> renormalise.
// ---------------------------------------------------------------------------------------------------------------
// One warp routes token t: lane j < k returns the j-th selected expert and its weight.
template <int kPerLane>
__device__ __forceinline__ void route_one_token(const void* logits, int dtype, int t, int E, int k, int lane, int& sel_i,
                                                float& weight) {
  float v[kPerLane];
#pragma unroll
  for (int i = 0; i < kPerLane; ++i) {
    const int eid = i * 32 + lane;
    const long long idx = (long long)t * E + eid;
    float x;
    if (eid >= E) x = -INFINITY;
    else if (dtype == 0) x = __half2float(reinterpret_cast<const __half*>(logits)[idx]);
    else if (dtype == 1) x = reinterpret_cast<const float*>(logits)[idx];
    else x = __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(logits)[idx]);
    // NaN never wins a comparison below; map it to -inf so the order stays total.
    v[i] = (x == x) ? x : -INFINITY;
  }
  float sel_v = 0.f;  // lane j keeps the j-th selected logit / id
  sel_i = 0;
  float vmax = 0.f;
  for (int j = 0; j < k; ++j) {
    float best = v[0];
    int bi = lane;
#pragma unroll
    for (int i = 1; i < kPerLane; ++i)
      if (v[i] > best) { best = v[i]; bi = i * 32 + lane; }  // strict >: lower index wins inside a lane
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    if (j == 0) vmax = best;
    if (lane == j) { sel_v = best; sel_i = bi; }
    if ((bi & 31) == lane) {
#pragma unroll
      for (int i = 0; i < kPerLane; ++i)
        if (i == (bi >> 5)) v[i] = -INFINITY;  // taken
    }
  }
  const float ex = lane < k ? expf(sel_v - vmax) : 0.f;
  float sum = ex;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  weight = ex / sum;
}

template <int kPerLane>
__global__ void __launch_bounds__(256) route_topk_kernel(const void* logits, int dtype, int T, int E, int k,
                                                         int32_t* __restrict__ ids, float* __restrict__ weights) {
  grid_dep_launch();
  grid_dep_wait();   // launched with launch_pdl(): the logits (and a previous layer's readers of ids/weights) are done past here
  const int lane = threadIdx.x & 31;
  const int t = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (t >= T) return;
  int sel_i;
  float w;
  route_one_token<kPerLane>(logits, dtype, t, E, k, lane, sel_i, w);
  if (lane < k) {
    ids[(long long)t * k + lane] = sel_i;
    weights[(long long)t * k + lane] = w;
  }
}

// Router AND bucketing for the smallest batches (T <= 8 and T * k <= 32: batch-1..8 decode, where a layer is launch latency, not
// work): ONE CTA of 8 warps, warp t routes token t, then warp 0 sorts the <= 32 pairs with a __match_any_sync (a pair's rank
// inside its expert = the matching lanes below it: ascending pair index, as in bucket_kernel) and scans the per-expert counts.  One
// dependent launch instead of two (~3.5 us of a 31 us layer at one token).  No token-tile list: this serves the GEMV regime only.
template <int kPerLane>
__global__ void __launch_bounds__(256) route_bucket_small_kernel(const void* logits, int dtype, int T, int E, int k,
                                                                 int32_t* __restrict__ ids, float* __restrict__ weights,
                                                                 int32_t* __restrict__ expert_offsets, int32_t* __restrict__ sorted_pairs,
                                                                 int32_t* __restrict__ active_experts, int32_t* __restrict__ meta) {
  __shared__ int s_id[32];
  __shared__ int s_cnt[256];
  __shared__ int s_off[256];
  grid_dep_launch();
  grid_dep_wait();
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  s_cnt[tid] = 0;
  if (warp < T) {
    int sel_i;
    float w;
    route_one_token<kPerLane>(logits, dtype, warp, E, k, lane, sel_i, w);
    if (lane < k) {
      ids[warp * k + lane] = sel_i;
      weights[warp * k + lane] = w;
      s_id[warp * k + lane] = sel_i;
    }
  }
  __syncthreads();
  if (warp != 0) return;
  const int n = T * k;
  const int id = lane < n ? s_id[lane] : -1;
  const bool valid = id >= 0 && id < E;
  const unsigned same = __match_any_sync(0xffffffffu, id);
  const int rank = __popc(same & ((1u << lane) - 1u));
  if (valid && rank == 0) s_cnt[id] = __popc(same);
  __syncwarp();
  int run_base = 0, act_base = 0, mx = 0;
  const int n_pass = (E + 31) >> 5;
#pragma unroll 1
  for (int j = 0; j < n_pass; ++j) {
    const int e = j * 32 + lane;
    const int c = e < E ? s_cnt[e] : 0;
    int inc = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int nb = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += nb;
    }
    const int off = run_base + inc - c;
    if (e < E) { s_off[e] = off; expert_offsets[e] = off; }
    const unsigned act = __ballot_sync(0xffffffffu, c > 0);
    if (c > 0) active_experts[act_base + __popc(act & ((1u << lane) - 1u))] = e;
    act_base += __popc(act);
    mx = max(mx, c);
    run_base += __shfl_sync(0xffffffffu, inc, 31);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  __syncwarp();
  if (valid) sorted_pairs[s_off[id] + rank] = lane;
#pragma unroll 1
  for (int i = act_base + lane; i < E; i += 32) active_experts[i] = -1;
  if (lane == 0) { expert_offsets[E] = run_base; meta[0] = act_base; meta[1] = run_base; meta[2] = 0; meta[3] = mx; meta[5] = 0; }
}

// Warp 0 of the bucket kernels: exclusive scan of the per-expert counts (E <= 256) into s_off, and -- in the same pass -- the
// compacted list of active experts, the token-tile list of the grouped GEMM and meta[0..3].  One expert per LANE per pass of 32
// experts (E = 32: a single pass).  Round 2 measured the earlier form -- 8 consecutive experts per lane, fully unrolled, three
// integer divisions each -- at 5.4 us (T = 1) to 8.9 us (T = 256) of a lone warp's dependent instructions, most of the bucket
// kernel and ~7 % of a decode layer (scripts/bucket_stamps.py); round 1's serial compaction on one thread was slower still.
// Token tiles: an expert's c rows split into ceil(c / tile_rows) near-equal tiles (multiples of 16 rows).
__device__ __forceinline__ void bucket_scan_and_lists(const int* s_count, int* s_off, int E, int lane, int32_t* expert_offsets,
                                                      int32_t* active_experts, int32_t* meta, int32_t* tile_list, int tile_rows,
                                                      bool write_global) {
  int run_base = 0, act_base = 0, tile_base = 0, mx = 0;
  const int n_pass = (E + 31) >> 5;
#pragma unroll 1
  for (int j = 0; j < n_pass; ++j) {
    const int e = j * 32 + lane;
    const int c = e < E ? s_count[e] : 0;
    int tsz = 0, nt = 0;
    if (c > 0) {
      // ceil(a / b) for 0 < a, b < 2^20 through one reciprocal and a +-1 correction: the integer-division sequences were most of
      // this function's code, and a lone warp running cold straight-line code pays an instruction-cache miss per 8 instructions
      auto cdiv = [](int a, int b) {
        int q = __float2int_rz(__fdividef((float)a, (float)b));
        q += (q + 1) * b <= a ? 1 : 0;
        q -= q * b > a ? 1 : 0;
        return q + (q * b < a ? 1 : 0);
      };
      const int ntl = cdiv(c, tile_rows);
      tsz = (cdiv(c, ntl) + 15) & ~15;
      nt = cdiv(c, tsz);
    }
    mx = max(mx, c);
    int inc = c, inc_a = c > 0 ? 1 : 0, inc_t = nt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int n = __shfl_up_sync(0xffffffffu, inc, o), na = __shfl_up_sync(0xffffffffu, inc_a, o), nn = __shfl_up_sync(0xffffffffu, inc_t, o);
      if (lane >= o) { inc += n; inc_a += na; inc_t += nn; }
    }
    const int off = run_base + inc - c;
    s_off[e] = off;                      // e <= 255
    if (write_global) {
      if (e < E) expert_offsets[e] = off;
      if (c > 0) {
        active_experts[act_base + inc_a - 1] = e;
        if (tile_list) {
          int t = tile_base + inc_t - nt;
#pragma unroll 1
          for (int r = 0; r < c; r += tsz, ++t) *reinterpret_cast<int4*>(tile_list + 4 * t) = make_int4(e, off + r, min(tsz, c - r), 0);
        }
      }
    }
    run_base += __shfl_sync(0xffffffffu, inc, 31);
    act_base += __shfl_sync(0xffffffffu, inc_a, 31);
    tile_base += __shfl_sync(0xffffffffu, inc_t, 31);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if (lane == 0) s_off[256] = run_base;
  if (!write_global) return;
  if (lane == 0) expert_offsets[E] = run_base;
#pragma unroll 1
  for (int i = act_base + lane; i < E; i += 32) active_experts[i] = -1;   // (unrolled 16x this was a fifth of the function's code)
  if (lane == 0) { meta[0] = act_base; meta[1] = run_base; meta[2] = tile_base; meta[3] = mx; meta[5] = tile_list ? tile_rows : 0; }
}

// ---------------------------------------------------------------------------------------------------------------
// (2) bucketing: ONE CTA of 1024 threads; ids staged in shared memory as bytes (E <= 256; 0xFF = dropped).
// Warp w owns experts w, w+32, ...: a counting pass and a placement pass over all pairs with ballots, so the
// order inside an expert is ascending pair index (deterministic).  Replaces the per-(token,k) `.item()` dispatch of
// P/unsloth_to_vllm_converter.py:298-305.
// ---------------------------------------------------------------------------------------------------------------
constexpr int kBucketThreads = 1024;
#ifdef NF4MOE_EXPERIMENTS
__device__ long long g_bucket_stamps[8];
#define BUCKET_STAMP(i) do { if (threadIdx.x == 0) { unsigned long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); g_bucket_stamps[i] = (long long)t_; } } while (0)
#else
#define BUCKET_STAMP(i) do { } while (0)
#endif

__global__ void __launch_bounds__(kBucketThreads) bucket_kernel(const int32_t* ids, int n_pairs, int E,
                                                                int32_t* __restrict__ expert_offsets,
                                                                int32_t* __restrict__ sorted_pairs,
                                                                int32_t* __restrict__ active_experts,
                                                                int32_t* __restrict__ meta,
                                                                int32_t* __restrict__ tile_list, int tile_rows) {
  extern __shared__ uint8_t s_ids[];  // n_pairs rounded up to 32
  __shared__ int s_count[256];
  __shared__ int s_off[257];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  BUCKET_STAMP(0);
  grid_dep_launch();
  grid_dep_wait();
  BUCKET_STAMP(1);
  const int n_pad = (n_pairs + 127) & ~127;
  for (int i = tid; i < n_pad; i += kBucketThreads) {
    int e = i < n_pairs ? ids[i] : -1;
    s_ids[i] = (e >= 0 && e < E) ? (uint8_t)e : (uint8_t)0xFF;
  }
  if (tid < 256) s_count[tid] = 0;
  __syncthreads();
  BUCKET_STAMP(2);
  // ids are read four at a time (one 32-bit shared-memory load per lane = 128 pairs per warp iteration); n_pad128 rounds
  // the scan up to whole iterations, the tail is 0xFF (never matches)
  const int n_pad128 = (n_pairs + 127) & ~127;
  for (int e = warp; e < E; e += 32) {
    const uint32_t pat = 0x01010101u * (uint32_t)e;
    int c = 0;
    for (int i = lane * 4; i < n_pad128; i += 128) {
      const uint32_t x = *reinterpret_cast<const uint32_t*>(s_ids + i) ^ pat;   // a zero byte = a match
      const uint32_t z = __vcmpeq4(x, 0u);                                        // 0xFF in matching bytes
      c += __popc(z) >> 3;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if (lane == 0) s_count[e] = c;
  }
  __syncthreads();
  BUCKET_STAMP(3);
  // Every warp scans the counts itself (one expert per lane and pass: 5 shuffles) and goes straight on to place its own experts'
  // pairs -- no second block barrier, no shared offsets.  The lists the later kernels read are written on the side by three
  // different warps (tile list + meta[2] by warp 0, active experts + meta[0] by warp 1, offsets + meta[1], meta[3] by warp 2), each
  // running only its own short piece of code: one warp doing all of it was 3.8 us of cold, dependent instructions on the critical
  // path of every layer (scripts/bucket_stamps.py).
  (void)s_off;
  int my_off[8];                       // offsets of this warp's experts e = warp + 32 j
  {
    int run_base = 0, act_base = 0, tile_base = 0, mx = 0;
    const int n_pass = (E + 31) >> 5;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      my_off[j] = 0;
      if (j < n_pass) {
        const int e = j * 32 + lane;
        const int c = e < E ? s_count[e] : 0;
        int inc = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const int n = __shfl_up_sync(0xffffffffu, inc, o);
          if (lane >= o) inc += n;
        }
        const int off = run_base + inc - c;
        my_off[j] = __shfl_sync(0xffffffffu, off, warp);
        if (warp == 0) {                 // token tiles: an expert's c rows split into near-equal tiles, multiples of 16 rows
          int tsz = 0, nt = 0;
          if (c > 0) {
            const int ntl = (c + tile_rows - 1) / tile_rows;
            tsz = (((c + ntl - 1) / ntl) + 15) & ~15;
            nt = (c + tsz - 1) / tsz;
          }
          int inc_t = nt;
#pragma unroll
          for (int o = 1; o < 32; o <<= 1) {
            const int n = __shfl_up_sync(0xffffffffu, inc_t, o);
            if (lane >= o) inc_t += n;
          }
          if (tile_list != nullptr && c > 0) {
            int t = tile_base + inc_t - nt;
#pragma unroll 1
            for (int r = 0; r < c; r += tsz, ++t) *reinterpret_cast<int4*>(tile_list + 4 * t) = make_int4(e, off + r, min(tsz, c - r), 0);
          }
          tile_base += __shfl_sync(0xffffffffu, inc_t, 31);
        } else if (warp == 1) {          // active experts, ascending
          int inc_a = c > 0 ? 1 : 0;
#pragma unroll
          for (int o = 1; o < 32; o <<= 1) {
            const int n = __shfl_up_sync(0xffffffffu, inc_a, o);
            if (lane >= o) inc_a += n;
          }
          if (c > 0) active_experts[act_base + inc_a - 1] = e;
          act_base += __shfl_sync(0xffffffffu, inc_a, 31);
        } else if (warp == 2) {          // offsets and the largest count
          if (e < E) expert_offsets[e] = off;
          mx = max(mx, c);
        }
        run_base += __shfl_sync(0xffffffffu, inc, 31);
      }
    }
    if (warp == 0) {
      if (lane == 0) { meta[2] = tile_base; meta[5] = tile_list ? tile_rows : 0; }   // [5]: no tile has more rows than this
    } else if (warp == 1) {
#pragma unroll 1
      for (int i = act_base + lane; i < E; i += 32) active_experts[i] = -1;
      if (lane == 0) meta[0] = act_base;
    } else if (warp == 2) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      if (lane == 0) { expert_offsets[E] = run_base; meta[1] = run_base; meta[3] = mx; }
    }
  }
  BUCKET_STAMP(4);
  for (int e = warp, j = 0; e < E; e += 32, ++j) {
    const uint32_t pat = 0x01010101u * (uint32_t)e;
    int run = my_off[j];
    for (int i = lane * 4; i < n_pad128; i += 128) {
      const uint32_t z = __vcmpeq4(*reinterpret_cast<const uint32_t*>(s_ids + i) ^ pat, 0u);
      const int mine = __popc(z) >> 3;                      // matches among this lane's four pairs
      int inc = mine;                                        // inclusive prefix over the lanes (pair order = lane-major)
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int nb = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += nb;
      }
      int pos = run + inc - mine;
#pragma unroll
      for (int jj = 0; jj < 4; ++jj)
        if ((z >> (8 * jj)) & 1u) sorted_pairs[pos++] = i + jj;
      run += __shfl_sync(0xffffffffu, inc, 31);
    }
  }
  BUCKET_STAMP(5);
}

// Prefill-sized batches: the same stable counting sort by ONE thread-block cluster of kBucketCluster CTAs (the single CTA takes
// 155 us for 8192 x 4 pairs, 8 % of the prefill layer).  CTA r sorts the contiguous chunk r of the pair list; the per-expert
// counts of all CTAs are exchanged through distributed shared memory, so a pair's position is
// offset[e] + (matches in lower chunks) + (rank inside its chunk): the order inside an expert stays ascending pair index.
constexpr int kBucketCluster = 8;
__global__ void __launch_bounds__(kBucketThreads) bucket_cluster_kernel(const int32_t* ids, int n_pairs, int E,
                                                                        int32_t* __restrict__ expert_offsets,
                                                                        int32_t* __restrict__ sorted_pairs,
                                                                        int32_t* __restrict__ active_experts,
                                                                        int32_t* __restrict__ meta,
                                                                        int32_t* __restrict__ tile_list, int tile_rows) {
  extern __shared__ uint8_t s_ids[];  // this CTA's chunk, padded to a multiple of 128 with 0xFF
  __shared__ int s_count[256];        // this chunk's matches per expert (read by the other CTAs of the cluster)
  __shared__ int s_tot[256];
  __shared__ int s_before[256];
  __shared__ int s_off[257];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  grid_dep_launch();
  grid_dep_wait();
  uint32_t rank, nc;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(nc));
  const int chunk = (((n_pairs + (int)nc - 1) / (int)nc) + 127) & ~127;
  const int c0 = min(n_pairs, (int)rank * chunk), c1 = min(n_pairs, c0 + chunk);
  const int n_loc = c1 - c0, n_pad128 = (n_loc + 127) & ~127;
  for (int i = tid; i < n_pad128; i += kBucketThreads) {
    const int e = i < n_loc ? ids[c0 + i] : -1;
    s_ids[i] = (e >= 0 && e < E) ? (uint8_t)e : (uint8_t)0xFF;
  }
  if (tid < 256) s_count[tid] = 0;
  __syncthreads();
  for (int e = warp; e < E; e += 32) {
    const uint32_t pat = 0x01010101u * (uint32_t)e;
    int c = 0;
    for (int i = lane * 4; i < n_pad128; i += 128) c += __popc(__vcmpeq4(*reinterpret_cast<const uint32_t*>(s_ids + i) ^ pat, 0u)) >> 3;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if (lane == 0) s_count[e] = c;
  }
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  if (tid < 256) {   // totals over the cluster and the matches in lower chunks, through distributed shared memory
    int tot = 0, before = 0;
    const uint32_t mine = smem_u32(&s_count[tid]);
    for (uint32_t r = 0; r < nc; ++r) {
      uint32_t remote, v;
      asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(mine), "r"(r));
      asm volatile("ld.shared::cluster.u32 %0, [%1];" : "=r"(v) : "r"(remote) : "memory");
      tot += (int)v;
      if (r < rank) before += (int)v;
    }
    s_tot[tid] = tot;
    s_before[tid] = before;
  }
  __syncthreads();
  if (warp == 0) bucket_scan_and_lists(s_tot, s_off, E, lane, expert_offsets, active_experts, meta, tile_list, tile_rows, rank == 0);
  __syncthreads();
  for (int e = warp; e < E; e += 32) {
    const uint32_t pat = 0x01010101u * (uint32_t)e;
    int run = s_off[e] + s_before[e];
    for (int i = lane * 4; i < n_pad128; i += 128) {
      const uint32_t z = __vcmpeq4(*reinterpret_cast<const uint32_t*>(s_ids + i) ^ pat, 0u);
      const int mine = __popc(z) >> 3;
      int inc = mine;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int nb = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += nb;
      }
      int pos = run + inc - mine;
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if ((z >> (8 * j)) & 1u) sorted_pairs[pos++] = c0 + i + j;
      run += __shfl_sync(0xffffffffu, inc, 31);
    }
  }
  // no CTA may exit while another can still read its s_count
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// x_sorted[s,:] = x[sorted_pairs[s]/k,:]; one CTA of 128 threads per row pass (a 2880-wide row = 360 16-byte chunks = three per
// thread, all in flight at once: round 1 walked a row with one warp, twelve dependent round trips, 5.8 us for 256 rows).
// Rows past meta[1] are zeroed so the grouped GEMM's last tile reads defined data.
__global__ void __launch_bounds__(128) permute_rows_kernel(const __half* x,
                                                           const int32_t* sorted_pairs,
                                                           const int32_t* meta, int n_rows, int k, int H,
                                                           __half* __restrict__ xs) {
  grid_dep_launch();
  grid_dep_wait();
  const int n_kept = ld_dep_s32(meta + 1);
  const int vec = H / 8;
  // rows past the kept pairs are only ever read as padding of the LAST token tile (<= 256 rows): zero that much, not a whole
  // mostly empty fixed-capacity buffer
  n_rows = min(n_rows, n_kept + 256);
  for (int s = blockIdx.x; s < n_rows; s += gridDim.x) {
    uint4* dst = reinterpret_cast<uint4*>(xs + (long long)s * H);
    if (s < n_kept) {
      const uint4* src = reinterpret_cast<const uint4*>(x + (long long)(sorted_pairs[s] / k) * H);
      for (int i0 = 0; i0 < vec; i0 += 4 * 128) {
        uint4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int i = i0 + u * 128 + (int)threadIdx.x;
          if (i < vec) v[u] = src[i];
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int i = i0 + u * 128 + (int)threadIdx.x;
          if (i < vec) dst[i] = v[u];
        }
      }
    } else {
      for (int i = threadIdx.x; i < vec; i += 128) dst[i] = make_uint4(0, 0, 0, 0);
    }
  }
}

// (5) out[t,:] = fp16( sum_slot w[t,slot] * y[t*k+slot,:] ), fp32, slot ascending
// (P/unsloth_to_vllm_converter.py:321-323).  One thread = 8 columns.
__global__ void __launch_bounds__(256) combine_kernel(const __half* y, const float* w,
                                                      const int32_t* ids, int E, int T, int k, int H,
                                                      __half* __restrict__ out, const __half* __restrict__ bias) {
  // bias (optional, needs ids): fp16 [E, H], the down-projection bias of the slot's expert, added to y in fp32 before weighting
  // (gpt-oss `w2_bias`, V/model_executor/models/gpt_oss.py:199)
  grid_dep_launch();
  grid_dep_wait();
  const int vec = H / 8;
  const long long total = (long long)T * vec;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int t = (int)(i / vec), c = (int)(i - (long long)t * vec);
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    // slots in chunks of 4 with every load of a chunk issued before its first use (round 1 walked the slots one dependent
    // round trip at a time); the sum itself stays strictly slot-ascending in fp32
    for (int s0 = 0; s0 < k; s0 += 4) {
      int e[4];
      float ws[4];
      uint4 v[4], bv[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int s = s0 + u;
        e[u] = -1;
        if (s < k) {
          e[u] = ids ? ids[(long long)t * k + s] : 0;
          ws[u] = w[(long long)t * k + s];
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const bool ok = e[u] >= 0 && e[u] < (ids ? E : 1);
        if (!ok) e[u] = -1;
        v[u] = bv[u] = make_uint4(0u, 0u, 0u, 0u);
        if (ok) {
          v[u] = reinterpret_cast<const uint4*>(y + ((long long)t * k + s0 + u) * H)[c];
          if (bias) bv[u] = __ldg(reinterpret_cast<const uint4*>(bias + (long long)e[u] * H) + c);
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (e[u] < 0) continue;
        const __half2* h = reinterpret_cast<const __half2*>(&v[u]);
        const __half2* bh = reinterpret_cast<const __half2*>(&bv[u]);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float2 f = __half22float2(h[j]);
          if (bias) { const float2 b = __half22float2(bh[j]); f.x = __fadd_rn(f.x, b.x); f.y = __fadd_rn(f.y, b.y); }
          acc[2 * j] = __fadd_rn(acc[2 * j], __fmul_rn(ws[u], f.x));
          acc[2 * j + 1] = __fadd_rn(acc[2 * j + 1], __fmul_rn(ws[u], f.y));
        }
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
