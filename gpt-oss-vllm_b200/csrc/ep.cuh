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
__global__ void __launch_bounds__(256) ep_pack_kernel(const __half* x, const int32_t* ids,
                                                      const int32_t* expert_offsets,
                                                      const int32_t* sorted_pairs, int k, int H, int world,
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
      for (int i = lane; i < vec; i += 32) dst[i] = src[i];
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
__global__ void __launch_bounds__(256) combine_indexed_kernel(const __half* y, const float* w,
                                                              const int32_t* pair_slot, const int32_t* ids,
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
      const uint4 v = reinterpret_cast<const uint4*>(y + (long long)row * H)[c];
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


// =====================================================================================================================
// Peer-memory exchange (NVLink / NVSwitch, no NCCL on the data path).
//
// Every rank owns, in memory its peers can address (torch symmetric memory; any peer-mapped allocation works):
//   recv_rows fp16 [W, cap, H]   chunk q = the rows rank q routed to this rank's experts      (written by q's dispatch kernel)
//   recv_ids  int32 [W, cap]     local expert id per row, -1 = unused slot                     (written by q's dispatch kernel)
//   back_rows fp16 [W, cap, H]   chunk p = expert outputs of the rows this rank sent to rank p (written by p's return kernel)
//   flags     u64  [2, W]        [0][q] = epoch of q's last completed dispatch to this rank, [1][p] = ... of p's last return
// and a private counter block ctr u32 [8] (tickets and epochs).  The dispatch kernel stores rows straight from x into the
// OWNER's recv_rows over NVLink (only rows that exist travel, so the worst-case capacity cap = T*k costs nothing on the wire
// and no pair is ever dropped -- round 1's fixed-capacity NCCL all-to-all moved W*cap rows each way and had to trade drops
// against bytes), then the LAST CTA to finish publishes the epoch to every peer (release at system scope).  The consumer
// side is a one-warp kernel that spins (acquire at system scope) until all W flags carry its own epoch; epochs live in
// device memory and advance once per launch, so the whole layer replays from a CUDA graph, in lock step on every rank.
// Buffer reuse needs no extra handshake: a rank starts the next layer's dispatch only after its combine, which waited for
// every owner's return, which every owner issued after its expert kernels had consumed the rows of this layer.
// =====================================================================================================================
constexpr int kEpMaxWorld = 16;
struct EpPeers {
  __half* rows[kEpMaxWorld];              // peer p's recv_rows (dispatch) or back_rows (return)
  int32_t* ids[kEpMaxWorld];              // peer p's recv_ids (dispatch only)
  unsigned long long* flags[kEpMaxWorld]; // peer p's flag row to write into (this rank's entry is [rank])
};

__device__ __forceinline__ void st_release_sys_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

// Last-CTA-done: every CTA fences its (peer) stores at system scope and takes a ticket; the CTA that draws the last one
// bumps the epoch and publishes it to all peers.  ctr[0] = ticket, ctr[1] = epoch (reset / advanced here).
// wait_flags / wait_ctr (optional): after publishing, the same last CTA also WAITS until every peer's flag for this rank has reached
// this rank's next wait epoch (wait_ctr[0], advanced here), so the kernel completes only when all peers' rows have landed and no
// separate one-warp wait launch sits between it and its consumer.  Only for one process per GPU: with several virtual ranks on one
// stream the peers' kernels are queued BEHIND this one and the wait would never end (the tests use the separate wait kernel).
__device__ __forceinline__ void ep_publish_epoch(unsigned int* ctr, const EpPeers& peers, int world, int rank,
                                                 const unsigned long long* wait_flags = nullptr, unsigned int* wait_ctr = nullptr,
                                                 int32_t* status = nullptr) {
  __shared__ unsigned int s_last;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    s_last = atomicAdd(ctr, 1u) == gridDim.x - 1 ? 1u : 0u;
  }
  __syncthreads();
  if (s_last) {
    __shared__ unsigned int s_wait_epoch;
    if (threadIdx.x == 0) {
      __threadfence_system();
      ctr[0] = 0u;
      ctr[1] += 1u;
      if (wait_flags != nullptr) { s_wait_epoch = wait_ctr[0] + 1u; wait_ctr[0] = s_wait_epoch; }
    }
    __syncthreads();
    const unsigned long long epoch = ctr[1];
    if ((int)threadIdx.x < world) st_release_sys_u64(peers.flags[threadIdx.x] + rank, epoch);
    if (wait_flags != nullptr) {
      if ((int)threadIdx.x < world) {
        const unsigned long long want = s_wait_epoch;
        unsigned long long spins = 0;
        while (ld_acquire_sys_u64(wait_flags + threadIdx.x) < want) {
          // a peer died or never launched: its rows will never land, so nothing downstream may run on them -- leave the bit for
          // post-mortems and kill the launch; the host's next synchronisation fails (as an NCCL abort would)
          if (++spins > (1ull << 27)) { if (status) atomicOr(status, 8); __threadfence_system(); __trap(); }
          __nanosleep(64);
        }
      }
      __threadfence_system();
    }
  }
}

// Dispatch: slot (p, j) of this rank's pair list (sorted by global expert id, so grouped by owner) -> owner p's chunk `rank`.
__global__ void __launch_bounds__(256) ep_dispatch_kernel(const __half* x, const int32_t* ids,
                                                          const int32_t* expert_offsets,
                                                          const int32_t* sorted_pairs, int k, int H, int world, int rank,
                                                          int e_local, int cap, const EpPeers peers, int32_t* __restrict__ pair_slot,
                                                          unsigned int* __restrict__ ctr, int32_t* __restrict__ status, int n_pairs,
                                                          const unsigned long long* wait_flags, unsigned int* wait_ctr) {
  grid_dep_launch();
  grid_dep_wait();
  const int lane = threadIdx.x & 31;
  const int n_slots = world * cap;
  const int vec = H / 8;
  // pairs whose expert id is outside [0, E) are not in the sorted list: mark them "not dispatched" here (every other pair gets its
  // slot below, a different word), instead of a memset node in front of every dispatch
  for (int pr = blockIdx.x * blockDim.x + threadIdx.x; pr < n_pairs; pr += gridDim.x * blockDim.x) {
    const int id = ids[pr];
    if (id < 0 || id >= world * e_local) pair_slot[pr] = -1;
  }
  for (int slot = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; slot < n_slots; slot += (gridDim.x * blockDim.x) >> 5) {
    const int p = slot / cap, j = slot - p * cap;
    const int start = expert_offsets[p * e_local];
    const int cnt = expert_offsets[(p + 1) * e_local] - start;
    if (j == 0 && cnt > cap) {   // capacity below the worst case: the pairs past it are dropped (flagged) and marked
      if (lane == 0) atomicOr(status, 1);
      for (int jj = cap + lane; jj < cnt; jj += 32) pair_slot[sorted_pairs[start + jj]] = -1;
    }
    const long long dst_slot = (long long)rank * cap + j;
    if (j < cnt) {
      const int pair = sorted_pairs[start + j];
      const uint4* src = reinterpret_cast<const uint4*>(x + (long long)(pair / k) * H);
      uint4* dst = reinterpret_cast<uint4*>(peers.rows[p] + dst_slot * H);
      for (int i0 = 0; i0 < vec; i0 += 128) {     // 4 chunks per lane in flight
        uint4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) { const int i = i0 + u * 32 + lane; if (i < vec) v[u] = src[i]; }
#pragma unroll
        for (int u = 0; u < 4; ++u) { const int i = i0 + u * 32 + lane; if (i < vec) dst[i] = v[u]; }
      }
      if (lane == 0) {
        peers.ids[p][dst_slot] = ids[pair] - p * e_local;
        pair_slot[pair] = slot;
      }
    } else if (lane == 0) {
      peers.ids[p][dst_slot] = -1;
    }
  }
  ep_publish_epoch(ctr, peers, world, rank, wait_flags, wait_ctr, status);
}

// Return: row (q, j) of this rank's expert outputs (valid where recv_ids >= 0) -> home rank q's back_rows chunk `rank`.
__global__ void __launch_bounds__(256) ep_return_kernel(const __half* y_rows, const int32_t* recv_ids, int H,
                                                        int world, int rank, int cap, const EpPeers peers,
                                                        unsigned int* __restrict__ ctr, const unsigned long long* wait_flags,
                                                        unsigned int* wait_ctr, int32_t* status) {
  grid_dep_launch();
  grid_dep_wait();
  const int lane = threadIdx.x & 31;
  const int n_slots = world * cap;
  const int vec = H / 8;
  for (int slot = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; slot < n_slots; slot += (gridDim.x * blockDim.x) >> 5) {
    if (recv_ids[slot] < 0) continue;
    const int q = slot / cap, j = slot - q * cap;
    const uint4* src = reinterpret_cast<const uint4*>(y_rows + (long long)slot * H);
    uint4* dst = reinterpret_cast<uint4*>(peers.rows[q] + ((long long)rank * cap + j) * H);
    for (int i0 = 0; i0 < vec; i0 += 128) {
      uint4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) { const int i = i0 + u * 32 + lane; if (i < vec) v[u] = src[i]; }
#pragma unroll
      for (int u = 0; u < 4; ++u) { const int i = i0 + u * 32 + lane; if (i < vec) dst[i] = v[u]; }
    }
  }
  ep_publish_epoch(ctr, peers, world, rank, wait_flags, wait_ctr, status);
}

// One warp: advance this rank's wait epoch (ctr[0]) and spin until every peer's flag has reached it.  A peer that never
// arrives (a crashed rank) raises NF4MOE_STATUS_EP_TIMEOUT after 2^27 polls (about half a minute) and traps -- the stream's next
// synchronisation reports a launch failure -- instead of hanging the GPU for ever or letting the layer run on stale rows.
__global__ void __launch_bounds__(32) ep_wait_kernel(const unsigned long long* __restrict__ flags, int world, unsigned int* __restrict__ ctr,
                                                     int32_t* __restrict__ status) {
  grid_dep_launch();
  grid_dep_wait();
  const int lane = threadIdx.x;
  unsigned int epoch = 0;
  if (lane == 0) { epoch = ctr[0] + 1u; ctr[0] = epoch; }
  epoch = __shfl_sync(0xffffffffu, epoch, 0);
  bool ok = true;
  if (lane < world) {
    unsigned long long spins = 0;
    while (ld_acquire_sys_u64(flags + lane) < (unsigned long long)epoch) {
      if (++spins > (1ull << 27)) { ok = false; break; }   // ~half a minute: a peer died or never launched
      __nanosleep(64);
    }
  }
  if (!ok) {
    atomicOr(status, 8);
    __threadfence_system();
    __trap();
  }
  __threadfence_system();
}

}  // namespace nf4moe
