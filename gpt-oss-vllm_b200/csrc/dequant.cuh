// NF4 dequantise / quantise kernels (HBM-bound elementwise work; CUDA cores, 128-bit accesses).
#pragma once
#include <cuda_bf16.h>
#include "common.cuh"

namespace nf4moe {

// ---------------------------------------------------------------------------------------------------------------
// dequant: one thread = 16 packed bytes = 32 weights (half a quantisation block) -> 4 x 16-byte fp16 stores.
// w16[i] = fp16_rn( NF4[q_i] * absmax[i/64] ) with a single fp32 multiply: bit-exact bnb dequantize_4bit semantics
// (structure: P/streaming_unsloth_bnb.py:199-202,228-229; nibble order: bnb, high nibble first).
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) dequant_kernel(QW q, int e0, long long n_chunks_per_expert, long long total_chunks,
                                                      __half* __restrict__ out) {
  __shared__ float lut[16];
  if (threadIdx.x < 16) lut[threadIdx.x] = kNF4[threadIdx.x];
  __syncthreads();
  for (long long c = (long long)blockIdx.x * blockDim.x + threadIdx.x; c < total_chunks;
       c += (long long)gridDim.x * blockDim.x) {
    const int el = (int)(c / n_chunks_per_expert);
    const long long cc = c - (long long)el * n_chunks_per_expert;  // 16-byte chunk inside the expert
    const int e = e0 + el;
    const uint4 p = ldg_nc_v4(q.packed + (long long)e * q.packed_stride + cc * 16);
    const float am = absmax_of(q, e, cc >> 1);
    const uint32_t w[4] = {p.x, p.y, p.z, p.w};
    uint4 o[4];
    uint32_t* ow = reinterpret_cast<uint32_t*>(o);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const uint32_t byte = (w[i] >> (8 * b)) & 0xFFu;
        const float hi = __fmul_rn(lut[byte >> 4], am);   // element 2j
        const float lo = __fmul_rn(lut[byte & 0xF], am);  // element 2j+1
        ow[i * 4 + b] = pack_half2(hi, lo);
      }
    }
    uint4* dst = reinterpret_cast<uint4*>(out + ((long long)el * q.packed_stride + cc * 16) * 2);
#pragma unroll
    for (int i = 0; i < 4; ++i) dst[i] = o[i];
  }
}

__global__ void __launch_bounds__(256) dequant_absmax_kernel(QW q, long long total, float* __restrict__ out) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int e = (int)(i / q.nb);
    out[i] = absmax_of(q, e, i - (long long)e * q.nb);
  }
}

// ---------------------------------------------------------------------------------------------------------------
// quantise (load-time): bnb quantize_4bit(blocksize=64, compress_statistics=True, 'nf4') as invoked at
// P/bnb_moe_4bit.py:86-91.  Pass 1: one warp per 64-block: absmax, nearest NF4 code (midpoints, ties -> lower index),
// pack high-nibble-first, and accumulate sum(absmax) per expert in fp64 for the offset.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) quantize_blocks_kernel(const __half* __restrict__ w, long long nb_per_expert,
                                                              long long total_blocks, uint8_t* __restrict__ packed,
                                                              float* __restrict__ absmax, double* __restrict__ sums) {
  __shared__ double mids[15];
  if (threadIdx.x < 15) mids[threadIdx.x] = ((double)kNF4[threadIdx.x] + (double)kNF4[threadIdx.x + 1]) * 0.5;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long b = warp; b < total_blocks; b += nwarps) {
    const __half2 v = reinterpret_cast<const __half2*>(w + b * 64)[lane];
    const float x0 = __low2float(v), x1 = __high2float(v);
    float m = fmaxf(fabsf(x0), fabsf(x1));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    const float safe = m > 0.f ? m : 1.f;
    const double n0 = (double)__fdiv_rn(x0, safe), n1 = (double)__fdiv_rn(x1, safe);
    int q0 = 0, q1 = 0;
#pragma unroll
    for (int i = 0; i < 15; ++i) {
      q0 += (mids[i] < n0);
      q1 += (mids[i] < n1);
    }
    packed[b * 32 + lane] = (uint8_t)((q0 << 4) | q1);
    if (lane == 0) {
      absmax[b] = m;
      atomicAdd(sums + b / nb_per_expert, (double)m);
    }
  }
}

// Pass 2: one CTA (256 threads) per nested block of 256 absmax values.
__global__ void __launch_bounds__(256) quantize_nested_kernel(const float* __restrict__ absmax, long long nb, long long n2,
                                                              const double* __restrict__ sums,
                                                              const float* __restrict__ code2,
                                                              uint8_t* __restrict__ absmax_u8,
                                                              float* __restrict__ absmax2, float* __restrict__ offset,
                                                              float* __restrict__ absmax_f32_out) {
  __shared__ double mids[255];
  __shared__ float c2[256];
  __shared__ float red[8];
  const int tid = threadIdx.x;
  c2[tid] = code2[tid];
  __syncthreads();
  if (tid < 255) mids[tid] = ((double)c2[tid] + (double)c2[tid + 1]) * 0.5;
  const long long blk = blockIdx.x;  // nested block over all experts
  const int e = (int)(blk / n2);
  const long long j = blk - (long long)e * n2;
  const float off = (float)(sums[e] / (double)nb);
  const long long i = j * 256 + tid;
  const bool valid = i < nb;
  const float centred = valid ? __fsub_rn(absmax[(long long)e * nb + i], off) : 0.f;
  float m = fabsf(centred);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((tid & 31) == 0) red[tid >> 5] = m;
  __syncthreads();
  m = red[0];
#pragma unroll
  for (int k = 1; k < 8; ++k) m = fmaxf(m, red[k]);
  const float safe = m > 0.f ? m : 1.f;
  const double nv = (double)__fdiv_rn(centred, safe);
  int lo = 0, hi = 255;  // number of midpoints strictly below nv
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (mids[mid] < nv) lo = mid + 1; else hi = mid;
  }
  if (valid) {
    absmax_u8[(long long)e * nb + i] = (uint8_t)lo;
    if (absmax_f32_out) absmax_f32_out[(long long)e * nb + i] = __fadd_rn(__fmul_rn(c2[lo], m), off);
  }
  if (tid == 0) {
    absmax2[blk] = m;
    if (j == 0) offset[e] = off;
  }
}

// fp16 <-> bf16 (elementwise).  bf16 -> fp16 is EXACT for every finite value inside the fp16 range (bf16 carries 8 significand
// bits, fp16 11) and saturates at +-65504 outside it; fp16 -> bf16 rounds to nearest even.  This is how bf16 callers are served:
// the kernels of this library take fp16 operands (the reference serves --dtype float16, P/gpt_oss_server.py:41; its config also
// lists bf16, P/bnb_moe_4bit.py:35-37), so a bf16 activation tensor is widened exactly, the layer runs as for fp16 (fp32
// accumulation) and the result is rounded to bf16 once at the end.
__global__ void __launch_bounds__(256) cast_f16_bf16_kernel(const uint16_t* __restrict__ src, uint16_t* __restrict__ dst, long long n,
                                                            int to_bf16) {
  grid_dep_launch();
  grid_dep_wait();
  auto one = [&](uint16_t v) -> uint16_t {
    if (to_bf16) return __bfloat16_as_ushort(__float2bfloat16_rn(__half2float(__ushort_as_half(v))));
    const float f0 = __bfloat162float(__ushort_as_bfloat16(v));
    const float f = f0 != f0 ? f0 : fminf(fmaxf(f0, -65504.f), 65504.f);   // NaN stays NaN (fmaxf / fminf would drop it)
    return __half_as_ushort(__float2half_rn(f));
  };
  const long long nv = n >> 3;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += (long long)gridDim.x * blockDim.x) {
    uint4 v = ld_dep_u4(reinterpret_cast<const uint4*>(src) + i);
    uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) w[j] = (uint32_t)one((uint16_t)(w[j] & 0xFFFFu)) | ((uint32_t)one((uint16_t)(w[j] >> 16)) << 16);
    reinterpret_cast<uint4*>(dst)[i] = make_uint4(w[0], w[1], w[2], w[3]);
  }
  for (long long i = (nv << 3) + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    uint16_t v;
    asm volatile("ld.global.u16 %0, [%1];" : "=h"(v) : "l"(src + i) : "memory");
    dst[i] = one(v);
  }
}

}  // namespace nf4moe
