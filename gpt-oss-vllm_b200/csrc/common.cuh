// Shared device helpers for libnf4moe (sm_100a only).  No CPU fallback exists anywhere in this library.
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include <atomic>

#include "../../include/nf4moe.h"

#ifndef __CUDA_ARCH_FEAT_SM100_ALL
#if defined(__CUDA_ARCH__)
#error "libnf4moe is written for sm_100a (compile with -gencode arch=compute_100a,code=sm_100a)"
#endif
#endif

namespace nf4moe {

constexpr int kBlock = 64;    // NF4 quantisation block (P/unsloth_bnb_loader.py:71,112)
constexpr int kNested = 256;  // double-quant block

// P/streaming_unsloth_bnb.py:101-106 (values only; exactly representable in fp32)
__device__ __constant__ float kNF4[16] = {-1.0f,
                                          -0.6961928009986877f,
                                          -0.5250730514526367f,
                                          -0.39491748809814453f,
                                          -0.28444138169288635f,
                                          -0.18477343022823334f,
                                          -0.09105003625154495f,
                                          0.0f,
                                          0.07958029955625534f,
                                          0.16093020141124725f,
                                          0.24611230194568634f,
                                          0.33791524171829224f,
                                          0.44070982933044434f,
                                          0.5626170039176941f,
                                          0.7229568362236023f,
                                          1.0f};

// Device-side view of nf4moe_qweight with the per-expert strides resolved.
struct QW {
  const uint8_t* packed;
  const float* absmax_f32;
  const uint8_t* absmax_u8;
  const float* absmax2;
  const float* code2;
  const float* offset;
  long long code2_stride;
  int rows, cols;
  int nkb;              // cols / 64
  long long nb;         // blocks per expert = rows * nkb
  long long n2;         // nested blocks per expert = ceil(nb / 256)
  long long packed_stride;  // bytes per expert
};

inline QW make_qw(const nf4moe_qweight& w) {
  QW q;
  q.packed = w.packed;
  q.absmax_f32 = w.absmax_f32;
  q.absmax_u8 = w.absmax_u8;
  q.absmax2 = w.absmax2;
  q.code2 = w.code2;
  q.offset = w.offset;
  q.code2_stride = w.code2_stride;
  q.rows = w.rows;
  q.cols = w.cols;
  q.nkb = w.cols / kBlock;
  q.nb = (long long)w.rows * q.nkb;
  q.n2 = (q.nb + kNested - 1) / kNested;
  q.packed_stride = (long long)w.rows * w.cols / 2;
  return q;
}

// absmax of block b of expert e, bnb double-quant semantics: two separately rounded fp32 ops
// (V/model_executor/model_loader/bitsandbytes_loader.py:599-601).
__device__ __forceinline__ float absmax_of(const QW& q, int e, long long b) {
  if (q.absmax_f32) return __ldg(q.absmax_f32 + (long long)e * q.nb + b);
  const uint8_t c = __ldg(q.absmax_u8 + (long long)e * q.nb + b);
  const float s = __ldg(q.absmax2 + (long long)e * q.n2 + (b >> 8));
  const float v = __ldg(q.code2 + (long long)e * q.code2_stride + c);
  return __fadd_rn(__fmul_rn(v, s), __ldg(q.offset + e));
}

// ---- PTX wrappers -------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// try_wait with a suspend-time hint: the waiting thread sleeps in hardware until the phase completes (or ~20 us pass)
// instead of returning to spin; a spinning warp steals issue slots from the working warps of its SM sub-partition
// (25 % of all instructions issued by the tcgen05 kernel were BRA/SYNCS/YIELD of such loops before this).
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity), "r"(20000u)
      : "memory");
  return ok != 0;
}
// Non-blocking probe (used to look at a barrier early and hide the probe's latency behind other work).
__device__ __forceinline__ bool mbar_test_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
// Programmatic dependent launch: a kernel launched with launch_pdl() may start (be scheduled, run its prologue) while the
// previous kernel of the stream is still draining; grid_dep_wait() blocks until that kernel has completed and its writes
// are visible.  Every thread executes it before touching anything a previous kernel produced (and before writing anything a
// previous kernel may still read).  Without the launch attribute it is a no-op.
__device__ __forceinline__ void grid_dep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
// Lets the NEXT kernel of the stream be scheduled as soon as every CTA of this one has started (it then blocks in its own
// grid_dep_wait() until this grid has completed): launch latency, CTA rasterisation and the next kernel's input-independent
// prologue overlap this kernel.  Called first thing by every kernel that is launched with launch_pdl().
__device__ __forceinline__ void grid_dep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
// A load of something the PREVIOUS kernel wrote, for kernels launched with launch_pdl().  Through a `const T* __restrict__`
// parameter (or __ldg) the compiler treats the location as invariant for the whole kernel and may hoist the load ABOVE
// grid_dep_wait() -- it did, for the bucket metadata in permute_rows_kernel (round 2: LDG.E.CONSTANT ahead of ACQBULK in the
// SASS, stale routing under back-to-back launches).  This form stays where it is written.
// RULE for every kernel in this library: a pointer to anything an earlier kernel of the stream may have written -- activations,
// logits, ids, offsets, sorted pairs, meta, tile lists, h, y -- is a plain `const T*` (no __restrict__) and is read with ordinary
// loads (no __ldg), which the "memory" clobber of grid_dep_wait() keeps behind the wait; `const T* __restrict__` / __ldg are for
// weights, absmax, code tables and biases only (written before any launch).  tests/test_abi_host.py scans the built library's
// SASS for global loads ahead of the wait as a second line of defence.
__device__ __forceinline__ uint4 ld_dep_u4(const uint4* p) {
  uint4 v;
  asm volatile("ld.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ int ld_dep_s32(const int32_t* p) {
  int v;
  asm volatile("ld.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}

__device__ __forceinline__ uint64_t l2_policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}

// TMA 1-D bulk copy global -> shared, completion counted on an mbarrier (SASS: UBLKCP).
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(dst),
      "l"(src), "r"(bytes), "r"(bar), "l"(policy)
      : "memory");
}

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// Shared-memory loads as PURE asm (not volatile, no memory clobber) so ptxas/nvcc may hoist and interleave them
// across the unrolled k-blocks.  Ordering against mbarrier waits is by DATA dependence: every address is derived from
// a value passed through pin() after the wait (pin is volatile, so it cannot move above the wait).
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
  uint32_t v;
  asm("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint2 lds64(uint32_t addr) {
  uint2 v;
  asm("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
  return v;
}
__device__ __forceinline__ float lds_f32(uint32_t addr) {
  float v;
  asm("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint32_t lds_u8(uint32_t addr) {
  uint32_t v;
  asm("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
// Keep a value in a register and order it after preceding volatile asm (mbarrier waits).
__device__ __forceinline__ uint32_t pin(uint32_t x) {
  asm volatile("mov.b32 %0, %0;" : "+r"(x));
  return x;
}
__device__ __forceinline__ void sts32(uint32_t addr, uint32_t v) {
  asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}

__device__ __forceinline__ uint4 ldg_nc_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

// D(16x8,f32) += A(16x16,f16,row) * B(16x8,f16,col)   (SASS: HMMA.16816.F32)
__device__ __forceinline__ void mma_16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                          uint32_t b0, uint32_t b1) {
  asm("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
  uint32_t r;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
  return r;
}

__device__ __forceinline__ uint32_t pack_half2(float lo, float hi) {
  __half2 h = __floats2half2_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}

// gpt-oss activation, V/model_executor/layers/activation.py:389-397; inputs already rounded to fp16 storage.
__device__ __forceinline__ float swiglu_oai(float g, float u) {
  g = fminf(g, 7.0f);
  u = fminf(fmaxf(u, -7.0f), 7.0f);
  const float glu = g * (1.0f / (1.0f + __expf(-1.702f * g)));
  return (u + 1.0f) * glu;
}
// Two (gate, up) pairs packed as half2 -> two results packed as half2.  The clamps are exact in fp16; everything after them
// is the fp32 arithmetic of swiglu_oai above, so the result is bit-identical to the scalar form.
__device__ __forceinline__ uint32_t swiglu_oai_h2(uint32_t g2, uint32_t u2) {
  const __half2 lim = __float2half2_rn(7.0f), nlim = __float2half2_rn(-7.0f);
  const __half2 g = __hmin2(*reinterpret_cast<const __half2*>(&g2), lim);
  const __half2 u = __hmax2(__hmin2(*reinterpret_cast<const __half2*>(&u2), lim), nlim);
  const float2 gf = __half22float2(g), uf = __half22float2(u);
  const float r0 = (uf.x + 1.0f) * (gf.x * (1.0f / (1.0f + __expf(-1.702f * gf.x))));
  const float r1 = (uf.y + 1.0f) * (gf.y * (1.0f / (1.0f + __expf(-1.702f * gf.y))));
  return pack_half2(r0, r1);
}
// the reference wrapper's own variant, P/unsloth_to_vllm_converter.py:312-315
__device__ __forceinline__ float silu_mul(float g, float u) { return g / (1.0f + __expf(-g)) * u; }

// ---- host: launch with the programmatic-stream-serialization attribute ----------------------------------------------------
// Tuning / A-B switches read from the environment exist only in builds made with -DNF4MOE_EXPERIMENTS (scripts/); the shipped
// library never consults the environment on its launch path.
inline bool pdl_enabled() {
#ifdef NF4MOE_EXPERIMENTS
  static const bool on = !(getenv("NF4MOE_PDL") && atoi(getenv("NF4MOE_PDL")) == 0);
  return on;
#else
  return true;
#endif
}
inline int experiment_env(const char* name, int dflt) {
#ifdef NF4MOE_EXPERIMENTS
  const char* s = getenv(name);
  return s ? atoi(s) : dflt;
#else
  (void)name;
  return dflt;
#endif
}
// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-DEVICE attribute of a kernel: set it once per (kernel, device ordinal).
template <auto Kern>
inline cudaError_t ensure_dyn_smem(int bytes) {
  static std::atomic<unsigned long long> done{0ull};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 64 && ((done.load(std::memory_order_acquire) >> dev) & 1ull)) return cudaSuccess;
  e = cudaFuncSetAttribute(Kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess && dev < 64) done.fetch_or(1ull << dev, std::memory_order_release);
  return e;
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl_cluster(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, int cluster,
                                      Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  int n = 0;
  if (pdl_enabled()) {
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  if (cluster > 1) {   // thread-block clusters of `cluster` CTAs along x (grid.x must be a multiple)
    attr[n].id = cudaLaunchAttributeClusterDimension;
    attr[n].val.clusterDim.x = (unsigned)cluster;
    attr[n].val.clusterDim.y = 1;
    attr[n].val.clusterDim.z = 1;
    ++n;
  }
  cfg.attrs = attr;
  cfg.numAttrs = n;
  return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args&&... args) {
  return launch_pdl_cluster(kern, grid, block, smem, stream, 1, static_cast<Args&&>(args)...);
}

}  // namespace nf4moe
