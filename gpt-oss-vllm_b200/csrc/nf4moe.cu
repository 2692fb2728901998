// libnf4moe.so -- C ABI over the sm_100a kernels (single translation unit).  See include/nf4moe.h for the contract.
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "common.cuh"
#include "dequant.cuh"
#include "ep.cuh"
#include "gemm_tc.cuh"
#include "gemv.cuh"
#include "gemm_dec.cuh"
#include "route.cuh"

using namespace nf4moe;

namespace {

thread_local char g_err[512] = "";
thread_local int g_launches = 0;
std::atomic<int> g_merged_decode{1};   // nf4moe_set_merged_decode

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

int cuda_fail(cudaError_t e, const char* what) {
  snprintf(g_err, sizeof(g_err), "%s: %s", what, cudaGetErrorString(e));
  return (int)e;
}

#define CHECK_LAUNCH(what)                                  \
  do {                                                      \
    ++g_launches;                                           \
    cudaError_t e__ = cudaGetLastError();                   \
    if (e__ != cudaSuccess) return cuda_fail(e__, what);    \
  } while (0)

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

struct DevInfo {
  int sms = 0;
  int cc_major = 0;
};

int device_info(DevInfo* out) {
  static thread_local int cached_dev = -1;
  static thread_local DevInfo cached;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return cuda_fail(e, "cudaGetDevice");
  if (dev != cached_dev) {
    e = cudaDeviceGetAttribute(&cached.sms, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return cuda_fail(e, "cudaDeviceGetAttribute");
    e = cudaDeviceGetAttribute(&cached.cc_major, cudaDevAttrComputeCapabilityMajor, dev);
    if (e != cudaSuccess) return cuda_fail(e, "cudaDeviceGetAttribute");
    cached_dev = dev;
  }
  if (cached.cc_major != 10) return fail(NF4MOE_EARCH, "libnf4moe needs an sm_100 device (got sm_%d0)", cached.cc_major);
  *out = cached;
  return 0;
}

int check_qw(const nf4moe_qweight* w, const char* name) {
  if (!w || !w->packed) return fail(NF4MOE_EINVAL, "%s: null weight", name);
  if (w->rows <= 0 || w->cols <= 0 || w->cols % kBlock) return fail(NF4MOE_EINVAL, "%s: cols must be a multiple of 64", name);
  if (!w->absmax_f32 && !(w->absmax_u8 && w->absmax2 && w->code2 && w->offset))
    return fail(NF4MOE_EINVAL, "%s: give absmax_f32 or the nested {absmax_u8, absmax2, code2, offset}", name);
  if (!aligned16(w->packed) || (w->absmax_f32 && !aligned16(w->absmax_f32)) || (w->absmax_u8 && !aligned16(w->absmax_u8)))
    return fail(NF4MOE_EALIGN, "%s: packed/absmax pointers must be 16-byte aligned", name);
  return 0;
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct Workspace {
  int32_t *status, *ids, *offsets, *sorted, *active, *meta, *tiles;
  float* weights;
  __half *h, *raw, *xs, *y;
  size_t total;
};

Workspace carve(void* base, int T, int E, int H, int I, int k) {
  Workspace w;
  char* p = reinterpret_cast<char*>(base);
  size_t off = 0;
  const size_t n = (size_t)T * k;
  const size_t n_pad = align_up(n, 256) + 256;
  auto take = [&](size_t bytes) {
    char* r = p ? p + off : nullptr;
    off += align_up(bytes, 256);
    return r;
  };
  w.status = (int32_t*)take(256);   // first word: NF4MOE_STATUS_* bits (nf4moe_workspace_status_offset() == 0)
  w.ids = (int32_t*)take(n * 4);
  w.weights = (float*)take(n * 4);
  w.offsets = (int32_t*)take((size_t)(E + 1) * 4);
  w.sorted = (int32_t*)take(n_pad * 4);
  w.active = (int32_t*)take((size_t)E * 4);
  w.meta = (int32_t*)take(64);
  w.tiles = (int32_t*)take((n / 16 + E + 2) * 16);
  w.h = (__half*)take(n_pad * I * 2);
  w.raw = (__half*)take(n_pad * 2 * I * 2);
  w.xs = (__half*)take(n_pad * H * 2);
  w.y = (__half*)take(n * H * 2);
  w.total = off;
  return w;
}

template <int NT, int KBW, bool W13, bool NESTED, int NKB_T>
int launch_gemv_inst(const GemvArgs& args, int sms, cudaStream_t stream) {
  cudaError_t e = ensure_dyn_smem<gemv_kernel<NT, KBW, W13, NESTED, NKB_T>>((int)kGemvSmemBytes);
  if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(gemv)");
  e = launch_pdl(gemv_kernel<NT, KBW, W13, NESTED, NKB_T>, dim3(sms), dim3(kGemvThreads), kGemvSmemBytes, stream, args);
  if (e != cudaSuccess) return cuda_fail(e, "gemv_kernel");
  CHECK_LAUNCH("gemv_kernel");
  return 0;
}

template <int NT, int KBW, bool W13, bool NESTED>
int launch_gemv_k(const GemvArgs& args, int sms, cudaStream_t stream) {
  // K = 2880 (gpt-oss H = I) gets the compile-time specialisation; any other K <= 2880 the runtime-K build
  if (args.w.nkb == 45) return launch_gemv_inst<NT, KBW, W13, NESTED, 45>(args, sms, stream);
  return launch_gemv_inst<NT, KBW, W13, NESTED, 0>(args, sms, stream);
}

// 0 = automatic; 19 = one token tile, 9 k-blocks/warp (three 5-warp teams); 13 / 23 = one / two tiles, 3 k-blocks/warp
// (the override exists in NF4MOE_EXPERIMENTS builds only)
int gemv_config_override() { return experiment_env("NF4MOE_GEMV_CFG", 0); }

template <bool W13>
int launch_gemv(const GemvArgs& args, int T, bool nested, int sms, cudaStream_t stream) {
  // every expert sees at most T rows (top-k experts of a token are distinct): T <= 8 needs one token tile
  int cfg = gemv_config_override();
  if (cfg == 0) cfg = T <= 8 ? 19 : 23;
  switch (cfg) {
    case 19: return nested ? launch_gemv_k<1, 9, W13, true>(args, sms, stream) : launch_gemv_k<1, 9, W13, false>(args, sms, stream);
    case 13: return nested ? launch_gemv_k<1, 3, W13, true>(args, sms, stream) : launch_gemv_k<1, 3, W13, false>(args, sms, stream);
    default: return nested ? launch_gemv_k<2, 3, W13, true>(args, sms, stream) : launch_gemv_k<2, 3, W13, false>(args, sms, stream);
  }
}

int check_gemv_shape(const nf4moe_qweight* w, const char* name) {
  if (int rc = check_qw(w, name)) return rc;
  if (w->cols > kGemvMaxK) return fail(NF4MOE_EINVAL, "%s: decode kernel supports cols <= %d", name, kGemvMaxK);
  if (w->rows % kGemvRows) return fail(NF4MOE_EINVAL, "%s: rows must be a multiple of 16", name);
  return 0;
}

int check_gemv_tokens(int T, int k, int E, const char* name) {
  // (expert, 8-token chunk) entries the kernel lists in shared memory; beyond this the tcgen05 regime is the right tool
  if ((long long)E + ((long long)T * k + 7) / 8 > kMaxVirt)
    return fail(NF4MOE_EINVAL, "%s: T*k = %lld is past the decode kernel's limit (E + T*k/8 <= %d); use the GEMM regime", name,
                (long long)T * k, kMaxVirt);
  return 0;
}

}  // namespace

extern "C" {

int nf4moe_version(void) { return NF4MOE_VERSION; }

constexpr long long kBucketClusterMinPairs = 2048;   // pairs from which the 8-CTA cluster bucket kernel replaces the single CTA

// What NF4MOE_REGIME_AUTO resolves to (measured on B200, gpt-oss-20b shapes, profiles/r1c_kernel_times.txt):
//   T <= 8                       mma.sync GEMV (few active experts, 3.3-3.5 TB/s; the tcgen05 pipeline's fill costs more)
//   token tiles of <= 32 rows    decode-regime tcgen05 kernel (gemm_dec.cuh; FAST operand), 3.1-3.3 TB/s vs the GEMV's 2.4-2.9
//   larger                       prefill tcgen05 grouped GEMM with the EXACT (bit-identical dequant) operand
int nf4moe_auto_regime(int T, int k, int E) { return nf4moe_auto_regime_for(T, k, E, 0); }
// nested_absmax != 0: both weight sets carry the nested (double-quant) absmax, so the decode kernel may use 64-row token tiles and
// its regime (FAST operand) extends to a mean of kDec64MaxMeanRows routed rows per expert (measured, E = 32: 156 us per layer at
// T = 384 and 204 us at T = 512 against 240 / 250 us through the prefill grouped GEMM)
int nf4moe_auto_regime_for(int T, int k, int E, int nested_absmax) {
  if (T <= kDecAutoMinTokens) return NF4MOE_REGIME_GEMV;
  if (dec_serves(T, k, E, NF4MOE_GEMM_FAST, nested_absmax != 0)) return NF4MOE_REGIME_GEMM_FAST;
  return NF4MOE_REGIME_GEMM;
}
// profiling hook (not part of the public header): 64 int64 cycle counters written by CTA 0 of the tcgen05 kernels
void nf4moe_debug_set_gemm_counters(long long* dev_ptr) { gemm_debug_buffer() = dev_ptr; }
#ifdef NF4MOE_EXPERIMENTS
extern "C" int nf4moe_debug_bucket_stamps(long long* out8) { return (int)cudaMemcpyFromSymbol(out8, g_bucket_stamps, 8 * sizeof(long long)); }
#endif
const char* nf4moe_last_error(void) { return g_err; }
int nf4moe_last_launch_count(void) { return g_launches; }
int nf4moe_set_merged_decode(int enabled) { return g_merged_decode.exchange(enabled ? 1 : 0, std::memory_order_relaxed); }

size_t nf4moe_workspace_bytes(int T, int E, int H, int I, int k) {
  if (T < 0 || E <= 0 || H <= 0 || I <= 0 || k <= 0) return 0;
  return carve(nullptr, T, E, H, I, k).total;
}

size_t nf4moe_workspace_status_offset(int T, int E, int H, int I, int k) {
  if (T < 0 || E <= 0 || H <= 0 || I <= 0 || k <= 0) return 0;
  return 0;   // the status word is the first word of the workspace for every geometry
}

int nf4moe_dequant(const nf4moe_qweight* w, int e0, int ne, void* out_f16, void* stream) {
  if (int rc = check_qw(w, "dequant")) return rc;
  if (!out_f16 || e0 < 0 || ne < 0) return fail(NF4MOE_EINVAL, "dequant: bad arguments");
  if (!aligned16(out_f16)) return fail(NF4MOE_EALIGN, "dequant: out must be 16-byte aligned");
  if (ne == 0) return 0;
  DevInfo di;
  if (int rc = device_info(&di)) return rc;
  const QW q = make_qw(*w);
  const long long per = q.packed_stride / 16, total = per * ne;
  const int grid = (int)std::min<long long>((total + 255) / 256, (long long)di.sms * 16);
  dequant_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(q, e0, per, total, (__half*)out_f16);
  CHECK_LAUNCH("dequant_kernel");
  return 0;
}

int nf4moe_dequant_absmax(const nf4moe_qweight* w, int E, float* out, void* stream) {
  if (!w || !out || E <= 0) return fail(NF4MOE_EINVAL, "dequant_absmax: bad arguments");
  if (!(w->absmax_u8 && w->absmax2 && w->code2 && w->offset)) return fail(NF4MOE_EINVAL, "dequant_absmax: nested state required");
  DevInfo di;
  if (int rc = device_info(&di)) return rc;
  nf4moe_qweight tmp = *w;
  tmp.absmax_f32 = nullptr;
  const QW q = make_qw(tmp);
  const long long total = q.nb * E;
  const int grid = (int)std::min<long long>((total + 255) / 256, (long long)di.sms * 16);
  dequant_absmax_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(q, total, out);
  CHECK_LAUNCH("dequant_absmax_kernel");
  return 0;
}

int nf4moe_quantize(const void* w_f16, int E, int rows, int cols, uint8_t* packed, uint8_t* absmax_u8, float* absmax2,
                    const float* code2, float* offset, float* absmax_f32_out, void* scratch, void* stream) {
  if (!w_f16 || !packed || !absmax_u8 || !absmax2 || !code2 || !offset || !scratch || E <= 0 || rows <= 0 || cols <= 0 ||
      cols % kBlock)
    return fail(NF4MOE_EINVAL, "quantize: bad arguments");
  DevInfo di;
  if (int rc = device_info(&di)) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const long long nb = (long long)rows * (cols / kBlock), n2 = (nb + kNested - 1) / kNested, total = nb * E;
  double* sums = reinterpret_cast<double*>(scratch);
  float* am = reinterpret_cast<float*>(reinterpret_cast<char*>(scratch) + align_up((size_t)E * 8, 256));
  cudaError_t e = cudaMemsetAsync(sums, 0, (size_t)E * 8, st);
  if (e != cudaSuccess) return cuda_fail(e, "cudaMemsetAsync");
  const int grid = (int)std::min<long long>((total + 7) / 8, (long long)di.sms * 32);
  quantize_blocks_kernel<<<grid, 256, 0, st>>>((const __half*)w_f16, nb, total, packed, am, sums);
  CHECK_LAUNCH("quantize_blocks_kernel");
  quantize_nested_kernel<<<(unsigned)(n2 * E), 256, 0, st>>>(am, nb, n2, sums, code2, absmax_u8, absmax2, offset,
                                                            absmax_f32_out);
  CHECK_LAUNCH("quantize_nested_kernel");
  return 0;
}

int nf4moe_cast_f16_bf16(const void* src, void* dst, long long n, int to_bf16, void* stream) {
  if (!src || !dst || n < 0) return fail(NF4MOE_EINVAL, "cast_f16_bf16: bad arguments");
  if (!aligned16(src) || !aligned16(dst)) return fail(NF4MOE_EALIGN, "cast_f16_bf16: src and dst must be 16-byte aligned");
  if (n == 0) return 0;
  DevInfo di;
  if (int rc = device_info(&di)) return rc;
  const int grid = (int)std::min<long long>((n / 8 + 255) / 256 + 1, (long long)di.sms * 8);
  launch_pdl(cast_f16_bf16_kernel, dim3(grid), dim3(256), 0, (cudaStream_t)stream, (const uint16_t*)src, (uint16_t*)dst, n, to_bf16 ? 1 : 0);
  CHECK_LAUNCH("cast_f16_bf16_kernel");
  return 0;
}

int nf4moe_route_topk(const void* logits, int dtype, int T, int E, int k, int32_t* ids, float* weights, void* stream) {
  if (!logits || !ids || !weights || T < 0 || k <= 0 || k > 32 || k > E || dtype < 0 || dtype > 2)
    return fail(NF4MOE_EINVAL, "route_topk: bad arguments");
  if (E <= 0 || E > 256) return fail(NF4MOE_EINVAL, "route_topk: E must be in 1..256 (got %d)", E);
  if (T == 0) return 0;
  const int grid = (T + 7) / 8;
  cudaStream_t st = (cudaStream_t)stream;
  const int per_lane = (E + 31) / 32;
  switch (per_lane <= 2 ? per_lane : per_lane <= 4 ? 4 : 8) {
    case 1: launch_pdl(route_topk_kernel<1>, dim3(grid), dim3(256), 0, st, logits, dtype, T, E, k, ids, weights); break;
    case 2: launch_pdl(route_topk_kernel<2>, dim3(grid), dim3(256), 0, st, logits, dtype, T, E, k, ids, weights); break;
    case 4: launch_pdl(route_topk_kernel<4>, dim3(grid), dim3(256), 0, st, logits, dtype, T, E, k, ids, weights); break;
    case 8: launch_pdl(route_topk_kernel<8>, dim3(grid), dim3(256), 0, st, logits, dtype, T, E, k, ids, weights); break;
    default: return fail(NF4MOE_EINVAL, "route_topk: unsupported E %d", E);
  }
  CHECK_LAUNCH("route_topk_kernel");
  return 0;
}

int nf4moe_bucket(const int32_t* ids, int T, int k, int E, int32_t* expert_offsets, int32_t* sorted_pairs,
                  int32_t* active_experts, int32_t* meta, int32_t* tile_list, int tile_rows, void* stream) {
  if (!ids || !expert_offsets || !sorted_pairs || !active_experts || !meta || T < 0 || k <= 0 || E <= 0 || E > 255)
    return fail(NF4MOE_EINVAL, "bucket: bad arguments (E must be <= 255)");
  if (tile_rows == 0) tile_rows = gemm_tile_tokens(T, k, E);   // the size that suits either absmax form
  if (tile_rows < 16 || tile_rows > 256 || tile_rows % 16)
    return fail(NF4MOE_EINVAL, "bucket: tile_rows must be a multiple of 16 in 16..256 (or 0 = automatic)");
  const long long n = (long long)T * k;
  const size_t smem = (size_t)((n + 127) & ~127ll);
  if (smem > 200 * 1024) return fail(NF4MOE_EINVAL, "bucket: T*k = %lld exceeds the limit of 204800 pairs", n);
  const bool cluster_on = experiment_env("NF4MOE_BUCKET_CLUSTER", 1) != 0;
  if (cluster_on && n >= kBucketClusterMinPairs) {   // prefill-sized: one cluster of 8 CTAs, a chunk of the pair list each
    const size_t chunk = (size_t)((((n + kBucketCluster - 1) / kBucketCluster) + 127) & ~127ll);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(kBucketCluster);
    cfg.blockDim = dim3(kBucketThreads);
    cfg.dynamicSmemBytes = chunk;
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = kBucketCluster; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 2 : 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, bucket_cluster_kernel, ids, (int)n, E, expert_offsets, sorted_pairs, active_experts, meta,
                                       tile_list, tile_rows);
    if (e != cudaSuccess) return cuda_fail(e, "bucket_cluster_kernel");
    CHECK_LAUNCH("bucket_cluster_kernel");
    return 0;
  }
  cudaError_t e = ensure_dyn_smem<bucket_kernel>(200 * 1024);
  if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(bucket)");
  launch_pdl(bucket_kernel, dim3(1), dim3(kBucketThreads), smem, (cudaStream_t)stream, ids, (int)n, E, expert_offsets, sorted_pairs,
             active_experts, meta, tile_list, tile_rows);
  CHECK_LAUNCH("bucket_kernel");
  return 0;
}

int nf4moe_permute_rows(const void* x, const int32_t* sorted_pairs, const int32_t* meta, int T, int k, int H,
                        void* x_sorted, void* stream) {
  if (!x || !sorted_pairs || !meta || !x_sorted || T < 0 || k <= 0 || H <= 0 || H % 8)
    return fail(NF4MOE_EINVAL, "permute_rows: bad arguments");
  if (!aligned16(x) || !aligned16(x_sorted)) return fail(NF4MOE_EALIGN, "permute_rows: 16-byte alignment required");
  if (T == 0) return 0;
  DevInfo di;
  if (int rc = device_info(&di)) return rc;
  const long long rows = (long long)T * k;
  const int grid = (int)std::min<long long>(rows, (long long)di.sms * 16);
  launch_pdl(permute_rows_kernel, dim3(grid), dim3(128), 0, (cudaStream_t)stream, (const __half*)x, sorted_pairs, meta, (int)rows, k, H,
             (__half*)x_sorted);
  CHECK_LAUNCH("permute_rows_kernel");
  return 0;
}

static int combine_impl(const void* y, const float* weights, const int32_t* ids, int E, int T, int k, int H, void* out,
                        const void* w2_bias, void* stream);

int nf4moe_combine(const void* y, const float* weights, const int32_t* ids, int E, int T, int k, int H, void* out,
                   void* stream) {
  return combine_impl(y, weights, ids, E, T, k, H, out, nullptr, stream);
}

static int combine_impl(const void* y, const float* weights, const int32_t* ids, int E, int T, int k, int H, void* out,
                        const void* w2_bias, void* stream) {
  if (!y || !weights || !out || T < 0 || k <= 0 || H <= 0 || H % 8) return fail(NF4MOE_EINVAL, "combine: bad arguments");
  if (w2_bias && (!ids || !aligned16(w2_bias))) return fail(NF4MOE_EINVAL, "combine: w2_bias needs topk ids and 16-byte alignment");
  if (!aligned16(y) || !aligned16(out)) return fail(NF4MOE_EALIGN, "combine: 16-byte alignment required");
  if (T == 0) return 0;
  DevInfo di;
  if (int rc = device_info(&di)) return rc;
  const long long total = (long long)T * (H / 8);
  const int grid = (int)std::min<long long>((total + 255) / 256, (long long)di.sms * 8);
  launch_pdl(combine_kernel, dim3(grid), dim3(256), 0, (cudaStream_t)stream, (const __half*)y, weights, ids, E, T, k, H, (__half*)out,
             (const __half*)w2_bias);
  CHECK_LAUNCH("combine_kernel");
  return 0;
}

int nf4moe_gemv_w13_swiglu(const void* x, const nf4moe_qweight* w13, const int32_t* expert_offsets,
                           const int32_t* sorted_pairs, const int32_t* active_experts, const int32_t* meta, int T, int k,
                           int E, int act, void* h_sorted, void* stream) {
  if (int rc = check_gemv_shape(w13, "gemv_w13")) return rc;
  if (!x || !expert_offsets || !sorted_pairs || !active_experts || !meta || !h_sorted || T < 0 || k <= 0 || E <= 0 ||
      (w13->rows & 1) || (act != NF4MOE_ACT_SWIGLU_OAI && act != NF4MOE_ACT_CHUNK_SILU))
    return fail(NF4MOE_EINVAL, "gemv_w13: bad arguments");
  if (int rc = check_gemv_tokens(T, k, E, "gemv_w13")) return rc;
  if (!aligned16(x)) return fail(NF4MOE_EALIGN, "gemv_w13: x must be 16-byte aligned");
  if (T == 0) return 0;
  DevInfo di;
  if (int rc = device_info(&di)) return rc;
  GemvArgs a;
  a.w = make_qw(*w13);
  a.x = (const __half*)x;
  a.out = h_sorted;
  a.expert_offsets = expert_offsets;
  a.sorted_pairs = sorted_pairs;
  a.active_experts = active_experts;
  a.meta = meta;
  a.k = k;
  a.act = act;
  return launch_gemv<true>(a, T, w13->absmax_f32 == nullptr, di.sms, (cudaStream_t)stream);
}

int nf4moe_gemv_w2(const void* h_sorted, const nf4moe_qweight* w2, const int32_t* expert_offsets,
                   const int32_t* sorted_pairs, const int32_t* active_experts, const int32_t* meta, int T, int k, int E,
                   void* y, void* stream) {
  if (int rc = check_gemv_shape(w2, "gemv_w2")) return rc;
  if (!h_sorted || !expert_offsets || !sorted_pairs || !active_experts || !meta || !y || T < 0 || k <= 0 || E <= 0)
    return fail(NF4MOE_EINVAL, "gemv_w2: bad arguments");
  if (int rc = check_gemv_tokens(T, k, E, "gemv_w2")) return rc;
  if (!aligned16(h_sorted)) return fail(NF4MOE_EALIGN, "gemv_w2: h must be 16-byte aligned");
  if (T == 0) return 0;
  DevInfo di;
  if (int rc = device_info(&di)) return rc;
  GemvArgs a;
  a.w = make_qw(*w2);
  a.x = (const __half*)h_sorted;
  a.out = y;
  a.expert_offsets = expert_offsets;
  a.sorted_pairs = sorted_pairs;
  a.active_experts = active_experts;
  a.meta = meta;
  a.k = k;
  a.act = 0;
  return launch_gemv<false>(a, T, w2->absmax_f32 == nullptr, di.sms, (cudaStream_t)stream);
}

// tile_rows: the token-tile size the tile list was built with (0 = what the library derives from (T, k, E)); <= 32 rows go to the
// decode-regime kernel, larger tiles to the prefill grouped GEMM
static int gemm_w13_any(const void* x_sorted, const nf4moe_qweight* w13, const int32_t* meta, const int32_t* tile_list, int T,
                        int k, int E, int act, int mode, void* h_sorted, void* stream, const void* bias, int32_t* status,
                        int tile_rows = 0) {
  if (tile_rows > 0 ? tile_rows <= kDecMaxTile : dec_serves(T, k, E, mode))
    return dec_launch<true>(x_sorted, w13, nullptr, meta, tile_list, T, k, E, act, mode, h_sorted, stream, g_err, sizeof(g_err),
                            &g_launches, bias, status, tile_rows > 0 ? tile_rows : 32);
  if (bias) return fail(NF4MOE_EINVAL, "gemm_w13: the fused bias exists in the decode-regime kernel only");
  return gemm_launch<true>(x_sorted, w13, nullptr, meta, tile_list, T, k, E, act, mode, h_sorted, stream, g_err, sizeof(g_err),
                           &g_launches, status, tile_rows);
}

static int gemm_w2_any(const void* h_sorted, const nf4moe_qweight* w2, const int32_t* sorted_pairs, const int32_t* meta,
                       const int32_t* tile_list, int T, int k, int E, int mode, void* y, void* stream, int32_t* status,
                       int tile_rows = 0) {
  if (tile_rows > 0 ? tile_rows <= kDecMaxTile : dec_serves(T, k, E, mode))
    return dec_launch<false>(h_sorted, w2, sorted_pairs, meta, tile_list, T, k, E, 0, mode, y, stream, g_err, sizeof(g_err),
                             &g_launches, nullptr, status, tile_rows > 0 ? tile_rows : 32);
  return gemm_launch<false>(h_sorted, w2, sorted_pairs, meta, tile_list, T, k, E, 0, mode, y, stream, g_err, sizeof(g_err),
                            &g_launches, status, tile_rows);
}

int nf4moe_gemm_w13_swiglu(const void* x_sorted, const nf4moe_qweight* w13, const int32_t* expert_offsets,
                           const int32_t* meta, const int32_t* tile_list, int T, int k, int E, int act, int mode,
                           void* h_sorted, void* stream) {
  (void)expert_offsets;
  return gemm_w13_any(x_sorted, w13, meta, tile_list, T, k, E, act, mode, h_sorted, stream, nullptr, nullptr);
}

int nf4moe_gemm_w2(const void* h_sorted, const nf4moe_qweight* w2, const int32_t* expert_offsets,
                   const int32_t* sorted_pairs, const int32_t* meta, const int32_t* tile_list, int T, int k, int E,
                   int mode, void* y, void* stream) {
  (void)expert_offsets;
  return gemm_w2_any(h_sorted, w2, sorted_pairs, meta, tile_list, T, k, E, mode, y, stream, nullptr);
}

int nf4moe_gemm_w13_w2(const void* x_sorted, const nf4moe_qweight* w13, const void* w13_bias, void* h_sorted,
                       const nf4moe_qweight* w2, void* y, const int32_t* sorted_pairs, const int32_t* meta, int32_t* tile_list,
                       int T, int k, int E, int act, int mode, void* stream) {
  g_launches = 0;
  if (w13_bias && !aligned16(w13_bias)) return fail(NF4MOE_EALIGN, "gemm_w13_w2: w13_bias must be 16-byte aligned");
  return dec_launch_both(x_sorted, w13, h_sorted, w2, y, sorted_pairs, meta, tile_list, T, k, E, act, mode, stream, g_err,
                         sizeof(g_err), &g_launches, w13_bias, nullptr);
}

// bucket -> (permute) -> w13 (+activation) -> w2 for the (row, slot) pairs of `ids` [T, k]; y fp16 [T*k, H] in pair order.
// policy_T: the token count the regime and the token-tile size are chosen for (an expert-parallel owner passes the rows it
// EXPECTS, not the capacity of its receive buffer); buffers and TMA maps are always sized by T.
static int run_experts(const void* x, const int32_t* ids, const nf4moe_qweight* w13, const nf4moe_qweight* w2, void* y, int T,
                       int E, int H, int I, int k, int act, int regime, const Workspace& ws, void* stream, int* launches,
                       const void* w13_bias = nullptr, int policy_T = 0, bool prebucketed = false) {
  cudaStream_t st = (cudaStream_t)stream;
  if (policy_T <= 0 || policy_T > T) policy_T = T;
  if (regime == NF4MOE_REGIME_AUTO) {
    // the tcgen05 kernels need 16-row / 64-column granularity and K <= 4096; other shapes stay on the GEMV
    const bool tc_ok = w13->rows % 16 == 0 && w2->rows % 16 == 0 && w13->cols % kBlock == 0 && w2->cols % kBlock == 0 &&
                       w13->cols <= 4096 && w2->cols <= 4096;
    regime = tc_ok ? nf4moe_auto_regime_for(policy_T, k, E, w13->absmax_f32 == nullptr && w2->absmax_f32 == nullptr && act == NF4MOE_ACT_SWIGLU_OAI)
                   : NF4MOE_REGIME_GEMV;
  }
  if (regime != NF4MOE_REGIME_GEMV && regime != NF4MOE_REGIME_GEMM && regime != NF4MOE_REGIME_GEMM_FAST)
    return fail(NF4MOE_EINVAL, "unknown regime %d", regime);
  const bool gemm = regime != NF4MOE_REGIME_GEMV;
  const int mode = regime == NF4MOE_REGIME_GEMM_FAST ? NF4MOE_GEMM_FAST : NF4MOE_GEMM_EXACT;
  // 64-row token tiles: decode kernel, FAST operand, nested absmax on both sides (what fits its shared memory); the EXACT regime keeps
  // round 1's 32 / 160 split, and so do the stand-alone nf4moe_bucket / nf4moe_gemm_* entry points (they cannot know the absmax form)
  const bool wide_ok = regime == NF4MOE_REGIME_GEMM_FAST && w13->absmax_f32 == nullptr && w2->absmax_f32 == nullptr &&
                       act == NF4MOE_ACT_SWIGLU_OAI;
  const int tile_rows = gemm ? gemm_tile_tokens(policy_T, k, E, wide_ok) : 0;
  if (prebucketed) {   // route_bucket_small_kernel already sorted the pairs (GEMV regime only: it writes no tile list)
    if (gemm) return fail(NF4MOE_EINVAL, "run_experts: the fused router + bucket launch serves the GEMV regime only");
  } else {
    if (int rc = nf4moe_bucket(ids, T, k, E, ws.offsets, ws.sorted, ws.active, ws.meta, gemm ? ws.tiles : nullptr, tile_rows, stream)) return rc;
    *launches += g_launches; g_launches = 0;
  }
  // with an up-projection bias the w13 kernels store their raw products and bias_act_kernel applies bias + activation
  // (the decode tcgen05 kernel adds the bias inside its fused SwiGLU epilogue instead)
  const bool bias_fused = w13_bias && gemm && act == NF4MOE_ACT_SWIGLU_OAI && tile_rows <= kDecMaxTile;
  const bool bias_split = w13_bias && !bias_fused;
  const int act_k = bias_split ? NF4MOE_ACT_CHUNK_SILU : act;   // CHUNK_SILU = "store raw [rows, 2I]" for the w13 kernels
  __half* h_dst = act_k == NF4MOE_ACT_CHUNK_SILU ? ws.raw : ws.h;
  if (gemm) {
    if (int rc = nf4moe_permute_rows(x, ws.sorted, ws.meta, T, k, H, ws.xs, stream)) return rc;
    *launches += g_launches; g_launches = 0;
    // decode regime with the fused activation: w13 and w2 in ONE persistent launch (gemm_dec.cuh, DecSide)
    if (tile_rows <= kDecMaxTile && act == NF4MOE_ACT_SWIGLU_OAI && !bias_split && (w13->absmax_f32 == nullptr) == (w2->absmax_f32 == nullptr) &&
        g_merged_decode.load(std::memory_order_relaxed) != 0 && experiment_env("NF4MOE_DEC_MERGED", 1) != 0) {
      if (int rc = dec_launch_both(ws.xs, w13, ws.h, w2, y, ws.sorted, ws.meta, ws.tiles, T, k, E, act, mode, stream, g_err, sizeof(g_err),
                                   &g_launches, bias_fused ? w13_bias : nullptr, ws.status, tile_rows)) return rc;
      *launches += g_launches; g_launches = 0;
      return 0;
    }
    if (int rc = gemm_w13_any(ws.xs, w13, ws.meta, ws.tiles, T, k, E, act_k, mode, h_dst, stream, bias_fused ? w13_bias : nullptr,
                              ws.status, tile_rows)) return rc;
  } else {
    if (int rc = nf4moe_gemv_w13_swiglu(x, w13, ws.offsets, ws.sorted, ws.active, ws.meta, T, k, E, act_k, h_dst, stream)) return rc;
  }
  *launches += g_launches; g_launches = 0;
  if (bias_split) {
    DevInfo di;
    if (int rc = device_info(&di)) return rc;
    launch_pdl(bias_act_kernel, dim3(di.sms * 4), dim3(256), 0, st, (const __half*)ws.raw, (const __half*)w13_bias, ids,
               (const int32_t*)ws.sorted, (const int32_t*)ws.meta, I, act, ws.h);
    CHECK_LAUNCH("bias_act_kernel");
    *launches += g_launches; g_launches = 0;
  } else if (act == NF4MOE_ACT_CHUNK_SILU) {
    DevInfo di;
    if (int rc = device_info(&di)) return rc;
    chunk_silu_kernel<<<di.sms * 4, 256, 0, st>>>(ws.raw, ws.meta, I, ws.h);
    CHECK_LAUNCH("chunk_silu_kernel");
    *launches += g_launches; g_launches = 0;
  }
  if (gemm) {
    if (int rc = gemm_w2_any(ws.h, w2, ws.sorted, ws.meta, ws.tiles, T, k, E, mode, y, stream, ws.status, tile_rows)) return rc;
  } else {
    if (int rc = nf4moe_gemv_w2(ws.h, w2, ws.offsets, ws.sorted, ws.active, ws.meta, T, k, E, y, stream)) return rc;
  }
  *launches += g_launches; g_launches = 0;
  return 0;
}

static int check_layer_shapes(const char* name, const nf4moe_qweight* w13, const nf4moe_qweight* w2, int H, int I) {
  if (w13->rows != 2 * I || w13->cols != H || w2->rows != H || w2->cols != I)
    return fail(NF4MOE_EINVAL, "%s: weight shapes must be w13 [2I,H] = [%d,%d], w2 [H,I] = [%d,%d]", name, 2 * I, H, H, I);
  return 0;
}

int nf4moe_moe_forward(const void* x, const void* router_logits, int logits_dtype, const int32_t* topk_ids,
                       const float* topk_weights, const nf4moe_qweight* w13, const nf4moe_qweight* w2, void* out, int T,
                       int E, int H, int I, int k, int act, int regime, void* workspace, size_t workspace_bytes,
                       int32_t* topk_ids_out, float* topk_weights_out, void* stream) {
  return nf4moe_moe_forward_bias(x, router_logits, logits_dtype, topk_ids, topk_weights, w13, w2, nullptr, nullptr, out, T, E, H, I, k,
                                 act, regime, workspace, workspace_bytes, topk_ids_out, topk_weights_out, stream);
}

int nf4moe_moe_forward_bias(const void* x, const void* router_logits, int logits_dtype, const int32_t* topk_ids,
                            const float* topk_weights, const nf4moe_qweight* w13, const nf4moe_qweight* w2, const void* w13_bias,
                            const void* w2_bias, void* out, int T, int E, int H, int I, int k, int act, int regime, void* workspace,
                            size_t workspace_bytes, int32_t* topk_ids_out, float* topk_weights_out, void* stream) {
  g_launches = 0;
  if ((w13_bias && !aligned16(w13_bias)) || (w2_bias && !aligned16(w2_bias)))
    return fail(NF4MOE_EALIGN, "moe_forward: biases must be 16-byte aligned");
  if (!x || !out || !w13 || !w2 || T < 0 || E <= 0 || H <= 0 || I <= 0 || k <= 0)
    return fail(NF4MOE_EINVAL, "moe_forward: bad arguments");
  if (!router_logits && !(topk_ids && topk_weights))
    return fail(NF4MOE_EINVAL, "moe_forward: give router_logits or (topk_ids, topk_weights)");
  if (int rc = check_layer_shapes("moe_forward", w13, w2, H, I)) return rc;
  if (T == 0) return 0;
  const Workspace ws = carve(workspace, T, E, H, I, k);
  if (!workspace || workspace_bytes < ws.total)
    return fail(NF4MOE_EWORKSPACE, "moe_forward: workspace needs %zu bytes, got %zu", ws.total, workspace_bytes);
  if (!aligned16(workspace)) return fail(NF4MOE_EALIGN, "moe_forward: workspace must be 16-byte aligned");
  int launches = 0;
  const int32_t* ids = topk_ids;
  const float* wts = topk_weights;
  bool prebucketed = false;
  if (router_logits) {
    int32_t* ids_w = topk_ids_out ? topk_ids_out : ws.ids;
    float* wts_w = topk_weights_out ? topk_weights_out : ws.weights;
    // batch-1..8 decode (always the GEMV regime under AUTO): router and bucketing in ONE single-CTA launch
    const bool tiny = T <= 8 && T * k <= 32 && k <= 32 && E <= 255 && (regime == NF4MOE_REGIME_AUTO || regime == NF4MOE_REGIME_GEMV) &&
                      logits_dtype >= 0 && logits_dtype <= 2 && experiment_env("NF4MOE_ROUTE_BUCKET", 1) != 0;
    if (tiny) {
      cudaStream_t st = (cudaStream_t)stream;
      const int per_lane = (E + 31) / 32;
#define NF4MOE_RB(P) launch_pdl(route_bucket_small_kernel<P>, dim3(1), dim3(256), 0, st, router_logits, logits_dtype, T, E, k, ids_w, wts_w, \
                                ws.offsets, ws.sorted, ws.active, ws.meta)
      switch (per_lane <= 2 ? per_lane : per_lane <= 4 ? 4 : 8) {
        case 1: NF4MOE_RB(1); break;
        case 2: NF4MOE_RB(2); break;
        case 4: NF4MOE_RB(4); break;
        default: NF4MOE_RB(8); break;
      }
#undef NF4MOE_RB
      CHECK_LAUNCH("route_bucket_small_kernel");
      prebucketed = true;
    } else {
      if (int rc = nf4moe_route_topk(router_logits, logits_dtype, T, E, k, ids_w, wts_w, stream)) return rc;
    }
    launches += g_launches; g_launches = 0;
    ids = ids_w;
    wts = wts_w;
  }
  if (int rc = run_experts(x, ids, w13, w2, ws.y, T, E, H, I, k, act, regime, ws, stream, &launches, w13_bias, 0, prebucketed)) return rc;
  if (int rc = combine_impl(ws.y, wts, ids, E, T, k, H, out, w2_bias, stream)) return rc;
  g_launches += launches;
  return 0;
}

int nf4moe_expert_rows(const void* rows, const int32_t* row_expert, const nf4moe_qweight* w13, const nf4moe_qweight* w2,
                       void* y_rows, int n_rows, int E, int H, int I, int act, int regime, void* workspace,
                       size_t workspace_bytes, void* stream) {
  return nf4moe_expert_rows_bias(rows, row_expert, w13, w2, nullptr, y_rows, n_rows, 0, E, H, I, act, regime, workspace, workspace_bytes,
                                 stream);
}

int nf4moe_expert_rows_bias(const void* rows, const int32_t* row_expert, const nf4moe_qweight* w13, const nf4moe_qweight* w2,
                            const void* w13_bias, void* y_rows, int n_rows, int rows_hint, int E, int H, int I, int act, int regime,
                            void* workspace, size_t workspace_bytes, void* stream) {
  g_launches = 0;
  if (w13_bias && !aligned16(w13_bias)) return fail(NF4MOE_EALIGN, "expert_rows: w13_bias must be 16-byte aligned");
  if (!rows || !row_expert || !y_rows || !w13 || !w2 || n_rows < 0 || E <= 0 || H <= 0 || I <= 0)
    return fail(NF4MOE_EINVAL, "expert_rows: bad arguments");
  if (int rc = check_layer_shapes("expert_rows", w13, w2, H, I)) return rc;
  if (n_rows == 0) return 0;
  const Workspace ws = carve(workspace, n_rows, E, H, I, 1);
  if (!workspace || workspace_bytes < ws.total)
    return fail(NF4MOE_EWORKSPACE, "expert_rows: workspace needs %zu bytes, got %zu", ws.total, workspace_bytes);
  if (!aligned16(workspace) || !aligned16(rows) || !aligned16(y_rows))
    return fail(NF4MOE_EALIGN, "expert_rows: rows, y_rows and workspace must be 16-byte aligned");
  int launches = 0;
  if (int rc = run_experts(rows, row_expert, w13, w2, y_rows, n_rows, E, H, I, 1, act, regime, ws, stream, &launches, w13_bias, rows_hint)) return rc;
  g_launches += launches;
  return 0;
}

int nf4moe_ep_pack(const void* x, const int32_t* ids, const int32_t* expert_offsets, const int32_t* sorted_pairs, int T, int k,
                   int E, int H, int world, int cap, void* send_rows, int32_t* send_ids, int32_t* pair_slot,
                   int32_t* status, void* stream) {
  if (!x || !ids || !expert_offsets || !sorted_pairs || !send_rows || !send_ids || !pair_slot || !status || T < 0 || k <= 0 ||
      E <= 0 || H <= 0 || H % 8 || world <= 0 || cap <= 0 || E % world)
    return fail(NF4MOE_EINVAL, "ep_pack: bad arguments (E must be a multiple of world, H of 8)");
  if (!aligned16(x) || !aligned16(send_rows)) return fail(NF4MOE_EALIGN, "ep_pack: 16-byte alignment required");
  DevInfo di;
  if (int rc = device_info(&di)) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t e = cudaMemsetAsync(pair_slot, 0xFF, (size_t)T * k * 4, st);   // -1: dropped / not dispatched
  if (e != cudaSuccess) return cuda_fail(e, "cudaMemsetAsync(pair_slot)");
  const long long slots = (long long)world * cap;
  const int grid = (int)std::min<long long>((slots + 7) / 8, (long long)di.sms * 8);
  ep_pack_kernel<<<grid, 256, 0, st>>>((const __half*)x, ids, expert_offsets, sorted_pairs, k, H, world, E / world, cap,
                                      (__half*)send_rows, send_ids, pair_slot, status);
  CHECK_LAUNCH("ep_pack_kernel");
  return 0;
}

int nf4moe_combine_indexed(const void* y, const float* weights, const int32_t* pair_slot, int T, int k, int H, void* out,
                           void* stream) {
  return nf4moe_combine_indexed_bias(y, weights, pair_slot, nullptr, nullptr, 0, T, k, H, out, stream);
}

int nf4moe_combine_indexed_bias(const void* y, const float* weights, const int32_t* pair_slot, const int32_t* ids,
                                const void* w2_bias, int E, int T, int k, int H, void* out, void* stream) {
  if (!y || !weights || !pair_slot || !out || T < 0 || k <= 0 || H <= 0 || H % 8)
    return fail(NF4MOE_EINVAL, "combine_indexed: bad arguments");
  if (w2_bias && (!ids || E <= 0 || !aligned16(w2_bias)))
    return fail(NF4MOE_EINVAL, "combine_indexed: w2_bias needs the GLOBAL topk ids, E and 16-byte alignment");
  if (!aligned16(y) || !aligned16(out)) return fail(NF4MOE_EALIGN, "combine_indexed: 16-byte alignment required");
  if (T == 0) return 0;
  DevInfo di;
  if (int rc = device_info(&di)) return rc;
  const long long total = (long long)T * (H / 8);
  const int grid = (int)std::min<long long>((total + 255) / 256, (long long)di.sms * 8);
  combine_indexed_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const __half*)y, weights, pair_slot, ids, (const __half*)w2_bias, E, T,
                                                                 k, H, (__half*)out);
  CHECK_LAUNCH("combine_indexed_kernel");
  return 0;
}

// ---- peer-memory exchange -----------------------------------------------------------------------------------------------
static int fill_peers(EpPeers* peers, void* const* rows, void* const* ids, void* const* flags, int world, bool want_ids, const char* name) {
  if (world <= 0 || world > kEpMaxWorld) return fail(NF4MOE_EINVAL, "%s: world must be in 1..%d", name, kEpMaxWorld);
  if (!rows || !flags || (want_ids && !ids)) return fail(NF4MOE_EINVAL, "%s: null peer pointer table", name);
  for (int p = 0; p < kEpMaxWorld; ++p) {
    peers->rows[p] = p < world ? (__half*)rows[p] : nullptr;
    peers->ids[p] = (p < world && want_ids) ? (int32_t*)ids[p] : nullptr;
    peers->flags[p] = p < world ? (unsigned long long*)flags[p] : nullptr;
    if (p < world && (!rows[p] || !flags[p] || (want_ids && !ids[p]) || !aligned16(rows[p])))
      return fail(NF4MOE_EINVAL, "%s: peer %d has a null or misaligned buffer", name, p);
  }
  return 0;
}

int nf4moe_ep_dispatch_p2p(const void* x, const int32_t* ids, const int32_t* expert_offsets, const int32_t* sorted_pairs, int T, int k,
                           int E, int H, int world, int rank, int cap, void* const* peer_recv_rows, void* const* peer_recv_ids,
                           void* const* peer_flags, int32_t* pair_slot, uint32_t* ctr, int32_t* status, const void* wait_flags,
                           uint32_t* wait_ctr, void* stream) {
  if ((wait_flags != nullptr) != (wait_ctr != nullptr)) return fail(NF4MOE_EINVAL, "ep_dispatch_p2p: give wait_flags and wait_ctr together");
  if (!x || !ids || !expert_offsets || !sorted_pairs || !pair_slot || !ctr || !status || T < 0 || k <= 0 || E <= 0 || H <= 0 || H % 8 ||
      cap <= 0 || rank < 0 || rank >= world || E % world)
    return fail(NF4MOE_EINVAL, "ep_dispatch_p2p: bad arguments (E must be a multiple of world, H of 8)");
  if (!aligned16(x)) return fail(NF4MOE_EALIGN, "ep_dispatch_p2p: x must be 16-byte aligned");
  EpPeers peers;
  if (int rc = fill_peers(&peers, peer_recv_rows, peer_recv_ids, peer_flags, world, true, "ep_dispatch_p2p")) return rc;
  DevInfo di;
  if (int rc = device_info(&di)) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const long long slots = (long long)world * cap;
  const int grid = (int)std::min<long long>((slots + 7) / 8, (long long)di.sms * 4);
  // (plain launches: these kernels read the router's and the bucket kernel's outputs through const __restrict__ pointers)
  ep_dispatch_kernel<<<grid, 256, 0, st>>>((const __half*)x, ids, expert_offsets, sorted_pairs, k, H, world, rank, E / world, cap, peers,
                                          pair_slot, (unsigned int*)ctr, status, T * k, (const unsigned long long*)wait_flags,
                                          (unsigned int*)wait_ctr);
  CHECK_LAUNCH("ep_dispatch_kernel");
  return 0;
}

int nf4moe_ep_return_p2p(const void* y_rows, const int32_t* recv_ids, int H, int world, int rank, int cap, void* const* peer_back_rows,
                         void* const* peer_flags, uint32_t* ctr, const void* wait_flags, uint32_t* wait_ctr, int32_t* status,
                         void* stream) {
  if ((wait_flags != nullptr) != (wait_ctr != nullptr)) return fail(NF4MOE_EINVAL, "ep_return_p2p: give wait_flags and wait_ctr together");
  if (!y_rows || !recv_ids || !ctr || H <= 0 || H % 8 || cap <= 0 || rank < 0 || rank >= world)
    return fail(NF4MOE_EINVAL, "ep_return_p2p: bad arguments");
  if (!aligned16(y_rows)) return fail(NF4MOE_EALIGN, "ep_return_p2p: y_rows must be 16-byte aligned");
  EpPeers peers;
  if (int rc = fill_peers(&peers, peer_back_rows, nullptr, peer_flags, world, false, "ep_return_p2p")) return rc;
  DevInfo di;
  if (int rc = device_info(&di)) return rc;
  const long long slots = (long long)world * cap;
  const int grid = (int)std::min<long long>((slots + 7) / 8, (long long)di.sms * 4);
  ep_return_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const __half*)y_rows, recv_ids, H, world, rank, cap, peers, (unsigned int*)ctr,
                                                          (const unsigned long long*)wait_flags, (unsigned int*)wait_ctr, status);
  CHECK_LAUNCH("ep_return_kernel");
  return 0;
}

int nf4moe_ep_wait(const void* flags, int world, uint32_t* ctr, int32_t* status, void* stream) {
  if (!flags || !ctr || !status || world <= 0 || world > kEpMaxWorld) return fail(NF4MOE_EINVAL, "ep_wait: bad arguments");
  ep_wait_kernel<<<1, 32, 0, (cudaStream_t)stream>>>((const unsigned long long*)flags, world, ctr, status);
  CHECK_LAUNCH("ep_wait_kernel");
  return 0;
}

}  // extern "C"

// ---- nf4moe_quantize_static: the quantiser for hosts that cannot allocate device memory --------------------------------------
// The quantiser's two passes talk through scratch (the per-expert fp64 absmax sums and the fp32 absmax of every block).  Here that
// scratch is a fixed array inside the library image instead of a caller buffer: each GPU that loads libnf4moe.so carries these
// NF4MOE_STATIC_SCRATCH_BYTES of HBM until the process exits, whether or not this entry point is ever called.
namespace {

__device__ __align__(256) unsigned char g_quant_scratch[NF4MOE_STATIC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
std::atomic<void*> g_quant_scratch_ptr[kMaxDevices];   // per device: where g_quant_scratch lives in THAT device's memory

// `g_quant_scratch` in host code is only the handle the runtime registered the array under, not somewhere a kernel can write: the
// device address has to be asked for, separately on every device (each holds its own copy), with that device current.
int quant_scratch_for_current_device(void** out) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return cuda_fail(e, "cudaGetDevice");
  if (dev < 0 || dev >= kMaxDevices) return fail(NF4MOE_EINVAL, "quantize_static: device ordinal %d is past %d", dev, kMaxDevices - 1);
  void* p = g_quant_scratch_ptr[dev].load(std::memory_order_acquire);
  if (!p) {
    e = cudaGetSymbolAddress(&p, g_quant_scratch);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(g_quant_scratch)");
    if (!p) return fail(NF4MOE_EINVAL, "quantize_static: the runtime returned no address for the scratch array");
    g_quant_scratch_ptr[dev].store(p, std::memory_order_release);   // two threads racing here store the same address
  }
  *out = p;
  return 0;
}

}  // namespace

extern "C" int nf4moe_quantize_static(const void* w_f16, int E, int rows, int cols, uint8_t* packed, uint8_t* absmax_u8, float* absmax2,
                                      const float* code2, float* offset, float* absmax_f32_out, void* stream) {
  if (E <= 0 || rows <= 0 || cols <= 0 || cols % kBlock) return fail(NF4MOE_EINVAL, "quantize_static: bad arguments");
  const unsigned long long nb = (unsigned long long)rows * (unsigned long long)(cols / kBlock);
  const unsigned long long need = align_up((size_t)E * 8, 256) + (unsigned long long)E * nb * 4ull;   // nf4moe_quantize's layout
  if (need > (unsigned long long)NF4MOE_STATIC_SCRATCH_BYTES)
    return fail(NF4MOE_EINVAL,
                "quantize_static: E=%d experts of %d x %d need %llu bytes of scratch, the built-in array holds %llu; quantise fewer "
                "experts per call or use nf4moe_quantize with a scratch buffer of your own",
                E, rows, cols, need, (unsigned long long)NF4MOE_STATIC_SCRATCH_BYTES);
  void* scratch = nullptr;
  if (int rc = quant_scratch_for_current_device(&scratch)) return rc;
  // nf4moe_quantize zeroes the sums on `stream` before its first pass and overwrites every absmax word it later reads, so what the
  // previous call left in the array is never seen
  return nf4moe_quantize(w_f16, E, rows, cols, packed, absmax_u8, absmax2, code2, offset, absmax_f32_out, scratch, stream);
}
