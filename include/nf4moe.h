/*
 * nf4moe.h -- C ABI of libnf4moe.so: the bitsandbytes-NF4 fused-MoE expert path on B200 (sm_100a).
 *
 * This is the drop-in boundary.  Each entry point names the reference interface it replaces
 * (P/ = toddllm/gpt-oss-vllm patches/, V/ = vLLM 0.22 site-packages/vllm/):
 *
 *   nf4moe_dequant            bitsandbytes.functional.dequantize_4bit as called from
 *                             V/model_executor/layers/quantization/bitsandbytes.py:588-603 and restated on CPU by
 *                             P/streaming_unsloth_bnb.py:109-230 (dequant_unsloth_expert_cpu)
 *   nf4moe_dequant_absmax     the double-quant de-nesting of V/model_executor/model_loader/bitsandbytes_loader.py:585-617
 *                             and P/unsloth_direct_loader.py:21-48 (dequantize_absmax)
 *   nf4moe_quantize           bnb quantize_4bit(blocksize=64, compress_statistics=True, quant_type='nf4') as invoked by
 *                             P/bnb_moe_4bit.py:86-91 and P/streaming_unsloth_loader.py:48-71
 *   nf4moe_route_topk         P/unsloth_to_vllm_converter.py:291-292 == V/.../fused_moe/router/fused_topk_router.py:69-102
 *   nf4moe_bucket             the per-(token,k) expert dispatch loop of P/unsloth_to_vllm_converter.py:298-305
 *                             (== moe_align_block_size inside V/.../fused_moe/fused_moe.py fused_experts)
 *   nf4moe_gemv_w13_swiglu    w13_experts[e](x) + split + activation, P/unsloth_to_vllm_converter.py:309-315,
 *                             activation per V/model_executor/layers/activation.py:389-397
 *   nf4moe_gemv_w2            w2_experts[e](intermediate), P/unsloth_to_vllm_converter.py:318
 *   nf4moe_gemm_w13_swiglu /
 *   nf4moe_gemm_w2            same two products at prefill sizes (vLLM: Triton fused_experts, V/.../fused_moe.py:1587-1617)
 *   nf4moe_combine            token_output += weight * expert_output, P/unsloth_to_vllm_converter.py:321-323
 *   nf4moe_moe_forward        UnslothMoEWrapper.forward (P/unsloth_to_vllm_converter.py:282-325) ==
 *                             BitsAndBytesMoEMethod.apply (V/.../quantization/bitsandbytes.py:478-506)
 *
 * Conventions
 *   - every pointer is a DEVICE pointer owned by the caller; the library allocates nothing and keeps no device state -- with one
 *     stated exception: the fixed NF4MOE_STATIC_SCRATCH_BYTES array behind nf4moe_quantize_static (see its comment);
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*), does no host<->device sync and is
 *     CUDA-graph capturable;
 *   - return value: 0 on success, a negative NF4MOE_E* code on a usage error, or a positive cudaError_t passed through;
 *     nf4moe_last_error() returns a thread-local description of the last failure;
 *   - activations are fp16 (the reference serves with --dtype float16, P/gpt_oss_server.py:41), accumulation is fp32;
 *   - packed weights: uint8, element 2j in the HIGH nibble, element 2j+1 in the LOW nibble (bitsandbytes order),
 *     row-major [rows, cols] per expert, experts concatenated: exactly vLLM's `w13_weight[E, rows*cols/2, 1]`.
 */
#ifndef NF4MOE_H_
#define NF4MOE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NF4MOE_VERSION 101 /* 0.1.1.  ABI notes: since 101 `meta` (nf4moe_bucket and every entry point that takes it) is 16 int32, not 4 --
                              [4] is the status word kernels OR into, [5] the token-tile size of the list; a caller that allocated 4
                              words for version 100's first drafts must allocate 16.  nf4moe_quantize_static is new in 101. */

#define NF4MOE_OK 0
#define NF4MOE_EINVAL (-1)      /* bad shape / null pointer / unsupported size */
#define NF4MOE_EALIGN (-2)      /* pointer not 16-byte aligned */
#define NF4MOE_EWORKSPACE (-3)  /* workspace too small */
#define NF4MOE_EARCH (-4)       /* not an sm_100 device */

/* activation selector for the w13 epilogue */
#define NF4MOE_ACT_SWIGLU_OAI 0 /* interleaved, clamped, alpha=1.702, limit=7 (gpt-oss; the product default) */
#define NF4MOE_ACT_CHUNK_SILU 1 /* gate=[:I], up=[I:], silu(gate)*up -- the reference wrapper's own variant */

/* regime selector for nf4moe_moe_forward */
#define NF4MOE_REGIME_AUTO 0
#define NF4MOE_REGIME_GEMV 1  /* decode: HBM-bound NF4 streaming kernel (mma.sync) */
#define NF4MOE_REGIME_GEMM 2  /* tcgen05 kernels with the NF4MOE_GEMM_EXACT operand at every T: the decode-regime kernel while
                                 token tiles stay <= 32 rows, the prefill grouped GEMM above */
#define NF4MOE_REGIME_GEMM_FAST 3 /* the same two kernels with the NF4MOE_GEMM_FAST operand */

/* how the tcgen05 kernels turn an NF4 block into the fp16 MMA operand */
#define NF4MOE_GEMM_EXACT 0 /* fp16_rn(fp32(code) * fp32(absmax)): bit-identical to bnb dequantize_4bit */
#define NF4MOE_GEMM_FAST 1  /* fp16(code) * fp16(absmax) in half2 arithmetic: ~2^-11 relative per weight, inside the MoE
                               output tolerance (1e-2 abs / 2e-3 rel), half the shared-memory traffic of EXACT */

/* One fused NF4 expert matrix set: E experts, each [rows, cols] row-major, cols % 64 == 0.
 * Quantisation blocks (64) run along the flat row-major index, so block b of expert e covers
 * row b / (cols/64), columns (b % (cols/64))*64 .. +63.
 * Exactly one absmax form is given:
 *   flat   : absmax_f32 != NULL, fp32 [E, rows*cols/64]       (what vLLM keeps after _dequantize_dq)
 *   nested : absmax_f32 == NULL; absmax_u8 [E, nb], absmax2 fp32 [E, ceil(nb/256)], code2 fp32 [256] (+ e*code2_stride),
 *            offset fp32 [E]:  absmax[b] = code2[absmax_u8[b]] * absmax2[b/256] + offset   (two roundings, no FMA)
 *            -- the on-disk Unsloth/bnb key set {.absmax, .nested_absmax, .nested_quant_map} of P/unsloth_bnb_loader.py:202-222. */
typedef struct nf4moe_qweight {
  const uint8_t* packed;
  const float* absmax_f32;
  const uint8_t* absmax_u8;
  const float* absmax2;
  const float* code2;
  const float* offset;
  int64_t code2_stride; /* 0: one table shared by all experts; 256: one per expert */
  int32_t rows;
  int32_t cols;
} nf4moe_qweight;

int nf4moe_version(void);
/* The regime NF4MOE_REGIME_AUTO resolves to for T tokens, top-k k, E experts (pure function of its arguments; replaces the
 * reference's absence of any regime choice: its forward is a per-(token, slot) Python loop, P/unsloth_to_vllm_converter.py:298-323). */
int nf4moe_auto_regime(int T, int k, int E);
/* The same for a layer whose w13 AND w2 carry the nested absmax (nested_absmax != 0) and the fused gpt-oss activation: the decode-regime
 * kernel then takes 64-row token tiles and serves up to a mean of 80 routed rows per expert.  nf4moe_moe_forward / nf4moe_expert_rows
 * with NF4MOE_REGIME_AUTO resolve through this function. */
int nf4moe_auto_regime_for(int T, int k, int E, int nested_absmax);
const char* nf4moe_last_error(void);

/* Scratch the fused forward needs for T tokens (bytes).  Layout is private to the library. */
size_t nf4moe_workspace_bytes(int T, int E, int H, int I, int k);

/* Byte offset, inside a workspace of that geometry, of an int32 STATUS word.  The kernels only ever OR bits into it and never
 * clear it, so the caller zeroes it once (e.g. allocates the workspace zero-filled) and may read it whenever it synchronises
 * anyway: a non-zero value means a kernel refused to run and the output is not valid.  Bits: NF4MOE_STATUS_*.  The host-side
 * checks of every entry point already reject the shapes that could raise a bit; the word exists so that "a kernel did nothing"
 * can never pass silently (the reference has no such failure mode to mirror: its per-pair Python loop raises,
 * P/unsloth_to_vllm_converter.py:298-323).  For the stand-alone nf4moe_gemm_* entry points the word is meta[4]. */
size_t nf4moe_workspace_status_offset(int T, int E, int H, int I, int k);
#define NF4MOE_STATUS_EP_OVERFLOW 1  /* nf4moe_ep_pack: a peer got more rows than its capacity (written to ITS status argument) */
#define NF4MOE_STATUS_DEC_NOFIT 2    /* decode tcgen05 kernel: shared-memory plan did not fit */
#define NF4MOE_STATUS_GEMM_NOFIT 4   /* prefill tcgen05 kernel: shared-memory plan did not fit */
#define NF4MOE_STATUS_DEP_TIMEOUT 16 /* merged w13+w2 decode launch: a w2 tile waited > 2^26 polls for its w13 row tiles (the
                                        grid was not fully resident); EP_TIMEOUT = 8 is defined with nf4moe_ep_wait below.
                                        The two TIMEOUT conditions are fatal: the kernel sets the bit and then TRAPS, so the
                                        caller's next synchronisation returns a launch failure and the context is lost -- a layer
                                        never continues on rows that did not arrive (what NCCL's abort does for a dead peer) */

/* ---- NF4 (de)quantisation ------------------------------------------------------------------------------------ */

/* out_f16[e, i] = fp16_rn( fp32(NF4[q_i]) * absmax[i/64] ) for experts [e0, e0+ne): bit-exact bnb semantics. */
int nf4moe_dequant(const nf4moe_qweight* w, int e0, int ne, void* out_f16, void* stream);

/* absmax_f32[e, b] = code2[absmax_u8[e,b]] * absmax2[e, b/256] + offset[e] for all E experts (nested form only). */
int nf4moe_dequant_absmax(const nf4moe_qweight* w, int E, float* absmax_f32_out, void* stream);

/* fp16 [E, rows, cols] -> packed NF4 + double-quantised absmax (also writes the de-nested fp32 absmax if non-NULL).
 * `code2` (256 sorted fp32) is an INPUT.  Two passes on `stream`; `scratch` needs E*(rows*cols/64)*4 + E*8 + 512 bytes. */
int nf4moe_quantize(const void* w_f16, int E, int rows, int cols, uint8_t* packed, uint8_t* absmax_u8, float* absmax2,
                    const float* code2, float* offset, float* absmax_f32_out, void* scratch, void* stream);

/* nf4moe_quantize for a host that must not, or cannot, allocate device memory.  It has no counterpart in the reference (which
 * quantises through bitsandbytes and lets torch's allocator find the temporaries); it exists for two kinds of caller of THIS library:
 *   - a weight loader running inside a serving process whose memory is already spoken for: once an engine such as vLLM has sized
 *     its KV-cache pool to the configured share of HBM, a cudaMalloc (or a torch.empty that misses torch's cache) for the
 *     quantiser's temporaries can fail with out-of-memory although the layer itself fits;
 *   - a C / ctypes / cgo host that binds this ABI without an allocator of its own, or that may not call cudaMalloc / cudaFree at
 *     that point because both can synchronise the whole device and stall the streams that are serving requests.
 * Same arguments, arithmetic and results as nf4moe_quantize, minus `scratch`: the two passes share a 256-byte-aligned array of
 * NF4MOE_STATIC_SCRATCH_BYTES that is part of the library image.  What that costs and implies:
 *   - every GPU on which the process loads libnf4moe.so gives up that much memory until the process exits, used or not;
 *   - limit: align256(E*8) + E*(rows*cols/64)*4 <= NF4MOE_STATIC_SCRATCH_BYTES, i.e. up to 4 gate/up matrices [5760, 2880] or 8 down
 *     matrices [2880, 2880] of gpt-oss per call; anything larger returns NF4MOE_EINVAL with the two sizes in nf4moe_last_error()
 *     and launches nothing -- quantise fewer experts per call (the results per expert do not depend on E);
 *   - the call itself allocates and frees nothing (no cudaMalloc / cudaFree, no library routine that would);
 *   - the first call on each device looks up where the array lives on THAT device (cudaGetSymbolAddress, with the caller's current
 *     device; the answer is kept per device).  Make that first call while no stream capture is in progress;
 *   - there is one array per device, shared by every caller and left holding the last call's temporaries.  The entry point clears
 *     the part that has to start at zero on `stream` itself, so back-to-back calls on one stream are fine; two calls that can run
 *     on the GPU at the same time (different streams, different host threads) overwrite each other's temporaries and both results
 *     are garbage -- serialise them, or give the concurrent ones nf4moe_quantize and their own buffers. */
#define NF4MOE_STATIC_SCRATCH_BYTES 4194304 /* 4 MiB */
int nf4moe_quantize_static(const void* w_f16, int E, int rows, int cols, uint8_t* packed, uint8_t* absmax_u8, float* absmax2,
                           const float* code2, float* offset, float* absmax_f32_out, void* stream);

/* ---- routing --------------------------------------------------------------------------------------------------- */

/* logits [T,E] (dtype: 0 = fp16, 1 = fp32, 2 = bf16) -> ids int32 [T,k] (descending logit, ties -> lowest index),
 * weights fp32 [T,k] = softmax over the k selected logits (== softmax -> top-k -> renormalise). 1 <= E <= 256. */
int nf4moe_route_topk(const void* logits, int logits_dtype, int T, int E, int k, int32_t* ids, float* weights,
                      void* stream);

/* fp16 <-> bf16, elementwise, n elements (to_bf16 != 0: src fp16 -> dst bf16, round to nearest even; else src bf16 -> dst fp16, exact
 * inside the fp16 range, saturating at +-65504).  How a bf16 caller is served (the reference's config lists bf16 beside fp16,
 * P/bnb_moe_4bit.py:35-37): activations are widened exactly, the layer runs with fp16 operands and fp32 accumulation, the output
 * is rounded to bf16 once. */
int nf4moe_cast_f16_bf16(const void* src, void* dst, long long n, int to_bf16, void* stream);

/* Stable counting sort of the T*k (token,slot) pairs by expert.  ids outside [0,E) are dropped (expert-parallel
 * callers mask non-local experts with -1).  Outputs (all int32):
 *   expert_offsets[E+1]  exclusive prefix of per-expert pair counts
 *   sorted_pairs[T*k]    pair index t*k+slot, grouped by expert, ascending within an expert
 *   active_experts[E]    experts with >=1 pair, ascending;  meta[0] = how many, meta[1] = total kept pairs,
 *                        meta[2] = number of token tiles, meta[3] = largest per-expert count; meta has 16 entries:
 *                        meta[4] is the status word the nf4moe_gemm_* kernels OR NF4MOE_STATUS_* bits into (never written
 *                        here: zero it once), meta[5] = the tile_rows this list was built with (0 without a tile list; no
 *                        tile has more rows), the rest is reserved
 *   tile_list[4*(T*k/16+E+2)]  (expert, first sorted row, rows, 0) per token tile of the grouped GEMM: an expert with c
 *                        rows gets ceil(c/tile_rows) near-equal tiles, each a multiple of 16 rows except the last.
 *                        tile_rows: multiple of 16 in 16..256, or 0 = what nf4moe_gemm_* expect for this (T, k, E).
 *                        tile_list may be NULL (decode regime). */
int nf4moe_bucket(const int32_t* ids, int T, int k, int E, int32_t* expert_offsets, int32_t* sorted_pairs,
                  int32_t* active_experts, int32_t* meta, int32_t* tile_list, int tile_rows, void* stream);

/* x_sorted[s, :] = x[sorted_pairs[s] / k, :]  (fp16, H % 8 == 0); rows s >= meta[1] are zero-filled. */
int nf4moe_permute_rows(const void* x, const int32_t* sorted_pairs, const int32_t* meta, int T, int k, int H,
                        void* x_sorted, void* stream);

/* out[t, :] = fp16( sum_slot weights[t,slot] * y[t*k+slot, :] ) in fp32, slot ascending; pairs whose id was dropped
 * contribute zero (ids may be NULL when none are).  y is fp16 [T*k, H]. */
int nf4moe_combine(const void* y, const float* weights, const int32_t* ids, int E, int T, int k, int H, void* out,
                   void* stream);

/* ---- expert products --------------------------------------------------------------------------------------------- */

/* Decode regime.  h_sorted[s, j] = act( x[sorted_pairs[s]/k, :] . W13[e]^T )  for every sorted row s of every active e.
 * w13->rows = 2*I, w13->cols = H (H % 64 == 0, H <= 2880, rows % 16 == 0). h_sorted is fp16 [T*k, I]. */
int nf4moe_gemv_w13_swiglu(const void* x, const nf4moe_qweight* w13, const int32_t* expert_offsets,
                           const int32_t* sorted_pairs, const int32_t* active_experts, const int32_t* meta, int T, int k,
                           int E, int act, void* h_sorted, void* stream);

/* y[sorted_pairs[s], n] = h_sorted[s, :] . W2[e]^T ; w2->rows = H, w2->cols = I.  y is fp16 [T*k, H]. */
int nf4moe_gemv_w2(const void* h_sorted, const nf4moe_qweight* w2, const int32_t* expert_offsets,
                   const int32_t* sorted_pairs, const int32_t* active_experts, const int32_t* meta, int T, int k, int E,
                   void* y, void* stream);

/* Prefill regime: the same two products as tcgen05 grouped GEMMs.  Weight rows are the MMA M dimension (128 per tile),
 * the token tiles of `tile_list` (built by nf4moe_bucket with tile_rows = 0) the N dimension; NF4 blocks are dequantised
 * (mode NF4MOE_GEMM_EXACT: fp32 multiply, one rounding to fp16, bit-exact; NF4MOE_GEMM_FAST: half2 arithmetic) straight
 * into TENSOR MEMORY as the A operand, activations arrive by TMA, accumulators live in TMEM.  x_sorted / h_sorted: fp16 [>= T*k rows, cols], rows grouped by expert. */
int nf4moe_gemm_w13_swiglu(const void* x_sorted, const nf4moe_qweight* w13, const int32_t* expert_offsets,
                           const int32_t* meta, const int32_t* tile_list, int T, int k, int E, int act, int mode,
                           void* h_sorted, void* stream);
int nf4moe_gemm_w2(const void* h_sorted, const nf4moe_qweight* w2, const int32_t* expert_offsets,
                   const int32_t* sorted_pairs, const int32_t* meta, const int32_t* tile_list, int T, int k, int E,
                   int mode, void* y, void* stream);

/* Decode regime (token tiles of at most 32 rows, i.e. a tile list built with tile_rows <= 32) with the fused gpt-oss activation:
 * both products in ONE persistent launch -- h_sorted[T*k, I] = SwiGLU(x_sorted W13^T (+ w13_bias)), y = h_sorted W2^T scattered to
 * (token, slot) rows.  Replaces the same two bnb Linear4bit calls per routed row as nf4moe_gemm_w13_swiglu + nf4moe_gemm_w2
 * (P/unsloth_to_vllm_converter.py:309-318); nf4moe_moe_forward / nf4moe_expert_rows use it whenever it applies.
 * `tile_list` is written to: the kernel counts the finished w13 row tiles of token tile t in tile_list[4 t + 3], which
 * nf4moe_bucket zeroes -- the list must come fresh from nf4moe_bucket for EVERY call.  Status bits as for the other tcgen05
 * entry points (meta[4]); NF4MOE_STATUS_DEP_TIMEOUT if the grid was not fully resident.  A tile list built for larger tiles (meta[5] >
 * 32, e.g. nf4moe_bucket(tile_rows = 0) at prefill sizes) is refused as a whole: NF4MOE_STATUS_DEC_NOFIT is raised, h and y stay untouched. */
int nf4moe_gemm_w13_w2(const void* x_sorted, const nf4moe_qweight* w13, const void* w13_bias, void* h_sorted,
                       const nf4moe_qweight* w2, void* y, const int32_t* sorted_pairs, const int32_t* meta, int32_t* tile_list,
                       int T, int k, int E, int act, int mode, void* stream);

/* Process-wide switch for GPUs this process does not own outright.  The merged launch above assigns its work items to CTAs
 * statically and lets w2 items spin until the w13 items of their token tile are counted, so it makes progress only if all of its
 * CTAs (one per SM) get to run at the same time.  On an exclusively owned GPU they do.  Under an MPS active-thread percentage, a
 * green context / SM partition, or next to a long-running kernel of another stream that itself waits for this one, some CTAs may
 * never be scheduled while the resident ones spin: the kernel then gives up after 2^26 polls (seconds) and traps
 * (NF4MOE_STATUS_DEP_TIMEOUT).  enabled = 0 makes nf4moe_moe_forward* / nf4moe_expert_rows* issue w13 and w2 as two launches of the
 * same kernel (no cross-CTA wait, any residency; bit-identical results; about 9 % slower at 64 tokens on a B200); enabled != 0
 * (the default) restores the single launch.  Returns the previous setting.  Takes effect for calls made afterwards -- set it
 * before capturing CUDA graphs.  No counterpart in the reference (its forward is a Python loop, P/unsloth_to_vllm_converter.py:298-323). */
int nf4moe_set_merged_decode(int enabled);

/* ---- the whole layer ---------------------------------------------------------------------------------------------- */

/* out[T,H] = MoE(x[T,H]).  Routing is computed from `router_logits` when non-NULL (the north-star / legacy
 * `apply(x, router_logits, top_k)` and vLLM `apply_monolithic` shape), else taken from topk_ids/topk_weights (vLLM 0.22
 * `apply`).  `topk_ids_out`/`topk_weights_out` (optional) receive the routing actually used. */
int nf4moe_moe_forward(const void* x, const void* router_logits, int logits_dtype, const int32_t* topk_ids,
                       const float* topk_weights, const nf4moe_qweight* w13, const nf4moe_qweight* w2, void* out, int T,
                       int E, int H, int I, int k, int act, int regime, void* workspace, size_t workspace_bytes,
                       int32_t* topk_ids_out, float* topk_weights_out, void* stream);

/* The same layer with the expert biases of gpt-oss (V/model_executor/models/gpt_oss.py:199; SURVEY 8(f) row 3 -- absent from the
 * reference wrapper and from vLLM's bnb method): w13_bias fp16 [E, 2I] is added to the up-projection before the activation,
 * w2_bias fp16 [E, H] to the down-projection before the routing weight.  Either may be NULL. */
int nf4moe_moe_forward_bias(const void* x, const void* router_logits, int logits_dtype, const int32_t* topk_ids,
                            const float* topk_weights, const nf4moe_qweight* w13, const nf4moe_qweight* w2, const void* w13_bias,
                            const void* w2_bias, void* out, int T, int E, int H, int I, int k, int act, int regime, void* workspace,
                            size_t workspace_bytes, int32_t* topk_ids_out, float* topk_weights_out, void* stream);

/* ---- expert parallelism (one process per GPU; experts [r*E/world, (r+1)*E/world) live on rank r) ------------------------
 * The reference is single-GPU (P/gpt_oss_server.py:25 pins CUDA_VISIBLE_DEVICES=0); these entry points are what the
 * per-(token,k) expert loop of P/unsloth_to_vllm_converter.py:298-323 becomes when the experts of that loop are sharded.
 * The all-to-all itself is issued by the host (NCCL via torch.distributed) between these calls. */

/* Dispatch side.  `expert_offsets` / `sorted_pairs` come from nf4moe_bucket over the GLOBAL expert ids.  Writes, for peer
 * p and its j-th routed pair:  send_rows[p*cap + j, :] = x[pair/k, :] (fp16 [world*cap, H]),  send_ids[p*cap + j] = expert id
 * local to p (-1 in unused slots),  pair_slot[pair] = p*cap + j (-1 for pairs not dispatched).  status[0] |= 1 if a peer
 * got more than `cap` pairs (the excess is dropped; cap = T*k can never overflow). */
int nf4moe_ep_pack(const void* x, const int32_t* ids, const int32_t* expert_offsets, const int32_t* sorted_pairs, int T, int k,
                   int E, int H, int world, int cap, void* send_rows, int32_t* send_ids, int32_t* pair_slot,
                   int32_t* status, void* stream);

/* Expert side.  y_rows[r, :] = W2[e] * act(W13[e] * rows[r, :]) with e = row_expert[r] in [0, E) (rows with any other id are
 * skipped and their y is left untouched).  E, w13, w2 are the LOCAL experts.  Workspace: nf4moe_workspace_bytes(n_rows, E,
 * H, I, 1). */
int nf4moe_expert_rows(const void* rows, const int32_t* row_expert, const nf4moe_qweight* w13, const nf4moe_qweight* w2,
                       void* y_rows, int n_rows, int E, int H, int I, int act, int regime, void* workspace,
                       size_t workspace_bytes, void* stream);

/* The same with gpt-oss's up-projection bias (w13_bias fp16 [E, 2I], LOCAL expert index; NULL = none), added before the
 * activation exactly as nf4moe_moe_forward_bias does.  rows_hint (0 = n_rows): how many of the n_rows slots are EXPECTED to hold a
 * row -- a fixed-capacity receive buffer is mostly empty, and the regime / token-tile size should follow the rows that exist
 * (round 1 chose them by capacity: a worst-case buffer ran the prefill pipeline over 2048 nearly empty slots). */
int nf4moe_expert_rows_bias(const void* rows, const int32_t* row_expert, const nf4moe_qweight* w13, const nf4moe_qweight* w2,
                            const void* w13_bias, void* y_rows, int n_rows, int rows_hint, int E, int H, int I, int act, int regime,
                            void* workspace, size_t workspace_bytes, void* stream);

/* Combine side.  out[t, :] = fp16( sum_slot weights[t,slot] * y[pair_slot[t*k+slot], :] ) -- nf4moe_combine's arithmetic and
 * order, reading each pair's row where the return all-to-all left it. */
int nf4moe_combine_indexed(const void* y, const float* weights, const int32_t* pair_slot, int T, int k, int H, void* out,
                           void* stream);

/* The same with the down-projection bias: w2_bias is the FULL table fp16 [E, H] (every rank keeps a copy: 184 KB for
 * gpt-oss-20b), indexed by the home rank's GLOBAL topk ids [T, k]; it is added at home in fp32 before the routing weight, so
 * the expert-parallel result equals nf4moe_moe_forward_bias bit for bit.  w2_bias == NULL: ids / E are ignored. */
int nf4moe_combine_indexed_bias(const void* y, const float* weights, const int32_t* pair_slot, const int32_t* ids,
                                const void* w2_bias, int E, int T, int k, int H, void* out, void* stream);

/* ---- the same exchange over PEER MEMORY (NVLink / NVSwitch), no collective library on the data path ------------------------
 * Each rank owns, in memory its peers can address, recv_rows fp16 [world, cap, H], recv_ids int32 [world, cap], back_rows fp16
 * [world, cap, H] and flags uint64 [2, world] (zeroed once), plus a private counter block ctr uint32 (zeroed once; 2 words per
 * entry point below).  The tables peer_*[p] are HOST arrays of `world` DEVICE pointers into rank p's copies.
 *   nf4moe_ep_dispatch_p2p   home rank: stores row (p, j) of its pair list straight into owner p's recv_rows[rank, j] and the local
 *                            expert id into recv_ids[rank, j] (-1 for unused slots), pair_slot[pair] = p*cap + j, then publishes a
 *                            new epoch to flags[0][rank] of every peer.  Only rows that exist cross the link, so cap = T*k (which
 *                            can never overflow) costs nothing: the exchange is drop-free.
 *   nf4moe_ep_wait           spins (system-scope acquire) until all `world` entries of one flag row carry this rank's next epoch;
 *                            if a peer never arrives (2^27 polls, roughly half a minute): status |= NF4MOE_STATUS_EP_TIMEOUT and the
 *                            kernel traps -- the next synchronisation fails; the same holds for the waits inside dispatch / return.
 *   nf4moe_ep_return_p2p     owner rank: stores y_rows[q*cap + j] (where recv_ids >= 0) into home rank q's back_rows[rank, j] and
 *                            publishes to flags[1][rank] of every peer; nf4moe_combine_indexed[_bias] then reads back_rows.
 * wait_flags / wait_ctr (both or neither; one process per GPU only): the dispatch / return kernel itself then waits, after publishing,
 * for THIS rank's flag row (flags[0] resp. flags[1]) to reach its next epoch -- the call returns a stream on which the peers' rows
 * have landed and the separate nf4moe_ep_wait launch is not needed.
 * Epochs live in device memory and advance once per launch, so a layer built from these calls replays from a CUDA graph.  Every
 * rank must issue the same sequence of calls (SPMD).  Replaces nothing in the reference (single GPU, P/gpt_oss_server.py:25). */
#define NF4MOE_STATUS_EP_TIMEOUT 8
int nf4moe_ep_dispatch_p2p(const void* x, const int32_t* ids, const int32_t* expert_offsets, const int32_t* sorted_pairs, int T, int k,
                           int E, int H, int world, int rank, int cap, void* const* peer_recv_rows, void* const* peer_recv_ids,
                           void* const* peer_flags, int32_t* pair_slot, uint32_t* ctr, int32_t* status, const void* wait_flags,
                           uint32_t* wait_ctr, void* stream);
int nf4moe_ep_return_p2p(const void* y_rows, const int32_t* recv_ids, int H, int world, int rank, int cap, void* const* peer_back_rows,
                         void* const* peer_flags, uint32_t* ctr, const void* wait_flags, uint32_t* wait_ctr, int32_t* status,
                         void* stream);
int nf4moe_ep_wait(const void* flags, int world, uint32_t* ctr, int32_t* status, void* stream);

/* Number of kernels the last nf4moe_moe_forward / nf4moe_expert_rows on this thread launched (bench.py's gpu_launches). */
int nf4moe_last_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* NF4MOE_H_ */
