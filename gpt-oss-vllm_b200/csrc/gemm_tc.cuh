// Prefill-regime grouped GEMM (tcgen05 / TMEM).  Placeholder until the kernel lands: the entry points fail loudly.
#pragma once
#include <cstdio>

#include "common.cuh"

namespace nf4moe {

constexpr int kGemmAutoThreshold = 64;  // tokens; above this nf4moe_moe_forward(AUTO) uses the tcgen05 path
inline bool gemm_available() { return false; }

inline int gemm_w13_entry(const void*, const nf4moe_qweight*, const int32_t*, const int32_t*, const int32_t*, int, int,
                          int, int, void*, void*, char* err, size_t err_len, int*) {
  snprintf(err, err_len, "gemm_w13: tcgen05 grouped GEMM not built in this version");
  return NF4MOE_EINVAL;
}
inline int gemm_w2_entry(const void*, const nf4moe_qweight*, const int32_t*, const int32_t*, const int32_t*,
                         const int32_t*, int, int, int, void*, void*, char* err, size_t err_len, int*) {
  snprintf(err, err_len, "gemm_w2: tcgen05 grouped GEMM not built in this version");
  return NF4MOE_EINVAL;
}

}  // namespace nf4moe
