// Shared helpers for libgeopurify_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/geopurify_b200.h"

#define GP_SM_COUNT 148

void gp_set_error(const char* fmt, ...);
int gp_l2_normalize_rows(float* x, int64_t rows, int E, cudaStream_t st);  // student.cu
// logits.cu: fp32 SIMT logits + argmax on the rows flagged in `only` (fix-up pass of gp_logits_argmax_tc)
int gp_logits_fix_rows(const float* feat, int64_t stride, int D, int64_t rows, const float* text, int C,
                       float logit_scale, int32_t* label, uint8_t* zero, const uint8_t* only, void* workspace,
                       size_t workspace_bytes, cudaStream_t st);
// capi.cu: optional event timing around the dominant kernels (gp_timing_enable); slot ids below
enum { GP_T_STUDENT_GEMM = 0, GP_T_STUDENT_REDUCE = 1, GP_T_POOL_APPLY = 2, GP_T_POOL_PLAN = 3 };
bool gp_timing_begin(int which, cudaStream_t st);
void gp_timing_end(int which, cudaStream_t st);

#define GP_CHECK_ARG(cond, ...)          \
  do {                                   \
    if (!(cond)) {                       \
      gp_set_error(__VA_ARGS__);         \
      return GP_ERR_ARG;                 \
    }                                    \
  } while (0)

#define GP_CUDA(expr)                                                                   \
  do {                                                                                  \
    cudaError_t e_ = (expr);                                                            \
    if (e_ != cudaSuccess) {                                                            \
      gp_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__); \
      return GP_ERR_CUDA;                                                               \
    }                                                                                   \
  } while (0)

#define GP_LAUNCH_CHECK()                                                               \
  do {                                                                                  \
    cudaError_t e_ = cudaGetLastError();                                                \
    if (e_ != cudaSuccess) {                                                            \
      gp_set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(e_), __FILE__, __LINE__); \
      return GP_ERR_CUDA;                                                               \
    }                                                                                   \
  } while (0)

static inline size_t gp_align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

// Bump allocator over a caller-provided workspace.
struct GpArena {
  char* base;
  size_t size, off;
  GpArena(void* p, size_t n) : base((char*)p), size(n), off(0) {}
  template <typename T>
  T* take(size_t count) {
    size_t bytes = gp_align_up(count * sizeof(T));
    T* r = (T*)(base + off);
    off += bytes;
    return r;
  }
  bool ok() const { return off <= size; }
};

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Read-only 128-bit load that does not allocate in L1 (streamed once).
__device__ __forceinline__ float4 ldg_stream_f4(const float4* p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p));
  return v;
}
__device__ __forceinline__ uint4 ldg_stream_u4(const uint4* p) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p));
  return v;
}
