// Shared host/device helpers for libasf_b200.so (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "asf_b200.h"

namespace asf {

void set_error(const char* fmt, ...);
void count_launch();  // every kernel launch of the library is counted (asf_launch_count)

#define ASF_CHECK_ARG(cond, ...)         \
  do {                                   \
    if (!(cond)) {                       \
      asf::set_error(__VA_ARGS__);       \
      return ASF_E_BADARG;               \
    }                                    \
  } while (0)

#define ASF_CUDA(call)                                                              \
  do {                                                                              \
    cudaError_t e__ = (call);                                                       \
    if (e__ != cudaSuccess) {                                                       \
      asf::set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
      return (int)e__;                                                              \
    }                                                                               \
  } while (0)

#define ASF_LAUNCH_CHECK()                                                          \
  do {                                                                              \
    asf::count_launch();                                                            \
    cudaError_t e__ = cudaGetLastError();                                           \
    if (e__ != cudaSuccess) {                                                       \
      asf::set_error("%s:%d launch -> %s", __FILE__, __LINE__, cudaGetErrorString(e__)); \
      return (int)e__;                                                              \
    }                                                                               \
  } while (0)

typedef __nv_bfloat16 bf16;

__device__ __forceinline__ float ldf(const float* p) { return *p; }
__device__ __forceinline__ float ldf(const bf16* p) { return __bfloat162float(*p); }
__device__ __forceinline__ void stf(float* p, float v) { *p = v; }
__device__ __forceinline__ void stf(bf16* p, float v) { *p = __float2bfloat16_rn(v); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

static inline int64_t cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }

// dispatch on a dtype code
#define ASF_DISPATCH_DTYPE(code, T, ...)                       \
  do {                                                         \
    if ((code) == ASF_F32) {                                   \
      typedef float T;                                         \
      __VA_ARGS__;                                             \
    } else if ((code) == ASF_BF16) {                           \
      typedef asf::bf16 T;                                     \
      __VA_ARGS__;                                             \
    } else {                                                   \
      asf::set_error("bad dtype code %d", (int)(code));        \
      return ASF_E_BADARG;                                     \
    }                                                          \
  } while (0)

}  // namespace asf
