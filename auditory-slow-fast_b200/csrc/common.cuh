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

// Programmatic dependent launch.  Every hot kernel starts with pdl_trigger() (its successor may be scheduled as soon
// as all CTAs of this grid have started), runs its prologue (barrier init, TMEM allocation, descriptor prefetch,
// index arithmetic -- nothing that touches global memory another kernel produces or consumes) and then pdl_wait(),
// which returns once every prerequisite grid has completed and its writes are visible.  In a stream (or a captured
// graph) of such launches the prologue / launch latency of kernel i+1 overlaps the tail of kernel i.
// ASF_PDL=0 in the environment launches without the attribute (both instructions are then no-ops).
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
bool pdl_enabled();

template <typename... KArgs, typename... Args>
static inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                     Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// 16-byte vector path: bf16 rows of C % 8 == 0 channels whose pitch and base keep every 8-channel group aligned
static inline bool al16(const void* p) { return (((uintptr_t)p) & 15) == 0; }
static inline bool ld8(int64_t ld) { return (ld & 7) == 0; }

// elementwise_bf16.cu
int bn_apply_v8(const bf16* x, int64_t ld_x, const bf16* res, int64_t ld_res, bf16* y, int64_t ld_y, int64_t rows,
                int C, const float* scale, const float* shift, int relu, cudaStream_t st);
int bn_train_fwd_v8(const bf16* x, int64_t ld_x, const bf16* res, int64_t ld_res, bf16* y, int64_t ld_y, int64_t rows,
                    int C, const double* stats, double count, const float* gamma, const float* beta, float eps,
                    float momentum, float* running_mean, float* running_var, int64_t* num_batches_tracked, float* scale,
                    float* shift, float* save_mean, float* save_invstd, int relu, cudaStream_t st);
int bn_bwd_apply_v8(const bf16* dz, int64_t ld_dz, const bf16* x, int64_t ld_x, bf16* dx, int64_t ld_dx, int64_t rows,
                    int C, const float* gamma, const float* mean, const float* invstd, const double* red,
                    double inv_count, float* dgamma, float* dbeta, int accumulate, cudaStream_t st);
template <bool kBwd>
int channel_reduce_v8(const bf16* a, int64_t ld_a, const bf16* x, int64_t ld_x, int64_t rows, int C, const float* mean,
                      const float* invstd, double* out, float* ws, unsigned* ticket, cudaStream_t st);
int relu_bwd_v8(const bf16* a, int64_t ld_a, const bf16* b, int64_t ld_b, const bf16* mask, int64_t ld_mask, bf16* y,
                int64_t ld_y, int64_t rows, int C, cudaStream_t st);
int avgpool_bwd_v8(const float* dfeat, int64_t ld_feat, int c_off, const bf16* mask, bf16* dx, int B, int P, int C,
                   cudaStream_t st);
int bn_relu_maxpool_fwd_v8(const bf16* x, bf16* y, int64_t ld_y, int B, int T, int F, int C, const float* scale,
                           const float* shift, uint32_t* argmax, cudaStream_t st);
int bn_relu_maxpool_bwd_v8(const bf16* x, const bf16* dy, int64_t ld_dy, bf16* dz, int B, int T, int F, int C,
                           const float* scale, const float* shift, const uint32_t* argmax, cudaStream_t st);
int stem_conv_fwd_bf16(const float* in, const float* w, bf16* out, const asf_conv_geom_t* g, cudaStream_t st);
int stem_wgrad_v9_blocks(const asf_conv_geom_t* g);
int stem_wgrad_v9(const float* in, const bf16* dout, float* dwp, const asf_conv_geom_t* g, cudaStream_t st);
int stem_wgrad_v10_rows(const asf_conv_geom_t* g);
int stem_wgrad_v10(const float* in, const bf16* dout, float* dwp, float* dw, int accumulate, unsigned* ticket,
                   const asf_conv_geom_t* g, cudaStream_t st);

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
