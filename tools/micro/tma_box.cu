// Microbenchmark: cp.async.bulk.tensor.2d box-load rate vs. channel extent / box width (is OOB zero-fill slow?).
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tma_box tma_box.cu -lcuda && ./tma_box
#include <cstdint>
#include <cstdio>
#include <cuda.h>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(32, 1) k(const __grid_constant__ CUtensorMap tm, int box_bytes, int rows_total,
                                           int iters, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar[4];
  if (threadIdx.x == 0) {
    for (int i = 0; i < 4; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar[i])));
    asm volatile("fence.mbarrier_init.release.cluster;");
    const long long t0 = clock64();
    int row = (blockIdx.x * 977) % (rows_total - 128);
    for (int i = 0; i < iters + 4; ++i) {
      const int s = i & 3;
      if (i >= 4) {
        uint32_t ok = 0;
        const uint32_t par = ((i >> 2) - 1) & 1;
        while (!ok)
          asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.b32 %0, 1, 0, p;\n}\n"
                       : "=r"(ok) : "r"(smem_u32(&bar[s])), "r"(par));
      }
      if (i < iters) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar[s])), "r"(box_bytes));
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                     ::"r"(smem_u32(smem + s * 16384)), "l"((uint64_t)&tm), "r"(smem_u32(&bar[s])), "r"(0), "r"(row) : "memory");
        row += 128 * 148;
        if (row > rows_total - 128) row -= rows_total - 128;
      }
    }
    out[blockIdx.x] = clock64() - t0;
  }
}

int main() {
  const int rows = 1 << 18;
  void* d;
  cudaMalloc(&d, (size_t)rows * 256 * 2);
  cudaMemset(d, 0, (size_t)rows * 256 * 2);
  long long* o;
  cudaMalloc(&o, 148 * 8);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  const int iters = 400;
  for (int C : {8, 16, 32, 64, 256})
    for (int narrow = 0; narrow < 2; ++narrow) {
      const int boxc = narrow ? (C < 64 ? C : 64) : 64;
      CUtensorMap tm;
      cuuint64_t gdim[2] = {(cuuint64_t)C, (cuuint64_t)rows}, gstr[1] = {(cuuint64_t)C * 2};
      cuuint32_t bdim[2] = {(cuuint32_t)boxc, 128}, estr[2] = {1, 1};
      CUtensorMapSwizzle sw = boxc == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : boxc == 32 ? CU_TENSOR_MAP_SWIZZLE_64B
                              : boxc == 16 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_NONE;
      CUresult r = cuTensorMapEncodeTiled(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, d, gdim, gstr, bdim, estr,
                                          CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); continue; }
      for (int grid : {1, 148}) {
        k<<<grid, 32, 64 * 1024>>>(tm, boxc * 128 * 2, rows, iters, o);
        cudaError_t e = cudaDeviceSynchronize();
        long long h[148], mx = 0;
        cudaMemcpy(h, o, grid * 8, cudaMemcpyDeviceToHost);
        for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
        printf("C %3d box %2d x128 rows grid %3d : %.0f cycles/box (4 in flight) %s\n", C, boxc, grid, (double)mx / iters,
               cudaGetErrorString(e));
      }
    }
  return 0;
}
