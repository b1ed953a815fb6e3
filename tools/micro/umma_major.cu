// Microbenchmark: tcgen05.mma 128xNx16 bf16 issue rate with K-major vs MN-major shared-memory operands.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o umma_major umma_major.cu && ./umma_major
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t make_sdesc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
__host__ __device__ constexpr uint32_t make_idesc(int M, int N, int a_mn, int b_mn) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__global__ void __launch_bounds__(128, 1) k(int N, int a_mn, int b_mn, int iters, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_base;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 48 * 1024 / 4; i += 128) ((uint32_t*)smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_base)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("fence.proxy.async.shared::cta;");
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  if (warp == 1 && lane == 0) {
    const uint32_t idesc = make_idesc(128, N, a_mn, b_mn);
    const uint32_t s0 = smem_u32(smem);
    // K-major: 8-row groups 1024 B apart (SBO), LBO unused; MN-major: 64-element chunks LBO = 8192 apart, K groups SBO 1024
    const uint64_t ad = a_mn ? make_sdesc(s0, 8192, 1024) : make_sdesc(s0, 16, 1024);
    const uint64_t bd = b_mn ? make_sdesc(s0 + 16384, 8192, 1024) : make_sdesc(s0 + 16384, 16, 1024);
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      const uint64_t ka = a_mn ? (uint64_t)((i & 3) * 128) : (uint64_t)((i & 3) * 2);
      const uint64_t kb = b_mn ? (uint64_t)((i & 3) * 128) : (uint64_t)((i & 3) * 2);
      asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}\n" ::"r"(
                       tmem_base),
                   "l"(ad + ka), "l"(bd + kb), "r"(idesc), "r"(i));
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)));
    uint32_t ok = 0;
    while (!ok)
      asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\nselp.b32 %0, 1, 0, p;\n}\n"
                   : "=r"(ok) : "r"(smem_u32(&bar)));
    out[blockIdx.x] = clock64() - t0;
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base));
}

int main() {
  long long* d;
  cudaMalloc(&d, 148 * 8);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  const int iters = 2000;
  for (int grid : {1, 148})
    for (int N : {64, 128, 256})
      for (int am = 0; am < 2; ++am)
        for (int bm = 0; bm < 2; ++bm) {
          k<<<grid, 128, 64 * 1024>>>(N, am, bm, iters, d);
          cudaError_t e = cudaDeviceSynchronize();
          long long h[148];
          cudaMemcpy(h, d, grid * 8, cudaMemcpyDeviceToHost);
          long long mx = 0;
          for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
          printf("grid %3d N %3d A %s B %s : %.1f cycles/MMA (%s)\n", grid, N, am ? "MN" : "K ", bm ? "MN" : "K ",
                 (double)mx / iters, cudaGetErrorString(e));
        }
  return 0;
}
