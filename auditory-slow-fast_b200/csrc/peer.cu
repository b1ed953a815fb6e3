// SyncBatchNorm statistics exchange over NVLink peer memory: one tiny kernel per exchange instead of an NCCL
// all-reduce.
//
// With NUM_GPUS > 1 the reference converts every BatchNorm to torch.nn.SyncBatchNorm
// (/root/reference/audio_slowfast/models/build.py:45-46): each train-mode layer needs the sum over ranks of its
// per-channel statistics in the forward pass (sum x, sum x^2) and again in the backward pass (sum dz, sum dz xhat) --
// 220 exchanges of 16 ... 4096 doubles per from-scratch training step, each on the critical path of the layer chain.
// A library all-reduce costs ~20 us of latency per call there (13.1 ms per step at 2 GPUs against 8.4 ms at one).
//
// Every rank owns a "symmetric" buffer mapped into all peers (torch.distributed._symmetric_memory; NVSwitch gives every
// GPU a direct path to every peer).  For one exchange a rank
//   1. copies its n doubles into its own slot (parity of the exchange's epoch: a slot is reused every second time),
//   2. after a system-scope fence, stores the epoch into flag[my rank] of EVERY peer (remote 4-byte stores),
//   3. spins until its own flags show the epoch from every peer,
//   4. reads every peer's slot with volatile remote loads and adds the ranks IN RANK ORDER in double -- the same order
//      on every rank, so all ranks hold bit-identical sums, reproducibly.
// Slot reuse is safe: a rank passes step 3 of epoch e + 1 only after every peer has signalled e + 1, which a peer does
// after it has finished epoch e (exchanges of one (layer, direction) are stream-ordered on every rank).
// The kernels of different ranks wait for one another, so the ranks must be different GPUs (they are: one process per
// GPU); on each GPU the waiting kernel is a single CTA.
#include "common.cuh"

namespace asf {

__device__ __forceinline__ unsigned ld_volatile_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ double ld_volatile_f64(const double* p) {
  double v;
  asm volatile("ld.volatile.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_volatile_u32(unsigned* p, unsigned v) {
  asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__global__ void __launch_bounds__(256) peer_allreduce_f64_kernel(double* __restrict__ data, int n,
                                                                 const unsigned long long* __restrict__ peer_base, int rank,
                                                                 int world, long long slot_off, long long flag_off,
                                                                 unsigned* __restrict__ epoch) {
  // NO early trigger of the dependent launch: the next kernel of this stream (a few hundred CTAs of the normalise
  // pass) would become resident and sit in its griddepcontrol.wait, holding registers and thread slots on every SM
  // while this kernel waits for the other GPUs -- enough to keep the OTHER pathway's tcgen05 CTAs (a whole SM each)
  // from being scheduled, i.e. to keep this rank from ever reaching the exchange its peer is waiting in.
  pdl_wait();
  const unsigned e = *epoch + 1u;
  const int parity = (int)(e & 1u);
  double* my_slot = (double*)((char*)peer_base[rank] + slot_off) + (long long)parity * n;
  for (int i = threadIdx.x; i < n; i += blockDim.x) my_slot[i] = data[i];
  __threadfence_system();
  __syncthreads();
  if ((int)threadIdx.x < world && (int)threadIdx.x != rank) {
    unsigned* peer_flags = (unsigned*)((char*)peer_base[threadIdx.x] + flag_off);
    st_volatile_u32(peer_flags + rank, e);
  }
  if ((int)threadIdx.x < world && (int)threadIdx.x != rank) {
    const unsigned* my_flags = (const unsigned*)((const char*)peer_base[rank] + flag_off);
    const long long t0 = clock64();
    // epochs only grow: ">= e" (as a wrapped difference) also accepts a peer that is already one exchange ahead
    while ((int)(ld_volatile_u32(my_flags + threadIdx.x) - e) < 0) {
      if (clock64() - t0 > 20000000000LL) {
        printf("asf peer exchange: rank %d timed out waiting for rank %d (epoch %u)\n", rank, (int)threadIdx.x, e);
        __trap();
      }
    }
  }
  __threadfence_system();
  __syncthreads();
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    double acc = 0.0;
    for (int r = 0; r < world; ++r) {
      const double* slot = (const double*)((const char*)peer_base[r] + slot_off) + (long long)parity * n;
      acc += (r == rank) ? my_slot[i] : ld_volatile_f64(slot + i);
    }
    data[i] = acc;
  }
  __syncthreads();
  if (threadIdx.x == 0) *epoch = e;
}

}  // namespace asf

using namespace asf;

extern "C" int asf_peer_allreduce_f64(double* data, int n, const unsigned long long* peer_base_dev, int rank, int world,
                                      long long slot_off_bytes, long long flag_off_bytes, unsigned* epoch, void* stream) {
  ASF_CHECK_ARG(data && peer_base_dev && epoch && n > 0 && world >= 2 && world <= 64 && rank >= 0 && rank < world,
                "asf_peer_allreduce_f64: bad arguments");
  ASF_CHECK_ARG(slot_off_bytes % 8 == 0 && flag_off_bytes % 4 == 0, "asf_peer_allreduce_f64: misaligned offsets");
  ASF_CUDA(launch_pdl(peer_allreduce_f64_kernel, dim3(1), dim3(256), 0, (cudaStream_t)stream, data, n, peer_base_dev, rank, world,
                      slot_off_bytes, flag_off_bytes, epoch));
  ASF_LAUNCH_CHECK();
  return 0;
}
