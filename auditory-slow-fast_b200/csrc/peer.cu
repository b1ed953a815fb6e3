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
//   1. PUSHES its n doubles into slot [parity][my rank] of EVERY rank's buffer (its own included): remote stores are
//      posted, they do not wait for a round trip the way remote loads do,
//   2. after a system-scope fence, stores the epoch into flag[my rank] of every peer,
//   3. spins on its OWN flags until every peer shows the epoch,
//   4. adds the world slots of its own buffer -- local memory -- IN RANK ORDER in double: the same order on every
//      rank, so all ranks hold bit-identical sums, reproducibly.
// (The first version pulled: every rank read the peers' slots with remote loads, 7 x 2C doubles through one CTA at 8
// GPUs -- 30 us per exchange against 9 us at 2 GPUs.)
// Slot reuse is safe with two parities: a rank writes epoch e + 2 into a peer's slot only after passing step 3 of
// epoch e + 1, i.e. after that peer has signalled e + 1, which it does after its kernel of epoch e -- reads included --
// has finished (exchanges of one (layer, direction) are stream-ordered on every rank).
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
  const long long my_off = ((long long)parity * world + rank) * n;  // this rank's slot inside any rank's buffer
  // n is even (2 C): 16-byte stores, peers interleaved so that consecutive threads hit consecutive addresses of one peer
  const int n2 = n >> 1;
  for (int idx = threadIdx.x; idx < n2 * world; idx += blockDim.x) {
    const int r = idx / n2, i = idx - r * n2;
    double2* dst = (double2*)((double*)((char*)peer_base[r] + slot_off) + my_off) + i;
    *dst = ((const double2*)data)[i];
  }
  __threadfence_system();
  __syncthreads();
  if ((int)threadIdx.x < world && (int)threadIdx.x != rank) {
    unsigned* peer_flags = (unsigned*)((char*)peer_base[threadIdx.x] + flag_off);
    st_volatile_u32(peer_flags + rank, e);
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
  const double* mine = (const double*)((const char*)peer_base[rank] + slot_off) + (long long)parity * world * n;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    double acc = 0.0;
    for (int r = 0; r < world; ++r) acc += ld_volatile_f64(mine + (long long)r * n + i);
    data[i] = acc;
  }
  __syncthreads();
  if (threadIdx.x == 0) *epoch = e;
}

}  // namespace asf

using namespace asf;

extern "C" int asf_peer_allreduce_f64(double* data, int n, const unsigned long long* peer_base_dev, int rank, int world,
                                      long long slot_off_bytes, long long flag_off_bytes, unsigned* epoch, void* stream) {
  ASF_CHECK_ARG(data && peer_base_dev && epoch && n > 0 && n % 2 == 0 && world >= 2 && world <= 64 && rank >= 0 && rank < world,
                "asf_peer_allreduce_f64: bad arguments");
  ASF_CHECK_ARG(slot_off_bytes % 16 == 0 && flag_off_bytes % 4 == 0, "asf_peer_allreduce_f64: misaligned offsets (slots 8-byte, flags 4-byte; use 16-byte slot offsets)");
  ASF_CUDA(launch_pdl(peer_allreduce_f64_kernel, dim3(1), dim3(256), 0, (cudaStream_t)stream, data, n, peer_base_dev, rank, world,
                      slot_off_bytes, flag_off_bytes, epoch));
  ASF_LAUNCH_CHECK();
  return 0;
}
