// Deterministic cross-CTA reduction of per-channel statistics (BatchNorm sum(x), sum(x^2), sum(dz), sum(dz * xhat)).
//
// No floating-point atomics: every producing CTA folds its rows in a fixed order, STORES its partial sums into its own
// row of a workspace and draws a ticket; the CTA that draws the last ticket adds the rows in ascending row order in
// double precision and writes the final sums as doubles (the consumer forms mean / variance from them in double, so the
// variance is not a difference of two rounded fp32 sums).  Which CTA finishes last varies from run to run, the order of
// the additions does not: results are bit-reproducible.
//
// Replaces the arrival-order atomics behind torch's batch_norm statistics
// (/root/reference/audio_slowfast/models/resnet_helper.py:181-207 -> nn.BatchNorm2d, training mode).
#pragma once
#include "common.cuh"

namespace asf {

// Called by ALL threads of the CTA after their partial-sum stores.  Returns true (in every thread) in the CTA that
// arrived last among `total` CTAs sharing `ticket`; that CTA may then read every partial row.  The counter resets
// itself, so a zero-initialised ticket serves every launch (launches sharing a ticket must be stream-ordered).
__device__ __forceinline__ bool stats_last_cta(unsigned* ticket, unsigned total, unsigned* s_flag) {
  __threadfence();  // this thread's partial sums are visible device-wide before the ticket is drawn
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned t = atomicAdd(ticket, 1u);
    const unsigned last = (t == total - 1u) ? 1u : 0u;
    if (last) *ticket = 0u;  // every other CTA has already drawn
    *s_flag = last;
  }
  __syncthreads();
  const bool last = *s_flag != 0u;
  if (last) __threadfence();
  return last;
}

// Ordered sum over P workspace rows for `slots` independent columns: slot s of row p is ws[p * row_stride + s].
// L lanes share a slot (lane l adds rows l, l + L, ... in ascending order; the lanes are then folded in ascending
// order through `scratch`, >= blockDim.x doubles of shared memory).  emit(s, sum) is called once per slot.
// Must be called by all threads of the CTA (it synchronises).
template <typename Emit>
__device__ __forceinline__ void stats_ordered_reduce(const float* ws, int P, long long row_stride, int slots,
                                                     double* scratch, Emit emit) {
  const int nthreads = (int)blockDim.x, tid = (int)threadIdx.x;
  int L = 1;
  while (L < 16 && slots * (L * 2) <= nthreads && L * 2 <= P) L *= 2;
  const int per = nthreads / L;  // slots handled per pass
  const int sl = tid % per, l = tid / per;
  for (int s0 = 0; s0 < slots; s0 += per) {
    const int s = s0 + sl;
    const bool act = s < slots && l < L;
    double acc = 0.0;
    if (act) {
      const float* src = ws + s;
      int p = l;
      for (; p + 3 * L < P; p += 4 * L) {  // four independent loads in flight, added in row order
        const float a = __ldcg(src + (long long)p * row_stride);
        const float b = __ldcg(src + (long long)(p + L) * row_stride);
        const float c = __ldcg(src + (long long)(p + 2 * L) * row_stride);
        const float d = __ldcg(src + (long long)(p + 3 * L) * row_stride);
        acc += (double)a;
        acc += (double)b;
        acc += (double)c;
        acc += (double)d;
      }
      for (; p < P; p += L) acc += (double)__ldcg(src + (long long)p * row_stride);
    }
    if (L > 1) {
      scratch[tid] = acc;
      __syncthreads();
      if (act && l == 0)
        for (int j = 1; j < L; ++j) acc += scratch[j * per + sl];
      __syncthreads();
    }
    if (act && l == 0) emit(s, acc);
  }
}

// the same over rows of doubles (fp32 parity mode keeps its partial sums in double)
template <typename Emit>
__device__ __forceinline__ void stats_ordered_reduce_f64(const double* ws, int P, long long row_stride, int slots,
                                                         Emit emit) {
  for (int s = (int)threadIdx.x; s < slots; s += (int)blockDim.x) {
    double acc = 0.0;
    for (int p = 0; p < P; ++p) acc += __ldcg(ws + (long long)p * row_stride + s);
    emit(s, acc);
  }
}

}  // namespace asf
