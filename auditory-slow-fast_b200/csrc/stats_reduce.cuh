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
// A thread owns FOUR adjacent slots (one 16-byte load per row; slots, row_stride multiples of 4, ws 16-byte aligned --
// every caller's layout) and L lanes share a slot group: lane l adds rows l, l + L, ... in ascending order with eight
// independent loads in flight, the lanes are then folded in ascending order through `scratch` (>= 4 * blockDim.x
// doubles of shared memory).  emit(s, sum) is called once per slot.  Must be called by all threads of the CTA.
template <typename Emit>
__device__ __forceinline__ void stats_ordered_reduce(const float* ws, int P, long long row_stride, int slots,
                                                     double* scratch, Emit emit) {
  const int nthreads = (int)blockDim.x, tid = (int)threadIdx.x;
  const int groups = slots >> 2;  // 4-slot groups
  int L = 1;
  while (L < 32 && groups * (L * 2) <= nthreads && L * 2 <= P) L *= 2;
  const int per = nthreads / L;  // slot groups handled per pass
  const int gl = tid % per, l = tid / per;
  for (int g0 = 0; g0 < groups; g0 += per) {
    const int g = g0 + gl;
    const bool act = g < groups && l < L;
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
    if (act) {
      const float* src = ws + 4 * g;
      int p = l;
      for (; p + 7 * L < P; p += 8 * L) {
        float4 v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = __ldcg((const float4*)(src + (long long)(p + j * L) * row_stride));
#pragma unroll
        for (int j = 0; j < 8; ++j) {  // added in row order
          a0 += (double)v[j].x;
          a1 += (double)v[j].y;
          a2 += (double)v[j].z;
          a3 += (double)v[j].w;
        }
      }
      for (; p < P; p += L) {
        const float4 v = __ldcg((const float4*)(src + (long long)p * row_stride));
        a0 += (double)v.x;
        a1 += (double)v.y;
        a2 += (double)v.z;
        a3 += (double)v.w;
      }
    }
    if (L > 1) {
      scratch[4 * tid] = a0;
      scratch[4 * tid + 1] = a1;
      scratch[4 * tid + 2] = a2;
      scratch[4 * tid + 3] = a3;
      __syncthreads();
      if (act && l == 0)
        for (int j = 1; j < L; ++j) {
          const double* q = scratch + 4 * (j * per + gl);
          a0 += q[0];
          a1 += q[1];
          a2 += q[2];
          a3 += q[3];
        }
      __syncthreads();
    }
    if (act && l == 0) {
      emit(4 * g, a0);
      emit(4 * g + 1, a1);
      emit(4 * g + 2, a2);
      emit(4 * g + 3, a3);
    }
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
