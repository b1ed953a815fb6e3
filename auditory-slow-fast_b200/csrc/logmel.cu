// Batched log-mel extraction: framing + Hann + DFT magnitude + HTK mel projection + log, one pass,
// waveform in -> [B,1,T,128] out, nothing intermediate in HBM.
//
// Replaces librosa.stft / filters.mel / np.dot / np.log of
//   /root/reference/audio_slowfast/datasets/audio_loader_epic.py:66-97
// and the "edge" padding of :151-154.
//
// Only win (=240) of the n_fft (=2048) samples of a frame are non-zero after the centred window, so
// the DFT is evaluated as a dense [frames x win] . [win x 2*bins] contraction against a precomputed
// windowed basis instead of a 2048-point FFT that is 88 % zeros.  Each CTA owns FR consecutive frames of
// one clip: the (overlapping) samples are staged once in shared memory, every thread sweeps the basis
// column(s) of its bins (coalesced float2 reads, L2-resident 2 MB table, reused by all CTAs), magnitudes
// go to shared memory, and the mel projection + log run from there using the per-filter non-zero range.
#include <math.h>

#include <vector>

#include "common.cuh"

namespace asf {

constexpr int LM_FR = 32;        // frames per CTA
constexpr int LM_THREADS = 256;
constexpr int LM_KB = LM_THREADS;  // bins per pass
constexpr int LM_MEL_PER_THREAD = 16;  // (frame, mel) outputs per thread: LM_FR * n_mels / LM_THREADS for n_mels = 128

// One CTA = LM_FR consecutive frames of one clip.  The (overlapping) samples of the frames are staged once; bins are
// processed in passes of 256 (one bin per thread, LM_FR frames in registers, four samples per step so that one
// LDS.128 per frame feeds 8 FMAs and one basis load serves 32 frames -- the 2 MB basis table is read once per CTA from
// L2).  After each pass the magnitudes of its 256 bins sit in shared memory and every thread advances the mel dot
// products of its fixed (frame, mel) outputs over the part of each filter that falls into the pass, in ascending bin
// order, so the summation order equals a single sweep.  Leftover bins that do not fill a pass (the Nyquist bin) are
// split over 8 lanes per frame instead of costing a pass for one column.
// HOP_C / WIN_C != 0: compile-time hop / window (the shipped 5 ms / 10 ms at 24 kHz), so every shared-memory sample
// address is a constant offset.  0 / 0: any hop / window.  (Packed fma.rn.f32x2 was measured slower here: it halves
// the FMA count but needs twice the LDS traffic for the duplicated samples.)
template <int HOP_C, int WIN_C>
__global__ void __launch_bounds__(LM_THREADS) logmel_kernel(const float* __restrict__ wave, int64_t wave_ld,
                                                            int n_samples, int T_valid, int T_out, int win, int hop,
                                                            int first_off, int n_bins, int nbp, int n_mels, float eps,
                                                            const float2* __restrict__ basisT,
                                                            const float* __restrict__ melT,
                                                            const int32_t* __restrict__ mel_range,
                                                            float* __restrict__ fast) {
  extern __shared__ float smem[];
  constexpr bool kFast = HOP_C != 0;
  if (kFast) {
    hop = HOP_C;
    win = WIN_C;
  }
  const int span = (LM_FR - 1) * hop + win;
  float* sx = smem;                        // [span]
  float* smag = smem + ((span + 3) & ~3);  // [LM_FR][LM_KB]
  const int b = blockIdx.y;
  const int t0 = blockIdx.x * LM_FR;
  const float* wv = wave + (int64_t)b * wave_ld;
  // frame t (t < T_valid) reads samples t*hop + first_off + j; rows t >= T_valid replicate frame T_valid-1
  // (np.pad "edge").  A CTA that holds at least one real frame computes all its frames from one staged span and maps
  // the padded rows onto its last real frame; a CTA that lies entirely in the padding recomputes frame T_valid-1.
  const int fbase = (t0 < T_valid) ? t0 : (T_valid - 1);
  const int base = fbase * hop + first_off;
  for (int i = threadIdx.x; i < span; i += LM_THREADS) {
    const int s = base + i;
    sx[i] = (s >= 0 && s < n_samples) ? wv[s] : 0.f;
  }
  float macc[LM_MEL_PER_THREAD];
#pragma unroll
  for (int q = 0; q < LM_MEL_PER_THREAD; ++q) macc[q] = 0.f;
  const bool vec4 = ((hop | win) & 3) == 0;
  const int full_bins = (n_bins / LM_KB) * LM_KB;
  __syncthreads();
  for (int kb = 0; kb < n_bins; kb += LM_KB) {
    if (kb < full_bins) {
      const int k = kb + threadIdx.x;
      float re[LM_FR], im[LM_FR];
#pragma unroll
      for (int f = 0; f < LM_FR; ++f) re[f] = im[f] = 0.f;
      const float2* bp = basisT + k;
      if (kFast) {
        static_assert((HOP_C % 4 == 0) && (WIN_C % 4 == 0), "constant-offset path: hop and window multiples of 4");
#pragma unroll 1
        for (int j = 0; j < WIN_C; j += 4) {
          float2 c[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) c[q] = bp[(int64_t)(j + q) * nbp];
          const float* sj = sx + j;
#pragma unroll
          for (int f = 0; f < LM_FR; ++f) {
            const float4 xs = *(const float4*)(sj + f * HOP_C);
            re[f] = fmaf(c[0].x, xs.x, re[f]); im[f] = fmaf(c[0].y, xs.x, im[f]);
            re[f] = fmaf(c[1].x, xs.y, re[f]); im[f] = fmaf(c[1].y, xs.y, im[f]);
            re[f] = fmaf(c[2].x, xs.z, re[f]); im[f] = fmaf(c[2].y, xs.z, im[f]);
            re[f] = fmaf(c[3].x, xs.w, re[f]); im[f] = fmaf(c[3].y, xs.w, im[f]);
          }
        }
      } else if (vec4) {
        for (int j = 0; j < win; j += 4) {
          float2 c[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) c[q] = bp[(int64_t)(j + q) * nbp];
#pragma unroll
          for (int f = 0; f < LM_FR; ++f) {
            const float4 xs = *(const float4*)(sx + f * hop + j);
            re[f] = fmaf(c[0].x, xs.x, re[f]); im[f] = fmaf(c[0].y, xs.x, im[f]);
            re[f] = fmaf(c[1].x, xs.y, re[f]); im[f] = fmaf(c[1].y, xs.y, im[f]);
            re[f] = fmaf(c[2].x, xs.z, re[f]); im[f] = fmaf(c[2].y, xs.z, im[f]);
            re[f] = fmaf(c[3].x, xs.w, re[f]); im[f] = fmaf(c[3].y, xs.w, im[f]);
          }
        }
      } else {
        for (int j = 0; j < win; ++j) {
          const float2 cs = bp[(int64_t)j * nbp];
#pragma unroll
          for (int f = 0; f < LM_FR; ++f) {
            const float xs = sx[f * hop + j];
            re[f] = fmaf(cs.x, xs, re[f]);
            im[f] = fmaf(cs.y, xs, im[f]);
          }
        }
      }
#pragma unroll
      for (int f = 0; f < LM_FR; ++f) smag[f * LM_KB + threadIdx.x] = sqrtf(re[f] * re[f] + im[f] * im[f]);
    } else {
      // leftover bins: 8 lanes per frame split the window and fold with shuffles
      static_assert(LM_THREADS == 8 * LM_FR, "leftover-bin path maps 8 lanes to each frame");
      const int f = threadIdx.x >> 3, part = threadIdx.x & 7;
      for (int k = kb; k < n_bins; ++k) {
        float re = 0.f, im = 0.f;
        for (int j = part; j < win; j += 8) {
          const float2 cs = basisT[(int64_t)j * nbp + k];
          const float xs = sx[f * hop + j];
          re = fmaf(cs.x, xs, re);
          im = fmaf(cs.y, xs, im);
        }
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) {
          re += __shfl_xor_sync(0xffffffffu, re, o);
          im += __shfl_xor_sync(0xffffffffu, im, o);
        }
        if (part == 0) smag[f * LM_KB + (k - kb)] = sqrtf(re * re + im * im);
      }
    }
    __syncthreads();
    // mel dot products over the bins of this pass (mel fastest across threads: melT reads coalesce)
    const int kend = min(kb + LM_KB, n_bins);
#pragma unroll
    for (int q = 0; q < LM_MEL_PER_THREAD; ++q) {
      const int i = threadIdx.x + q * LM_THREADS;
      if (i < LM_FR * n_mels) {
        const int f = i / n_mels, m = i - f * n_mels;
        const int k0 = max(mel_range[2 * m], kb), k1 = min(mel_range[2 * m + 1], kend);
        float acc = macc[q];
        for (int k = k0; k < k1; ++k) acc = fmaf(melT[(int64_t)k * n_mels + m], smag[f * LM_KB + (k - kb)], acc);
        macc[q] = acc;
      }
    }
    __syncthreads();
  }
  // frame f of this CTA serves row t0 + f, and every padded row after the last real frame
#pragma unroll
  for (int q = 0; q < LM_MEL_PER_THREAD; ++q) {
    const int i = threadIdx.x + q * LM_THREADS;
    if (i < LM_FR * n_mels) {
      const int f = i / n_mels, m = i - f * n_mels;
      const float v = logf(macc[q] + eps);
      const int t = fbase + f;
      if (t0 < T_valid) {
        if (t < T_valid && t < T_out) fast[((int64_t)b * T_out + t) * n_mels + m] = v;
        if (t == T_valid - 1)
          for (int tp = T_valid; tp < min(t0 + LM_FR, T_out); ++tp) fast[((int64_t)b * T_out + tp) * n_mels + m] = v;
      } else if (f == 0) {
        for (int tp = t0; tp < min(t0 + LM_FR, T_out); ++tp) fast[((int64_t)b * T_out + tp) * n_mels + m] = v;
      }
    }
  }
}

}  // namespace asf

using namespace asf;

extern "C" {

int asf_logmel_tables(float* basisT, float* melT, int32_t* mel_range, int sr, int n_fft, int win, int n_mels,
                      int n_bins_pad, void* stream) {
  ASF_CHECK_ARG(basisT && melT && mel_range && sr > 0 && n_fft > 0 && win > 0 && win <= n_fft && n_mels > 0 &&
                    n_bins_pad >= n_fft / 2 + 1,
                "asf_logmel_tables: bad arguments");
  const int n_bins = n_fft / 2 + 1, nbp = n_bins_pad;
  const int lpad = (n_fft - win) / 2;
  const double PI = 3.14159265358979323846;
  std::vector<float> hb((size_t)win * nbp * 2, 0.f), hm((size_t)nbp * n_mels, 0.f);
  std::vector<int32_t> hr((size_t)n_mels * 2, 0);
  for (int j = 0; j < win; ++j) {
    const double w = 0.5 - 0.5 * cos(2.0 * PI * (double)j / (double)win);  // periodic Hann (fftbins=True)
    for (int k = 0; k < n_bins; ++k) {
      const int64_t ph = ((int64_t)k * (lpad + j)) % n_fft;                 // exact phase reduction
      const double ang = 2.0 * PI * (double)ph / (double)n_fft;
      hb[((size_t)j * nbp + k) * 2 + 0] = (float)(w * cos(ang));
      hb[((size_t)j * nbp + k) * 2 + 1] = (float)(-w * sin(ang));
    }
  }
  // librosa.filters.mel(htk=True, norm=None)
  const double mel_max = 2595.0 * log10(1.0 + (0.5 * sr) / 700.0);
  std::vector<double> mel_f(n_mels + 2);
  for (int i = 0; i < n_mels + 2; ++i) {
    double m = (i == n_mels + 1) ? mel_max : (mel_max / (double)(n_mels + 1)) * (double)i;
    mel_f[i] = 700.0 * (pow(10.0, m / 2595.0) - 1.0);
  }
  for (int m = 0; m < n_mels; ++m) {
    int k0 = n_bins, k1 = 0;
    for (int k = 0; k < n_bins; ++k) {
      const double fk = (double)k * (double)sr / (double)n_fft;
      const double lower = (fk - mel_f[m]) / (mel_f[m + 1] - mel_f[m]);
      const double upper = (mel_f[m + 2] - fk) / (mel_f[m + 2] - mel_f[m + 1]);
      double v = lower < upper ? lower : upper;
      if (v < 0.0) v = 0.0;
      const float vf = (float)v;
      hm[(size_t)k * n_mels + m] = vf;
      if (vf > 0.f) {
        if (k < k0) k0 = k;
        if (k + 1 > k1) k1 = k + 1;
      }
    }
    if (k1 <= k0) k0 = k1 = 0;
    hr[2 * m] = k0;
    hr[2 * m + 1] = k1;
  }
  cudaStream_t st = (cudaStream_t)stream;
  ASF_CUDA(cudaMemcpyAsync(basisT, hb.data(), hb.size() * sizeof(float), cudaMemcpyHostToDevice, st));
  ASF_CUDA(cudaMemcpyAsync(melT, hm.data(), hm.size() * sizeof(float), cudaMemcpyHostToDevice, st));
  ASF_CUDA(cudaMemcpyAsync(mel_range, hr.data(), hr.size() * sizeof(int32_t), cudaMemcpyHostToDevice, st));
  ASF_CUDA(cudaStreamSynchronize(st));  // host staging vectors die at return
  return 0;
}

int asf_logmel_fwd(const float* wave, int64_t wave_ld, int B, int n_samples, int n_fft, int win, int hop,
                   int n_mels, float eps, const float* basisT, const float* melT, const int32_t* mel_range,
                   int n_bins_pad, float* fast, int T_out, void* stream) {
  ASF_CHECK_ARG(wave && basisT && melT && mel_range && fast, "asf_logmel_fwd: null pointer");
  ASF_CHECK_ARG(B > 0 && n_samples >= 0 && hop > 0 && win > 0 && win <= n_fft && n_mels > 0 && wave_ld >= n_samples,
                "asf_logmel_fwd: bad sizes");
  const int T_valid = 1 + n_samples / hop;
  if (T_out <= 0) T_out = T_valid;
  ASF_CHECK_ARG(T_out >= T_valid, "asf_logmel_fwd: T_out (%d) smaller than the %d frames of the clip (no truncation)",
                T_out, T_valid);
  const int n_bins = n_fft / 2 + 1;
  const int first_off = -(n_fft / 2) + (n_fft - win) / 2;
  const int span = (LM_FR - 1) * hop + win;
  const int fixed = (hop == 120 && win == 240) ? 1 : 0;
  size_t smem = (size_t)(((span + 3) & ~3) + LM_FR * LM_KB) * sizeof(float);
  ASF_CHECK_ARG((int64_t)LM_FR * n_mels <= (int64_t)LM_MEL_PER_THREAD * LM_THREADS,
                "asf_logmel_fwd: n_mels %d exceeds the %d mel bands the kernel keeps in registers", n_mels,
                LM_MEL_PER_THREAD * LM_THREADS / LM_FR);
  ASF_CHECK_ARG(smem <= 227 * 1024, "asf_logmel_fwd: window/hop too large for the shared-memory staging (%zu B)", smem);
  static size_t smem_set[2] = {0, 0};  // raise the dynamic shared-memory limit only when a larger request appears
  if (smem > smem_set[fixed]) {
    if (fixed)
      ASF_CUDA(cudaFuncSetAttribute(logmel_kernel<120, 240>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    else
      ASF_CUDA(cudaFuncSetAttribute(logmel_kernel<0, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    smem_set[fixed] = smem;
  }
  dim3 grid((unsigned)cdiv(T_out, LM_FR), B);
  if (fixed)
    logmel_kernel<120, 240><<<grid, LM_THREADS, smem, (cudaStream_t)stream>>>(
        wave, wave_ld, n_samples, T_valid, T_out, win, hop, first_off, n_bins, n_bins_pad, n_mels, eps,
        (const float2*)basisT, melT, mel_range, fast);
  else
    logmel_kernel<0, 0><<<grid, LM_THREADS, smem, (cudaStream_t)stream>>>(
        wave, wave_ld, n_samples, T_valid, T_out, win, hop, first_off, n_bins, n_bins_pad, n_mels, eps,
        (const float2*)basisT, melT, mel_range, fast);
  ASF_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
