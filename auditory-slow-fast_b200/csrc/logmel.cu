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

constexpr int LM_FR = 16;        // frames per CTA
constexpr int LM_THREADS = 256;

__global__ void __launch_bounds__(LM_THREADS) logmel_kernel(const float* __restrict__ wave, int64_t wave_ld,
                                                            int n_samples, int T_valid, int T_out, int win, int hop,
                                                            int first_off, int n_bins, int nbp, int n_mels, float eps,
                                                            const float2* __restrict__ basisT,
                                                            const float* __restrict__ melT,
                                                            const int32_t* __restrict__ mel_range,
                                                            float* __restrict__ fast) {
  extern __shared__ float smem[];
  const int span = (LM_FR - 1) * hop + win;
  float* sx = smem;                 // [span]
  float* smag = smem + ((span + 3) & ~3);  // [LM_FR][nbp]
  const int b = blockIdx.y;
  const int t0 = blockIdx.x * LM_FR;
  const float* wv = wave + (int64_t)b * wave_ld;

  // frame t (t < T_valid) reads samples t*hop + first_off + j ; frames >= T_valid replicate frame T_valid-1.
  // When the whole CTA is within the valid range the frames overlap regularly and one span is staged;
  // otherwise (tail CTA) each frame is staged separately through the same code with a per-frame base.
  const bool regular = (t0 + LM_FR <= T_valid);
  const int nbins_pass = (n_bins + LM_THREADS - 1) / LM_THREADS;

  if (regular) {
    const int base = t0 * hop + first_off;
    for (int i = threadIdx.x; i < span; i += LM_THREADS) {
      int s = base + i;
      sx[i] = (s >= 0 && s < n_samples) ? wv[s] : 0.f;
    }
    __syncthreads();
    for (int pass = 0; pass < nbins_pass; ++pass) {
      const int k = pass * LM_THREADS + threadIdx.x;
      if (k < n_bins) {
        float re[LM_FR], im[LM_FR];
#pragma unroll
        for (int f = 0; f < LM_FR; ++f) re[f] = im[f] = 0.f;
        for (int j = 0; j < win; ++j) {
          const float2 cs = basisT[(int64_t)j * nbp + k];
#pragma unroll
          for (int f = 0; f < LM_FR; ++f) {
            const float xs = sx[f * hop + j];
            re[f] = fmaf(cs.x, xs, re[f]);
            im[f] = fmaf(cs.y, xs, im[f]);
          }
        }
#pragma unroll
        for (int f = 0; f < LM_FR; ++f) smag[f * nbp + k] = sqrtf(re[f] * re[f] + im[f] * im[f]);
      }
    }
  } else {
    for (int f = 0; f < LM_FR; ++f) {
      int t = t0 + f;
      if (t >= T_out) break;
      if (t >= T_valid) t = T_valid - 1;
      const int base = t * hop + first_off;
      __syncthreads();
      for (int i = threadIdx.x; i < win; i += LM_THREADS) {
        int s = base + i;
        sx[i] = (s >= 0 && s < n_samples) ? wv[s] : 0.f;
      }
      __syncthreads();
      for (int pass = 0; pass < nbins_pass; ++pass) {
        const int k = pass * LM_THREADS + threadIdx.x;
        if (k < n_bins) {
          float re = 0.f, im = 0.f;
          for (int j = 0; j < win; ++j) {
            const float2 cs = basisT[(int64_t)j * nbp + k];
            re = fmaf(cs.x, sx[j], re);
            im = fmaf(cs.y, sx[j], im);
          }
          smag[f * nbp + k] = sqrtf(re * re + im * im);
        }
      }
    }
  }
  __syncthreads();
  // mel projection + log: thread -> (frame, mel), mel fastest so melT reads and the output store coalesce
  for (int i = threadIdx.x; i < LM_FR * n_mels; i += LM_THREADS) {
    const int f = i / n_mels, m = i - f * n_mels;
    const int t = t0 + f;
    if (t >= T_out) continue;
    const int k0 = mel_range[2 * m], k1 = mel_range[2 * m + 1];
    float acc = 0.f;
    for (int k = k0; k < k1; ++k) acc = fmaf(melT[(int64_t)k * n_mels + m], smag[f * nbp + k], acc);
    fast[((int64_t)b * T_out + t) * n_mels + m] = logf(acc + eps);
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
  size_t smem = (size_t)(((span + 3) & ~3) + LM_FR * n_bins_pad) * sizeof(float);
  ASF_CHECK_ARG(smem <= 227 * 1024, "asf_logmel_fwd: window/hop too large for the shared-memory staging (%zu B)", smem);
  ASF_CUDA(cudaFuncSetAttribute(logmel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((unsigned)cdiv(T_out, LM_FR), B);
  logmel_kernel<<<grid, LM_THREADS, smem, (cudaStream_t)stream>>>(wave, wave_ld, n_samples, T_valid, T_out, win, hop,
                                                                  first_off, n_bins, n_bins_pad, n_mels, eps,
                                                                  (const float2*)basisT, melT, mel_range, fast);
  ASF_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
