// SpecAugment on the GPU, batched, between the log-mel kernel and the stems.
//
// Replaces, for a whole batch in two launches, what the reference applies per sample on the CPU in its dataset
// (/root/reference/audio_slowfast/datasets/epickitchens.py:160-168):
//   spec_augment = time_mask(freq_mask(time_warp(x)))       datasets/spec_augment.py:144-191
//   time_warp -> sparse_image_warp (one control point)      spec_augment.py:9-44, datasets/sparse_image_warp.py:5-125
//   dense_image_warp / interpolate_bilinear                 sparse_image_warp.py:213-354
//   freq_mask (2 masks, width < 27), time_mask (2, < 25), filled with the CURRENT mean    spec_augment.py:47-141
//
// What the reference's one-point polyharmonic warp really computes is an affine flow along time,
//   flow_x(f, t) = d * (u . (f, t, 1) - 1e-10 K) / (u . b),   u = randn(3,3)^-1 b,  b = (F/2, x_c + d, 1),
// with x_c the spectrogram VALUE at (t = i, f = F/2) and K = 0.5 r0 ln r0, r0 = sum over the grid of (f^2 + t^2)
// (derivation and numerical pin against the reference functions: oracle/spec_augment_np.py,
// tests/test_oracle_specaug.py).  asf_specaug_params draws (i, d, the 3x3 normal block, the mask intervals) with a
// counter-based generator (Philox4x32-10: distribution parity with Python's random / torch.randn, not bit parity) and
// reduces them to 3 flow coefficients + 4 intervals per sample; asf_specaug_apply does the resampling, the running
// means and the fills for one clip per CTA with the whole T x F clip in shared memory.
#include "common.cuh"

namespace asf {

// ------------------------------------------------------------------------------------------------ Philox4x32-10
struct Philox {
  uint32_t c[4], k[2];
  __device__ __forceinline__ void round_() {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
    const uint32_t hi0 = __umulhi(M0, c[0]), lo0 = M0 * c[0];
    const uint32_t hi1 = __umulhi(M1, c[2]), lo1 = M1 * c[2];
    const uint32_t n0 = hi1 ^ c[1] ^ k[0], n1 = lo1, n2 = hi0 ^ c[3] ^ k[1], n3 = lo0;
    c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
  }
  // 4 x 32 random bits for (counter, key)
  __device__ __forceinline__ void generate(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1,
                                           uint32_t out[4]) {
    c[0] = c0; c[1] = c1; c[2] = c2; c[3] = c3;
    k[0] = k0; k[1] = k1;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
      round_();
      k[0] += 0x9E3779B9u;
      k[1] += 0xBB67AE85u;
    }
    out[0] = c[0]; out[1] = c[1]; out[2] = c[2]; out[3] = c[3];
  }
};

// stream of 32-bit draws of one sample: counter = (sample, block index, step lo, step hi), key = seed
struct DrawStream {
  uint32_t sample, step_lo, step_hi, k0, k1, blk, have;
  uint32_t buf[4];
  __device__ __forceinline__ uint32_t next() {
    if (have == 0) {
      Philox p;
      p.generate(sample, blk++, step_lo, step_hi, k0, k1, buf);
      have = 4;
    }
    return buf[4 - have--];
  }
  // Python's random.randrange(a, b): uniform integer in [a, b)
  __device__ __forceinline__ int randrange(int a, int b) {
    return a + (int)__umulhi(next(), (uint32_t)(b - a));
  }
  __device__ __forceinline__ double uniform() {  // (0, 1]
    return ((double)next() + 1.0) * (1.0 / 4294967296.0);
  }
  __device__ __forceinline__ double normal() {  // Box-Muller (one value per two draws: nine normals per clip)
    const double u1 = uniform(), u2 = uniform();
    return sqrt(-2.0 * log(u1)) * cos(6.283185307179586 * u2);
  }
};

// One thread per clip (grid-stride).  `counter`: device uint64 advanced by one per launch by the last thread, so a
// captured CUDA graph draws fresh numbers on every replay.
__global__ void __launch_bounds__(256) specaug_params_kernel(const float* __restrict__ spec, int B, int T, int F,
                                                             unsigned long long seed, unsigned long long* counter, int W,
                                                             int Fmax, int Tmax, int nfm, int ntm, double radial_k,
                                                             asf_specaug_params_t* __restrict__ out) {
  const unsigned long long step = counter ? *counter : 0ull;
  for (int b = threadIdx.x; b < B; b += blockDim.x) {
    DrawStream ds;
    ds.sample = (uint32_t)b;
    ds.step_lo = (uint32_t)step;
    ds.step_hi = (uint32_t)(step >> 32);
    ds.k0 = (uint32_t)seed;
    ds.k1 = (uint32_t)(seed >> 32);
    ds.blk = 0;
    ds.have = 0;
    asf_specaug_params_t p;
    // time_warp: spec_augment.py:35-38
    const int i = ds.randrange(W, T - W);
    const int d = ds.randrange(-W, W);
    const double xc = (double)spec[((long long)b * T + i) * F + F / 2];
    double E[9];
#pragma unroll
    for (int j = 0; j < 9; ++j) E[j] = ds.normal();
    const double bv[3] = {(double)(F / 2), xc + (double)d, 1.0};
    // u = E^-1 b up to the common factor 1 / det(E), which cancels in u / (u . b): u = adj(E) b
    const double a00 = E[4] * E[8] - E[5] * E[7], a01 = E[2] * E[7] - E[1] * E[8], a02 = E[1] * E[5] - E[2] * E[4];
    const double a10 = E[5] * E[6] - E[3] * E[8], a11 = E[0] * E[8] - E[2] * E[6], a12 = E[2] * E[3] - E[0] * E[5];
    const double a20 = E[3] * E[7] - E[4] * E[6], a21 = E[1] * E[6] - E[0] * E[7], a22 = E[0] * E[4] - E[1] * E[3];
    const double det = E[0] * a00 + E[1] * a10 + E[2] * a20;
    const double u0 = a00 * bv[0] + a01 * bv[1] + a02 * bv[2];
    const double u1 = a10 * bv[0] + a11 * bv[1] + a12 * bv[2];
    const double u2 = a20 * bv[0] + a21 * bv[1] + a22 * bv[2];
    const double den = u0 * bv[0] + u1 * bv[1] + u2 * bv[2];  // (u . b) * det
    double c0 = 0.0, c1 = 0.0, c2 = 0.0;
    if (d != 0 && den != 0.0) {
      c0 = d * u0 / den;
      c1 = d * u1 / den;
      c2 = d * (u2 - 1e-10 * radial_k * det) / den;  // the radial term K w of the spline, a constant offset
    }
    p.coef[0] = (float)c0;
    p.coef[1] = (float)c1;
    p.coef[2] = (float)c2;
    // freq_mask / time_mask: spec_augment.py:78-91, :128-141 (a zero width RETURNS, skipping the remaining masks)
#pragma unroll
    for (int which = 0; which < 2; ++which) {
      const int size = which == 0 ? F : T, mx = which == 0 ? Fmax : Tmax, cnt = which == 0 ? nfm : ntm;
      bool stop = false;
      for (int m = 0; m < ASF_SPECAUG_MAX_MASKS; ++m) {
        int z = 0, e = 0;
        if (m < cnt && !stop) {
          const int w = ds.randrange(0, mx);
          z = ds.randrange(0, size - w);
          if (w == 0) {
            stop = true;
            z = 0;
          } else {
            e = ds.randrange(z, z + w);
          }
        }
        if (which == 0) {
          p.fmask[m][0] = z;
          p.fmask[m][1] = e;
        } else {
          p.tmask[m][0] = z;
          p.tmask[m][1] = e;
        }
      }
    }
    out[b] = p;
  }
  __syncthreads();
  if (threadIdx.x == 0 && counter) *counter = step + 1ull;
}

// mean of the clip held in shared memory: per-thread strided partial sums, xor-shuffle tree, warp partials added in
// warp order in double -- a fixed order, so the augmented batch is bit-reproducible for given parameters.
__device__ __forceinline__ float clip_mean(const float* s, int n, double* warp_part, float* bcast) {
  float acc = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) acc += s[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) warp_part[threadIdx.x >> 5] = (double)acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += warp_part[w];
    *bcast = (float)(t / (double)n);
  }
  __syncthreads();
  return *bcast;
}

// One CTA per clip; the clip ([T][F] fp32, F fastest: the layout of the extractor and of the network input) lives in
// shared memory from the warp to the last fill.
//   warp:  out[t][f] = lerp(in[x0][f], in[x0 + 1][f], a),  q = t - flow(f, t),  x0 = clamp(floor(q), 0, T - 2),
//          a = clamp(q - x0, 0, 1)                          (interpolate_bilinear, sparse_image_warp.py:293-354; the
//          frequency coordinate is never displaced, so the bilinear stencil degenerates to two taps)
//   masks: rows f in [z, e) of every frame, then frames t in [z, e), each filled with the mean of the clip as it is
//          at that moment (spec_augment.py:88-89, :138-139)
__global__ void __launch_bounds__(1024) specaug_apply_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                             const asf_specaug_params_t* __restrict__ params, int T,
                                                             int F) {
  extern __shared__ __align__(16) float s_clip[];
  __shared__ double warp_part[32];
  __shared__ float bcast;
  pdl_trigger();
  pdl_wait();
  const int b = blockIdx.x, n = T * F;
  const asf_specaug_params_t p = params[b];
  const float* src = in + (long long)b * n;
  const float c0 = p.coef[0], c1 = p.coef[1], c2 = p.coef[2];
  const bool warp = c0 != 0.f || c1 != 0.f || c2 != 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int t = i / F, f = i - t * F;
    float v;
    if (warp) {
      const float flow = c0 * (float)f + c1 * (float)t + c2;
      float q = (float)t - flow;
      if (!(q == q)) q = 0.f;
      const float x0 = fminf(fmaxf(floorf(q), 0.f), (float)(T - 2));
      const float a = fminf(fmaxf(q - x0, 0.f), 1.f);
      const int j = (int)x0;
      const float l = src[j * F + f], r = src[(j + 1) * F + f];
      v = a * (r - l) + l;
    } else {
      v = src[i];
    }
    s_clip[i] = v;
  }
  __syncthreads();
#pragma unroll 1
  for (int m = 0; m < 2 * ASF_SPECAUG_MAX_MASKS; ++m) {
    const bool freq = m < ASF_SPECAUG_MAX_MASKS;
    const int z = freq ? p.fmask[m][0] : p.tmask[m - ASF_SPECAUG_MAX_MASKS][0];
    const int e = freq ? p.fmask[m][1] : p.tmask[m - ASF_SPECAUG_MAX_MASKS][1];
    if (e <= z) continue;  // uniform across the CTA
    const float mean = clip_mean(s_clip, n, warp_part, &bcast);
    if (freq) {
      const int w = e - z;
      for (int i = threadIdx.x; i < T * w; i += blockDim.x) {
        const int t = i / w, f = z + (i - t * w);
        s_clip[t * F + f] = mean;
      }
    } else {
      for (int i = z * F + threadIdx.x; i < e * F; i += blockDim.x) s_clip[i] = mean;
    }
    __syncthreads();
  }
  float* dst = out + (long long)b * n;
  for (int i = threadIdx.x; i < n; i += blockDim.x) dst[i] = s_clip[i];
}

}  // namespace asf

using namespace asf;

extern "C" {

int asf_specaug_params(const float* spec, int B, int T, int F, unsigned long long seed, unsigned long long* counter,
                       int W, int Fmax, int Tmax, int num_freq_masks, int num_time_masks, asf_specaug_params_t* params,
                       void* stream) {
  ASF_CHECK_ARG(spec && params && B > 0 && T > 2 * W && F > 1 && W > 0, "asf_specaug_params: bad arguments");
  ASF_CHECK_ARG(Fmax > 0 && Fmax <= F && Tmax > 0 && Tmax <= T, "asf_specaug_params: mask widths exceed the clip");
  ASF_CHECK_ARG(num_freq_masks >= 0 && num_freq_masks <= ASF_SPECAUG_MAX_MASKS && num_time_masks >= 0 &&
                    num_time_masks <= ASF_SPECAUG_MAX_MASKS,
                "asf_specaug_params: at most %d masks of a kind", ASF_SPECAUG_MAX_MASKS);
  // K = phi_2(r0), r0 = sum over the F x T grid of (f^2 + t^2): the constant radial term of the reference's spline
  const double Fd = F, Td = T;
  const double r0 = Td * (Fd - 1) * Fd * (2 * Fd - 1) / 6.0 + Fd * (Td - 1) * Td * (2 * Td - 1) / 6.0;
  const double radial_k = 0.5 * r0 * log(r0);
  specaug_params_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(spec, B, T, F, seed, counter, W, Fmax, Tmax, num_freq_masks,
                                                           num_time_masks, radial_k, params);
  ASF_LAUNCH_CHECK();
  return 0;
}

int asf_specaug_apply(const float* in, float* out, const asf_specaug_params_t* params, int B, int T, int F,
                      void* stream) {
  ASF_CHECK_ARG(in && out && params && B > 0 && T >= 2 && F >= 1, "asf_specaug_apply: bad arguments");
  ASF_CHECK_ARG(in != out, "asf_specaug_apply: the warp gathers along time, in and out must not alias");
  const size_t smem = sizeof(float) * (size_t)T * F;
  if (smem > 220 * 1024) {
    set_error("asf_specaug_apply: a %d x %d clip does not fit in shared memory (%zu bytes)", T, F, smem);
    return ASF_E_UNSUPPORTED;
  }
  static size_t smem_set = 48 * 1024;
  if (smem > smem_set) {
    ASF_CUDA(cudaFuncSetAttribute(specaug_apply_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    smem_set = smem;
  }
  ASF_CUDA(launch_pdl(specaug_apply_kernel, dim3((unsigned)B), dim3(1024), smem, (cudaStream_t)stream, in, out, params, T, F));
  ASF_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
