// CUDA-core direct convolution (any geometry, any dtype mix).  Two uses:
//   * the stems (C_in = 1, 7 or 35 taps: not a tensor-core shape, pure bandwidth) and their wgrad;
//   * the fp32 parity mode / bring-up cross-check of the tcgen05 tap-GEMM path.
// Thread mapping keeps the output channel as the fastest index so stores and weight-row reads coalesce.
#include "common.cuh"

namespace asf {

// accumulator type: fp32 inputs (the parity mode, an accuracy reference) accumulate in double, bf16 inputs in fp32
template <typename T> struct AccT { typedef float type; };
template <> struct AccT<float> { typedef double type; };

template <typename T>
__device__ __forceinline__ float apply_epilogue(float acc, int64_t row, int n, const asf_epilogue_t& e, T* out,
                                                int64_t ld_out) {
  float v = acc;
  if (e.scale) v *= e.scale[n];
  if (e.shift) v += e.shift[n];
  if (e.res) v += ldf((const T*)e.res + row * e.ld_res + n);
  if (e.accumulate) v += ldf(out + row * ld_out + n);
  if (e.relu) v = fmaxf(v, 0.f);
  if (e.mask && !(ldf((const T*)e.mask + row * e.ld_mask + n) > 0.f)) v = 0.f;
  return v;
}

template <typename TI, typename TW, typename TO>
__global__ void __launch_bounds__(256) conv_fwd_simt_kernel(const TI* __restrict__ in, const TW* __restrict__ w,
                                                            TO* __restrict__ out, asf_conv_geom_t g,
                                                            asf_epilogue_t e) {
  const int64_t total = (int64_t)g.B * g.Tout * g.Fout * g.Cout;
  const int taps = g.kt * g.kf;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int n = (int)(i % g.Cout);
    int64_t row = i / g.Cout;
    int fo = (int)(row % g.Fout);
    int64_t r2 = row / g.Fout;
    int to = (int)(r2 % g.Tout);
    int b = (int)(r2 / g.Tout);
    typename AccT<TI>::type acc = 0;
    const TW* wn = w + (int64_t)n * taps * g.Cin;
    for (int it = 0; it < g.kt; ++it) {
      int t = to * g.st + it - g.pt;
      if (t < 0 || t >= g.Tin) continue;
      for (int jf = 0; jf < g.kf; ++jf) {
        int f = fo * g.sf + jf - g.pf;
        if (f < 0 || f >= g.Fin) continue;
        const TI* ip = in + (((int64_t)b * g.Tin + t) * g.Fin + f) * g.ld_in;
        const TW* wp = wn + (int64_t)(it * g.kf + jf) * g.Cin;
        for (int c = 0; c < g.Cin; ++c) acc += (typename AccT<TI>::type)ldf(ip + c) * (typename AccT<TI>::type)ldf(wp + c);
      }
    }
    float v = apply_epilogue<TO>((float)acc, row, n, e, out, g.ld_out);
    stf(out + row * g.ld_out + n, v);
  }
}

// din[b,ti,fi,c] = sum over (tap, n) with to*st + it - pt == ti of dout[b,to,fo,n] * w[n][tap][c]
template <typename TI, typename TW, typename TO>
__global__ void __launch_bounds__(256) conv_dgrad_simt_kernel(const TI* __restrict__ dout, const TW* __restrict__ w,
                                                              TO* __restrict__ din, asf_conv_geom_t g,
                                                              asf_epilogue_t e) {
  const int64_t total = (int64_t)g.B * g.Tin * g.Fin * g.Cin;
  const int taps = g.kt * g.kf;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int c = (int)(i % g.Cin);
    int64_t row = i / g.Cin;
    int fi = (int)(row % g.Fin);
    int64_t r2 = row / g.Fin;
    int ti = (int)(r2 % g.Tin);
    int b = (int)(r2 / g.Tin);
    typename AccT<TI>::type acc = 0;
    for (int it = 0; it < g.kt; ++it) {
      int tt = ti + g.pt - it;
      if (tt < 0 || tt % g.st != 0) continue;
      int to = tt / g.st;
      if (to >= g.Tout) continue;
      for (int jf = 0; jf < g.kf; ++jf) {
        int ff = fi + g.pf - jf;
        if (ff < 0 || ff % g.sf != 0) continue;
        int fo = ff / g.sf;
        if (fo >= g.Fout) continue;
        const TI* dp = dout + (((int64_t)b * g.Tout + to) * g.Fout + fo) * g.ld_out;
        const TW* wp = w + (int64_t)(it * g.kf + jf) * g.Cin + c;
        for (int n = 0; n < g.Cout; ++n)
          acc += (typename AccT<TI>::type)ldf(dp + n) * (typename AccT<TI>::type)ldf(wp + (int64_t)n * taps * g.Cin);
      }
    }
    float v = apply_epilogue<TO>((float)acc, row, c, e, din, g.ld_in);
    stf(din + row * g.ld_in + c, v);
  }
}

// One block per (n, tap) pair; threads over c; the block walks a slice of the output rows (gridDim.y slices) and STORES
// its partial sums to ws[slice][n][tap][c]; asf_unpack_wgrad adds the slices in order (no atomics).
template <typename TI, typename TD>
__global__ void __launch_bounds__(128) conv_wgrad_simt_kernel(const TI* __restrict__ in, const TD* __restrict__ dout,
                                                              float* __restrict__ ws, asf_conv_geom_t g) {
  const int taps = g.kt * g.kf;
  const int n = blockIdx.x / taps, tap = blockIdx.x % taps;
  const int it = tap / g.kf, jf = tap % g.kf;
  const int64_t rows = (int64_t)g.B * g.Tout * g.Fout;
  const int64_t per = (rows + gridDim.y - 1) / gridDim.y;
  const int64_t r0 = (int64_t)blockIdx.y * per;
  const int64_t r1 = (r0 + per < rows) ? r0 + per : rows;
  for (int c0 = 0; c0 < g.Cin; c0 += blockDim.x) {
    int c = c0 + threadIdx.x;
    typename AccT<TI>::type acc = 0;
    for (int64_t row = r0; row < r1; ++row) {
      int fo = (int)(row % g.Fout);
      int64_t r2 = row / g.Fout;
      int to = (int)(r2 % g.Tout);
      int b = (int)(r2 / g.Tout);
      int t = to * g.st + it - g.pt, f = fo * g.sf + jf - g.pf;
      if (t < 0 || t >= g.Tin || f < 0 || f >= g.Fin) continue;
      typename AccT<TI>::type d = ldf(dout + row * g.ld_out + n);
      if (c < g.Cin) acc += d * (typename AccT<TI>::type)ldf(in + (((int64_t)b * g.Tin + t) * g.Fin + f) * g.ld_in + c);
    }
    if (c < g.Cin) ws[(((int64_t)blockIdx.y * g.Cout + n) * taps + tap) * g.Cin + c] = (float)acc;
  }
}

static int wgrad_simt_ysplit(const asf_conv_geom_t* g) {
  const int taps = g->kt * g->kf;
  const int64_t rows = (int64_t)g->B * g->Tout * g->Fout;
  const int64_t blocks_xy = (int64_t)g->Cout * taps;
  int ysplit = (int)(148 * 8 / blocks_xy);
  if (ysplit < 1) ysplit = 1;
  if (ysplit > 256) ysplit = 256;
  if ((int64_t)ysplit > rows) ysplit = (int)rows;
  return ysplit;
}

static int check_geom(const asf_conv_geom_t* g) {
  ASF_CHECK_ARG(g && g->B > 0 && g->Cin > 0 && g->Cout > 0 && g->kt > 0 && g->kf > 0 && g->st > 0 && g->sf > 0,
                "conv geometry: non-positive size");
  ASF_CHECK_ARG(g->Tout == (g->Tin + 2 * g->pt - g->kt) / g->st + 1 && g->Fout == (g->Fin + 2 * g->pf - g->kf) / g->sf + 1,
                "conv geometry: Tout/Fout inconsistent with Tin/Fin, kernel, stride, padding");
  ASF_CHECK_ARG(g->ld_in >= g->Cin && g->ld_out >= g->Cout, "conv geometry: row pitch smaller than channels");
  return 0;
}

static inline unsigned grid_for(int64_t total) {
  int64_t gsz = cdiv(total, 256), cap = 148 * 32;
  return (unsigned)(gsz > cap ? cap : (gsz < 1 ? 1 : gsz));
}

}  // namespace asf

using namespace asf;

#define ASF_DISPATCH3(a, TA, b, TB, c, TC, ...) \
  ASF_DISPATCH_DTYPE(a, TA, ASF_DISPATCH_DTYPE(b, TB, ASF_DISPATCH_DTYPE(c, TC, __VA_ARGS__)))

extern "C" {

int asf_conv_fwd_simt(const void* in, int in_dtype, const void* w, int w_dtype, void* out, int out_dtype,
                      const asf_conv_geom_t* g, const asf_epilogue_t* epi, void* stream) {
  ASF_CHECK_ARG(in && w && out && epi, "asf_conv_fwd_simt: null pointer");
  ASF_CHECK_ARG(epi->stats == nullptr, "asf_conv_fwd_simt: statistics are not produced here (use asf_channel_stats)");
  int rc = check_geom(g);
  if (rc) return rc;
  int64_t total = (int64_t)g->B * g->Tout * g->Fout * g->Cout;
  // the stems: C_in = 1, fp32 samples and weights, bf16 output, no fused epilogue -> dedicated vector kernel
  if (g->Cin == 1 && in_dtype == ASF_F32 && w_dtype == ASF_F32 && out_dtype == ASF_BF16 && g->Cout % 8 == 0 &&
      g->ld_in == 1 && ld8(g->ld_out) && al16(out) && !epi->scale && !epi->shift && !epi->res && !epi->mask &&
      !epi->stats && !epi->relu && !epi->accumulate && g->kt * g->kf * g->Cout * sizeof(float) <= 48 * 1024)
    return stem_conv_fwd_bf16((const float*)in, (const float*)w, (bf16*)out, g, (cudaStream_t)stream);
  ASF_DISPATCH3(in_dtype, TI, w_dtype, TW, out_dtype, TO, {
    conv_fwd_simt_kernel<TI, TW, TO><<<grid_for(total), 256, 0, (cudaStream_t)stream>>>((const TI*)in, (const TW*)w,
                                                                                     (TO*)out, *g, *epi);
    ASF_LAUNCH_CHECK();
    return 0;
  });
}

int asf_conv_dgrad_simt(const void* dout, int dout_dtype, const void* w, int w_dtype, void* din, int din_dtype,
                        const asf_conv_geom_t* g, const asf_epilogue_t* epi, void* stream) {
  ASF_CHECK_ARG(dout && w && din && epi, "asf_conv_dgrad_simt: null pointer");
  int rc = check_geom(g);
  if (rc) return rc;
  ASF_CHECK_ARG(epi->stats == nullptr, "asf_conv_dgrad_simt: stats not supported");
  int64_t total = (int64_t)g->B * g->Tin * g->Fin * g->Cin;
  ASF_DISPATCH3(dout_dtype, TI, w_dtype, TW, din_dtype, TO, {
    conv_dgrad_simt_kernel<TI, TW, TO><<<grid_for(total), 256, 0, (cudaStream_t)stream>>>(
        (const TI*)dout, (const TW*)w, (TO*)din, *g, *epi);
    ASF_LAUNCH_CHECK();
    return 0;
  });
}

int asf_conv_wgrad_simt_ksplit(const asf_conv_geom_t* g) {
  int rc = check_geom(g);
  if (rc) return rc;
  return wgrad_simt_ysplit(g);
}

int asf_conv_wgrad_simt(const void* in, int in_dtype, const void* dout, int dout_dtype, float* dw_oihw,
                        const asf_conv_geom_t* g, const float* rscale, int accumulate, float* ws, void* stream) {
  ASF_CHECK_ARG(in && dout && dw_oihw && ws, "asf_conv_wgrad_simt: null pointer");
  int rc = check_geom(g);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  int taps = g->kt * g->kf;
  const int ysplit = wgrad_simt_ysplit(g);
  dim3 grid((unsigned)((int64_t)g->Cout * taps), ysplit);
  int threads = g->Cin >= 128 ? 128 : (g->Cin >= 64 ? 64 : 32);
  ASF_DISPATCH_DTYPE(in_dtype, TI, ASF_DISPATCH_DTYPE(dout_dtype, TD, {
    conv_wgrad_simt_kernel<TI, TD><<<grid, threads, 0, st>>>((const TI*)in, (const TD*)dout, ws, *g);
    ASF_LAUNCH_CHECK();
    return asf_unpack_wgrad(ws, dw_oihw, g->Cout, g->Cin, taps, rscale, accumulate, ysplit, stream);
  }));
}

}  // extern "C"
