// BatchNorm statistics / apply / backward, pools, head helpers, weight (re)packing.
// All kernels are HBM-bound streaming passes: channel index is the fastest-varying thread index so every
// warp touches contiguous addresses; reductions go warp/block-level first, every block stores its partial sums and
// the last block to arrive adds them in a fixed order (stats_reduce.cuh): no floating-point atomics.
#include <stdarg.h>
#include <stdlib.h>

#include <mutex>

#include "common.cuh"
#include "stats_reduce.cuh"

namespace asf {

static thread_local char g_err[512] = "";
static long long g_launches = 0;

void count_launch() { __atomic_add_fetch(&g_launches, 1, __ATOMIC_RELAXED); }
bool pdl_enabled() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("ASF_PDL");
    on = (e && e[0] == '0') ? 0 : 1;
  }
  return on == 1;
}

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

// ----------------------------------------------------------------------------------------------- stats
// Scalar path (fp32 parity mode and ragged layouts).  block = 256 threads = RP row-lanes x CB channels (CB = min(C,256));
// grid.x over row chunks, grid.y over channel blocks.  Each thread streams rows with stride RP accumulating in DOUBLE,
// the block folds its row-lanes in a fixed order and stores one partial row (doubles) into the workspace; the last CTA
// of a channel block (ticket) adds the partial rows in ascending order.  No atomics on data: bit-reproducible.
template <typename T, bool kBwd>
__global__ void __launch_bounds__(256) channel_reduce_kernel(const T* __restrict__ a, int64_t ld_a,
                                                             const T* __restrict__ x, int64_t ld_x, int64_t rows,
                                                             int C, int CB, int64_t rows_per_block,
                                                             const float* __restrict__ mean,
                                                             const float* __restrict__ invstd, double* out, double* ws,
                                                             unsigned* ticket) {
  __shared__ double s0[256], s1[256];
  __shared__ unsigned s_flag;
  const int RP = 256 / CB;
  const int cl = threadIdx.x % CB, rl = threadIdx.x / CB;
  const int c = blockIdx.y * CB + cl;
  double acc0 = 0.0, acc1 = 0.0;
  if (rl < RP && c < C) {
    const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
    int64_t r1 = r0 + rows_per_block;
    if (r1 > rows) r1 = rows;
    float mu = 0.f, is = 1.f;
    if (kBwd) {
      mu = mean[c];
      is = invstd[c];
    }
    for (int64_t r = r0 + rl; r < r1; r += RP) {
      const float v = ldf(a + r * ld_a + c);
      if (kBwd) {
        const float xh = (ldf(x + r * ld_x + c) - mu) * is;
        acc0 += (double)v;
        acc1 += (double)v * (double)xh;
      } else {
        acc0 += (double)v;
        acc1 += (double)v * (double)v;
      }
    }
  }
  s0[threadIdx.x] = acc0;
  s1[threadIdx.x] = acc1;
  __syncthreads();
  // partial row of this CTA: ws[(blockIdx.y * gridDim.x + blockIdx.x)][2][CB]
  double* my = ws + ((int64_t)blockIdx.y * gridDim.x + blockIdx.x) * 2 * CB;
  if (threadIdx.x < CB) {
    double t0 = 0.0, t1 = 0.0;
    for (int k = 0; k < RP; ++k) {
      t0 += s0[threadIdx.x + k * CB];
      t1 += s1[threadIdx.x + k * CB];
    }
    my[threadIdx.x] = t0;
    my[CB + threadIdx.x] = t1;
  }
  if (!stats_last_cta(ticket + blockIdx.y, gridDim.x, &s_flag)) return;
  const double* col = ws + (int64_t)blockIdx.y * gridDim.x * 2 * CB;
  const int P = (int)gridDim.x, c0 = blockIdx.y * CB;
  stats_ordered_reduce_f64(col, P, 2 * CB, 2 * CB, [&](int s, double v) {
    const int half = s / CB, cc = c0 + (s - half * CB);
    if (cc < C) out[half * C + cc] = v;
  });
}

constexpr int SCALAR_REDUCE_MAX_X = 64;  // row chunks of the scalar path (partial rows per channel block)

template <typename T, bool kBwd>
static int launch_channel_reduce(const T* a, int64_t ld_a, const T* x, int64_t ld_x, int64_t rows, int C,
                                 const float* mean, const float* invstd, double* out, float* ws, unsigned* ticket,
                                 cudaStream_t st) {
  int CB = C < 256 ? C : 256;
  // CB must divide 256 for the row-lane layout; otherwise fall back to one row-lane of 256 channels
  if (256 % CB != 0) CB = 256;
  int cblocks = (int)cdiv(C, CB);
  ASF_CHECK_ARG(cblocks <= ASF_STATS_TICKETS, "channel reduce: %d channel blocks exceed the %d tickets", cblocks, ASF_STATS_TICKETS);
  int64_t target_blocks = SCALAR_REDUCE_MAX_X;
  int64_t rpb = cdiv(rows, target_blocks);
  int RP = 256 / CB;
  if (rpb < RP * 4) rpb = RP * 4;
  dim3 grid((unsigned)cdiv(rows, rpb), cblocks);
  // partial rows are doubles: grid.x * cblocks * 2 * CB doubles <= 64 * 2 * (C + 255) * 2 floats < ASF_STATS_WS_FLOATS(C)
  ASF_CHECK_ARG(((uintptr_t)ws & 7) == 0, "channel reduce: workspace must be 8-byte aligned");
  channel_reduce_kernel<T, kBwd><<<grid, 256, 0, st>>>(a, ld_a, x, ld_x, rows, C, CB, rpb, mean, invstd, out, (double*)ws, ticket);
  ASF_LAUNCH_CHECK();
  return 0;
}

__global__ void bn_finalize_kernel(const double* __restrict__ stats, int C, double count,
                                   const float* __restrict__ gamma, const float* __restrict__ beta, float eps,
                                   float momentum, float* running_mean, float* running_var,
                                   int64_t* num_batches_tracked, float* scale, float* shift, float* save_mean,
                                   float* save_invstd) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c == 0 && num_batches_tracked) *num_batches_tracked += 1;
  if (c >= C) return;
  double mean = stats[c] / count;
  double var = stats[C + c] / count - mean * mean;
  if (var < 0.0) var = 0.0;
  float invstd = (float)(1.0 / sqrt(var + (double)eps));
  float g = gamma ? gamma[c] : 1.f, b = beta ? beta[c] : 0.f;
  float sc = g * invstd;
  scale[c] = sc;
  shift[c] = b - (float)mean * sc;
  if (save_mean) save_mean[c] = (float)mean;
  if (save_invstd) save_invstd[c] = invstd;
  if (running_mean) {
    double unbiased = count > 1.0 ? var * count / (count - 1.0) : var;
    running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * (float)mean;
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)unbiased;
  }
}

__global__ void bn_fold_kernel(const float* __restrict__ gamma, const float* __restrict__ beta,
                               const float* __restrict__ rm, const float* __restrict__ rv, float eps, int C,
                               float* scale, float* shift) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  float sc = gamma[c] * rsqrtf(rv[c] + eps);
  // rsqrtf is 2 ulp; use the exact form to match torch's 1/sqrt(var+eps)
  sc = gamma[c] / sqrtf(rv[c] + eps);
  scale[c] = sc;
  shift[c] = beta[c] - rm[c] * sc;
}

template <typename T>
__global__ void bn_apply_kernel(const T* __restrict__ x, int64_t ld_x, const T* __restrict__ res, int64_t ld_res,
                                T* __restrict__ y, int64_t ld_y, int64_t rows, int C,
                                const float* __restrict__ scale, const float* __restrict__ shift, int relu) {
  int64_t total = rows * C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = i / C;
    int c = (int)(i - r * C);
    float v = ldf(x + r * ld_x + c) * scale[c] + shift[c];
    if (res) v += ldf(res + r * ld_res + c);
    if (relu) v = fmaxf(v, 0.f);
    stf(y + r * ld_y + c, v);
  }
}

template <typename T>
__global__ void bn_bwd_apply_kernel(const T* __restrict__ dz, int64_t ld_dz, const T* __restrict__ x, int64_t ld_x,
                                    T* __restrict__ dx, int64_t ld_dx, int64_t rows, int C,
                                    const float* __restrict__ gamma, const float* __restrict__ mean,
                                    const float* __restrict__ invstd, const double* __restrict__ red, double inv_count) {
  int64_t total = rows * C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = i / C;
    int c = (int)(i - r * C);
    float is = invstd[c];
    float xh = (ldf(x + r * ld_x + c) - mean[c]) * is;
    float g = gamma ? gamma[c] : 1.f;
    const float k0 = (float)(red[c] * inv_count), k1 = (float)(red[C + c] * inv_count);
    float v = g * is * (ldf(dz + r * ld_dz + c) - k0 - xh * k1);
    stf(dx + r * ld_dx + c, v);
  }
}

__global__ void copy_dgamma_dbeta_kernel(const double* __restrict__ red, int C, float* dgamma, float* dbeta,
                                         int accumulate) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  if (dbeta) dbeta[c] = (accumulate ? dbeta[c] : 0.f) + (float)red[c];
  if (dgamma) dgamma[c] = (accumulate ? dgamma[c] : 0.f) + (float)red[C + c];
}

template <typename T>
__global__ void relu_bwd_kernel(const T* __restrict__ a, int64_t ld_a, const T* __restrict__ b, int64_t ld_b,
                                const T* __restrict__ mask, int64_t ld_mask, T* __restrict__ y, int64_t ld_y,
                                int64_t rows, int C) {
  int64_t total = rows * C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = i / C;
    int c = (int)(i - r * C);
    float v = ldf(a + r * ld_a + c);
    if (b) v += ldf(b + r * ld_b + c);
    if (mask && !(ldf(mask + r * ld_mask + c) > 0.f)) v = 0.f;
    stf(y + r * ld_y + c, v);
  }
}

// ----------------------------------------------------------------------------------------------- pools
template <typename T>
__global__ void bn_relu_maxpool_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, int64_t ld_y, int B, int Tn,
                                           int F, int C, int To, int Fo, const float* __restrict__ scale,
                                           const float* __restrict__ shift) {
  int64_t total = (int64_t)B * To * Fo * C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int c = (int)(i % C);
    int64_t r = i / C;
    int fo = (int)(r % Fo);
    r /= Fo;
    int to = (int)(r % To);
    int b = (int)(r / To);
    float sc = scale[c], sh = shift[c];
    float m = -INFINITY;
    for (int dt = -1; dt <= 1; ++dt) {
      int t = 2 * to + dt;
      if (t < 0 || t >= Tn) continue;
      for (int df = -1; df <= 1; ++df) {
        int f = 2 * fo + df;
        if (f < 0 || f >= F) continue;
        float v = fmaxf(ldf(x + (((int64_t)b * Tn + t) * F + f) * C + c) * sc + sh, 0.f);
        // the pooled activation is what the stored dtype can hold
        m = fmaxf(m, v);
      }
    }
    stf(y + (i / C) * ld_y + c, m);
  }
}

template <typename T>
__device__ __forceinline__ bool window_argmax_is(const T* __restrict__ x, int b, int Tn, int F, int C, int c, int to,
                                                 int fo, float sc, float sh, int t_self, int f_self) {
  float m = -INFINITY;
  int at = -1, af = -1;
  for (int dt = -1; dt <= 1; ++dt) {
    int t = 2 * to + dt;
    if (t < 0 || t >= Tn) continue;
    for (int df = -1; df <= 1; ++df) {
      int f = 2 * fo + df;
      if (f < 0 || f >= F) continue;
      float v = fmaxf(ldf(x + (((int64_t)b * Tn + t) * F + f) * C + c) * sc + sh, 0.f);
      if (v > m) {  // strict: first occurrence in (t, f) scan order wins, as in torch
        m = v;
        at = t;
        af = f;
      }
    }
  }
  return at == t_self && af == f_self;
}

template <typename T>
__global__ void bn_relu_maxpool_bwd_kernel(const T* __restrict__ x, const T* __restrict__ dy, int64_t ld_dy,
                                           T* __restrict__ dz, int B, int Tn, int F, int C, int To, int Fo,
                                           const float* __restrict__ scale, const float* __restrict__ shift) {
  int64_t total = (int64_t)B * Tn * F * C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int c = (int)(i % C);
    int64_t r = i / C;
    int f = (int)(r % F);
    r /= F;
    int t = (int)(r % Tn);
    int b = (int)(r / Tn);
    float sc = scale[c], sh = shift[c];
    float a = ldf(x + i) * sc + sh;
    float g = 0.f;
    if (a > 0.f) {
      int to0 = t >> 1, to1 = (t + 1) >> 1;  // windows containing t: ceil((t-1)/2) .. floor((t+1)/2)
      int fo0 = f >> 1, fo1 = (f + 1) >> 1;
      if (t & 1) to0 = (t - 1) >> 1; else to1 = to0;
      if (f & 1) fo0 = (f - 1) >> 1; else fo1 = fo0;
      for (int to = to0; to <= to1; ++to) {
        if (to >= To) continue;
        for (int fo = fo0; fo <= fo1; ++fo) {
          if (fo >= Fo) continue;
          if (window_argmax_is(x, b, Tn, F, C, c, to, fo, sc, sh, t, f))
            g += ldf(dy + (((int64_t)b * To + to) * Fo + fo) * ld_dy + c);
        }
      }
    }
    stf(dz + i, g);
  }
}

template <typename T>
__global__ void avgpool_fwd_kernel(const T* __restrict__ x, int B, int P, int C, float* __restrict__ feat,
                                   int64_t ld_feat, int c_off) {
  // block (32 channels x 8 position-lanes): coalesced along channels, position-lanes folded in smem
  __shared__ float s[8][33];
  int c = blockIdx.x * 32 + threadIdx.x;
  int b = blockIdx.y;
  float acc = 0.f;
  if (c < C)
    for (int p = threadIdx.y; p < P; p += 8) acc += ldf(x + ((int64_t)b * P + p) * C + c);
  s[threadIdx.y][threadIdx.x] = acc;
  __syncthreads();
  if (threadIdx.y == 0 && c < C) {
    float t = 0.f;
    for (int k = 0; k < 8; ++k) t += s[k][threadIdx.x];
    feat[(int64_t)b * ld_feat + c_off + c] = t / (float)P;
  }
}

template <typename T>
__global__ void avgpool_bwd_kernel(const float* __restrict__ dfeat, int64_t ld_feat, int c_off,
                                   const T* __restrict__ mask, T* __restrict__ dx, int B, int P, int C) {
  int64_t total = (int64_t)B * P * C;
  float inv = 1.f / (float)P;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int c = (int)(i % C);
    int b = (int)(i / ((int64_t)P * C));
    float v = dfeat[(int64_t)b * ld_feat + c_off + c] * inv;
    if (mask && !(ldf(mask + i) > 0.f)) v = 0.f;
    stf(dx + i, v);
  }
}

// ----------------------------------------------------------------------------------------------- head
// tb = 1: both operands contiguous along K -> one warp per output, lanes stride K (coalesced).
__global__ void sgemm_nt_kernel(const float* __restrict__ A, const float* __restrict__ Bm, float* __restrict__ C,
                                int M, int N, int K, int ta, int64_t lda, int64_t ldb, int64_t ldc, float alpha,
                                float beta, const float* __restrict__ bias) {
  int warp = (int)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (warp >= M * N) return;
  int m = warp / N, n = warp - m * N;
  float acc = 0.f;
  for (int k = lane; k < K; k += 32) {
    float a = ta ? A[(int64_t)k * lda + m] : A[(int64_t)m * lda + k];
    acc += a * Bm[(int64_t)n * ldb + k];
  }
  acc = warp_sum(acc);
  if (lane == 0) {
    float v = alpha * acc + (bias ? bias[n] : 0.f);
    if (beta != 0.f) v += beta * C[(int64_t)m * ldc + n];
    C[(int64_t)m * ldc + n] = v;
  }
}
// tb = 0: B row-major [K][N] -> one thread per output, n fastest (coalesced on B and C, A broadcast).
__global__ void sgemm_nn_kernel(const float* __restrict__ A, const float* __restrict__ Bm, float* __restrict__ C,
                                int M, int N, int K, int ta, int64_t lda, int64_t ldb, int64_t ldc, float alpha,
                                float beta, const float* __restrict__ bias) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)M * N) return;
  int m = (int)(i / N), n = (int)(i - (int64_t)m * N);
  float acc = 0.f;
  for (int k = 0; k < K; ++k) {
    float a = ta ? A[(int64_t)k * lda + m] : A[(int64_t)m * lda + k];
    acc += a * Bm[(int64_t)k * ldb + n];
  }
  float v = alpha * acc + (bias ? bias[n] : 0.f);
  if (beta != 0.f) v += beta * C[(int64_t)m * ldc + n];
  C[(int64_t)m * ldc + n] = v;
}

__global__ void softmax_rows_kernel(float* x, int rows, int n, int64_t ld) {
  int warp = (int)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (warp >= rows) return;
  float* p = x + (int64_t)warp * ld;
  float m = -INFINITY;
  for (int i = lane; i < n; i += 32) m = fmaxf(m, p[i]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  float s = 0.f;
  for (int i = lane; i < n; i += 32) s += expf(p[i] - m);
  s = warp_sum(s);
  float inv = 1.f / s;
  for (int i = lane; i < n; i += 32) p[i] = expf(p[i] - m) * inv;
}

__global__ void scale_mask_kernel(const float* __restrict__ x, const float* __restrict__ mask, float* y, int64_t n,
                                  float inv_keep) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) y[i] = x[i] * mask[i] * inv_keep;
}

__global__ void gather_rows_kernel(const float* __restrict__ src, float* __restrict__ dst,
                                   const int32_t* __restrict__ idx, int B, int T_src, int n_idx, int F) {
  int64_t total = (int64_t)B * n_idx * F;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int f = (int)(i % F);
    int64_t r = i / F;
    int j = (int)(r % n_idx);
    int b = (int)(r / n_idx);
    dst[i] = src[((int64_t)b * T_src + idx[j]) * F + f];
  }
}

// ----------------------------------------------------------------------------------------------- weights
template <typename T>
__global__ void pack_weight_kernel(const float* __restrict__ w, T* __restrict__ p, int O, int I, int taps,
                                   const float* __restrict__ rscale, int transpose) {
  int64_t total = (int64_t)O * I * taps;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    // i indexes the DESTINATION so writes are coalesced
    int o, c, tp;
    if (!transpose) {  // [O][taps][I]
      c = (int)(i % I);
      int64_t r = i / I;
      tp = (int)(r % taps);
      o = (int)(r / taps);
    } else {  // [I][taps][O]
      o = (int)(i % O);
      int64_t r = i / O;
      tp = (int)(r % taps);
      c = (int)(r / taps);
    }
    float v = w[((int64_t)o * I + c) * taps + tp];
    if (rscale) v *= rscale[o];
    stf(p + i, v);
  }
}

// All convs of the network in ONE launch: block b handles PACK_CHUNK destination indices of the unit whose chunk range
// contains b (binary search over the descriptor table), for both the [O][taps][I] and the [I][taps][O] copy.
constexpr int PACK_CHUNK = 2048;
constexpr int PACK_TILE = 32;      // output channels x input channels per shared-memory tile
constexpr int PACK_MAX_TAPS = 7;   // tiled path; more taps (the 5x7 stems, C_in = 1) take the element-wise loop
// A block owns the tiles t = b - chunk0, + nb, ... of its unit (nb = the unit's chunk count, as laid out by the caller).
// The OIHW master is read in whole rows (I x taps contiguous floats per output channel, scaled by the folded BN
// factor), staged in shared memory, and leaves as 64-byte runs of both packed layouts.  The first version gathered
// every destination element on its own (4-byte reads at a stride of I*taps floats for the transposed copy): 176 us
// for the 26.9 M parameters, in front of the stems of every training step.
__global__ void __launch_bounds__(256) pack_weights_multi_kernel(const asf_pack_desc_t* __restrict__ d, int n) {
  __shared__ float tile[PACK_TILE][PACK_TILE * PACK_MAX_TAPS + 1];
  int lo = 0, hi = n - 1;
  const long long b = blockIdx.x;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (d[mid].chunk0 <= b) lo = mid; else hi = mid - 1;
  }
  const asf_pack_desc_t u = d[lo];
  const long long total = (long long)u.O * u.I * u.taps;
  const int nb = (int)((total + PACK_CHUNK - 1) / PACK_CHUNK);
  const int bl = (int)(b - u.chunk0);
  const bool bf = u.dtype == ASF_BF16;
  if (u.taps > PACK_MAX_TAPS) {
    const long long i0 = (long long)bl * PACK_CHUNK;
    for (int k = threadIdx.x; k < PACK_CHUNK; k += 256) {
      const long long i = i0 + k;
      if (i >= total) break;
      {  // [O][taps][I]
        const int c = (int)(i % u.I);
        const long long r = i / u.I;
        const int tp = (int)(r % u.taps), o = (int)(r / u.taps);
        float v = u.w[((long long)o * u.I + c) * u.taps + tp];
        if (u.rscale) v *= u.rscale[o];
        if (bf) ((bf16*)u.packed)[i] = __float2bfloat16_rn(v); else ((float*)u.packed)[i] = v;
      }
      if (u.packed_t) {  // [I][taps][O]
        const int o = (int)(i % u.O);
        const long long r = i / u.O;
        const int tp = (int)(r % u.taps), c = (int)(r / u.taps);
        float v = u.w[((long long)o * u.I + c) * u.taps + tp];
        if (u.rscale) v *= u.rscale[o];
        if (bf) ((bf16*)u.packed_t)[i] = __float2bfloat16_rn(v); else ((float*)u.packed_t)[i] = v;
      }
    }
    return;
  }
  const int O = u.O, I = u.I, taps = u.taps;
  const int itl = (I + PACK_TILE - 1) / PACK_TILE, nt = ((O + PACK_TILE - 1) / PACK_TILE) * itl;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int t = bl; t < nt; t += nb) {
    const int o0 = (t / itl) * PACK_TILE, i0 = (t % itl) * PACK_TILE;
    const int ow = min(PACK_TILE, O - o0), iw = min(PACK_TILE, I - i0);
    const int seg = iw * taps;
    for (int o = warp; o < ow; o += 8) {
      const float* src = u.w + ((long long)(o0 + o) * I + i0) * taps;
      const float sc = u.rscale ? u.rscale[o0 + o] : 1.f;
      for (int k = lane; k < seg; k += 32) tile[o][k] = src[k] * sc;
    }
    __syncthreads();
    // bf16 with 4-aligned tile widths (every conv but the stems): 8-byte stores of four consecutive channels, a warp
    // covers four 64-byte runs per instruction
    const bool v4 = bf && !(iw & 3) && !(ow & 3) && !(I & 3) && !(O & 3);
    if (v4) {
      for (int idx = threadIdx.x; idx < ow * taps * 8; idx += 256) {  // [O][taps][I]: 8 lanes x 4 channels along I
        const int i = (idx & 7) * 4, r = idx >> 3;
        const int tp = r % taps, o = r / taps;
        if (i < iw) {
          const float* tr = &tile[o][i * taps + tp];
          __nv_bfloat162 lo2 = __floats2bfloat162_rn(tr[0], tr[taps]), hi2 = __floats2bfloat162_rn(tr[2 * taps], tr[3 * taps]);
          uint2 q;
          q.x = *(const uint32_t*)&lo2;
          q.y = *(const uint32_t*)&hi2;
          *(uint2*)((bf16*)u.packed + ((long long)(o0 + o) * taps + tp) * I + i0 + i) = q;
        }
      }
      if (u.packed_t) {
        for (int idx = threadIdx.x; idx < iw * taps * 8; idx += 256) {  // [I][taps][O]: 8 lanes x 4 channels along O
          const int o = (idx & 7) * 4, r = idx >> 3;
          const int tp = r % taps, i = r / taps;
          if (o < ow) {
            const int k = i * taps + tp;
            __nv_bfloat162 lo2 = __floats2bfloat162_rn(tile[o][k], tile[o + 1][k]),
                           hi2 = __floats2bfloat162_rn(tile[o + 2][k], tile[o + 3][k]);
            uint2 q;
            q.x = *(const uint32_t*)&lo2;
            q.y = *(const uint32_t*)&hi2;
            *(uint2*)((bf16*)u.packed_t + ((long long)(i0 + i) * taps + tp) * O + o0 + o) = q;
          }
        }
      }
    } else {
    for (int idx = threadIdx.x; idx < ow * taps * PACK_TILE; idx += 256) {  // [O][taps][I]: lanes along I
      const int i = idx & 31, r = idx >> 5;
      const int tp = r % taps, o = r / taps;
      if (i < iw) {
        const float v = tile[o][i * taps + tp];
        const long long dst = ((long long)(o0 + o) * taps + tp) * I + i0 + i;
        if (bf) ((bf16*)u.packed)[dst] = __float2bfloat16_rn(v); else ((float*)u.packed)[dst] = v;
      }
    }
    if (u.packed_t) {
      for (int idx = threadIdx.x; idx < iw * taps * PACK_TILE; idx += 256) {  // [I][taps][O]: lanes along O
        const int o = idx & 31, r = idx >> 5;
        const int tp = r % taps, i = r / taps;
        if (o < ow) {
          const float v = tile[o][i * taps + tp];
          const long long dst = ((long long)(i0 + i) * taps + tp) * O + o0 + o;
          if (bf) ((bf16*)u.packed_t)[dst] = __float2bfloat16_rn(v); else ((float*)u.packed_t)[dst] = v;
        }
      }
    }
    }
    __syncthreads();
  }
}

// SGD (momentum, optional nesterov, weight decay) over every parameter tensor in ONE launch; block b owns
// SGD_CHUNK elements of the tensor whose chunk range contains b.  lr is read from device memory (one float per
// parameter group) so a captured CUDA graph follows the per-iteration learning-rate schedule.
constexpr int SGD_CHUNK = 4096;
__global__ void __launch_bounds__(256) sgd_multi_kernel(const asf_sgd_desc_t* __restrict__ d, int n,
                                                        const float* __restrict__ lr_dev, float momentum,
                                                        float dampening, int nesterov, int first_step) {
  int lo = 0, hi = n - 1;
  const long long b = blockIdx.x;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (d[mid].chunk0 <= b) lo = mid; else hi = mid - 1;
  }
  const asf_sgd_desc_t u = d[lo];
  const float lr = lr_dev[u.group];
  const long long i0 = (b - u.chunk0) * SGD_CHUNK;
  const long long i1 = (i0 + SGD_CHUNK < u.n) ? i0 + SGD_CHUNK : u.n;
  const bool vec = ((((uintptr_t)u.p | (uintptr_t)u.g | (uintptr_t)u.m) & 15) == 0);
  if (vec) {
    const long long v1 = i0 + ((i1 - i0) & ~3ll);
    for (long long i = i0 + 4 * threadIdx.x; i < v1; i += 4 * 256) {
      float4 p = *(float4*)(u.p + i);
      const float4 g = *(const float4*)(u.g + i);
      float4 m = u.m ? *(float4*)(u.m + i) : make_float4(0.f, 0.f, 0.f, 0.f);
      float* pp = (float*)&p;
      const float* gg = (const float*)&g;
      float* mm = (float*)&m;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float dd = gg[j] + u.wd * pp[j];
        if (u.m) {
          mm[j] = first_step ? dd : momentum * mm[j] + (1.f - dampening) * dd;
          dd = nesterov ? dd + momentum * mm[j] : mm[j];
        }
        pp[j] -= lr * dd;
      }
      *(float4*)(u.p + i) = p;
      if (u.m) *(float4*)(u.m + i) = m;
    }
    for (long long i = v1 + threadIdx.x; i < i1; i += 256) {
      float dd = u.g[i] + u.wd * u.p[i];
      if (u.m) {
        const float mv = first_step ? dd : momentum * u.m[i] + (1.f - dampening) * dd;
        u.m[i] = mv;
        dd = nesterov ? dd + momentum * mv : mv;
      }
      u.p[i] -= lr * dd;
    }
  } else {
    for (long long i = i0 + threadIdx.x; i < i1; i += 256) {
      float dd = u.g[i] + u.wd * u.p[i];
      if (u.m) {
        const float mv = first_step ? dd : momentum * u.m[i] + (1.f - dampening) * dd;
        u.m[i] = mv;
        dd = nesterov ? dd + momentum * mv : mv;
      }
      u.p[i] -= lr * dd;
    }
  }
}

// Split-K reduction + re-layout.  dwp is [ksplit][E] with E = O*taps*I elements in (o, tap, c) order.  A block owns
// 256 / SL consecutive elements; SL "split lanes" share the ksplit partial tiles of an element (lane j adds tiles
// j, j + SL, ... in ascending order, four loads in flight), then the lanes are folded in shared memory in a fixed
// order -- bit-reproducible, and narrow layers with a deep split (8 x 8 x 3 weights, 148 partial tiles) no longer
// serialise hundreds of dependent loads in 64 threads.
__global__ void __launch_bounds__(256) unpack_wgrad_kernel(const float* __restrict__ dwp, float* __restrict__ dw,
                                                           int O, int I, int taps, const float* __restrict__ rscale,
                                                           int accumulate, int ksplit, int sl_log2) {
  __shared__ float part[256];
  pdl_trigger();
  const int SL = 1 << sl_log2, EL = 256 >> sl_log2;
  const int el = threadIdx.x & (EL - 1), sl = threadIdx.x >> (8 - sl_log2);
  const int64_t E = (int64_t)O * taps * I;
  const int64_t e = (int64_t)blockIdx.x * EL + el;
  pdl_wait();
  float v0 = 0.f, v1 = 0.f, v2 = 0.f, v3 = 0.f;
  if (e < E) {
    const float* src = dwp + e;
    int k = sl;
    for (; k + 3 * SL < ksplit; k += 4 * SL) {
      v0 += src[(int64_t)k * E];
      v1 += src[(int64_t)(k + SL) * E];
      v2 += src[(int64_t)(k + 2 * SL) * E];
      v3 += src[(int64_t)(k + 3 * SL) * E];
    }
    for (; k < ksplit; k += SL) v0 += src[(int64_t)k * E];
  }
  float v = (v0 + v1) + (v2 + v3);
  if (SL > 1) {
    part[threadIdx.x] = v;
    __syncthreads();
    if (sl == 0)
      for (int j = 1; j < SL; ++j) v += part[j * EL + el];
  }
  if (sl == 0 && e < E) {
    const int c = (int)(e % I);
    const int64_t ot = e / I;
    const int tp = (int)(ot % taps), o = (int)(ot / taps);
    if (rscale) v *= rscale[o];
    const int64_t d = ((int64_t)o * I + c) * taps + tp;
    if (accumulate) v += dw[d];
    dw[d] = v;
  }
}

// The same reduction for MANY layers in one launch (the engine defers the unpack of a backward segment: ~100 small
// launches on the weight-gradient streams become a handful).  Block b owns `chunk` consecutive elements of the layer
// whose chunk range contains b.  The 256 threads are EL element lanes (one float4 each) x SL split lanes; SL follows the
// layer's split depth -- 1 for ksplit <= 2 (the big s4 / s5 layers: every thread streams elements), 4 up to 31, 16 from
// 32 on (the narrow layers and the stems leave 100..300 partial rows of a few hundred weights: 16 lanes share them
// and the block is 256 instead of 1024 elements, so that the reduction does not sit on one CTA for 70 dependent loads).
// Split lane j adds partial rows j, j + SL, ... in ascending order, four loads in flight; the lanes are then folded in
// ascending order: bit-reproducible.
__global__ void __launch_bounds__(256) unpack_wgrad_multi_kernel(const asf_unpack_desc_t* __restrict__ d, int n) {
  __shared__ float4 part[256];
  pdl_trigger();
  int lo = 0, hi = n - 1;
  const long long b = blockIdx.x;
  pdl_wait();
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (d[mid].chunk0 <= b) lo = mid; else hi = mid - 1;
  }
  const asf_unpack_desc_t u = d[lo];
  const int sl_log2 = u.ksplit <= 2 ? 0 : (u.ksplit < 32 ? 2 : 4);
  const int SL = 1 << sl_log2, EL = 256 >> sl_log2;
  const int chunk = u.ksplit < 32 ? ASF_UNPACK_CHUNK : ASF_UNPACK_CHUNK_DEEP;
  const int el = threadIdx.x & (EL - 1), sl = threadIdx.x >> (8 - sl_log2);
  const int64_t E = (int64_t)u.O * u.taps * u.I;
  const int64_t e0 = (b - u.chunk0) * chunk;
  const int passes = chunk / (EL * 4);
#pragma unroll 1
  for (int pass = 0; pass < passes; ++pass) {
    const int64_t e = e0 + (int64_t)pass * EL * 4 + el * 4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (e < E) {
      const float* src = u.dwp + e;
      int k = sl;
      for (; k + 3 * SL < u.ksplit; k += 4 * SL) {
        const float4 a0 = *(const float4*)(src + (int64_t)k * E);
        const float4 a1 = *(const float4*)(src + (int64_t)(k + SL) * E);
        const float4 a2 = *(const float4*)(src + (int64_t)(k + 2 * SL) * E);
        const float4 a3 = *(const float4*)(src + (int64_t)(k + 3 * SL) * E);
        v.x = (((v.x + a0.x) + a1.x) + a2.x) + a3.x;
        v.y = (((v.y + a0.y) + a1.y) + a2.y) + a3.y;
        v.z = (((v.z + a0.z) + a1.z) + a2.z) + a3.z;
        v.w = (((v.w + a0.w) + a1.w) + a2.w) + a3.w;
      }
      for (; k < u.ksplit; k += SL) {
        const float4 a0 = *(const float4*)(src + (int64_t)k * E);
        v.x += a0.x; v.y += a0.y; v.z += a0.z; v.w += a0.w;
      }
    }
    if (SL > 1) {
      if (pass) __syncthreads();  // the previous pass has finished reading `part`
      part[threadIdx.x] = v;
      __syncthreads();
    }
    if (sl == 0 && e < E) {
      float r[4] = {v.x, v.y, v.z, v.w};
      for (int j = 1; j < SL; ++j) {
        const float4 q = part[j * EL + el];
        r[0] += q.x; r[1] += q.y; r[2] += q.z; r[3] += q.w;
      }
      const int c = (int)(e % u.I);  // I % 4 == 0: the four elements share (o, tap)
      const int64_t ot = e / u.I;
      const int tp = (int)(ot % u.taps), o = (int)(ot / u.taps);
      const float sc = u.rscale ? u.rscale[o] : 1.f;
      float* dst = u.dw + ((int64_t)o * u.I + c) * u.taps + tp;
      if (u.taps == 1 && (((uintptr_t)u.dw) & 15) == 0) {  // OIHW with one tap is the packed order: one 16-byte store
        float4 x = make_float4(r[0] * sc, r[1] * sc, r[2] * sc, r[3] * sc);
        if (u.accumulate) {
          const float4 y = *(const float4*)dst;
          x.x += y.x; x.y += y.y; x.z += y.z; x.w += y.w;
        }
        *(float4*)dst = x;
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float x = r[j] * sc;
          if (u.accumulate) x += dst[(int64_t)j * u.taps];
          dst[(int64_t)j * u.taps] = x;
        }
      }
    }
  }
}

// Stem weight gradient (C_in = 1), scalar path (fp32 parity mode, ragged shapes):
//   dwp[block][n][tap] = sum over the block's rows of dout[row][n] * in[row shifted by tap].
// block = 256 threads = RL row-lanes x Cout channels; each thread keeps all taps (<= 35) in registers (double
// accumulators) while it streams its rows (dout coalesced along n, the input sample is a warp broadcast); row-lanes are
// folded in shared memory in a fixed order and the block STORES its partial sums (asf_unpack_wgrad adds the blocks).
constexpr int STEM_MAX_TAPS = 35;
template <typename TD>
__global__ void __launch_bounds__(256) stem_wgrad_kernel(const float* __restrict__ in, const TD* __restrict__ dout,
                                                         float* __restrict__ dwp, asf_conv_geom_t g,
                                                         int64_t rows_per_block) {
  __shared__ double red[256];
  const int taps = g.kt * g.kf;
  const int n = threadIdx.x % g.Cout, rl = threadIdx.x / g.Cout, RL = 256 / g.Cout;
  const int64_t rows = (int64_t)g.B * g.Tout * g.Fout;
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
  const int64_t r1 = (r0 + rows_per_block < rows) ? r0 + rows_per_block : rows;
  double acc[STEM_MAX_TAPS];
#pragma unroll
  for (int i = 0; i < STEM_MAX_TAPS; ++i) acc[i] = 0.0;
  if (rl < RL) {
    for (int64_t row = r0 + rl; row < r1; row += RL) {
      const int fo = (int)(row % g.Fout);
      const int64_t r2 = row / g.Fout;
      const int to = (int)(r2 % g.Tout);
      const int b = (int)(r2 / g.Tout);
      const double d = (double)ldf(dout + row * g.ld_out + n);
      const float* ib = in + (int64_t)b * g.Tin * g.Fin;
#pragma unroll
      for (int tp = 0; tp < STEM_MAX_TAPS; ++tp) {
        if (tp < taps) {
          const int it = tp / g.kf, jf = tp - it * g.kf;
          const int t = to * g.st + it - g.pt, f = fo * g.sf + jf - g.pf;
          if (t >= 0 && t < g.Tin && f >= 0 && f < g.Fin) acc[tp] += d * (double)ib[(int64_t)t * g.Fin + f];
        }
      }
    }
  }
  float* my = dwp + (int64_t)blockIdx.x * g.Cout * taps;
#pragma unroll
  for (int tp = 0; tp < STEM_MAX_TAPS; ++tp) {
    if (tp < taps) {
      __syncthreads();
      red[threadIdx.x] = acc[tp];
      __syncthreads();
      if (threadIdx.x < g.Cout) {
        double t = 0.0;
        for (int k = 0; k < RL; ++k) t += red[threadIdx.x + k * g.Cout];
        my[(int64_t)threadIdx.x * taps + tp] = (float)t;
      }
    }
  }
}

static inline unsigned ew_grid(int64_t total) {
  int64_t g = cdiv(total, 256);
  int64_t cap = 148 * 16;
  return (unsigned)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace asf

using namespace asf;

extern "C" {

const char* asf_last_error(void) { return asf::g_err; }
int asf_version(void) { return 101; }
long long asf_launch_count(void) { return __atomic_load_n(&asf::g_launches, __ATOMIC_RELAXED); }

int asf_device_is_sm100(void) {
  int dev = 0, major = 0;  // (cudaGetDeviceProperties costs milliseconds; this is called on every public entry)
  ASF_CUDA(cudaGetDevice(&dev));
  ASF_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  return major == 10 ? 1 : 0;
}

int asf_channel_stats(const void* x, int dtype, int64_t rows, int C, int64_t ld, double* stats, float* ws,
                      uint32_t* ticket, void* stream) {
  ASF_CHECK_ARG(x && stats && ws && ticket && rows > 0 && C > 0 && ld >= C, "asf_channel_stats: bad arguments");
  if (dtype == ASF_BF16 && C % 8 == 0 && ld8(ld) && al16(x))
    return channel_reduce_v8<false>((const bf16*)x, ld, nullptr, 0, rows, C, nullptr, nullptr, stats, ws, ticket,
                                    (cudaStream_t)stream);
  ASF_DISPATCH_DTYPE(dtype, T, return (launch_channel_reduce<T, false>((const T*)x, ld, (const T*)nullptr, 0, rows, C,
                                                                      nullptr, nullptr, stats, ws, ticket,
                                                                      (cudaStream_t)stream)));
}

// ---- allocation-free variant (no counterpart in the reference) --------
// The partial-sum workspace and the arrival tickets of the reduction live in file-scope device arrays, so a caller that
// may not allocate (CUDA graph capture, serving loops) passes inputs and outputs only.  Every GPU that loads the library
// holds these ASF_NOALLOC_SCRATCH_BYTES until exit.  In HOST code the array names below are stand-ins, not device
// addresses: the device addresses are fetched once per device with cudaGetSymbolAddress and passed to the kernel.
static __device__ __align__(256) float g_noalloc_ws[ASF_STATS_WS_FLOATS(ASF_NOALLOC_MAX_CHANNELS)];
static __device__ __align__(256) uint32_t g_noalloc_ticket[ASF_STATS_TICKETS];
static_assert(sizeof(float) * ASF_STATS_WS_FLOATS(ASF_NOALLOC_MAX_CHANNELS) + sizeof(uint32_t) * ASF_STATS_TICKETS ==
                  ASF_NOALLOC_SCRATCH_BYTES, "ASF_NOALLOC_SCRATCH_BYTES in asf_b200.h is out of date");

static int noalloc_scratch(float** ws, uint32_t** ticket) {
  constexpr int MAX_DEV = 64;
  static std::mutex mu;
  static float* ws_of[MAX_DEV] = {};
  static uint32_t* tk_of[MAX_DEV] = {};
  int dev = 0;
  ASF_CUDA(cudaGetDevice(&dev));
  ASF_CHECK_ARG(dev >= 0 && dev < MAX_DEV, "asf_channel_stats_noalloc: device index %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (!ws_of[dev]) {  // once per device, with that device current
    float* w = nullptr;
    uint32_t* t = nullptr;
    ASF_CUDA(cudaGetSymbolAddress((void**)&w, g_noalloc_ws));
    ASF_CUDA(cudaGetSymbolAddress((void**)&t, g_noalloc_ticket));
    ASF_CHECK_ARG(w && t, "asf_channel_stats_noalloc: no device address for the scratch arrays");
    ws_of[dev] = w;
    tk_of[dev] = t;
  }
  *ws = ws_of[dev];
  *ticket = tk_of[dev];
  return 0;
}

int asf_channel_stats_noalloc(const void* x, int dtype, int64_t rows, int C, int64_t ld, double* stats, void* stream) {
  ASF_CHECK_ARG(x && stats && rows > 0 && C > 0 && ld >= C, "asf_channel_stats_noalloc: bad arguments");
  ASF_CHECK_ARG(C <= ASF_NOALLOC_MAX_CHANNELS,
                "asf_channel_stats_noalloc: %d channels exceed the built-in scratch (ASF_NOALLOC_MAX_CHANNELS = %d); use "
                "asf_channel_stats with a caller-owned workspace", C, ASF_NOALLOC_MAX_CHANNELS);
  float* ws = nullptr;
  uint32_t* ticket = nullptr;
  int rc = noalloc_scratch(&ws, &ticket);
  if (rc) return rc;
  // The array keeps the previous call's contents.  The partial sums are all written before they are read; the tickets
  // must start at zero (they reset themselves, but a call that was aborted mid-way would leave them counting).
  ASF_CUDA(cudaMemsetAsync(ticket, 0, sizeof(uint32_t) * ASF_STATS_TICKETS, (cudaStream_t)stream));
  return asf_channel_stats(x, dtype, rows, C, ld, stats, ws, ticket, stream);
}

int asf_bn_finalize(const double* stats, int C, double count, const float* gamma, const float* beta, float eps,
                    float momentum, float* running_mean, float* running_var, int64_t* num_batches_tracked,
                    float* scale, float* shift, float* save_mean, float* save_invstd, void* stream) {
  ASF_CHECK_ARG(stats && scale && shift && C > 0 && count > 0, "asf_bn_finalize: bad arguments");
  bn_finalize_kernel<<<(unsigned)cdiv(C, 128), 128, 0, (cudaStream_t)stream>>>(
      stats, C, count, gamma, beta, eps, momentum, running_mean, running_var, num_batches_tracked, scale, shift,
      save_mean, save_invstd);
  ASF_LAUNCH_CHECK();
  return 0;
}

int asf_bn_fold(const float* gamma, const float* beta, const float* running_mean, const float* running_var,
                float eps, int C, float* scale, float* shift, void* stream) {
  ASF_CHECK_ARG(gamma && beta && running_mean && running_var && scale && shift && C > 0, "asf_bn_fold: bad arguments");
  bn_fold_kernel<<<(unsigned)cdiv(C, 128), 128, 0, (cudaStream_t)stream>>>(gamma, beta, running_mean, running_var, eps,
                                                                          C, scale, shift);
  ASF_LAUNCH_CHECK();
  return 0;
}

int asf_bn_apply(const void* x, int64_t ld_x, const void* res, int64_t ld_res, void* y, int64_t ld_y, int dtype,
                 int64_t rows, int C, const float* scale, const float* shift, int relu, void* stream) {
  ASF_CHECK_ARG(x && y && scale && shift && rows > 0 && C > 0, "asf_bn_apply: bad arguments");
  if (dtype == ASF_BF16 && C % 8 == 0 && ld8(ld_x) && ld8(ld_y) && al16(x) && al16(y) && (!res || (ld8(ld_res) && al16(res))))
    return bn_apply_v8((const bf16*)x, ld_x, (const bf16*)res, ld_res, (bf16*)y, ld_y, rows, C, scale, shift, relu,
                       (cudaStream_t)stream);
  ASF_DISPATCH_DTYPE(dtype, T, {
    bn_apply_kernel<T><<<ew_grid(rows * C), 256, 0, (cudaStream_t)stream>>>((const T*)x, ld_x, (const T*)res, ld_res,
                                                                          (T*)y, ld_y, rows, C, scale, shift, relu);
    ASF_LAUNCH_CHECK();
    return 0;
  });
}

int asf_bn_train_fwd(const void* x, int64_t ld_x, const void* res, int64_t ld_res, void* y, int64_t ld_y, int dtype,
                     int64_t rows, int C, const double* stats, double count, const float* gamma, const float* beta,
                     float eps, float momentum, float* running_mean, float* running_var, int64_t* num_batches_tracked,
                     float* scale, float* shift, float* save_mean, float* save_invstd, int relu, void* stream) {
  ASF_CHECK_ARG(x && stats && scale && shift && rows > 0 && C > 0 && count > 0, "asf_bn_train_fwd: bad arguments");
  if (dtype == ASF_BF16 && C % 8 == 0 && ld8(ld_x) && al16(x) && (!y || (ld8(ld_y) && al16(y))) &&
      (!res || (ld8(ld_res) && al16(res))))
    return bn_train_fwd_v8((const bf16*)x, ld_x, (const bf16*)res, ld_res, (bf16*)y, ld_y, rows, C, stats, count, gamma,
                           beta, eps, momentum, running_mean, running_var, num_batches_tracked, scale, shift, save_mean,
                           save_invstd, relu, (cudaStream_t)stream);
  int rc = asf_bn_finalize(stats, C, count, gamma, beta, eps, momentum, running_mean, running_var, num_batches_tracked,
                           scale, shift, save_mean, save_invstd, stream);
  if (rc || !y) return rc;
  return asf_bn_apply(x, ld_x, res, ld_res, y, ld_y, dtype, rows, C, scale, shift, relu, stream);
}

int asf_bn_bwd_reduce(const void* dz, int64_t ld_dz, const void* x, int64_t ld_x, int dtype, int64_t rows, int C,
                      const float* mean, const float* invstd, double* red, float* ws, uint32_t* ticket, void* stream) {
  ASF_CHECK_ARG(dz && x && mean && invstd && red && ws && ticket && rows > 0 && C > 0, "asf_bn_bwd_reduce: bad arguments");
  if (dtype == ASF_BF16 && C % 8 == 0 && ld8(ld_dz) && ld8(ld_x) && al16(dz) && al16(x))
    return channel_reduce_v8<true>((const bf16*)dz, ld_dz, (const bf16*)x, ld_x, rows, C, mean, invstd, red, ws, ticket,
                                   (cudaStream_t)stream);
  ASF_DISPATCH_DTYPE(dtype, T, return (launch_channel_reduce<T, true>((const T*)dz, ld_dz, (const T*)x, ld_x, rows, C,
                                                                     mean, invstd, red, ws, ticket,
                                                                     (cudaStream_t)stream)));
}

int asf_bn_bwd_apply(const void* dz, int64_t ld_dz, const void* x, int64_t ld_x, void* dx, int64_t ld_dx, int dtype,
                     int64_t rows, int C, const float* gamma, const float* mean, const float* invstd,
                     const double* red, double count, float* dgamma, float* dbeta, int accumulate, void* stream) {
  ASF_CHECK_ARG(dz && x && dx && mean && invstd && red && rows > 0 && C > 0 && count > 0, "asf_bn_bwd_apply: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == ASF_BF16 && C % 8 == 0 && ld8(ld_dz) && ld8(ld_x) && ld8(ld_dx) && al16(dz) && al16(x) && al16(dx))
    return bn_bwd_apply_v8((const bf16*)dz, ld_dz, (const bf16*)x, ld_x, (bf16*)dx, ld_dx, rows, C, gamma, mean, invstd,
                           red, 1.0 / count, dgamma, dbeta, accumulate, st);
  if (dgamma || dbeta) {
    copy_dgamma_dbeta_kernel<<<(unsigned)cdiv(C, 128), 128, 0, st>>>(red, C, dgamma, dbeta, accumulate);
    ASF_LAUNCH_CHECK();
  }
  ASF_DISPATCH_DTYPE(dtype, T, {
    bn_bwd_apply_kernel<T><<<ew_grid(rows * C), 256, 0, st>>>((const T*)dz, ld_dz, (const T*)x, ld_x, (T*)dx, ld_dx,
                                                            rows, C, gamma, mean, invstd, red, 1.0 / count);
    ASF_LAUNCH_CHECK();
    return 0;
  });
}

int asf_relu_bwd(const void* a, int64_t ld_a, const void* b, int64_t ld_b, const void* mask, int64_t ld_mask,
                 void* y, int64_t ld_y, int dtype, int64_t rows, int C, void* stream) {
  ASF_CHECK_ARG(a && y && rows > 0 && C > 0, "asf_relu_bwd: bad arguments");
  if (dtype == ASF_BF16 && C % 8 == 0 && ld8(ld_a) && ld8(ld_y) && al16(a) && al16(y) && (!b || (ld8(ld_b) && al16(b))) &&
      (!mask || (ld8(ld_mask) && al16(mask))))
    return relu_bwd_v8((const bf16*)a, ld_a, (const bf16*)b, ld_b, (const bf16*)mask, ld_mask, (bf16*)y, ld_y, rows, C,
                       (cudaStream_t)stream);
  ASF_DISPATCH_DTYPE(dtype, T, {
    relu_bwd_kernel<T><<<ew_grid(rows * C), 256, 0, (cudaStream_t)stream>>>(
        (const T*)a, ld_a, (const T*)b, ld_b, (const T*)mask, ld_mask, (T*)y, ld_y, rows, C);
    ASF_LAUNCH_CHECK();
    return 0;
  });
}

int asf_bn_relu_maxpool_fwd(const void* x, void* y, int64_t ld_y, int dtype, int B, int T, int F, int C,
                            const float* scale, const float* shift, uint32_t* argmax, void* stream) {
  ASF_CHECK_ARG(x && y && scale && shift && B > 0 && T > 0 && F > 0 && C > 0, "asf_bn_relu_maxpool_fwd: bad arguments");
  int To = (T + 2 - 3) / 2 + 1, Fo = (F + 2 - 3) / 2 + 1;
  if (dtype == ASF_BF16 && C % 8 == 0 && ld8(ld_y) && al16(x) && al16(y))
    return bn_relu_maxpool_fwd_v8((const bf16*)x, (bf16*)y, ld_y, B, T, F, C, scale, shift, argmax, (cudaStream_t)stream);
  ASF_CHECK_ARG(!argmax, "asf_bn_relu_maxpool_fwd: arg-max codes need the bf16 vector path (C %% 8 == 0, aligned)");
  ASF_DISPATCH_DTYPE(dtype, Tt, {
    bn_relu_maxpool_fwd_kernel<Tt><<<ew_grid((int64_t)B * To * Fo * C), 256, 0, (cudaStream_t)stream>>>(
        (const Tt*)x, (Tt*)y, ld_y, B, T, F, C, To, Fo, scale, shift);
    ASF_LAUNCH_CHECK();
    return 0;
  });
}

int asf_bn_relu_maxpool_bwd(const void* x, const void* dy, int64_t ld_dy, void* dz, int dtype, int B, int T, int F,
                            int C, const float* scale, const float* shift, const uint32_t* argmax, void* stream) {
  ASF_CHECK_ARG(dy && dz && (argmax || (x && scale && shift)) && B > 0 && T > 0 && F > 0 && C > 0,
                "asf_bn_relu_maxpool_bwd: bad arguments");
  int To = (T + 2 - 3) / 2 + 1, Fo = (F + 2 - 3) / 2 + 1;
  if (dtype == ASF_BF16 && C % 8 == 0 && ld8(ld_dy) && (argmax || al16(x)) && al16(dy) && al16(dz))
    return bn_relu_maxpool_bwd_v8((const bf16*)x, (const bf16*)dy, ld_dy, (bf16*)dz, B, T, F, C, scale, shift, argmax,
                                  (cudaStream_t)stream);
  ASF_CHECK_ARG(!argmax, "asf_bn_relu_maxpool_bwd: arg-max codes need the bf16 vector path (C %% 8 == 0, aligned)");
  ASF_DISPATCH_DTYPE(dtype, Tt, {
    bn_relu_maxpool_bwd_kernel<Tt><<<ew_grid((int64_t)B * T * F * C), 256, 0, (cudaStream_t)stream>>>(
        (const Tt*)x, (const Tt*)dy, ld_dy, (Tt*)dz, B, T, F, C, To, Fo, scale, shift);
    ASF_LAUNCH_CHECK();
    return 0;
  });
}

int asf_avgpool_fwd(const void* x, int dtype, int B, int P, int C, float* feat, int64_t ld_feat, int c_off,
                    void* stream) {
  ASF_CHECK_ARG(x && feat && B > 0 && P > 0 && C > 0, "asf_avgpool_fwd: bad arguments");
  dim3 grid((unsigned)cdiv(C, 32), B), block(32, 8);
  ASF_DISPATCH_DTYPE(dtype, T, {
    avgpool_fwd_kernel<T><<<grid, block, 0, (cudaStream_t)stream>>>((const T*)x, B, P, C, feat, ld_feat, c_off);
    ASF_LAUNCH_CHECK();
    return 0;
  });
}

int asf_avgpool_bwd(const float* dfeat, int64_t ld_feat, int c_off, const void* mask, void* dx, int dtype, int B,
                    int P, int C, void* stream) {
  ASF_CHECK_ARG(dfeat && dx && B > 0 && P > 0 && C > 0, "asf_avgpool_bwd: bad arguments");
  if (dtype == ASF_BF16 && C % 8 == 0 && al16(dx) && (!mask || al16(mask)) && al16(dfeat) && (ld_feat % 4 == 0) && (c_off % 4 == 0))
    return avgpool_bwd_v8(dfeat, ld_feat, c_off, (const bf16*)mask, (bf16*)dx, B, P, C, (cudaStream_t)stream);
  ASF_DISPATCH_DTYPE(dtype, T, {
    avgpool_bwd_kernel<T><<<ew_grid((int64_t)B * P * C), 256, 0, (cudaStream_t)stream>>>(dfeat, ld_feat, c_off,
                                                                                       (const T*)mask, (T*)dx, B, P, C);
    ASF_LAUNCH_CHECK();
    return 0;
  });
}

int asf_sgemm(const float* A, const float* Bm, float* C, int M, int N, int K, int ta, int tb, int64_t lda,
              int64_t ldb, int64_t ldc, float alpha, float beta, const float* bias, void* stream) {
  ASF_CHECK_ARG(A && Bm && C && M > 0 && N > 0 && K > 0, "asf_sgemm: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  if (tb) {
    int64_t threads = (int64_t)M * N * 32;
    sgemm_nt_kernel<<<(unsigned)cdiv(threads, 256), 256, 0, st>>>(A, Bm, C, M, N, K, ta, lda, ldb, ldc, alpha, beta,
                                                                  bias);
  } else {
    sgemm_nn_kernel<<<(unsigned)cdiv((int64_t)M * N, 256), 256, 0, st>>>(A, Bm, C, M, N, K, ta, lda, ldb, ldc, alpha,
                                                                         beta, bias);
  }
  ASF_LAUNCH_CHECK();
  return 0;
}

int asf_softmax_rows(float* x, int rows, int n, int64_t ld, void* stream) {
  ASF_CHECK_ARG(x && rows > 0 && n > 0, "asf_softmax_rows: bad arguments");
  softmax_rows_kernel<<<(unsigned)cdiv((int64_t)rows * 32, 256), 256, 0, (cudaStream_t)stream>>>(x, rows, n, ld);
  ASF_LAUNCH_CHECK();
  return 0;
}

int asf_scale_mask(const float* x, const float* mask, float* y, int64_t n, float inv_keep, void* stream) {
  ASF_CHECK_ARG(x && mask && y && n > 0, "asf_scale_mask: bad arguments");
  scale_mask_kernel<<<(unsigned)cdiv(n, 256), 256, 0, (cudaStream_t)stream>>>(x, mask, y, n, inv_keep);
  ASF_LAUNCH_CHECK();
  return 0;
}

int asf_gather_rows(const float* src, float* dst, const int32_t* idx, int B, int T_src, int n_idx, int F,
                    void* stream) {
  ASF_CHECK_ARG(src && dst && idx && B > 0 && n_idx > 0 && F > 0, "asf_gather_rows: bad arguments");
  gather_rows_kernel<<<ew_grid((int64_t)B * n_idx * F), 256, 0, (cudaStream_t)stream>>>(src, dst, idx, B, T_src,
                                                                                      n_idx, F);
  ASF_LAUNCH_CHECK();
  return 0;
}

int asf_pack_weight(const float* w_oihw, void* packed, int dtype, int O, int I, int taps, const float* rscale,
                    void* stream) {
  ASF_CHECK_ARG(w_oihw && packed && O > 0 && I > 0 && taps > 0, "asf_pack_weight: bad arguments");
  ASF_DISPATCH_DTYPE(dtype, T, {
    pack_weight_kernel<T><<<ew_grid((int64_t)O * I * taps), 256, 0, (cudaStream_t)stream>>>(w_oihw, (T*)packed, O, I,
                                                                                         taps, rscale, 0);
    ASF_LAUNCH_CHECK();
    return 0;
  });
}

int asf_pack_weight_t(const float* w_oihw, void* packed, int dtype, int O, int I, int taps, const float* rscale,
                      void* stream) {
  ASF_CHECK_ARG(w_oihw && packed && O > 0 && I > 0 && taps > 0, "asf_pack_weight_t: bad arguments");
  ASF_DISPATCH_DTYPE(dtype, T, {
    pack_weight_kernel<T><<<ew_grid((int64_t)O * I * taps), 256, 0, (cudaStream_t)stream>>>(w_oihw, (T*)packed, O, I,
                                                                                         taps, rscale, 1);
    ASF_LAUNCH_CHECK();
    return 0;
  });
}

int asf_pack_weights_multi(const asf_pack_desc_t* descs_dev, int n, long long total_chunks, void* stream) {
  ASF_CHECK_ARG(descs_dev && n > 0 && total_chunks > 0 && total_chunks < (1ll << 31), "asf_pack_weights_multi: bad arguments");
  pack_weights_multi_kernel<<<(unsigned)total_chunks, 256, 0, (cudaStream_t)stream>>>(descs_dev, n);
  ASF_LAUNCH_CHECK();
  return 0;
}

int asf_sgd_multi(const asf_sgd_desc_t* descs_dev, int n, long long total_chunks, const float* lr_dev, float momentum,
                  float dampening, int nesterov, int first_step, void* stream) {
  ASF_CHECK_ARG(descs_dev && lr_dev && n > 0 && total_chunks > 0 && total_chunks < (1ll << 31), "asf_sgd_multi: bad arguments");
  sgd_multi_kernel<<<(unsigned)total_chunks, 256, 0, (cudaStream_t)stream>>>(descs_dev, n, lr_dev, momentum, dampening,
                                                                            nesterov, first_step);
  ASF_LAUNCH_CHECK();
  return 0;
}

// scalar-path geometry: blocks of rows (also the number of partial rows)
static void stem_scalar_geom(const asf_conv_geom_t* g, int64_t* blocks, int64_t* rpb) {
  const int64_t rows = (int64_t)g->B * g->Tout * g->Fout;
  int64_t b = 148 * 2;
  int64_t r = cdiv(rows, b);
  const int RL = 256 / g->Cout;
  if (r < RL) r = RL;
  *rpb = r;
  *blocks = cdiv(rows, r);
}
static bool stem_vec_ok(const asf_conv_geom_t* g, int dout_dtype) {
  return dout_dtype == ASF_BF16 && g->Cout % 8 == 0 && g->kf <= 8 && (g->Cout / 8) * g->kt <= 256 && ld8(g->ld_out);
}

int asf_stem_wgrad_ksplit(const asf_conv_geom_t* g, int dout_dtype) {
  ASF_CHECK_ARG(g && g->Cin == 1 && g->kt * g->kf <= STEM_MAX_TAPS && g->Cout > 0 && g->Cout <= 256 && 256 % g->Cout == 0,
                "asf_stem_wgrad: needs Cin == 1, <= %d taps, Cout dividing 256", STEM_MAX_TAPS);
  if (stem_vec_ok(g, dout_dtype)) return stem_wgrad_v9_blocks(g);
  int64_t blocks, rpb;
  stem_scalar_geom(g, &blocks, &rpb);
  return (int)blocks;
}

int asf_stem_wgrad(const float* in, const void* dout, int dout_dtype, float* dwp, const asf_conv_geom_t* g,
                   void* stream) {
  ASF_CHECK_ARG(in && dout && dwp && g, "asf_stem_wgrad: null pointer");
  ASF_CHECK_ARG(g->Cin == 1 && g->kt * g->kf <= STEM_MAX_TAPS && g->Cout <= 256 && 256 % g->Cout == 0,
                "asf_stem_wgrad: needs Cin == 1, <= %d taps, Cout dividing 256", STEM_MAX_TAPS);
  cudaStream_t st = (cudaStream_t)stream;
  if (stem_vec_ok(g, dout_dtype) && al16(dout)) return stem_wgrad_v9(in, (const bf16*)dout, dwp, g, st);
  ASF_CHECK_ARG(!stem_vec_ok(g, dout_dtype), "asf_stem_wgrad: dout must be 16-byte aligned");
  int64_t blocks, rpb;
  stem_scalar_geom(g, &blocks, &rpb);
  ASF_DISPATCH_DTYPE(dout_dtype, TD, {
    stem_wgrad_kernel<TD><<<(unsigned)blocks, 256, 0, st>>>(in, (const TD*)dout, dwp, *g, rpb);
    ASF_LAUNCH_CHECK();
    return 0;
  });
}

int asf_stem_wgrad_fused_rows(const asf_conv_geom_t* g, int dout_dtype) {
  if (!g || g->Cin != 1 || g->kt * g->kf > STEM_MAX_TAPS || !stem_vec_ok(g, dout_dtype)) return 0;
  return stem_wgrad_v10_rows(g);
}

int asf_stem_wgrad_fused(const float* in, const void* dout, int dout_dtype, float* dwp, float* dw, int accumulate,
                         uint32_t* ticket, const asf_conv_geom_t* g, void* stream) {
  ASF_CHECK_ARG(in && dout && dwp && dw && ticket && g, "asf_stem_wgrad_fused: null pointer");
  ASF_CHECK_ARG(asf_stem_wgrad_fused_rows(g, dout_dtype) > 0 && al16(dout) && al16(dwp),
                "asf_stem_wgrad_fused: shape / dtype / alignment not supported (see asf_stem_wgrad_fused_rows)");
  return stem_wgrad_v10(in, (const bf16*)dout, dwp, dw, accumulate, ticket, g, (cudaStream_t)stream);
}

int asf_unpack_wgrad_multi(const asf_unpack_desc_t* descs_dev, int n, long long total_chunks, void* stream) {
  ASF_CHECK_ARG(descs_dev && n > 0 && total_chunks > 0 && total_chunks < (1ll << 31), "asf_unpack_wgrad_multi: bad arguments");
  // (every layer: I % 4 == 0 and 16-byte aligned dwp -- the tap-GEMM wgrad already requires I % 8 == 0)
  ASF_CUDA(launch_pdl(unpack_wgrad_multi_kernel, dim3((unsigned)total_chunks), dim3(256), 0, (cudaStream_t)stream, descs_dev, n));
  ASF_LAUNCH_CHECK();
  return 0;
}

int asf_unpack_wgrad(const float* dwp, float* dw_oihw, int O, int I, int taps, const float* rscale,
                     int accumulate, int ksplit, void* stream) {
  ASF_CHECK_ARG(dwp && dw_oihw && O > 0 && I > 0 && taps > 0 && ksplit >= 1, "asf_unpack_wgrad: bad arguments");
  const int sl_log2 = ksplit < 4 ? 0 : (ksplit < 32 ? 2 : 4);
  const int64_t E = (int64_t)O * I * taps;
  ASF_CUDA(launch_pdl(unpack_wgrad_kernel, dim3((unsigned)cdiv(E, 256 >> sl_log2)), dim3(256), 0, (cudaStream_t)stream,
                      dwp, dw_oihw, O, I, taps, rscale, accumulate, ksplit, sl_log2));
  ASF_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
