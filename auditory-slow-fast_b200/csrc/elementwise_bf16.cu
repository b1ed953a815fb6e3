// Vectorised bf16 streaming kernels (8 channels = one 16-byte access per thread) for the HBM-bound passes of the
// hot path: training-mode BatchNorm apply / backward, per-channel reductions, ReLU/avg-pool/max-pool backward and
// the C_in = 1 stems.  elementwise.cu dispatches here whenever the tensor is bf16 with C % 8 == 0 and 16-byte
// aligned rows (every activation of the network); the scalar kernels there remain for fp32 parity mode and
// ragged layouts.
//
// Replaces the separate cuDNN BatchNorm / ATen elementwise / pooling kernels behind
//   /root/reference/audio_slowfast/models/stem_helper.py:150-168, resnet_helper.py:181-225, :335-341.
#include "common.cuh"
#include "stats_reduce.cuh"

namespace asf {

struct F8 {
  float v[8];
};

__device__ __forceinline__ F8 unpack8(const uint4& q) {
  F8 r;
  const __nv_bfloat162* h = (const __nv_bfloat162*)&q;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float2 f = __bfloat1622float2(h[j]);
    r.v[2 * j] = f.x;
    r.v[2 * j + 1] = f.y;
  }
  return r;
}
__device__ __forceinline__ uint4 pack8(const F8& a) {
  uint4 q;
  __nv_bfloat162* h = (__nv_bfloat162*)&q;
#pragma unroll
  for (int j = 0; j < 4; ++j) h[j] = __floats2bfloat162_rn(a.v[2 * j], a.v[2 * j + 1]);
  return q;
}
__device__ __forceinline__ F8 ldf8(const float* p) {
  F8 r;
  const float4 a = *(const float4*)p, b = *(const float4*)(p + 4);
  r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w;
  r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
  return r;
}

static inline unsigned vgrid(int64_t nvec) {
  int64_t g = cdiv(nvec, 256), cap = 148 * 16;
  return (unsigned)(g < 1 ? 1 : (g > cap ? cap : g));
}

// ------------------------------------------------------------------------------------------- column streaming
// The training-BN passes are "column-owned": a block is CVB vector-columns (8 channels each) x RP row lanes; a thread
// keeps its 8 channels' parameters in registers and streams rows r0 + rl, + RP, ... four at a time (four independent
// 16-byte loads per tensor in flight).  Consecutive threads touch consecutive 16-byte vectors, rows are contiguous.
struct ColGeom {
  int CVB, RP, cblocks;
  int64_t rows_per_block;
  dim3 grid;
};
static ColGeom col_geom(int64_t rows, int C, int blocks_per_sm) {
  ColGeom g;
  const int CV = C / 8;
  g.CVB = CV < 256 ? CV : 256;
  g.cblocks = (int)cdiv(CV, g.CVB);
  g.RP = 256 / g.CVB;
  int64_t target = 148 * blocks_per_sm / g.cblocks;
  if (target < 1) target = 1;
  int64_t rpb = cdiv(rows, target);
  const int64_t quantum = (int64_t)g.RP * 4;
  rpb = cdiv(rpb, quantum) * quantum;
  g.rows_per_block = rpb;
  g.grid = dim3((unsigned)cdiv(rows, rpb), (unsigned)g.cblocks);
  return g;
}

// ------------------------------------------------------------------------------------------- BN apply (train)
// FINALIZE: scale / shift are derived here from the conv epilogue's sums (stats[c], stats[C + c]); the blocks of the
// first row range also publish scale / shift / mean / invstd and update the running statistics (what bn_finalize did
// as a separate launch per layer).
template <bool FINALIZE>
__global__ void __launch_bounds__(256) bn_apply_v8_kernel(const bf16* __restrict__ x, int64_t ld_x,
                                                          const bf16* __restrict__ res, int64_t ld_res,
                                                          bf16* __restrict__ y, int64_t ld_y, int64_t rows, int C,
                                                          int CVB, int64_t rows_per_block, float* scale, float* shift,
                                                          int relu, const double* __restrict__ stats, double count,
                                                          const float* __restrict__ gamma,
                                                          const float* __restrict__ beta, float eps, float momentum,
                                                          float* running_mean, float* running_var,
                                                          long long* num_batches_tracked, float* save_mean,
                                                          float* save_invstd) {
  pdl_trigger();
  pdl_wait();
  const int RP = 256 / CVB;
  const int cvl = threadIdx.x % CVB, rl = threadIdx.x / CVB;
  const int c = (blockIdx.y * CVB + cvl) * 8;
  if (rl >= RP || c >= C) return;
  F8 sc, sh;
  if (FINALIZE) {
    const bool publish = blockIdx.x == 0 && rl == 0;
    if (publish && c == 0 && num_batches_tracked) *num_batches_tracked += 1;
    // the sums are doubles (ordered two-stage reduction): mean and E[x^2] - mean^2 are formed in double, so the
    // variance does not lose digits to the cancellation; only the final rsqrt runs in fp32 in the streaming blocks
    double s0[8], s1[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      s0[j] = stats[c + j];
      s1[j] = stats[C + c + j];
    }
    F8 ga, be;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      ga.v[j] = gamma ? gamma[c + j] : 1.f;
      be.v[j] = beta ? beta[c + j] : 0.f;
    }
    if (!publish) {
      const double inv_n = 1.0 / count;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const double mean = s0[j] * inv_n;
        const float var = fmaxf((float)fma(-mean, mean, s1[j] * inv_n), 0.f);
        const float mf = (float)mean;
        sc.v[j] = ga.v[j] * rsqrtf(var + eps);
        sh.v[j] = fmaf(-mf, sc.v[j], be.v[j]);
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (publish) {
        const double mean = s0[j] / count;
        double var = s1[j] / count - mean * mean;
        if (var < 0.0) var = 0.0;
        const float invstd = (float)(1.0 / sqrt(var + (double)eps));
        sc.v[j] = ga.v[j] * invstd;
        sh.v[j] = be.v[j] - (float)mean * sc.v[j];
        scale[c + j] = sc.v[j];
        shift[c + j] = sh.v[j];
        if (save_mean) save_mean[c + j] = (float)mean;
        if (save_invstd) save_invstd[c + j] = invstd;
        if (running_mean) {
          const double unbiased = count > 1.0 ? var * count / (count - 1.0) : var;
          running_mean[c + j] = (1.f - momentum) * running_mean[c + j] + momentum * (float)mean;
          running_var[c + j] = (1.f - momentum) * running_var[c + j] + momentum * (float)unbiased;
        }
      }
    }
  } else {
    sc = ldf8(scale + c);
    sh = ldf8(shift + c);
  }
  if (!y) return;
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
  int64_t r1 = r0 + rows_per_block;
  if (r1 > rows) r1 = rows;
  for (int64_t r = r0 + rl; r < r1; r += (int64_t)RP * 4) {
    uint4 xv[4], rv[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int64_t rr = r + (int64_t)k * RP;
      if (rr < r1) {
        xv[k] = *(const uint4*)(x + rr * ld_x + c);
        if (res) rv[k] = *(const uint4*)(res + rr * ld_res + c);
      }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int64_t rr = r + (int64_t)k * RP;
      if (rr < r1) {
        F8 a = unpack8(xv[k]);
#pragma unroll
        for (int j = 0; j < 8; ++j) a.v[j] = fmaf(a.v[j], sc.v[j], sh.v[j]);
        if (res) {
          const F8 q = unpack8(rv[k]);
#pragma unroll
          for (int j = 0; j < 8; ++j) a.v[j] += q.v[j];
        }
        if (relu) {
#pragma unroll
          for (int j = 0; j < 8; ++j) a.v[j] = fmaxf(a.v[j], 0.f);
        }
        *(uint4*)(y + rr * ld_y + c) = pack8(a);
      }
    }
  }
}

int bn_apply_v8(const bf16* x, int64_t ld_x, const bf16* res, int64_t ld_res, bf16* y, int64_t ld_y, int64_t rows,
                int C, const float* scale, const float* shift, int relu, cudaStream_t st) {
  const ColGeom g = col_geom(rows, C, 4);
  ASF_CUDA(launch_pdl(bn_apply_v8_kernel<false>, dim3(g.grid), dim3(256), 0, st, x, ld_x, res, ld_res, y, ld_y, rows, C, g.CVB, g.rows_per_block,
                                                   (float*)scale, (float*)shift, relu, nullptr, 1.0, nullptr, nullptr,
                                                   0.f, 0.f, nullptr, nullptr, nullptr, nullptr, nullptr));
  ASF_LAUNCH_CHECK();
  return 0;
}

int bn_train_fwd_v8(const bf16* x, int64_t ld_x, const bf16* res, int64_t ld_res, bf16* y, int64_t ld_y, int64_t rows,
                    int C, const double* stats, double count, const float* gamma, const float* beta, float eps,
                    float momentum, float* running_mean, float* running_var, int64_t* num_batches_tracked, float* scale,
                    float* shift, float* save_mean, float* save_invstd, int relu, cudaStream_t st) {
  const ColGeom g = col_geom(rows, C, 4);
  ASF_CUDA(launch_pdl(bn_apply_v8_kernel<true>, dim3(g.grid), dim3(256), 0, st, x, ld_x, res, ld_res, y, ld_y, rows, C, g.CVB, g.rows_per_block,
                                                  scale, shift, relu, stats, count, gamma, beta, eps, momentum,
                                                  running_mean, running_var, (long long*)num_batches_tracked, save_mean,
                                                  save_invstd));
  ASF_LAUNCH_CHECK();
  return 0;
}

// ------------------------------------------------------------------------------------------- BN backward apply
// dx = gamma*invstd * (dz - red[c]/count - xhat * red[C+c]/count); the first row range also publishes dgamma / dbeta.
__global__ void __launch_bounds__(256) bn_bwd_apply_v8_kernel(const bf16* __restrict__ dz, int64_t ld_dz,
                                                              const bf16* __restrict__ x, int64_t ld_x,
                                                              bf16* __restrict__ dx, int64_t ld_dx, int64_t rows, int C,
                                                              int CVB, int64_t rows_per_block,
                                                              const float* __restrict__ gamma,
                                                              const float* __restrict__ mean,
                                                              const float* __restrict__ invstd,
                                                              const double* __restrict__ red, double inv_count,
                                                              float* dgamma, float* dbeta, int accumulate) {
  pdl_trigger();
  pdl_wait();
  const int RP = 256 / CVB;
  const int cvl = threadIdx.x % CVB, rl = threadIdx.x / CVB;
  const int c = (blockIdx.y * CVB + cvl) * 8;
  if (rl >= RP || c >= C) return;
  const F8 mu = ldf8(mean + c), is = ldf8(invstd + c);
  double r0v[8], r1v[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    r0v[j] = red[c + j];
    r1v[j] = red[C + c + j];
  }
  if (blockIdx.x == 0 && rl == 0) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (dbeta) dbeta[c + j] = (accumulate ? dbeta[c + j] : 0.f) + (float)r0v[j];
      if (dgamma) dgamma[c + j] = (accumulate ? dgamma[c + j] : 0.f) + (float)r1v[j];
    }
  }
  // o = a * (g - k0 - (x - mu) * is * k1) = a * g - (a * k0 - a * is * k1 * mu) - (a * is * k1) * x
  F8 ca, cb, cc;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float ga = gamma ? gamma[c + j] : 1.f;
    ca.v[j] = ga * is.v[j];
    cb.v[j] = (float)(r0v[j] * inv_count);
    cc.v[j] = (float)(r1v[j] * inv_count);
  }
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
  int64_t r1 = r0 + rows_per_block;
  if (r1 > rows) r1 = rows;
  for (int64_t r = r0 + rl; r < r1; r += (int64_t)RP * 4) {
    uint4 gv[4], xv[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int64_t rr = r + (int64_t)k * RP;
      if (rr < r1) {
        gv[k] = *(const uint4*)(dz + rr * ld_dz + c);
        xv[k] = *(const uint4*)(x + rr * ld_x + c);
      }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int64_t rr = r + (int64_t)k * RP;
      if (rr < r1) {
        const F8 g = unpack8(gv[k]), xx = unpack8(xv[k]);
        F8 o;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float xh = (xx.v[j] - mu.v[j]) * is.v[j];
          o.v[j] = ca.v[j] * (g.v[j] - cb.v[j] - xh * cc.v[j]);
        }
        *(uint4*)(dx + rr * ld_dx + c) = pack8(o);
      }
    }
  }
}

int bn_bwd_apply_v8(const bf16* dz, int64_t ld_dz, const bf16* x, int64_t ld_x, bf16* dx, int64_t ld_dx, int64_t rows,
                    int C, const float* gamma, const float* mean, const float* invstd, const double* red,
                    double inv_count, float* dgamma, float* dbeta, int accumulate, cudaStream_t st) {
  const ColGeom g = col_geom(rows, C, 4);
  ASF_CUDA(launch_pdl(bn_bwd_apply_v8_kernel, dim3(g.grid), dim3(256), 0, st, dz, ld_dz, x, ld_x, dx, ld_dx, rows, C, g.CVB, g.rows_per_block, gamma,
                                                mean, invstd, red, inv_count, dgamma, dbeta, accumulate));
  ASF_LAUNCH_CHECK();
  return 0;
}

// ------------------------------------------------------------------------------------------- channel reductions
// out[c] = sum_rows a ; out[C + c] = sum_rows (kBwd ? a * (x - mean) * invstd : a * a), as doubles, DETERMINISTIC:
//   thread:  streams its rows (four 16-byte loads in flight) into 8 + 8 fp32 accumulators;
//   warp:    row lanes that share a column inside a warp are folded with xor-shuffles (a fixed tree);
//   block:   the 8 warp partials of a column are added in warp order and STORED as this block's row of the workspace;
//   grid:    the last block of a channel block to finish (ticket) adds the rows in ascending order in double.
// A block covers CVB <= 32 vector columns (<= 256 channels) so that the final pass of one block stays short and the
// channel blocks finish in parallel.  Workspace row of block (x, y): ws[(y * gridDim.x + x)][2][CVB * 8].
template <bool kBwd>
__global__ void __launch_bounds__(256) channel_reduce_v8_kernel(const bf16* __restrict__ a, int64_t ld_a,
                                                                const bf16* __restrict__ x, int64_t ld_x,
                                                                int64_t rows, int C, int CVB, int64_t rows_per_block,
                                                                const float* __restrict__ mean,
                                                                const float* __restrict__ invstd, double* out,
                                                                float* ws, unsigned* ticket) {
  pdl_trigger();
  pdl_wait();
  __shared__ __align__(16) float part[2][8][256];  // [sum | second sum][warp][channel of the block]
  __shared__ unsigned s_flag;
  const int RP = 256 / CVB;
  const int cvl = threadIdx.x % CVB, rl = threadIdx.x / CVB;
  const int c = (blockIdx.y * CVB + cvl) * 8;
  const int warp = threadIdx.x >> 5;
  float acc0[8], acc1[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc0[j] = acc1[j] = 0.f;
  const bool active = c < C;
  if (active) {
    const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
    int64_t r1 = r0 + rows_per_block;
    if (r1 > rows) r1 = rows;
    F8 mu, is;
    if (kBwd) {
      mu = ldf8(mean + c);
      is = ldf8(invstd + c);
    }
    for (int64_t r = r0 + rl; r < r1; r += (int64_t)RP * 4) {
      uint4 av[4], xv[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int64_t rr = r + (int64_t)k * RP;
        if (rr < r1) {
          av[k] = *(const uint4*)(a + rr * ld_a + c);
          if (kBwd) xv[k] = *(const uint4*)(x + rr * ld_x + c);
        }
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int64_t rr = r + (int64_t)k * RP;
        if (rr < r1) {
          const F8 v = unpack8(av[k]);
          if (kBwd) {
            const F8 xx = unpack8(xv[k]);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              acc0[j] += v.v[j];
              acc1[j] = fmaf(v.v[j], (xx.v[j] - mu.v[j]) * is.v[j], acc1[j]);
            }
          } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              acc0[j] += v.v[j];
              acc1[j] = fmaf(v.v[j], v.v[j], acc1[j]);
            }
          }
        }
      }
    }
  }
  // lanes of one warp that own the same column: lane ^ CVB, ^ 2 CVB, ... (CVB is a power of two <= 32)
  for (int o = 16; o >= CVB; o >>= 1) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      acc0[j] += __shfl_xor_sync(0xffffffffu, acc0[j], o);
      acc1[j] += __shfl_xor_sync(0xffffffffu, acc1[j], o);
    }
  }
  if ((int)(threadIdx.x & 31) < CVB) {  // one holder per (warp, column); CVB == 32: a warp is one row lane
    *(float4*)&part[0][warp][cvl * 8] = make_float4(acc0[0], acc0[1], acc0[2], acc0[3]);
    *(float4*)&part[0][warp][cvl * 8 + 4] = make_float4(acc0[4], acc0[5], acc0[6], acc0[7]);
    *(float4*)&part[1][warp][cvl * 8] = make_float4(acc1[0], acc1[1], acc1[2], acc1[3]);
    *(float4*)&part[1][warp][cvl * 8 + 4] = make_float4(acc1[4], acc1[5], acc1[6], acc1[7]);
  }
  __syncthreads();
  const int CB = CVB * 8;
  float* my = ws + ((int64_t)blockIdx.y * gridDim.x + blockIdx.x) * 2 * CB;
  for (int i = threadIdx.x; i < 2 * CB; i += 256) {
    const int half = i / CB, ch = i - half * CB;
    float t = part[half][0][ch];
#pragma unroll
    for (int w = 1; w < 8; ++w) t += part[half][w][ch];
    my[i] = t;
  }
  if (!stats_last_cta(ticket + blockIdx.y, gridDim.x, &s_flag)) return;
  const float* col = ws + (int64_t)blockIdx.y * gridDim.x * 2 * CB;
  const int c0 = blockIdx.y * CB;
  stats_ordered_reduce(col, (int)gridDim.x, 2 * CB, 2 * CB, (double*)&part[0][0][0], [&](int s, double v) {
    const int half = s / CB, cc = c0 + (s - half * CB);
    if (cc < C) out[half * C + cc] = v;
  });
}

template <bool kBwd>
int channel_reduce_v8(const bf16* a, int64_t ld_a, const bf16* x, int64_t ld_x, int64_t rows, int C, const float* mean,
                      const float* invstd, double* out, float* ws, unsigned* ticket, cudaStream_t st) {
  const int CV = C / 8;
  int CVB = 1;
  while (CVB < 32 && CVB < CV) CVB *= 2;
  const int cblocks = (int)cdiv(CV, CVB);
  ASF_CHECK_ARG(cblocks <= ASF_STATS_TICKETS, "channel reduce: %d channel blocks exceed the %d tickets", cblocks, ASF_STATS_TICKETS);
  const int RP = 256 / CVB;
  int64_t target = 148 * 2 / cblocks;  // <= 296 blocks in all: the partial rows fit ASF_STATS_WS_FLOATS(C)
  if (target < 1) target = 1;
  int64_t rpb = cdiv(rows, target);
  const int64_t quantum = (int64_t)RP * 4;
  rpb = cdiv(rpb, quantum) * quantum;
  dim3 grid((unsigned)cdiv(rows, rpb), (unsigned)cblocks);
  ASF_CUDA(launch_pdl(channel_reduce_v8_kernel<kBwd>, grid, dim3(256), 0, st, a, ld_a, x, ld_x, rows, C, CVB, rpb, mean,
                      invstd, out, ws, ticket));
  ASF_LAUNCH_CHECK();
  return 0;
}
template int channel_reduce_v8<false>(const bf16*, int64_t, const bf16*, int64_t, int64_t, int, const float*,
                                      const float*, double*, float*, unsigned*, cudaStream_t);
template int channel_reduce_v8<true>(const bf16*, int64_t, const bf16*, int64_t, int64_t, int, const float*,
                                     const float*, double*, float*, unsigned*, cudaStream_t);

// ------------------------------------------------------------------------------------------- ReLU backward / add
__global__ void __launch_bounds__(256) relu_bwd_v8_kernel(const bf16* __restrict__ a, int64_t ld_a,
                                                          const bf16* __restrict__ b, int64_t ld_b,
                                                          const bf16* __restrict__ mask, int64_t ld_mask,
                                                          bf16* __restrict__ y, int64_t ld_y, int64_t nvec, int CV) {
  pdl_trigger();
  pdl_wait();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / CV;
    const int c = (int)(i - r * CV) * 8;
    F8 v = unpack8(*(const uint4*)(a + r * ld_a + c));
    if (b) {
      const F8 q = unpack8(*(const uint4*)(b + r * ld_b + c));
#pragma unroll
      for (int j = 0; j < 8; ++j) v.v[j] += q.v[j];
    }
    if (mask) {
      const F8 m = unpack8(*(const uint4*)(mask + r * ld_mask + c));
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (!(m.v[j] > 0.f)) v.v[j] = 0.f;
    }
    *(uint4*)(y + r * ld_y + c) = pack8(v);
  }
}

int relu_bwd_v8(const bf16* a, int64_t ld_a, const bf16* b, int64_t ld_b, const bf16* mask, int64_t ld_mask, bf16* y,
                int64_t ld_y, int64_t rows, int C, cudaStream_t st) {
  const int64_t nvec = rows * (C / 8);
  ASF_CUDA(launch_pdl(relu_bwd_v8_kernel, dim3(vgrid(nvec)), dim3(256), 0, st, a, ld_a, b, ld_b, mask, ld_mask, y, ld_y, nvec, C / 8));
  ASF_LAUNCH_CHECK();
  return 0;
}

// ------------------------------------------------------------------------------------------- avg-pool backward
__global__ void __launch_bounds__(256) avgpool_bwd_v8_kernel(const float* __restrict__ dfeat, int64_t ld_feat,
                                                             int c_off, const bf16* __restrict__ mask,
                                                             bf16* __restrict__ dx, int64_t nvec, int P, int CV,
                                                             float inv) {
  pdl_trigger();
  pdl_wait();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / CV;
    const int c = (int)(i - r * CV) * 8;
    const int b = (int)(r / P);
    F8 v = ldf8(dfeat + (int64_t)b * ld_feat + c_off + c);
#pragma unroll
    for (int j = 0; j < 8; ++j) v.v[j] *= inv;
    if (mask) {
      const F8 m = unpack8(*(const uint4*)(mask + i * 8));
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (!(m.v[j] > 0.f)) v.v[j] = 0.f;
    }
    *(uint4*)(dx + i * 8) = pack8(v);
  }
}

int avgpool_bwd_v8(const float* dfeat, int64_t ld_feat, int c_off, const bf16* mask, bf16* dx, int B, int P, int C,
                   cudaStream_t st) {
  const int64_t nvec = (int64_t)B * P * (C / 8);
  ASF_CUDA(launch_pdl(avgpool_bwd_v8_kernel, dim3(vgrid(nvec)), dim3(256), 0, st, dfeat, ld_feat, c_off, mask, dx, nvec, P, C / 8, 1.f / (float)P));
  ASF_LAUNCH_CHECK();
  return 0;
}

// ------------------------------------------------------------------------------------------- stem max-pool
// y[b,to,fo,c] = max over the 3x3 / stride 2 / pad 1 window of relu(x*scale + shift)
// argmax != NULL: also records, per (window, 8-channel group), one nibble per channel = position (dt+1)*3 + (df+1) of the
// FIRST strict maximum of the positive values in scan order (15: none positive, nothing flows back through the ReLU);
// the backward then never re-reads x.
__global__ void __launch_bounds__(256) bn_relu_maxpool_fwd_v8_kernel(const bf16* __restrict__ x, bf16* __restrict__ y,
                                                                     int64_t ld_y, int64_t nvec, int Tn, int F, int CV,
                                                                     int To, int Fo, const float* __restrict__ scale,
                                                                     const float* __restrict__ shift,
                                                                     uint32_t* __restrict__ argmax) {
  pdl_trigger();
  pdl_wait();
  const int C = CV * 8;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % CV) * 8;
    int64_t r = i / CV;
    const int64_t orow = r;
    const int fo = (int)(r % Fo);
    r /= Fo;
    const int to = (int)(r % To);
    const int b = (int)(r / To);
    const F8 sc = ldf8(scale + c), sh = ldf8(shift + c);
    F8 m;
#pragma unroll
    for (int j = 0; j < 8; ++j) m.v[j] = 0.f;  // relu output is >= 0 and every window holds a valid element
    uint32_t code = 0xffffffffu;
#pragma unroll
    for (int dt = -1; dt <= 1; ++dt) {
      const int t = 2 * to + dt;
      if (t < 0 || t >= Tn) continue;
#pragma unroll
      for (int df = -1; df <= 1; ++df) {
        const int f = 2 * fo + df;
        if (f < 0 || f >= F) continue;
        const F8 v = unpack8(*(const uint4*)(x + (((int64_t)b * Tn + t) * F + f) * C + c));
        const uint32_t pos = (uint32_t)((dt + 1) * 3 + (df + 1));
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float a = fmaf(v.v[j], sc.v[j], sh.v[j]);
          if (a > m.v[j]) {
            m.v[j] = a;
            code = (code & ~(15u << (4 * j))) | (pos << (4 * j));
          }
        }
      }
    }
    *(uint4*)(y + orow * ld_y + c) = pack8(m);
    if (argmax) argmax[i] = code;
  }
}

int bn_relu_maxpool_fwd_v8(const bf16* x, bf16* y, int64_t ld_y, int B, int T, int F, int C, const float* scale,
                           const float* shift, uint32_t* argmax, cudaStream_t st) {
  const int To = (T + 2 - 3) / 2 + 1, Fo = (F + 2 - 3) / 2 + 1;
  const int64_t nvec = (int64_t)B * To * Fo * (C / 8);
  ASF_CUDA(launch_pdl(bn_relu_maxpool_fwd_v8_kernel, dim3(vgrid(nvec)), dim3(256), 0, st, x, y, ld_y, nvec, T, F, C / 8, To, Fo, scale, shift, argmax));
  ASF_LAUNCH_CHECK();
  return 0;
}

// dz[b,t,f,c] = (a > 0) * sum over the (<= 4) windows whose FIRST arg-max (t,f scan order, as torch) is (t,f) of dy
__global__ void __launch_bounds__(256) bn_relu_maxpool_bwd_v8_kernel(const bf16* __restrict__ x,
                                                                     const bf16* __restrict__ dy, int64_t ld_dy,
                                                                     bf16* __restrict__ dz, int64_t nvec, int Tn, int F,
                                                                     int CV, int To, int Fo,
                                                                     const float* __restrict__ scale,
                                                                     const float* __restrict__ shift) {
  pdl_trigger();
  pdl_wait();
  const int C = CV * 8;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % CV) * 8;
    int64_t r = i / CV;
    const int f = (int)(r % F);
    r /= F;
    const int t = (int)(r % Tn);
    const int b = (int)(r / Tn);
    const F8 sc = ldf8(scale + c), sh = ldf8(shift + c);
    const F8 xs = unpack8(*(const uint4*)(x + i * 8));
    F8 a, g;
    bool any = false;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      a.v[j] = fmaf(xs.v[j], sc.v[j], sh.v[j]);
      g.v[j] = 0.f;
      any = any || (a.v[j] > 0.f);
    }
    if (any) {
      // windows containing t: to in {t/2} if t even, {(t-1)/2, (t+1)/2} if odd (same along f)
      const int to0 = (t & 1) ? (t - 1) >> 1 : t >> 1, to1 = (t & 1) ? (t + 1) >> 1 : t >> 1;
      const int fo0 = (f & 1) ? (f - 1) >> 1 : f >> 1, fo1 = (f & 1) ? (f + 1) >> 1 : f >> 1;
      for (int to = to0; to <= to1; ++to) {
        if (to >= To) continue;
        for (int fo = fo0; fo <= fo1; ++fo) {
          if (fo >= Fo) continue;
          // is (t, f) the first arg-max of window (to, fo)?  an element scanned earlier must be strictly smaller...
          // torch keeps the first maximum: (t,f) wins iff every earlier element is < a and every later one is <= a
          bool win[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) win[j] = a.v[j] > 0.f;
          for (int dt = -1; dt <= 1; ++dt) {
            const int tt = 2 * to + dt;
            if (tt < 0 || tt >= Tn) continue;
            for (int df = -1; df <= 1; ++df) {
              const int ff = 2 * fo + df;
              if (ff < 0 || ff >= F || (tt == t && ff == f)) continue;
              const bool earlier = (tt < t) || (tt == t && ff < f);
              const F8 v = unpack8(*(const uint4*)(x + (((int64_t)b * Tn + tt) * F + ff) * C + c));
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const float o = fmaxf(fmaf(v.v[j], sc.v[j], sh.v[j]), 0.f);
                if (earlier ? (o >= a.v[j]) : (o > a.v[j])) win[j] = false;
              }
            }
          }
          const F8 d = unpack8(*(const uint4*)(dy + (((int64_t)b * To + to) * Fo + fo) * ld_dy + c));
#pragma unroll
          for (int j = 0; j < 8; ++j)
            if (win[j]) g.v[j] += d.v[j];
        }
      }
    }
    *(uint4*)(dz + i * 8) = pack8(g);
  }
}

// Backward from the forward's arg-max codes: dz[b,t,f,c] = sum of dy over the (<= 4) windows whose code names (t,f).
__global__ void __launch_bounds__(256) maxpool_bwd_argmax_v8_kernel(const uint32_t* __restrict__ argmax,
                                                                    const bf16* __restrict__ dy, int64_t ld_dy,
                                                                    bf16* __restrict__ dz, int64_t nvec, int Tn, int F,
                                                                    int CV, int To, int Fo) {
  pdl_trigger();
  pdl_wait();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
    const int cv = (int)(i % CV);
    int64_t r = i / CV;
    const int f = (int)(r % F);
    r /= F;
    const int t = (int)(r % Tn);
    const int b = (int)(r / Tn);
    // windows containing t: to in {t/2} if t even, {(t-1)/2, (t+1)/2} if odd (same along f)
    const int to0 = t >> 1, nto = (t & 1) ? 2 : 1;
    const int fo0 = f >> 1, nfo = (f & 1) ? 2 : 1;
    uint32_t code[4];
    uint4 dv[4];
    uint32_t pos[4];
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const int to = to0 + a, fo = fo0 + q;
        const bool ok = a < nto && q < nfo && to < To && fo < Fo;
        const int64_t w = ((int64_t)b * To + to) * Fo + fo;
        code[a * 2 + q] = ok ? argmax[w * CV + cv] : 0xffffffffu;
        dv[a * 2 + q] = ok ? *(const uint4*)(dy + w * ld_dy + cv * 8) : make_uint4(0, 0, 0, 0);
        pos[a * 2 + q] = (uint32_t)((t - (2 * to - 1)) * 3 + (f - (2 * fo - 1)));
      }
    F8 g;
#pragma unroll
    for (int j = 0; j < 8; ++j) g.v[j] = 0.f;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      const F8 d = unpack8(dv[w]);
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (((code[w] >> (4 * j)) & 15u) == pos[w]) g.v[j] += d.v[j];
    }
    *(uint4*)(dz + i * 8) = pack8(g);
  }
}

int bn_relu_maxpool_bwd_v8(const bf16* x, const bf16* dy, int64_t ld_dy, bf16* dz, int B, int T, int F, int C,
                           const float* scale, const float* shift, const uint32_t* argmax, cudaStream_t st) {
  const int To = (T + 2 - 3) / 2 + 1, Fo = (F + 2 - 3) / 2 + 1;
  const int64_t nvec = (int64_t)B * T * F * (C / 8);
  if (argmax) {
    ASF_CUDA(launch_pdl(maxpool_bwd_argmax_v8_kernel, dim3(vgrid(nvec)), dim3(256), 0, st, argmax, dy, ld_dy, dz, nvec, T, F,
                        C / 8, To, Fo));
    ASF_LAUNCH_CHECK();
    return 0;
  }
  ASF_CUDA(launch_pdl(bn_relu_maxpool_bwd_v8_kernel, dim3(vgrid(nvec)), dim3(256), 0, st, x, dy, ld_dy, dz, nvec, T, F, C / 8, To, Fo, scale, shift));
  ASF_LAUNCH_CHECK();
  return 0;
}

// ------------------------------------------------------------------------------------------- stems (C_in = 1)
// out[b,to,fo,n] = sum_taps in[b, to*st+it-pt, fo*sf+jf-pf] * w[n][tap]   (fp32 in / weights, bf16 out)
// one thread = one output position x 8 output channels; the weights sit transposed in shared memory.
__global__ void __launch_bounds__(256) stem_conv_fwd_kernel(const float* __restrict__ in, const float* __restrict__ w,
                                                            bf16* __restrict__ out, asf_conv_geom_t g, int64_t nvec) {
  pdl_trigger();
  pdl_wait();
  extern __shared__ float sw[];  // [taps][Cout]
  const int taps = g.kt * g.kf, CG = g.Cout / 8;
  for (int i = threadIdx.x; i < taps * g.Cout; i += blockDim.x) {
    const int n = i / taps, tp = i - n * taps;
    sw[tp * g.Cout + n] = w[i];
  }
  __syncthreads();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
    const int cg = (int)(i % CG);
    int64_t row = i / CG;
    const int64_t orow = row;
    const int fo = (int)(row % g.Fout);
    row /= g.Fout;
    const int to = (int)(row % g.Tout);
    const int b = (int)(row / g.Tout);
    const float* ib = in + (int64_t)b * g.Tin * g.Fin;
    F8 acc;
#pragma unroll
    for (int j = 0; j < 8; ++j) acc.v[j] = 0.f;
    for (int it = 0; it < g.kt; ++it) {
      const int t = to * g.st + it - g.pt;
      if (t < 0 || t >= g.Tin) continue;
      for (int jf = 0; jf < g.kf; ++jf) {
        const int f = fo * g.sf + jf - g.pf;
        if (f < 0 || f >= g.Fin) continue;
        const float xv = ib[(int64_t)t * g.Fin + f];
        const float* wp = sw + (it * g.kf + jf) * g.Cout + cg * 8;
#pragma unroll
        for (int j = 0; j < 8; ++j) acc.v[j] = fmaf(xv, wp[j], acc.v[j]);
      }
    }
    *(uint4*)(out + orow * g.ld_out + cg * 8) = pack8(acc);
  }
}

int stem_conv_fwd_bf16(const float* in, const float* w, bf16* out, const asf_conv_geom_t* g, cudaStream_t st) {
  const int64_t nvec = (int64_t)g->B * g->Tout * g->Fout * (g->Cout / 8);
  const size_t smem = sizeof(float) * (size_t)g->kt * g->kf * g->Cout;
  ASF_CUDA(launch_pdl(stem_conv_fwd_kernel, dim3(vgrid(nvec)), dim3(256), smem, st, in, w, out, *g, nvec));
  ASF_LAUNCH_CHECK();
  return 0;
}

// dwp[block][n][it*kf + jf] = sum over the block's rows of dout[row][n] * in[row shifted by tap]  (stem, C_in = 1).
// A WARP owns one (8-channel group, kernel row, row slice) task, lanes walk consecutive rows (dout loads of a warp are
// contiguous, two rows per lane in flight, 32-bit index arithmetic), the 8 x kf accumulators are folded by xor-shuffles
// (a fixed tree) and lane 0 stores them into the task's own shared-memory slot; the slices of a group are then added in
// slice order and the block STORES its partial sums -- no atomics, bit-reproducible; the cross-block sum is
// asf_unpack_wgrad(_multi)'s fixed-order split-K reduction.
constexpr int STEM_KF_MAX = 8;
__global__ void __launch_bounds__(256) stem_wgrad_v9_kernel(const float* __restrict__ in, const bf16* __restrict__ dout,
                                                            float* __restrict__ dwp, asf_conv_geom_t g, int rows_per_block,
                                                            int nslice) {
  pdl_trigger();
  pdl_wait();
  extern __shared__ float sacc[];  // [nslice][ngroups][8][STEM_KF_MAX]
  const int taps = g.kt * g.kf, CG = g.Cout / 8;
  const int ngroups = CG * g.kt;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rows = g.B * g.Tout * g.Fout;
  const int r0 = blockIdx.x * rows_per_block;
  const int r1 = min(r0 + rows_per_block, rows);
  const int slice_rows = (rows_per_block + nslice - 1) / nslice;
  const unsigned Fout = (unsigned)g.Fout, Tout = (unsigned)g.Tout;
  for (int task = warp; task < ngroups * nslice; task += 8) {
    const int grp = task % ngroups, sl = task / ngroups;
    const int cg = grp % CG, it = grp / CG;
    float acc[8][STEM_KF_MAX];
#pragma unroll
    for (int j = 0; j < 8; ++j)
#pragma unroll
      for (int k = 0; k < STEM_KF_MAX; ++k) acc[j][k] = 0.f;
    const int s0 = r0 + sl * slice_rows, s1 = min(s0 + slice_rows, r1);
    for (int base = s0 + lane; base < s1; base += 64) {
      uint4 dq[2];
      float xv[2][STEM_KF_MAX];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int row = base + 32 * h;
        dq[h] = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
        for (int k = 0; k < STEM_KF_MAX; ++k) xv[h][k] = 0.f;
        if (row < s1) {
          const unsigned r2 = (unsigned)row / Fout;
          const int fo = (int)((unsigned)row - r2 * Fout);
          const unsigned b = r2 / Tout;
          const int to = (int)(r2 - b * Tout);
          const int t = to * g.st + it - g.pt;
          if (t >= 0 && t < g.Tin) {
            dq[h] = *(const uint4*)(dout + (int64_t)row * g.ld_out + cg * 8);
            const float* ip = in + ((int64_t)b * g.Tin + t) * g.Fin;
            const int f0 = fo * g.sf - g.pf;
#pragma unroll
            for (int k = 0; k < STEM_KF_MAX; ++k) {
              const int f = f0 + k;
              if (k < g.kf && f >= 0 && f < g.Fin) xv[h][k] = ip[f];
            }
          }
        }
      }
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const F8 d = unpack8(dq[h]);
#pragma unroll
        for (int k = 0; k < STEM_KF_MAX; ++k)
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[j][k] = fmaf(d.v[j], xv[h][k], acc[j][k]);
      }
    }
    float* slot = sacc + (size_t)task * 8 * STEM_KF_MAX;  // task = sl * ngroups + grp
#pragma unroll
    for (int j = 0; j < 8; ++j)
#pragma unroll
      for (int k = 0; k < STEM_KF_MAX; ++k) {
        if (k < g.kf) {  // uniform across the warp
          float v = acc[j][k];
#pragma unroll
          for (int sft = 16; sft >= 1; sft >>= 1) v += __shfl_xor_sync(0xffffffffu, v, sft);
          if (lane == 0) slot[j * STEM_KF_MAX + k] = v;
        }
      }
  }
  __syncthreads();
  float* my = dwp + (size_t)blockIdx.x * g.Cout * taps;
  for (int i = threadIdx.x; i < g.Cout * taps; i += 256) {
    const int n = i / taps, tp = i - n * taps;
    const int it = tp / g.kf, k = tp - it * g.kf;
    const int grp = it * CG + (n >> 3), j = n & 7;
    float t = 0.f;
    for (int sl = 0; sl < nslice; ++sl) t += sacc[((size_t)(sl * ngroups + grp) * 8 + j) * STEM_KF_MAX + k];
    my[i] = t;
  }
}

// Stem weight gradient, third layout ("v10"): the last kernel of the backward pass, i.e. the serial tail of the step.
// v9 read the input straight from global memory (7 scalar loads + one 16-byte load per row and kernel row, two rows in
// flight per lane: 9 % of the FMA rate, 74 us for the fast stem).  Here a block walks chunks of SW10_LC output lines of one
// clip: the (SW10_LC - 1) * st + kt input lines a chunk touches are staged once in shared memory WITH their zero padding
// (no bounds tests in the inner loop), every warp owns one (8-channel group, kernel row) pair and a share of the
// chunk's lines for the whole launch -- its 8 x kf accumulators live in registers across chunks -- and the only global
// load in the inner loop, the row's 16 bytes of dout, is fetched four iterations ahead; the next chunk's lines arrive
// by cp.async into the other buffer while this chunk is being consumed (one barrier per chunk).
// With dw != nullptr the cross-block sum is folded in: blocks store their partial rows, the block that draws the last
// ticket adds them in block order (stats_reduce.cuh) and writes (or accumulates into) the gradient -- no separate
// split-K reduction launch behind the last kernel of the step.  Bit-reproducible either way.
constexpr int SW10_LC = 8;
template <int THREADS, int MINB, int DQ>
__global__ void __launch_bounds__(THREADS, MINB) stem_wgrad_v10_kernel(const float* __restrict__ in, const bf16* __restrict__ dout,
                                                             float* __restrict__ dwp, float* __restrict__ dw,
                                                             int accumulate, unsigned* __restrict__ ticket,
                                                             asf_conv_geom_t g, int nsl, int NL, int W, int cpc,
                                                             int n_chunks) {
  pdl_trigger();
  pdl_wait();
  extern __shared__ __align__(16) float s10[];  // staged input lines [NL][W]; later task slots, then reduction scratch
  const int taps = g.kt * g.kf, CG = g.Cout / 8, ngroups = CG * g.kt;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const int grp = warp % ngroups, sl = warp / ngroups;
  const int cg = grp % CG, it = grp / CG;
  const int lines_per_sl = SW10_LC / nsl;
  const int fiters = (g.Fout + 31) >> 5;
  const int fl = (fiters & (fiters - 1)) == 0 ? 31 - __clz(fiters) : -1;  // log2(fiters) when it is a power of two
  const int nq = lines_per_sl * fiters;
  float acc[8][STEM_KF_MAX];
#pragma unroll
  for (int j = 0; j < 8; ++j)
#pragma unroll
    for (int k = 0; k < STEM_KF_MAX; ++k) acc[j][k] = 0.f;
  // staging of one chunk's input lines (with their zero padding) into buffer `buf`: 4-byte cp.async, zero-filled outside
  auto stage = [&](int chunk, int buf) {
    if (chunk < n_chunks) {
      const int b = chunk / cpc, to0 = (chunk - b * cpc) * SW10_LC;
      const int t0 = to0 * g.st - g.pt;
      float* dstb = s10 + (size_t)buf * NL * W;
      for (int l = warp; l < NL; l += nwarps) {
        const int t = t0 + l;
        const bool tv = t >= 0 && t < g.Tin;
        const float* src = in + ((int64_t)b * g.Tin + (tv ? t : 0)) * g.Fin;
        for (int i = lane; i < W; i += 32) {
          const int f = i - g.pf;
          const bool v = tv && f >= 0 && f < g.Fin;
          asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"((uint32_t)__cvta_generic_to_shared(dstb + l * W + i)), "l"(src + (v ? f : 0)),
                       "r"(v ? 4 : 0)
                       : "memory");
        }
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  stage(blockIdx.x, 0);
  int cbuf = 0;
  for (int chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x, cbuf ^= 1) {
    const int b = chunk / cpc, to0 = (chunk - b * cpc) * SW10_LC;
    // this warp's rows of the chunk: row q -> output line sl * lines_per_sl + q / fiters, column 32 (q % fiters) + lane
    // (fiters is a power of two for every shipped stem: shifts instead of divisions in the inner loop)
    const bf16* dbase = dout + ((int64_t)b * g.Tout + to0) * g.Fout * g.ld_out + cg * 8;
    const int ld_o = (int)g.ld_out;
    auto row_meta = [&](int q, int& line, int& fo) -> bool {
      const int ql = fl >= 0 ? (q >> fl) : q / fiters;
      line = sl * lines_per_sl + ql;
      fo = (q - ql * fiters) * 32 + lane;
      return q < nq && to0 + line < g.Tout && fo < g.Fout;
    };
    auto row_load = [&](int q) -> uint4 {
      int line, fo;
      if (!row_meta(q, line, fo)) return make_uint4(0u, 0u, 0u, 0u);
      return *(const uint4*)(dbase + (int64_t)((line * g.Fout + fo) * ld_o));
    };
    uint4 dq[DQ];  // DQ rows of dout in flight per lane, the first ones across the staging barrier
#pragma unroll
    for (int u = 0; u < DQ; ++u) dq[u] = row_load(u);
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();  // this chunk's lines are visible; every warp is done with the other buffer
    stage(chunk + gridDim.x, cbuf ^ 1);
    const float* sbuf = s10 + (size_t)cbuf * NL * W;
    for (int q0 = 0; q0 < nq; q0 += DQ) {
#pragma unroll
      for (int u = 0; u < DQ; ++u) {
        const uint4 dcur = dq[u];
        dq[u] = row_load(q0 + u + DQ);
        int line_c, fo_c;
        if (!row_meta(q0 + u, line_c, fo_c)) continue;  // (zeros would add nothing)
        const float* xs = sbuf + (line_c * g.st + it) * W + fo_c * g.sf;
        float xv[STEM_KF_MAX];
#pragma unroll
        for (int k = 0; k < STEM_KF_MAX; ++k) xv[k] = k < g.kf ? xs[k] : 0.f;
        const F8 d = unpack8(dcur);
#pragma unroll
        for (int k = 0; k < STEM_KF_MAX; ++k)
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[j][k] = fmaf(d.v[j], xv[k], acc[j][k]);
      }
    }
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();  // the staged lines are dead: their space becomes the per-warp task slots [nwarps][8][STEM_KF_MAX]
#pragma unroll
  for (int j = 0; j < 8; ++j)
#pragma unroll
    for (int k = 0; k < STEM_KF_MAX; ++k) {
      if (k < g.kf) {  // uniform across the warp
        float v = acc[j][k];
#pragma unroll
        for (int sft = 16; sft >= 1; sft >>= 1) v += __shfl_xor_sync(0xffffffffu, v, sft);
        if (lane == 0) s10[(warp * 8 + j) * STEM_KF_MAX + k] = v;
      }
    }
  __syncthreads();
  const int n_out = g.Cout * taps;
  float* my = dwp + (size_t)blockIdx.x * n_out;
  for (int i = threadIdx.x; i < n_out; i += blockDim.x) {
    const int n = i / taps, tp = i - n * taps;
    const int it_ = tp / g.kf, k = tp - it_ * g.kf;
    const int gq = it_ * CG + (n >> 3), j = n & 7;
    float t = 0.f;
    for (int s_ = 0; s_ < nsl; ++s_) t += s10[((s_ * ngroups + gq) * 8 + j) * STEM_KF_MAX + k];  // slices in order
    my[i] = t;
  }
  if (dw == nullptr) return;
  __shared__ unsigned s_flag;
  if (!stats_last_cta(ticket, gridDim.x, &s_flag)) return;
  stats_ordered_reduce(dwp, (int)gridDim.x, n_out, n_out, (double*)s10, [&](int s_, double v) {
    dw[s_] = accumulate ? dw[s_] + (float)v : (float)v;
  });
}

static void stem_v9_geom(const asf_conv_geom_t* g, int* nslice, int64_t* blocks, int64_t* rpb) {
  const int64_t rows = (int64_t)g->B * g->Tout * g->Fout;
  const int ngroups = (g->Cout / 8) * g->kt;
  // (group, slice) tasks are dealt round robin to the 8 warps of a block: the smallest slice count that balances them
  int ns = 1;
  while ((ngroups * ns) % 8) ++ns;
  int64_t b = 148 * 2;
  int64_t r = cdiv(rows, b);
  const int64_t min_rpb = 64LL * 4 * ns;  // at least four double-row rounds per task
  if (r < min_rpb) r = min_rpb;
  *nslice = ns;
  *rpb = r;
  *blocks = cdiv(rows, r);
}

int stem_wgrad_v9_blocks(const asf_conv_geom_t* g) {
  int ns;
  int64_t blocks, rpb;
  stem_v9_geom(g, &ns, &blocks, &rpb);
  return (int)blocks;
}

int stem_wgrad_v9(const float* in, const bf16* dout, float* dwp, const asf_conv_geom_t* g, cudaStream_t st) {
  const int64_t rows = (int64_t)g->B * g->Tout * g->Fout;
  ASF_CHECK_ARG(rows < (1LL << 31) && (int64_t)g->B * g->Tin * g->Fin < (1LL << 31), "asf_stem_wgrad: tensor too large");
  const int ngroups = (g->Cout / 8) * g->kt;
  int nslice;
  int64_t blocks, rpb;
  stem_v9_geom(g, &nslice, &blocks, &rpb);
  const size_t smem = sizeof(float) * (size_t)nslice * ngroups * 8 * STEM_KF_MAX;
  ASF_CHECK_ARG(smem <= 48 * 1024, "asf_stem_wgrad: %zu bytes of task slots exceed 48 KB", smem);
  ASF_CUDA(launch_pdl(stem_wgrad_v9_kernel, dim3((unsigned)blocks), dim3(256), smem, st, in, dout, dwp, *g, (int)rpb,
                      nslice));
  ASF_LAUNCH_CHECK();
  return 0;
}

// v10 launch geometry; rows == 0: this shape stays on v9 (more than 16 (group, kernel row) pairs, or lines too long)
struct StemV10 {
  int nsl, NL, W, cpc, n_chunks, threads, grid, per_sm;
  size_t smem;
};
static bool stem_v10_geom(const asf_conv_geom_t* g, StemV10* v) {
  const int ngroups = (g->Cout / 8) * g->kt;
  if (g->Cout % 8 || g->kf > STEM_KF_MAX || ngroups < 1 || ngroups > 16 || (g->Cout * g->kt * g->kf) % 4) return false;
  int nsl = 1;
  while (nsl * 2 <= SW10_LC && ngroups * nsl * 2 <= 16) nsl *= 2;
  v->nsl = nsl;
  v->NL = (SW10_LC - 1) * g->st + g->kt;
  v->W = (g->Fout - 1) * g->sf + g->kf;
  v->cpc = (int)cdiv(g->Tout, SW10_LC);
  const int64_t chunks = (int64_t)g->B * v->cpc;
  if (chunks >= (1LL << 31)) return false;
  v->n_chunks = (int)chunks;
  v->threads = 32 * ngroups * nsl;
  size_t bytes = 2 * sizeof(float) * (size_t)v->NL * v->W;  // double buffered
  const size_t slots = sizeof(float) * (size_t)(v->threads / 32) * 8 * STEM_KF_MAX;
  const size_t scratch = sizeof(double) * 4 * (size_t)v->threads;
  if (slots > bytes) bytes = slots;
  if (scratch > bytes) bytes = scratch;
  if (bytes > 48 * 1024) return false;
  v->smem = bytes;
  // two blocks per SM when a block is <= 320 threads (the <320, 2, 2> instantiation: <= 102 registers); ASF_STEM_V10_PER_SM=1
  // keeps one
  static int per_sm_env = -1;
  if (per_sm_env < 0) {
    const char* e = getenv("ASF_STEM_V10_PER_SM");
    per_sm_env = e ? atoi(e) : 2;
  }
  v->per_sm = (v->threads <= 320 && per_sm_env == 2) ? 2 : 1;
  v->grid = (int)(chunks < 148 * v->per_sm ? chunks : 148 * v->per_sm);
  return true;
}

int stem_wgrad_v10_rows(const asf_conv_geom_t* g) {
  StemV10 v;
  return stem_v10_geom(g, &v) ? v.grid : 0;
}

int stem_wgrad_v10(const float* in, const bf16* dout, float* dwp, float* dw, int accumulate, unsigned* ticket,
                   const asf_conv_geom_t* g, cudaStream_t st) {
  StemV10 v;
  ASF_CHECK_ARG(stem_v10_geom(g, &v), "asf_stem_wgrad_fused: geometry not supported (asf_stem_wgrad_fused_rows == 0)");
  ASF_CHECK_ARG((int64_t)g->B * g->Tout * g->Fout < (1LL << 31) && (int64_t)g->B * g->Tin * g->Fin < (1LL << 31),
                "asf_stem_wgrad_fused: tensor too large");
  if (v.threads <= 320 && v.per_sm == 2) {
    ASF_CUDA(launch_pdl(stem_wgrad_v10_kernel<320, 2, 2>, dim3((unsigned)v.grid), dim3((unsigned)v.threads), v.smem, st, in, dout,
                        dwp, dw, accumulate, ticket, *g, v.nsl, v.NL, v.W, v.cpc, v.n_chunks));
  } else {
    ASF_CUDA(launch_pdl(stem_wgrad_v10_kernel<512, 1, 4>, dim3((unsigned)v.grid), dim3((unsigned)v.threads), v.smem, st, in, dout,
                        dwp, dw, accumulate, ticket, *g, v.nsl, v.NL, v.W, v.cpc, v.n_chunks));
  }
  ASF_LAUNCH_CHECK();
  return 0;
}

}  // namespace asf
