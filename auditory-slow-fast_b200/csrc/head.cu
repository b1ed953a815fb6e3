// The classifier head and the loss as a handful of launches.
//
// Between the last convolution of the forward pass and the first of the backward pass the training step is a serial
// chain of tiny operations -- dropout, one fully connected layer per task, two cross-entropies, their gradients, three
// small GEMMs and two bias sums: ~40 launches, ~200 us during which the GPU runs one small kernel at a time
// (profiles/r02_ncu_step_frozen.txt).  Here they are four:
//   dropout_apply_kernel   feat * (u < keep) / keep, mask kept for the backward pass
//   head_fwd_kernel        logits of BOTH task heads (+ bias) in one launch
//   softmax_ce_kernel      both cross-entropies (mean over the batch), their mean, and d loss / d logits
//   head_bwd_kernel        dW, db of both heads and d feat (dropout mask applied) in one launch
//
// Replaces   ResNetBasicHead.forward: dropout + nn.Linear per task     /root/reference/audio_slowfast/models/head_helper.py:70-109
//            compute_loss: F.cross_entropy(verb), F.cross_entropy(noun), mean  tools/train_utils.py:76-101
//            and their autograd backward.
// All sums run in a fixed order (warp trees / ordered loops): results are bit-reproducible.
#include "common.cuh"
#include "stats_reduce.cuh"

namespace asf {

struct HeadDesc {
  const float* W[2];  // [N][K] row-major fp32 masters (nn.Linear.weight)
  const float* bias[2];
  float* out[2];      // forward: logits [B][ld_out]
  int64_t ld_out[2];
  int N[2];
  int heads;
};

__global__ void __launch_bounds__(256) dropout_apply_kernel(const float* __restrict__ x, const float* __restrict__ u,
                                                            float* __restrict__ y, float* __restrict__ mask, int64_t n,
                                                            float keep, float inv_keep) {
  pdl_trigger();
  pdl_wait();
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const float m = u[i] < keep ? 1.f : 0.f;
    mask[i] = m;
    y[i] = x[i] * m * inv_keep;
  }
}

// One CTA = HF_JT consecutive classes of one head x HF_BG samples (blockIdx.y).  The JT weight rows sit in shared
// memory; a warp takes HF_SW samples at once and its lanes stride over K in float4 steps, so that HF_SW x 6 independent
// 16-byte loads are in flight per lane (the whole layer is 58 MFLOP: it is latency, not arithmetic, that has to be hidden).
constexpr int HF_JT = 4, HF_SW = 2, HF_BG = 8 * HF_SW;  // 8 warps x 2 samples per CTA
__global__ void __launch_bounds__(256) head_fwd_kernel(HeadDesc d, const float* __restrict__ feat, int64_t ld_feat, int B,
                                                       int K, int tiles0) {
  extern __shared__ float4 hf_w4[];  // [JT][K / 4]
  pdl_trigger();
  const int h = (int)blockIdx.x < tiles0 ? 0 : 1;
  const int j0 = ((int)blockIdx.x - (h ? tiles0 : 0)) * HF_JT;
  const int N = d.N[h];
  const int K4 = K >> 2;
  const float4* __restrict__ W4 = (const float4*)d.W[h];
  for (int i0 = 0; i0 < HF_JT * K4; i0 += 8 * 256) {  // parameters; 8 independent loads in flight per thread
    float4 v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int i = i0 + u * 256 + (int)threadIdx.x;
      const int jj = i / K4, k = i - jj * K4;
      v[u] = (i < HF_JT * K4 && j0 + jj < N) ? W4[(int64_t)(j0 + jj) * K4 + k] : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int i = i0 + u * 256 + (int)threadIdx.x;
      if (i < HF_JT * K4) hf_w4[i] = v[u];
    }
  }
  pdl_wait();
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b0 = (int)blockIdx.y * HF_BG + warp * HF_SW;
  if (b0 >= B) return;
  const float4* f4[HF_SW];
#pragma unroll
  for (int s = 0; s < HF_SW; ++s) f4[s] = (const float4*)(feat + (int64_t)min(b0 + s, B - 1) * ld_feat);
  float acc[HF_SW][HF_JT];
#pragma unroll
  for (int s = 0; s < HF_SW; ++s)
#pragma unroll
    for (int jj = 0; jj < HF_JT; ++jj) acc[s][jj] = 0.f;
#pragma unroll 6
  for (int k = lane; k < K4; k += 32) {
    float4 x[HF_SW];
#pragma unroll
    for (int s = 0; s < HF_SW; ++s) x[s] = f4[s][k];
#pragma unroll
    for (int jj = 0; jj < HF_JT; ++jj) {
      const float4 w = hf_w4[jj * K4 + k];
#pragma unroll
      for (int s = 0; s < HF_SW; ++s)
        acc[s][jj] = fmaf(x[s].x, w.x, fmaf(x[s].y, w.y, fmaf(x[s].z, w.z, fmaf(x[s].w, w.w, acc[s][jj]))));
    }
  }
#pragma unroll
  for (int s = 0; s < HF_SW; ++s)
#pragma unroll
    for (int jj = 0; jj < HF_JT; ++jj) acc[s][jj] = warp_sum(acc[s][jj]);
  if (lane < HF_SW * HF_JT) {
    const int s = lane / HF_JT, jj = lane % HF_JT;
    float v = 0.f;
#pragma unroll
    for (int a = 0; a < HF_SW; ++a)
#pragma unroll
      for (int c = 0; c < HF_JT; ++c) v = (a == s && c == jj) ? acc[a][c] : v;
    if (b0 + s < B && j0 + jj < N)
      d.out[h][(int64_t)(b0 + s) * d.ld_out[h] + j0 + jj] = v + (d.bias[h] ? d.bias[h][j0 + jj] : 0.f);
  }
}

// Single CTA (B x (N0 + N1) numbers): a warp per sample row and head.  *loss = mean over heads of the batch-mean
// cross-entropies, loss_heads[h] = head h's; dl_h[b][j] = (softmax_j - [j == label]) / (B * heads).
struct CeDesc {
  const float* logits[2];
  const int64_t* labels[2];
  float* dl[2];
  int64_t ld[2], ld_dl[2];
  int N[2];
  int heads;
};
constexpr int CE_THREADS = 256, CE_REG = 16;  // a lane keeps up to 16 logits of its row: rows up to 512 classes in registers
// A warp per (head, sample) row, 8 rows per CTA; per-row losses go to `scratch`, the CTA drawing the last ticket adds them
// in index order (csrc/stats_reduce.cuh).
__global__ void __launch_bounds__(CE_THREADS) softmax_ce_kernel(CeDesc d, int B, float* __restrict__ loss,
                                                                float* __restrict__ loss_heads,
                                                                float* __restrict__ ce_rows, unsigned* __restrict__ ticket) {
  __shared__ unsigned s_flag;
  pdl_trigger();
  pdl_wait();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = (blockDim.x >> 5) * gridDim.x;
  const float gscale = 1.f / ((float)B * (float)d.heads);
  for (int r = blockIdx.x * (blockDim.x >> 5) + warp; r < B * d.heads; r += nwarps) {
    const int h = r / B, b = r - h * B;
    const int N = d.N[h];
    const float* __restrict__ x = d.logits[h] + (int64_t)b * d.ld[h];
    const int label = (int)d.labels[h][b];
    const bool ok = label >= 0 && label < N;
    float* __restrict__ g = d.dl[h] ? d.dl[h] + (int64_t)b * d.ld_dl[h] : nullptr;
    if (N <= 32 * CE_REG) {  // one pass over global memory
      float v[CE_REG];
      float mx = -INFINITY;
#pragma unroll
      for (int i = 0; i < CE_REG; ++i) {
        const int j = lane + 32 * i;
        v[i] = j < N ? x[j] : -INFINITY;
        mx = fmaxf(mx, v[i]);
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      float se = 0.f;
#pragma unroll
      for (int i = 0; i < CE_REG; ++i) se += (lane + 32 * i < N) ? expf(v[i] - mx) : 0.f;
      se = warp_sum(se);
      const float lse = mx + logf(se);
      float xl = 0.f;
#pragma unroll
      for (int i = 0; i < CE_REG; ++i) xl = (lane + 32 * i == label) ? v[i] : xl;
      xl = warp_sum(xl);  // exactly one lane holds the label's logit
      if (lane == 0) ce_rows[h * B + b] = ok ? lse - xl : 0.f;
      if (g) {
#pragma unroll
        for (int i = 0; i < CE_REG; ++i) {
          const int j = lane + 32 * i;
          if (j < N) g[j] = ok ? (expf(v[i] - lse) - (j == label ? 1.f : 0.f)) * gscale : 0.f;
        }
      }
    } else {
      float mx = -INFINITY;
      for (int j = lane; j < N; j += 32) mx = fmaxf(mx, x[j]);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      float se = 0.f;
      for (int j = lane; j < N; j += 32) se += expf(x[j] - mx);
      se = warp_sum(se);
      const float lse = mx + logf(se);
      if (lane == 0) ce_rows[h * B + b] = ok ? lse - x[label] : 0.f;
      if (g)
        for (int j = lane; j < N; j += 32) g[j] = ok ? (expf(x[j] - lse) - (j == label ? 1.f : 0.f)) * gscale : 0.f;
    }
  }
  if (!stats_last_cta(ticket, gridDim.x, &s_flag)) return;
  if (warp < d.heads) {  // per-head batch means: fixed tree (lanes stride the batch, then the shuffle tree)
    float s = 0.f;
    for (int b = lane; b < B; b += 32) s += __ldcg(ce_rows + warp * B + b);
    s = warp_sum(s);
    if (lane == 0) loss_heads[warp] = s / (float)B;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int h = 0; h < d.heads; ++h) s += loss_heads[h];
    *loss = s / (float)d.heads;
  }
}

// One CTA = HB_CT (16) consecutive feature columns and one of the two jobs (blockIdx.y: 0 = dW and db, 1 = dfeat), so
// that two CTAs share an SM; samples in chunks of 64 (dW) / 32 (dfeat).  Shared memory: dl of both heads
// for the chunk [HB_BC][NT], the chunk's feat columns [HB_BC][HB_CT] and the CTA's weight columns [NT][HB_CT] -- the
// inner loops touch shared memory only, with register tiles so that they are FMA- and not load-bound:
//   dW[j][c]    (+)= sum_b dl[b][j] feat[b][c]     thread = (4 columns) x (HB_JR consecutive classes), 64 class groups
//   dfeat[b][c]   = m[b][c] inv_keep sum_j dl[b][j] W[j][c]          thread = (4 columns) x (one sample of the chunk)
//   db[j]       (+)= sum_b dl[b][j]                                   (CTA 0)
constexpr int HB_CT = 16, HB_BC_W = 64, HB_BC_F = 32, HB_JR = 8, HB_JG = 64;  // HB_JR * HB_JG = 512 classes per pass
struct HeadBwdDesc {
  const float* dl[2];
  int64_t ld_dl[2];
  const float* W[2];
  float* dW[2];
  float* db[2];
  int N[2];
  int acc_w[2], acc_b[2];
  int heads;
};
__global__ void __launch_bounds__(256) head_bwd_kernel(HeadBwdDesc d, const float* __restrict__ feat, int64_t ld_feat,
                                                       const float* __restrict__ mask, float inv_keep, int B, int K,
                                                       float* __restrict__ dfeat, int64_t ld_dfeat) {
  extern __shared__ float4 hb_s4[];
  float* hb_s = (float*)hb_s4;
  pdl_trigger();
  const int N0 = d.N[0];
  const int NT = N0 + (d.heads > 1 ? d.N[1] : 0);
  const int tid = threadIdx.x;
  const int cg = tid & 3, rg = tid >> 2;      // 4 column groups of 4 columns; 64 row groups (classes / samples x halves)
  const bool job_w = blockIdx.y == 0, job_f = blockIdx.y == 1 && dfeat != nullptr;
  float* s_f = hb_s;                          // [64][HB_CT]   (16-byte aligned rows)
  float* s_w = s_f + HB_BC_W * HB_CT;         // [NT][HB_CT]  (dfeat job only; the dW job's sample rows start here)
  float* s_dl = job_w ? s_w : s_w + NT * HB_CT;  // [BC][NT]
  const int c0 = blockIdx.x * HB_CT;
  const int BC = job_w ? HB_BC_W : HB_BC_F;   // samples per chunk
  // the CTA's weight columns (parameters: not produced by the previous kernel); 8 independent loads in flight per thread
  if (job_f) {
    for (int h = 0; h < d.heads; ++h) {
      const float* __restrict__ Wh = d.W[h];
      const int Nh = d.N[h], joff = h ? N0 : 0;
      for (int i0 = 0; i0 < Nh * HB_CT; i0 += 8 * 256) {
        float v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int i = i0 + u * 256 + tid;
          const int j = i / HB_CT, cc = i - j * HB_CT;
          v[u] = (i < Nh * HB_CT && c0 + cc < K) ? Wh[(int64_t)j * K + c0 + cc] : 0.f;
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int i = i0 + u * 256 + tid;
          if (i < Nh * HB_CT) s_w[joff * HB_CT + i] = v[u];
        }
      }
    }
  }
  pdl_wait();
  for (int b0 = 0; b0 < B; b0 += BC) {
    const int nb = min(BC, B - b0);
    __syncthreads();
    // a warp copies sample rows (zeros past the batch), 8 independent loads in flight per lane
    for (int h = 0; h < d.heads; ++h) {
      const float* __restrict__ dlh = d.dl[h];
      const int64_t ldh = d.ld_dl[h];
      const int Nh = d.N[h], joff = h ? N0 : 0;
      for (int b = tid >> 5; b < BC; b += 8) {
        const float* __restrict__ src = dlh + (int64_t)(b0 + b) * ldh;
        float* dst = s_dl + b * NT + joff;
        for (int jb = 0; jb < Nh; jb += 8 * 32) {
          float v[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const int j = jb + u * 32 + (tid & 31);
            v[u] = (b < nb && j < Nh) ? src[j] : 0.f;
          }
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const int j = jb + u * 32 + (tid & 31);
            if (j < Nh) dst[j] = v[u];
          }
        }
      }
    }
    if (job_w) {
      for (int i0 = 0; i0 < BC * HB_CT; i0 += 4 * 256) {
        float v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int i = i0 + u * 256 + tid;
          const int b = i / HB_CT, cc = i - b * HB_CT;
          v[u] = (i < BC * HB_CT && b < nb && c0 + cc < K) ? feat[(int64_t)(b0 + b) * ld_feat + c0 + cc] : 0.f;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int i = i0 + u * 256 + tid;
          if (i < BC * HB_CT) s_f[i] = v[u];
        }
      }
    }
    __syncthreads();
    // ---- dW
    for (int jp = 0; job_w && jp < NT; jp += HB_JR * HB_JG) {
      const int j0 = jp + rg * HB_JR;
      if (j0 < NT) {
        float acc[HB_JR][4];
#pragma unroll
        for (int i = 0; i < HB_JR; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.f;
        for (int b = 0; b < HB_BC_W; ++b) {
          const float4 f = *(const float4*)(s_f + b * HB_CT + 4 * cg);
          const float* row = s_dl + b * NT + j0;
#pragma unroll
          for (int i = 0; i < HB_JR; ++i) {
            const float g = (j0 + i < NT) ? row[i] : 0.f;
            acc[i][0] = fmaf(g, f.x, acc[i][0]);
            acc[i][1] = fmaf(g, f.y, acc[i][1]);
            acc[i][2] = fmaf(g, f.z, acc[i][2]);
            acc[i][3] = fmaf(g, f.w, acc[i][3]);
          }
        }
#pragma unroll
        for (int i = 0; i < HB_JR; ++i) {
          const int j = j0 + i;
          if (j >= NT) break;
          const int h = j < N0 ? 0 : 1;
          if (!d.dW[h]) continue;
          float* dst = d.dW[h] + (int64_t)(j - (h ? N0 : 0)) * K + c0 + 4 * cg;
          const bool add = b0 > 0 || d.acc_w[h];
#pragma unroll
          for (int q = 0; q < 4; ++q)
            if (c0 + 4 * cg + q < K) dst[q] = add ? dst[q] + acc[i][q] : acc[i][q];
        }
      }
    }
    // ---- dfeat: sample (rg & 31) of the chunk, 4 columns, half (rg >> 5) of the classes; halves added in order
    if (job_f) {
      const int sb = rg & (HB_BC_F - 1), half = rg >> 5;
      const int jm = (NT + 1) >> 1;
      const int ja = half ? jm : 0, jz = half ? NT : jm;
      float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
      const float* row = s_dl + sb * NT;
#pragma unroll 4
      for (int j = ja; j < jz; ++j) {
        const float4 w = *(const float4*)(s_w + j * HB_CT + 4 * cg);
        const float g = row[j];
        a0 = fmaf(g, w.x, a0);
        a1 = fmaf(g, w.y, a1);
        a2 = fmaf(g, w.z, a2);
        a3 = fmaf(g, w.w, a3);
      }
      __syncthreads();  // every thread is done with s_f's readers (none in this job): reuse it for the upper halves
      if (half) *(float4*)(s_f + sb * HB_CT + 4 * cg) = make_float4(a0, a1, a2, a3);
      __syncthreads();
      if (!half && sb < nb) {
        const float4 u = *(const float4*)(s_f + sb * HB_CT + 4 * cg);
        const float v[4] = {a0 + u.x, a1 + u.y, a2 + u.z, a3 + u.w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int c = c0 + 4 * cg + q;
          if (c < K) {
            float x = v[q];
            if (mask) x *= mask[(int64_t)(b0 + sb) * ld_feat + c] * inv_keep;
            dfeat[(int64_t)(b0 + sb) * ld_dfeat + c] = x;
          }
        }
      }
    }
    if (blockIdx.x == 0 && job_w) {
      for (int j = tid; j < NT; j += blockDim.x) {
        const int h = j < N0 ? 0 : 1;
        if (!d.db[h]) continue;
        float acc = 0.f;
        for (int b = 0; b < nb; ++b) acc += s_dl[b * NT + j];
        float* dst = d.db[h] + (j - (h ? N0 : 0));
        *dst = (b0 > 0 || d.acc_b[h]) ? *dst + acc : acc;
      }
    }
  }
}

}  // namespace asf

using namespace asf;

extern "C" {

int asf_dropout_apply(const float* x, const float* u, float* y, float* mask, int64_t n, float keep, void* stream) {
  ASF_CHECK_ARG(x && u && y && mask && n > 0 && keep > 0.f && keep <= 1.f, "asf_dropout_apply: bad arguments");
  ASF_CUDA(launch_pdl(dropout_apply_kernel, dim3((unsigned)cdiv(n, 256)), dim3(256), 0, (cudaStream_t)stream, x, u, y, mask, n,
                      keep, 1.f / keep));
  ASF_LAUNCH_CHECK();
  return 0;
}

int asf_head_fwd(const float* feat, int64_t ld_feat, int B, int K, int heads, const float* const* W,
                 const float* const* bias, const int* N, float* const* out, const int64_t* ld_out, void* stream) {
  ASF_CHECK_ARG(feat && W && N && out && ld_out && B > 0 && K > 0 && (heads == 1 || heads == 2) && ld_feat >= K,
                "asf_head_fwd: bad arguments");
  HeadDesc d = {};
  d.heads = heads;
  int tiles[2] = {0, 0};
  for (int h = 0; h < heads; ++h) {
    ASF_CHECK_ARG(W[h] && out[h] && N[h] > 0 && ld_out[h] >= N[h], "asf_head_fwd: head %d", h);
    d.W[h] = W[h];
    d.bias[h] = bias ? bias[h] : nullptr;
    d.out[h] = out[h];
    d.ld_out[h] = ld_out[h];
    d.N[h] = N[h];
    tiles[h] = (int)cdiv(N[h], HF_JT);
  }
  ASF_CHECK_ARG(K % 4 == 0 && ld_feat % 4 == 0, "asf_head_fwd: K and the feature pitch must be multiples of 4");
  const size_t smem = (size_t)HF_JT * K * sizeof(float);
  ASF_CHECK_ARG(smem <= 200 * 1024, "asf_head_fwd: K = %d too large", K);
  static size_t smem_set = 0;
  if (smem > smem_set) {
    ASF_CUDA(cudaFuncSetAttribute(head_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    smem_set = 200 * 1024;
  }
  ASF_CUDA(launch_pdl(head_fwd_kernel, dim3(tiles[0] + tiles[1], (unsigned)cdiv(B, HF_BG)), dim3(256), smem, (cudaStream_t)stream, d, feat, ld_feat, B, K,
                      tiles[0]));
  ASF_LAUNCH_CHECK();
  return 0;
}

int asf_softmax_ce(int heads, const float* const* logits, const int64_t* ld, const int* N, const int64_t* const* labels,
                   int B, float* loss, float* loss_heads, float* const* dl, const int64_t* ld_dl, float* row_scratch,
                   unsigned* ticket, void* stream) {
  ASF_CHECK_ARG(logits && ld && N && labels && loss && loss_heads && row_scratch && ticket && B > 0 && (heads == 1 || heads == 2), "asf_softmax_ce: bad arguments");
  CeDesc d = {};
  d.heads = heads;
  for (int h = 0; h < heads; ++h) {
    ASF_CHECK_ARG(logits[h] && labels[h] && N[h] > 0 && ld[h] >= N[h], "asf_softmax_ce: head %d", h);
    d.logits[h] = logits[h];
    d.labels[h] = labels[h];
    d.ld[h] = ld[h];
    d.N[h] = N[h];
    d.dl[h] = dl ? dl[h] : nullptr;
    d.ld_dl[h] = (dl && ld_dl) ? ld_dl[h] : 0;
    ASF_CHECK_ARG(!d.dl[h] || d.ld_dl[h] >= N[h], "asf_softmax_ce: gradient pitch of head %d", h);
  }
  const int64_t ctas = cdiv((int64_t)heads * B, CE_THREADS / 32);
  ASF_CUDA(launch_pdl(softmax_ce_kernel, dim3((unsigned)(ctas < 1184 ? ctas : 1184)), dim3(CE_THREADS), 0, (cudaStream_t)stream, d,
                      B, loss, loss_heads, row_scratch, ticket));
  ASF_LAUNCH_CHECK();
  return 0;
}

int asf_head_bwd(int heads, const float* const* dl, const int64_t* ld_dl, const int* N, const float* const* W,
                 const float* feat, int64_t ld_feat, const float* mask, float inv_keep, int B, int K, float* const* dW,
                 const int* acc_w, float* const* db, const int* acc_b, float* dfeat, int64_t ld_dfeat, void* stream) {
  ASF_CHECK_ARG(dl && ld_dl && N && W && feat && B > 0 && K > 0 && (heads == 1 || heads == 2) && ld_feat >= K,
                "asf_head_bwd: bad arguments");
  ASF_CHECK_ARG(!dfeat || ld_dfeat >= K, "asf_head_bwd: dfeat pitch");
  HeadBwdDesc d = {};
  d.heads = heads;
  int NT = 0;
  for (int h = 0; h < heads; ++h) {
    ASF_CHECK_ARG(dl[h] && W[h] && N[h] > 0 && ld_dl[h] >= N[h], "asf_head_bwd: head %d", h);
    d.dl[h] = dl[h];
    d.ld_dl[h] = ld_dl[h];
    d.W[h] = W[h];
    d.N[h] = N[h];
    d.dW[h] = dW ? dW[h] : nullptr;
    d.db[h] = db ? db[h] : nullptr;
    d.acc_w[h] = acc_w ? acc_w[h] : 0;
    d.acc_b[h] = acc_b ? acc_b[h] : 0;
    NT += N[h];
  }
  const size_t smem_w = ((size_t)HB_BC_W * NT + (size_t)HB_BC_W * HB_CT) * sizeof(float);
  const size_t smem_f = ((size_t)HB_BC_F * NT + (size_t)HB_BC_W * HB_CT + (size_t)NT * HB_CT) * sizeof(float);
  const size_t smem = smem_w > smem_f ? smem_w : smem_f;
  ASF_CHECK_ARG(smem <= 200 * 1024, "asf_head_bwd: %d classes do not fit", NT);
  static size_t smem_set = 0;
  if (smem > smem_set) {
    ASF_CUDA(cudaFuncSetAttribute(head_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    smem_set = 200 * 1024;
  }
  ASF_CUDA(launch_pdl(head_bwd_kernel, dim3((unsigned)cdiv(K, HB_CT), 2), dim3(256), smem, (cudaStream_t)stream, d, feat, ld_feat,
                      mask, inv_keep, B, K, dfeat, ld_dfeat));
  ASF_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
