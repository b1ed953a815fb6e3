// Forward / dgrad of the narrow layers (fast pathway: 8..64 channels at 10^5 output positions), included by
// tapgemm_sm100.cu inside namespace asf.  Same contract as tap-GEMM mode 0 with a bf16 output:
//
//   out[row(m)][n] = epi( sum_tap sum_c A[coord(m) + a_off[tap]][c] * Bw[n][b_tap[tap]][c] )
//
// At N, K <= 64 a 128-row tcgen05 tile carries 1..8 KB of operands: that kernel is bound by the TMA row rate of
// 16..128-byte rows and by its per-tile barrier round trips (20..55 us per launch where the data moves in < 5 us), and
// a CUDA-core FMA loop is bound by the shared-memory broadcast of the weights (measured: 69 % LSU data pipe, IPC 1.4).
// These layers are HBM/L2-bound streaming work with a tiny dense contraction per position, which is what warp-level
// mma.sync.m16n8k16 (bf16 in, fp32 accumulate) is for: a warp owns 16 consecutive output positions and all N channels.
//
// Operand feeding without any shuffle or ldmatrix: the contraction axis of the MMA may be permuted freely as long as
// A and B agree.  Flatten (tap, channel octet) into 16-byte segments; in step s lane (g = lane / 4, t = lane % 4) loads
// segment 4 s + t of rows g and g + 8 straight from global memory (one LDG.128 each, 64 contiguous bytes per row and
// quad) and the same segment of weight row n = 8 j + g from shared memory.  Registers x, y of a 16-byte vector feed
// MMA 1 (k slots 2t, 2t+1 and 2t+8, 2t+9), z, w feed MMA 2: both operands use the identical slot assignment, so the
// products pair up correctly.  Weights sit in shared memory in fragment order [step][j][lane] (conflict-free LDS.128).
#pragma once

constexpr int NC_THREADS = 128;
constexpr int NC_MAX_W = 8192;  // N * K * taps

struct NcParams {
  const bf16* a;
  const bf16* bw;
  bf16* out;
  long long ld_out;
  int a_dim[4];
  long long a_stride[4];
  int m_ext[3], m_to_a[3];
  long long o_stride[3];
  int ntaps, ntaps_w, N, K;
  int kv_log2;  // log2(K / 8)
  int segs;     // ntaps * K / 8
  int steps;    // ceil(segs / 4)
  int a_off[ASF_MAX_TAPS][4];
  long long tap_off[ASF_MAX_TAPS];  // sum_d a_off[tap][d] * a_stride[d]
  int b_tap[ASF_MAX_TAPS];
  int m_total;
  int tiles;  // ceil(m_total / 16)
  asf_epilogue_t epi;
};

__device__ __forceinline__ void nc_mma(float* d, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                       uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

__device__ __forceinline__ float2 nc_bf2(uint32_t v) {
  return make_float2(__uint_as_float(v << 16), __uint_as_float(v & 0xffff0000u));
}

// NO = N / 8 output-channel blocks per warp.
template <int NO, bool STATS>
__global__ void __launch_bounds__(NC_THREADS) narrow_conv_kernel(const __grid_constant__ NcParams p) {
  extern __shared__ __align__(16) uint8_t nc_smem[];
  pdl_trigger();
  uint4* wfrag = (uint4*)nc_smem;
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int steps = p.steps, K = p.K, KVm = (K >> 3) - 1;
  // weights are constant across the step: stage them before the PDL wait (safe under programmatic dependent launch
  // because the kernels that write packed weights never trigger early)
  const int G = steps * NO * 32;
  for (int i0 = 0; i0 < G; i0 += 4 * NC_THREADS) {
    uint4 q[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = i0 + u * NC_THREADS + threadIdx.x;
      q[u] = make_uint4(0u, 0u, 0u, 0u);
      if (i < G) {
        const int li = i & 31, j = (i >> 5) % NO, st = (i >> 5) / NO;
        const int seg = 4 * st + (li & 3), n = 8 * j + (li >> 2);
        if (seg < p.segs) {
          const int tap = seg >> p.kv_log2, v = seg & KVm;
          q[u] = *(const uint4*)(p.bw + ((long long)n * p.ntaps_w + p.b_tap[tap]) * K + v * 8);
        }
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = i0 + u * NC_THREADS + threadIdx.x;
      if (i < G) wfrag[i] = q[u];
    }
  }
  float* sstat = (float*)(wfrag + G);  // STATS: [4 warps][2][N] warp partial sums, then 4 x 128 doubles + a flag
  __syncthreads();
  pdl_wait();

  const int gwarp = blockIdx.x * (NC_THREADS / 32) + (threadIdx.x >> 5);
  const int nwarps = gridDim.x * (NC_THREADS / 32);
  const unsigned e0 = (unsigned)p.m_ext[0], e1 = (unsigned)p.m_ext[1];
  float2 shift[NO];
#pragma unroll
  for (int j = 0; j < NO; ++j)
    shift[j] = p.epi.shift ? *(const float2*)(p.epi.shift + 8 * j + 2 * t) : make_float2(0.f, 0.f);
  float ssum[NO][2], ssq[NO][2];
#pragma unroll
  for (int j = 0; j < NO; ++j) ssum[j][0] = ssum[j][1] = ssq[j][0] = ssq[j][1] = 0.f;

  for (int tile = gwarp; tile < p.tiles; tile += nwarps) {
    int coord[2][4];
    long long lin[2], orow[2];
    bool live[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const int pos = tile * 16 + g + 8 * r;
      live[r] = pos < p.m_total;
      const unsigned up = (unsigned)(live[r] ? pos : 0);
      const unsigned q1 = up / e0;
      const int m[3] = {(int)(up - q1 * e0), (int)(q1 % e1), (int)(q1 / e1)};
      lin[r] = 0;
#pragma unroll
      for (int d = 0; d < 4; ++d) {
        int c0 = 0;
#pragma unroll
        for (int jj = 0; jj < 3; ++jj)
          if (p.m_to_a[jj] == d + 1) c0 += m[jj];
        coord[r][d] = c0;
        lin[r] += (long long)c0 * p.a_stride[d];
      }
      orow[r] = m[0] * p.o_stride[0] + m[1] * p.o_stride[1] + m[2] * p.o_stride[2];
    }
    float acc[NO][4];
#pragma unroll
    for (int j = 0; j < NO; ++j) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f;

    auto load_seg = [&](int st, uint4* x) {
      const int seg = 4 * st + t;
      const bool sv = seg < p.segs;
      const int tap = sv ? seg >> p.kv_log2 : 0, v = seg & KVm;
      const long long off = p.tap_off[tap] + v * 8;
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        bool ok = sv && live[r];
#pragma unroll
        for (int d = 0; d < 4; ++d) ok = ok && (unsigned)(coord[r][d] + p.a_off[tap][d]) < (unsigned)p.a_dim[d];
        x[r] = ok ? *(const uint4*)(p.a + lin[r] + off) : make_uint4(0u, 0u, 0u, 0u);
      }
    };
    for (int s0 = 0; s0 < steps; s0 += 2) {
      uint4 x0[2], x1[2];
      load_seg(s0, x0);
      const bool two = s0 + 1 < steps;
      if (two) load_seg(s0 + 1, x1);
      const uint4* wf = wfrag + (s0 * NO) * 32 + lane;
#pragma unroll
      for (int j = 0; j < NO; ++j) {
        const uint4 w = wf[j * 32];
        nc_mma(acc[j], x0[0].x, x0[1].x, x0[0].y, x0[1].y, w.x, w.y);
        nc_mma(acc[j], x0[0].z, x0[1].z, x0[0].w, x0[1].w, w.z, w.w);
      }
      if (two) {
#pragma unroll
        for (int j = 0; j < NO; ++j) {
          const uint4 w = wf[(NO + j) * 32];
          nc_mma(acc[j], x1[0].x, x1[1].x, x1[0].y, x1[1].y, w.x, w.y);
          nc_mma(acc[j], x1[0].z, x1[1].z, x1[0].w, x1[1].w, w.z, w.w);
        }
      }
    }
    // epilogue on the accumulator fragment: lane holds channels 8 j + 2 t, + 1 of rows g (acc 0, 1) and g + 8 (acc 2, 3).
    // shift, residual / accumulate, relu, mask (the order of asf_epilogue_t), one rounding to bf16, statistics of the
    // rounded values
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      if (!live[r]) continue;
      uint32_t vres[NO], vacc[NO], vmask[NO];
#pragma unroll
      for (int j = 0; j < NO; ++j) {
        const int n = 8 * j + 2 * t;
        if (p.epi.res) vres[j] = *(const uint32_t*)((const bf16*)p.epi.res + orow[r] * p.epi.ld_res + n);
        if (p.epi.accumulate) vacc[j] = *(const uint32_t*)(p.out + orow[r] * p.ld_out + n);
        if (p.epi.mask) vmask[j] = *(const uint32_t*)((const bf16*)p.epi.mask + orow[r] * p.epi.ld_mask + n);
      }
#pragma unroll
      for (int j = 0; j < NO; ++j) {
        float v0 = acc[j][2 * r] + shift[j].x, v1 = acc[j][2 * r + 1] + shift[j].y;
        if (p.epi.res) {
          const float2 f = nc_bf2(vres[j]);
          v0 += f.x;
          v1 += f.y;
        }
        if (p.epi.accumulate) {
          const float2 f = nc_bf2(vacc[j]);
          v0 += f.x;
          v1 += f.y;
        }
        if (p.epi.relu) {
          v0 = fmaxf(v0, 0.f);
          v1 = fmaxf(v1, 0.f);
        }
        if (p.epi.mask) {
          const float2 f = nc_bf2(vmask[j]);
          v0 = f.x > 0.f ? v0 : 0.f;
          v1 = f.y > 0.f ? v1 : 0.f;
        }
        const __nv_bfloat162 o = __floats2bfloat162_rn(v0, v1);
        const uint32_t ob = *(const uint32_t*)&o;
        *(uint32_t*)(p.out + orow[r] * p.ld_out + 8 * j + 2 * t) = ob;
        if (STATS) {
          const float2 f = nc_bf2(ob);
          ssum[j][0] += f.x;
          ssum[j][1] += f.y;
          ssq[j][0] = fmaf(f.x, f.x, ssq[j][0]);
          ssq[j][1] = fmaf(f.y, f.y, ssq[j][1]);
        }
      }
    }
  }
  if (STATS) {
    // deterministic statistics: fixed xor-shuffle tree over the 8 row groups of a warp, one slot per warp in shared
    // memory, warps added in order, one workspace row per CTA, ordered final pass by the last CTA (stats_reduce.cuh)
    constexpr int N = 8 * NO;
    const int warp = threadIdx.x >> 5;
#pragma unroll
    for (int j = 0; j < NO; ++j) {
#pragma unroll
      for (int c = 0; c < 2; ++c) {
#pragma unroll
        for (int sft = 4; sft <= 16; sft <<= 1) {
          ssum[j][c] += __shfl_xor_sync(0xffffffffu, ssum[j][c], sft);
          ssq[j][c] += __shfl_xor_sync(0xffffffffu, ssq[j][c], sft);
        }
        if (g == 0) {
          sstat[warp * 2 * N + 8 * j + 2 * t + c] = ssum[j][c];
          sstat[warp * 2 * N + N + 8 * j + 2 * t + c] = ssq[j][c];
        }
      }
    }
    __syncthreads();
    float* my = p.epi.stats_ws + (long long)blockIdx.x * 2 * N;
    for (int i = threadIdx.x; i < 2 * N; i += NC_THREADS)
      my[i] = ((sstat[i] + sstat[2 * N + i]) + sstat[4 * N + i]) + sstat[6 * N + i];
    double* scratch = (double*)(sstat + 8 * N);
    unsigned* flag = (unsigned*)(scratch + 4 * NC_THREADS);
    if (stats_last_cta(p.epi.stats_ticket, gridDim.x, flag)) {
      double* out = p.epi.stats;
      const int stride = p.epi.stats_stride;
      stats_ordered_reduce(p.epi.stats_ws, (int)gridDim.x, 2 * N, 2 * N, scratch, [&](int s_, double v) {
        const int half = s_ / N;
        out[half * stride + (s_ - half * N)] = v;
      });
    }
  }
}

// Which mode-0 launches take this kernel.  ASF_NARROW_CONV=0 sends everything back to the tensor-core kernel.
static bool narrow_conv_ok(const asf_tapgemm_t* q) {
  const char* e = getenv("ASF_NARROW_CONV");  // read per call: the parity tests flip it between two launches
  if (e && e[0] == '0') return false;  // (an SM budget, max_ctas, does not apply: these CTAs share SMs with anything)
  const int N = q->N, K = q->K;
  // K up to 128: the dgrad of the fast pathway's 1x1 expansion convs (s4: 128 -> 32 channels) has 2.7 us of traffic and
  // took 45..54 us on the tcgen05 kernel (profiles/r01_tapgemm_layers_v16.txt)
  if (N < 8 || K < 8 || N > 64 || K > 128 || (N & (N - 1)) || (K & (K - 1))) return false;
  if ((long long)N * K * q->ntaps > NC_MAX_W) return false;
  if (q->out_dtype != ASF_BF16 || q->epi.scale) return false;
  const long long m_total = (long long)q->m_ext[0] * q->m_ext[1] * q->m_ext[2];
  if (m_total < 16384 || m_total > 0x7fffffffLL) return false;
  if (q->ld_out % 8 || ((uintptr_t)q->out & 15) || ((uintptr_t)q->a & 15) || ((uintptr_t)q->bw & 15)) return false;
  if (q->epi.res && (q->epi.ld_res % 8 || ((uintptr_t)q->epi.res & 15))) return false;
  if (q->epi.mask && (q->epi.ld_mask % 8 || ((uintptr_t)q->epi.mask & 15))) return false;
  for (int d = 0; d < q->nd; ++d)
    if (q->a_stride[d] % 8) return false;
  return true;
}

template <int NO>
static int narrow_conv_launch_no(const NcParams& p, int grid, int smem, cudaStream_t st) {
  if (p.epi.stats) {
    ASF_CUDA(launch_pdl(narrow_conv_kernel<NO, true>, dim3(grid), dim3(NC_THREADS), smem, st, p));
  } else {
    ASF_CUDA(launch_pdl(narrow_conv_kernel<NO, false>, dim3(grid), dim3(NC_THREADS), smem, st, p));
  }
  return 0;
}

static long long g_narrow_conv_launches = 0;

static int narrow_conv_launch(const asf_tapgemm_t* q, cudaStream_t st) {
  ++g_narrow_conv_launches;
  NcParams p;
  memset(&p, 0, sizeof(p));
  p.a = (const bf16*)q->a;
  p.bw = (const bf16*)q->bw;
  p.out = (bf16*)q->out;
  p.ld_out = q->ld_out;
  for (int d = 0; d < 4; ++d) {
    p.a_dim[d] = d < q->nd ? (int)q->a_dim[d] : 1;
    p.a_stride[d] = d < q->nd ? q->a_stride[d] : 0;
  }
  for (int j = 0; j < 3; ++j) {
    p.m_ext[j] = q->m_ext[j];
    p.m_to_a[j] = q->m_to_a[j];
    p.o_stride[j] = q->o_stride[j];
  }
  p.ntaps = q->ntaps;
  p.ntaps_w = q->ntaps_w;
  p.N = q->N;
  p.K = q->K;
  p.kv_log2 = ilog2_exact(q->K / 8);
  p.segs = q->ntaps * (q->K / 8);
  p.steps = (int)cdiv(p.segs, 4);
  for (int t = 0; t < q->ntaps; ++t) {
    p.tap_off[t] = 0;
    for (int d = 0; d < 4; ++d) {
      p.a_off[t][d] = d < q->nd ? q->a_off[t][d] : 0;
      p.tap_off[t] += (long long)p.a_off[t][d] * p.a_stride[d];
    }
    p.b_tap[t] = q->b_tap[t];
  }
  p.m_total = q->m_ext[0] * q->m_ext[1] * q->m_ext[2];
  p.tiles = (int)cdiv(p.m_total, 16);
  p.epi = q->epi;
  const int no = q->N / 8, wpc = NC_THREADS / 32;
  const int smem = p.steps * no * 32 * 16 + (q->epi.stats ? 8 * q->N * 4 + 4 * NC_THREADS * 8 + 16 : 0);
  long long grid = cdiv(p.tiles, wpc);
  // resident CTAs per SM by register count (64 / 72 / 94 / 128 registers); a warp then walks its tiles.
  // With statistics every CTA leaves one workspace row for the ordered final pass: at most 4 CTAs per SM.
  const long long cap = (long long)num_sms() * (q->epi.stats ? 4 : (no <= 2 ? 8 : no == 4 ? 5 : 4));
  if (grid > cap) grid = cap;
  if (q->epi.stats)
    ASF_CHECK_ARG(q->epi.stats_ws && q->epi.stats_ticket && grid * 2 * q->N <= ASF_STATS_WS_FLOATS(q->N),
                  "asf_tapgemm: statistics need stats_ws / stats_ticket");
  switch (no) {
    case 1: return narrow_conv_launch_no<1>(p, (int)grid, smem, st);
    case 2: return narrow_conv_launch_no<2>(p, (int)grid, smem, st);
    case 4: return narrow_conv_launch_no<4>(p, (int)grid, smem, st);
    case 8: return narrow_conv_launch_no<8>(p, (int)grid, smem, st);
    default: break;
  }
  return -1;
}
