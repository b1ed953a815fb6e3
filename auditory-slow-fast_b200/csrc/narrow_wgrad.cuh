// Weight gradient of the narrow layers (fast pathway: 8..64 channels, 10^5 output positions), included by
// tapgemm_sm100.cu inside namespace asf.
//
//   dW[n][tap][c] = sum_rows Y[row][n] * X[row + off(tap)][c]              (same contract as tap-GEMM mode 1)
//
// With N * K * taps <= 3072 results and ~10^5 rows this is an HBM-bound reduction, not a GEMM: on the tcgen05 path a
// CTA spends its time waiting for TMA boxes whose rows are 16..64 bytes (the TMA row rate, ~10 cycles per row, bounds
// those launches at 40..90 us).  Here a CTA owns a contiguous range of output positions, streams them in tiles of 64
// through shared memory with 16-byte loads (next tile prefetched into registers) and keeps N*K*taps accumulators in
// registers, c fastest across threads (conflict-free x reads, broadcast y reads).  Partial sums leave in the same
// [ksplit][N][taps][K] workspace layout as the tensor-core kernel, so asf_unpack_wgrad(_multi) finishes both.
#pragma once

constexpr int NW_THREADS = 256;
constexpr int NW_ROWS = 128;       // positions per tile
constexpr int NW_MAX_VEC = 8;      // 16-byte loads per thread per tile (<= 16 vectors per position)
constexpr int NW_PARTS = NW_THREADS / NW_ROWS;  // threads per position in the load role
constexpr int NW_MAX_BLK = 256;    // 2 (n) x 4 (c) accumulator blocks -> N * K * taps <= 2048

struct NwParams {
  const bf16* x;
  const bf16* y;
  float* dwp;
  int nd, y_nd;
  long long a_dim[4], a_stride[4], y_stride[4];
  int m_ext[3], m_to_a[3], y_m_to_a[3];
  int ntaps, ntaps_w, N, K;
  int a_off[ASF_MAX_TAPS][4];
  int b_tap[ASF_MAX_TAPS];
  int m_total;
  int rows_per_cta;  // multiple of NW_ROWS
  int groups;        // row groups sharing an accumulator block (power of two, nblk * groups <= 256)
};

// Shared memory: sy[64][N + 4], sx[taps][64][K + 4] (+ 8 floats per tap): the +4 pitches make the 16-byte stores of
// consecutive positions and the float4 / float2 reads of the FMA loop conflict-free.
// Thread t < nblk * groups owns a 2 (n) x 4 (c) block of one tap for the positions r = g (mod groups): per position one
// LDS.64 (y) + one LDS.128 (x) feed 8 FMAs.  Row groups are folded through shared memory in a fixed order at the end.
__global__ void __launch_bounds__(NW_THREADS) narrow_wgrad_kernel(const __grid_constant__ NwParams p) {
  extern __shared__ __align__(16) float nw_smem[];
  pdl_trigger();
  const int N = p.N, K = p.K, T = p.ntaps;
  const int KV = K >> 3, NV = N >> 3, KQ = K >> 2;
  const int NP = N + 4, KP = K + 4, xpitch = NW_ROWS * KP + 8;
  float* sy = nw_smem;
  float* sx = nw_smem + NW_ROWS * NP;
  const int vpr = NV + T * KV;  // vectors per position
  const int nblk = (N >> 1) * T * KQ;
  const int G = p.groups;
  const int t = threadIdx.x;
  // compute role
  const bool worker = t < nblk * G;
  const int g = t / nblk, blk = t - g * nblk;
  const int cq = blk % KQ, bt = blk / KQ;
  const int tap = bt % T, np = bt / T;
  const float* yp = sy + 2 * np;
  const float* xp = sx + tap * xpitch + 4 * cq;
  float acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
  // load role: position row_l of the tile, vectors part_l, part_l + NW_PARTS, ...
  const int row_l = t & (NW_ROWS - 1), part_l = t / NW_ROWS;
  const int r_begin = blockIdx.x * p.rows_per_cta;
  const int r_end = min(r_begin + p.rows_per_cta, p.m_total);
  const unsigned e0 = (unsigned)p.m_ext[0], e1 = (unsigned)p.m_ext[1];

  auto load_tile = [&](int tile0, uint4* pre) {
    const int pnt = tile0 + row_l;
    const bool live = pnt < r_end;
    const unsigned up = (unsigned)(live ? pnt : 0);
    const unsigned q1 = up / e0;
    const long long m[3] = {(long long)(up - q1 * e0), (long long)(q1 % e1), (long long)(q1 / e1)};
    long long yoff = 0;
#pragma unroll
    for (int j = 0; j < 3; ++j)
      if (p.y_m_to_a[j] > 0) yoff += m[j] * p.y_stride[p.y_m_to_a[j] - 1];
    long long base[4];  // tensor-map coordinates of the position (before the tap offset)
#pragma unroll
    for (int d = 0; d < 4; ++d) {
      long long c0 = 0;
#pragma unroll
      for (int j = 0; j < 3; ++j)
        if (p.m_to_a[j] == d + 1) c0 += m[j];
      base[d] = c0;
    }
#pragma unroll
    for (int u = 0; u < NW_MAX_VEC; ++u) {
      const int v = part_l + NW_PARTS * u;
      uint4 q = make_uint4(0u, 0u, 0u, 0u);
      if (live && v < vpr) {
        if (v < NV) {
          q = *(const uint4*)(p.y + yoff + v * 8);
        } else {
          const int w = v - NV, tp = w / KV, cv = w - tp * KV;
          long long off = cv * 8;
          bool ok = true;
#pragma unroll
          for (int d = 0; d < 4; ++d) {
            if (d < p.nd) {
              const long long coord = base[d] + p.a_off[tp][d];
              ok = ok && coord >= 0 && coord < p.a_dim[d];
              off += coord * p.a_stride[d];
            }
          }
          if (ok) q = *(const uint4*)(p.x + off);
        }
      }
      pre[u] = q;
    }
  };
  auto store_tile = [&](const uint4* pre) {
#pragma unroll
    for (int u = 0; u < NW_MAX_VEC; ++u) {
      const int v = part_l + NW_PARTS * u;
      if (v < vpr) {
        float* dst;
        if (v < NV) {
          dst = sy + row_l * NP + v * 8;
        } else {
          const int w = v - NV, tp = w / KV, cv = w - tp * KV;
          dst = sx + tp * xpitch + row_l * KP + cv * 8;
        }
        const uint32_t wd[4] = {pre[u].x, pre[u].y, pre[u].z, pre[u].w};
        float f[8];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          f[2 * i] = __uint_as_float(wd[i] << 16);
          f[2 * i + 1] = __uint_as_float(wd[i] & 0xffff0000u);
        }
        *(float4*)dst = make_float4(f[0], f[1], f[2], f[3]);
        *(float4*)(dst + 4) = make_float4(f[4], f[5], f[6], f[7]);
      }
    }
  };

  pdl_wait();
  uint4 pre[NW_MAX_VEC];
  load_tile(r_begin, pre);
  for (int tile0 = r_begin; tile0 < r_end; tile0 += NW_ROWS) {
    __syncthreads();  // the previous tile has been consumed
    store_tile(pre);
    __syncthreads();
    if (tile0 + NW_ROWS < r_end) load_tile(tile0 + NW_ROWS, pre);
    if (worker) {
#pragma unroll 4
      for (int r = g; r < NW_ROWS; r += G) {
        const float2 yv = *(const float2*)(yp + r * NP);
        const float4 xv = *(const float4*)(xp + r * KP);
        acc[0] = fmaf(yv.x, xv.x, acc[0]); acc[1] = fmaf(yv.x, xv.y, acc[1]);
        acc[2] = fmaf(yv.x, xv.z, acc[2]); acc[3] = fmaf(yv.x, xv.w, acc[3]);
        acc[4] = fmaf(yv.y, xv.x, acc[4]); acc[5] = fmaf(yv.y, xv.y, acc[5]);
        acc[6] = fmaf(yv.y, xv.z, acc[6]); acc[7] = fmaf(yv.y, xv.w, acc[7]);
      }
    }
  }
  __syncthreads();
  float* red = nw_smem;  // [G][nblk][8]
  if (worker && g > 0) {
#pragma unroll
    for (int i = 0; i < 8; ++i) red[(g * nblk + blk) * 8 + i] = acc[i];
  }
  __syncthreads();
  if (worker && g == 0) {
    for (int gg = 1; gg < G; ++gg) {
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[i] += red[(gg * nblk + blk) * 8 + i];
    }
#pragma unroll
    for (int dn = 0; dn < 2; ++dn) {
      const int n = 2 * np + dn;
      float* dst = p.dwp + (((long long)blockIdx.x * N + n) * p.ntaps_w + p.b_tap[tap]) * K + 4 * cq;
      *(float4*)dst = make_float4(acc[4 * dn], acc[4 * dn + 1], acc[4 * dn + 2], acc[4 * dn + 3]);
    }
  }
}

// positions per CTA and number of CTAs (= the ksplit the caller sizes the workspace for) of a narrow-layer launch
static void narrow_wgrad_rows(const asf_tapgemm_t* q, int* ksplit, int* rows_per_cta) {
  const long long m_total = (long long)q->m_ext[0] * q->m_ext[1] * q->m_ext[2];
  int ctas = 2 * sm_budget(q);
  long long rpc = (cdiv(m_total, ctas) + NW_ROWS - 1) / NW_ROWS * NW_ROWS;
  if (rpc < 4 * NW_ROWS) rpc = 4 * NW_ROWS;
  *rows_per_cta = (int)rpc;
  *ksplit = (int)cdiv(m_total, rpc);
}

// which launches take the CUDA-core path
static bool narrow_wgrad_plan(const asf_tapgemm_t* q, int* ksplit, int* rows_per_cta) {
  const long long m_total = (long long)q->m_ext[0] * q->m_ext[1] * q->m_ext[2];
  const long long nblk = (long long)(q->N / 2) * q->ntaps * (q->K / 4);
  const long long vpr = q->N / 8 + (long long)q->ntaps * (q->K / 8);
  static int enabled = -1;
  if (enabled < 0) {
    const char* e = getenv("ASF_NARROW_WGRAD");
    enabled = (e && e[0] == '0') ? 0 : 1;
  }
  if (!enabled || q->N % 8 || q->K % 8 || q->N > 64 || q->K > 64 || nblk > NW_MAX_BLK || vpr > NW_PARTS * NW_MAX_VEC ||
      m_total < 16384 || m_total >= (1ll << 31))
    return false;
  // every tensor-map stride keeps 8-channel vectors 16-byte aligned (checked by the launcher); all taps of a
  // layer share b_tap space ntaps_w
  narrow_wgrad_rows(q, ksplit, rows_per_cta);
  return true;
}

static int narrow_wgrad_launch(const asf_tapgemm_t* q, int ksplit, int rows_per_cta, cudaStream_t st) {
  NwParams p;
  memset(&p, 0, sizeof(p));
  p.x = (const bf16*)q->a;
  p.y = (const bf16*)q->y;
  p.dwp = q->dwp;
  p.nd = q->nd;
  p.y_nd = q->y_nd;
  ASF_CHECK_ARG(al16(q->a) && al16(q->y), "asf_tapgemm: wgrad operands must be 16-byte aligned");
  for (int d = 0; d < 4; ++d) {
    p.a_dim[d] = d < q->nd ? q->a_dim[d] : 1;
    p.a_stride[d] = d < q->nd ? q->a_stride[d] : 0;
    p.y_stride[d] = d < q->y_nd ? q->y_stride[d] : 0;
    ASF_CHECK_ARG(p.a_stride[d] % 8 == 0 && p.y_stride[d] % 8 == 0, "asf_tapgemm: strides must be multiples of 8 elements");
  }
  for (int j = 0; j < 3; ++j) {
    p.m_ext[j] = q->m_ext[j];
    p.m_to_a[j] = q->m_to_a[j];
    p.y_m_to_a[j] = q->y_m_to_a[j];
  }
  p.ntaps = q->ntaps;
  p.ntaps_w = q->ntaps_w;
  p.N = q->N;
  p.K = q->K;
  for (int t = 0; t < q->ntaps; ++t) {
    for (int d = 0; d < 4; ++d) p.a_off[t][d] = q->a_off[t][d];
    p.b_tap[t] = q->b_tap[t];
  }
  p.m_total = q->m_ext[0] * q->m_ext[1] * q->m_ext[2];
  p.rows_per_cta = rows_per_cta;
  const int nblk = (q->N / 2) * q->ntaps * (q->K / 4);
  int groups = 1;
  while (groups < 8 && nblk * groups * 2 <= NW_THREADS) groups *= 2;
  p.groups = groups;
  size_t smem = sizeof(float) * ((size_t)NW_ROWS * (q->N + 4 + q->ntaps * (q->K + 4)) + 8 * q->ntaps);
  const size_t red = sizeof(float) * (size_t)groups * nblk * 8;
  if (red > smem) smem = red;
  static size_t smem_set = 48 * 1024;
  if (smem > smem_set) {
    ASF_CUDA(cudaFuncSetAttribute(narrow_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    smem_set = smem;
  }
  ASF_CUDA(launch_pdl(narrow_wgrad_kernel, dim3((unsigned)ksplit), dim3(NW_THREADS), smem, st, p));
  return 0;
}
