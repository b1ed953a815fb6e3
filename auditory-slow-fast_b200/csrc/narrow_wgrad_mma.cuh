// Weight gradient of the narrow layers on warp-level tensor cores (mma.sync.m16n8k16, bf16 in, fp32 accumulate),
// included by tapgemm_sm100.cu inside namespace asf.
//
//   dW[n][tap][c] = sum_rows Y[row][n] * X[row + off(tap)][c]              (same contract as tap-GEMM mode 1)
//
// GEMM view: M = output channels n (padded to 16), N = j = tap * K + c, K = rows.  Both operands sit in memory with the
// REDUCTION index (row) slowest, so they are staged row-major in shared memory by cp.async (zero-fill for padding taps
// and rows past the CTA's range) and fed to the MMA through ldmatrix.trans, which delivers exactly the transposed 8 x 8
// fragments the m16n8k16 A (row-major) and B (column-major) operands want.  A CTA owns a contiguous range of output
// positions (the split-K partial `blockIdx.x`), streams it in double-buffered tiles of 128 (or 64) positions; warp w
// owns the 8-column blocks jb = w, w + 8, ... of j for ALL n blocks, so every B fragment feeds up to four MMAs.
// Partial sums leave as plain stores in the [ksplit][N][taps][K] workspace that asf_unpack_wgrad(_multi) reduces.
//
// History: the first (CUDA-core FMA) version of this kernel took 18..44 us per launch for 7..16 MB of operands
// (profiles/r02_narrow_wgrad_ncu.txt); the layers with N * K * taps > 2048 ran on the tcgen05 kernel at 30..70 us,
// bound by the TMA row rate of their 16..64-byte rows.
#pragma once

constexpr int NWM_THREADS = 256;
constexpr int NW_ROWS = 128;  // granularity of a CTA's range of positions

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

constexpr int NWM_MAX_JBW = 4;   // 8-column blocks of j per warp -> taps * K <= 256

struct NwmParams {
  NwParams g;     // geometry / pointers
  int rows_tile;  // positions per pipeline stage: 128 or 64
  int py, px;     // shared-memory pitches (elements) of the Y and X tiles
  int n_pad;      // N rounded up to 16
};

__device__ __forceinline__ void nwm_ldsm_x4_t(uint32_t* r, const void* p) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(smem_u32(p)));
}
__device__ __forceinline__ void nwm_ldsm_x2_t(uint32_t* r, const void* p) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.trans.shared.b16 {%0,%1}, [%2];" : "=r"(r[0]), "=r"(r[1]) : "r"(smem_u32(p)));
}
__device__ __forceinline__ void nwm_mma(float* d, const uint32_t* a, const uint32_t* b) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
// 16 bytes global -> shared; src_bytes = 0 writes zeros (padding taps, positions past the range)
__device__ __forceinline__ void nwm_cp16(void* dst, const void* src, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(dst)), "l"(src), "r"(src_bytes) : "memory");
}

template <int NB>  // 16-row blocks of n: 1, 2 or 4
__global__ void __maxnreg__(96) narrow_wgrad_mma_kernel(const __grid_constant__ NwmParams q) {
  extern __shared__ __align__(16) uint8_t nwm_smem[];
  pdl_trigger();
  const NwParams& p = q.g;
  const int N = p.N, K = p.K, T = p.ntaps;
  const int J = T * K, JB = J >> 3;
  const int NV = N >> 3, KV = K >> 3;
  const int R = q.rows_tile, PY = q.py, PX = q.px;
  const int stage_elems = R * (PY + PX);
  bf16* sbase = (bf16*)nwm_smem;
  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  const int r_begin = blockIdx.x * p.rows_per_cta;
  const int r_end = min(r_begin + p.rows_per_cta, p.m_total);
  const unsigned e0 = (unsigned)p.m_ext[0], e1 = (unsigned)p.m_ext[1];
  // Load role.  A tile is R positions x (1 + T) "units": unit 0 = the position's N output gradients, unit 1 + tap = the
  // K input channels that tap reads.  A unit is CONTIGUOUS in global and in shared memory, so a thread does the
  // coordinate arithmetic once per unit (two 32-bit divisions, the tap's bounds test) and then issues plain 16-byte
  // cp.async with constant offsets.  (The first version decoded every 16-byte vector separately: 1200 instructions
  // per warp and tile, 16.6 M warp instructions per launch -- the kernel was bound by address generation, 38 us for
  // 16 MB.)
  const int r_log2 = R == 128 ? 7 : 6;
  auto load_tile = [&](int tile0, int stage) {
    bf16* sy = sbase + (size_t)stage * stage_elems;
    bf16* sx = sy + R * PY;
    for (int uidx = t; uidx < R * (T + 1); uidx += NWM_THREADS) {
      const int row = uidx & (R - 1), unit = uidx >> r_log2;
      const int pnt = tile0 + row;
      const bool live = pnt < r_end;
      const unsigned up = (unsigned)(live ? pnt : 0);
      const unsigned q1 = up / e0;
      const unsigned q2 = q1 / e1;
      const int m[3] = {(int)(up - q1 * e0), (int)(q1 - q2 * e1), (int)q2};
      const bf16* src;
      bf16* dst;
      int nvec;
      bool ok = live;
      if (unit == 0) {
        long long yoff = 0;
#pragma unroll
        for (int j = 0; j < 3; ++j)
          if (p.y_m_to_a[j] > 0) yoff += (long long)m[j] * p.y_stride[p.y_m_to_a[j] - 1];
        src = p.y + yoff;
        dst = sy + row * PY;
        nvec = NV;
      } else {
        const int tp = unit - 1;
        long long off = 0;
#pragma unroll
        for (int d = 0; d < 4; ++d) {
          if (d < p.nd) {
            int c0 = p.a_off[tp][d];
#pragma unroll
            for (int j = 0; j < 3; ++j)
              if (p.m_to_a[j] == d + 1) c0 += m[j];
            ok = ok && c0 >= 0 && c0 < (int)p.a_dim[d];
            off += (long long)c0 * p.a_stride[d];
          }
        }
        src = p.x + off;
        dst = sx + row * PX + tp * K;
        nvec = KV;
      }
      if (!ok) src = p.x;  // any valid address: zero bytes are read, sixteen zero bytes are written
      const int nbytes = ok ? 16 : 0;
      for (int v = 0; v < nvec; ++v) nwm_cp16(dst + v * 8, src + v * 8, nbytes);
    }
    cp_async_commit();
  };

  // the padding rows of n (N = 8 -> 16) are never loaded: zero them once in both stages
  if (q.n_pad > N) {
    for (int i = t; i < 2 * R; i += NWM_THREADS) {
      bf16* sy = sbase + (size_t)(i / R) * stage_elems + (i % R) * PY + N;
      *(uint4*)sy = make_uint4(0u, 0u, 0u, 0u);
    }
  }
  float acc[NB][NWM_MAX_JBW][4];
#pragma unroll
  for (int a = 0; a < NB; ++a)
#pragma unroll
    for (int b = 0; b < NWM_MAX_JBW; ++b) acc[a][b][0] = acc[a][b][1] = acc[a][b][2] = acc[a][b][3] = 0.f;

  pdl_wait();
  const int n_tiles = (r_end - r_begin + R - 1) / R;
  if (n_tiles > 0) load_tile(r_begin, 0);
  // ldmatrix lane roles: matrix = lane >> 3, row inside it = lane & 7
  const int a_row = (lane & 7) + ((lane >> 4) & 1) * 8, a_col = ((lane >> 3) & 1) * 8;
  const int b_row = (lane & 7) + ((lane >> 3) & 1) * 8;
  for (int i = 0; i < n_tiles; ++i) {
    const int stage = i & 1;
    if (i + 1 < n_tiles) {
      load_tile(r_begin + (i + 1) * R, stage ^ 1);  // the other stage was consumed before the barrier closing tile i - 1
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    const bf16* sy = sbase + (size_t)stage * stage_elems;
    const bf16* sx = sy + R * PY;
    for (int k0 = 0; k0 < R; k0 += 16) {
      uint32_t a[NB][4];
#pragma unroll
      for (int nb = 0; nb < NB; ++nb) nwm_ldsm_x4_t(a[nb], sy + (k0 + a_row) * PY + nb * 16 + a_col);
#pragma unroll
      for (int jj = 0; jj < NWM_MAX_JBW; ++jj) {
        const int jb = warp + 8 * jj;
        if (jb < JB) {  // uniform across the warp
          uint32_t b[2];
          nwm_ldsm_x2_t(b, sx + (k0 + b_row) * PX + jb * 8);
#pragma unroll
          for (int nb = 0; nb < NB; ++nb) nwm_mma(acc[nb][jj], a[nb], b);
        }
      }
    }
    __syncthreads();  // everybody is done with this stage before the tile after next overwrites it
  }
  // accumulator fragment: rows g, g + 8 of the n block, columns 2 tq, 2 tq + 1 of the j block
  const int g = lane >> 2, tq = lane & 3;
#pragma unroll
  for (int jj = 0; jj < NWM_MAX_JBW; ++jj) {
    const int jb = warp + 8 * jj;
    if (jb >= JB) continue;
    const int j = jb * 8 + 2 * tq;
    const int tap = j / K, c = j - tap * K;
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) {
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int n = nb * 16 + g + 8 * h;
        if (n < N) {
          float* dst = p.dwp + (((long long)blockIdx.x * N + n) * p.ntaps_w + p.b_tap[tap]) * K + c;
          *(float2*)dst = make_float2(acc[nb][jj][2 * h], acc[nb][jj][2 * h + 1]);
        }
      }
    }
  }
}

// positions per CTA and number of CTAs (= the ksplit the caller sizes the workspace for) of a narrow-layer launch
static void narrow_wgrad_rows(const asf_tapgemm_t* q, int* ksplit, int* rows_per_cta) {
  const long long m_total = (long long)q->m_ext[0] * q->m_ext[1] * q->m_ext[2];
  // CTAs per SM: every CTA leaves one partial copy of the layer's gradient for the split-K reduction to read
  static int per_sm = -1;
  if (per_sm < 0) {
    const char* e = getenv("ASF_NWM_CTAS_PER_SM");
    per_sm = e ? atoi(e) : 2;
    if (per_sm < 1 || per_sm > 4) per_sm = 2;
  }
  int ctas = per_sm * sm_budget(q);
  long long rpc = (cdiv(m_total, ctas) + NW_ROWS - 1) / NW_ROWS * NW_ROWS;
  if (rpc < 4 * NW_ROWS) rpc = 4 * NW_ROWS;
  *rows_per_cta = (int)rpc;
  *ksplit = (int)cdiv(m_total, rpc);
}

// which layers this kernel takes
static bool narrow_wgrad_mma_ok(const asf_tapgemm_t* q) {
  static int enabled = -1;
  if (enabled < 0) {
    const char* e = getenv("ASF_NARROW_WGRAD_MMA");
    enabled = (e && e[0] == '0') ? 0 : 1;
  }
  const long long m_total = (long long)q->m_ext[0] * q->m_ext[1] * q->m_ext[2];
  const int J = q->ntaps * q->K;
  if (!enabled || q->N % 8 || q->K % 8 || q->N > 64 || q->N < 8 || q->K < 8 || J > 8 * 8 * NWM_MAX_JBW) return false;
  if (m_total < 16384 || m_total >= (1ll << 31)) return false;
  return true;
}

static void nwm_geometry(const asf_tapgemm_t* q, NwmParams* w) {
  const int J = q->ntaps * q->K;
  w->n_pad = (q->N + 15) / 16 * 16;
  // pitch / 8 odd: the eight 16-byte rows of an ldmatrix 8 x 8 tile fall into distinct bank groups
  w->py = w->n_pad + 8;
  w->px = J + (((J >> 3) & 1) ? 16 : 8);
  int R = 128;
  static int r_env = -1;
  if (r_env < 0) {
    const char* e = getenv("ASF_NWM_ROWS");
    r_env = e ? atoi(e) : 0;
  }
  if (r_env == 64) R = 64;
  if ((size_t)2 * 128 * (w->py + w->px) * 2 > 96 * 1024) R = 64;
  w->rows_tile = R;
}

static int narrow_wgrad_mma_launch(const asf_tapgemm_t* q, int ksplit, int rows_per_cta, cudaStream_t st) {
  NwmParams w;
  memset(&w, 0, sizeof(w));
  NwParams& p = w.g;
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
  nwm_geometry(q, &w);
  const size_t smem = (size_t)2 * w.rows_tile * (w.py + w.px) * sizeof(bf16);
  ASF_CHECK_ARG(smem <= 200 * 1024, "asf_tapgemm: narrow wgrad tile does not fit in shared memory");
  const int nb = w.n_pad / 16;
  static size_t smem_set[5] = {48 * 1024, 48 * 1024, 48 * 1024, 48 * 1024, 48 * 1024};
  if (smem > smem_set[nb]) {
    if (nb == 1) ASF_CUDA(cudaFuncSetAttribute(narrow_wgrad_mma_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (nb == 2) ASF_CUDA(cudaFuncSetAttribute(narrow_wgrad_mma_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (nb == 3 || nb == 4) ASF_CUDA(cudaFuncSetAttribute(narrow_wgrad_mma_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    smem_set[nb] = smem;
  }
  if (nb == 1) ASF_CUDA(launch_pdl(narrow_wgrad_mma_kernel<1>, dim3((unsigned)ksplit), dim3(NWM_THREADS), smem, st, w));
  else if (nb == 2) ASF_CUDA(launch_pdl(narrow_wgrad_mma_kernel<2>, dim3((unsigned)ksplit), dim3(NWM_THREADS), smem, st, w));
  else ASF_CUDA(launch_pdl(narrow_wgrad_mma_kernel<4>, dim3((unsigned)ksplit), dim3(NWM_THREADS), smem, st, w));
  return 0;
}
