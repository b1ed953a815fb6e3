// tap-GEMM weight gradient, second generation (included by tapgemm_sm100.cu inside namespace asf).
//
//   dW[n][tap][c] = sum_rows Y[row][n] * X[row + off(tap)][c]
//
// GEMM view per tap: M = output channels n, N = input channels c, K = rows.  Both operands are read straight from
// the channels-last activations as MN-major UMMA operands (TMA boxes of 64 channels x 64 rows, 128B swizzle).
// The first generation used 128 x 128 tiles: one K = 16 step moved 8 KB through L2 -> smem for 64 tensor cycles
// (128 B/cycle/SM, three times what the L2 can feed 148 SMs) and finished every CTA with 16 K fp32 atomics.
// Here a CTA owns up to 256 (n) x 256 (c) of one tap = two M = 128 halves x N <= 256 accumulators filling TMEM's
// 512 columns: 64 B/cycle/SM, and the split-K partial tiles are written with plain stores into a
// [ksplit][N][taps][K] fp32 workspace that asf_unpack_wgrad sums while it converts to the OIHW master layout
// (no atomics, no memset).
#pragma once

constexpr int WG2_ROWS = 64;                    // rows (GEMM K) per pipeline stage
constexpr int WG2_BOX_BYTES = WG2_ROWS * 128;   // one TMA box: 64 rows x 64 channels bf16
constexpr int WG2_THREADS = 320;                // warp 0 TMA, warp 1 MMA, warps 2..5 / 6..9 epilogue of half 0 / 1

__global__ void __launch_bounds__(WG2_THREADS, 1) tapgemm_wgrad2_kernel(const __grid_constant__ TgParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  pdl_trigger();
  uint8_t* smem = smem_raw;
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  const int ny = p.wg_ny, nx = p.wg_nx;
  const int stage_bytes = (ny + nx) * WG2_BOX_BYTES;
  TgShared* sh = (TgShared*)(smem + (size_t)p.stages * stage_bytes);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // tile decode: blockIdx.x = ((r_tile * ntaps + tap) * c_tiles + c_tile) * ksplit + ks
  int idx = blockIdx.x;
  const int ks = idx % p.ksplit;
  idx /= p.ksplit;
  const int c_tile = idx % p.c_tiles;
  idx /= p.c_tiles;
  const int tap = idx % p.ntaps;
  const int r_tile = idx / p.ntaps;
  const int n_iters = (p.m_tiles_total - ks + p.ksplit - 1) / p.ksplit;  // row tiles ks, ks + ksplit, ...
  const int TM = p.wg_halves * 128, TN = p.wg_tn;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tm_a);
    tma_prefetch_desc(&p.tm_b);
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(&sh->full[s], 1);
      mbar_init(&sh->empty[s], 1);
    }
    mbar_init(&sh->tmem_full[0], 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(&sh->tmem_base, (uint32_t)p.wg_tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = sh->tmem_base;
  pdl_wait();  // everything above touched only shared memory, TMEM and kernel parameters

  if (warp != 1 && lane == 0) {
    // ===== TMA issue, spread over nine threads: a thread spends ~250 issue cycles per cp.async.bulk.tensor, and a
    // stage has up to eight boxes for 1024 tensor cycles.  Warp 0 posts the expected bytes; the epilogue warps (idle
    // until the accumulators are complete) each stream one box of every stage: warp 2 + i owns box i. =====
    const int my_box = warp - 2;  // -1 for warp 0
    if (warp == 0 || my_box < ny + nx) {
      int s = 0;
      uint32_t ph = 0;
      for (int i = 0; i < n_iters; ++i) {
        const int m_tile = ks + i * p.ksplit;
        mbar_wait(&sh->empty[s], ph ^ 1, 11);
        uint8_t* sy = smem + (size_t)s * stage_bytes;
        if (warp == 0) {
          mbar_expect_tx(&sh->full[s], (uint32_t)stage_bytes);
        } else {
          int st[3];
          tile_coords(p, m_tile, st);
          const bool is_y = my_box < ny;
          int c1 = 0, c2 = 0, c3 = 0, c4 = 0;
#pragma unroll
          for (int j = 0; j < 3; ++j) {
            const int d = is_y ? p.y_m_to_a[j] : p.m_to_a[j];
            c1 = d == 1 ? st[j] : c1; c2 = d == 2 ? st[j] : c2; c3 = d == 3 ? st[j] : c3; c4 = d == 4 ? st[j] : c4;
          }
          if (is_y) {
            tma_load(&p.tm_b, &sh->full[s], sy + my_box * WG2_BOX_BYTES, p.rank_b, r_tile * TM + my_box * 64, c1, c2, c3,
                     c4);
          } else {
            c1 += p.a_off[tap][0]; c2 += p.a_off[tap][1]; c3 += p.a_off[tap][2]; c4 += p.a_off[tap][3];
            tma_load(&p.tm_a, &sh->full[s], sy + my_box * WG2_BOX_BYTES, p.rank_a, c_tile * TN + (my_box - ny) * 64, c1,
                     c2, c3, c4);
          }
        }
        if (++s == p.stages) { s = 0; ph ^= 1; }
      }
    }
  }
  __syncwarp();
  if (warp == 0) {
    // nothing else to do
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = make_idesc(128, TN, 1, 1);
      const uint32_t smem0 = smem_u32(smem);
      // MN-major, 128B swizzle: 64-channel chunks LBO = 8192 B apart, 8-row (K) groups SBO = 1024 B apart
      const uint64_t ydesc0 = make_sdesc(smem0, WG2_BOX_BYTES, 1024);
      const uint64_t xdesc0 = make_sdesc(smem0 + ny * WG2_BOX_BYTES, WG2_BOX_BYTES, 1024);
      const uint64_t dstage = (uint64_t)(stage_bytes >> 4);
      const uint64_t dhalf = (uint64_t)((2 * WG2_BOX_BYTES) >> 4);  // second M half = Y chunks 2, 3
      int s = 0;
      uint32_t ph = 0;
      uint32_t acc = 0;
      for (int i = 0; i < n_iters; ++i) {
        mbar_wait(&sh->full[s], ph, 12);
        tc_fence_after();
        const uint64_t yd = ydesc0 + dstage * (uint64_t)s, xd = xdesc0 + dstage * (uint64_t)s;
#pragma unroll
        for (int k = 0; k < WG2_ROWS / 16; ++k) {  // 16 rows = 2048 B = 128 x 16B units per UMMA_K step
          umma_bf16(tmem_base, yd + (uint64_t)(k * 128), xd + (uint64_t)(k * 128), idesc, acc | (uint32_t)k);
          if (p.wg_halves == 2)
            umma_bf16(tmem_base + (uint32_t)TN, yd + dhalf + (uint64_t)(k * 128), xd + (uint64_t)(k * 128), idesc,
                      acc | (uint32_t)k);
        }
        acc = 1;
        umma_commit(&sh->empty[s]);
        if (++s == p.stages) { s = 0; ph ^= 1; }
      }
      umma_commit(&sh->tmem_full[0]);
    }
  } else {
    const int half = (warp - 2) >> 2;
    const int quarter = warp & 3;
    if (half < p.wg_halves) {
      mbar_wait_warp(&sh->tmem_full[0], 0, 13, 128);
      tc_fence_after();
      const uint32_t taddr0 = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(half * TN);
      const int wt = p.b_tap[tap];
      // 32 rows x 32 columns per step: TMEM -> registers (one row per lane) -> this warp's patch of the (now idle)
      // operand ring, pitch 36 floats -> read back 4 rows x 128 contiguous bytes per store instruction.
      float* patch = (float*)smem + (warp - 2) * (32 * 36);
      const int n0 = r_tile * TM + half * 128 + quarter * 32;
      const int sub_r = lane >> 3, sub_c = (lane & 7) * 4;
      for (int c0 = 0; c0 < TN; c0 += 32) {
        uint32_t raw[32];
        tmem_ld32_issue(taddr0 + (uint32_t)c0, raw);
        tmem_ld_wait();
#pragma unroll
        for (int g = 0; g < 8; ++g) {
          float4 v;
          v.x = n_iters > 0 ? __uint_as_float(raw[4 * g]) : 0.f;
          v.y = n_iters > 0 ? __uint_as_float(raw[4 * g + 1]) : 0.f;
          v.z = n_iters > 0 ? __uint_as_float(raw[4 * g + 2]) : 0.f;
          v.w = n_iters > 0 ? __uint_as_float(raw[4 * g + 3]) : 0.f;
          *(float4*)(patch + lane * 36 + 4 * g) = v;
        }
        __syncwarp();
        const int c = c_tile * TN + c0 + sub_c;
        const bool c_ok = c0 + sub_c < TN && c + 4 <= p.K;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int rr = 4 * i + sub_r;
          const int nn = n0 + rr;
          if (c_ok && nn < p.N) {
            const float4 v = *(const float4*)(patch + rr * 36 + sub_c);
            *(float4*)(p.dwp + (((long long)ks * p.N + nn) * p.ntaps_w + wt) * p.K + c) = v;
          }
        }
        __syncwarp();
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, (uint32_t)p.wg_tmem_cols);
  }
}
