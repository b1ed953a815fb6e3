// tap-GEMM forward / dgrad, second generation (included by tapgemm_sm100.cu inside namespace asf).
//
// Same mainloop contract as tapgemm_fwd_kernel (TMA -> smem ring -> tcgen05.mma -> TMEM, persistent CTAs), with
//   * a tight single-thread MMA issue loop (no integer division, four unrolled K=16 steps per 64-channel chunk),
//   * an all-asynchronous epilogue: residual / accumulate / mask operands arrive by per-thread cp.async into swizzled
//     smem slabs (prefetched one slab ahead, each thread its own row), results leave through smem + TMA store, one
//     named barrier per slab; no thread ever waits on a global load.
//     OOB rows / channels of a ragged tile are clipped by the TMA store, so the math runs unpredicated,
//   * training-BN statistics summed column-wise from the bf16 slab that is being stored (what BN will read).
//
// Slab = 128 rows x 32 channels bf16 = 8 KB, CU_TENSOR_MAP_SWIZZLE_64B: 16-byte chunk c of row r lives at
//   r*64 + ((c ^ ((r >> 1) & 3)) << 4).
#pragma once

constexpr int EPI_SLAB_COLS = 32;
constexpr int EPI_SLAB_BYTES = TG_BLOCK_M * EPI_SLAB_COLS * 2;  // 8192
constexpr int EPI_SHIFT_FLOATS = 256;
// training-BN statistics (STATS launches only): per epilogue group [2][256] column accumulators + [4 warps][64] slab
// partials.  Deterministic: no atomics, fixed summation order (see epi_slab_stats_part / epi_stats_accumulate).
constexpr int EPI_STAT_ACC_FLOATS = 2 * 256;
constexpr int EPI_STAT_SCR_HALF = 4 * 64;                   // [4 warps][64] partials of one slab
constexpr int EPI_STAT_SCR_FLOATS = 2 * EPI_STAT_SCR_HALF;  // double buffered: one barrier per slab
constexpr int EPI_STAT_FLOATS = 2 * (EPI_STAT_ACC_FLOATS + EPI_STAT_SCR_FLOATS);
constexpr int TGT_THREADS = 352;  // warp 0: TMA (activations), 1: MMA, 2..9: epilogue, 10: TMA (weights)

__device__ __forceinline__ uint32_t slab_off(int r, int c) { return (uint32_t)(r * 64 + ((c ^ ((r >> 1) & 3)) << 4)); }

__device__ __forceinline__ void bf16x8_to_f32(const uint4& q, float* f) {
  const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    f[2 * i] = __uint_as_float(w[i] << 16);
    f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
  }
}

// One slab of one row: raw fp32 accumulators -> epilogue math -> bf16 into the out slab.
template <bool SHIFT, int NADD, bool MASK, bool RELU>
__device__ __forceinline__ void epi_slab_math(const uint32_t* raw, const float* s_shift, const uint8_t* op0,
                                              const uint8_t* op1, uint8_t* out, int r, bool zero) {
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    float w[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) w[j] = __uint_as_float(raw[8 * c + j]);
    const uint32_t off = slab_off(r, c);
    if (SHIFT) {
      const float4 s0 = *(const float4*)(s_shift + 8 * c), s1 = *(const float4*)(s_shift + 8 * c + 4);
      w[0] += s0.x; w[1] += s0.y; w[2] += s0.z; w[3] += s0.w; w[4] += s1.x; w[5] += s1.y; w[6] += s1.z; w[7] += s1.w;
    }
    if (NADD) {
      float a[8];
      bf16x8_to_f32(*(const uint4*)(op0 + off), a);
#pragma unroll
      for (int j = 0; j < 8; ++j) w[j] += a[j];
    }
    if (RELU) {
#pragma unroll
      for (int j = 0; j < 8; ++j) w[j] = fmaxf(w[j], 0.f);
    }
    if (MASK) {
      float m[8];
      bf16x8_to_f32(*(const uint4*)((NADD ? op1 : op0) + off), m);
#pragma unroll
      for (int j = 0; j < 8; ++j) w[j] = (m[j] > 0.f) ? w[j] : 0.f;
    }
    if (zero) {
#pragma unroll
      for (int j = 0; j < 8; ++j) w[j] = 0.f;
    }
    uint4 o;
    __nv_bfloat162* ob = (__nv_bfloat162*)&o;
#pragma unroll
    for (int j = 0; j < 4; ++j) ob[j] = __floats2bfloat162_rn(w[2 * j], w[2 * j + 1]);
    *(uint4*)(out + off) = o;
  }
}

// Column sums / sums of squares of one finished bf16 slab (rows outside the tensor were written as zeros).
// t = thread index inside the 128-thread epilogue group.  Thread t owns channel pair (2w, 2w+1), w = t & 15, of rows
// (t >> 4) + 8 i: a warp reads two adjacent rows per step = one conflict-free 128-byte wavefront.  The two row groups of
// a warp are folded with one xor-shuffle; lanes 0..15 then STORE the warp's partial sums into the warp's own 64 floats
// of `wscr` ([sum 32 | sum of squares 32]).  epi_stats_accumulate adds the four warps in warp order.
__device__ __forceinline__ void epi_slab_stats_part(const uint8_t* slab, int t, float* wscr) {
  const int w = t & 15, rg = t >> 4;
  float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const int r = rg + 8 * i;
    const uint32_t v = *(const uint32_t*)(slab + r * 64 + (((w >> 2) ^ ((r >> 1) & 3)) << 4) + ((w & 3) << 2));
    const float a = __uint_as_float(v << 16), b = __uint_as_float(v & 0xffff0000u);
    s0 += a; q0 = fmaf(a, a, q0);
    s1 += b; q1 = fmaf(b, b, q1);
  }
  s0 += __shfl_xor_sync(0xffffffffu, s0, 16);
  s1 += __shfl_xor_sync(0xffffffffu, s1, 16);
  q0 += __shfl_xor_sync(0xffffffffu, q0, 16);
  q1 += __shfl_xor_sync(0xffffffffu, q1, 16);
  if ((t & 16) == 0) {
    float* d = wscr + (t >> 5) * 64;
    *(float2*)(d + 2 * w) = make_float2(s0, s1);
    *(float2*)(d + 32 + 2 * w) = make_float2(q0, q1);
  }
}
// threads t < 64 of a group: slot t = (half, channel) of the pending slab; the group's accumulator of that column is
// owned by this one thread, so the order of additions over slabs and tiles is the program order.
__device__ __forceinline__ void epi_stats_accumulate(float* acc, const float* wscr, int t, int col0) {
  const int half = t >> 5, ch = t & 31;
  const float v = ((wscr[t] + wscr[64 + t]) + wscr[128 + t]) + wscr[192 + t];
  acc[half * 256 + col0 + ch] += v;
}

// ------------------------------------------------------------------------------------------------ the kernel
// warp 0: TMA producer (mainloop operands), warp 1: MMA issuer + TMEM owner,
// warps 2..5 / 6..9: two epilogue groups (group g drains TMEM buffer g, i.e. every other tile of this CTA).
template <bool STATS>
__global__ void __launch_bounds__(TGT_THREADS, 1) tapgemm_fwd_tma_kernel(const __grid_constant__ TgParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  pdl_trigger();
  unsigned long long gt_entry = 0;
  if (p.dbg && threadIdx.x == 0) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(gt_entry));
  uint8_t* smem = smem_raw;  // 1024-byte aligned (128B-swizzle atoms); checked below
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  const int b_bytes = p.block_n * 128;
  const int stage_bytes = TG_A_BYTES + b_bytes;
  const int n_ops = p.n_add + p.has_mask;
  uint8_t* epi_base = smem + p.epi_smem_off;
  const int OD = p.op_depth;  // operand ring depth
  const int group_bytes = (2 + OD * n_ops) * EPI_SLAB_BYTES;
  float* shift_base = (float*)(epi_base + 2 * group_bytes);
  float* stat_base = shift_base + 2 * EPI_SHIFT_FLOATS;  // STATS: [2 groups][acc 512 | scratch 256]
  TgShared* sh = (TgShared*)(stat_base + (STATS ? EPI_STAT_FLOATS : 0));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total_tiles = p.m_tiles_total * p.n_tiles;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tm_a);
    tma_prefetch_desc(&p.tm_b);
    tma_prefetch_desc(&p.tm_out);
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(&sh->full[s], 1);
      mbar_init(&sh->empty[s], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&sh->tmem_full[i], 1);
      mbar_init(&sh->tmem_empty[i], 4);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(&sh->tmem_base, 512);
  // Training-BN statistics: the launcher sizes the grid as a multiple of n_tiles, so this CTA keeps ONE N tile for all
  // its M tiles; each epilogue group sums its slabs' columns into its own shared-memory accumulators.
  if (STATS && threadIdx.x >= 64 && threadIdx.x < 320)
    for (int i = (int)threadIdx.x - 64; i < EPI_STAT_FLOATS; i += 256) stat_base[i] = 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = sh->tmem_base;
  pdl_wait();  // everything above touched only shared memory, TMEM and kernel parameters
  Probe pb{p.dbg, 0}, pb_all{p.dbg, 0};
  if (threadIdx.x == 0) pb_all.begin();
  if (p.dbg && threadIdx.x == 0) {
    unsigned long long g1;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(g1));
    atomicAdd(p.dbg + 9, g1 - gt_entry);
    atomicMin(p.dbg + 11, gt_entry);
  }

  if (warp == 0 || warp == 10) {
    // ===== TMA producers: warp 0 streams the activation boxes, warp 10 the weight boxes.  One thread issuing both
    // spends ~250 issue cycles per cp.async.bulk.tensor (coordinates, descriptor, barrier) -- as long as the four MMAs
    // of a 128 x 128 k-chunk -- so the two streams run in parallel; warp 0 also posts the expected byte count. =====
    if (lane == 0) {
      const bool act = warp == 0;
      int s = 0;
      uint32_t ph = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int n_tile = tile % p.n_tiles, m_tile = tile / p.n_tiles;
        int st[3];
        tile_coords(p, m_tile, st);
        int b1 = 0, b2 = 0, b3 = 0, b4 = 0;
#pragma unroll
        for (int j = 0; j < 3; ++j) {
          const int d = p.m_to_a[j];
          b1 = d == 1 ? st[j] : b1;
          b2 = d == 2 ? st[j] : b2;
          b3 = d == 3 ? st[j] : b3;
          b4 = d == 4 ? st[j] : b4;
        }
        const int nrow = n_tile * p.block_n;
        for (int tap = 0; tap < p.ntaps; ++tap) {
          const int c1 = b1 + p.a_off[tap][0], c2 = b2 + p.a_off[tap][1], c3 = b3 + p.a_off[tap][2],
                    c4 = b4 + p.a_off[tap][3];
          const int wt = p.b_tap[tap];
          for (int kc = 0; kc < p.kchunks; ++kc) {
            if (act) pb.begin();
            mbar_wait(&sh->empty[s], ph ^ 1, 1);
            if (act) pb.end(1);
            uint8_t* sa = smem + (size_t)s * stage_bytes;
            if (act) {
              mbar_expect_tx(&sh->full[s], (uint32_t)stage_bytes);
              tma_load(&p.tm_a, &sh->full[s], sa, p.rank_a, kc * TG_KC, c1, c2, c3, c4);
            } else {
              tma_load(&p.tm_b, &sh->full[s], sa + TG_A_BYTES, 3, kc * TG_KC, wt, nrow, 0, 0);
            }
            if (++s == p.stages) { s = 0; ph ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread) =====
    if (lane == 0) {
      const uint32_t idesc = make_idesc(TG_BLOCK_M, p.block_n, 0, 0);
      const int kiters = p.ntaps * p.kchunks;
      const int ksteps_last = min(TG_KC / 16, (p.K - (p.kchunks - 1) * TG_KC + 15) >> 4);
      const bool full4 = ksteps_last == 4;
      const uint32_t smem0 = smem_u32(smem);
      const uint64_t adesc0 = make_sdesc(smem0, 16, 1024);
      const uint64_t bdesc0 = make_sdesc(smem0 + TG_A_BYTES, 16, 1024);
      const uint64_t dstage = (uint64_t)(stage_bytes >> 4);
      int s = 0;
      uint32_t ph = 0;
      int it = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
        const int buf = it & 1;
        pb.begin();
        mbar_wait(&sh->tmem_empty[buf], ((it >> 1) & 1) ^ 1, 2);
        pb.end(3);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(buf * 256);
        uint32_t acc = 0;
        int kc = 0;
        for (int ki = 0; ki < kiters; ++ki) {
          pb.begin();
          mbar_wait(&sh->full[s], ph, 3);
          pb.end(2);
          tc_fence_after();
          const uint64_t adesc = adesc0 + dstage * (uint64_t)s, bdesc = bdesc0 + dstage * (uint64_t)s;
          if (full4 || kc + 1 < p.kchunks) {
            umma_bf16(d_tmem, adesc, bdesc, idesc, acc);
            umma_bf16(d_tmem, adesc + 2, bdesc + 2, idesc, 1);
            umma_bf16(d_tmem, adesc + 4, bdesc + 4, idesc, 1);
            umma_bf16(d_tmem, adesc + 6, bdesc + 6, idesc, 1);
          } else {  // short last chunk (K % 64 != 0): skip the zero-filled K tail
            for (int k = 0; k < ksteps_last; ++k)
              umma_bf16(d_tmem, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, acc | (uint32_t)k);
          }
          acc = 1;
          umma_commit(&sh->empty[s]);
          if (++kc == p.kchunks) kc = 0;
          if (++s == p.stages) { s = 0; ph ^= 1; }
        }
        umma_commit(&sh->tmem_full[buf]);
      }
    }
  } else {
    // ===== epilogue groups =====
    const int group = (warp - 2) >> 2;
    const int quarter = warp & 3;                 // TMEM lane quarter this warp may read
    const int r = quarter * 32 + lane;            // row of the tile = TMEM lane
    const int t = (int)threadIdx.x - 64 - group * 128;  // 0..127 inside the group
    const bool leader = t == 0;
    const int bar_id = 1 + group;
    uint8_t* gbase = epi_base + group * group_bytes;
    uint8_t* out_buf = gbase;                             // [2] slabs
    uint8_t* op_buf = gbase + 2 * EPI_SLAB_BYTES;         // [n_ops][OD] slabs
    float* s_shift = shift_base + group * EPI_SHIFT_FLOATS;
    const float* g_shift = p.epi.shift;
    const int variant = (g_shift ? 8 : 0) | (p.n_add ? 4 : 0) | (p.has_mask ? 2 : 0) | (p.epi.relu ? 1 : 0);
    const int r1 = r & ((1 << p.m_log2[0]) - 1);
    const int r2 = (r >> p.m_log2[0]) & ((1 << p.m_log2[1]) - 1);
    const int r3 = r >> (p.m_log2[0] + p.m_log2[1]);
    uint32_t gs = 0;  // slabs processed by this group so far (selects buffers and mbarrier parities)
    float* st_acc = stat_base + group * (EPI_STAT_ACC_FLOATS + EPI_STAT_SCR_FLOATS);
    float* st_scr = st_acc + EPI_STAT_ACC_FLOATS;
    // column offsets of the (at most two) slabs whose warp partials wait in the two halves of st_scr (uniform across
    // the group): with ONE barrier per slab the partials of slab g are complete only after the barrier that closes
    // slab g + 1, so they are added during slab g + 2, just before that half is written again
    int st_pend = -1, st_pend_old = -1;
    // Operand prefetch runs one slab ahead of the math ACROSS tile boundaries (leader thread only): the sequence of
    // (tile, slab) pairs of a group is known up front, and for narrow layers (one slab per tile) the TMA latency of
    // the residual / mask slab would otherwise sit in every tile's critical path.
    // The epilogue of a CTA's LAST tile has nothing to overlap with (no next mainloop), and most launches of this
    // network are 1..3 tiles per CTA: there the group that does not own the tile's TMEM buffer takes the upper half of
    // its slabs, so the exposed tail is half as long.  seg(): the slab range [s0, s1) of tile `it_` for this group.
    const int n_it = (total_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
    auto seg = [&](int it_, int nslab_, int& s0, int& s1) -> bool {
      const bool owner = (it_ & 1) == group;
      if (!(p.epi_split && it_ == n_it - 1 && nslab_ >= 2)) {
        s0 = 0;
        s1 = nslab_;
        return owner;
      }
      const int h = (nslab_ + 1) >> 1;
      s0 = owner ? 0 : h;
      s1 = owner ? h : nslab_;
      return true;
    };
    // Residual / accumulate / mask operands: every thread fetches ITS OWN row of a slab (4 x 16 bytes per operand) with
    // cp.async straight into the swizzled slot it will read itself -- no TMA (a 128-row x 64-byte operand box costs
    // ~1.3 kcycles of TMA row rate, 2.7 kcycles per slab with two operands, against ~0.1 for the LSU path), no
    // mbarrier, no cross-thread hand-off, no registers held across the wait.  The prefetch runs OD - 1 slabs ahead
    // ACROSS tile boundaries: the (tile, slab) sequence of a group is known up front.
    const uint8_t* opsrc[2] = {nullptr, nullptr};
    long long opld[2] = {0, 0};
    {
      int no = 0;
      if (p.n_add) {
        opsrc[no] = (const uint8_t*)(p.epi.accumulate ? p.out : p.epi.res);
        opld[no] = p.epi.accumulate ? p.ld_out : p.epi.ld_res;
        ++no;
      }
      if (p.has_mask) {
        opsrc[no] = (const uint8_t*)p.epi.mask;
        opld[no] = p.epi.ld_mask;
      }
    }
    int pf_it = -1, pf_s = 0, pf_s1 = 0, pf_n0 = 0;
    int pf_ob = 0;  // operand ring slot of the next prefetch
    // The loads of a warp are laid out for the LSU, not for the consumer: lane l fetches 16-byte chunk (l & 3) of rows
    // (l >> 2) + 8 i, i = 0..3, of the warp's 32 rows, so one instruction touches 8 rows x 64 contiguous bytes = 8 cache
    // lines.  (Every thread fetching its OWN row made each instruction touch 32 lines for 16 bytes each; at one tag
    // lookup per line and clock the two epilogue groups kept the LSU busy ~1 kcycle per slab and operand -- the whole
    // difference between an epilogue with a residual and one without.)  The consumer of a row is another lane of the
    // same warp: __syncwarp() after the wait, and the one that ends every slab iteration, order the hand-over.
    int pf_rows[4] = {0, 0, 0, 0};  // output-row index of row (lane >> 2) + 8 i of this warp in the prefetch tile, -1 = outside
    const int pf_c = lane & 3, pf_rl = quarter * 32 + (lane >> 2);
    auto pf_seek = [&]() {  // next tile of this CTA in which this group has slabs; the warp's rows in it
      for (++pf_it; pf_it < n_it; ++pf_it) {
        const int tile_ = blockIdx.x + pf_it * gridDim.x;
        const int n0_ = (tile_ % p.n_tiles) * p.block_n;
        const int nslab_ = (min(p.block_n, p.N - n0_) + EPI_SLAB_COLS - 1) / EPI_SLAB_COLS;
        if (seg(pf_it, nslab_, pf_s, pf_s1)) {
          int st_[3];
          tile_coords(p, tile_ / p.n_tiles, st_);
          const int m1 = st_[0] + r1, m2 = st_[1] + r2, m3 = st_[2] + r3;
          const bool ok = (m1 < p.m_ext[0]) && (m2 < p.m_ext[1]) && (m3 < p.m_ext[2]);
          const long long row = (long long)m1 * p.o_stride[0] + (long long)m2 * p.o_stride[1] + (long long)m3 * p.o_stride[2];
          const int mine = ok ? (int)row : -1;  // (row indices fit 31 bits: the launcher bounds the output space)
#pragma unroll
          for (int i = 0; i < 4; ++i) pf_rows[i] = __shfl_sync(0xffffffffu, mine, (lane >> 2) + 8 * i);
          pf_n0 = n0_;
          return;
        }
      }
    };
    pf_seek();
    auto prefetch_next = [&]() {
      if (pf_it < n_it) {
        const int nb = pf_n0 + pf_s * 32;
        if (nb + 8 * pf_c + 8 <= p.N) {
          for (int o = 0; o < n_ops; ++o) {
            const uint8_t* src = opsrc[o] + ((long long)nb + 8 * pf_c) * 2;
            uint8_t* dst = op_buf + (o * OD + pf_ob) * EPI_SLAB_BYTES;
#pragma unroll
            for (int i = 0; i < 4; ++i)
              if (pf_rows[i] >= 0)
                cp_async16(dst + slab_off(pf_rl + 8 * i, pf_c), src + (long long)pf_rows[i] * opld[o] * 2);
          }
        }
        if (++pf_ob == OD) pf_ob = 0;
        if (++pf_s == pf_s1) pf_seek();
      }
      cp_async_commit();  // one group per call, also when empty: the wait counts stay aligned with the slab index
    };
    if (n_ops > 0)
      for (int i = 0; i < OD - 1; ++i) prefetch_next();  // the first OD - 1 slabs of this group's sequence
    int oslot = 0;     // operand ring slot of the slab being processed
    for (int it = 0; it < n_it; ++it) {
      const int tile = blockIdx.x + it * gridDim.x;
      const int buf = it & 1;
      const int n_tile = tile % p.n_tiles, m_tile = tile / p.n_tiles;
      const int n0 = n_tile * p.block_n;
      const int nslab = (min(p.block_n, p.N - n0) + EPI_SLAB_COLS - 1) / EPI_SLAB_COLS;
      int s0, s1;
      if (!seg(it, nslab, s0, s1)) continue;
      const bool owner = buf == group;
      int st[3];
      tile_coords(p, m_tile, st);
      int o1 = 0, o2 = 0, o3 = 0;
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        const int d = p.o_dim[j];
        o1 = d == 1 ? st[j] : o1;
        o2 = d == 2 ? st[j] : o2;
        o3 = d == 3 ? st[j] : o3;
      }
      const bool valid = (st[0] + r1 < p.m_ext[0]) && (st[1] + r2 < p.m_ext[1]) && (st[2] + r3 < p.m_ext[2]);
      if (g_shift) {
        for (int i = t; i < p.block_n; i += 128) s_shift[i] = (n0 + i < p.N) ? g_shift[n0 + i] : 0.f;
      }
      named_bar(bar_id, 128);  // s_shift visible to the group
      if (leader) pb.begin();
      mbar_wait_warp(&sh->tmem_full[buf], (it >> 1) & 1, 4);
      if (leader) { pb.end(4); pb.begin(); }
      tc_fence_after();
      const uint32_t taddr0 = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(buf * 256);
      for (int s = s0; s < s1; ++s, ++gs) {
        const int b = gs & 1;
        uint32_t raw[32];
        tmem_ld32_issue(taddr0 + (uint32_t)(s * 32), raw);
        // this thread's operand rows of slab gs + OD - 1 (possibly of a later tile) into the ring slot that slab
        // gs - 1 used -- read by this same thread one iteration ago, so plain program order protects it
        if (n_ops > 0) prefetch_next();
        tmem_ld_wait();
        // accumulators are in registers: the owning group hands the TMEM buffer back to the MMA warp (on a shared last
        // tile the other group may still be reading it, but no mainloop follows the last tile)
        if (s == s1 - 1 && owner) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&sh->tmem_empty[buf]);
        }
        if (n_ops > 0) {
          if (OD == 2) cp_async_wait<1>(); else if (OD == 3) cp_async_wait<2>(); else cp_async_wait<3>();
          __syncwarp();  // this thread's operand row was fetched by other lanes of the warp
        }
        if (STATS) {
          // two barriers ago every warp had stored slab gs - 2's partials into scratch half gs & 1; nobody writes that
          // half again before the barrier that closes THIS slab
          if (st_pend_old >= 0 && t < 64) epi_stats_accumulate(st_acc, st_scr + (gs & 1) * EPI_STAT_SCR_HALF, t, st_pend_old);
          st_pend_old = st_pend;
          st_pend = -1;
        }
        // out_buf[b] is free: the leader drained the store that last read it before it arrived at the previous barrier
        uint8_t* ob = out_buf + b * EPI_SLAB_BYTES;
        const uint8_t* op0 = op_buf + oslot * EPI_SLAB_BYTES;
        const uint8_t* op1 = op_buf + (OD + oslot) * EPI_SLAB_BYTES;
        if (++oslot == OD) oslot = 0;
        const float* ss = s_shift + s * 32;
        const bool zero = STATS && !valid;
        switch (variant) {
          case 0: epi_slab_math<false, 0, false, false>(raw, ss, op0, op1, ob, r, zero); break;
          case 1: epi_slab_math<false, 0, false, true>(raw, ss, op0, op1, ob, r, zero); break;
          case 2: epi_slab_math<false, 0, true, false>(raw, ss, op0, op1, ob, r, zero); break;
          case 3: epi_slab_math<false, 0, true, true>(raw, ss, op0, op1, ob, r, zero); break;
          case 4: epi_slab_math<false, 1, false, false>(raw, ss, op0, op1, ob, r, zero); break;
          case 5: epi_slab_math<false, 1, false, true>(raw, ss, op0, op1, ob, r, zero); break;
          case 6: epi_slab_math<false, 1, true, false>(raw, ss, op0, op1, ob, r, zero); break;
          case 7: epi_slab_math<false, 1, true, true>(raw, ss, op0, op1, ob, r, zero); break;
          case 8: epi_slab_math<true, 0, false, false>(raw, ss, op0, op1, ob, r, zero); break;
          case 9: epi_slab_math<true, 0, false, true>(raw, ss, op0, op1, ob, r, zero); break;
          case 10: epi_slab_math<true, 0, true, false>(raw, ss, op0, op1, ob, r, zero); break;
          case 11: epi_slab_math<true, 0, true, true>(raw, ss, op0, op1, ob, r, zero); break;
          case 12: epi_slab_math<true, 1, false, false>(raw, ss, op0, op1, ob, r, zero); break;
          case 13: epi_slab_math<true, 1, false, true>(raw, ss, op0, op1, ob, r, zero); break;
          case 14: epi_slab_math<true, 1, true, false>(raw, ss, op0, op1, ob, r, zero); break;
          default: epi_slab_math<true, 1, true, true>(raw, ss, op0, op1, ob, r, zero); break;
        }
        fence_proxy_async();
        if (leader) {
          // ONE barrier per slab: before arriving, the leader makes sure the store issued one slab ago has finished
          // reading the OTHER output buffer, so everybody may overwrite it right after this barrier
          Probe pw{p.dbg, 0};
          pw.begin();
          bulk_wait_read<0>();
          pw.end(6);
        }
        named_bar(bar_id, 128);  // slab complete; the other output buffer is free
        if (leader) {
          tma_store(&p.tm_out, ob, p.rank_o, n0 + s * 32, o1, o2, o3);
          bulk_commit();
        }
        if (STATS) {
          epi_slab_stats_part(ob, t, st_scr + (gs & 1) * EPI_STAT_SCR_HALF);
          st_pend = s * 32;
        }
        __syncwarp();
      }
      if (leader) pb.end(5);  // tmem_full seen -> last slab of the tile handed to the store engine
    }
    if (STATS) {
      named_bar(bar_id, 128);
      if (st_pend_old >= 0 && t < 64) epi_stats_accumulate(st_acc, st_scr + (gs & 1) * EPI_STAT_SCR_HALF, t, st_pend_old);
      if (st_pend >= 0 && t < 64) epi_stats_accumulate(st_acc, st_scr + ((gs + 1) & 1) * EPI_STAT_SCR_HALF, t, st_pend);
    }
    if (leader) {
      pb.begin();
      bulk_wait_all();
      pb.end(7);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 0) {
    pb_all.end(0);
    if (p.dbg) atomicAdd(p.dbg + 8, 1ull);
  }
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
  if (p.dbg && threadIdx.x == 0) {
    unsigned long long g3;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(g3));
    atomicMax(p.dbg + 12, g3);
  }
  if (STATS) {
    // this CTA's row of the workspace: ws[n_tile][row][2][block_n]; the last CTA of the column (ticket) adds the rows
    // in ascending order in double and writes the final sums (stats_reduce.cuh)
    const int n_tile = (int)blockIdx.x % p.n_tiles, prow = (int)blockIdx.x / p.n_tiles, P = (int)gridDim.x / p.n_tiles;
    const int bn = p.block_n;
    float* col = p.epi.stats_ws + (long long)n_tile * P * 2 * bn;
    float* my = col + (long long)prow * 2 * bn;
    const float* g0 = stat_base, * g1 = stat_base + EPI_STAT_ACC_FLOATS + EPI_STAT_SCR_FLOATS;
    for (int i = (int)threadIdx.x; i < 2 * bn; i += (int)blockDim.x) {
      const int half = i >= bn ? 1 : 0, ch = i - half * bn;
      my[i] = g0[half * 256 + ch] + g1[half * 256 + ch];
    }
    if (stats_last_cta(p.epi.stats_ticket + n_tile, (unsigned)P, &sh->stat_flag)) {
      double* out = p.epi.stats;
      const int stride = p.epi.stats_stride, N = p.N, n0 = n_tile * bn;
      stats_ordered_reduce(col, P, 2 * bn, 2 * bn, (double*)smem, [&](int s_, double v) {
        const int half = s_ >= bn ? 1 : 0, n = n0 + s_ - half * bn;
        if (n < N) out[half * stride + n] = v;
      });
    }
  }
}
