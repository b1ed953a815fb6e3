// Log-mel extraction on the tensor cores (tcgen05 / TMEM), the shipped configuration of the reference:
// 24 kHz, n_fft 2048, 10 ms Hann window (240 samples), 5 ms hop (120 samples), 128 HTK mel bands
//   /root/reference/audio_slowfast/datasets/audio_loader_epic.py:66-97 (+ edge padding :151-154).
//
// The STFT of a frame is a dense contraction of its 240 non-zero windowed samples with a [240 x 2*1025] basis.
// Because win = 2*hop, frame t is the concatenation of two consecutive 120-sample "chunk rows" of the clip, so
//
//     DFT[frame t] = chunk_row[t] . B_top  +  chunk_row[t+1] . B_bot            (K = 120 each)
//
// and ONE staged copy of 129 chunk rows serves both halves of 128 frames: the A operand is kept in the canonical
// no-swizzle K-major layout with 8-row groups packed back to back (SBO = 128 B), which makes a row a 16-byte pitch
// and "the same matrix shifted by one row" simply the descriptor start address + 16.
//
// fp32 accuracy (log-mel within 1e-4 of the reference, bins 80 dB below the peak included) on TF32 tensor cores
// comes from the 3-term split  a.b ~= a_hi.b_hi + a_lo.b_hi + a_hi.b_lo  (hi = round-to-nearest TF32, lo = TF32 of
// the remainder): three kind::tf32 MMAs per K = 8 step accumulate into the same fp32 TMEM accumulator.
//
// One CTA per SM, persistent over tiles of 128 frame rows of the flattened (clip, frame) space:
//   warp 0      streams the basis (pre-split, pre-arranged as the shared-memory image) with cp.async.bulk
//   warp 1      issues tcgen05.mma: per tile 10 bin chunks x 2 halves x 5 slabs x 3 K-steps x 3 split terms
//   warps 2-5   drain TMEM (double-buffered 208-column accumulators = re | im of 104 bins): magnitude, then the
//               HTK mel filterbank as a running two-filter accumulation (a bin touches the descending side of
//               filter j-1 and the ascending side of filter j), log, store -- one thread per frame
//   warps 6-9   stage the next tile's chunk rows: coalesced loads, hi/lo split, conflict-free smem stores
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "common.cuh"
#include "sm100_ptx.cuh"

namespace asf {

constexpr int LT_HOP = 120;
constexpr int LT_K4H = LT_HOP / 4;         // 16-byte K groups per half
constexpr int LT_NBC = 104;                // bins per chunk
constexpr int LT_NCOL = 2 * LT_NBC;        // accumulator columns: re | im
constexpr int LT_KS = 24;                  // K floats per pipeline slab (3 MMA K-steps; a tcgen05.commit costs a ~200-cycle
                                           // bubble of the tensor pipe, so slabs are as large as shared memory allows)
constexpr int LT_SLABS = LT_HOP / LT_KS;   // 5 per half
constexpr int LT_NR = 137;                 // padded row count of the A image (row pitch 16 B, K-group pitch 137*16)
constexpr int LT_A_LBO = LT_NR * 16;       // 2192 B between K groups: 548 words = 4 (mod 32) -> conflict-free staging
constexpr int LT_A_BYTES = LT_K4H * LT_A_LBO;          // 65,760
constexpr int LT_A_STRIDE = 65792;                     // hi / lo images, 128-byte aligned
constexpr int LT_B_LBO = LT_NCOL * 16;                 // 3328
constexpr int LT_B_HALF = (LT_KS / 4) * LT_B_LBO;      // 19,968: hi (or lo) part of a stage
constexpr int LT_B_STAGE = 2 * LT_B_HALF;              // 39,936
constexpr int LT_STAGES = 2;
constexpr int LT_MAX_BINS = 1040;                      // 10 chunks (n_fft <= 2048)
constexpr int LT_OFF_B = 2 * LT_A_STRIDE;
constexpr int LT_OFF_W = LT_OFF_B + LT_STAGES * LT_B_STAGE;
constexpr int LT_OFF_J = LT_OFF_W + LT_MAX_BINS * 8;
constexpr int LT_OFF_BAR = LT_OFF_J + LT_MAX_BINS * 4;
constexpr int LT_SMEM = LT_OFF_BAR + 24 * 8;
constexpr int LT_THREADS = 320;
static_assert(LT_A_BYTES <= LT_A_STRIDE && LT_SMEM <= 227 * 1024, "log-mel tile does not fit in shared memory");

struct LtBars {
  uint64_t b_full[LT_STAGES], b_empty[LT_STAGES];
  uint64_t a_full, a_free;
  uint64_t tmem_full[2], tmem_empty[2];
  uint32_t tmem_base;
};
static_assert(sizeof(LtBars) <= 24 * 8, "barrier block");

// no-swizzle K-major operand: core matrices (8 rows x 16 B) LBO bytes apart along K, SBO bytes apart along M/N
__device__ __forceinline__ uint64_t lt_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  return d;
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}\n"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst)), "l"((uint64_t)src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld8_issue(uint32_t taddr, uint32_t* r) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];\n"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr));
}
// a finished mel band of one frame: log, store (and the np.pad(..., "edge") copies of the clip's last frame)
__device__ __noinline__ void lt_emit(float* orow, int m, float v, float eps, bool ok, int n_rep, int n_mels) {
  if (ok) {
    const float l = logf(v + eps);
    orow[m] = l;
    for (int q = 1; q <= n_rep; ++q) orow[(long long)q * n_mels + m] = l;
  }
}
__device__ __forceinline__ void lt_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ uint32_t tf32_rna(float x) {
  uint32_t u;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(x));
  return u;
}

__global__ void __launch_bounds__(LT_THREADS, 1)
logmel_tc_kernel(const float* __restrict__ wave, long long wave_ld, int n_samples, int T_valid, int T_out,
                 int n_chunks, int n_mels, float eps, const float* __restrict__ btab,
                 const float2* __restrict__ melw, const int* __restrict__ melj, float* __restrict__ out,
                 long long total_rows, int n_tiles, int dbg, unsigned long long* probe) {
  extern __shared__ __align__(128) uint8_t smem[];
  pdl_trigger();
  LtBars* bars = (LtBars*)(smem + LT_OFF_BAR);
  float2* sw = (float2*)(smem + LT_OFF_W);
  int* sj = (int*)(smem + LT_OFF_J);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int RP = T_valid + 1;  // chunk rows (= frame rows, the last one unused) per clip
  if (threadIdx.x == 0) {
    for (int s = 0; s < LT_STAGES; ++s) {
      mbar_init(&bars->b_full[s], 1);
      mbar_init(&bars->b_empty[s], 1);
    }
    mbar_init(&bars->a_full, 128);
    mbar_init(&bars->a_free, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bars->tmem_full[i], 1);
      mbar_init(&bars->tmem_empty[i], 4);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(&bars->tmem_base, 512);
  for (int i = threadIdx.x; i < n_chunks * LT_NBC; i += LT_THREADS) {  // constant tables (never written by a kernel)
    sw[i] = melw[i];
    sj[i] = melj[i];
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;
  pdl_wait();
  long long pt = 0;
#define LT_PB() do { if (probe) pt = clock64(); } while (0)
#define LT_PE(slot) do { if (probe) atomicAdd(probe + (slot), (unsigned long long)(clock64() - pt)); } while (0)
  const long long t_start = clock64();

  if (warp == 0) {
    // ===== basis producer =====
    if (lane == 0) {
      int st = 0;
      uint32_t ph = 0;
      for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const uint8_t* src = (const uint8_t*)btab;
        for (int i = 0; i < n_chunks * 2 * LT_SLABS; ++i) {
          LT_PB();
          mbar_wait(&bars->b_empty[st], ph ^ 1, 21);
          LT_PE(1);
          const uint32_t nbytes = (dbg & 1) ? 16u : (uint32_t)LT_B_STAGE;  // dbg bit 0: no basis traffic (wrong results)
          mbar_expect_tx(&bars->b_full[st], nbytes);
          bulk_g2s(smem + LT_OFF_B + st * LT_B_STAGE, src, nbytes, &bars->b_full[st]);
          src += LT_B_STAGE;
          if (++st == LT_STAGES) { st = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (lane == 0) {
      const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(LT_NCOL >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
      const uint32_t s0 = smem_u32(smem);
      const uint64_t a_hi0 = lt_desc(s0, LT_A_LBO, 128), a_lo0 = lt_desc(s0 + LT_A_STRIDE, LT_A_LBO, 128);
      const uint64_t b0 = lt_desc(s0 + LT_OFF_B, LT_B_LBO, 128);
      int st = 0;
      uint32_t ph = 0;
      uint32_t g = 0;  // accumulator use counter
      int it = 0;
      for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
        LT_PB();
        mbar_wait(&bars->a_full, (uint32_t)(it & 1), 22);
        LT_PE(2);
        tc_fence_after();
        for (int c = 0; c < n_chunks; ++c, ++g) {
          const uint32_t buf = g & 1;
          LT_PB();
          mbar_wait(&bars->tmem_empty[buf], ((g >> 1) & 1) ^ 1, 23);
          LT_PE(3);
          tc_fence_after();
          const uint32_t d = tmem_base + buf * 256;
          uint32_t acc = 0;
          for (int h = 0; h < 2; ++h) {
            for (int s = 0; s < LT_SLABS; ++s) {
              LT_PB();
              mbar_wait(&bars->b_full[st], ph, 24);
              LT_PE(4);
              tc_fence_after();
              const uint64_t bh = b0 + (uint64_t)((st * LT_B_STAGE) >> 4), bl = bh + (uint64_t)(LT_B_HALF >> 4);
#pragma unroll
              for (int j = 0; j < LT_KS / 8; ++j) {
                const uint64_t ao = (uint64_t)((((s * (LT_KS / 4) + 2 * j) * LT_A_LBO) + h * 16) >> 4);
                const uint64_t bo = (uint64_t)((2 * j * LT_B_LBO) >> 4);
                umma_tf32(d, a_lo0 + ao, bh + bo, idesc, acc);
                umma_tf32(d, a_hi0 + ao, bl + bo, idesc, 1);
                umma_tf32(d, a_hi0 + ao, bh + bo, idesc, 1);
                acc = 1;
              }
              umma_commit(&bars->b_empty[st]);
              if (++st == LT_STAGES) { st = 0; ph ^= 1; }
            }
          }
          umma_commit(&bars->tmem_full[buf]);
        }
        umma_commit(&bars->a_free);  // every MMA that reads this tile's chunk rows has completed
      }
    }
  } else if (warp < 6) {
    // ===== epilogue: one thread per frame =====
    const int quarter = warp & 3;
    const int r = quarter * 32 + lane;
    uint32_t g = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
      const int R = tile * 128 + r;
      const int b = R / RP, t = R - b * RP;
      const bool valid = (long long)R < total_rows && t < T_valid;
      float* orow = out + ((long long)b * T_out + t) * n_mels;
      const int n_rep = (valid && t == T_valid - 1) ? (T_out - T_valid) : 0;  // np.pad(..., "edge") rows
      float cur = 0.f, nxt = 0.f;
      int j_cur = 0;
      auto emit = [&](int m, float v) { lt_emit(orow, m, v, eps, valid && m >= 0 && m < n_mels, n_rep, n_mels); };
      for (int c = 0; c < n_chunks; ++c, ++g) {
        const uint32_t buf = g & 1;
        if (threadIdx.x == 64) LT_PB();
        mbar_wait_warp(&bars->tmem_full[buf], (g >> 1) & 1, 25, 0);
        if (threadIdx.x == 64) { LT_PE(5); LT_PB(); }
        tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + buf * 256;
#pragma unroll 1
        for (int j0 = 0; j0 < ((dbg & 2) ? 8 : LT_NBC); j0 += 8) {  // dbg bit 1: epilogue touches one column group
          uint32_t re[8], im[8];
          tmem_ld8_issue(taddr + (uint32_t)j0, re);
          tmem_ld8_issue(taddr + (uint32_t)(LT_NBC + j0), im);
          const int k0 = c * LT_NBC + j0;
          // tables of the eight bins (independent of the accumulators: issued under the TMEM load latency)
          const int4 ja = *(const int4*)(sj + k0), jb = *(const int4*)(sj + k0 + 4);
          const float4 wa = *(const float4*)(sw + k0), wb = *(const float4*)(sw + k0 + 2),
                       wc = *(const float4*)(sw + k0 + 4), wd = *(const float4*)(sw + k0 + 6);
          const int jk[8] = {ja.x, ja.y, ja.z, ja.w, jb.x, jb.y, jb.z, jb.w};
          const float wdn[8] = {wa.x, wa.z, wb.x, wb.z, wc.x, wc.z, wd.x, wd.z};
          const float wup[8] = {wa.y, wa.w, wb.y, wb.w, wc.y, wc.w, wd.y, wd.w};
          lt_ld_wait();
          // magnitudes, branch-free and independent of each other: s * rsqrt(s) + one Newton step (~1 ulp)
          float mag[8];
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const float x = __uint_as_float(re[q]), y = __uint_as_float(im[q]);
            const float s2 = fmaf(x, x, y * y);
            float rs;
            asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(rs) : "f"(s2));
            const float m0 = s2 * rs;
            const float m1 = fmaf(fmaf(-m0, m0, s2), 0.5f * rs, m0);
            mag[q] = s2 >= 1e-36f ? m1 : 0.f;
          }
          if (jk[7] == j_cur) {  // no filter boundary inside the group (uniform): two short FMA chains
#pragma unroll
            for (int q = 0; q < 8; ++q) {
              cur = fmaf(wdn[q], mag[q], cur);
              nxt = fmaf(wup[q], mag[q], nxt);
            }
          } else {
#pragma unroll
            for (int q = 0; q < 8; ++q) {
              while (j_cur < jk[q]) {  // uniform across the CTA: filter j_cur - 1 is complete
                emit(j_cur - 1, cur);
                cur = nxt;
                nxt = 0.f;
                ++j_cur;
              }
              cur = fmaf(wdn[q], mag[q], cur);
              nxt = fmaf(wup[q], mag[q], nxt);
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars->tmem_empty[buf]);
        if (threadIdx.x == 64) LT_PE(6);
      }
      while (j_cur <= n_mels) {
        emit(j_cur - 1, cur);
        cur = nxt;
        nxt = 0.f;
        ++j_cur;
      }
    }
  } else {
    // ===== chunk-row stager (128 threads) =====
    // Warp w of the stagers owns chunk rows w, w + 4, ... (33 at most); a lane holds samples lane, lane + 32, ... of
    // each row.  The NEXT tile's samples are loaded into registers while the current tile computes, so that once the
    // MMAs release the image only the hi/lo split and the (bank-conflict-free) stores remain.
    const int sidx = (int)threadIdx.x - 192;
    const int sw_ = (int)(threadIdx.x >> 5) - 6;
    constexpr int LT_RPW = 33;
    float xr[LT_RPW][4];
    auto load_tile = [&](int tile) {
      const int row0 = tile * 128;  // < 2^31 (checked by the launcher)
#pragma unroll
      for (int i = 0; i < LT_RPW; ++i) {
        const int r2 = sw_ + 4 * i;
        const int R = row0 + r2;
        const int b = R / RP, t = R - b * RP;
        const bool row_ok = tile < n_tiles && r2 < 129 && (long long)R < total_rows && t >= 1;
        const float* src = wave + (long long)b * wave_ld + (long long)(t - 1) * LT_HOP;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int k = lane + 32 * u;
          xr[i][u] = (row_ok && k < LT_HOP && (t - 1) * LT_HOP + k < n_samples) ? src[k] : 0.f;
        }
      }
    };
    load_tile(blockIdx.x);
    int it = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
      if (sidx == 0) LT_PB();
      mbar_wait_warp(&bars->a_free, (uint32_t)((it & 1) ^ 1), 26, 256);
      if (sidx == 0) { LT_PE(7); LT_PB(); }
#pragma unroll
      for (int i = 0; i < LT_RPW; ++i) {
        const int r2 = sw_ + 4 * i;
        if (r2 < 129) {
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int k = lane + 32 * u;
            if (k < LT_HOP) {
              const uint32_t hi = tf32_rna(xr[i][u]);
              const uint32_t lo = tf32_rna(xr[i][u] - __uint_as_float(hi));
              const int off = (k >> 2) * LT_A_LBO + r2 * 16 + (k & 3) * 4;
              *(uint32_t*)(smem + off) = hi;
              *(uint32_t*)(smem + LT_A_STRIDE + off) = lo;
            }
          }
        }
      }
      fence_proxy_async();
      mbar_arrive(&bars->a_full);
      if (sidx == 0) LT_PE(8);
      load_tile(tile + (int)gridDim.x);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (probe && threadIdx.x == 0) {
    atomicAdd(probe + 0, (unsigned long long)(clock64() - t_start));
    atomicAdd(probe + 9, 1ull);
  }
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

static int lt_dbg() {  // ASF_LOGMEL_DBG: timing experiments only (bit 0 / bit 1 switch off basis traffic / epilogue)
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("ASF_LOGMEL_DBG");
    v = e ? atoi(e) : 0;
  }
  return v;
}

static inline float host_tf32_rna(float x) {
  uint32_t u;
  memcpy(&u, &x, 4);
  u = (u + 0x1000u) & 0xFFFFE000u;
  float r;
  memcpy(&r, &u, 4);
  return r;
}

}  // namespace asf

using namespace asf;

extern "C" {

int asf_logmel_tc_supported(int n_fft, int win, int hop) {
  return (win == 2 * LT_HOP && hop == LT_HOP && n_fft >= win && (n_fft / 2 + 1 + LT_NBC - 1) / LT_NBC * LT_NBC <= LT_MAX_BINS)
             ? 1 : 0;
}

long long asf_logmel_tc_btab_floats(int n_fft) {
  const int n_chunks = (n_fft / 2 + 1 + LT_NBC - 1) / LT_NBC;
  return (long long)n_chunks * 2 * LT_SLABS * (LT_B_STAGE / 4);
}

int asf_logmel_tc_tables(float* btab, float* melw, int32_t* melj, int sr, int n_fft, int win, int hop, int n_mels,
                         void* stream) {
  ASF_CHECK_ARG(btab && melw && melj && sr > 0 && n_mels > 0, "asf_logmel_tc_tables: bad arguments");
  ASF_CHECK_ARG(asf_logmel_tc_supported(n_fft, win, hop), "asf_logmel_tc_tables: only win = 240, hop = 120, n_fft <= 2048");
  const int n_bins = n_fft / 2 + 1;
  const int n_chunks = (n_bins + LT_NBC - 1) / LT_NBC, nbp = n_chunks * LT_NBC;
  const int lpad = (n_fft - win) / 2;
  const double PI = 3.14159265358979323846;
  std::vector<float> hb((size_t)asf_logmel_tc_btab_floats(n_fft), 0.f);
  for (int c = 0; c < n_chunks; ++c)
    for (int h = 0; h < 2; ++h)
      for (int s = 0; s < LT_SLABS; ++s) {
        float* stage = hb.data() + (size_t)((c * 2 + h) * LT_SLABS + s) * (LT_B_STAGE / 4);
        for (int kl = 0; kl < LT_KS; ++kl) {
          const int j = h * LT_HOP + s * LT_KS + kl;  // sample inside the window
          const double w = 0.5 - 0.5 * cos(2.0 * PI * (double)j / (double)win);  // periodic Hann
          for (int n = 0; n < LT_NCOL; ++n) {
            const int bin = c * LT_NBC + (n % LT_NBC);
            float v = 0.f;
            if (bin < n_bins) {
              const int64_t phs = ((int64_t)bin * (lpad + j)) % n_fft;  // exact phase reduction
              const double ang = 2.0 * PI * (double)phs / (double)n_fft;
              v = (n < LT_NBC) ? (float)(w * cos(ang)) : (float)(-w * sin(ang));
            }
            const float hi = host_tf32_rna(v), lo = host_tf32_rna(v - hi);
            const size_t idx = ((size_t)(kl >> 2) * LT_NCOL + n) * 4 + (kl & 3);
            stage[idx] = hi;
            stage[LT_B_HALF / 4 + idx] = lo;
          }
        }
      }
  // librosa.filters.mel(htk=True, norm=None) as a per-bin pair of adjacent filters
  const double mel_max = 2595.0 * log10(1.0 + (0.5 * sr) / 700.0);
  std::vector<double> mel_f(n_mels + 2);
  for (int i = 0; i < n_mels + 2; ++i) {
    double m = (i == n_mels + 1) ? mel_max : (mel_max / (double)(n_mels + 1)) * (double)i;
    mel_f[i] = 700.0 * (pow(10.0, m / 2595.0) - 1.0);
  }
  std::vector<float> hw((size_t)nbp * 2, 0.f);
  std::vector<int32_t> hj((size_t)nbp, 0);
  int jprev = 0;
  for (int k = 0; k < nbp; ++k) {
    if (k >= n_bins) {
      hj[k] = jprev;
      continue;
    }
    const double fk = (double)k * (double)sr / (double)n_fft;
    int j = -1;
    for (int i = 0; i < n_mels + 2; ++i)
      if (mel_f[i] <= fk) j = i;
    if (j < 0) j = 0;
    if (j > n_mels) j = n_mels;
    if (j < jprev) j = jprev;
    for (int m = 0; m < n_mels; ++m) {
      const double lower = (fk - mel_f[m]) / (mel_f[m + 1] - mel_f[m]);
      const double upper = (mel_f[m + 2] - fk) / (mel_f[m + 2] - mel_f[m + 1]);
      double v = lower < upper ? lower : upper;
      if (v < 0.0) v = 0.0;
      const float vf = (float)v;
      if (m == j - 1) hw[2 * k] = vf;
      else if (m == j) hw[2 * k + 1] = vf;
      else
        ASF_CHECK_ARG(vf == 0.f, "asf_logmel_tc_tables: bin %d touches mel filter %d outside its adjacent pair (%d, %d)",
                      k, m, j - 1, j);
    }
    hj[k] = j;
    jprev = j;
  }
  cudaStream_t st = (cudaStream_t)stream;
  ASF_CUDA(cudaMemcpyAsync(btab, hb.data(), hb.size() * sizeof(float), cudaMemcpyHostToDevice, st));
  ASF_CUDA(cudaMemcpyAsync(melw, hw.data(), hw.size() * sizeof(float), cudaMemcpyHostToDevice, st));
  ASF_CUDA(cudaMemcpyAsync(melj, hj.data(), hj.size() * sizeof(int32_t), cudaMemcpyHostToDevice, st));
  ASF_CUDA(cudaStreamSynchronize(st));  // host staging vectors die at return
  return 0;
}

int asf_logmel_fwd_tc(const float* wave, int64_t wave_ld, int B, int n_samples, int n_fft, int win, int hop,
                      int n_mels, float eps, const float* btab, const float* melw, const int32_t* melj, float* fast,
                      int T_out, void* stream) {
  ASF_CHECK_ARG(wave && btab && melw && melj && fast, "asf_logmel_fwd_tc: null pointer");
  ASF_CHECK_ARG(asf_logmel_tc_supported(n_fft, win, hop), "asf_logmel_fwd_tc: only win = 240, hop = 120, n_fft <= 2048");
  ASF_CHECK_ARG(B > 0 && n_samples >= 0 && n_mels > 0 && wave_ld >= n_samples, "asf_logmel_fwd_tc: bad sizes");
  const int T_valid = 1 + n_samples / hop;
  if (T_out <= 0) T_out = T_valid;
  ASF_CHECK_ARG(T_out >= T_valid, "asf_logmel_fwd_tc: T_out (%d) smaller than the %d frames of the clip (no truncation)",
                T_out, T_valid);
  const int n_chunks = (n_fft / 2 + 1 + LT_NBC - 1) / LT_NBC;
  const long long total_rows = (long long)B * (T_valid + 1);
  const long long n_tiles = cdiv(total_rows, 128);
  ASF_CHECK_ARG(total_rows + 256 < (1ll << 31), "asf_logmel_fwd_tc: batch too large (B * frames must stay below 2^31)");
  static bool attr_set = false;
  if (!attr_set) {
    ASF_CUDA(cudaFuncSetAttribute(logmel_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, LT_SMEM));
    attr_set = true;
  }
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  {
    // ASF_LOGMEL_CTAS: SM budget of the extraction when it runs as a side branch of the training step
    // (TrainStep(pipeline=True)): a full-chip persistent launch holds every SM for its whole duration
    const char* e = getenv("ASF_LOGMEL_CTAS");
    const int cap = e ? atoi(e) : 0;
    if (cap > 0 && cap < sms) sms = cap;
  }
  const int grid = (int)(n_tiles < sms ? n_tiles : sms);
  unsigned long long* probe = nullptr;
  if (lt_dbg() & 32) {  // ASF_LOGMEL_DBG bit 5: print the stall counters of this launch (synchronises)
    static unsigned long long* dprobe = nullptr;
    if (!dprobe) ASF_CUDA(cudaMalloc(&dprobe, 16 * sizeof(unsigned long long)));
    ASF_CUDA(cudaMemsetAsync(dprobe, 0, 16 * sizeof(unsigned long long), (cudaStream_t)stream));
    probe = dprobe;
  }
  ASF_CUDA(launch_pdl(logmel_tc_kernel, dim3(grid), dim3(LT_THREADS), (size_t)LT_SMEM, (cudaStream_t)stream, wave,
                      (long long)wave_ld, n_samples, T_valid, T_out, n_chunks, n_mels, eps, btab, (const float2*)melw,
                      (const int*)melj, fast, total_rows, (int)n_tiles, lt_dbg(), probe));
  ASF_LAUNCH_CHECK();
  if (probe) {
    unsigned long long h[16];
    ASF_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    ASF_CUDA(cudaMemcpy(h, probe, sizeof(h), cudaMemcpyDeviceToHost));
    const double n = h[9] ? (double)h[9] : 1.0;
    fprintf(stderr, "logmel_tc probe (kcycles per CTA, %d tiles on %d CTAs): body %.1f | producer wait stage %.1f | mma wait A %.1f, "
            "wait acc %.1f, wait basis %.1f | epilogue wait acc %.1f, busy %.1f | stager wait free %.1f, busy %.1f\n",
            (int)n_tiles, grid, h[0] / n / 1e3, h[1] / n / 1e3, h[2] / n / 1e3, h[3] / n / 1e3, h[4] / n / 1e3, h[5] / n / 1e3,
            h[6] / n / 1e3, h[7] / n / 1e3, h[8] / n / 1e3);
  }
  return 0;
}

}  // extern "C"
