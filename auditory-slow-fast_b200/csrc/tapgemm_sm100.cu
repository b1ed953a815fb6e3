// tap-GEMM: the implicit-GEMM convolution engine of the hot path, written directly against the Blackwell
// (sm_100a) execution model -- TMA (cp.async.bulk.tensor) stages bf16 operand tiles into 128B-swizzled shared
// memory, ONE elected thread issues tcgen05.mma (UMMA 128xNx16, kind::f16) with fp32 accumulators in TMEM, and
// four epilogue warps drain TMEM with tcgen05.ld and apply the fused BN/ReLU/residual/mask/stats epilogue.
//
//   mode 0 (forward conv, dgrad):  D[m][n] = sum_tap sum_c A[coord(m)+off(tap)][c] * W[n][tap][c]
//        persistent CTAs (one per SM), 128 x BLOCK_N output tiles, double-buffered TMEM accumulators so the
//        epilogue of tile i overlaps the MMAs of tile i+1, S-stage smem ring (full/empty mbarriers).
//        Convolution padding = TMA out-of-bounds zero fill; taps = coordinate offsets on the tensor map;
//        stride-2 / stride-4 convs = a parity / phase dimension in the tensor map.  No im2col, no halo copies.
//   mode 1 (wgrad): dW[n][tap][c] = sum_rows Y[row][n] * X[row+off(tap)][c]; both operands are read straight from
//        the channels-last activations as MN-major UMMA operands (rows are the GEMM K), split over row tiles
//        across CTAs; every CTA STORES its partial tile into the layer's split-K workspace and asf_unpack_wgrad(_multi)
//        adds the partials in a fixed order (no floating-point atomics anywhere: results are bit-reproducible).
//   Narrow layers (<= 64 output channels) take the warp-MMA kernels of narrow_conv.cuh / narrow_wgrad_mma.cuh behind
//   the same entry point.
//
// Replaces cuDNN conv fwd/dgrad/wgrad + the separate BN / ReLU / add / cat kernels behind
//   /root/reference/audio_slowfast/models/resnet_helper.py:170-225, :310-341 and audio_model_builder.py:90-111.
#include <cuda.h>

#include "common.cuh"
#include "sm100_ptx.cuh"
#include "stats_reduce.cuh"

namespace asf {

// ------------------------------------------------------------------------------------------------ parameters
constexpr int TG_MAX_STAGES = 8;
constexpr int TG_BLOCK_M = 128;
constexpr int TG_KC = 64;                 // bf16 elements per 128B swizzle row
constexpr int TG_A_BYTES = TG_BLOCK_M * 128;

struct TgParams {
  CUtensorMap tm_a;  // mode 0: activations (A operand).   mode 1: X (activations, UMMA B operand)
  CUtensorMap tm_b;  // mode 0: weights [K][taps_w][N].     mode 1: Y (output grads, UMMA A operand)
  int rank_a, rank_b;
  int m_ext[3], m_log2[3], m_to_a[3], m_tiles[3], y_m_to_a[3];
  int m_tiles_total;
  int ntaps;
  int a_off[ASF_MAX_TAPS][4];
  int b_tap[ASF_MAX_TAPS];
  int K, N, kchunks, block_n, n_tiles, stages;
  // mode 0 output
  void* out;
  int out_bf16;
  int vec_ok;
  long long ld_out;
  long long o_stride[3];
  asf_epilogue_t epi;
  // mode 0, TMA epilogue: output / operand tensor maps over the output M-space, 32-channel boxes, 64B swizzle
  CUtensorMap tm_out;
  int rank_o;            // rank of tm_out / tm_op (1 + mapped M dims)
  int o_dim[3];          // M-dim j -> coordinate slot (1..3) of the output maps, 0 = unmapped
  int n_add, has_mask;   // operands: n_add in {0,1} additive, has_mask in {0,1}
  int epi_smem_off;      // byte offset of the epilogue staging region inside dynamic smem
  int op_depth;          // residual / mask slabs in flight per epilogue group (2..4): op_depth - 1 slabs of TMA prefetch
  int epi_split;         // both epilogue groups share the slabs of a CTA's LAST tile (nothing left to overlap it with)
  // mode 1
  float* dwp;
  int ntaps_w, ksplit, c_tiles, r_tiles;
  int wg_ny, wg_nx, wg_halves, wg_tn, wg_tmem_cols;  // second-generation wgrad tile (tapgemm_wgrad.cuh)
  unsigned long long* dbg;  // diagnostic stall counters (asf_tapgemm_probe), normally NULL
};

// stall probe: cycles a role spent inside a wait, summed over CTAs into dbg[slot]
struct Probe {
  unsigned long long* d;
  long long t;
  __device__ __forceinline__ void begin() { if (d) t = clock64(); }
  __device__ __forceinline__ void end(int slot) { if (d) atomicAdd(d + slot, (unsigned long long)(clock64() - t)); }
};

struct TgShared {
  uint64_t full[TG_MAX_STAGES];
  uint64_t empty[TG_MAX_STAGES];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint32_t tmem_base;
  uint32_t stat_flag;  // "this CTA drew the last statistics ticket" (stats_reduce.cuh)
};

__device__ __forceinline__ void tile_coords(const TgParams& p, int m_tile, int start[3]) {
  int i0 = m_tile % p.m_tiles[0];
  int r = m_tile / p.m_tiles[0];
  int i1 = r % p.m_tiles[1];
  int i2 = r / p.m_tiles[1];
  start[0] = i0 << p.m_log2[0];
  start[1] = i1 << p.m_log2[1];
  start[2] = i2 << p.m_log2[2];
}

// ------------------------------------------------------------------------------------------------ epilogue helpers
__device__ __forceinline__ void tmem_ld32_issue(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

#include "tapgemm_fwd_tma.cuh"

#include "tapgemm_wgrad.cuh"

// ------------------------------------------------------------------------------------------------ host side
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode() {
  static PFN_encodeTiled fn = nullptr;
  if (!fn) {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (PFN_encodeTiled)f;
  }
  return fn;
}

// bf16 tensor map: dim0 = channels (extent c_ext, box 64 = one 128B swizzle row), dims 1.. = (ext, stride) pairs
static int encode_map(CUtensorMap* tm, const void* base, int c_ext, int nd, const int64_t* ext, const int64_t* stride,
                      const int* box, int inner_box = TG_KC, CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_128B) {
  PFN_encodeTiled enc = get_encode();
  if (!enc) {
    set_error("cuTensorMapEncodeTiled entry point not available (no CUDA driver?)");
    return ASF_E_DRIVER;
  }
  cuuint64_t gdim[5], gstr[4];
  cuuint32_t bdim[5], estr[5];
  gdim[0] = (cuuint64_t)c_ext;
  bdim[0] = (cuuint32_t)inner_box;
  estr[0] = 1;
  for (int i = 0; i < nd; ++i) {
    gdim[i + 1] = (cuuint64_t)ext[i];
    gstr[i] = (cuuint64_t)stride[i] * 2;
    bdim[i + 1] = (cuuint32_t)box[i];
    estr[i + 1] = 1;
    if ((gstr[i] & 15) != 0) {
      set_error("tensor map: stride %lld elements is not a multiple of 16 bytes", (long long)stride[i]);
      return ASF_E_BADARG;
    }
  }
  if (((uintptr_t)base & 15) != 0) {
    set_error("tensor map: base pointer not 16-byte aligned");
    return ASF_E_BADARG;
  }
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)(nd + 1), const_cast<void*>(base), gdim, gstr, bdim,
                   estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (rank %d, c_ext %d)", (int)r, nd + 1, c_ext);
    return ASF_E_DRIVER;
  }
  return 0;
}

static int ilog2_exact(int v) {
  int l = 0;
  while ((1 << l) < v) ++l;
  return ((1 << l) == v) ? l : -1;
}

static int num_sms() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}
// SMs this launch may occupy: the engine runs the slow and the fast pathway on two streams and gives each a share of
// the SMs (asf_tapgemm_t::max_ctas), so that their persistent CTAs run side by side instead of queueing.
static int sm_budget(const asf_tapgemm_t* q) {
  const int n = num_sms();
  return (q->max_ctas > 0 && q->max_ctas < n) ? q->max_ctas : n;
}

// ---- optional per-launch timing (bench.py's roofline leg): cudaEvent pairs around every tap-GEMM launch
struct ProfRec {
  cudaEvent_t e0, e1;
  double flops;
  int mode;
  int M, N, K, taps, block_n, grid;
  float ms;
};
static ProfRec* g_prof = nullptr;
static int g_prof_cap = 0, g_prof_n = 0, g_prof_last = 0;
static bool g_prof_on = false;

static ProfRec* prof_begin(cudaStream_t st, int mode, double flops) {
  if (!g_prof_on || g_prof_n >= g_prof_cap) return nullptr;
  ProfRec* r = &g_prof[g_prof_n++];
  r->mode = mode;
  r->flops = flops;
  cudaEventRecord(r->e0, st);
  return r;
}
static void prof_end(ProfRec* r, cudaStream_t st) {
  if (r) cudaEventRecord(r->e1, st);
}

static unsigned long long* g_probe = nullptr;
static int g_debug_sync = -1;
static int debug_sync(cudaStream_t st) {
  if (g_debug_sync < 0) {
    const char* e = getenv("ASF_DEBUG_SYNC");
    g_debug_sync = (e && e[0] == '1') ? 1 : 0;
  }
  if (g_debug_sync) ASF_CUDA(cudaStreamSynchronize(st));
  return 0;
}

#include "narrow_wgrad_mma.cuh"
#include "narrow_conv.cuh"

// wgrad tiling: 64-row boxes (half of the 128-row M tile, split on its slowest boxed dim), output-channel tile of one
// or two 128-row halves, input-channel tile <= 256, split-K so that about one wave of CTAs is in flight.
struct WgPlan {
  int box[3];
  int halves, tn, ny, nx, r_tiles, c_tiles, tmem_cols, stages, ksplit, m_tiles;
  int narrow, narrow_rows;  // warp-MMA path for the narrow layers (narrow_wgrad_mma.cuh): ksplit = its CTA count
};
static int wgrad_plan(const asf_tapgemm_t* q, WgPlan* w) {
  w->narrow = 0;
  {
    int ks = 0, rpc = 0;  // a caller-chosen ksplit other than the plan's keeps the layer on the tensor-core kernel
    if (narrow_wgrad_mma_ok(q)) {  // warp-MMA kernel: every narrow layer (N <= 64, taps * K <= 256)
      narrow_wgrad_rows(q, &ks, &rpc);
      if (q->ksplit <= 0 || q->ksplit == ks) {
        w->narrow = 2;
        w->narrow_rows = rpc;
        w->ksplit = ks;
      }
    }
  }
  int prod = 1;
  for (int j = 0; j < 3; ++j) {
    w->box[j] = q->m_box[j];
    prod *= q->m_box[j];
  }
  ASF_CHECK_ARG(prod == TG_BLOCK_M, "asf_tapgemm: product of m_box must be %d (got %d)", TG_BLOCK_M, prod);
  for (int j = 2; j >= 0; --j)
    if (w->box[j] > 1) {
      w->box[j] /= 2;
      break;
    }
  w->m_tiles = 1;
  for (int j = 0; j < 3; ++j) {
    ASF_CHECK_ARG(ilog2_exact(w->box[j]) >= 0 && q->m_ext[j] >= 1, "asf_tapgemm: m_box must be powers of two");
    w->m_tiles *= (int)cdiv(q->m_ext[j], w->box[j]);
  }
  w->halves = q->N > 128 ? 2 : 1;
  const int TM = w->halves * 128;
  w->r_tiles = (int)cdiv(q->N, TM);
  const int k16 = (int)cdiv(q->K, 16) * 16;
  w->tn = k16 < 256 ? k16 : 256;
  w->c_tiles = (int)cdiv(q->K, w->tn);
  w->ny = (int)cdiv(q->N < TM ? q->N : TM, 64);
  w->nx = (int)cdiv(w->tn, 64);
  const int need = (w->halves - 1) * w->tn + (int)cdiv(w->tn, 32) * 32;
  int cols = 32;
  while (cols < need) cols *= 2;
  w->tmem_cols = cols;
  const int stage_bytes = (w->ny + w->nx) * WG2_BOX_BYTES;
  int stages = (227 * 1024 - (int)sizeof(TgShared)) / stage_bytes;
  if (stages > TG_MAX_STAGES) stages = TG_MAX_STAGES;
  {
    static int cap = -1;
    if (cap < 0) {
      const char* e = getenv("ASF_WG_STAGES");
      cap = e ? atoi(e) : 0;
    }
    if (cap >= 2 && stages > cap) stages = cap;
  }
  ASF_CHECK_ARG(stages >= 2, "asf_tapgemm: wgrad stage does not fit twice in shared memory");
  w->stages = stages;
  const int64_t base_tiles = (int64_t)w->r_tiles * w->c_tiles * q->ntaps;
  // CTAs a weight-gradient launch aims for (ASF_WG_CTAS, 0 = every SM).  tcgen05 CTAs of two kernels never share an SM
  // (each allocates its TMEM columns and most of the shared memory), so a wgrad grid over all 148 SMs makes the dgrad
  // chain beside it on the pathway stream queue behind it.  A quarter of the chip per wgrad launch -- four of them run
  // on the four weight-gradient streams -- leaves the rest to the critical chain and cuts the split-K partial traffic
  // 4x: measured 5.95 -> 5.35 ms per step (148: 5.95, 96: 5.60, 64: 5.47, 48: 5.36, 32: 5.35, 24: 5.42, 16: 5.47).
  static int wg_ctas = -1;
  if (wg_ctas < 0) {
    const char* e = getenv("ASF_WG_CTAS");
    wg_ctas = e ? atoi(e) : 32;
  }
  int budget = sm_budget(q);
  if (wg_ctas > 0 && wg_ctas < budget) budget = wg_ctas;
  int ks = (int)cdiv(budget, base_tiles);
  if (ks > w->m_tiles / 4) ks = w->m_tiles / 4;  // at least four 64-row stages of MMA work per partial tile
  if (ks < 1) ks = 1;
  if (!w->narrow) w->ksplit = ks;
  return 0;
}

}  // namespace asf

using namespace asf;

extern "C" int asf_tapgemm_probe(unsigned long long* dev_counters) {
  g_probe = dev_counters;
  return 0;
}

extern "C" long long asf_narrow_conv_launches(void) { return g_narrow_conv_launches; }

extern "C" int asf_tapgemm(const asf_tapgemm_t* q, int mode, void* stream) {
  ASF_CHECK_ARG(q && (mode == 0 || mode == 1), "asf_tapgemm: bad mode");
  ASF_CHECK_ARG(q->a && q->nd >= 1 && q->nd <= 4 && q->K > 0 && q->K % 8 == 0 && q->N > 0,
                "asf_tapgemm: bad A description (nd %d, K %d, N %d)", q->nd, q->K, q->N);
  ASF_CHECK_ARG(q->ntaps >= 1 && q->ntaps <= ASF_MAX_TAPS && q->ntaps_w >= 1, "asf_tapgemm: bad tap count");
  if (mode == 0 && q->bw && q->out && q->ld_out >= q->N && narrow_conv_ok(q))
    return narrow_conv_launch(q, (cudaStream_t)stream);
  static TgParams p;  // large POD; filled per call (the launch copies it)
  memset(&p, 0, sizeof(p));
  p.dbg = g_probe;
  int box_a[4] = {1, 1, 1, 1}, box_y[4] = {1, 1, 1, 1};
  int prod = 1, last = 0;
  p.m_tiles_total = 1;
  for (int j = 0; j < 3; ++j) {
    int l = ilog2_exact(q->m_box[j]);
    ASF_CHECK_ARG(l >= 0 && q->m_ext[j] >= 1, "asf_tapgemm: m_box[%d]=%d must be a power of two", j, q->m_box[j]);
    p.m_ext[j] = q->m_ext[j];
    p.m_log2[j] = l;
    p.m_to_a[j] = q->m_to_a[j];
    p.m_tiles[j] = (int)cdiv(q->m_ext[j], q->m_box[j]);
    p.m_tiles_total *= p.m_tiles[j];
    prod *= q->m_box[j];
    if (q->m_to_a[j] > 0) {
      ASF_CHECK_ARG(q->m_to_a[j] <= q->nd && q->m_to_a[j] > last, "asf_tapgemm: m_to_a must be increasing and <= nd");
      last = q->m_to_a[j];
      box_a[q->m_to_a[j] - 1] = q->m_box[j];
    } else {
      ASF_CHECK_ARG(q->m_box[j] == 1 && q->m_ext[j] == 1, "asf_tapgemm: unmapped M dim %d must have extent/box 1", j);
    }
  }
  ASF_CHECK_ARG(prod == TG_BLOCK_M, "asf_tapgemm: product of m_box must be %d (got %d)", TG_BLOCK_M, prod);
  p.ntaps = q->ntaps;
  for (int t = 0; t < q->ntaps; ++t) {
    for (int d = 0; d < 4; ++d) p.a_off[t][d] = q->a_off[t][d];
    p.b_tap[t] = q->b_tap[t];
    ASF_CHECK_ARG(q->b_tap[t] >= 0 && q->b_tap[t] < q->ntaps_w, "asf_tapgemm: b_tap out of range");
  }
  p.K = q->K;
  p.N = q->N;
  p.kchunks = (int)cdiv(q->K, TG_KC);
  p.rank_a = q->nd + 1;
  p.ntaps_w = q->ntaps_w;
  int rc = encode_map(&p.tm_a, q->a, q->K, q->nd, q->a_dim, q->a_stride, box_a);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;

  if (mode == 0) {
    ASF_CHECK_ARG(q->bw && q->out && q->ld_out >= q->N, "asf_tapgemm: mode 0 needs bw/out");
    ASF_CHECK_ARG(q->out_dtype == ASF_BF16 || q->out_dtype == ASF_F32, "asf_tapgemm: bad out dtype");
    p.out_bf16 = q->out_dtype == ASF_BF16;
    bool vec = p.out_bf16 && (q->ld_out % 8 == 0) && (((uintptr_t)q->out & 15) == 0);
    if (q->epi.res) vec = vec && (q->epi.ld_res % 8 == 0) && (((uintptr_t)q->epi.res & 15) == 0);
    if (q->epi.mask) vec = vec && (q->epi.ld_mask % 8 == 0) && (((uintptr_t)q->epi.mask & 15) == 0);
    p.vec_ok = vec ? 1 : 0;
    // one kernel (TMA epilogue): bf16 output, 16-byte aligned rows, BN scale folded into the weights, and at most one
    // additive operand (residual OR accumulate).  Everything the network launches has this form; anything else is
    // refused (the CUDA-core kernels of conv_simt.cu take arbitrary layouts).
    if (!vec || q->epi.scale || (q->epi.res && q->epi.accumulate)) {
      set_error("asf_tapgemm: mode 0 needs a bf16 output with 16-byte aligned rows (out / res / mask), no per-channel "
                "scale and not residual + accumulate together");
      return ASF_E_UNSUPPORTED;
    }
    const bool use_tma = true;
    const int n_add = (q->epi.res || q->epi.accumulate) ? 1 : 0, has_mask = q->epi.mask ? 1 : 0;
    // Operand slabs (residual / mask) arrive by TMA ahead of the math through a ring of op_depth slabs per operand and
    // epilogue group.  Measured on the training step (ASF_OP_DEPTH): 2 -> 5.21 ms, 3 -> 5.22 ms, 4 -> 5.75 ms -- every
    // extra slab of lookahead costs 16..32 KB of shared memory, i.e. a mainloop stage or half the N tile, which loses
    // more than the operand latency it hides.  Default 2.
    static int op_depth_env = -1;
    if (op_depth_env < 0) {
      const char* e = getenv("ASF_OP_DEPTH");
      op_depth_env = e ? atoi(e) : 0;
    }
    int op_depth = 2;
    if (op_depth_env >= 2 && op_depth_env <= 4) op_depth = op_depth_env;
    const int group_bytes = (2 + op_depth * (n_add + has_mask)) * EPI_SLAB_BYTES;
    const int stat_bytes = q->epi.stats ? EPI_STAT_FLOATS * 4 : 0;
    const int tma_fixed = 2 * group_bytes + 2 * EPI_SHIFT_FLOATS * 4 + stat_bytes + (int)sizeof(TgShared);
    const int smem_cap = 227 * 1024;
    int bn = q->block_n;
    if (bn <= 0) {
      int n16 = (int)cdiv(q->N, 16) * 16;
      bn = n16 <= 256 ? n16 : 256;
      if (n16 > 256 && n16 % 256 != 0 && n16 % 128 == 0) bn = 128;
      if (use_tma) {
        // N = 256 per MMA keeps the single issuing thread ahead of the tensor pipe; halve only when that at least
        // doubles the number of busy SMs, or when the operand slabs leave fewer than 3 ring stages
        while (bn > 64 && bn % 32 == 0 && (int64_t)p.m_tiles_total * cdiv(q->N, bn) * 2 <= sm_budget(q)) bn /= 2;
        while (bn > 64 && bn % 32 == 0 && (smem_cap - tma_fixed) / (TG_A_BYTES + bn * 128) < 3) bn /= 2;
        {
          // Wave quantisation (ASF_TG_WAVE = fixed cost of a tile in units of output columns, 0 = off): tiles are dealt
          // round robin to the persistent CTAs, so a launch takes ceil(tiles / CTAs) rounds of (fixed + block_n); e.g.
          // 200 tiles of 256 columns on 148 SMs are 2 rounds with the second a third full, 400 tiles of 128 are 3
          // rounds of half the length.
          static int wave = -1;
          if (wave < 0) {
            const char* e = getenv("ASF_TG_WAVE");
            wave = e ? atoi(e) : 0;
          }
          if (wave > 0 && !q->epi.stats) {
            int best = bn;
            double best_cost = 1e30;
            for (int c = bn; c >= 64 && c % 16 == 0; c /= 2) {
              if ((smem_cap - tma_fixed) / (TG_A_BYTES + c * 128) < 3) continue;
              const int64_t tiles = (int64_t)p.m_tiles_total * cdiv(q->N, c);
              const double cost = (double)cdiv(tiles, sm_budget(q)) * (wave + c);
              if (cost < best_cost * 0.97) {  // prefer the wider tile unless the narrower one is clearly ahead
                best_cost = cost;
                best = c;
              }
              if (c % 32) break;
            }
            bn = best;
          }
        }
      }
    }
    ASF_CHECK_ARG(bn >= 16 && bn <= 256 && bn % 16 == 0, "asf_tapgemm: block_n %d must be a multiple of 16 in [16,256]", bn);
    p.block_n = bn;
    p.n_tiles = (int)cdiv(q->N, bn);
    // weights [N][taps_w][K] viewed as a rank-3 map (K, taps_w, N), box (64, 1, block_n)
    int64_t wext[2] = {q->ntaps_w, q->N}, wstr[2] = {q->K, (int64_t)q->ntaps_w * q->K};
    int wbox[2] = {1, bn};
    rc = encode_map(&p.tm_b, q->bw, q->K, 2, wext, wstr, wbox);
    if (rc) return rc;
    p.rank_b = 3;
    const int stage_bytes = TG_A_BYTES + bn * 128;
    int stages = 0;
    p.out = q->out;
    p.ld_out = q->ld_out;
    for (int j = 0; j < 3; ++j) p.o_stride[j] = q->o_stride[j];
    p.epi = q->epi;
    int64_t total = (int64_t)p.m_tiles_total * p.n_tiles;
    int grid = (int)(total < sm_budget(q) ? total : sm_budget(q));
    if (q->epi.stats) {
      // statistics: a CTA keeps one N tile for all its M tiles (grid a multiple of n_tiles), sums its columns in shared
      // memory and leaves one workspace row; see tapgemm_fwd_tma.cuh
      ASF_CHECK_ARG(q->epi.stats_ws && q->epi.stats_ticket && p.n_tiles <= ASF_STATS_TICKETS && grid >= p.n_tiles,
                    "asf_tapgemm: statistics need stats_ws / stats_ticket (n_tiles %d)", p.n_tiles);
      grid = grid / p.n_tiles * p.n_tiles;
    }
    const double flops = 2.0 * (double)q->m_ext[0] * q->m_ext[1] * q->m_ext[2] * q->N * q->K * q->ntaps;
    if (use_tma) {
      stages = (smem_cap - tma_fixed) / stage_bytes;
      if (stages > TG_MAX_STAGES) stages = TG_MAX_STAGES;
      {
        // ring depth cap (ASF_TG_STAGES): shared memory a tcgen05 CTA does not take is what lets the CUDA-core /
        // warp-MMA CTAs of the narrow layers run on the same SM beside it
        static int cap = -1;
        if (cap < 0) {
          const char* e = getenv("ASF_TG_STAGES");
          cap = e ? atoi(e) : 0;
        }
        if (cap >= 2 && stages > cap) stages = cap;
      }
      ASF_CHECK_ARG(stages >= 2, "asf_tapgemm: not enough shared memory for 2 stages");
      p.stages = stages;
      p.n_add = n_add;
      p.has_mask = has_mask;
      p.epi_smem_off = stages * stage_bytes;
      p.op_depth = op_depth;
      {
        static int split = -1;
        if (split < 0) {
          const char* e = getenv("ASF_EPI_SPLIT");
          split = (e && e[0] == '0') ? 0 : 1;
        }
        p.epi_split = split;
      }
      // output-space tensor maps: dim 0 = channels, then every M dim that moves the output row
      int64_t oext[3], ostr_rows[3];
      int obox[3], nd_o = 0;
      for (int j = 0; j < 3; ++j) {
        p.o_dim[j] = 0;
        if (q->m_ext[j] > 1 || q->m_box[j] > 1) {
          ASF_CHECK_ARG(q->o_stride[j] > 0, "asf_tapgemm: M dim %d has extent > 1 but no output stride", j);
          oext[nd_o] = q->m_ext[j];
          ostr_rows[nd_o] = q->o_stride[j];
          obox[nd_o] = q->m_box[j];
          p.o_dim[j] = ++nd_o;
        }
      }
      ASF_CHECK_ARG(nd_o >= 1, "asf_tapgemm: empty output space");
      p.rank_o = nd_o + 1;
      int64_t ostr[3];
      for (int i = 0; i < nd_o; ++i) ostr[i] = ostr_rows[i] * q->ld_out;
      rc = encode_map(&p.tm_out, q->out, q->N, nd_o, oext, ostr, obox, EPI_SLAB_COLS, CU_TENSOR_MAP_SWIZZLE_64B);
      if (rc) return rc;
      const size_t smem_t = (size_t)stages * stage_bytes + tma_fixed;
      static size_t smem_set_t[2] = {0, 0};
      const int si = q->epi.stats ? 1 : 0;
      if (smem_t > smem_set_t[si]) {
        if (si)
          ASF_CUDA(cudaFuncSetAttribute(tapgemm_fwd_tma_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_cap));
        else
          ASF_CUDA(cudaFuncSetAttribute(tapgemm_fwd_tma_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_cap));
        smem_set_t[si] = smem_cap;
      }
      ProfRec* pr = prof_begin(st, 0, flops);
      if (pr) {
        pr->M = q->m_ext[0] * q->m_ext[1] * q->m_ext[2]; pr->N = q->N; pr->K = q->K; pr->taps = q->ntaps;
        pr->block_n = bn; pr->grid = grid;
      }
      if (si)
        ASF_CUDA(launch_pdl(tapgemm_fwd_tma_kernel<true>, dim3(grid), dim3(TGT_THREADS), smem_t, st, p));
      else
        ASF_CUDA(launch_pdl(tapgemm_fwd_tma_kernel<false>, dim3(grid), dim3(TGT_THREADS), smem_t, st, p));
      prof_end(pr, st);
      ASF_LAUNCH_CHECK();
      return debug_sync(st);
    }
    return ASF_E_UNSUPPORTED;
  }

  // mode 1: wgrad (second generation: 64-row stages, up to 256 x 256 per CTA, split-K partials by plain stores)
  ASF_CHECK_ARG(q->y && q->dwp && q->y_nd >= 1 && q->y_nd <= 4, "asf_tapgemm: mode 1 needs y/dwp");
  WgPlan wp;
  rc = wgrad_plan(q, &wp);
  if (rc) return rc;
  if (wp.narrow) {
    const double nflops = 2.0 * (double)q->m_ext[0] * q->m_ext[1] * q->m_ext[2] * q->N * q->K * q->ntaps;
    ProfRec* npr = prof_begin(st, 1, nflops);
    if (npr) {
      npr->M = q->m_ext[0] * q->m_ext[1] * q->m_ext[2]; npr->N = q->N; npr->K = q->K; npr->taps = q->ntaps;
      npr->block_n = wp.ksplit; npr->grid = wp.ksplit;
    }
    rc = narrow_wgrad_mma_launch(q, wp.ksplit, wp.narrow_rows, st);
    if (rc) return rc;
    prof_end(npr, st);
    ASF_LAUNCH_CHECK();
    return debug_sync(st);
  }
  // re-tile the M-space with the 64-row boxes of the plan
  p.m_tiles_total = 1;
  for (int j = 0; j < 3; ++j) {
    p.m_log2[j] = ilog2_exact(wp.box[j]);
    p.m_tiles[j] = (int)cdiv(q->m_ext[j], wp.box[j]);
    p.m_tiles_total *= p.m_tiles[j];
  }
  for (int j = 0; j < 4; ++j) box_a[j] = 1;
  for (int j = 0; j < 3; ++j)
    if (q->m_to_a[j] > 0) box_a[q->m_to_a[j] - 1] = wp.box[j];
  rc = encode_map(&p.tm_a, q->a, q->K, q->nd, q->a_dim, q->a_stride, box_a);
  if (rc) return rc;
  last = 0;
  for (int j = 0; j < 3; ++j) {
    p.y_m_to_a[j] = q->y_m_to_a[j];
    if (q->y_m_to_a[j] > 0) {
      ASF_CHECK_ARG(q->y_m_to_a[j] <= q->y_nd && q->y_m_to_a[j] > last, "asf_tapgemm: y_m_to_a must be increasing");
      last = q->y_m_to_a[j];
      box_y[q->y_m_to_a[j] - 1] = wp.box[j];
    } else {
      ASF_CHECK_ARG(q->m_box[j] == 1, "asf_tapgemm: unmapped Y M-dim must have box 1");
    }
  }
  rc = encode_map(&p.tm_b, q->y, q->N, q->y_nd, q->y_dim, q->y_stride, box_y);
  if (rc) return rc;
  p.rank_b = q->y_nd + 1;
  p.dwp = q->dwp;
  p.c_tiles = wp.c_tiles;
  p.r_tiles = wp.r_tiles;
  p.wg_halves = wp.halves;
  p.wg_tn = wp.tn;
  p.wg_ny = wp.ny;
  p.wg_nx = wp.nx;
  p.wg_tmem_cols = wp.tmem_cols;
  p.stages = wp.stages;
  int ks = q->ksplit > 0 ? q->ksplit : wp.ksplit;
  ASF_CHECK_ARG(ks >= 1 && ks <= p.m_tiles_total, "asf_tapgemm: ksplit %d out of range (row tiles %d)", ks, p.m_tiles_total);
  p.ksplit = ks;
  const size_t smem = (size_t)wp.stages * (wp.ny + wp.nx) * WG2_BOX_BYTES + sizeof(TgShared);
  static size_t smem_set_w = 0;
  if (smem > smem_set_w) {
    ASF_CUDA(cudaFuncSetAttribute(tapgemm_wgrad2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    smem_set_w = 227 * 1024;
  }
  int64_t grid = (int64_t)wp.r_tiles * wp.c_tiles * p.ntaps * ks;
  ASF_CHECK_ARG(grid < (1ll << 31), "asf_tapgemm: wgrad grid too large");
  const double flops = 2.0 * (double)q->m_ext[0] * q->m_ext[1] * q->m_ext[2] * q->N * q->K * q->ntaps;
  ProfRec* pr = prof_begin(st, 1, flops);
  if (pr) {
    pr->M = q->m_ext[0] * q->m_ext[1] * q->m_ext[2]; pr->N = q->N; pr->K = q->K; pr->taps = q->ntaps;
    pr->block_n = ks; pr->grid = (int)grid;
  }
  ASF_CUDA(launch_pdl(tapgemm_wgrad2_kernel, dim3((unsigned)grid), dim3(WG2_THREADS), smem, st, p));
  prof_end(pr, st);
  ASF_LAUNCH_CHECK();
  return debug_sync(st);
}

extern "C" int asf_tapgemm_wgrad_ksplit(const asf_tapgemm_t* q) {
  ASF_CHECK_ARG(q != nullptr, "asf_tapgemm_wgrad_ksplit: null");
  WgPlan wp;
  int rc = wgrad_plan(q, &wp);
  if (rc) return rc;
  return wp.ksplit;
}

extern "C" int asf_prof_start(int max_records) {
  ASF_CHECK_ARG(max_records > 0, "asf_prof_start: max_records must be positive");
  if (g_prof_cap < max_records) {
    ProfRec* n = (ProfRec*)realloc(g_prof, sizeof(ProfRec) * (size_t)max_records);
    ASF_CHECK_ARG(n != nullptr, "asf_prof_start: out of host memory");
    g_prof = n;
    for (int i = g_prof_cap; i < max_records; ++i) {
      ASF_CUDA(cudaEventCreate(&g_prof[i].e0));
      ASF_CUDA(cudaEventCreate(&g_prof[i].e1));
    }
    g_prof_cap = max_records;
  }
  g_prof_n = 0;
  g_prof_on = true;
  return 0;
}

extern "C" int asf_prof_stop(double* ms, double* flops, long long* launches) {
  g_prof_on = false;
  for (int m = 0; m < 2; ++m) {
    ms[m] = 0.0;
    flops[m] = 0.0;
    launches[m] = 0;
  }
  for (int i = 0; i < g_prof_n; ++i) {
    ASF_CUDA(cudaEventSynchronize(g_prof[i].e1));
    float t = 0.f;
    ASF_CUDA(cudaEventElapsedTime(&t, g_prof[i].e0, g_prof[i].e1));
    const int m = g_prof[i].mode ? 1 : 0;
    g_prof[i].ms = t;
    ms[m] += t;
    flops[m] += g_prof[i].flops;
    launches[m] += 1;
  }
  g_prof_last = g_prof_n;
  g_prof_n = 0;
  return 0;
}

extern "C" int asf_prof_record(int i, double* out) {
  ASF_CHECK_ARG(i >= 0 && i < g_prof_last && out, "asf_prof_record: index %d out of range (%d records)", i, g_prof_last);
  const ProfRec& r = g_prof[i];
  out[0] = r.mode; out[1] = r.ms; out[2] = r.flops; out[3] = r.M; out[4] = r.N; out[5] = r.K; out[6] = r.taps;
  out[7] = r.block_n; out[8] = r.grid;
  return 0;
}
