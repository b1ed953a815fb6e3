/* asf_b200.h -- C ABI of libasf_b200.so: the B200 (sm_100a) kernels behind the Auditory SlowFast hot path.
 *
 * The reference (ClementSicard/auditory-slow-fast) has NO native / FFI interface of its own: its hot
 * path is Python calling librosa/numpy (CPU) and torch.nn (cuDNN/cuBLAS).  The drop-in boundary is
 * therefore the reference's Python call signatures (mirrored 1:1 by auditory-slow-fast_b200/asf_b200),
 * and this header is the C ABI *below* that mirror: one entry point per library kernel the reference
 * dispatches to today.  Each declaration cites the reference call site it replaces
 * (paths relative to /root/reference).
 *
 * Conventions
 *  - plain pointers and sizes only; no torch types.  Every pointer is a DEVICE pointer unless named h_*.
 *  - the caller owns every buffer (including workspaces); kernels never allocate or free device memory.
 *  - every launch is asynchronous on `stream` (a cudaStream_t passed as void*; NULL = legacy default).
 *  - return value: 0 on success, otherwise a cudaError_t / negative ASF_E_* code; asf_last_error()
 *    gives a thread-local message.  There is no CPU fallback anywhere.
 *  - activations are channels-last "rows x channels": row = (b*T + t)*F + f, channel stride 1,
 *    row pitch `ld` elements (>= C, so a tensor can be a channel slice of a wider buffer).
 *  - dtype codes: ASF_F32 = 0, ASF_BF16 = 1.
 */
#ifndef ASF_B200_H_
#define ASF_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ASF_F32 0
#define ASF_BF16 1

#define ASF_E_BADARG (-1)
#define ASF_E_UNSUPPORTED (-2)
#define ASF_E_DRIVER (-3)

#define ASF_MAX_TAPS 8

const char* asf_last_error(void);
int asf_version(void);
/* number of kernel launches the library has issued in this process (every launch, every entry point) */
long long asf_launch_count(void);
/* Per-launch timing of the tap-GEMM kernels (measurement aid for bench.py's roofline leg; not used by the
 * product path).  asf_prof_start arms up to max_records cudaEvent pairs; every asf_tapgemm launch until
 * asf_prof_stop is bracketed by one pair on its stream.  asf_prof_stop synchronises and returns, for
 * mode 0 (forward/dgrad) in [0] and mode 1 (wgrad) in [1]: total kernel milliseconds, total algorithmic
 * FLOPs (2 * M * N * K * taps of each launch) and the number of launches. */
int asf_prof_start(int max_records);
int asf_prof_stop(double* ms, double* flops, long long* launches);
/* record i of the last asf_prof_stop: out[9] = {mode, ms, flops, M, N, K, taps, block_n (mode 1: ksplit), grid} */
int asf_prof_record(int i, double* out);
/* 1 if the current device is sm_100 (tcgen05/TMA kernels can run), 0 otherwise, <0 on error. */
int asf_device_is_sm100(void);

/* ------------------------------------------------------------------------------------------------
 * Feature extraction.  Replaces librosa.stft + filters.mel + np.dot + np.log in
 *   audio_slowfast/datasets/audio_loader_epic.py:66-97  (_log_specgram; VGG twin audio_loader_vggsound.py:57-69)
 * the edge padding of _extract_sound_feature (audio_loader_epic.py:151-154) and the slow-pathway
 * gather of pack_pathway_output (audio_slowfast/datasets/utils.py:59-63), batched.
 * ---------------------------------------------------------------------------------------------- */

/* Builds the constant tables once per configuration, on the host in float64 rounded once to float32,
 * and uploads them (device buffers owned by the caller; synchronises `stream`):
 *   basisT    float [win][n_bins_pad][2]  (w[j] cos(2 pi k (lpad+j)/n_fft), -w[j] sin(..)), w = periodic Hann,
 *                                          lpad = (n_fft - win)/2  (librosa util.pad_center of the window)
 *   melT      float [n_bins_pad][n_mels]  librosa.filters.mel(sr, n_fft, n_mels, htk=True, norm=None) transposed
 *   mel_range int32 [n_mels][2]           [first, last+1) non-zero bin of every filter
 * n_bins = n_fft/2 + 1 <= n_bins_pad. */
int asf_logmel_tables(float* basisT, float* melT, int32_t* mel_range, int sr, int n_fft, int win, int n_mels,
                      int n_bins_pad, void* stream);

/* wave [B][n_samples] fp32 (row pitch wave_ld) -> fast [B][1][T_out][n_mels] fp32 where the
 * T = 1 + n_samples/hop frames of librosa.stft(center=True, pad_mode="constant") are computed and rows
 * T..T_out-1 replicate frame T-1 (np.pad "edge", audio_loader_epic.py:151-154).  T_out <= 0 means T.
 * T_out < T is an error (the reference never truncates). */
int asf_logmel_fwd(const float* wave, int64_t wave_ld, int B, int n_samples, int n_fft, int win, int hop,
                   int n_mels, float eps, const float* basisT, const float* melT, const int32_t* mel_range,
                   int n_bins_pad, float* fast, int T_out, void* stream);

/* Tensor-core (tcgen05 / TMEM, 3-term TF32 split = fp32-grade accuracy) version of asf_logmel_fwd for the shipped
 * configuration win = 240, hop = 120 (= win / 2), n_fft <= 2048: same inputs, same output, same padding rules.
 *   asf_logmel_tc_supported   1 when (n_fft, win, hop) can run on this path
 *   asf_logmel_tc_btab_floats number of floats of the pre-split basis image for n_fft
 *   asf_logmel_tc_tables      fills btab [asf_logmel_tc_btab_floats], melw [nbp][2] (descending / ascending weight of
 *                             the two mel filters a bin touches) and melj [nbp] (index of the ascending filter), with
 *                             nbp = ceil((n_fft/2+1) / 104) * 104; fails if a bin touches non-adjacent filters
 *   asf_logmel_fwd_tc         wave [B][wave_ld] fp32 -> fast [B][T_out][n_mels] fp32 */
int asf_logmel_tc_supported(int n_fft, int win, int hop);
long long asf_logmel_tc_btab_floats(int n_fft);
int asf_logmel_tc_tables(float* btab, float* melw, int32_t* melj, int sr, int n_fft, int win, int hop, int n_mels,
                         void* stream);
int asf_logmel_fwd_tc(const float* wave, int64_t wave_ld, int B, int n_samples, int n_fft, int win, int hop,
                      int n_mels, float eps, const float* btab, const float* melw, const int32_t* melj, float* fast,
                      int T_out, void* stream);

/* SpecAugment, batched on the device.  Replaces datasets/spec_augment.py:144-191 (time_warp through
 * datasets/sparse_image_warp.py:5-354, 2 frequency masks of width < Fmax, 2 time masks of width < Tmax, each filled with
 * the current mean of the clip) as called per training sample at datasets/epickitchens.py:160-168.
 *   asf_specaug_params  draws, per clip, the reference's random quantities with a counter-based generator (Philox4x32-10
 *                       keyed by `seed`, counter = (clip, draw, *counter); *counter, a DEVICE uint64 that may be NULL, is
 *                       advanced by one per launch so that a replayed CUDA graph keeps drawing fresh numbers) and
 *                       reduces them to the affine time flow  flow(f, t) = coef[0] f + coef[1] t + coef[2]  that the
 *                       reference's one-control-point spline amounts to, plus the half-open mask intervals.
 *                       spec is read at (t = i, f = F/2): the reference uses that VALUE as a control coordinate.
 *   asf_specaug_apply   out[b][t][f] = augmented in[b][t][f]  (fp32 [B][T][F], in != out): time resampling with linear
 *                       interpolation clamped at the borders, then the fills.  Bit-reproducible for given params. */
#define ASF_SPECAUG_MAX_MASKS 2
typedef struct {
  float coef[3];
  int32_t fmask[ASF_SPECAUG_MAX_MASKS][2]; /* [start, end) along frequency; end <= start: nothing */
  int32_t tmask[ASF_SPECAUG_MAX_MASKS][2]; /* [start, end) along time */
  int32_t pad_;
} asf_specaug_params_t;
int asf_specaug_params(const float* spec, int B, int T, int F, unsigned long long seed, unsigned long long* counter,
                       int W, int Fmax, int Tmax, int num_freq_masks, int num_time_masks, asf_specaug_params_t* params,
                       void* stream);
int asf_specaug_apply(const float* in, float* out, const asf_specaug_params_t* params, int B, int T, int F,
                      void* stream);

/* ------------------------------------------------------------------------------------------------
 * Convolution family.  Replaces every nn.Conv2d (+ the BatchNorm2d / ReLU / residual add / torch.cat
 * that follow it) of
 *   audio_slowfast/models/stem_helper.py:150-168, resnet_helper.py:170-225 and :310-341,
 *   audio_model_builder.py:90-111 (FuseFastToSlow)
 * and their autograd backward (cuDNN dgrad / wgrad).
 * ---------------------------------------------------------------------------------------------- */

/* Fused epilogue applied to the fp32 accumulator `acc` of output element (row, n):
 *   v = acc * (scale ? scale[n] : 1) + (shift ? shift[n] : 0)
 *   if (res)        v += res[row*ld_res + n]
 *   if (accumulate) v += out[row*ld_out + n]
 *   if (relu)       v = max(v, 0)
 *   if (mask)       v = mask[row*ld_mask + n] > 0 ? v : 0
 *   if (stats)      stats[n] = sum_rows v ; stats[stats_stride + n] = sum_rows v*v   (valid rows only; see below)
 *   out[row*ld_out + n] = (out dtype) v
 * res / mask have the out dtype.
 *
 * Statistics are DETERMINISTIC: no floating-point atomics anywhere.  Every CTA stores its partial sums into its own
 * row of `stats_ws`, the CTA that draws the last ticket adds the rows in ascending order in double precision and
 * WRITES (not accumulates) the final sums as doubles.  stats_ws: >= ASF_STATS_WS_FLOATS(channels) floats, contents
 * irrelevant; stats_ticket: ASF_STATS_TICKETS uint32 counters, zero before the first launch (they reset themselves);
 * launches that share a workspace / ticket must be stream-ordered. */
typedef struct {
  const float* scale;
  const float* shift;
  const void* res;
  int64_t ld_res;
  const void* mask;
  int64_t ld_mask;
  double* stats;
  int32_t stats_stride; /* offset of the sum-of-squares half (= channels of the stats buffer) */
  int32_t relu;
  int32_t accumulate;
  float* stats_ws;
  uint32_t* stats_ticket;
} asf_epilogue_t;
#define ASF_STATS_TICKETS 128
#define ASF_STATS_WS_FLOATS(C) (640ll * ((C) > 256 ? (C) : 256))

/* Dense direct convolution geometry, channels-last. */
typedef struct {
  int32_t B, Tin, Fin, Cin;
  int32_t Tout, Fout, Cout;
  int32_t kt, kf, st, sf, pt, pf;
  int64_t ld_in, ld_out;
} asf_conv_geom_t;

/* CUDA-core reference/stem path (any geometry incl. Cin = 1; dtype codes per tensor).
 *   w   : packed weights [Cout][kt*kf][Cin] in dtype w_dtype
 *   fwd : out[b,to,fo,n] = epi( sum in[b,to*st+it-pt,fo*sf+if-pf,c] * w[n][it*kf+if][c] )
 *   dgrad: din[b,ti,fi,c] = epi( sum dout[b,to,fo,n] * w[n][tap][c] )  (epilogue indexed by (in-row, c))
 *   wgrad: dw[n][c][it][if] (OIHW fp32) = (accumulate ? dw : 0) + rscale[n] * sum dout[row,n] * in[row_tap,c]
 * fp32 inputs accumulate in double (the fp32 parity mode is an accuracy reference, not a fast path); the forward
 * epilogue does not produce statistics (epi->stats must be NULL: use asf_channel_stats). */
int asf_conv_fwd_simt(const void* in, int in_dtype, const void* w, int w_dtype, void* out, int out_dtype,
                      const asf_conv_geom_t* g, const asf_epilogue_t* epi, void* stream);
int asf_conv_dgrad_simt(const void* dout, int dout_dtype, const void* w, int w_dtype, void* din, int din_dtype,
                        const asf_conv_geom_t* g, const asf_epilogue_t* epi, void* stream);
/* wgrad: the rows are split over asf_conv_wgrad_simt_ksplit(g) slices whose partial sums are STORED to
 * ws [ksplit][Cout][taps][Cin] fp32 and added in slice order (no atomics). */
int asf_conv_wgrad_simt_ksplit(const asf_conv_geom_t* g);
int asf_conv_wgrad_simt(const void* in, int in_dtype, const void* dout, int dout_dtype, float* dw_oihw,
                        const asf_conv_geom_t* g, const float* rscale, int accumulate, float* ws, void* stream);

/* tcgen05 / TMEM / TMA implicit-GEMM ("tap-GEMM"), bf16 operands, fp32 accumulation in TMEM.
 *
 *   D[m][n] = sum_{tap < ntaps} sum_{c < K}  A[ coord(m) + a_off[tap] ][c] * Bw[n][ b_tap[tap] ][c]
 *
 * A is addressed through a rank-(1+nd) TMA tensor map over a bf16 tensor: dim 0 = channels (extent K,
 * contiguous), dims 1..nd described by (a_dim, a_stride) in elements.  The GEMM M-space is a 3-D grid
 * (m1 fastest) of extents m_ext tiled by boxes m_box (powers of two, product 128); M-dim j lives on tensor-map
 * dim m_to_a[j] (1..nd) or is absent (0: extent 1).  Coordinates outside the tensor are zero-filled by TMA,
 * which is how every conv padding is realised.  Bw is a bf16 tensor [N][ntaps_w][K] (K contiguous).
 * Output row of M-coordinate (m1,m2,m3) is  m1*o_stride[0] + m2*o_stride[1] + m3*o_stride[2]  (in rows).
 *
 * mode 0 (forward / dgrad): as above, out dtype out_dtype with the fused epilogue.
 * mode 1 (wgrad): D[n][c] = sum_rows Y[row][n] * A[row + a_off[tap]][c]; Y uses the same M-space through its
 *         own tensor map (y_*).  The rows are split over `ksplit` groups of CTAs; group k WRITES (no atomics, no
 *         pre-zeroing) its partial sums to dwp[k][N][ntaps_w][K] at tap b_tap[tap]; asf_unpack_wgrad adds the
 *         groups.  ksplit = 0 lets the launcher choose -- query it first with asf_tapgemm_wgrad_ksplit to size dwp.
 *
 * Narrow layers (8..64 output channels, >= 16384 output positions: the fast pathway) do not fill a 128-row tcgen05
 * tile; both modes route them to warp-level tensor-core kernels (mma.sync) behind this same entry point (mode 0:
 * csrc/narrow_conv.cuh when N, K are powers of two, K <= 128, N*K*ntaps <= 8192 and the output is bf16; mode 1:
 * csrc/narrow_wgrad_mma.cuh when N <= 64 and ntaps*K <= 256, ksplit = its CTA count).
 * asf_narrow_conv_launches() counts the mode-0 launches that took that route (tests use it to prove coverage).
 */
typedef struct {
  const void* a;      /* bf16 */
  int32_t nd;         /* number of non-channel tensor-map dims (1..4) */
  int64_t a_dim[4];   /* extents of tensor-map dims 1..nd */
  int64_t a_stride[4];/* strides of tensor-map dims 1..nd, elements (multiple of 8) */
  int32_t K;          /* channels (extent of dim 0) */
  int32_t m_ext[3], m_box[3], m_to_a[3];
  int32_t ntaps;
  int32_t a_off[ASF_MAX_TAPS][4]; /* per-tap coordinate offset on tensor-map dims 1..nd */
  int32_t b_tap[ASF_MAX_TAPS];    /* per-tap index into the tap dimension of Bw / dwp */
  const void* bw;     /* mode 0: bf16 [N][ntaps_w][K] */
  int32_t N, ntaps_w;
  /* mode 0 output */
  void* out;
  int32_t out_dtype;
  int64_t ld_out;
  int64_t o_stride[3];
  int32_t zero_invalid; /* rows outside m_ext are skipped (0) -- reserved */
  asf_epilogue_t epi;
  /* mode 1 */
  const void* y;      /* bf16 [rows][N] viewed through the same M-space */
  int64_t y_dim[4];
  int64_t y_stride[4];
  int32_t y_nd, y_m_to_a[3];
  float* dwp;         /* fp32 [ksplit][N][ntaps_w][K] workspace, written */
  int32_t ksplit;
  int32_t block_n;    /* 0 = choose */
  int32_t max_ctas;   /* 0 = all SMs; otherwise the launch sizes its persistent grid / split-K for this many SMs */
} asf_tapgemm_t;

int asf_tapgemm(const asf_tapgemm_t* p, int mode, void* stream);
/* Diagnostic: device array of >= 16 uint64 counters that later mode-0 launches add stall cycles to (summed over
 * CTAs): [0] kernel body, [1] producer waiting for a free stage, [2] MMA waiting for operands, [3] MMA waiting for
 * a free accumulator, [4] epilogue waiting for accumulators, [5] epilogue busy per tile, [6] epilogue waiting for
 * its store buffer, [7] final store drain, [8] CTAs.  NULL switches the probe off (default). */
int asf_tapgemm_probe(unsigned long long* dev_counters);
long long asf_narrow_conv_launches(void);
/* the ksplit asf_tapgemm(p, 1, ...) uses when p->ksplit == 0 (>= 1), or a negative error code */
int asf_tapgemm_wgrad_ksplit(const asf_tapgemm_t* p);

/* Stem weight gradient (Cin = 1, <= 35 taps): in fp32 [B][Tin][Fin], dout [rows][Cout] -> row-range partial sums
 * dwp [ksplit][Cout][taps] fp32 (plain stores, fixed summation order inside a CTA: bit-reproducible), to be summed by
 * asf_unpack_wgrad(dwp, dw, O = 1, I = Cout*taps, taps = 1, NULL, accumulate, ksplit) or as one entry of
 * asf_unpack_wgrad_multi.  asf_stem_wgrad_ksplit returns the number of partial rows the launch writes (>= 1; < 0 on
 * error) -- size dwp with it.  Replaces cuDNN wgrad of the stem convs (stem_helper.py:150-157). */
int asf_stem_wgrad_ksplit(const asf_conv_geom_t* g, int dout_dtype);
int asf_stem_wgrad(const float* in, const void* dout, int dout_dtype, float* dwp, const asf_conv_geom_t* g,
                   void* stream);
/* The same with the cross-block sum folded in (bf16 dout): blocks store partial rows into dwp (>= rows x Cout*kt*kf
 * floats, rows = asf_stem_wgrad_fused_rows(), 0 when the shape is not supported), the block drawing the last `ticket`
 * (one zero-initialised, self-resetting uint32) adds them in block order and writes dw[Cout][1][kt][kf]
 * (accumulate: dw += ...).  Input lines are staged in shared memory; this is the last kernel of the backward pass
 * (stem_helper.py:150-157 -> autograd's conv weight gradient), so no reduction launch follows it. */
int asf_stem_wgrad_fused_rows(const asf_conv_geom_t* g, int dout_dtype);
int asf_stem_wgrad_fused(const float* in, const void* dout, int dout_dtype, float* dwp, float* dw, int accumulate,
                         uint32_t* ticket, const asf_conv_geom_t* g, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Weight (re)packing between the reference's state_dict layout (OIHW fp32) and kernel layouts.
 * ---------------------------------------------------------------------------------------------- */
/* w_oihw [O][I][kt][kf] fp32 -> packed [O][kt*kf][I] (dtype), each output row scaled by rscale[o] if given */
int asf_pack_weight(const float* w_oihw, void* packed, int dtype, int O, int I, int taps, const float* rscale,
                    void* stream);
/* -> transposed packed [I][kt*kf][O] (dtype), scaled by rscale[o] */
int asf_pack_weight_t(const float* w_oihw, void* packed, int dtype, int O, int I, int taps, const float* rscale,
                      void* stream);
/* Every conv of the network in one launch.  descs_dev: DEVICE array of n descriptors sorted by chunk0, where unit i
 * owns the blocks [chunk0_i, chunk0_{i+1}) of 2048 destination elements each (chunks_i = ceil(O*I*taps / 2048));
 * total_chunks = sum of chunks_i.  packed_t may be NULL. */
typedef struct {
  const float* w;      /* OIHW fp32 master weights */
  void* packed;        /* [O][taps][I] */
  void* packed_t;      /* [I][taps][O] or NULL */
  const float* rscale; /* per-output-channel scale (folded BN) or NULL */
  int32_t O, I, taps, dtype;
  int64_t chunk0;
} asf_pack_desc_t;
#define ASF_PACK_CHUNK 2048
int asf_pack_weights_multi(const asf_pack_desc_t* descs_dev, int n, long long total_chunks, void* stream);
/* ------------------------------------------------------------------------------------------------
 * Optimiser step.  torch.optim.SGD(momentum, dampening, nesterov, weight_decay) as configured by
 * audio_slowfast/models/optimizer.py:55-62, for all parameter tensors in one launch:
 *   d = g + wd*p;  m = first_step ? d : momentum*m + (1-dampening)*d;  d = nesterov ? d + momentum*m : m;  p -= lr*d
 * (m == NULL: plain SGD).  lr_dev[group] is read on the device so a captured CUDA graph follows the per-iteration
 * schedule of utils/lr_policy.py.  descs_dev: DEVICE array sorted by chunk0; unit i owns ceil(n_i / 4096) blocks.
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
  float* p;
  const float* g;
  float* m;
  int64_t n;
  float wd;
  int32_t group;
  int64_t chunk0;
} asf_sgd_desc_t;
#define ASF_SGD_CHUNK 4096
int asf_sgd_multi(const asf_sgd_desc_t* descs_dev, int n, long long total_chunks, const float* lr_dev, float momentum,
                  float dampening, int nesterov, int first_step, void* stream);
/* dwp [ksplit][O][taps][I] fp32 (the split-K partial tiles of asf_tapgemm mode 1)
 *   -> dw_oihw [O][I][taps] = (accumulate ? dw : 0) + rscale[o] * sum_k dwp[k] */
int asf_unpack_wgrad(const float* dwp, float* dw_oihw, int O, int I, int taps, const float* rscale,
                     int accumulate, int ksplit, void* stream);
/* The same for many layers in one launch.  descs_dev: DEVICE array of n descriptors sorted by chunk0, where unit i owns
 * the blocks [chunk0_i, chunk0_{i+1}) of C_i destination elements each, C_i = ASF_UNPACK_CHUNK for ksplit < 32 and
 * ASF_UNPACK_CHUNK_DEEP from 32 on (chunks_i = ceil(O*I*taps / C_i)); I % 4 == 0, dw and dwp 16-byte aligned. */
typedef struct {
  const float* dwp;    /* [ksplit][O][taps][I] partial tiles */
  float* dw;           /* OIHW fp32 gradient */
  const float* rscale; /* per-output-channel scale or NULL */
  int32_t O, I, taps, ksplit, accumulate, pad_;
  int64_t chunk0;
} asf_unpack_desc_t;
#define ASF_UNPACK_CHUNK 1024
#define ASF_UNPACK_CHUNK_DEEP 256
int asf_unpack_wgrad_multi(const asf_unpack_desc_t* descs_dev, int n, long long total_chunks, void* stream);

/* ------------------------------------------------------------------------------------------------
 * BatchNorm2d (audio_slowfast/models/batchnorm_helper.py:16-31 -> nn.BatchNorm2d; train / eval / frozen,
 * audio_model_builder.py:327-344) as per-channel statistics + fused elementwise passes.
 * ---------------------------------------------------------------------------------------------- */
/* stats[c] = sum_rows x, stats[C + c] = sum_rows x^2 as doubles, written (not accumulated) by a deterministic
 * two-stage reduction: per-CTA partial sums in `ws` (>= ASF_STATS_WS_FLOATS(C) floats), ordered final pass by the last
 * CTA (`ticket`: ASF_STATS_TICKETS zero-initialised uint32, self-resetting).  See asf_epilogue_t. */
int asf_channel_stats(const void* x, int dtype, int64_t rows, int C, int64_t ld, double* stats, float* ws,
                      uint32_t* ticket, void* stream);
/* Allocation-free variant of asf_channel_stats.  No counterpart in the reference: added
 * for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  Same result, bit for bit; the caller passes inputs and outputs only.  The
 * reduction's workspace and tickets are file-scope device arrays inside the library (ASF_NOALLOC_SCRATCH_BYTES per GPU
 * that loads it, held until exit), so C <= ASF_NOALLOC_MAX_CHANNELS; more channels are rejected with ASF_E_BADARG.
 * The call allocates nothing (no cudaMalloc / cudaFree / allocating library routine).  EVERY call on a device shares
 * the one scratch array: calls that overlap on the GPU (different streams, concurrent graph branches) race -- order
 * them on one stream or with events. */
#define ASF_NOALLOC_MAX_CHANNELS 256
#define ASF_NOALLOC_SCRATCH_BYTES (4ll * ASF_STATS_WS_FLOATS(ASF_NOALLOC_MAX_CHANNELS) + 4ll * ASF_STATS_TICKETS)
int asf_channel_stats_noalloc(const void* x, int dtype, int64_t rows, int C, int64_t ld, double* stats, void* stream);
/* train-mode finalize: mean/var from stats over `count` rows (in double: var = E[x^2] - mean^2 is formed from the
 * double sums); scale = gamma*invstd, shift = beta - mean*scale;
 * running_mean/var momentum update (unbiased var) and num_batches_tracked += 1 when running_mean != NULL. */
int asf_bn_finalize(const double* stats, int C, double count, const float* gamma, const float* beta, float eps,
                    float momentum, float* running_mean, float* running_var, int64_t* num_batches_tracked,
                    float* scale, float* shift, float* save_mean, float* save_invstd, void* stream);
/* train-mode BatchNorm forward in one pass over the raw conv output: asf_bn_finalize (same side effects, computed
 * per block from `stats`) fused with asf_bn_apply; y == NULL publishes the statistics only.  Replaces the forward of
 * nn.BatchNorm2d in training mode (+ ReLU / residual add: resnet_helper.py:181-225, :335-341). */
int asf_bn_train_fwd(const void* x, int64_t ld_x, const void* res, int64_t ld_res, void* y, int64_t ld_y, int dtype,
                     int64_t rows, int C, const double* stats, double count, const float* gamma, const float* beta,
                     float eps, float momentum, float* running_mean, float* running_var, int64_t* num_batches_tracked,
                     float* scale, float* shift, float* save_mean, float* save_invstd, int relu, void* stream);
/* eval/frozen: scale = gamma / sqrt(running_var + eps), shift = beta - running_mean*scale */
int asf_bn_fold(const float* gamma, const float* beta, const float* running_mean, const float* running_var,
                float eps, int C, float* scale, float* shift, void* stream);
/* y = [relu]( x*scale[c] + shift[c] (+ res) ) */
int asf_bn_apply(const void* x, int64_t ld_x, const void* res, int64_t ld_res, void* y, int64_t ld_y, int dtype,
                 int64_t rows, int C, const float* scale, const float* shift, int relu, void* stream);
/* backward reductions: red[c] = sum dz, red[C + c] = sum dz * (x - mean[c]) * invstd[c]  (doubles, written; same
 * deterministic two-stage reduction, workspace and ticket as asf_channel_stats) */
int asf_bn_bwd_reduce(const void* dz, int64_t ld_dz, const void* x, int64_t ld_x, int dtype, int64_t rows, int C,
                      const float* mean, const float* invstd, double* red, float* ws, uint32_t* ticket, void* stream);
/* dx = gamma*invstd * (dz - red[c]/count - xhat * red[C+c]/count);  dgamma (+)= red[C+c], dbeta (+)= red[c]
 * (+= when accumulate != 0, the autograd convention for an existing .grad) */
int asf_bn_bwd_apply(const void* dz, int64_t ld_dz, const void* x, int64_t ld_x, void* dx, int64_t ld_dx, int dtype,
                     int64_t rows, int C, const float* gamma, const float* mean, const float* invstd,
                     const double* red, double count, float* dgamma, float* dbeta, int accumulate, void* stream);
/* elementwise helper: y = (mask > 0) ? (a (+ b)) : 0  -- ReLU backward / gradient accumulation */
int asf_relu_bwd(const void* a, int64_t ld_a, const void* b, int64_t ld_b, const void* mask, int64_t ld_mask,
                 void* y, int64_t ld_y, int dtype, int64_t rows, int C, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Pools and head.  nn.MaxPool2d([3,3],[2,2],[1,1]) of stem_helper.py:161; nn.AvgPool2d + cat + Dropout +
 * Linear (+ Softmax + mean) of audio_slowfast/models/head_helper.py:70-109.
 * ---------------------------------------------------------------------------------------------- */
/* y[b,tp,fp,c] = max over the 3x3/s2/p1 window of relu(x*scale+shift);  x [B][T][F][C], y [B][To][Fo][C].
 * argmax (optional, bf16 with C % 8 == 0 only): [B][To][Fo][C/8] uint32, one nibble per channel = position
 * (dt+1)*3 + (df+1) of the window's first positive maximum (15 = none) -- what torch's MaxPool2d keeps as indices. */
int asf_bn_relu_maxpool_fwd(const void* x, void* y, int64_t ld_y, int dtype, int B, int T, int F, int C,
                            const float* scale, const float* shift, uint32_t* argmax, void* stream);
/* dz[b,t,f,c] = (a>0) * sum over windows whose (first) argmax is (t,f) of dy;  a = relu(x*scale+shift).
 * With the forward's argmax codes the kernel reads only the codes and dy (x, scale, shift are not touched);
 * argmax == NULL recomputes the windows from x. */
int asf_bn_relu_maxpool_bwd(const void* x, const void* dy, int64_t ld_dy, void* dz, int dtype, int B, int T, int F,
                            int C, const float* scale, const float* shift, const uint32_t* argmax, void* stream);
/* feat[b][c_off + c] = mean over P positions of x[b][p][c]   (fp32 out, row pitch ld_feat) */
int asf_avgpool_fwd(const void* x, int dtype, int B, int P, int C, float* feat, int64_t ld_feat, int c_off,
                    void* stream);
/* dx[b][p][c] = dfeat[b][c_off + c] / P, masked by (mask > 0) when mask != NULL */
int asf_avgpool_bwd(const float* dfeat, int64_t ld_feat, int c_off, const void* mask, void* dx, int dtype, int B,
                    int P, int C, void* stream);
/* ---- the classifier head and the loss in four launches (csrc/head.cu)
 * Replaces ResNetBasicHead.forward (dropout + one nn.Linear per task, audio_slowfast/models/head_helper.py:70-109),
 * compute_loss (tools/train_utils.py:76-101: F.cross_entropy per task, mean over tasks) and their backward passes.
 * Arrays indexed by head are HOST arrays of `heads` (1 or 2) entries holding device pointers / sizes. */
/* y = x * (u < keep) / keep, mask = (u < keep) as floats (u: uniform [0,1) draws, torch.rand_like) */
int asf_dropout_apply(const float* x, const float* u, float* y, float* mask, int64_t n, float keep, void* stream);
/* out[h][b][j] = sum_k feat[b][k] W[h][j][k] + bias[h][j]   (bias or bias[h] may be NULL) */
int asf_head_fwd(const float* feat, int64_t ld_feat, int B, int K, int heads, const float* const* W,
                 const float* const* bias, const int* N, float* const* out, const int64_t* ld_out, void* stream);
/* loss_heads[h] = mean_b CE(logits[h][b], labels[h][b]), *loss = mean_h loss_heads[h];
 * dl[h][b][j] = d loss / d logits[h][b][j] = (softmax_j - [j == label]) / (B * heads)  (dl or dl[h] may be NULL).
 * row_scratch: heads * B floats; ticket: one zero-initialised uint32 (self-resetting; launches sharing it must be
 * stream-ordered). */
int asf_softmax_ce(int heads, const float* const* logits, const int64_t* ld, const int* N, const int64_t* const* labels,
                   int B, float* loss, float* loss_heads, float* const* dl, const int64_t* ld_dl, float* row_scratch,
                   unsigned* ticket, void* stream);
/* dW[h][j][k] (+)= sum_b dl[h][b][j] feat[b][k];  db[h][j] (+)= sum_b dl[h][b][j];
 * dfeat[b][k] = (mask ? mask[b][k] * inv_keep : 1) * sum_{h,j} dl[h][b][j] W[h][j][k]   (dW/db/dfeat entries may be NULL) */
int asf_head_bwd(int heads, const float* const* dl, const int64_t* ld_dl, const int* N, const float* const* W,
                 const float* feat, int64_t ld_feat, const float* mask, float inv_keep, int B, int K, float* const* dW,
                 const int* acc_w, float* const* db, const int* acc_b, float* dfeat, int64_t ld_dfeat, void* stream);
/* small fp32 GEMM on CUDA cores: C[M][N] = alpha * op(A) op(B) + beta*C (+ bias[n]); row-major, ta/tb transpose */
int asf_sgemm(const float* A, const float* Bm, float* C, int M, int N, int K, int ta, int tb, int64_t lda,
              int64_t ldb, int64_t ldc, float alpha, float beta, const float* bias, void* stream);
/* in-place row softmax over [rows][n] fp32 */
int asf_softmax_rows(float* x, int rows, int n, int64_t ld, void* stream);
/* y = x * mask * inv_keep (dropout with a caller-provided {0,1} mask); in place allowed */
int asf_scale_mask(const float* x, const float* mask, float* y, int64_t n, float inv_keep, void* stream);

/* gather rows: dst[b][i][:] = src[b][idx[i]][:]  (fp32; pack_pathway_output index_select) */
int asf_gather_rows(const float* src, float* dst, const int32_t* idx, int B, int T_src, int n_idx, int F,
                    void* stream);

/* ------------------------------------------------------------------------------------------------
 * Waveform staging: clip slicing on the device.  Replaces the host-side whole-video read + slice of
 * audio_slowfast/datasets/audio_loader_epic.py:45, :138 and the per-view crops of get_start_end_idx
 * (datasets/utils.py:12-40; 10 test views: datasets/epickitchens.py:46-47).
 *   out[b][j] = samples[starts[b] + j] for j < n while starts[b] + j < limits[b] (absolute end of that clip's video
 *   inside `samples`), else 0;  valid[b] (optional) = number of real samples of clip b.
 * ---------------------------------------------------------------------------------------------- */
int asf_gather_clips(const float* samples, const int64_t* starts, const int64_t* limits, int B, int n, float* out,
                     int64_t out_ld, int32_t* valid, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Metrics and the multi-view test ensemble on the device (no per-iteration host sync).  Replaces
 * audio_slowfast/utils/metrics.py:11-41, :80-108 (topks_correct, multitask_topks_correct), their nine .item() syncs
 * in tools/train_net.py:497-539, and EPICTestMeter.update_stats / finalize_metrics (utils/meters.py:1641-1755) fed by
 * tools/test_net.py:194-239.
 *   asf_topk_counts       counts[0..1] += samples whose label_a ranks inside the top k1 / k2 of preds_a; with a second
 *                         task (preds_b != NULL) counts[2..3] the same for b and counts[4..5] for both at once (the
 *                         "action" accuracy).  Rank ties go to the lower class index.  uint64 integer counters.
 *   asf_ensemble_scatter  clip id = video * num_views + view: stores the clip's scores in clips[video][view][:], its
 *                         label in vid_labels[video], seen[video][view] = 1.
 *   asf_ensemble_reduce   ens[video][:] = sum (mode 0) / max(0, .) (mode 1) over the views, in view order.
 * ---------------------------------------------------------------------------------------------- */
int asf_topk_counts(const float* preds_a, int Ca, int64_t lda, const int64_t* labels_a, const float* preds_b, int Cb,
                    int64_t ldb, const int64_t* labels_b, int B, int k1, int k2, unsigned long long* counts,
                    void* stream);
int asf_ensemble_scatter(const float* preds, int C, int64_t ld, const int64_t* labels, const int64_t* clip_ids, int B,
                         int num_views, int num_videos, float* clips, int64_t* vid_labels, int32_t* seen,
                         void* stream);
int asf_ensemble_reduce(const float* clips, int C, int num_views, int num_videos, int mode, float* ens, void* stream);

/* ------------------------------------------------------------------------------------------------
 * SyncBatchNorm statistics exchange over NVLink peer memory (one process per GPU).  Replaces the all-reduce that
 * torch.nn.SyncBatchNorm issues per layer and pass (the reference converts to it at models/build.py:45-46).
 *   data[0..n) (doubles, this rank's sums) <- sum over ranks, added in rank order (bit-identical on every rank).
 * peer_base_dev: DEVICE array of `world` addresses, the base of every rank's symmetric buffer as mapped into THIS
 * process; slot_off_bytes: offset (16-byte aligned) of this exchange's 2 x world x n doubles (every rank pushes its n
 * values, n even, into slot [epoch parity][its rank] of every rank's buffer), flag_off_bytes: offset of its `world` uint32
 * flags (both inside the symmetric buffer, zero-initialised, the same on every rank); epoch: local device counter of this
 * exchange, zero-initialised.  Every rank must call it the same number of times, stream-ordered per exchange.
 * ---------------------------------------------------------------------------------------------- */
int asf_peer_allreduce_f64(double* data, int n, const unsigned long long* peer_base_dev, int rank, int world,
                           long long slot_off_bytes, long long flag_off_bytes, unsigned* epoch, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ASF_B200_H_ */
