// Accuracy metrics and the multi-view test ensemble on the device: no per-iteration host synchronisation.
//
// Replaces, without the nine .item() syncs per training iteration and the per-batch .cpu() copies of the test loop,
//   topks_correct / topk_accuracies / multitask_topks_correct       /root/reference/audio_slowfast/utils/metrics.py:11-41, :80-108
//   the metric block of train_epoch / eval_epoch                    tools/train_net.py:497-539
//   EPICTestMeter.update_stats / finalize_metrics (10-view "sum" / "max" ensemble)   utils/meters.py:1641-1755
//   the gather + update of perform_test                             tools/test_net.py:194-239
//
// Rank of the label's score: rank = #{j : p_j > p_label} + #{j < label : p_j == p_label} (ties broken towards the lower
// class index, a fixed rule where torch.topk leaves the order of equal scores unspecified); "top-k correct" = rank < k.
// Multi-task (action) correctness at k = every task's label inside its own top-k, which is what
// multitask_topks_correct's  ge(sum over the first k ranks of the per-task hits, task_count)  evaluates.
// Counters are 64-bit INTEGERS (integer addition is associative: atomics stay deterministic); scores are never summed
// with floating-point atomics -- the ensemble keeps every view's scores and adds them in view order.
#include "common.cuh"

namespace asf {

// warp-cooperative rank of row[label] inside row[0..C)
__device__ __forceinline__ int warp_label_rank(const float* __restrict__ row, int C, int label, int lane) {
  if (label < 0 || label >= C) return C;  // invalid label: never correct
  const float pl = row[label];
  int cnt = 0;
  for (int j = lane; j < C; j += 32) {
    const float p = row[j];
    cnt += (p > pl) || (p == pl && j < label);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  return cnt;
}

// counts[0..1] += #(rank_a < k1), #(rank_a < k2); with a second task: counts[2..3] the same for b and counts[4..5]
// for both tasks at once.  One warp per sample.
__global__ void __launch_bounds__(256) topk_counts_kernel(const float* __restrict__ pa, int Ca, int64_t lda,
                                                          const int64_t* __restrict__ la, const float* __restrict__ pb,
                                                          int Cb, int64_t ldb, const int64_t* __restrict__ lb, int B, int k1,
                                                          int k2, unsigned long long* __restrict__ counts) {
  __shared__ unsigned int s_cnt[6];
  if (threadIdx.x < 6) s_cnt[threadIdx.x] = 0u;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int warp = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5), nwarps = (int)((gridDim.x * blockDim.x) >> 5);
  for (int b = warp; b < B; b += nwarps) {
    const int ra = warp_label_rank(pa + (int64_t)b * lda, Ca, (int)la[b], lane);
    int rb = 0;
    if (pb) rb = warp_label_rank(pb + (int64_t)b * ldb, Cb, (int)lb[b], lane);
    if (lane == 0) {
      if (ra < k1) atomicAdd(&s_cnt[0], 1u);
      if (ra < k2) atomicAdd(&s_cnt[1], 1u);
      if (pb) {
        if (rb < k1) atomicAdd(&s_cnt[2], 1u);
        if (rb < k2) atomicAdd(&s_cnt[3], 1u);
        if (ra < k1 && rb < k1) atomicAdd(&s_cnt[4], 1u);
        if (ra < k2 && rb < k2) atomicAdd(&s_cnt[5], 1u);
      }
    }
  }
  __syncthreads();
  if (threadIdx.x < 6 && s_cnt[threadIdx.x]) atomicAdd(counts + threadIdx.x, (unsigned long long)s_cnt[threadIdx.x]);
}

// Test-time bookkeeping of one batch: clip id -> (video, view); the view's scores are STORED in its own slot
// clips[video][view][:], the label in labels[video], seen[video][view] = 1.
__global__ void __launch_bounds__(256) ensemble_scatter_kernel(const float* __restrict__ preds, int C, int64_t ld,
                                                               const int64_t* __restrict__ labels,
                                                               const int64_t* __restrict__ clip_ids, int B, int num_views,
                                                               int num_videos, float* __restrict__ clips,
                                                               int64_t* __restrict__ vid_labels,
                                                               int32_t* __restrict__ seen) {
  const int64_t total = (int64_t)B * C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int b = (int)(i / C), c = (int)(i - (int64_t)b * C);
    const int64_t id = clip_ids[b];
    const int64_t vid = id / num_views, view = id - vid * num_views;
    if (vid < 0 || vid >= num_videos) continue;
    clips[(vid * num_views + view) * C + c] = preds[(int64_t)b * ld + c];
    if (c == 0) {
      vid_labels[vid] = labels[b];
      seen[vid * num_views + view] = 1;
    }
  }
}

// ensemble[video][c] = sum (mode 0) or max (mode 1) over the views of a video, in view order
__global__ void __launch_bounds__(256) ensemble_reduce_kernel(const float* __restrict__ clips, int C, int num_views,
                                                              int num_videos, int mode, float* __restrict__ ens) {
  const int64_t total = (int64_t)num_videos * C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t vid = i / C;
    const int c = (int)(i - vid * C);
    const float* p = clips + vid * num_views * C + c;
    float acc = mode ? fmaxf(0.f, p[0]) : p[0];  // the reference's accumulators start at zero
    for (int v = 1; v < num_views; ++v) acc = mode ? fmaxf(acc, p[(int64_t)v * C]) : acc + p[(int64_t)v * C];
    ens[i] = acc;
  }
}

}  // namespace asf

using namespace asf;

extern "C" {

int asf_topk_counts(const float* preds_a, int Ca, int64_t lda, const int64_t* labels_a, const float* preds_b, int Cb,
                    int64_t ldb, const int64_t* labels_b, int B, int k1, int k2, unsigned long long* counts,
                    void* stream) {
  ASF_CHECK_ARG(preds_a && labels_a && counts && B > 0 && Ca > 0 && lda >= Ca && k1 > 0 && k2 > 0,
                "asf_topk_counts: bad arguments");
  ASF_CHECK_ARG(!preds_b || (labels_b && Cb > 0 && ldb >= Cb), "asf_topk_counts: second task needs labels / sizes");
  int64_t blocks = cdiv((int64_t)B * 32, 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  topk_counts_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(preds_a, Ca, lda, labels_a, preds_b, Cb, ldb,
                                                                        labels_b, B, k1, k2, counts);
  ASF_LAUNCH_CHECK();
  return 0;
}

int asf_ensemble_scatter(const float* preds, int C, int64_t ld, const int64_t* labels, const int64_t* clip_ids, int B,
                         int num_views, int num_videos, float* clips, int64_t* vid_labels, int32_t* seen,
                         void* stream) {
  ASF_CHECK_ARG(preds && labels && clip_ids && clips && vid_labels && seen && B > 0 && C > 0 && ld >= C && num_views > 0 &&
                    num_videos > 0,
                "asf_ensemble_scatter: bad arguments");
  int64_t blocks = cdiv((int64_t)B * C, 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  ensemble_scatter_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(preds, C, ld, labels, clip_ids, B, num_views,
                                                                             num_videos, clips, vid_labels, seen);
  ASF_LAUNCH_CHECK();
  return 0;
}

int asf_ensemble_reduce(const float* clips, int C, int num_views, int num_videos, int mode, float* ens, void* stream) {
  ASF_CHECK_ARG(clips && ens && C > 0 && num_views > 0 && num_videos > 0 && (mode == 0 || mode == 1),
                "asf_ensemble_reduce: bad arguments");
  int64_t blocks = cdiv((int64_t)num_videos * C, 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  ensemble_reduce_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(clips, C, num_views, num_videos, mode, ens);
  ASF_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
