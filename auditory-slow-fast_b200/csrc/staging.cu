// Clip slicing on the device: B fixed-length windows cut out of waveforms that are already resident in HBM.
//
// Replaces the per-item host work of /root/reference/audio_slowfast/datasets/audio_loader_epic.py:16-63 (pack_audio
// reads the WHOLE video array from HDF5 for every clip, :45, then slices samples[start:end], :138): the raw waveform of
// a video is staged once (pinned host ring -> device, asf_b200/datasets/staging.py) and every training crop, each of
// the 10 test views (datasets/epickitchens.py:46-47, get_start_end_idx datasets/utils.py:12-40) and every sliding
// window is a strided read of it.
//   out[b][j] = samples[starts[b] + j]   for 0 <= j < n and starts[b] + j < limits[b], else 0
// (Python's slice clamps at the end of the array; the samples that the reference would not have are zeros here and
// `valid[b]` reports how many were real, so that the caller can edge-pad frames exactly like audio_loader_epic.py:151-154.)
#include "common.cuh"

namespace asf {

__global__ void __launch_bounds__(256) gather_clips_kernel(const float* __restrict__ samples,
                                                           const int64_t* __restrict__ starts,
                                                           const int64_t* __restrict__ limits, int B, int n,
                                                           int64_t out_ld, float* __restrict__ out,
                                                           int32_t* __restrict__ valid) {
  pdl_trigger();
  pdl_wait();
  const int b = blockIdx.y;
  const int64_t s = starts[b], lim = limits[b];
  const float* src = samples + s;
  float* dst = out + (int64_t)b * out_ld;
  // 16-byte vector path when both sides are aligned (start offsets are arbitrary sample indices: usually not)
  const bool vec = ((((uintptr_t)src) | ((uintptr_t)dst)) & 15) == 0;
  const int64_t avail = lim - s < 0 ? 0 : (lim - s < n ? lim - s : n);
  if (vec) {
    const int nv = n >> 2;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += gridDim.x * blockDim.x) {
      const int j = i << 2;
      float4 v;
      if (j + 4 <= avail) {
        v = *(const float4*)(src + j);
      } else {
        v.x = j < avail ? src[j] : 0.f;
        v.y = j + 1 < avail ? src[j + 1] : 0.f;
        v.z = j + 2 < avail ? src[j + 2] : 0.f;
        v.w = j + 3 < avail ? src[j + 3] : 0.f;
      }
      *(float4*)(dst + j) = v;
    }
    for (int j = (nv << 2) + blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x)
      dst[j] = j < avail ? src[j] : 0.f;
  } else {
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x)
      dst[j] = j < avail ? src[j] : 0.f;
  }
  if (valid && blockIdx.x == 0 && threadIdx.x == 0) valid[b] = (int32_t)avail;
}

}  // namespace asf

using namespace asf;

extern "C" int asf_gather_clips(const float* samples, const int64_t* starts, const int64_t* limits, int B, int n,
                                float* out, int64_t out_ld, int32_t* valid, void* stream) {
  ASF_CHECK_ARG(samples && starts && limits && out && B > 0 && n > 0 && out_ld >= n, "asf_gather_clips: bad arguments");
  int gx = (int)cdiv(n, 256 * 4);
  if (gx < 1) gx = 1;
  if (gx > 64) gx = 64;
  ASF_CUDA(launch_pdl(gather_clips_kernel, dim3((unsigned)gx, (unsigned)B), dim3(256), 0, (cudaStream_t)stream, samples,
                      starts, limits, B, n, out_ld, out, valid));
  ASF_LAUNCH_CHECK();
  return 0;
}
