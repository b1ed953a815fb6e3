#!/usr/bin/env bash
# Builds libasf_b200.so (all CUDA kernels + the C ABI) for sm_100a, in-tree.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
OUT="$HERE/asf_b200/libasf_b200.so"
mkdir -p "$HERE/build"
FLAGS=(-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -I"$HERE/../include" -I"$HERE/csrc")
objs=()
pids=()
for f in elementwise elementwise_bf16 conv_simt logmel logmel_tc specaug evalmetrics staging peer head tapgemm_sm100; do
  src="$HERE/csrc/$f.cu"; obj="$HERE/build/$f.o"
  stale=0
  for dep in "$src" "$HERE"/csrc/*.cuh "$HERE/../include/asf_b200.h"; do
    [[ ! -f "$obj" || "$dep" -nt "$obj" ]] && stale=1
  done
  if [[ $stale == 1 ]]; then
    "$NVCC" "${FLAGS[@]}" ${ASF_PTXAS_V:+-Xptxas -v} -c "$src" -o "$obj" &
    pids+=($!)
  fi
  objs+=("$obj")
done
for pid in "${pids[@]:-}"; do [[ -z "$pid" ]] || wait "$pid"; done
"$NVCC" -shared -o "$OUT" "${objs[@]}" -cudart static
echo "built $OUT"
