#!/usr/bin/env bash
# Runs the GPU test-suite in separate processes (a faulting kernel must not poison the other groups) and
# leaves full logs under gpurun_out/.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,memory.total,clocks.max.sm --format=csv | tee gpurun_out/gpu.txt
run() { # name, timeout, args...
  local name=$1 to=$2; shift 2
  echo "=== $name"
  timeout "$to" python -m pytest -m gpu -q --tb=short -p no:cacheprovider -s "$@" > "gpurun_out/$name.log" 2>&1
  echo "exit $? : $(tail -n 3 gpurun_out/$name.log | tr '\n' ' ')"
}
run logmel 600 tests/test_gpu_logmel.py
run kernels_simt 600 tests/test_gpu_kernels.py -k "not tapgemm"
run model_fp32 900 tests/test_gpu_model.py -k "fp32 or float32"
ASF_DEBUG_SYNC=1 run tapgemm 900 tests/test_gpu_kernels.py -k "tapgemm"
run model_bf16 900 tests/test_gpu_model.py -k "not (fp32 or float32)"
run specaug 600 tests/test_gpu_specaug.py
run pipeline 600 tests/test_gpu_pipeline.py
grep -h -E "FAILED|ERROR|passed|failed" gpurun_out/*.log | sort | uniq -c | sort -rn | head -60
