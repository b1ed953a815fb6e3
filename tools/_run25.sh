python -m pytest tests/test_gpu_model.py -m gpu -q -x -k "pipelined or graph_replay or reproducible" -s > gpurun_out/r2_t25.log 2>&1; tail -5 gpurun_out/r2_t25.log
for cfg in "0 0" "0 1" "1 0" "1 1"; do set -- $cfg
  python bench.py --no-cpu-baseline --no-stock --pipeline $1 --segmented $2 > gpurun_out/r2_bench25_p$1_s$2.json 2> gpurun_out/r2_bench25_p$1_s$2.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2_bench25_p$1_s$2.json').read().strip().splitlines()[-1])
    print('pipeline $1 segmented $2: %.3f ms  %.0f clips/s  e2e %.0f (%.3f ms)'%(d['ms_per_step'],d['value'],d['e2e']['value'],d['e2e']['ms_per_step']))
except Exception as e:
    print('pipeline $1 segmented $2 FAILED', e); print(open('gpurun_out/r2_bench25_p$1_s$2.err').read()[-1500:])
PY
done
for after in start s3_fuse s1_fuse; do
  ASF_EXTRACT_AFTER=$after python bench.py --no-cpu-baseline --no-stock > gpurun_out/r2_bench25_after_$after.json 2>/dev/null
  python -c "
import json
d=json.loads(open('gpurun_out/r2_bench25_after_$after.json').read().strip().splitlines()[-1])
print('extract after $after: %.3f ms e2e %.3f ms'%(d['ms_per_step'],d['e2e']['ms_per_step']))"
done
