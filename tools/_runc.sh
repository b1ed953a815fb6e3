python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k tapgemm > gpurun_out/r2_c_tests.log 2>&1; tail -2 gpurun_out/r2_c_tests.log
python tools/tg_probe.py s2_c,s3_a --modes fwd_frozen,fwd_plain,dgrad_res 2>&1 | cut -c1-200
python bench.py --no-cpu-baseline --no-stock > gpurun_out/r2_benchc.json 2>/dev/null
python -c "
import json
d=json.loads(open('gpurun_out/r2_benchc.json').read().strip().splitlines()[-1])
print('train: %.3f ms  fwd kernel %.3f ms/step e2e %.0f'%(d['ms_per_step'], d['roofline']['kernel_ms_per_step'], d['e2e']['value']))"
python bench.py --workload infer --no-cpu-baseline --no-stock > gpurun_out/r2_benchc_infer.json 2>/dev/null
python -c "
import json
d=json.loads(open('gpurun_out/r2_benchc_infer.json').read().strip().splitlines()[-1])
print('infer: %.3f ms %.0f'%(d['ms_per_step'], d['value']))"
