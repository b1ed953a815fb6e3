python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "stem" > gpurun_out/r2_st_tests.log 2>&1; tail -2 gpurun_out/r2_st_tests.log
python bench.py --no-cpu-baseline --no-stock > gpurun_out/r2_benchst.json 2>/dev/null; python -c "
import json
d=json.loads(open('gpurun_out/r2_benchst.json').read().strip().splitlines()[-1])
print('train: %.3f ms'%(d['ms_per_step']))"
ASF_PROFILER_RANGE=1 timeout 300 ncu --metrics gpu__time_duration.sum -k regex:"stem_wgrad" --clock-control none --profile-from-start off --csv --log-file gpurun_out/r2_ncu_st.csv python bench.py --no-graph --steps 1 --warmup 5 --no-cpu-baseline --no-stock > /dev/null 2>&1
