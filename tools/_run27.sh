for c in 0 74 48 32 24; do
 for after in start s2_fuse; do
  ASF_LOGMEL_CTAS=$c ASF_EXTRACT_AFTER=$after python bench.py --no-cpu-baseline --no-stock --pipeline 1 --segmented 1 > gpurun_out/r2_bench27_$c_$after.json 2>/dev/null
  python -c "
import json
d=json.loads(open('gpurun_out/r2_bench27_$c_$after.json').read().strip().splitlines()[-1])
print('logmel ctas $c after $after: %.3f ms  e2e %.3f'%(d['ms_per_step'], d['e2e']['ms_per_step']))"
 done
done
