for c in 0 132 116 100; do
  ASF_DGRAD_CTAS=$c python bench.py --no-cpu-baseline --no-stock --pipeline 0 --segmented 0 > gpurun_out/r2_bench26_$c.json 2>/dev/null
  python -c "
import json
d=json.loads(open('gpurun_out/r2_bench26_$c.json').read().strip().splitlines()[-1])
print('dgrad ctas $c: %.3f ms'%(d['ms_per_step']))"
done
