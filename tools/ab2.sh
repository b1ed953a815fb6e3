# A/B of environment knobs on the default bench: tools/ab2.sh "label ENV=.. ENV=.." ...
for spec in "$@"; do
  set -- $spec; label=$1; shift
  env "$@" python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_$label.json 2> gpurun_out/bench_$label.err || tail -5 gpurun_out/bench_$label.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_$label.json").read().strip().splitlines()[-1])
    print("$label", round(d["value"],1), "clips/s", round(d["ms_per_step"],3), "ms; e2e", round(d["e2e"]["value"],1))
except Exception as e: print("$label failed", e)
PY
done
