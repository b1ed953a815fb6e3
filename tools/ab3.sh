# A/B of environment knobs on the default bench: tools/ab3.sh "label ENV=.. ENV=.." ...   (results: gpurun_out/ab_<label>.json)
for spec in "$@"; do
  set -- $spec; label=$1; shift
  env "$@" python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-stock > gpurun_out/ab_$label.json 2> gpurun_out/ab_$label.err || tail -5 gpurun_out/ab_$label.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/ab_$label.json").read().strip().splitlines()[-1])
    print("$label", round(d["value"],1), "clips/s", round(d["ms_per_step"],3), "ms; e2e", round(d["e2e"]["value"],1))
except Exception as e: print("$label failed", e)
PY
done
