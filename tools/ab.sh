run() { # label, env...
  label=$1; shift
  env "$@" python bench.py --steps 20 --warmup 5 > gpurun_out/bench_$label.json 2> gpurun_out/bench_$label.err || tail -5 gpurun_out/bench_$label.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_$label.json").read().strip().splitlines()[-1])
    print("$label", round(d["value"],1), "clips/s", round(d["ms_per_step"],3), "ms; e2e", round(d["e2e"]["value"],1))
except Exception as e: print("$label failed", e)
PY
}
