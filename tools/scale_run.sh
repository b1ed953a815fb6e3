#!/usr/bin/env bash
# usage: tools/scale_run.sh <N> <tag>   -- bench lines of the data-parallel step on N GPUs of this box (gpurun_out/)
N=$1; TAG=$2; P=29540
run() { # name, env..., -- bench args
  local name=$1; shift
  local envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  P=$((P+1))
  env "${envs[@]}" timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P \
    bench.py --gpus $N --no-cpu-baseline --no-stock "$@" > gpurun_out/${TAG}_n${N}_$name.json 2> gpurun_out/${TAG}_n${N}_$name.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${TAG}_n${N}_$name.json').read().strip().splitlines()[-1])
    print('N=$N $name: %.3f ms  %.0f clips/s  e2e %.0f'%(d['ms_per_step'],d['value'],d['e2e']['value']))
except Exception as e:
    print('N=$N $name FAILED', e); print(open('gpurun_out/${TAG}_n${N}_$name.err').read()[-800:])
PY
}
run frozen A=1 --
run frozen_seg A=1 -- --segmented 1
run frozen_ncclcta4 NCCL_MAX_CTAS=4 --
run fromscratch A=1 -- --bn-freeze 0 --classes 309
run fromscratch_nccl ASF_PEER_SYNCBN=0 -- --bn-freeze 0 --classes 309
