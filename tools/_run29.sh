N=$1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29561 tools/ddp_check.py > gpurun_out/r2_ddp_check29_n$N.txt 2>&1; grep "freeze=" gpurun_out/r2_ddp_check29_n$N.txt | cut -c1-180
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29562 bench.py --gpus $N --no-cpu-baseline --no-stock --bn-freeze 0 --classes 309 > gpurun_out/r2_bench29_n${N}_fromscratch.json 2> gpurun_out/r2_bench29.err
python -c "
import json
d=json.loads(open('gpurun_out/r2_bench29_n${N}_fromscratch.json').read().strip().splitlines()[-1])
print('N=$N fromscratch: %.3f ms  %.0f clips/s'%(d['ms_per_step'],d['value']))" || tail -c 600 gpurun_out/r2_bench29.err
