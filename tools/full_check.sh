#!/usr/bin/env bash
# Everything the driver runs at round end, plus the launch list of the bench command under ncu (gpurun_out/).
python -m pytest tests -m gpu -x -q > gpurun_out/final_gpu_suite.log 2>&1; echo "suite rc $?"; tail -2 gpurun_out/final_gpu_suite.log
python __graft_entry__.py smoke > gpurun_out/final_smoke.txt 2>&1; echo "smoke rc $?"; tail -1 gpurun_out/final_smoke.txt
python bench.py > gpurun_out/final_bench_default.json 2> gpurun_out/final_bench_default.err; echo "bench rc $?"
python bench.py --steps 2 --warmup 5 --no-cpu-baseline --no-stock > gpurun_out/final_bench_short.json 2>&1 && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --graph-profiling node -c 3000 --csv --log-file gpurun_out/final_launches_train_frozen.csv python bench.py --steps 2 --warmup 5 --no-cpu-baseline --no-stock > gpurun_out/final_ncu.log 2>&1
echo "ncu rc $?"; wc -l gpurun_out/final_launches_train_frozen.csv
