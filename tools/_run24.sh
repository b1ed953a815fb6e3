python bench.py --no-graph --steps 1 --warmup 5 --no-cpu-baseline --no-stock > gpurun_out/r2_bench24_eager.json 2>&1 || exit 1
bash tools/ncu_step_table.sh gpurun_out/r2_ncu_step_frozen.csv > gpurun_out/r2_ncu24.log 2>&1
tail -2 gpurun_out/r2_ncu24.log; wc -l gpurun_out/r2_ncu_step_frozen.csv
