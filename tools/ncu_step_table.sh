#!/usr/bin/env bash
# Per-launch counters of ONE eager training step at the bench shape (the same kernels, in the same order, as the captured
# graph): duration, DRAM bytes, tcgen05 activity (sm__mem_tensor_cycles_active tracks UTCHMMA; sm__pipe_tensor_* does not
# on sm_100), SM / L2 / DRAM throughput.    bash tools/ncu_step_table.sh <out.csv> [bench args...]
out=$1; shift
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum
M=$M,sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,sm__throughput.avg.pct_of_peak_sustained_elapsed
M=$M,lts__throughput.avg.pct_of_peak_sustained_elapsed,dram__throughput.avg.pct_of_peak_sustained_elapsed
M=$M,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,launch__grid_size,smsp__inst_executed.sum
ASF_PROFILER_RANGE=1 ncu --metrics $M --clock-control none --profile-from-start off --csv --page raw --log-file "$out" \
  python bench.py --no-graph --steps 1 --warmup 5 --no-cpu-baseline --no-stock "$@"
