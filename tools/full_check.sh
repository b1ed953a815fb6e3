bash tools/gpu_check.sh > gpurun_out/r2_check23.txt 2>&1
python __graft_entry__.py smoke > gpurun_out/r2_smoke23.txt 2>&1; echo smoke rc $? >> gpurun_out/r2_smoke23.txt
python bench.py > gpurun_out/r2_bench23_default.json 2> gpurun_out/r2_bench23_default.err; echo rc $?
python bench.py --impl reference > gpurun_out/r2_bench23_reference.json 2> gpurun_out/r2_bench23_reference.err; echo rc $?
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-stock > gpurun_out/r2_bench23_short.json 2>&1 && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --graph-profiling node -c 3000 --csv --log-file gpurun_out/r2_launches_train_frozen.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-stock > gpurun_out/r2_ncu23.log 2>&1
tail -3 gpurun_out/r2_check23.txt; tail -2 gpurun_out/r2_smoke23.txt
