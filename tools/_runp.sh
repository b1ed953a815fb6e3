python tools/tg_probe.py s2_c,s3_c,s4_c,s3_a,s4_b --modes fwd_frozen,fwd_plain,dgrad,dgrad_res > gpurun_out/r2_probe.txt 2>&1
cat gpurun_out/r2_probe.txt | cut -c1-220
python tools/tg_bench.py s2_c,s3_c,s4_c,s3_a,s4_b --modes fwd_frozen,dgrad > gpurun_out/r2_tgbench.txt 2>&1; cat gpurun_out/r2_tgbench.txt
