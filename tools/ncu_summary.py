"""Summarises an `ncu --set full` report (.ncu-rep) per captured launch: duration, DRAM traffic, tensor-pipe activity,
L2 / shared-memory pressure, registers, dynamic shared memory.   python tools/ncu_summary.py <file.ncu-rep>"""
import csv
import subprocess
import sys

METRICS = [
    ("gpu__time_duration.sum", "time"),
    ("dram__bytes_read.sum", "dram_rd"),
    ("dram__bytes_write.sum", "dram_wr"),
    ("TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed", "tensor_pipe_active_%"),
    ("sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor_mem_active_%"),
    ("l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "smem_tc_wavefronts_%"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2_throughput_%"),
    ("FBSP.TriageCompute.dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram_throughput_%"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_throughput_%"),
    ("launch__registers_per_thread", "regs"),
    ("launch__shared_mem_per_block_dynamic", "dyn_smem"),
    ("launch__grid_size", "grid"),
]
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h, units = rows[0], rows[1]
idx = {name: h.index(name) for name, _ in METRICS if name in h}
ki = h.index("Kernel Name")
print("%-44s " % "kernel" + " ".join("%20s" % short for name, short in METRICS if name in idx))
for r in rows[2:]:
    cells = []
    for name, short in METRICS:
        if name in idx:
            u = units[idx[name]]
            cells.append("%20s" % (r[idx[name]][:12] + (" " + u if u and u != "%" else "")))
    print("%-44s " % r[ki].replace("void ", "").split("(")[0][:44] + " ".join(cells))
