#!/bin/bash
# Developer tool, run HERE after a gpurun capture:  tools/k1_profile.sh gpurun_out/NAME.ncu-rep [kernel-mangled-name]
# prints time / stall / shared-memory shares per "//@ph" phase of local_uncut_kernel.
REP=$1
K=${2:-_ZN2pb18local_uncut_kernelILi3ELi2ELi0ELi1ELi6ELi8ELi32ELi8EEEvNS_9LocalArgsE}
T=$(mktemp -d)
(cd $T && cuobjdump -xelf all /root/repo/proton_b200/libproton_b200.so > /dev/null && nvdisasm -gi *.cubin > all.sass 2>/dev/null)
ncu -i $REP --page source --csv --print-source sass 2>/dev/null > $T/src.csv
ncu -i $REP --page raw --csv 2>/dev/null > $T/raw.csv
python - $T/raw.csv <<'PY'
import csv, sys
rows = list(csv.reader(open(sys.argv[1]))); h, v = rows[0], rows[2]
want = ["gpu__time_duration.sum", "sm__issue_active.avg.pct_of_peak_sustained_elapsed", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "gcc__cache_requests_type_instruction.sum.pct_of_peak_sustained_elapsed",
        "smsp__inst_executed.sum", "smsp__warps_active.avg.per_cycle_active", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread"]
for a, b in zip(h, v):
    if a in want: print(f"{a:90s} {b}")
PY
python /root/repo/tools/sass_lines.py $T/all.sass $K $T/src.csv pb_local.cuh /root/repo/proton_b200/csrc/pb_local.cuh | grep -A40 "^phase"
rm -rf $T
