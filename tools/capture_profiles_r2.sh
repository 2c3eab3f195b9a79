#!/bin/bash
# Run ON THE GPU BOX (through gpurun):  tools/capture_profiles_r2.sh TAG
# 1. the plain commands must exit 0 without ncu; 2. ncu launch list of the bench command; 3. one --set full capture per hot kernel
# (each only after its own command has passed).  Summaries are written on the box (the merge back is capped at 64 MiB).
TAG=$1
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-cg-tol --no-interface --no-obstacle --no-cut"
CUT="python tools/cut_time.py 4096 2"
mkdir -p gpurun_out
$CMD > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err || { echo "plain bench failed"; tail -5 gpurun_out/${TAG}_plain.err; exit 1; }
$CUT > gpurun_out/${TAG}_cut_plain.txt 2>&1 || { echo "plain cut run failed"; tail -5 gpurun_out/${TAG}_cut_plain.txt; exit 1; }
cat gpurun_out/${TAG}_cut_plain.txt
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_launch.log 2>&1
for K in local_uncut_kernel scatter_tile_kernel pcg_spmv_blk_kernel mg_cheb_kernel mg_stencil_apply_kernel; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$K -s 3 -c 1 -o gpurun_out/${TAG}_$K -f $CMD > gpurun_out/${TAG}_ncu_$K.log 2>&1
done
# the BASELINE target's instances: K1 with the Dunavant fan + naive stabilisation, K2 on the cut cells, at 4096^2
timeout 600 ncu --set full --clock-control none --import-source on -k regex:local_uncut_kernel -s 2 -c 1 -o gpurun_out/${TAG}_cut4096_local_uncut_kernel -f $CUT > gpurun_out/${TAG}_ncu_cutk1.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:local_cut_kernel -s 2 -c 1 -o gpurun_out/${TAG}_cut4096_local_cut_kernel -f $CUT > gpurun_out/${TAG}_ncu_cutk2.log 2>&1
python tools/profile_summaries.py $TAG gpurun_out
ls -la gpurun_out/*.ncu-rep
rm -f gpurun_out/${TAG}_*.ncu-rep
ls gpurun_out | grep $TAG
