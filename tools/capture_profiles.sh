#!/bin/bash
# Run ON THE GPU BOX (through gpurun):  tools/capture_profiles.sh TAG
# 1. the plain bench must exit 0; 2. ncu launch list of the same command; 3. one --set full capture per hot kernel.
# Outputs go to gpurun_out/TAG_*; summarise them afterwards HERE with tools/profile_summaries.py TAG.
TAG=$1
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-cg-tol --no-interface"
mkdir -p gpurun_out
$CMD > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err || { echo "plain bench failed"; tail -5 gpurun_out/${TAG}_plain.err; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_launch.log 2>&1
for K in local_uncut_kernel scatter_kernel_t pcg_spmv_blk_kernel; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$K -s 3 -c 1 -o gpurun_out/${TAG}_$K -f $CMD > gpurun_out/${TAG}_ncu_$K.log 2>&1
done
# the merge back is capped at 64 MiB: summarise on the box, keep only the K1 report
python tools/profile_summaries.py $TAG gpurun_out
rm -f gpurun_out/${TAG}_scatter_kernel_t.ncu-rep gpurun_out/${TAG}_pcg_spmv_blk_kernel.ncu-rep
ls -la gpurun_out | grep $TAG
