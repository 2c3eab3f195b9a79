TAG=r3e; mkdir -p gpurun_out
python -m pytest tests -q -m gpu -x 2>&1 | tail -5 > gpurun_out/${TAG}_gpu_tests.txt; cat gpurun_out/${TAG}_gpu_tests.txt
python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE OK')" 2>&1 | tail -2
python bench.py > gpurun_out/${TAG}_bench_n1.json 2> gpurun_out/${TAG}_bench_n1.err; echo "bench rc=$?"
python bench.py --impl reference > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err; echo "ref rc=$?"
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-cg-tol --no-interface --no-obstacle --no-cut"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_launch.log 2>&1
for K in local_uncut_kernel scatter_tile_kernel; do
  timeout 400 ncu --set full --clock-control none --import-source on -k regex:$K -s 3 -c 1 -o gpurun_out/${TAG}_$K -f $CMD > gpurun_out/${TAG}_ncu_$K.log 2>&1
done
timeout 400 ncu --set full --clock-control none --import-source on -k regex:mg_line_solve_kernel -s 20 -c 1 -o gpurun_out/${TAG}_mg_line_solve_kernel -f python tools/mg_time.py 4096 2 cut > gpurun_out/${TAG}_ncu_lines.log 2>&1
python tools/profile_summaries.py $TAG gpurun_out > /dev/null
rm -f gpurun_out/${TAG}_*.ncu-rep
ls gpurun_out | grep $TAG
