mkdir -p gpurun_out; T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port"
timeout 700 $T 29572 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/r3g_bench_n8.json 2> gpurun_out/r3g_bench_n8.err; echo "bench rc=$?"
PB_MG_LINES=1 timeout 200 $T 29573 tools/mg_time_dist.py 4096 2 cut 2>&1 | grep "N=" > gpurun_out/r3g_mg_dist8_lines.txt; cat gpurun_out/r3g_mg_dist8_lines.txt
timeout 200 $T 29574 tools/mg_time_dist.py 4096 2 cut 2>&1 | grep "N=" > gpurun_out/r3g_mg_dist8_default.txt; cat gpurun_out/r3g_mg_dist8_default.txt
timeout 300 $T 29571 tests/dist_gpu_check.py 2>&1 | grep -E "multigrid|DIST|differs" > gpurun_out/r3g_dist8_check.txt; tail -4 gpurun_out/r3g_dist8_check.txt
