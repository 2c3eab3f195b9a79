mkdir -p gpurun_out; T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port"
timeout 600 $T 29571 tests/dist_gpu_check.py 2>&1 | grep -v "^\[W\|OMP_NUM\|\*\*\*\*\|^$" > gpurun_out/r3c_dist8_check.txt; grep -E "multigrid|interface|DIST" gpurun_out/r3c_dist8_check.txt | tail -12
timeout 900 $T 29572 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r3c_bench_n8.json 2> gpurun_out/r3c_bench_n8.err; echo "bench rc=$?"; tail -3 gpurun_out/r3c_bench_n8.err
python -c "
import json; d=json.loads([l for l in open('gpurun_out/r3c_bench_n8.json') if l.startswith('{')][-1]); print(d['dist_check']); print(d['e2e']['value'], d['value'], d['ms_per_step']); t=d['target_cut']; print(t['preprocess_ms'], t['assemble_ms'], t['solve'], t['pcg_ms_per_iter'], t['roofline']['frac']); print(d['cg']['multigrid'], d['cg']['ms_per_iter']); print(d['interface'])"
timeout 300 $T 29573 tools/mg_time_dist.py 4096 2 cut 2>&1 | grep "N=" > gpurun_out/r3c_mg_dist8.txt; cat gpurun_out/r3c_mg_dist8.txt
PB_MG_LINES=0 timeout 300 $T 29574 tools/mg_time_dist.py 4096 2 cut 2>&1 | grep "N=" > gpurun_out/r3c_mg_dist8_nolines.txt; cat gpurun_out/r3c_mg_dist8_nolines.txt
timeout 300 $T 29575 tools/multi_pcg_time.py 4096 2>&1 | grep "world" > gpurun_out/r3c_pcg_graph8.txt; cat gpurun_out/r3c_pcg_graph8.txt
