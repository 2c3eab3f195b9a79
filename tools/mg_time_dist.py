"""Developer timing of the multigrid-preconditioned CG on several GPUs (torchrun): one mesh sharded by cell rows.
  torchrun ... tools/mg_time_dist.py N k [cut] [jacobi]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import proton_b200 as pb
from proton_b200 import PB
from proton_b200.dist import make_rank_context

rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); lr = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
N = int(sys.argv[1]); k = int(sys.argv[2]); cut = "cut" in sys.argv
ctx = make_rank_context(pb.mesh_init_params(N, N), PB.QUAD_POLY_DUNAVANT if cut else PB.QUAD_TENSOR_GAUSS, rank, world, lr, dist if world > 1 else None)
hdi = pb.hho_degree_info(k + 1, k)
if cut:
    ctx.cut_levelset(0, 0.35, 0.5, 0.5); ctx.cut_preprocess(4)
    ctx.assemble(hdi, stab_kind=PB.STAB_CUT, problem=PB.PROBLEM_FICTDOM)
else:
    ctx.assemble(hdi, fun=PB.FUN_TANH)
def sync():
    if world > 1: dist.barrier()
    torch.cuda.synchronize(lr)
tol = 1e-9 if cut else 1e-12
for rep in range(2):
    sync(); t0 = time.perf_counter()
    r = ctx.conjugated_gradient(pb.cg_params(tol, 1e8, 5000, 3))
    sync(); wall = (time.perf_counter() - t0) * 1e3
    if rank == 0:
        print(f"N={N} k={k} cut={cut} world={world} rep {rep}: mg {r['iterations']} its, wall {wall:.1f} ms, device {ctx.timings()['ms_pcg']:.1f} ms, {wall/(r['iterations']+1):.2f} ms/iter", flush=True)
if "jacobi" in sys.argv:
    ctx.conjugated_gradient(pb.cg_params(0.0, 1e300, 20, True)); sync()
    r = ctx.conjugated_gradient(pb.cg_params(0.0, 1e300, 100, True))
    if rank == 0: print(f"   jacobi-pcg {ctx.timings()['ms_pcg']/101:.3f} ms/iter", flush=True)
ctx.close()
if world > 1: dist.destroy_process_group()
