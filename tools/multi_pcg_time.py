"""Developer timing (torchrun): PCG ms/iteration of the 4096^2 cut target on the ranks, plain launches vs PB_PCG_GRAPH_MULTI=1."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import proton_b200 as pb
from proton_b200 import PB
from proton_b200.dist import make_rank_context
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
ctx = make_rank_context(pb.mesh_init_params(N, N), PB.QUAD_POLY_DUNAVANT, rank, world, lr, dist)
ctx.cut_levelset(0, 1.0 / 3.0, 0.5, 0.5); ctx.cut_preprocess(4)
ctx.assemble(pb.hho_degree_info(3, 2), stab_kind=PB.STAB_CUT, problem=PB.PROBLEM_FICTDOM)
for mode in ("plain", "graph", "plain", "graph"):
    if mode == "graph": os.environ["PB_PCG_GRAPH_MULTI"] = "1"
    else: os.environ.pop("PB_PCG_GRAPH_MULTI", None)
    ctx.conjugated_gradient(pb.cg_params(0.0, 1e300, 40, True))
    dist.barrier(); torch.cuda.synchronize()
    r = ctx.conjugated_gradient(pb.cg_params(0.0, 1e300, 400, True))
    t = torch.tensor([ctx.timings()["ms_pcg"] / (r["iterations"] + 1)], device=f"cuda:{lr}", dtype=torch.float64); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0: print(f"world {world} N {N} {mode}: {float(t):.4f} ms/iter, relres {r['relative_residual']:.6e}", flush=True)
ctx.close(); dist.destroy_process_group()
