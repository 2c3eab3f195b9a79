import os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import torch
import proton_b200 as pb
from proton_b200 import PB
N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
for rep in range(4):
    ctx = pb.Context(0)
    ctx.mesh(pb.mesh_init_params(N, N), PB.QUAD_POLY_DUNAVANT)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    ctx.cut_levelset(0, 0.35, 0.5, 0.5); ctx.cut_preprocess(4)
    torch.cuda.synchronize(); t1 = time.perf_counter()
    ctx.cut_preprocess(4)
    torch.cuda.synchronize(); t2 = time.perf_counter()
    print(f"N={N} fresh context {rep}: first preprocess {1e3*(t1-t0):.2f} ms, second {1e3*(t2-t1):.2f} ms", flush=True)
    ctx.close()
