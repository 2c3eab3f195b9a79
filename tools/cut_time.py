"""Developer timing of the fictitious-domain assembly (K1 Dunavant instance + K2 cut cells + K3)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import proton_b200 as pb
from proton_b200 import PB

N = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
k = int(sys.argv[2]) if len(sys.argv) > 2 else 2
ctx = pb.Context(0)
ctx.mesh(pb.mesh_init_params(N, N), quad_kind=PB.QUAD_POLY_DUNAVANT)
ctx.cut_levelset(0, 0.35, 0.5, 0.5)
ctx.cut_preprocess(4)
hdi = pb.hho_degree_info(k + 1, k)
for i in range(4):
    ctx.assemble(hdi, stab_kind=PB.STAB_CUT, problem=PB.PROBLEM_FICTDOM)
t = ctx.timings()
print(f"cut k={k} N={N}: local {t['ms_local']:.3f} ms, scatter {t['ms_scatter']:.3f} ms, cut cells {ctx.cut_counts()[0]}", flush=True)
