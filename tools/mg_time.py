"""Developer timing of the multigrid-preconditioned CG (apply_preconditioner = 3) against the reference's Jacobi-PCG.
  python tools/mg_time.py N k [cut] [jacobi]     (jacobi: also run the diagonal preconditioner to tolerance)"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import proton_b200 as pb
from proton_b200 import PB

N = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
k = int(sys.argv[2]) if len(sys.argv) > 2 else 2
cut = "cut" in sys.argv; jac = "jacobi" in sys.argv
tol = 1e-12 if not cut else 1e-9
ctx = pb.Context(0)
ctx.mesh(pb.mesh_init_params(N, N), PB.QUAD_POLY_DUNAVANT if cut else PB.QUAD_TENSOR_GAUSS)
hdi = pb.hho_degree_info(k + 1, k)
if cut:
    ctx.cut_levelset(0, 0.35, 0.5, 0.5); ctx.cut_preprocess(4)
    ctx.assemble(hdi, stab_kind=PB.STAB_CUT, problem=PB.PROBLEM_FICTDOM)
else:
    ctx.assemble(hdi, fun=PB.FUN_TANH if "tanh" in sys.argv else PB.FUN_SINSIN)
nl, lmax, setup = ctx.mg_info()
print(f"N={N} k={k} cut={cut}: unknowns {ctx.n_rows}, mg levels {nl}, lmax {lmax:.3f}, setup {setup:.2f} ms", flush=True)
for nu in ([1, 2, 3] if "sweep" in sys.argv else ["default"]):
    if nu != "default": os.environ["PB_MG_NU"] = str(nu)
    for rep in range(2):
        r = ctx.conjugated_gradient(pb.cg_params(tol, 1e8, 5000, 3))
    t = ctx.timings()["ms_pcg"]
    e = ctx.error_norms()
    print(f"  mg nu={nu}: {r['iterations']} iterations, exit {r['exit_reason']}, relres {r['relative_residual']:.2e}, {t:.1f} ms ({t/(r['iterations']+1):.2f} ms/iter), error norms {np.sqrt(np.abs(e))}", flush=True)
if jac:
    r = ctx.conjugated_gradient(pb.cg_params(tol, 1e8, 3 * ctx.n_rows, True))
    t = ctx.timings()["ms_pcg"]; e = ctx.error_norms()
    print(f"  jacobi: {r['iterations']} iterations, exit {r['exit_reason']}, {t:.1f} ms ({t/(r['iterations']+1):.3f} ms/iter), error norms {np.sqrt(np.abs(e))}", flush=True)
