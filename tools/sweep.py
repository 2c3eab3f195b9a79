"""BASELINE config 2: convergence_test sweep k = 0..3 on uniform quads up to 512 x 512 (uncut HHO diffusion), through the
C ABI: assembly, PCG (reference settings: Jacobi, tol 1e-12; divergence guard widened as in apps/convergence_test.cpp),
back-substitution, the driver's error norms."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import proton_b200 as pb
from proton_b200 import PB

ctx = pb.Context(0)
print("k    N   assemble ms  pcg its   pcg ms   ms/it    L2 error      energy error   order(L2) order(en)")
for k in range(4):
    prev = None
    for N in (32, 64, 128, 256, 512):
        hdi = pb.hho_degree_info(k + 1, k)
        ctx.mesh(pb.mesh_init_params(N, N))
        ctx.assemble(hdi); ctx.assemble(hdi)
        t = ctx.timings(); asm = t["ms_local"] + t["ms_scatter"]
        res = ctx.conjugated_gradient(pb.cg_params(1e-12, 1e8, 3 * ctx.n_rows, True))
        ms = ctx.timings()["ms_pcg"]
        ctx.recover()
        e = ctx.error_norms()
        l2, en = float(np.sqrt(e[0])), float(np.sqrt(e[2]))
        o = ("%9.2f %9.2f" % (np.log2(prev[0] / l2), np.log2(prev[1] / en))) if prev else ""
        print(f"{k}  {N:4d}  {asm:10.3f}  {res['iterations']:7d}  {ms:8.2f}  {ms / max(res['iterations'], 1):6.3f}   {l2:.6e}  {en:.6e} {o}", flush=True)
        prev = (l2, en)
