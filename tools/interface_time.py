"""Developer timing of the two-material interface problem (cuthho_square -i): K1 (uncut cells) + K2i (cut cells), the
system build (repeated assembly: structure cached), PCG.  PB_IFC_SORT_ALL=1 selects the first version (all triplets sorted,
scalar-CSR SpMV)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import proton_b200 as pb
from proton_b200 import PB

k = int(sys.argv[1]) if len(sys.argv) > 1 else 2
for N in [int(v) for v in (sys.argv[2:] or ["512", "1024", "2048"])]:
    ctx = pb.Context(0)
    ctx.mesh(pb.mesh_init_params(N, N), PB.QUAD_POLY_DUNAVANT)
    ctx.cut_levelset(0, 0.35, 0.5, 0.5); ctx.cut_preprocess(4)
    hdi = pb.hho_degree_info(k + 1, k)
    for _ in range(3):
        ctx.assemble(hdi, stab_kind=PB.STAB_CUT, problem=PB.PROBLEM_INTERFACE)
    t = ctx.timings()
    ctx.conjugated_gradient(pb.cg_params(0.0, 1e300, 20, True))
    r = ctx.conjugated_gradient(pb.cg_params(0.0, 1e300, 200, True))
    ms_it = ctx.timings()["ms_pcg"] / (r["iterations"] + 1)
    F = k + 1
    by = (8.0 * ctx.nnz + 64.0 * ctx.n_rows / F + 16.0 * ctx.n_rows if F > 1 else 12.0 * ctx.nnz + 20.0 * ctx.n_rows) + 11 * 8.0 * ctx.n_rows
    line = (f"interface k={k} N={N}: cut cells {ctx.cut_counts()[0]}, unknowns {ctx.n_rows}, nnz {ctx.nnz}; local operators (K1 + K2i) {t['ms_local']:.2f} ms, "
            f"system (K3 on the regular rows + sorted triplets of the cut cells' rows) {t['ms_scatter']:.1f} ms -> {N*N/(t['ms_local']+t['ms_scatter'])/1e3:.1f} Mcells/s; PCG (block SpMV + scalar pass over the cut cells' rows) {ms_it:.3f} ms/iter = {by/ms_it/1e6:.0f} GB/s")
    if N <= 512:
        t0 = time.time(); r = ctx.conjugated_gradient(pb.cg_params(1e-12, 1e8, 20 * ctx.n_rows, True)); e = np.sqrt(ctx.error_norms()[0])
        line += f"; solve to 1e-12: {r['iterations']} iterations, {time.time()-t0:.2f} s, H1 error {e:.6e}"
    print(line, flush=True)
    ctx.close()
