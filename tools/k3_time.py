"""Developer timing of K3 (scatter into the CSR values): tile kernel against the one-warp-per-row kernel (PB_SCATTER_ROWS=1)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import proton_b200 as pb
N = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
k = int(sys.argv[2]) if len(sys.argv) > 2 else 2
ctx = pb.Context(0)
ctx.mesh(pb.mesh_init_params(N, N))
hdi = pb.hho_degree_info(k + 1, k)
vals = {}
for mode in ("tile", "rows", "tile"):
    if mode == "rows": os.environ["PB_SCATTER_ROWS"] = "1"
    else: os.environ.pop("PB_SCATTER_ROWS", None)
    ms = []
    for i in range(6):
        ctx.assemble(hdi); t = ctx.timings()
        if i >= 2: ms.append(t["ms_scatter"])
    F = k + 1; alg = ctx.ncells_local * (16 * F * F + 4 * F) * 8 + ctx.nnz * 8 + ctx.n_rows * 8
    print(f"k={k} N={N} scatter={mode}: {np.mean(ms):.3f} ms (min {min(ms):.3f}), algorithmic {alg/1e9:.2f} GB -> {alg/np.mean(ms)/1e6:.0f} GB/s; K1 {t['ms_local']:.2f} ms", flush=True)
    if N <= 512:
        rp, col, val, rhs = ctx.csr(); vals[mode] = (val.copy(), rhs.copy())
if "rows" in vals:
    print("bit-identical to the row kernel:", np.array_equal(vals["tile"][0], vals["rows"][0]) and np.array_equal(vals["tile"][1], vals["rows"][1]))
