"""Developer timing of a PCG iteration (fixed-iteration run): implicit-block SpMV vs the scalar-CSR one (PB_SPMV_CSR=1)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import proton_b200 as pb

N = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
k = int(sys.argv[2]) if len(sys.argv) > 2 else 2
ctx = pb.Context(0)
ctx.mesh(pb.mesh_init_params(N, N)).assemble(pb.hho_degree_info(k + 1, k))
n, nnz, F = ctx.n_rows, ctx.nnz, k + 1
for mode in ("blk", "csr"):
    if mode == "csr": os.environ["PB_SPMV_CSR"] = "1"
    else:
        os.environ.pop("PB_SPMV_CSR", None); os.environ["PB_SPMV_VAR"] = mode[3:] or "0"
    ctx.conjugated_gradient(pb.cg_params(0.0, 1e300, 20, True))
    r = ctx.conjugated_gradient(pb.cg_params(0.0, 1e300, 200, True))
    ms = ctx.timings()["ms_pcg"] / (r["iterations"] + 1)
    spmv = (12.0 * nnz + 20.0 * n) if (mode == "csr" or F == 1) else (8.0 * nnz + 64.0 * n / F + 16.0 * n)
    by = spmv + 11 * 8.0 * n
    print(f"k={k} N={N} spmv={mode}: {ms:.3f} ms/iter, {by/1e9:.2f} GB/iter -> {by/ms/1e6:.0f} GB/s", flush=True)
t = ctx.timings()
print("assemble: K1 %.3f ms, scatter %.3f ms" % (t["ms_local"], t["ms_scatter"]))
