"""Developer timing of K1 (pb_assemble local-operator stage) for the lane mappings selectable through PB_K1_GS."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import proton_b200 as pb

N = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
k = int(sys.argv[2]) if len(sys.argv) > 2 else 2
ctx = pb.Context(0)
ctx.mesh(pb.mesh_init_params(N, N))
hdi = pb.hho_degree_info(k + 1, k)
for gs in ([""] + os.environ.get("K1_VARIANTS", "").split()):
    if gs: os.environ["PB_K1_GS"] = gs
    else: os.environ.pop("PB_K1_GS", None)
    ts = []
    for i in range(6):
        ctx.assemble(hdi); ts.append(ctx.timings()["ms_local"])
    best = min(ts[2:])
    print(f"k={k} N={N} GS={gs or 'default'} K1 ms: {best:.3f}  ({N*N/best/1e3:.1f} Mcells/s) scatter {ctx.timings()['ms_scatter']:.3f}", flush=True)
