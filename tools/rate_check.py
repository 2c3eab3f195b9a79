import sys, time, numpy as np
sys.path.insert(0, '/root/repo')
import proton_b200 as pb
from proton_b200 import PB
ctx = pb.Context(0)
k = 2
for N in (512, 1024, 2048):
    ctx.mesh(pb.mesh_init_params(N, N)).assemble(pb.hho_degree_info(k + 1, k))
    t0 = time.time()
    r = ctx.conjugated_gradient(pb.cg_params(1e-13, 1e8, 3 * ctx.n_rows, 2))
    e = np.sqrt(ctx.error_norms())
    print(N, r["iterations"], r["exit_reason"], "%.1f s" % (time.time() - t0), e, flush=True)
