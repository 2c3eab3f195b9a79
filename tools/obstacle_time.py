"""Developer timing of BASELINE config 5: HHO obstacle problem, face degree k, N x N quads on [-1,1]^2, active-set loop."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import proton_b200 as pb
from proton_b200 import PB

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
k = int(sys.argv[2]) if len(sys.argv) > 2 else 1
ctx = pb.Context(0)
ctx.mesh(pb.mesh_init_params(N, N, -1.0, 1.0, -1.0, 1.0))
t0 = time.perf_counter()
ctx.obstacle_begin(k, 1)
t_begin = time.perf_counter() - t0
tot_cg = 0; it = 0; log = []
t0 = time.perf_counter()
while it < 50:
    t1 = time.perf_counter()
    d, na, cgit = ctx.obstacle_iterate(pb.cg_params(1e-13, 1e300, 1000000, True))
    log.append((d, na, cgit, time.perf_counter() - t1)); tot_cg += cgit
    if d < 1e-7: break
    it += 1
t_loop = time.perf_counter() - t0
err = ctx.obstacle_error()
print(f"obstacle k={k} N={N}: begin (local operators, generic kernel) {t_begin*1e3:.1f} ms; {len(log)} outer iterations, {tot_cg} PCG iterations, loop {t_loop:.2f} s "
      f"({1e3*t_loop/max(tot_cg,1):.3f} ms per PCG iteration incl. condense/scatter/update); error {err:.6e}; unknowns {ctx.n_rows}", flush=True)
for l in log: print("   delta %.3e  active %d  pcg %d  %.2f s" % l, flush=True)
