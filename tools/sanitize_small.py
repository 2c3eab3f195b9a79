"""Small run of every round-2 code path for compute-sanitizer (memcheck): windowed cut preprocessing on strips, K3 tile kernel,
multigrid set-up / V-cycle / graph replay, static error norms, keep_local = 2, local solution."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import proton_b200 as pb
from proton_b200 import PB

for k in (1, 2, 3):
    ctx = pb.Context(0)
    ctx.mesh(pb.mesh_init_params(21, 18)).assemble(pb.hho_degree_info(k + 1, k), fun=PB.FUN_TANH)
    r = ctx.conjugated_gradient(pb.cg_params(1e-12, 1e8, 500, 3)); assert r["exit_reason"] == 0, r
    a = ctx.error_norms(); b = ctx.error_norms_static(); assert np.allclose(a, b, rtol=1e-10)
    ctx.recover(); ctx.close()
for (rb, re) in ((0, 7), (7, 13), (13, 20)):
    c = pb.Context(0)
    c.mesh(pb.mesh_init_params(20, 20), quad_kind=PB.QUAD_POLY_DUNAVANT, row_begin=rb, row_end=re)
    c.cut_levelset(0, 0.35, 0.5, 0.5); c.cut_preprocess(4)
    c.assemble(pb.hho_degree_info(3, 2), stab_kind=PB.STAB_CUT, problem=PB.PROBLEM_FICTDOM)
    c.cut_download(); c.csr(); c.close()
c = pb.Context(0)
c.mesh(pb.mesh_init_params(24, 24), quad_kind=PB.QUAD_POLY_DUNAVANT)
c.cut_levelset(0, 0.35, 0.5, 0.5); c.cut_preprocess(4)
c.assemble(pb.hho_degree_info(3, 2), stab_kind=PB.STAB_CUT, problem=PB.PROBLEM_FICTDOM, keep_local=2)
r = c.conjugated_gradient(pb.cg_params(1e-12, 1e8, 2000, 3)); assert r["exit_reason"] == 0, r
c.error_norms_static(); c.local(7, 0, 10)
c.assemble(pb.hho_degree_info(3, 2), stab_kind=PB.STAB_CUT, problem=PB.PROBLEM_INTERFACE)
r = c.conjugated_gradient(pb.cg_params(1e-10, 1e8, 20000, True)); assert r["exit_reason"] == 0, r
c.recover(); c.close()
c = pb.Context(0)
c.mesh(pb.mesh_init_params(16, 16, -1.0, 1.0, -1.0, 1.0))
print("obstacle", c.obstacle_solve(1, cgp=pb.cg_params(1e-13, 1e300, 100000, 3))[:2]); c.close()
print("sanitize script done")
