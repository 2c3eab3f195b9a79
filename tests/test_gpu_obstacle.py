"""GPU parity tests of the generic local-operator kernel (cd != rd, literal full HHO stabilisation) and of the obstacle
problem: the reference's own convergence table (apps/obstacle/results/convergence.txt) through the GPU path."""
import numpy as np
import pytest
import scipy.sparse.linalg as spla

pytestmark = pytest.mark.gpu

import proton_b200 as pb
from proton_b200 import PB
from oracle import orc


def frob_rel(a, b):
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300)


@pytest.fixture(scope="module")
def ctx():
    c = pb.Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("cd,fd", [(0, 0), (0, 1), (1, 1), (2, 2), (1, 2)])
@pytest.mark.parametrize("stab", [PB.STAB_FANCY, PB.STAB_NAIVE])
def test_generic_local_operators(ctx, cd, fd, stab):
    N = 5
    m = orc.Mesh(N)
    hdi = pb.hho_degree_info(cd, fd)
    ctx.mesh(pb.mesh_init_params(N, N)).assemble(hdi, stab_kind=stab, keep_local=True)
    lc = ctx.local(PB.LOCAL_LC); oper = ctx.local(PB.LOCAL_OPER); rhs = ctx.local(PB.LOCAL_RHS)
    worst = 0.0
    for c in range(m.ncells):
        r = m.local(c, cd, fd, stab=stab)
        worst = max(worst, frob_rel(lc[c], r["data"] + r["stab"]), frob_rel(oper[c][: hdi.rbs - 1], r["oper"]), frob_rel(rhs[c], r["rhs"]))
    assert worst < 1e-11, worst


@pytest.mark.parametrize("cd,fd", [(0, 1), (1, 1), (2, 2)])
def test_generic_condensed_solution(ctx, cd, fd):
    N = 8
    m = orc.Mesh(N)
    S = m.assemble(cd, fd)
    sol = spla.splu(S.matrix().tocsc()).solve(S.rhs)
    ctx.mesh(pb.mesh_init_params(N, N)).assemble(pb.hho_degree_info(cd, fd))
    res = ctx.conjugated_gradient(pb.cg_params(1e-13, 100.0, 100000, True))
    assert res["exit_reason"] == PB.CG_CONVERGED
    assert frob_rel(ctx.recover(), sol) < 1e-8


@pytest.mark.parametrize("N", [8, 16, 32])
@pytest.mark.parametrize("k", [0, 1])
def test_obstacle_matches_reference_table_and_oracle(ctx, golden, N, k):
    ctx.mesh(pb.mesh_init_params(N, N, -1.0, 1.0, -1.0, 1.0))
    err, iters, log = ctx.obstacle_solve(k)
    ref = golden["obstacle"][str(N)][f"k{k}"]
    assert abs(err - ref) / ref < 5e-6                    # the table prints 6 digits
    err_o, iters_o, alpha_o = orc.obstacle_solve(N, k)
    assert iters == iters_o                               # same active-set iterations as the saddle-system formulation
    assert abs(err - err_o) / err_o < 1e-8
    a, b, ia = ctx.obstacle_state()
    assert frob_rel(a, alpha_o[: N * N]) < 1e-8


@pytest.mark.parametrize("N,k", [(64, 0), (64, 1), (128, 1)])
def test_obstacle_table_at_64_and_128(ctx, golden, N, k):
    """apps/obstacle/results/convergence.txt rows N = 64, 128; outer iteration counts of the oracle's saddle-system loop at N = 64
    (15 at k = 1); the multigrid-preconditioned inner solve walks through the same active sets"""
    ctx.mesh(pb.mesh_init_params(N, N, -1.0, 1.0, -1.0, 1.0))
    err, iters, log = ctx.obstacle_solve(k)
    ref = golden["obstacle"][str(N)][f"k{k}"]
    assert abs(err - ref) / ref < 5e-6
    if N == 64:
        err_o, iters_o, _ = orc.obstacle_solve(N, k)
        assert iters == iters_o and abs(err - err_o) / err_o < 1e-8
    err_m, iters_m, log_m = ctx.obstacle_solve(k, cgp=pb.cg_params(1e-13, 1e300, 100000, 3))
    assert iters_m == iters and abs(err_m - err) <= 1e-8 * err
    assert [l[1] for l in log_m] == [l[1] for l in log]                 # same active-set sizes in every outer iteration
    assert sum(l[2] for l in log_m) < sum(l[2] for l in log) / 4        # and far fewer inner iterations
