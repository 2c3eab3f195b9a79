"""Multigrid-preconditioned CG (apply_preconditioner = 3) on the GPU, through the C ABI: the hierarchy and one
application of the preconditioner against the scipy restatement (tests/mg_ref.py) on the oracle's condensed system, the
solve against the reference's Jacobi-PCG and the oracle's direct solve, iteration counts independent of the mesh size."""
import os
import sys

import numpy as np
import pytest
import scipy.sparse.linalg as spla

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import mg_ref
import proton_b200 as pb
from oracle import orc
from proton_b200 import PB

pytestmark = pytest.mark.gpu


def gpu_system(Nx, Ny, k, cut, fun=PB.FUN_SINSIN):
    ctx = pb.Context(0)
    ctx.mesh(pb.mesh_init_params(Nx, Ny), PB.QUAD_POLY_DUNAVANT if cut else PB.QUAD_TENSOR_GAUSS)
    hdi = pb.hho_degree_info(k + 1, k)
    if cut:
        ctx.cut_levelset(0, 0.35, 0.5, 0.5); ctx.cut_preprocess(4)
        ctx.assemble(hdi, stab_kind=PB.STAB_CUT, problem=PB.PROBLEM_FICTDOM, fun=fun)
    else:
        ctx.assemble(hdi, fun=fun)
    return ctx


@pytest.mark.parametrize("Nx,Ny,k,cut", [(12, 9, 1, False), (8, 8, 2, False), (7, 10, 0, False), (9, 9, 3, False), (12, 12, 1, True), (16, 16, 2, True)])
@pytest.mark.parametrize("values", ["fp64", "fp32-copy"])
def test_hierarchy_and_apply_match_restatement(Nx, Ny, k, cut, values, monkeypatch):
    # the V-cycle's own SpMVs stream a single-precision COPY of the matrix values by default (vectors and sums in double);
    # PB_MG_FP64 keeps them on the double values: that variant is the restatement to 1e-8, the default one to fp32 rounding
    if values == "fp64":
        monkeypatch.setenv("PB_MG_FP64", "1")
    else:
        monkeypatch.delenv("PB_MG_FP64", raising=False)
    ctx = gpu_system(Nx, Ny, k, cut)
    m, S, Sc, g = mg_ref.condensed_system(Nx, k, cut=cut, Ny=Ny)
    info = m.cut_info() if cut else None
    nlev, lmax, _ = ctx.mg_info()
    # the library's defaults: 1 Chebyshev step per sweep, on the LINE-block Jacobi (pb_mg_lines.cuh) for the fictitious-domain problem, on the face blocks otherwise
    ref = mg_ref.Preconditioner(m, S, Sc, k, info, nu=1, lmax=lmax, lines=cut)
    assert nlev == len(ref.h.levels)
    assert abs(lmax - mg_ref.power_lmax(ref.A, ref.Dinv)) < 1e-6 * lmax      # same start vector, same iteration
    for l in range(nlev):
        L = ctx.mg_level(l)
        assert (L["Nx"], L["Ny"]) == (ref.h.levels[l]["Nx"], ref.h.levels[l]["Ny"])
        W = L["Nx"] + 1; inner = mg_ref.interior_mask(L["Nx"], L["Ny"])
        st = L["stencil"]; want = ref.h.stencils(l)
        for v in np.flatnonzero(inner):          # entries that point at boundary vertices multiply zeros on the device
            for dj in (-1, 0, 1):
                for di in (-1, 0, 1):
                    if not inner[v + dj * W + di]:
                        st[v, (dj + 1) * 3 + di + 1] = 0.0
        assert np.abs(st[inner] - want[inner]).max() < 1e-9 * np.abs(want).max(), f"level {l}"      # (the top boundary row of level 0 is nobody's)
    rng = np.random.default_rng(5)
    for _ in range(2):
        r = rng.standard_normal(ctx.n_rows)
        z = ctx.mg_apply(r); zr = ref(r)
        # (cut cells: the 1/area-scaled stabilisation makes D^-1 A badly conditioned, which amplifies the 6e-8 rounding of the copy)
        err = np.linalg.norm(z - zr) / np.linalg.norm(zr)
        assert err < (1e-8 if values == "fp64" else (1e-3 if cut else 5e-6)), err
        print(f"mg_apply {values} {Nx}x{Ny} k={k} cut={cut}: relative difference to the restatement {err:.2e}")
    ctx.close()


@pytest.mark.parametrize("N,k,cut,fun", [(24, 0, False, PB.FUN_TANH), (32, 1, False, PB.FUN_TANH), (32, 2, False, PB.FUN_TANH), (24, 3, False, PB.FUN_TANH),
                                         (32, 2, False, PB.FUN_SINSIN), (33, 1, False, PB.FUN_SINSIN), (32, 1, True, PB.FUN_SINSIN), (40, 2, True, PB.FUN_SINSIN),
                                         (20, 0, True, PB.FUN_SINSIN)])
def test_solution_equals_reference_solver(N, k, cut, fun):
    ctx = gpu_system(N, N, k, cut, fun)
    rj = ctx.conjugated_gradient(pb.cg_params(1e-13, 1e8, 200000, True)); xj = ctx.face_solution(); ej = ctx.error_norms()
    ctx.pcg_history(4096)
    rm = ctx.conjugated_gradient(pb.cg_params(1e-13, 1e8, 200000, 3)); xm = ctx.face_solution(); em = ctx.error_norms()
    assert rj["exit_reason"] == PB.CG_CONVERGED and rm["exit_reason"] == PB.CG_CONVERGED
    assert rm["relative_residual"] < 1e-13
    assert np.linalg.norm(xm - xj) < 1e-8 * np.linalg.norm(xj)
    assert abs(em[0] - ej[0]) <= 1e-8 * abs(ej[0])
    assert rm["iterations"] <= (60 if cut else 25) and rm["iterations"] <= max(rj["iterations"], 20)      # (32^2 with the smooth load: Jacobi-PCG itself needs only 14)
    h = ctx.history()                                     # histfile semantics unchanged: iterations + 2 lines, last one < tol
    assert len(h) == rm["iterations"] + 2 and h[0] == 1.0 and h[-1] < 1e-13
    ctx.close()


def test_solution_equals_oracle_direct_solve():
    N, k = 16, 2
    ctx = gpu_system(N, N, k, False)
    assert ctx.conjugated_gradient(pb.cg_params(1e-14, 1e8, 1000, 3))["exit_reason"] == PB.CG_CONVERGED
    full = ctx.recover()
    S = orc.Mesh(N).assemble(k + 1, k)
    sol = spla.splu(S.matrix().tocsc()).solve(S.rhs)
    assert np.linalg.norm(full - sol) < 1e-8 * np.linalg.norm(sol)
    ctx.close()


@pytest.mark.parametrize("k", [1, 2, 3])
def test_iterations_do_not_grow_with_the_mesh(k):
    its = []
    for N in (64, 128, 256, 512):
        ctx = gpu_system(N, N, k, False, PB.FUN_TANH)
        r = ctx.conjugated_gradient(pb.cg_params(1e-12, 1e8, 1000, 3))
        assert r["exit_reason"] == PB.CG_CONVERGED
        its.append(r["iterations"]); ctx.close()
    assert max(its) <= min(its) + 3 and max(its) <= 20, its


@pytest.mark.parametrize("Nx,Ny,k", [(260, 257, 1), (300, 40, 2)])
def test_long_lines_take_the_warp_per_line_solve(Nx, Ny, k, monkeypatch):
    """lines of >= 256 faces on a mesh with < 2048 of them are solved by one warp each (mg_line_solve_warp: the case of the rows of
    vertical faces of a narrow strip); one application of the preconditioner against the restatement"""
    monkeypatch.setenv("PB_MG_FP64", "1")
    ctx = gpu_system(Nx, Ny, k, True)
    m, S, Sc, g = mg_ref.condensed_system(Nx, k, cut=True, Ny=Ny, threads=8)
    nlev, lmax, _ = ctx.mg_info()
    ref = mg_ref.Preconditioner(m, S, Sc, k, m.cut_info(), nu=1, lmax=lmax, lines=True)
    assert abs(lmax - mg_ref.power_lmax(ref.A, ref.Dinv)) < 1e-6 * lmax
    r = np.random.default_rng(11).standard_normal(ctx.n_rows)
    z = ctx.mg_apply(r); zr = ref(r)
    # (k = 2 on the 7.5 : 1 cells of the 300 x 40 mesh: the cut cells' blocks are conditioned like 1e9; measured 1.3e-7)
    assert np.linalg.norm(z - zr) < (1e-8 if k == 1 else 1e-6) * np.linalg.norm(zr)
    ctx.close()


def test_line_smoother_on_the_area_scaled_stabilisations(monkeypatch):
    """naive / cut stabilisation (scaled by 1/area): with the face-block smoother the count grows like sqrt(N) (modes that are
    constant along a mesh line and alternate from line to line); the line-block Jacobi of pb_mg_lines.cuh, the default there, removes
    them.  Same solution either way."""
    hdi = pb.hho_degree_info(2, 1)
    counts = {}
    for lines in ("1", "0"):
        monkeypatch.setenv("PB_MG_LINES", lines)
        for N in (64, 256):
            ctx = pb.Context(0)
            ctx.mesh(pb.mesh_init_params(N, N)).assemble(hdi, stab_kind=PB.STAB_NAIVE, fun=PB.FUN_TANH)
            r = ctx.conjugated_gradient(pb.cg_params(1e-12, 1e8, 1000, 3))
            assert r["exit_reason"] == PB.CG_CONVERGED
            counts[(lines, N)] = (r["iterations"], ctx.face_solution()); ctx.close()
    monkeypatch.delenv("PB_MG_LINES")
    for N in (64, 256):
        a, b = counts[("1", N)][1], counts[("0", N)][1]
        assert np.linalg.norm(a - b) < 1e-8 * np.linalg.norm(b)
    its = {k: v[0] for k, v in counts.items()}
    assert its[("1", 256)] <= its[("1", 64)] + 3, its                       # flat with the lines
    assert its[("0", 256)] >= its[("0", 64)] + 4 and its[("1", 256)] < its[("0", 256)], its
    # the default on the fictitious-domain problem is the line smoother: far fewer iterations than the face-block smoother
    res = {}
    for lines in (None, "0"):
        if lines is None: monkeypatch.delenv("PB_MG_LINES", raising=False)
        else: monkeypatch.setenv("PB_MG_LINES", lines)
        ctx = gpu_system(256, 256, 1, True)
        r = ctx.conjugated_gradient(pb.cg_params(1e-9, 1e8, 1000, 3))
        assert r["exit_reason"] == PB.CG_CONVERGED
        res[lines] = (r["iterations"], ctx.face_solution()); ctx.close()
    monkeypatch.delenv("PB_MG_LINES", raising=False)
    assert res[None][0] <= 18 and res[None][0] < 0.7 * res["0"][0], (res[None][0], res["0"][0])
    assert np.linalg.norm(res[None][1] - res["0"][1]) < 1e-6 * np.linalg.norm(res["0"][1])      # both solved to 1e-9 only


def test_rebuilds_after_reassembly_and_other_degree():
    ctx = pb.Context(0)
    ctx.mesh(pb.mesh_init_params(24, 24))
    sols = {}
    for k in (1, 2, 1):
        ctx.assemble(pb.hho_degree_info(k + 1, k), fun=PB.FUN_TANH)
        assert ctx.conjugated_gradient(pb.cg_params(1e-13, 1e8, 1000, 3))["exit_reason"] == PB.CG_CONVERGED
        x = ctx.face_solution()
        if k in sols:
            assert np.array_equal(x, sols[k])           # deterministic: same kernels, same reductions
        sols[k] = x
    # a second solve on the same assembly reuses the hierarchy
    assert ctx.mg_info()[2] < 0.5          # milliseconds: nothing is rebuilt
    ctx.close()


def test_workspace_follows_the_system_size():
    """pb_pcg after pb_set_face_solution on a fresh assembly, and after re-uploading another cut, must run on a
    workspace of the right size"""
    ctx = pb.Context(0)
    ctx.mesh(pb.mesh_init_params(12, 12)).assemble(pb.hho_degree_info(2, 1))
    ctx.set_face_solution(np.ones(ctx.n_rows))
    r = ctx.conjugated_gradient(pb.cg_params(1e-13, 1e8, 10000, True))
    assert r["exit_reason"] == PB.CG_CONVERGED
    x1 = ctx.face_solution()
    ctx.assemble(pb.hho_degree_info(2, 1))
    assert ctx.conjugated_gradient(pb.cg_params(1e-13, 1e8, 10000, True))["exit_reason"] == PB.CG_CONVERGED
    assert np.allclose(ctx.face_solution(), x1, rtol=0, atol=1e-12)
    # interface problem: a larger cut after a smaller one changes the number of unknowns without a new mesh
    ctx.mesh(pb.mesh_init_params(20, 20), PB.QUAD_POLY_DUNAVANT)
    hdi = pb.hho_degree_info(2, 1); sizes = []
    for radius in (0.2, 0.4):
        ctx.cut_levelset(0, radius, 0.5, 0.5); ctx.cut_preprocess(4)
        ctx.assemble(hdi, stab_kind=PB.STAB_CUT, problem=PB.PROBLEM_INTERFACE)
        assert ctx.conjugated_gradient(pb.cg_params(1e-12, 1e8, 100000, True))["exit_reason"] == PB.CG_CONVERGED
        rp, col, val, rhs = ctx.csr()
        import scipy.sparse as sp
        A = sp.csr_matrix((val, col, rp), shape=(ctx.n_rows, ctx.n_rows))
        x = ctx.face_solution()
        assert np.linalg.norm(A @ x - rhs) < 1e-10 * np.linalg.norm(rhs)
        sizes.append(ctx.n_rows)
    assert sizes[1] > sizes[0]
    ctx.close()


@pytest.mark.parametrize("Nx,Ny", [(2, 2), (2, 3), (3, 2), (2, 9), (5, 3)])
@pytest.mark.parametrize("k", [0, 2])
def test_tiny_meshes(Nx, Ny, k):
    """degenerate hierarchies: a single level, one interior vertex, one coarsening direction"""
    ctx = pb.Context(0)
    ctx.mesh(pb.mesh_init_params(Nx, Ny)).assemble(pb.hho_degree_info(k + 1, k), fun=PB.FUN_TANH)
    rj = ctx.conjugated_gradient(pb.cg_params(1e-13, 1e8, 10000, True)); xj = ctx.face_solution()
    rm = ctx.conjugated_gradient(pb.cg_params(1e-13, 1e8, 10000, 3)); xm = ctx.face_solution()
    assert rj["exit_reason"] == PB.CG_CONVERGED and rm["exit_reason"] == PB.CG_CONVERGED
    assert np.linalg.norm(xm - xj) < 1e-8 * np.linalg.norm(xj)
    ctx.close()
