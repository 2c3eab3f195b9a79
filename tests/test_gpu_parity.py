"""GPU parity tests (run with -m gpu on a B200): the CUDA path, called through the C ABI, against the CPU oracle.
Bars (BASELINE.json north_star): local matrices 1e-11 relative (Frobenius), numbering and sparsity pattern
bit-exact, discrete solution 1e-8 relative."""
import numpy as np
import pytest
import scipy.sparse as sp
import scipy.sparse.linalg as spla

pytestmark = pytest.mark.gpu

import proton_b200 as pb
from proton_b200 import PB
from oracle import orc

TOL_LOCAL = 1e-11
TOL_SOL = 1e-8


def frob_rel(a, b):
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300)


@pytest.fixture(scope="module")
def ctx():
    c = pb.Context(0)
    yield c
    c.close()


def oracle_locals(m, cd, fd, stab, problem=0, fun=0, rhs_di=0):
    out = [m.local(c, cd, fd, stab=stab, problem=problem, fun=fun, rhs_di=rhs_di) for c in range(m.ncells)]
    return out


def schur_of(S):
    """Condensed face system of the oracle's full system: K_FF - K_FT K_TT^-1 K_TF and the matching rhs."""
    K = S.matrix().tocsc()
    nT = S.cbs * S.mesh.ncells
    KTT, KTF, KFT, KFF = K[:nT, :nT], K[:nT, nT:], K[nT:, :nT], K[nT:, nT:]
    lu = spla.splu(KTT)
    X = lu.solve(KTF.toarray())
    Sc = KFF.toarray() - KFT @ X
    g = S.rhs[nT:] - KFT @ lu.solve(S.rhs[:nT])
    return Sc, g, KFF.tocsr()


@pytest.mark.parametrize("k", [0, 1, 2, 3])
@pytest.mark.parametrize("stab", [PB.STAB_FANCY, PB.STAB_NAIVE])
def test_local_operators_tensor(ctx, k, stab):
    N = 6
    m = orc.Mesh(N)
    hdi = pb.hho_degree_info(k + 1, k)
    ctx.mesh(pb.mesh_init_params(N, N)).assemble(hdi, stab_kind=stab, keep_local=True)
    lc = ctx.local(PB.LOCAL_LC); oper = ctx.local(PB.LOCAL_OPER); rhs = ctx.local(PB.LOCAL_RHS)
    ref = oracle_locals(m, k + 1, k, stab)
    worst = 0.0
    for c in range(m.ncells):
        r = ref[c]
        worst = max(worst, frob_rel(lc[c], r["data"] + r["stab"]), frob_rel(oper[c][: hdi.rbs - 1], r["oper"]), frob_rel(rhs[c], r["rhs"]))
    assert worst < TOL_LOCAL, worst


def test_numbering_bit_exact(ctx):
    Nx, Ny = 7, 5
    m = orc.Mesh(Nx, Ny)
    ctx.mesh(pb.mesh_init_params(Nx, Ny))
    assert np.array_equal(ctx.cell_faces(), m.cells()[1])
    S = m.assemble(1, 0)
    _, dirichlet = m.faces()
    ct = ctx.compress_table()
    assert np.array_equal(ct[dirichlet == 0], S.compress[dirichlet == 0])
    assert np.all(ct[dirichlet == 1] == -1)
    assert np.array_equal(ctx.points(), m.points())


@pytest.mark.parametrize("k,N", [(0, 8), (1, 8), (2, 8), (3, 6), (1, (9, 5))])
def test_condensed_system_pattern_and_values(ctx, k, N):
    Nx, Ny = N if isinstance(N, tuple) else (N, N)
    m = orc.Mesh(Nx, Ny)
    S = m.assemble(k + 1, k)
    Sc, g, KFF = schur_of(S)
    ctx.mesh(pb.mesh_init_params(Nx, Ny)).assemble(pb.hho_degree_info(k + 1, k))
    rp, col, val, rhs = ctx.csr()
    # pattern == face-face block of the reference LHS pattern, bit-exact (explicit zeros included)
    KFF.sort_indices()
    assert np.array_equal(rp, KFF.indptr.astype(np.int64))
    assert np.array_equal(col, KFF.indices.astype(np.int32))
    A = sp.csr_matrix((val, col, rp), shape=Sc.shape).toarray()
    assert frob_rel(A, Sc) < TOL_LOCAL
    assert frob_rel(rhs, g) < 1e-10


@pytest.mark.parametrize("k,N", [(0, 16), (1, 16), (2, 16), (3, 8)])
def test_pcg_solution_matches_full_system_direct_solve(ctx, k, N):
    m = orc.Mesh(N)
    S = m.assemble(k + 1, k)
    sol = spla.splu(S.matrix().tocsc()).solve(S.rhs)
    ctx.mesh(pb.mesh_init_params(N, N)).assemble(pb.hho_degree_info(k + 1, k))
    res = ctx.conjugated_gradient(pb.cg_params(convergence_threshold=1e-13, max_iter=3 * ctx.n_rows, apply_preconditioner=True))
    assert res["exit_reason"] == PB.CG_CONVERGED
    full = ctx.recover()
    nT = S.cbs * m.ncells
    assert frob_rel(full[nT:], sol[nT:]) < TOL_SOL
    assert frob_rel(full[:nT], sol[:nT]) < TOL_SOL
    # error norms of convergence_test.cpp:254-306 evaluated by the oracle on the GPU solution
    e_gpu = np.sqrt(np.array(S.errors_convergence(full))); e_ref = np.sqrt(np.array(S.errors_convergence(sol)))
    assert np.max(np.abs(e_gpu - e_ref) / e_ref) < TOL_SOL
    # the same norms computed on the device (K7, pb_error_norms)
    e_dev = np.sqrt(ctx.error_norms())
    assert np.max(np.abs(e_dev - e_ref) / e_ref) < TOL_SOL, (e_dev, e_ref)


@pytest.mark.parametrize("k,N", [(1, 24), (2, 24), (3, 16)])
def test_block_jacobi_extension_same_solution_fewer_iterations(ctx, k, N):
    """apply_preconditioner = 2 (fbs x fbs block-Jacobi, no reference counterpart): same discrete solution as the
    reference's diagonal preconditioner to 1e-8, never more iterations (the tanh data make the solve a generic one)."""
    ctx.mesh(pb.mesh_init_params(N, N))
    ctx.assemble(pb.hho_degree_info(k + 1, k), fun=PB.FUN_TANH)
    r1 = ctx.conjugated_gradient(pb.cg_params(1e-13, 1e6, 20000, 1)); x1 = ctx.recover()
    r2 = ctx.conjugated_gradient(pb.cg_params(1e-13, 1e6, 20000, 2)); x2 = ctx.recover()
    assert r1["exit_reason"] == PB.CG_CONVERGED and r2["exit_reason"] == PB.CG_CONVERGED
    assert np.linalg.norm(x1 - x2) / np.linalg.norm(x1) < TOL_SOL
    assert r2["iterations"] <= r1["iterations"]
    if k >= 2:
        assert r2["iterations"] < 0.9 * r1["iterations"], (r1["iterations"], r2["iterations"])


def _cg_history_numpy(A, b, tol, max_iter, precond):
    """the lines conjugated_gradient writes to cg_params::histfile (solver_cg.hpp:63-144), restated with numpy"""
    x = np.zeros_like(b); iA = 1.0 / A.diagonal()
    r = b - A @ x; d = iA * r if precond else r.copy(); nr0 = nr = np.linalg.norm(r); hist = []; it = 0
    while True:
        hist.append(nr / nr0)
        y = A @ d; iMr = iA * r if precond else r; rho = r @ iMr; alpha = rho / (d @ y)
        x = x + alpha * d; r = r - alpha * y; nr = np.linalg.norm(r)
        if nr / nr0 < tol or it > max_iter or nr / nr0 > 1e300:
            break
        iMr = iA * r if precond else r; d = iMr + (r @ iMr) / rho * d; it += 1
    hist.append(nr / nr0)
    return np.array(hist), it


@pytest.mark.parametrize("precond", [True, False])
def test_residual_history_matches_histfile_lines(ctx, precond):
    N, k = 20, 2
    ctx.mesh(pb.mesh_init_params(N, N)).assemble(pb.hho_degree_info(k + 1, k), fun=PB.FUN_TANH)
    rp, col, val, rhs = ctx.csr()
    A = sp.csr_matrix((val, col, rp), shape=(len(rhs), len(rhs)))
    want, it = _cg_history_numpy(A, rhs, 1e-10, 3 * len(rhs), precond)
    ctx.pcg_history(4 * len(rhs))
    res = ctx.conjugated_gradient(pb.cg_params(1e-10, 1e300, 3 * len(rhs), precond))
    got = ctx.history()
    ctx.pcg_history(0)
    assert res["exit_reason"] == PB.CG_CONVERGED and abs(res["iterations"] - it) <= 1
    assert len(got) == res["iterations"] + 2 and got[0] == 1.0 and got[-1] == res["relative_residual"]
    m = min(len(got), len(want)) - 2          # the last lines sit at the tolerance, where rounding decides the count
    rel = np.abs(got[:m] - want[:m]) / want[:m]
    assert np.max(rel[:30]) < 1e-8            # same recurrence: identical to rounding while the iterates still agree
    assert np.max(rel) < 0.25                 # rounding differences grow along the Krylov recurrence, the curve stays the same
    assert len(ctx.history()) == 0


@pytest.mark.parametrize("k,poly", [(0, False), (1, False), (2, False), (3, False), (1, True), (2, True)])
def test_solution_samples_match_the_driver_dat_rows(ctx, k, poly):
    """pb_sample_solution = rows of solution_N_*_k_*.dat (convergence_test.cpp:280-286): oracle quadrature points
    (integrate(msh, cl, 2*cd)), scaled monomial basis of bases.hpp:70-195 restated here, dofs from pb_recover."""
    N, cd = 6, k + 1
    qk = PB.QUAD_POLY_DUNAVANT if poly else PB.QUAD_TENSOR_GAUSS
    m = orc.Mesh(N, poly=poly)
    ctx.mesh(pb.mesh_init_params(N, N), qk).assemble(pb.hho_degree_info(cd, k), stab_kind=PB.STAB_NAIVE if poly else PB.STAB_FANCY)
    ctx.conjugated_gradient(pb.cg_params(1e-13, 1e6, 20000, True))
    full = ctx.recover(); C_ = (cd + 1) * (cd + 2) // 2
    got = ctx.sample_solution()
    bar = m.barycenters(); h = np.hypot(1.0 / N, 1.0 / N)
    rows = []
    for c in range(m.ncells):
        pts, _ = m.quadrature(c, 2 * cd)
        bx = (pts[:, 0] - bar[c, 0]) / (0.5 * h); by = (pts[:, 1] - bar[c, 1]) / (0.5 * h)
        phi = np.stack([bx ** (d - i) * by ** i for d in range(cd + 1) for i in range(d + 1)], axis=1)
        rows.append(np.column_stack([pts, phi @ full[c * C_:(c + 1) * C_]]))
    want = np.vstack(rows)
    assert got.shape == want.shape
    assert np.max(np.abs(got[:, :2] - want[:, :2])) < 1e-14
    assert np.max(np.abs(got[:, 2] - want[:, 2])) < 1e-12 * max(1.0, np.max(np.abs(want[:, 2])))


def test_pcg_iteration_count_matches_reference_cg(ctx):
    """Same algorithm => same iteration count and residual history end as solver_cg.hpp on the same matrix."""
    N, k = 16, 1
    ctx.mesh(pb.mesh_init_params(N, N)).assemble(pb.hho_degree_info(k + 1, k))
    rp, col, val, rhs = ctx.csr()
    for precond in (True, False):
        x_ref, e_ref, it_ref, rr_ref, _ = orc.cg(rp, col, val, rhs, tol=1e-12, max_iter=3 * len(rhs), precond=precond)
        res = ctx.conjugated_gradient(pb.cg_params(convergence_threshold=1e-12, max_iter=3 * len(rhs), apply_preconditioner=precond))
        assert res["exit_reason"] == e_ref
        assert abs(res["iterations"] - it_ref) <= 1
        assert frob_rel(ctx.face_solution(), x_ref) < 1e-9


@pytest.mark.parametrize("k", [0, 1, 2])
def test_local_operators_poly_dunavant(ctx, k):
    """uncut cells of the cut driver: Dunavant fan quadrature + naive stabilisation."""
    N = 5
    m = orc.Mesh(N, poly=True)
    hdi = pb.hho_degree_info(k + 1, k)
    ctx.mesh(pb.mesh_init_params(N, N), quad_kind=PB.QUAD_POLY_DUNAVANT).assemble(hdi, stab_kind=PB.STAB_NAIVE, keep_local=True)
    lc = ctx.local(PB.LOCAL_LC); rhs = ctx.local(PB.LOCAL_RHS)
    worst = 0.0
    for c in range(m.ncells):
        r = m.local(c, k + 1, k, stab=orc.STAB_NAIVE)
        worst = max(worst, frob_rel(lc[c], r["data"] + r["stab"]), frob_rel(rhs[c], r["rhs"]))
    assert worst < TOL_LOCAL, worst


def test_distorted_quads(ctx):
    """general (non-rectangular) quads: displaced interior nodes, as move_nodes produces near the interface."""
    N, k = 6, 2
    m = orc.Mesh(N)
    ctx.mesh(pb.mesh_init_params(N, N))
    xy = ctx.points()
    rng = np.random.default_rng(0)
    interior = np.ones(len(xy), bool)
    interior[(xy[:, 0] < 1e-12) | (xy[:, 0] > 1 - 1e-12) | (xy[:, 1] < 1e-12) | (xy[:, 1] > 1 - 1e-12)] = False
    xy[interior] += rng.uniform(-0.2 / N, 0.2 / N, size=(interior.sum(), 2))
    ctx.set_points(xy)
    import ctypes as C
    from oracle.orc import lib as olib
    # push the same points into the oracle mesh
    olib().orc_mesh_set_points(m.h, xy.ctypes.data_as(C.POINTER(C.c_double)))
    hdi = pb.hho_degree_info(k + 1, k)
    ctx.assemble(hdi, keep_local=True)
    lc = ctx.local(PB.LOCAL_LC)
    worst = max(frob_rel(lc[c], (lambda r: r["data"] + r["stab"])(m.local(c, k + 1, k))) for c in range(m.ncells))
    assert worst < TOL_LOCAL, worst


def test_pipelined_steps_match_the_serial_step(ctx):
    """pb_assemble_pipelined over a sequence of meshes (a different vertex set per step, page-locked buffers, copies on
    their own streams) gives bit for bit the rhs and CSR values of set_points + pb_assemble + pb_get_csr per mesh."""
    import torch
    N, k = 48, 2
    hdi = pb.hho_degree_info(k + 1, k)
    ctx.mesh(pb.mesh_init_params(N, N))
    base = ctx.points()
    rng = np.random.default_rng(1)
    interior = (base[:, 0] > 1e-12) & (base[:, 0] < 1 - 1e-12) & (base[:, 1] > 1e-12) & (base[:, 1] < 1 - 1e-12)
    meshes = []
    for _ in range(5):
        xy = base.copy(); xy[interior] += rng.uniform(-0.15 / N, 0.15 / N, size=(interior.sum(), 2)); meshes.append(xy)
    want = []
    for xy in meshes:
        ctx.set_points(xy); ctx.assemble(hdi)
        _, _, val, rhs = ctx.csr()
        want.append((val.copy(), rhs.copy()))
    n = ctx.n_rows
    pts = [torch.from_numpy(xy.copy()).pin_memory() for xy in meshes]
    out = [torch.zeros(n, dtype=torch.float64).pin_memory() for _ in meshes]
    for p, o in zip(pts, out):
        ctx.assemble_pipelined(hdi, p.data_ptr(), o.data_ptr())
    ctx.pipeline_sync()
    for (val, rhs), o in zip(want, out):
        assert np.array_equal(o.numpy(), rhs)
    _, _, val, rhs = ctx.csr()            # the device holds the last step
    assert np.array_equal(val, want[-1][0]) and np.array_equal(rhs, want[-1][1])
    t = ctx.timings()
    assert t["ms_local"] > 0 and t["ms_scatter"] > 0


@pytest.mark.parametrize("k,Nx,Ny,cuts", [(1, 8, 8, (0, 4, 8)), (2, 7, 9, (0, 2, 5, 9)), (0, 6, 6, (0, 1, 6))])
def test_strip_contexts_assemble_the_same_rows(k, Nx, Ny, cuts):
    """multi-GPU partition, stages 1+2: every strip (with its ghost cell row) reproduces its rows of the whole-mesh
    system bit for bit, without communication."""
    hdi = pb.hho_degree_info(k + 1, k)
    whole = pb.Context(0)
    whole.mesh(pb.mesh_init_params(Nx, Ny)).assemble(hdi)
    rp, col, val, rhs = whole.csr()
    for rb, re in zip(cuts[:-1], cuts[1:]):
        c = pb.Context(0)
        c.mesh(pb.mesh_init_params(Nx, Ny), row_begin=rb, row_end=re).assemble(hdi)
        L = pb.strip_layout(Nx, Ny, rb, re, hdi.fbs)
        assert c.n_rows == L["n_rows"]
        rp_s, col_s, val_s, rhs_s = c.csr()
        r0 = L["cf_begin"] * hdi.fbs
        lo, hi = rp[r0], rp[r0 + c.n_rows]
        assert np.array_equal(rp_s, rp[r0 : r0 + c.n_rows + 1] - lo)
        assert np.array_equal(col_s, col[lo:hi])
        assert np.array_equal(val_s, val[lo:hi])
        assert np.array_equal(rhs_s, rhs[r0 : r0 + c.n_rows])
        assert col_s.min() >= L["col_lo"] * hdi.fbs and col_s.max() < L["col_hi"] * hdi.fbs
        c.close()
    whole.close()
