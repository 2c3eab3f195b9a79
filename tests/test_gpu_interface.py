"""GPU parity of the two-material interface problem (`cuthho_square -i`) against oracle/orc_interface.hpp,
which is pinned against the reference's "Interface" table (tests/test_oracle_golden.py):
  * local matrices and loads of the cut cells (make_hho_laplacian_interface + both cut stabilisations), 1e-11;
  * face_table (doubled cut faces) bit-exact, sizes of the condensed system;
  * condensed CSR (pattern bit-exact, values 1e-10) and rhs against the Schur complement of the oracle's full system;
  * PCG solution of the condensed system against the oracle's direct solve (face part), 1e-8."""
import numpy as np
import pytest
import scipy.sparse as sp
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


def _setup(ctx, N, k, r=4, keep_local=False):
    m = orc.Mesh(N, poly=True)
    m.cut_preprocess(refsteps=r)
    ctx.mesh(pb.mesh_init_params(N, N), PB.QUAD_POLY_DUNAVANT)
    ctx.cut_levelset(0, 0.35, 0.5, 0.5)
    ctx.cut_preprocess(r)
    ctx.assemble(pb.hho_degree_info(k + 1, k), stab_kind=PB.STAB_CUT, problem=PB.PROBLEM_INTERFACE, keep_local=keep_local)
    return m


@pytest.mark.parametrize("k,N", [(0, 10), (1, 10), (2, 10), (2, 17)])
def test_interface_cut_cell_local_matrices(ctx, k, N):
    m = _setup(ctx, N, k, keep_local=True)
    cut = np.flatnonzero(m.cut_info()["cell_loc"] == 2)
    lc = ctx.interface_local(2); f = ctx.interface_local(3)
    assert len(cut) == len(lc) > 0
    worst = wf = 0.0
    for s, c in enumerate(cut):
        ref, fref = orc.interface_local(m, int(c), k)
        worst = max(worst, frob_rel(lc[s], ref)); wf = max(wf, frob_rel(f[s], fref))
    assert worst < 1e-11 and wf < 1e-12, (worst, wf)


@pytest.mark.parametrize("k,N", [(0, 10), (1, 12), (2, 10), (2, 20)])
def test_interface_condensed_system_and_solve(ctx, k, N):
    m = _setup(ctx, N, k)
    S = orc.InterfaceSystem(m, k)
    F = k + 1; cbs = (k + 2) * (k + 3) // 2
    # numbering: face_table bit-exact (Dirichlet faces are -1 here, unused there)
    ft = ctx.interface_face_table()
    dirichlet = m.faces()[1].astype(bool)
    assert np.array_equal(ft[~dirichlet], S.face_table[~dirichlet]) and np.all(ft[dirichlet] == -1)
    nT = cbs * S.num_all_cells
    assert ctx.n_rows == S.n - nT
    # condensed system = Schur complement of the oracle's full system on the cell unknowns
    K = S.matrix().tocsc()
    KTT, KTF, KFT, KFF = K[:nT, :nT], K[:nT, nT:], K[nT:, :nT], K[nT:, nT:]
    lu = spla.splu(KTT)
    Sc = KFF - KFT @ sp.csc_matrix(lu.solve(KTF.toarray()))
    bc = S.rhs[nT:] - KFT @ lu.solve(S.rhs[:nT])
    rp, col, val, rhs = ctx.csr()
    A = sp.csr_matrix((val, col, rp), shape=(ctx.n_rows, ctx.n_rows))
    Sd = Sc.toarray(); Ad = A.toarray()
    assert np.max(np.abs(Ad - Sd)) < 1e-10 * np.max(np.abs(Sd))
    # pattern: union of the local cliques with explicit zeros kept (as the reference's triplets)
    stored = np.zeros_like(Ad, dtype=bool); stored[np.repeat(np.arange(ctx.n_rows), np.diff(rp)), col] = True
    assert np.all(np.abs(Sd[~stored]) < 1e-9 * np.max(np.abs(Sd)))      # nothing of the Schur complement falls outside the pattern
    assert all(np.all(np.diff(col[rp[i]:rp[i + 1]]) > 0) for i in range(ctx.n_rows))        # sorted, no duplicates
    assert frob_rel(rhs, bc) < 1e-10
    # solve: the reference's CG settings of the driver (diagonal preconditioner), against the direct solve
    sol = spla.splu(K).solve(S.rhs)
    res = ctx.conjugated_gradient(pb.cg_params(1e-13, 1e8, 50 * ctx.n_rows, True))
    assert res["exit_reason"] == PB.CG_CONVERGED
    x = ctx.face_solution()
    assert frob_rel(x, sol[nT:]) < 1e-8
    if k >= 1:      # the block-Jacobi extension on the interface numbering (cut faces: two blocks each)
        rb = ctx.conjugated_gradient(pb.cg_params(1e-13, 1e8, 50 * ctx.n_rows, 2))
        assert rb["exit_reason"] == PB.CG_CONVERGED and frob_rel(ctx.face_solution(), sol[nT:]) < 1e-8
        ctx.conjugated_gradient(pb.cg_params(1e-13, 1e8, 50 * ctx.n_rows, True))
    # back-substitution on the device, in the layout of the reference's solution vector (cell_table: cut cells twice)
    full = ctx.recover()
    assert full.shape == sol.shape and frob_rel(full, sol) < 1e-8
    # the driver's H1 error: oracle on the GPU solution, and the device kernel
    e_ref = S.h1_error(sol)
    assert abs(S.h1_error(full) - e_ref) < 1e-8 * e_ref
    e_dev = np.sqrt(ctx.error_norms()[0])
    assert abs(e_dev - e_ref) < 1e-8 * e_ref, (e_dev, e_ref)


@pytest.mark.parametrize("k,N", [(0, 20), (1, 20), (2, 10), (2, 40)])
def test_interface_golden_through_the_gpu_path(ctx, golden, k, N):
    """reference table values (cuthho.xlsx, Interface block) through the whole GPU path: assembly, PCG, back-substitution, H1 error"""
    m = _setup(ctx, N, k)
    S = orc.InterfaceSystem(m, k)
    cbs = (k + 2) * (k + 3) // 2; nT = cbs * S.num_all_cells
    K = S.matrix().tocsc()
    res = ctx.conjugated_gradient(pb.cg_params(1e-13, 1e8, 50 * ctx.n_rows, True))
    assert res["exit_reason"] == PB.CG_CONVERGED
    ref = golden["cuthho_interface"][f"k{k}_N{N}_r4"]
    assert abs(np.sqrt(ctx.error_norms()[0]) - ref) / ref < 1e-5
    assert abs(S.h1_error(ctx.recover()) - ref) / ref < 1e-5


@pytest.mark.parametrize("k,N", [(1, 64), (2, 96)])
def test_interface_hybrid_system_equals_all_sorted_system(k, N):
    """the default build (K3 + block rows in interface order for the rows of uncut faces, sorted triplets for the cut
    cells' faces) gives the pattern of the all-sorted first version (PB_IFC_SORT_ALL=1) bit for bit and its values to
    rounding (K3 reads the transposed entry of the symmetric cell block); both symmetric; same PCG iteration count +-1."""
    import os
    got = {}
    for mode in ("hybrid", "sorted"):
        if mode == "sorted":
            os.environ["PB_IFC_SORT_ALL"] = "1"
        try:
            c = pb.Context(0)
            c.mesh(pb.mesh_init_params(N, N), PB.QUAD_POLY_DUNAVANT)
            c.cut_levelset(0, 0.35, 0.5, 0.5); c.cut_preprocess(4)
            for _ in range(2):          # the second assembly takes the cached-structure path
                c.assemble(pb.hho_degree_info(k + 1, k), stab_kind=PB.STAB_CUT, problem=PB.PROBLEM_INTERFACE)
            rp, col, val, rhs = c.csr()
            res = c.conjugated_gradient(pb.cg_params(1e-12, 1e8, 50 * c.n_rows, True))
            got[mode] = (rp, col, val, rhs, res["iterations"], c.face_solution())
            c.close()
        finally:
            os.environ.pop("PB_IFC_SORT_ALL", None)
    (rp, col, val, rhs, it, x), (rp2, col2, val2, rhs2, it2, x2) = got["hybrid"], got["sorted"]
    assert np.array_equal(rp, rp2) and np.array_equal(col, col2)
    scale = np.max(np.abs(val2))
    assert np.max(np.abs(val - val2)) < 1e-13 * scale and np.max(np.abs(rhs - rhs2)) <= 1e-15 * np.max(np.abs(rhs2))
    A = sp.csr_matrix((val, col, rp), shape=(len(rhs), len(rhs)))
    assert abs(A - A.T).max() < 1e-11 * scale
    assert abs(it - it2) <= 1 and frob_rel(x, x2) < 1e-8


def test_interface_full_size_properties():
    """1024 x 1024, k=2 (6.3 M unknowns): sizes from the cut data, sorted CSR, constants in the kernel of every row away
    from the Dirichlet boundary (gradient, consistency, penalty and stabilisation terms all vanish on the common constant),
    the solver runs on it."""
    import ctypes as C
    N, k = 1024, 2; F = k + 1
    c = pb.Context(0)
    try:
        c.mesh(pb.mesh_init_params(N, N), PB.QUAD_POLY_DUNAVANT)
        c.cut_levelset(0, 0.35, 0.5, 0.5); c.cut_preprocess(4)
        c.assemble(pb.hho_degree_info(k + 1, k), stab_kind=PB.STAB_CUT, problem=PB.PROBLEM_INTERFACE)
        face_loc = np.zeros(c.nfaces, np.int32)
        assert pb.lib().pb_cut_download(c._h, None, face_loc.ctypes.data_as(C.c_void_p), None, None, None, None, None) == 0
        ft = c.interface_face_table()
        ncut_faces = int(((face_loc == PB.ON_INTERFACE) & (ft >= 0)).sum())
        assert c.n_rows == (2 * N * (N - 1) + ncut_faces) * F
        ncut = c.cut_counts()[0]
        assert abs(ncut - 8 * 0.35 * N) < 0.01 * 8 * 0.35 * N and ncut_faces == ncut          # a closed curve cuts as many faces as cells
        rp, col, val, rhs = c.csr()
        assert rp[-1] == c.nnz and np.all(np.diff(rp) > 0)
        A = sp.csr_matrix((val, col, rp), shape=(c.n_rows, c.n_rows))
        e0 = np.zeros(c.n_rows); e0[::F] = 1.0
        y = A @ e0
        # rows with a full stencil (no Dirichlet face among the faces of their two cells; the cut cells' rows have more blocks)
        interior = np.diff(rp) >= 7 * F
        assert interior.sum() > 0.99 * c.n_rows
        assert np.max(np.abs(y[interior])) < 1e-9 * np.max(np.abs(val)), np.max(np.abs(y[interior]))
        res = c.conjugated_gradient(pb.cg_params(0.0, 1e300, 50, True))
        assert res["iterations"] >= 50 and np.all(np.isfinite(c.face_solution()))
    finally:
        c.close()


def test_interface_refuses_cut_cells_on_the_dirichlet_boundary(ctx):
    """interface_assembler::assemble_cut throws "Dirichlet boundary on cut cell not supported." (cuthho_square.cpp:1292-1293)"""
    N, k = 12, 1
    ctx.mesh(pb.mesh_init_params(N, N), PB.QUAD_POLY_DUNAVANT)
    ctx.cut_levelset(0, 0.48, 0.5, 0.5)          # the circle reaches into the boundary cells
    ctx.cut_preprocess(4)
    with pytest.raises(pb.ProtonError, match="Dirichlet boundary on cut cell"):
        ctx.assemble(pb.hho_degree_info(k + 1, k), stab_kind=PB.STAB_CUT, problem=PB.PROBLEM_INTERFACE)


@pytest.mark.parametrize("k,N", [(1, 10), (2, 12)])
def test_interface_solution_samples_match_the_driver_dat_rows(ctx, k, N):
    """pb_sample_solution on the interface problem = rows of interface_uT.dat (cuthho_square.cpp:1775-1833): oracle
    quadrature points (uncut cells, then both parts of the cut cells: negative first), scaled monomial basis restated here,
    cell unknowns from pb_recover (reference layout: cut cells carry [u-, u+])."""
    m = _setup(ctx, N, k)
    S = orc.InterfaceSystem(m, k)
    res = ctx.conjugated_gradient(pb.cg_params(1e-13, 1e8, 50 * ctx.n_rows, True))
    assert res["exit_reason"] == PB.CG_CONVERGED
    full = ctx.recover(); got = ctx.sample_solution()
    cd = k + 1; C_ = (cd + 1) * (cd + 2) // 2
    info = m.cut_info(); bar = m.barycenters(); h = np.hypot(1.0 / N, 1.0 / N)
    pts_all = m.points(); cells = m.cells()[0]
    rows = []
    for c in range(m.ncells):
        # cell basis: barycenter and diameter of the (possibly displaced) full cell
        v = pts_all[cells[c]]
        hc = max(np.linalg.norm(v[i] - v[j]) for i in range(4) for j in range(i + 1, 4))
        parts = [(m.quadrature(c, 2 * cd)[0], 0)] if info["cell_loc"][c] != 2 else \
                [(m.quadrature(c, 2 * cd, where_aware=1)[0], 0), (m.quadrature(c, 2 * cd, where_aware=2)[0], 1)]
        for pts, sd in parts:
            o = (S.cell_table[c] + sd) * C_
            bx = (pts[:, 0] - bar[c, 0]) / (0.5 * hc); by = (pts[:, 1] - bar[c, 1]) / (0.5 * hc)
            phi = np.stack([bx ** (d - i) * by ** i for d in range(cd + 1) for i in range(d + 1)], axis=1)
            rows.append(np.column_stack([pts, phi @ full[o:o + C_]]))
    want = np.vstack(rows)
    assert got.shape == want.shape
    assert np.max(np.abs(got[:, :2] - want[:, :2])) < 1e-13
    assert np.max(np.abs(got[:, 2] - want[:, 2])) < 1e-11 * max(1.0, np.max(np.abs(want[:, 2])))


@pytest.mark.parametrize("k,N,cuts", [(1, 20, (0, 7, 13, 20)), (2, 16, (0, 3, 8, 16)), (0, 33, (0, 5, 17, 18, 33)), (1, 24, (0, 12, 24)), (2, 20, (0, 4, 5, 6, 20))])
def test_interface_strip_contexts_assemble_the_same_rows(k, N, cuts):
    """multi-GPU partition of the interface problem (SURVEY 8e on 8f.3): face_table is monotone in the face id, so every strip of
    cell rows owns a contiguous range of slots.  Each strip context preprocesses its window of the level set, is told the cut
    faces of the strips below it (what the ranks of a communicator all-reduce), assembles its rows — regular ones by K3,
    the cut cells' faces from sorted triplets — and reproduces the whole-mesh rows, columns, values and rhs bit for bit."""
    hdi = pb.hho_degree_info(k + 1, k); F = hdi.fbs
    whole = pb.Context(0)
    whole.mesh(pb.mesh_init_params(N, N), PB.QUAD_POLY_DUNAVANT)
    whole.cut_levelset(0, 0.35, 0.5, 0.5); whole.cut_preprocess(4)
    whole.assemble(hdi, stab_kind=PB.STAB_CUT, problem=PB.PROBLEM_INTERFACE)
    rp, col, val, rhs = whole.csr()
    ft = whole.interface_face_table()
    LW = whole.interface_strip_layout()
    assert LW["slot_begin"] == 0 and LW["slot_end"] == LW["slots"] == whole.n_rows // F
    strips = []
    for rb, re in zip(cuts[:-1], cuts[1:]):
        c = pb.Context(0)
        c.mesh(pb.mesh_init_params(N, N), quad_kind=PB.QUAD_POLY_DUNAVANT, row_begin=rb, row_end=re)
        c.cut_levelset(0, 0.35, 0.5, 0.5); c.cut_preprocess(4)
        strips.append((c, rb, re, c.interface_cut_faces_owned()))
    total = sum(s[3] for s in strips)
    assert total == LW["slots"] - pb.strip_layout(N, N, 0, N, F)["cf_end"]            # every cut face is owned by exactly one strip
    # a strip that is not told about the others refuses to number its rows
    with pytest.raises(pb.ProtonError):
        strips[1][0].assemble(hdi, stab_kind=PB.STAB_CUT, problem=PB.PROBLEM_INTERFACE)
    before = 0; next_slot = 0
    for c, rb, re, mine in strips:
        c.interface_set_cut_offsets(before, total)
        c.assemble(hdi, stab_kind=PB.STAB_CUT, problem=PB.PROBLEM_INTERFACE)
        L = c.interface_strip_layout()
        assert L["slot_begin"] == next_slot and L["slots"] == LW["slots"]
        assert L["slot_lo"] <= L["slot_begin"] <= L["slot_end"] <= L["slot_hi"]
        next_slot = L["slot_end"]
        rp_s, col_s, val_s, rhs_s = c.csr()
        r0 = L["slot_begin"] * F
        assert c.n_rows == (L["slot_end"] - L["slot_begin"]) * F
        lo, hi = rp[r0], rp[r0 + c.n_rows]
        assert np.array_equal(rp_s, rp[r0 : r0 + c.n_rows + 1] - lo)
        assert np.array_equal(col_s, col[lo:hi])
        assert np.array_equal(val_s, val[lo:hi])
        assert np.array_equal(rhs_s, rhs[r0 : r0 + c.n_rows])
        # every referenced column lies in the halo'd range the solver vectors hold
        assert col_s.min() >= L["slot_lo"] * F and col_s.max() < L["slot_hi"] * F
        # face_table of the strip = the whole one inside its window, -2 outside
        fts = c.interface_face_table()
        inside = fts != -2
        assert np.array_equal(fts[inside], ft[inside]) and inside.sum() > 0
        # second assembly reuses the structure and gives the same values
        c.assemble(hdi, stab_kind=PB.STAB_CUT, problem=PB.PROBLEM_INTERFACE)
        assert np.array_equal(c.csr()[2], val_s)
        before += mine
        c.close()
    assert next_slot == LW["slots"]
    whole.close()
