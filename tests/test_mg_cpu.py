"""The index logic of the multigrid preconditioner (proton_b200/csrc/pb_mg.cuh: transfer operators, element matrices,
Galerkin products, stencils) compiled for the HOST by tests/mg_host_check.cu and checked against scipy's P^T A P on the
oracle's condensed system.  No GPU needed; the same functions run inside the CUDA kernels."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest
import scipy.sparse as sp

from oracle import orc
import mg_ref

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def H():
    src = os.path.join(HERE, "mg_host_check.cu"); out = os.path.join(HERE, "build", "libmg_host_check.so")
    hdr = os.path.join(HERE, "..", "proton_b200", "csrc", "pb_mg.cuh")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    hdr2 = os.path.join(HERE, "..", "proton_b200", "csrc", "pb_mg_lines.cuh")
    if not os.path.exists(out) or max(os.path.getmtime(src), os.path.getmtime(hdr), os.path.getmtime(hdr2)) > os.path.getmtime(out):
        subprocess.check_call(["nvcc", "-O1", "-std=c++17", "--expt-relaxed-constexpr", "-Wno-deprecated-gpu-targets", "-Xcompiler", "-fPIC", "-shared", "-o", out, src])
    return C.CDLL(out)


def p(a, t=C.c_double):
    return None if a is None else a.ctypes.data_as(C.POINTER(t))


def cut_args(m, info):
    if info is None:
        return [None, None, None, None], []
    keep = [np.ascontiguousarray(info["node_loc"], np.int32), np.ascontiguousarray(info["face_loc"], np.int32),
            np.ascontiguousarray(info["face_ip"], np.float64), np.ascontiguousarray(m.points(), np.float64)]
    return [p(keep[0], C.c_int), p(keep[1], C.c_int), p(keep[2]), p(keep[3])], keep


def assemble_nodal(E, Nx, Ny):
    """all-vertex matrix from 4x4 element matrices (corners (i,j) (i+1,j) (i+1,j+1) (i,j+1))"""
    rows, cols, vals = [], [], []
    for c in range(Nx * Ny):
        cj, ci = divmod(c, Nx); p0 = cj * (Nx + 1) + ci
        pc = [p0, p0 + 1, p0 + Nx + 2, p0 + Nx + 1]
        for a in range(4):
            for b in range(4):
                rows.append(pc[a]); cols.append(pc[b]); vals.append(E[c, a * 4 + b])
    n = (Nx + 1) * (Ny + 1)
    return sp.csr_matrix((vals, (rows, cols)), shape=(n, n))


def drop_boundary_couplings(st, Nx, Ny):
    """the device keeps the stencil entries that point at boundary vertices (they multiply zeros); the restatement drops them"""
    st = st.copy(); W = Nx + 1; inner = mg_ref.interior_mask(Nx, Ny)
    for v in np.flatnonzero(inner):
        for dj in (-1, 0, 1):
            for di in (-1, 0, 1):
                if not inner[v + dj * W + di]:
                    st[v, (dj + 1) * 3 + di + 1] = 0.0
    return st


CASES = [(6, 5, 1, False), (7, 7, 2, False), (5, 8, 0, False), (12, 12, 1, True), (10, 10, 2, True), (10, 10, 0, True)]


@pytest.mark.parametrize("Nx,Ny,k,cut", CASES)
def test_level0_operator_is_PtAP(H, Nx, Ny, k, cut):
    m, S, Sc, g = mg_ref.condensed_system(Nx, k, cut=cut, Ny=Ny)
    info = m.cut_info() if cut else None
    F = k + 1
    P = mg_ref.face_prolongation(m, S.compress, k, info)
    A0 = (P.T @ Sc @ P).toarray()
    blocks = np.ascontiguousarray(mg_ref.cell_schur_blocks(m, k, cut=cut))
    ca, keep = cut_args(m, info)
    E = np.zeros((m.ncells, 16))
    H.mgh_elements(Nx, Ny, F, p(blocks), *ca, p(E))
    A0h = assemble_nodal(E, Nx, Ny).toarray()
    inner = mg_ref.interior_mask(Nx, Ny)
    scale = np.abs(A0).max()
    assert np.abs(A0h[np.ix_(inner, inner)] - A0[np.ix_(inner, inner)]).max() < 1e-10 * scale
    # stencils gathered from the element matrices = rows of the same matrix
    st = np.zeros(((Nx + 1) * (Ny + 1), 9))
    H.mgh_stencils(Nx, Ny, 0, Ny + 1, p(E), p(st))
    ref = mg_ref.Hierarchy(sp.csr_matrix(A0), Nx, Ny).stencils(0)
    assert np.abs(drop_boundary_couplings(st, Nx, Ny) - ref).max() < 1e-10 * scale
    # face <-> vertex transfers are P and P^T
    rng = np.random.default_rng(1)
    x0 = rng.standard_normal((Nx + 1) * (Ny + 1)) * inner
    z = np.zeros(Sc.shape[0]); H.mgh_face_prolong(Nx, Ny, F, p(x0), *ca, p(z))
    assert np.abs(z - P @ x0).max() < 1e-13
    r = rng.standard_normal(Sc.shape[0]); b0 = np.zeros_like(x0)
    H.mgh_face_restrict(Nx, Ny, F, p(r), *ca, p(b0))
    assert np.abs(b0 - inner * (P.T @ r)).max() < 1e-12


@pytest.mark.parametrize("Nx,Ny", [(8, 8), (7, 5), (9, 3), (12, 2), (3, 3), (16, 11)])
def test_coarse_levels_are_galerkin(H, Nx, Ny):
    rng = np.random.default_rng(Nx * 100 + Ny)
    # any symmetric element matrices will do for the algebra
    E = rng.standard_normal((Nx * Ny, 4, 4)); E = (E + E.transpose(0, 2, 1)).reshape(Nx * Ny, 16).copy()
    A = assemble_nodal(E, Nx, Ny)
    h = mg_ref.Hierarchy(A, Nx, Ny)
    Ef, nx, ny = E, Nx, Ny
    for l in range(1, len(h.levels)):
        d = (C.c_int * 2)(); H.mgh_coarse_dims(nx, ny, d); ncx, ncy = d[0], d[1]
        assert (ncx, ncy) == (h.levels[l]["Nx"], h.levels[l]["Ny"])
        Ec = np.zeros((ncx * ncy, 16)); H.mgh_galerkin(nx, ny, p(Ef), 0, ny, p(Ec))
        # strips: partial products over disjoint fine cell rows add up to the whole (the all-reduce of level 1)
        parts = np.zeros_like(Ec); cutrow = ny // 2
        for lo, hi in ((0, cutrow), (cutrow, ny)):
            t = np.zeros_like(Ec); H.mgh_galerkin(nx, ny, p(Ef), lo, hi, p(t)); parts += t
        assert np.abs(parts - Ec).max() < 1e-12 * np.abs(Ec).max()
        inner = mg_ref.interior_mask(ncx, ncy)
        got = assemble_nodal(Ec, ncx, ncy).toarray()[np.ix_(inner, inner)]
        ref = h.levels[l]["A"].toarray()[np.ix_(inner, inner)]
        assert np.abs(got - ref).max() < 1e-11 * max(1.0, np.abs(ref).max()), f"level {l}"
        # transfers
        Pm = h.levels[l - 1]["P"]
        xc = rng.standard_normal((ncx + 1) * (ncy + 1)) * inner
        xf = np.zeros((nx + 1) * (ny + 1)); H.mgh_prolong(nx, ny, p(xc), p(xf))
        assert np.abs(xf - Pm @ xc).max() < 1e-13
        rf = rng.standard_normal((nx + 1) * (ny + 1)) * mg_ref.interior_mask(nx, ny)
        rc = np.zeros_like(xc); H.mgh_restrict(nx, ny, 0, ny + 1, p(rf), p(rc))
        assert np.abs(rc - Pm.T @ rf).max() < 1e-12
        acc = np.zeros_like(rc)
        for lo, hi in ((0, cutrow), (cutrow, ny + 1)):
            t = np.zeros_like(rc); H.mgh_restrict(nx, ny, lo, hi, p(rf), p(t)); acc += t
        assert np.abs(acc - rc).max() < 1e-12
        Ef, nx, ny = Ec, ncx, ncy


def test_strip_stencils_match_whole(H):
    Nx, Ny = 6, 9
    rng = np.random.default_rng(3)
    E = rng.standard_normal((Nx * Ny, 16))
    whole = np.zeros(((Nx + 1) * (Ny + 1), 9)); H.mgh_stencils(Nx, Ny, 0, Ny + 1, p(E), p(whole))
    for rb, re in ((0, 4), (4, 7), (7, 9)):
        crow0 = max(rb - 1, 0)
        st = np.zeros(((re - rb) * (Nx + 1), 9))
        H.mgh_stencils(Nx, Ny, rb, re, p(np.ascontiguousarray(E[crow0 * Nx:re * Nx])), p(st))
        assert np.array_equal(st, whole[rb * (Nx + 1):re * (Nx + 1)])


@pytest.mark.parametrize("k,cut,N", [(1, False, 16), (2, False, 16), (1, True, 20)])
def test_restated_preconditioner_gives_flat_iteration_counts(k, cut, N):
    """the method itself (scipy restatement): PCG with the V-cycle reaches 1e-12 in a few iterations, same solution as Jacobi-PCG"""
    its = []
    for n in (N, 2 * N):
        m, S, Sc, g = mg_ref.condensed_system(n, k, cut=cut, fun=orc.FUN_SINSIN if cut else orc.FUN_TANH, threads=4)
        M = mg_ref.Preconditioner(m, S, Sc, k, m.cut_info() if cut else None)
        x, it, rel = mg_ref.pcg(Sc, g, M)
        d = Sc.diagonal(); xj, itj, _ = mg_ref.pcg(Sc, g, lambda r: r / d, maxit=20000)
        assert np.linalg.norm(x - xj) <= 1e-9 * np.linalg.norm(xj)
        its.append(it)
        assert it < itj / 3
    assert its[1] <= its[0] + (8 if cut else 2), its


@pytest.mark.parametrize("Nx,Ny,rb,re", [(6, 5, 0, 5), (7, 9, 0, 4), (7, 9, 4, 7), (7, 9, 7, 9), (5, 6, 2, 3), (3, 3, 0, 1), (300, 40, 10, 30), (2, 2, 0, 2)])
def test_mesh_lines_cover_the_owned_faces_once(H, Nx, Ny, rb, re):
    """the line sets of the line-block Jacobi smoother (pb_mg_lines.cuh): every owned free face lies on exactly one line, at the step that
    the closed-form numbering says: columns of horizontal faces h(i, j) (line i, step j - j0), rows of vertical faces v(i, j) (line j - rb, step i - 1)"""
    per = 2 * Nx - 1
    row_start = lambda j: 0 if j == 0 else (Nx - 1) + (j - 1) * per
    cf_begin = row_start(rb); cf_end = row_start(re) if re < Ny else (Nx - 1) + (Ny - 1) * per
    n = cf_end - cf_begin
    owner = np.full(n, -1, np.int64); hits = np.zeros(n, np.int32); dims = np.zeros(6, np.int32)
    H.mgh_line_sets(Nx, Ny, rb, re, p(owner, C.c_longlong), p(hits, C.c_int), p(dims, C.c_int))
    assert np.all(hits == 1)
    j0 = max(rb, 1)
    assert tuple(dims[:4]) == (Nx, max(re - j0, 0), re - rb, Nx - 1)
    for j in range(rb, re):
        for i in range(Nx):
            if j >= 1:        # h(i, j): compress_h
                br = row_start(j) + (0 if i == 0 else 2 * i - 1) - cf_begin
                assert owner[br] == (0 << 40) + (i << 20) + (j - j0)
            if i >= 1:        # v(i, j): compress_v
                br = (i - 1 if j == 0 else row_start(j) + 2 * i) - cf_begin
                assert owner[br] == (1 << 40) + ((j - rb) << 20) + (i - 1)
    # few long lines go to the warp-per-line solve
    assert dims[5] == (1 if (re - rb < 2048 and Nx - 1 >= 256) else 0)


def test_restated_line_smoother_removes_the_growth_on_the_area_scaled_stabilisation():
    """the method of pb_mg_lines.cuh on the scipy restatement, no GPU: with the naive stabilisation (scaled by 1/area) the
    face-block smoother leaves modes that are constant along a mesh line and alternate from line to line, and the PCG count grows with N;
    block-tridiagonal solves along the lines remove them.  Same solution either way."""
    point, lines = [], []
    for N in (32, 64, 128):
        m, S, Sc, g = mg_ref.condensed_system(N, 1, stab=orc.STAB_NAIVE, fun=orc.FUN_TANH, threads=4)
        xs = []
        for use_lines, out in ((False, point), (True, lines)):
            M = mg_ref.Preconditioner(m, S, Sc, 1, None, nu=1, lines=use_lines)
            x, it, rel = mg_ref.pcg(Sc, g, M, tol=1e-10, maxit=500)
            out.append(it); xs.append(x)
        assert np.linalg.norm(xs[0] - xs[1]) <= 1e-7 * np.linalg.norm(xs[1])
    assert lines[2] <= lines[0] + 2, (point, lines)
    assert point[2] >= point[0] + 4 and lines[2] < point[2], (point, lines)
