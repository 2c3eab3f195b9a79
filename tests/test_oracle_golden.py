"""Pins the CPU oracle (oracle/) against every known-answer table the reference ships for the path
(SURVEY §8c): apps/obstacle/results/convergence.txt and the "F.D." block of apps/cuthho/cuthho.xlsx,
extracted verbatim into tests/golden/reference_golden.json by tests/golden/make_reference_golden.py.
The tables carry 6 significant digits, so the bar is 5e-6 relative."""
import numpy as np
import pytest
import scipy.sparse.linalg as spla

from oracle import orc

TOL = 5e-6


@pytest.mark.parametrize("N", [8, 16, 32, 64])
@pytest.mark.parametrize("k", [0, 1])
def test_obstacle_table(golden, N, k):
    err, iters, _ = orc.obstacle_solve(N, k)
    ref = golden["obstacle"][str(N)][f"k{k}"]
    assert abs(err - ref) / ref < TOL
    assert iters < 49  # converged, not capped (obstacle.cpp:119)


def _fictdom(N, k, r):
    m = orc.Mesh(N, poly=True)
    m.cut_preprocess(refsteps=r)
    S = m.assemble(k + 1, k, stab=orc.STAB_CUT, problem=1)
    sol = spla.splu(S.matrix().tocsc()).solve(S.rhs)
    return S.error_fictdom(sol), S, m


@pytest.mark.parametrize("k,N,r", [(k, N, 4) for k in (0, 1, 2) for N in (10, 20, 40)] +
                         [(1, 20, 5), (1, 20, 6), (1, 20, 7), (2, 10, 5), (2, 10, 7), (2, 20, 7), (0, 10, 6), (0, 20, 5)])
def test_cuthho_fictdom_table(golden, k, N, r):
    err, _, _ = _fictdom(N, k, r)
    ref = golden["cuthho_fictdom"][f"k{k}_N{N}_r{r}"]
    assert abs(err - ref) / ref < TOL


@pytest.mark.parametrize("k,N,r", [(k, N, 4) for k in (0, 1, 2) for N in (10, 20, 40)] + [(1, 10, 5), (2, 10, 6), (1, 20, 7)])
def test_cuthho_interface_table(golden, k, N, r):
    """`cuthho_square -i` (two-material interface problem; oracle/orc_interface.hpp): the "Interface" block of
    cuthho.xlsx.  The driver solves with CG; a direct solve of the same system gives the same 6 digits.  1e-5: the k=2 column
    of the table was produced with CG at the default 1e-9 tolerance and carries one digit less."""
    m = orc.Mesh(N, poly=True)
    m.cut_preprocess(refsteps=r)
    S = orc.InterfaceSystem(m, k)
    A = S.matrix()
    assert abs(A - A.T).max() < 1e-10 * abs(A).max()
    # doubled unknowns: two sets on every cut cell and cut face (cuthho_square.cpp:1143-1161)
    info = m.cut_info()
    ncut_cells = int((info["cell_loc"] == 2).sum()); ncut_faces = int((info["face_loc"] == 2).sum())
    assert S.num_all_cells == m.ncells + ncut_cells and S.num_all_faces == m.nfaces + ncut_faces
    sol = spla.splu(A.tocsc()).solve(S.rhs)
    err = S.h1_error(sol)
    ref = golden["cuthho_interface"][f"k{k}_N{N}_r{r}"]
    assert abs(err - ref) / ref < 1e-5, (err, ref)


def test_config1_anchor():
    """BASELINE config 1: cuthho_square -f -M 32 -N 32 -k 1 (survey-derived anchor)."""
    err, S, m = _fictdom(32, 1, 4)
    ci = m.cut_info()
    assert S.n == 10112
    assert int((ci["cell_loc"] == orc.CUT).sum()) == 92
    assert int(ci["node_displaced"].sum()) == 48
    assert abs(err - 1.134467728e-3) / 1.134467728e-3 < 1e-8


def test_dunavant_off_by_one_and_empty_rule():
    """degree d uses rule d+1; degree 8 yields zero points."""
    m = orc.Mesh(4, poly=True)
    counts = [len(m.quadrature(5, d)[1]) for d in range(0, 9)]
    assert counts == [4 * c for c in (3, 3, 4, 6, 7, 12, 13, 16, 0)]
    for d in range(1, 8):
        assert abs(m.quadrature(5, d)[1].sum() - 1.0 / 16) < 1e-15


def test_convergence_rates_direct():
    """convergence_test.cpp:312-325 expectation for HHO(k+1,k): L2 order k+2, energy order k+1."""
    for k in range(0, 3):
        errs = []
        for N in (4, 8, 16):
            m = orc.Mesh(N)
            S = m.assemble(k + 1, k)
            sol = spla.splu(S.matrix().tocsc()).solve(S.rhs)
            errs.append(np.sqrt(np.array(S.errors_convergence(sol))))
        r_l2 = np.log2(errs[1][0] / errs[2][0]); r_en = np.log2(errs[1][2] / errs[2][2])
        assert abs(r_l2 - (k + 2)) < 0.15 and abs(r_en - (k + 1)) < 0.1


def test_numbering_closed_form():
    """faces sorted by (ptid0, ptid1); cell 5 of the 4x4 mesh has faces {11,14,20,12}."""
    m = orc.Mesh(4)
    assert m.cells()[1][5].tolist() == [11, 14, 20, 12]
    Nx, Ny = 5, 3
    m = orc.Mesh(Nx, Ny)
    _, fids = m.cells()
    hf = lambda i, j: j * (2 * Nx + 1) + 2 * i if j < Ny else Ny * (2 * Nx + 1) + i
    vf = lambda i, j: j * (2 * Nx + 1) + 2 * i + 1 if i < Nx else j * (2 * Nx + 1) + 2 * Nx
    for j in range(Ny):
        for i in range(Nx):
            assert fids[j * Nx + i].tolist() == [hf(i, j), vf(i + 1, j), hf(i, j + 1), vf(i, j)]
