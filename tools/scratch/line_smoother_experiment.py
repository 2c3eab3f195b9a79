"""Scratch experiment (CPU, scipy): does a LINE-block Jacobi smoother (lines of horizontal faces along y, lines of vertical faces
along x) inside the same Chebyshev / vertex-space V-cycle make the iteration count of the fictitious-domain problem (1/area-scaled
stabilisation) independent of N?   python tools/scratch/line_smoother_experiment.py k cut|naive N1 N2 ..."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
import numpy as np, scipy.sparse as sp, scipy.sparse.linalg as spla
import mg_ref
from oracle import orc

k = int(sys.argv[1]); mode = sys.argv[2]; Ns = [int(a) for a in sys.argv[3:]]
F = k + 1
for N in Ns:
    if mode == "cut":
        m, S, Sc, g = mg_ref.condensed_system(N, k, cut=True, threads=8); info = m.cut_info()
    else:
        m = orc.Mesh(N, N); S = m.assemble(k + 1, k, stab=orc.STAB_NAIVE, threads=8); info = None
        A = S.matrix().tocsr(); b = S.rhs; C = S.cbs; nT = m.ncells * C
        TT = A[:nT, :nT].tocoo(); blocks = np.zeros((m.ncells, C, C)); blocks[TT.row // C, TT.row % C, TT.col % C] = TT.data
        TTinv = sp.bsr_matrix((np.linalg.inv(blocks), np.arange(m.ncells), np.arange(m.ncells + 1)), shape=(nT, nT)).tocsr()
        Sc = (A[nT:, nT:] - A[nT:, :nT] @ (TTinv @ A[:nT, nT:])).tocsr(); g = b[nT:] - A[nT:, :nT] @ (TTinv @ b[:nT])
    n = Sc.shape[0]
    # line id of every compressed face: horizontal face (i, j) -> line i; vertical face (i, j) -> line Nx + j
    fpt, fd = m.faces(); pts = m.points(); comp = S.compress
    line = np.zeros(n // F, np.int64)
    for f in range(m.nfaces):
        if fd[f]: continue
        p0, p1 = fpt[f]; horiz = abs(pts[p0][1] - pts[p1][1]) < 1e-12 * 1 if info is None else (abs(p1 - p0) == 1)
        i0, j0 = min(p0, p1) % (N + 1), min(p0, p1) // (N + 1)
        line[comp[f]] = (j0 if horiz else (N + 1) + i0) if os.environ.get("SWAP") else (i0 if horiz else (N + 1) + j0)
    ldof = np.repeat(line, F)
    Ac = Sc.tocoo(); keep = ldof[Ac.row] == ldof[Ac.col]
    L = sp.csc_matrix((Ac.data[keep], (Ac.row[keep], Ac.col[keep])), shape=(n, n))
    Llu = spla.splu(L)
    res = {}
    for name, Dinv in (("point-block", None), ("line-block", Llu.solve)):
        for nu in (1, 2):
            M = mg_ref.Preconditioner(m, S, Sc, k, info, nu=nu)
            if Dinv is not None:
                M.Dinv = Dinv; M.lmax = mg_ref.power_lmax(M.A, M.Dinv, its=20)
            x, it, rel = mg_ref.pcg(Sc, g, M, tol=1e-9, maxit=2000)
            res[(name, nu)] = (it, M.lmax)
    print(f"k={k} {mode} N={N}: " + "  ".join(f"{a} nu={b}: {v[0]} its (lmax {v[1]:.2f})" for (a, b), v in res.items()), flush=True)
