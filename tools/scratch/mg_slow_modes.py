"""Scratch: which modes does the V-cycle (tests/mg_ref.py) leave to CG on the naive (1/area) stabilisation?  Dense spectrum of M^-1 A."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
import numpy as np, scipy.sparse as sp, scipy.linalg as sla
import mg_ref
from oracle import orc
k = int(sys.argv[1]); N = int(sys.argv[2]); stab = orc.STAB_NAIVE if len(sys.argv) < 4 or sys.argv[3] == "naive" else orc.STAB_FANCY
F = k + 1
m = orc.Mesh(N, N); S = m.assemble(k + 1, k, stab=stab, threads=8)
A = S.matrix().tocsr(); C = S.cbs; nT = m.ncells * C
TT = A[:nT, :nT].tocoo(); blocks = np.zeros((m.ncells, C, C)); blocks[TT.row // C, TT.row % C, TT.col % C] = TT.data
TTinv = sp.bsr_matrix((np.linalg.inv(blocks), np.arange(m.ncells), np.arange(m.ncells + 1)), shape=(nT, nT)).tocsr()
Sc = (A[nT:, nT:] - A[nT:, :nT] @ (TTinv @ A[:nT, nT:])).tocsr()
n = Sc.shape[0]
M = mg_ref.Preconditioner(m, S, Sc, k, None, nu=2)
Minv = np.column_stack([M(e) for e in np.eye(n)])
Minv = 0.5 * (Minv + Minv.T)
Ad = Sc.toarray()
w, V = sla.eigh(Ad, np.linalg.inv(Minv))      # A v = w M v
print(f"k={k} N={N}: n={n}, eig(M^-1 A) min {w[0]:.4f} max {w[-1]:.4f}, cond {w[-1]/w[0]:.1f}; smallest 8: {np.round(w[:8],4)}")
fpt, fd = m.faces(); pts = m.points(); comp = S.compress
horiz = np.zeros(n // F, bool); fx = np.zeros(n // F); fy = np.zeros(n // F)
for f in range(m.nfaces):
    if fd[f]: continue
    p0, p1 = fpt[f]; horiz[comp[f]] = abs(p1 - p0) == 1; c = 0.5 * (pts[p0] + pts[p1]); fx[comp[f]] = c[0]; fy[comp[f]] = c[1]
for e in range(4):
    v = V[:, e].reshape(-1, F)
    en = (v ** 2).sum(axis=0) / (v ** 2).sum()
    print(f" mode {e} (lambda {w[e]:.4f}): energy per face dof {np.round(en,3)}, horizontal share {((v[horiz]**2).sum()/(v**2).sum()):.3f}")
    # dof-0 field of the horizontal faces as a grid (rows j = 1..N-1, cols i = 0..N-1)
    h0 = np.zeros((N - 1, N)); 
    for q in np.flatnonzero(horiz):
        i = int(round(fx[q] * N - 0.5)); j = int(round(fy[q] * N)) - 1; h0[j, i] = v[q, F - 1 if len(sys.argv) > 4 else 0]
    np.set_printoptions(linewidth=250, precision=2, suppress=True)
    s = h0 / np.abs(h0).max()
    w0 = int(os.environ.get('WIN', '0'))
    print(s[:: max(1, N // 8), :: max(1, N // 8)] if not w0 else s[N//2-w0//2:N//2+w0//2, N//2-w0//2:N//2+w0//2])
