"""Scratch: slow modes of M^-1 A on the CUT (fictitious-domain) problem, V-cycle of tests/mg_ref.py with the line-block smoother:
where do they live (distance of the face to the interface, cut faces or not, inside or outside)?   python ... k N"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
import numpy as np, scipy.sparse as sp, scipy.linalg as sla, scipy.sparse.linalg as spla
import mg_ref
from oracle import orc
k = int(sys.argv[1]); N = int(sys.argv[2]); F = k + 1
m, S, Sc, g = mg_ref.condensed_system(N, k, cut=True, threads=8); info = m.cut_info()
n = Sc.shape[0]
fpt, fd = m.faces(); pts = m.points(); comp = S.compress
line = np.zeros(n // F, np.int64); fc = np.zeros((n // F, 2)); fcut = np.zeros(n // F, bool)
for f in range(m.nfaces):
    if fd[f]: continue
    p0, p1 = fpt[f]; horiz = abs(p1 - p0) == 1
    i0, j0 = min(p0, p1) % (N + 1), min(p0, p1) // (N + 1)
    line[comp[f]] = i0 if horiz else (N + 1) + j0
    fc[comp[f]] = 0.5 * (pts[p0] + pts[p1]); fcut[comp[f]] = info["face_loc"][f] == orc.CUT
ldof = np.repeat(line, F)
Ac = Sc.tocoo(); keep = ldof[Ac.row] == ldof[Ac.col]
Llu = spla.splu(sp.csc_matrix((Ac.data[keep], (Ac.row[keep], Ac.col[keep])), shape=(n, n)))
M = mg_ref.Preconditioner(m, S, Sc, k, info, nu=1)
if "point" not in sys.argv:
    M.Dinv = Llu.solve; M.lmax = mg_ref.power_lmax(M.A, M.Dinv, its=20)
Minv = np.column_stack([M(e) for e in np.eye(n)]); Minv = 0.5 * (Minv + Minv.T)
w, V = sla.eigh(Sc.toarray(), np.linalg.inv(Minv))
print(f"k={k} N={N} cut: n={n}, eig(M^-1 A) min {w[0]:.4f} max {w[-1]:.4f}; smallest 12: {np.round(w[:12],4)}; below 0.5: {(w<0.5).sum()}, below 0.25: {(w<0.25).sum()}")
dist = np.abs(np.hypot(fc[:, 0] - 0.5, fc[:, 1] - 0.5) - 0.35) * N        # distance to the circle in cell widths
inside = np.hypot(fc[:, 0] - 0.5, fc[:, 1] - 0.5) < 0.35
for e in range(6):
    v2 = (V[:, e].reshape(-1, F) ** 2); tot = v2.sum(); pf = v2.sum(axis=1)
    print(f" mode {e} lambda {w[e]:.4f}: dof shares {np.round(v2.sum(axis=0)/tot,3)}, on cut faces {pf[fcut].sum()/tot:.3f}, within 1.5 cells of the interface {pf[dist<1.5].sum()/tot:.3f}, "
          f"inside {pf[inside].sum()/tot:.3f}, faces holding 90%: {(np.cumsum(np.sort(pf)[::-1]) < 0.9*tot).sum()+1}")
np.set_printoptions(linewidth=250, precision=2, suppress=True)
for e in (0, 1, 3):
    v = V[:, e].reshape(-1, F)[:, 0]
    G = np.zeros((N + 1, N))
    for q in range(n // F):
        if line[q] <= N and abs(fc[q, 1] * N - round(fc[q, 1] * N)) < 1e-9:
            G[int(round(fc[q, 1] * N)), int(fc[q, 0] * N)] = v[q]
    G /= np.abs(G).max()
    print(f"mode {e}: dof 0 of the horizontal faces, every 3rd row / column"); print(G[::3, ::3])
