"""Scratch (CPU, scipy): the fictitious-domain problem decouples into the inside and the outside of the interface; the vertex
hierarchy of the V-cycle couples them again on its coarse levels (bilinear interpolation across the interface).  Variant: vertices
outside the domain are inactive (like boundary vertices) on every level.   python ... k N1 N2 ..."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
import numpy as np, scipy.sparse as sp
import mg_ref
from oracle import orc


class MaskedHierarchy(mg_ref.Hierarchy):
    def __init__(self, A0_all, Nx, Ny, active0):
        act = active0 & mg_ref.interior_mask(Nx, Ny)
        m0 = sp.diags(act.astype(float))
        self.levels = [dict(A=(m0 @ A0_all @ m0).tocsr(), Nx=Nx, Ny=Ny, act=act)]
        while True:
            L = self.levels[-1]
            if L["Nx"] <= 2 and L["Ny"] <= 2:
                break
            nx, ny = L["Nx"], L["Ny"]; ncx, ncy = mg_ref.coarse_cells(nx), mg_ref.coarse_cells(ny)
            cj, ci = np.divmod(np.arange((ncx + 1) * (ncy + 1)), ncx + 1)
            fi = np.minimum(2 * ci, nx) if nx > 2 else ci; fj = np.minimum(2 * cj, ny) if ny > 2 else cj
            actc = L["act"][fj * (nx + 1) + fi] & mg_ref.interior_mask(ncx, ncy)
            P = sp.kron(mg_ref.interp1d(ny), mg_ref.interp1d(nx)).tocsr()
            P = (sp.diags(L["act"].astype(float)) @ P @ sp.diags(actc.astype(float))).tocsr()
            L["P"] = P
            self.levels.append(dict(A=(P.T @ L["A"] @ P).tocsr(), Nx=ncx, Ny=ncy, act=actc))
        for i, L in enumerate(self.levels):
            d = L["A"].diagonal(); om = 1.0 if i == len(self.levels) - 1 else mg_ref.OMEGA
            L["idiag"] = np.where(d > 0, om / np.where(d > 0, d, 1.0), 0.0) * L["act"]


k = int(sys.argv[1])
for N in [int(a) for a in sys.argv[2:]]:
    m, S, Sc, g = mg_ref.condensed_system(N, k, cut=True, threads=8); info = m.cut_info()
    out = []
    for nu in (1, 2):
        M = mg_ref.Preconditioner(m, S, Sc, k, info, nu=nu)
        x0, it0, _ = mg_ref.pcg(Sc, g, M, tol=1e-9, maxit=2000)
        inside = info["node_loc"] == orc.NEG
        M.h = MaskedHierarchy(M.P.T @ M.A @ M.P, m.Nx, m.Ny, inside); M.mask = (inside & mg_ref.interior_mask(m.Nx, m.Ny)).astype(float)
        x1, it1, _ = mg_ref.pcg(Sc, g, M, tol=1e-9, maxit=2000)
        if nu == 1:
            import scipy.sparse.linalg as spla
            F = k + 1; n = Sc.shape[0]; fpt, fd = m.faces(); line = np.zeros(n // F, np.int64)
            for f in range(m.nfaces):
                if fd[f]: continue
                p0, p1 = fpt[f]; hz = abs(p1 - p0) == 1; i0, j0 = min(p0, p1) % (N + 1), min(p0, p1) // (N + 1)
                line[S.compress[f]] = i0 if hz else (N + 1) + j0
            ld = np.repeat(line, F); Ac = Sc.tocoo(); keep = ld[Ac.row] == ld[Ac.col]
            Llu = spla.splu(sp.csc_matrix((Ac.data[keep], (Ac.row[keep], Ac.col[keep])), shape=(n, n)))
        M.Dinv = Llu.solve; M.lmax = mg_ref.power_lmax(M.A, M.Dinv, its=20)
        x2, it2, _ = mg_ref.pcg(Sc, g, M, tol=1e-9, maxit=2000)
        out.append(f"   + line smoother {it2} its")
        # implicit variant: only the face prolongation drops the outside vertices; the vertex hierarchy is the standard one
        M3 = mg_ref.Preconditioner(m, S, Sc, k, info, nu=nu)
        M3.P = (M3.P @ sp.diags(inside.astype(float))).tocsr()
        M3.h = mg_ref.Hierarchy(M3.P.T @ M3.A @ M3.P, m.Nx, m.Ny)
        x3, it3, _ = mg_ref.pcg(Sc, g, M3, tol=1e-9, maxit=2000)
        M3.Dinv = Llu.solve; M3.lmax = mg_ref.power_lmax(M3.A, M3.Dinv, its=20)
        x4, it4, _ = mg_ref.pcg(Sc, g, M3, tol=1e-9, maxit=2000)
        out.append(f"   implicit (P only) {it3} its, + line smoother {it4} its")
        # lines of horizontal faces cut at the boundaries of W strips of cell rows (what a rank can solve without communication)
        for W in (2, 8):
            line2 = line.copy(); fpt2, fd2 = m.faces()
            for f in range(m.nfaces):
                if fd2[f]: continue
                p0, p1 = fpt2[f]
                if abs(p1 - p0) == 1:
                    j0 = min(p0, p1) // (N + 1); line2[S.compress[f]] = line[S.compress[f]] + (2 * N + 2) * (1 + (j0 * W) // N)
            ld2 = np.repeat(line2, k + 1); keep2 = ld2[Ac.row] == ld2[Ac.col]
            L2 = spla.splu(sp.csc_matrix((Ac.data[keep2], (Ac.row[keep2], Ac.col[keep2])), shape=(n, n)))
            M3.Dinv = L2.solve; M3.lmax = mg_ref.power_lmax(M3.A, M3.Dinv, its=20)
            x5, it5, _ = mg_ref.pcg(Sc, g, M3, tol=1e-9, maxit=2000)
            out.append(f"   lines cut into {W} strips: {it5} its")
        out.append(f"nu={nu}: all vertices {it0} its, inside vertices only {it1} its (solutions differ by {np.linalg.norm(x1-x0)/np.linalg.norm(x0):.1e})")
    print(f"k={k} cut N={N}: " + "; ".join(out), flush=True)
