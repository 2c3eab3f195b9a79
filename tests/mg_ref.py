"""TEST INFRASTRUCTURE — scipy restatement of the multigrid preconditioner of proton_b200/csrc/pb_mg.cuh
(apply_preconditioner = 3): the same operators written as sparse matrices (P^T A P Galerkin products, Kronecker
interpolation), against which the tests hold the CUDA path (pb_mg_level / pb_mg_apply) and the host build of its index
logic (tests/mg_host_check.cu).  Uses the CPU oracle for the condensed system; never imported by the product."""
import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spla

from oracle import orc

CHEB_LO, OMEGA = 0.15, 0.8


def condensed_system(N, k, cut=False, fun=orc.FUN_SINSIN, Ny=None, radius=0.35, threads=1, stab=None):
    """(mesh, full System, condensed matrix, condensed rhs): Schur complement of the oracle's full cell+face system"""
    Ny = N if Ny is None else Ny
    m = orc.Mesh(N, Ny, poly=cut)
    if cut:
        m.cut_preprocess(radius=radius, refsteps=4)
        S = m.assemble(k + 1, k, stab=orc.STAB_CUT, problem=1, fun=fun, threads=threads)
    elif stab is not None:
        S = m.assemble(k + 1, k, stab=stab, fun=fun, threads=threads)
    else:
        S = m.assemble(k + 1, k, fun=fun, threads=threads)
    A = S.matrix().tocsr(); b = S.rhs
    C = S.cbs; nc = m.ncells; nT = nc * C
    TT = A[:nT, :nT].tocoo()
    blocks = np.zeros((nc, C, C))
    blocks[TT.row // C, TT.row % C, TT.col % C] = TT.data
    inv = np.linalg.inv(blocks)
    TTinv = sp.bsr_matrix((inv, np.arange(nc), np.arange(nc + 1)), shape=(nT, nT)).tocsr()
    TF = A[:nT, nT:]; FT = A[nT:, :nT]; FF = A[nT:, nT:]
    Sc = (FF - FT @ (TTinv @ TF)).tocsr()
    g = b[nT:] - FT @ (TTinv @ b[:nT])
    return m, S, Sc, g


def cell_schur_blocks(m, k, cut=False, fun=orc.FUN_SINSIN):
    """per-cell condensed blocks S_T (4F x 4F) in the device layout [cell][col][row] (column-major)"""
    cbs, fbs, rbs, n = orc.dims(k + 1, k)
    out = np.zeros((m.ncells, 4 * fbs, 4 * fbs))
    for c in range(m.ncells):
        r = m.local(c, k + 1, k, stab=orc.STAB_CUT, problem=1, fun=fun) if cut else m.local(c, k + 1, k, fun=fun)
        lc = r["data"] + r["stab"]
        TT, TF, FT, FF = lc[:cbs, :cbs], lc[:cbs, cbs:], lc[cbs:, :cbs], lc[cbs:, cbs:]
        if np.abs(TT).max() == 0.0:          # cells the reference never touches do not occur; keep the guard explicit
            continue
        S = FF - FT @ np.linalg.solve(TT, TF)
        out[c] = S.T
    return out


def face_prolongation(m, compress, k, info=None):
    """P: vertex values (ALL (Nx+1)(Ny+1) vertices, boundary columns kept) -> face unknowns (compressed numbering)"""
    F = k + 1
    fpt, fd = m.faces(); pts = m.points()
    rows, cols, vals = [], [], []
    for f in range(m.nfaces):
        if fd[f]:
            continue
        cf = compress[f]; p0, p1 = fpt[f]
        w = {p0: [0.5, -0.5], p1: [0.5, 0.5]}
        if info is not None and info["face_loc"][f] == orc.CUT:
            ip = info["face_ip"][f]
            t = np.linalg.norm(ip - pts[p0]) / np.linalg.norm(pts[p1] - pts[p0]); xi = 2 * t - 1
            if info["node_loc"][p0] == orc.NEG:
                w = {p0: [0.5 if F == 1 else xi / (xi + 1), -1 / (xi + 1)]}
            else:
                w = {p1: [0.5 if F == 1 else -xi / (1 - xi), 1 / (1 - xi)]}
        elif info is not None:               # vertices outside the domain carry no coarse unknown
            w = {p: c for p, c in w.items() if info["node_loc"][p] == orc.NEG}
        for p, (c0, c1) in w.items():
            rows.append(cf * F); cols.append(p); vals.append(c0)
            if F > 1:
                rows.append(cf * F + 1); cols.append(p); vals.append(c1)
    nf = (m.nfaces - int(fd.sum())) * F
    return sp.csr_matrix((vals, (rows, cols)), shape=(nf, m.npoints))


def interior_mask(Nx, Ny):
    j, i = np.divmod(np.arange((Nx + 1) * (Ny + 1)), Nx + 1)
    return (i > 0) & (i < Nx) & (j > 0) & (j < Ny)


def coarse_cells(N):
    return (N + 1) // 2 if N > 2 else N


def interp1d(N):
    """(N+1) x (Nc+1) linear interpolation; coarse vertex j sits at fine min(2j, N); identity when N <= 2"""
    if N <= 2:
        return sp.identity(N + 1, format="csr")
    Nc = (N + 1) // 2
    cpos = [min(2 * j, N) for j in range(Nc + 1)]
    P = sp.lil_matrix((N + 1, Nc + 1))
    for j, c in enumerate(cpos):
        P[c, j] = 1.0
    for i in range(N + 1):
        if i not in cpos:
            j = (i - 1) // 2
            P[i, j] = 0.5; P[i, j + 1] = 0.5
    return P.tocsr()


def interp2d(Nx, Ny):
    """all fine vertices x all coarse vertices, rows/columns of boundary vertices zeroed"""
    P = sp.kron(interp1d(Ny), interp1d(Nx)).tocsr()
    mf = interior_mask(Nx, Ny).astype(float); mc = interior_mask(coarse_cells(Nx), coarse_cells(Ny)).astype(float)
    return (sp.diags(mf) @ P @ sp.diags(mc)).tocsr()


class Hierarchy:
    """vertex levels: A_l on all vertices (boundary rows/cols zero), interpolation P_l between them"""

    def __init__(self, A0_all, Nx, Ny):
        m0 = sp.diags(interior_mask(Nx, Ny).astype(float))
        self.levels = [dict(A=(m0 @ A0_all @ m0).tocsr(), Nx=Nx, Ny=Ny)]
        while True:
            L = self.levels[-1]
            if L["Nx"] <= 2 and L["Ny"] <= 2:
                break
            P = interp2d(L["Nx"], L["Ny"]); L["P"] = P
            self.levels.append(dict(A=(P.T @ L["A"] @ P).tocsr(), Nx=coarse_cells(L["Nx"]), Ny=coarse_cells(L["Ny"])))
        for i, L in enumerate(self.levels):
            d = L["A"].diagonal()
            om = 1.0 if i == len(self.levels) - 1 else OMEGA
            L["idiag"] = np.where(d > 0, om / np.where(d > 0, d, 1.0), 0.0) * interior_mask(L["Nx"], L["Ny"])

    def stencils(self, l):
        """[vertex][9] like pb_mg_level (boundary vertices: identity)"""
        L = self.levels[l]; Nx, Ny = L["Nx"], L["Ny"]; A = L["A"].tocsr(); W = Nx + 1
        st = np.zeros(((Nx + 1) * (Ny + 1), 9))
        inner = interior_mask(Nx, Ny)
        for v in range(st.shape[0]):
            if not inner[v]:
                st[v, 4] = 1.0; continue
            for dj in (-1, 0, 1):
                for di in (-1, 0, 1):
                    st[v, (dj + 1) * 3 + di + 1] = A[v, v + dj * W + di]
        return st

    def vcycle(self, b, l=0):
        L = self.levels[l]
        x = L["idiag"] * b
        if l == len(self.levels) - 1:
            return x
        r = b - L["A"] @ x
        x = x + L["P"] @ self.vcycle(L["P"].T @ r, l + 1)
        return x + L["idiag"] * (b - L["A"] @ x)


def block_diag_inverse(Sc, F):
    n = Sc.shape[0]; nb = n // F
    B = np.zeros((nb, F, F)); idx = np.arange(nb) * F
    for a in range(F):
        for c in range(F):
            B[:, a, c] = np.asarray(Sc[idx + a, idx + c]).ravel()
    Binv = np.linalg.inv(B)
    return lambda r: np.einsum("bac,bc->ba", Binv, r.reshape(nb, F)).ravel()


def line_block_inverse(m, compress, Sc, F, row_cuts=None):
    """r -> L^-1 r, L = the block-tridiagonal parts of Sc along the mesh lines (pb_mg_lines.cuh): columns of horizontal faces (cut at the
    strip boundaries `row_cuts`, as the ranks of a communicator solve them) and rows of vertical faces"""
    fpt, fd = m.faces(); N1 = m.Nx + 1
    line = np.zeros(Sc.shape[0] // F, np.int64)
    for f in range(m.nfaces):
        if fd[f]:
            continue
        p0, p1 = sorted(fpt[f]); i0, j0 = p0 % N1, p0 // N1
        if p1 - p0 == 1:                                       # horizontal face h(i0, j0): owned by the strip of cell row j0
            strip = 0 if row_cuts is None else int(np.searchsorted(row_cuts, j0, side="right"))
            line[compress[f]] = i0 + N1 * strip
        else:
            line[compress[f]] = -1 - j0
    ld = np.repeat(line, F); A = Sc.tocoo(); keep = ld[A.row] == ld[A.col]
    lu = spla.splu(sp.csc_matrix((A.data[keep], (A.row[keep], A.col[keep])), shape=Sc.shape))
    return lu.solve


def power_lmax(Sc, Dinv, its=14):
    """the device's estimate: fixed hashed start vector, `its` steps of v <- D^-1 A v / |.|"""
    n = Sc.shape[0]
    h = (np.arange(n, dtype=np.uint64) * np.uint64(0x9E3779B97F4A7C15))
    h ^= h >> np.uint64(29); h *= np.uint64(0xBF58476D1CE4E5B9); h ^= h >> np.uint64(32)
    v = (h & np.uint64(0xFFFFF)).astype(np.float64) / 1048576.0 + 0.25
    v /= np.linalg.norm(v); lam = 0.0
    for _ in range(its):
        w = Dinv(Sc @ v); lam = np.linalg.norm(w); v = w / lam
    return 1.1 * lam


class Preconditioner:
    """z = M^-1 r exactly as mg_apply composes it"""

    def __init__(self, m, S, Sc, k, info=None, nu=2, lmax=None, lines=False, row_cuts=None):
        self.A = Sc.tocsr(); self.F = k + 1; self.nu = nu
        self.P = face_prolongation(m, S.compress, k, info)
        self.h = Hierarchy(self.P.T @ self.A @ self.P, m.Nx, m.Ny)
        self.mask = interior_mask(m.Nx, m.Ny).astype(float)
        self.Dinv = line_block_inverse(m, S.compress, self.A, self.F, row_cuts) if lines else block_diag_inverse(self.A, self.F)
        self.lmax = power_lmax(self.A, self.Dinv) if lmax is None else lmax

    def _cheb(self, rhs, x):
        lmax = self.lmax; lmin = CHEB_LO * lmax
        theta, delta = 0.5 * (lmax + lmin), 0.5 * (lmax - lmin); sigma = theta / delta; rho = 1.0 / sigma
        res = rhs.copy() if x is None else rhs - self.A @ x
        d = self.Dinv(res) / theta
        x = d.copy() if x is None else x + d
        for _ in range(1, self.nu):
            res = res - self.A @ d
            rn = 1.0 / (2 * sigma - rho)
            d = rn * rho * d + (2 * rn / delta) * self.Dinv(res)
            x = x + d; rho = rn
        return x, res - self.A @ d

    def __call__(self, r):
        z, res = self._cheb(r, None)
        b0 = self.mask * (self.P.T @ res)
        z = z + self.P @ self.h.vcycle(b0)
        z, _ = self._cheb(r, z)
        return z


def pcg(A, b, M, tol=1e-12, maxit=1000):
    """conjugated_gradient (solver_cg.hpp:63-144) with z = M(r); returns (x, iterations, relres)"""
    x = np.zeros_like(b); r = b.copy(); z = M(r); d = z.copy(); rho = r @ z; nr0 = np.linalg.norm(r); it = 0
    while True:
        y = A @ d; alpha = rho / (d @ y); x += alpha * d; r -= alpha * y
        rel = np.linalg.norm(r) / nr0
        if rel < tol or it > maxit:
            return x, it, rel
        z = M(r); rn = r @ z; d = z + (rn / rho) * d; rho = rn; it += 1
