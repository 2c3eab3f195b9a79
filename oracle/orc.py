"""ORACLE — TEST INFRASTRUCTURE ONLY.

ctypes front-end of the CPU restatement in oracle/*.hpp (see orc_core.hpp for the parity statement).
Importable only from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs; proton_b200/ must never import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "build", "liborc.so")

NEG, POS, CUT, UNDEF = 0, 1, 2, 3
STAB_NAIVE, STAB_FANCY, STAB_CUT = 0, 1, 2
FUN_SINSIN, FUN_OBSTACLE, FUN_TANH = 0, 1, 2


def build(force=False):
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".hpp", ".cpp"))]
    if force or not os.path.exists(_SO) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        L.orc_last_error.restype = C.c_char_p
        for f in ("orc_mesh_new", "orc_assemble", "orc_assemble_mt", "orc_obstacle_assemble"):
            getattr(L, f).restype = C.c_void_p
        L.orc_mesh_new.argtypes = [C.c_longlong, C.c_longlong, C.c_double, C.c_double, C.c_double, C.c_double, C.c_int]
        L.orc_cell_quadrature.restype = C.c_longlong
        L.orc_obstacle_error.restype = C.c_double
        _lib = L
    return _lib


def _p(a, t=C.c_double):
    return None if a is None else a.ctypes.data_as(C.POINTER(t))


def set_finalize(kind):
    """0: restatement of Eigen's linear setFromTriplets (default); 1: the earlier std::sort-based build (timing comparison)"""
    lib().orc_set_finalize(int(kind))


def dims(cd, fd):
    out = (C.c_int * 4)()
    lib().orc_dims(cd, fd, out)
    return tuple(out)  # cbs, fbs, rbs, n


class Mesh:
    def __init__(self, Nx, Ny=None, lo=(0.0, 0.0), hi=(1.0, 1.0), poly=False):
        Ny = Nx if Ny is None else Ny
        self.h = C.c_void_p(lib().orc_mesh_new(Nx, Ny, lo[0], hi[0], lo[1], hi[1], int(poly)))
        s = (C.c_longlong * 3)()
        lib().orc_mesh_sizes(self.h, s)
        self.ncells, self.nfaces, self.npoints = int(s[0]), int(s[1]), int(s[2])
        self.Nx, self.Ny, self.poly = Nx, Ny, poly

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_mesh_free(self.h)
            self.h = None

    def points(self):
        xy = np.zeros((self.npoints, 2))
        lib().orc_mesh_points(self.h, _p(xy))
        return xy

    def cells(self):
        pt = np.zeros((self.ncells, 4), dtype=np.int64)
        fc = np.zeros((self.ncells, 4), dtype=np.int64)
        lib().orc_mesh_cells(self.h, _p(pt, C.c_longlong), _p(fc, C.c_longlong))
        return pt, fc

    def faces(self):
        pt = np.zeros((self.nfaces, 2), dtype=np.int64)
        d = np.zeros(self.nfaces, dtype=np.int32)
        lib().orc_mesh_faces(self.h, _p(pt, C.c_longlong), _p(d, C.c_int))
        return pt, d

    def barycenters(self):
        xy = np.zeros((self.ncells, 2))
        lib().orc_cell_barycenters(self.h, _p(xy))
        return xy

    def cut_preprocess(self, radius=0.35, cx=0.5, cy=0.5, refsteps=4, ls_kind=0, cut_y=0.5):
        rc = lib().orc_cut_preprocess(self.h, ls_kind, C.c_double(radius), C.c_double(cx), C.c_double(cy), C.c_double(cut_y), refsteps)
        if rc:
            raise RuntimeError(lib().orc_last_error().decode())

    def cut_info(self):
        nl = np.zeros(self.npoints, np.int32); nd = np.zeros(self.npoints, np.int32)
        fl = np.zeros(self.nfaces, np.int32); fip = np.zeros((self.nfaces, 2))
        cl = np.zeros(self.ncells, np.int32); off = np.zeros(self.ncells + 1, np.int64)
        a = (self.h, _p(nl, C.c_int), _p(nd, C.c_int), _p(fl, C.c_int), _p(fip), _p(cl, C.c_int), _p(off, C.c_longlong))
        lib().orc_cut_info(*a, None)
        pts = np.zeros((int(off[-1]), 2))
        lib().orc_cut_info(*a, _p(pts))
        return dict(node_loc=nl, node_displaced=nd, face_loc=fl, face_ip=fip, cell_loc=cl, iface_off=off, iface_pts=pts)

    def local(self, cell, cd, fd, stab=STAB_FANCY, problem=0, fun=FUN_SINSIN, rhs_di=0):
        cbs, fbs, rbs, n = dims(cd, fd)
        oper = np.zeros(rbs * n); data = np.zeros(n * n); st = np.zeros(n * n); rhs = np.zeros(cbs)
        rows = C.c_int(0)
        rc = lib().orc_local(self.h, C.c_longlong(cell), cd, fd, stab, problem, fun, rhs_di, _p(oper), C.byref(rows), _p(data), _p(st), _p(rhs))
        if rc:
            raise RuntimeError(lib().orc_last_error().decode())
        r = rows.value
        return dict(oper=oper[: r * n].reshape(n, r).T.copy(), data=data.reshape(n, n).T.copy(), stab=st.reshape(n, n).T.copy(), rhs=rhs)

    def quadrature(self, cell, degree, kind=0, where_aware=False):
        nq = lib().orc_cell_quadrature(self.h, C.c_longlong(cell), degree, kind, int(where_aware), None, None)
        if nq < 0:
            raise RuntimeError(lib().orc_last_error().decode())
        pts = np.zeros((nq, 2)); w = np.zeros(nq)
        lib().orc_cell_quadrature(self.h, C.c_longlong(cell), degree, kind, int(where_aware), _p(pts), _p(w))
        return pts, w

    @staticmethod
    def omp_enabled():
        return bool(lib().orc_omp_enabled())

    def assemble(self, cd, fd, stab=STAB_FANCY, problem=0, fun=FUN_SINSIN, rhs_di=0, threads=1):
        """threads > 1: local operators by that many OpenMP threads over the cells (the reference loop is serial)"""
        return System(self, cd, fd, stab, problem, fun, rhs_di, threads)


class System:
    """Full (uncondensed) cell+face system of the reference assembler (hho.hpp:252-463)."""

    def __init__(self, mesh, cd, fd, stab, problem, fun, rhs_di, threads=1):
        self.mesh = mesh
        tl = C.c_double(0); tg = C.c_double(0)
        if threads > 1:
            h = lib().orc_assemble_mt(mesh.h, cd, fd, stab, problem, fun, rhs_di, int(threads), C.byref(tl), C.byref(tg))
        else:
            h = lib().orc_assemble(mesh.h, cd, fd, stab, problem, fun, rhs_di, C.byref(tl), C.byref(tg))
        if not h:
            raise RuntimeError(lib().orc_last_error().decode())
        self.h = C.c_void_p(h)
        self.seconds_local, self.seconds_global = tl.value, tg.value
        s = (C.c_longlong * 2)()
        lib().orc_sys_sizes(self.h, s)
        self.n, self.nnz = int(s[0]), int(s[1])
        self.cbs, self.fbs, self.rbs, self.nloc = dims(cd, fd)
        self.rowptr = np.zeros(self.n + 1, np.int64); self.col = np.zeros(self.nnz, np.int32)
        self.val = np.zeros(self.nnz); self.rhs = np.zeros(self.n); self.compress = np.zeros(mesh.nfaces, np.int64)
        lib().orc_sys_get(self.h, _p(self.rowptr, C.c_longlong), _p(self.col, C.c_int), _p(self.val), _p(self.rhs), _p(self.compress, C.c_longlong))

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_sys_free(self.h)
            self.h = None

    def matrix(self):
        import scipy.sparse as sp
        return sp.csr_matrix((self.val, self.col, self.rowptr), shape=(self.n, self.n))

    def errors_convergence(self, sol):
        out = (C.c_double * 3)()
        sol = np.ascontiguousarray(sol, dtype=np.float64)
        if lib().orc_errors_convergence(self.h, _p(sol), out):
            raise RuntimeError(lib().orc_last_error().decode())
        return tuple(out)

    def error_fictdom(self, sol):
        out = C.c_double(0)
        sol = np.ascontiguousarray(sol, dtype=np.float64)
        if lib().orc_error_fictdom(self.h, _p(sol), C.byref(out)):
            raise RuntimeError(lib().orc_last_error().decode())
        return float(np.sqrt(out.value))


def cg(rowptr, col, val, b, tol=1e-9, max_iter=1000, precond=False):
    """Reference conjugated_gradient (solver_cg.hpp:63-144). Returns (x, exit_reason, iters, relres, seconds)."""
    n = len(b)
    rowptr = np.ascontiguousarray(rowptr, np.int64); col = np.ascontiguousarray(col, np.int32)
    val = np.ascontiguousarray(val, np.float64); b = np.ascontiguousarray(b, np.float64)
    x = np.zeros(n); it = C.c_longlong(0); rr = C.c_double(0); sec = C.c_double(0)
    e = lib().orc_cg(C.c_longlong(n), _p(rowptr, C.c_longlong), _p(col, C.c_int), _p(val), _p(b), C.c_double(tol), C.c_longlong(max_iter),
                     int(precond), _p(x), C.byref(it), C.byref(rr), C.byref(sec))
    return x, e, it.value, rr.value, sec.value


def obstacle_solve(N, k, max_outer=50, solver=None):
    """apps/obstacle/obstacle.cpp:47-218 with scipy's SuperLU standing in for Eigen::SparseLU.
    Returns (error, outer_iterations, alpha)."""
    import scipy.sparse as sp
    import scipy.sparse.linalg as spla
    m = Mesh(N, N, lo=(-1.0, -1.0), hi=(1.0, 1.0))
    fbs = k + 1
    nc = m.ncells
    alpha = np.zeros(nc + fbs * m.nfaces); beta = np.ones(nc); gamma = np.zeros(nc); c = 1.0
    it = 0
    while it < max_outer:
        diff = beta + c * (alpha[:nc] - gamma)
        in_A = (diff < 0).astype(np.uint8)
        h = C.c_void_p(lib().orc_obstacle_assemble(m.h, k, _p(in_A, C.c_ubyte), _p(gamma)))
        s = (C.c_longlong * 2)(); lib().orc_obstacle_sizes(h, s)
        n, nnz = int(s[0]), int(s[1])
        rp = np.zeros(n + 1, np.int64); col = np.zeros(nnz, np.int32); val = np.zeros(nnz); rhs = np.zeros(n)
        lib().orc_obstacle_get(h, _p(rp, C.c_longlong), _p(col, C.c_int), _p(val), _p(rhs))
        A = sp.csr_matrix((val, col, rp), shape=(n, n)).tocsc()
        sol = spla.splu(A).solve(rhs)
        alpha_prev = alpha.copy()
        lib().orc_obstacle_expand(h, _p(sol), _p(gamma), _p(alpha), _p(beta))
        lib().orc_obstacle_free(h)
        if np.linalg.norm(alpha_prev - alpha) < 1e-7:
            break
        it += 1
    err = lib().orc_obstacle_error(m.h, k, _p(alpha))
    return float(err), it, alpha


class InterfaceSystem:
    """The two-material interface problem of `cuthho_square -i` (cuthho_square.cpp:1625-1846) on a preprocessed cut mesh:
    full system (cell unknowns, then face unknowns; cut cells and cut faces carry two sets) and the driver's H1 error."""

    def __init__(self, mesh, degree, fun=FUN_SINSIN):
        lib().orc_interface_assemble.restype = C.c_void_p
        self.mesh, self.degree = mesh, degree
        self.h = C.c_void_p(lib().orc_interface_assemble(mesh.h, degree, fun))
        if not self.h:
            raise RuntimeError(lib().orc_last_error().decode())
        s = (C.c_longlong * 4)()
        lib().orc_interface_sizes(self.h, s)
        self.n, self.nnz, self.num_all_cells, self.num_all_faces = int(s[0]), int(s[1]), int(s[2]), int(s[3])
        self.rowptr = np.zeros(self.n + 1, np.int64); self.col = np.zeros(self.nnz, np.int32); self.val = np.zeros(self.nnz); self.rhs = np.zeros(self.n)
        self.cell_table = np.zeros(mesh.ncells, np.int64); self.face_table = np.zeros(mesh.nfaces, np.int64)
        lib().orc_interface_get(self.h, _p(self.rowptr, C.c_longlong), _p(self.col, C.c_int), _p(self.val), _p(self.rhs),
                                _p(self.cell_table, C.c_longlong), _p(self.face_table, C.c_longlong))

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_interface_free(self.h); self.h = None

    def matrix(self):
        import scipy.sparse as sp
        return sp.csr_matrix((self.val, self.col, self.rowptr), shape=(self.n, self.n))

    def h1_error(self, sol):
        sol = np.ascontiguousarray(sol, np.float64); out = C.c_double(0)
        if lib().orc_interface_error(self.h, _p(sol), C.byref(out)):
            raise RuntimeError(lib().orc_last_error().decode())
        return float(np.sqrt(out.value))


def interface_local(mesh, cell, degree):
    """local matrix ([cell-, cell+, faces-, faces+]) and load of a cut cell of the interface problem"""
    cbs, fbs, rbs, n = dims(degree + 1, degree)
    lc = np.zeros(4 * n * n); rhs = np.zeros(2 * cbs)
    if lib().orc_interface_local(mesh.h, C.c_longlong(cell), degree, _p(lc), _p(rhs)):
        raise RuntimeError(lib().orc_last_error().decode())
    return lc.reshape(2 * n, 2 * n).T.copy(), rhs
