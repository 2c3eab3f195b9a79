"""ctypes binding of include/proton_b200.h.  Names follow the reference's host API where one exists
(mesh_init_params: basic_mesh.hpp:178-197; hho_degree_info: utils.hpp:62-111; cg_params: solver_cg.hpp:44-61)."""
import ctypes as C
import os
import re

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)


class ProtonError(RuntimeError):
    pass


class PB:
    OK, ERR_INVALID, ERR_CUDA, ERR_STATE, ERR_UNSUPPORTED = range(5)
    QUAD_TENSOR_GAUSS, QUAD_POLY_DUNAVANT = 0, 1
    STAB_NAIVE, STAB_FANCY, STAB_CUT = 0, 1, 2
    PROBLEM_HHO, PROBLEM_FICTDOM, PROBLEM_INTERFACE = 0, 1, 2
    FUN_SINSIN, FUN_OBSTACLE, FUN_TANH = 0, 1, 2
    IN_NEGATIVE_SIDE, IN_POSITIVE_SIDE, ON_INTERFACE, UNDEF = 0, 1, 2, 3
    CG_CONVERGED, CG_DIVERGED, CG_MAX_ITER_REACHED = 0, 1, 2
    LOCAL_LC, LOCAL_OPER, LOCAL_RHS, LOCAL_SCHUR, LOCAL_SCHUR_RHS, LOCAL_REC_A, LOCAL_REC_B, LOCAL_DATA = range(8)


class _AssemblyParams(C.Structure):
    _fields_ = [("cell_degree", C.c_int), ("face_degree", C.c_int), ("stab_kind", C.c_int), ("problem", C.c_int), ("fun", C.c_int),
                ("rhs_degree_increase", C.c_int), ("eta", C.c_double), ("keep_local", C.c_int)]


class _CgParams(C.Structure):
    _fields_ = [("convergence_threshold", C.c_double), ("divergence_threshold", C.c_double), ("max_iter", C.c_longlong),
                ("apply_preconditioner", C.c_int)]


class _CgResult(C.Structure):
    _fields_ = [("iterations", C.c_longlong), ("relative_residual", C.c_double), ("exit_reason", C.c_int), ("seconds", C.c_double)]


def lib_path():
    return os.path.join(_HERE, "libproton_b200.so")


def abi_symbols():
    """Entry points declared in include/proton_b200.h (parsed from the header)."""
    txt = open(os.path.join(_ROOT, "include", "proton_b200.h")).read()
    return sorted(set(re.findall(r"\b(pb_[a-z0-9_]+)\s*\(", txt)))


_lib = None


def lib():
    """Load libproton_b200.so.  Fails loudly if the CUDA extension has not been built: there is no fallback."""
    global _lib
    if _lib is None:
        p = lib_path()
        if not os.path.exists(p):
            raise ProtonError(f"{p} is missing: build it with `make -C proton_b200/csrc` (or __graft_entry__.build()); "
                              "proton_b200 has no CPU fallback")
        L = C.CDLL(p)
        L.pb_last_error.restype = C.c_char_p
        L.pb_last_error.argtypes = [C.c_void_p]
        L.pb_mesh_cartesian.argtypes = [C.c_void_p, C.c_longlong, C.c_longlong, C.c_double, C.c_double, C.c_double, C.c_double, C.c_int,
                                        C.c_longlong, C.c_longlong]
        L.pb_get_local.argtypes = [C.c_void_p, C.c_int, C.c_longlong, C.c_longlong, C.c_void_p]
        L.pb_assemble_pipelined.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.pb_pipeline_sync.argtypes = [C.c_void_p]
        L.pb_interface_get_local.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.pb_interface_face_table.argtypes = [C.c_void_p, C.c_void_p]
        L.pb_interface_cut_faces_owned.argtypes = [C.c_void_p, C.c_void_p]
        L.pb_interface_set_cut_offsets.argtypes = [C.c_void_p, C.c_longlong, C.c_longlong]
        L.pb_interface_strip_layout.argtypes = [C.c_void_p, C.c_void_p]
        L.pb_pcg_history.argtypes = [C.c_void_p, C.c_longlong]
        L.pb_pcg_get_history.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong, C.c_void_p]
        L.pb_sample_solution.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong, C.c_void_p]
        L.pb_static_scratch_bytes.restype = C.c_longlong
        L.pb_mg_info.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.pb_mg_level.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.pb_mg_apply.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def strip_layout(Nx, Ny, row_begin, row_end, fbs):
    """pb_strip_layout: host arithmetic of the 1-D strip partition (no GPU needed)."""
    out = (C.c_longlong * 8)()
    rc = lib().pb_strip_layout(C.c_longlong(Nx), C.c_longlong(Ny), C.c_longlong(row_begin), C.c_longlong(row_end), int(fbs), out)
    if rc != PB.OK:
        raise ProtonError(f"pb_strip_layout: invalid arguments ({Nx}, {Ny}, {row_begin}, {row_end}, {fbs})")
    keys = ("cf_begin", "cf_end", "col_lo", "col_hi", "cell_row0", "n_rows", "halo_lo", "halo_hi")
    return dict(zip(keys, [int(v) for v in out]))


def strip_rows(Ny, world, rank):
    """cell rows [begin, end) of `rank`: as even as possible, lower ranks take the remainder"""
    base, rem = divmod(Ny, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def dist_unique_id():
    buf = (C.c_ubyte * 128)()
    rc = lib().pb_dist_unique_id(buf)
    if rc != PB.OK:
        raise ProtonError(lib().pb_last_error(None).decode())
    return bytes(buf)


class mesh_init_params:
    def __init__(self, Nx=4, Ny=4, min_x=0.0, max_x=1.0, min_y=0.0, max_y=1.0):
        self.Nx, self.Ny, self.min_x, self.max_x, self.min_y, self.max_y = Nx, Ny, min_x, max_x, min_y, max_y


class hho_degree_info:
    def __init__(self, cd, fd=None):
        if fd is None:
            cd, fd = cd, cd
        c1 = fd > 0 and cd in (fd - 1, fd, fd + 1)
        c2 = fd == 0 and cd in (fd, fd + 1)
        if not (c1 or c2):
            cd = fd
        self.cd, self.fd, self.rd = cd, fd, fd + 1

    def cell_degree(self): return self.cd
    def face_degree(self): return self.fd
    def reconstruction_degree(self): return self.rd
    @property
    def cbs(self): return (self.cd + 1) * (self.cd + 2) // 2
    @property
    def fbs(self): return self.fd + 1
    @property
    def rbs(self): return (self.rd + 1) * (self.rd + 2) // 2


class cg_params:
    def __init__(self, convergence_threshold=1e-9, divergence_threshold=100.0, max_iter=1000, apply_preconditioner=False):
        self.convergence_threshold, self.divergence_threshold = convergence_threshold, divergence_threshold
        self.max_iter, self.apply_preconditioner = max_iter, apply_preconditioner


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Context:
    """One pb_ctx: a Cartesian mesh strip on one CUDA device and the condensed system assembled on it."""

    def __init__(self, device=0):
        self._h = C.c_void_p()
        rc = lib().pb_ctx_create(int(device), C.byref(self._h))
        if rc != PB.OK:
            raise ProtonError(lib().pb_last_error(None).decode())

    def close(self):
        if getattr(self, "_h", None):
            lib().pb_ctx_destroy(self._h)
            self._h = None

    __del__ = close

    def _ck(self, rc):
        if rc != PB.OK:
            raise ProtonError(f"[{rc}] " + lib().pb_last_error(self._h).decode())

    # ---- mesh
    def mesh(self, mip, quad_kind=PB.QUAD_TENSOR_GAUSS, row_begin=0, row_end=None):
        row_end = mip.Ny if row_end is None else row_end
        self._ck(lib().pb_mesh_cartesian(self._h, mip.Nx, mip.Ny, mip.min_x, mip.max_x, mip.min_y, mip.max_y, quad_kind, row_begin, row_end))
        self.mip, self.quad_kind = mip, quad_kind
        self.npoints = (mip.Nx + 1) * (mip.Ny + 1)
        self.nfaces = 2 * mip.Nx * mip.Ny + mip.Nx + mip.Ny
        self.ncells = mip.Nx * mip.Ny
        return self

    def dist_init(self, rank, world, unique_id):
        buf = (C.c_ubyte * 128).from_buffer_copy(unique_id)
        self._ck(lib().pb_dist_init(self._h, int(rank), int(world), buf))
        self.rank, self.world = rank, world

    def recover_strip(self, row_begin, row_end):
        """(cell dofs of the owned cell rows, owned face dofs).  Interface problem: the cell part is interface_assembler's packed layout of
        the owned cells (a cut cell carries its negative, then its positive set); the ranks' parts concatenate to the vector of recover()"""
        ncell = (row_end - row_begin) * self.mip.Nx
        if getattr(self, "problem", None) == PB.PROBLEM_INTERFACE:
            ncell += self.cut_counts_owned()[1]
        cells = np.zeros(ncell * self.hdi.cbs); faces = np.zeros(self.n_rows)
        self._ck(lib().pb_recover_strip(self._h, _ptr(cells), _ptr(faces)))
        return cells, faces

    def points(self):
        xy = np.zeros((self.npoints, 2)); self._ck(lib().pb_mesh_get_points(self._h, _ptr(xy))); return xy

    def set_points(self, xy):
        xy = np.ascontiguousarray(xy, np.float64); assert xy.shape == (self.npoints, 2)
        self._ck(lib().pb_mesh_set_points(self._h, _ptr(xy)))

    def cell_faces(self):
        f = np.zeros((self.ncells, 4), np.int64); self._ck(lib().pb_mesh_cell_faces(self._h, _ptr(f))); return f

    def compress_table(self):
        t = np.zeros(self.nfaces, np.int64); self._ck(lib().pb_mesh_compress_table(self._h, _ptr(t))); return t

    def cut_upload(self, node_loc, face_loc, face_ip, cell_loc, iface_off, iface_pts):
        a = [np.ascontiguousarray(node_loc, np.int32), np.ascontiguousarray(face_loc, np.int32), np.ascontiguousarray(face_ip, np.float64),
             np.ascontiguousarray(cell_loc, np.int32), np.ascontiguousarray(iface_off, np.int64), np.ascontiguousarray(iface_pts, np.float64)]
        self._ck(lib().pb_cut_upload(self._h, *[_ptr(x) for x in a]))

    def cut_levelset(self, kind=0, radius=0.35, cx=0.5, cy=0.5, cut_y=0.5):
        self._ck(lib().pb_cut_levelset(self._h, int(kind), C.c_double(radius), C.c_double(cx), C.c_double(cy), C.c_double(cut_y)))

    def cut_preprocess(self, refine_steps=4):
        self._ck(lib().pb_cut_preprocess(self._h, int(refine_steps)))

    def cut_download(self):
        nl = np.zeros(self.npoints, np.int32); fl = np.zeros(self.nfaces, np.int32); fip = np.zeros((self.nfaces, 2))
        cl = np.zeros(self.ncells, np.int32); ncut = C.c_longlong(); ni = C.c_int()
        self._ck(lib().pb_cut_download(self._h, _ptr(nl), _ptr(fl), _ptr(fip), _ptr(cl), C.byref(ncut), C.byref(ni), None))
        pts = np.zeros((ncut.value, ni.value, 2))
        self._ck(lib().pb_cut_download(self._h, None, None, None, None, None, None, _ptr(pts)))
        return dict(node_loc=nl, face_loc=fl, face_ip=fip, cell_loc=cl, iface_pts=pts)

    def cut_counts_owned(self):
        """(cut cells this context holds, ghost row included; cut cells of its OWNED rows)"""
        a = C.c_longlong(); b = C.c_longlong()
        self._ck(lib().pb_cut_counts(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def cut_counts(self):
        """(number of cut cells, interface points per cut cell) of the last preprocessing"""
        ncut = C.c_longlong(); ni = C.c_int()
        self._ck(lib().pb_cut_download(self._h, None, None, None, None, C.byref(ncut), C.byref(ni), None))
        return ncut.value, ni.value

    # ---- stages 1+2
    def assemble(self, hdi, stab_kind=PB.STAB_FANCY, problem=PB.PROBLEM_HHO, fun=PB.FUN_SINSIN, rhs_degree_increase=0, eta=5.0, keep_local=False):
        p = _AssemblyParams(hdi.cd, hdi.fd, stab_kind, problem, fun, rhs_degree_increase, eta, int(keep_local))
        self._ck(lib().pb_assemble(self._h, C.byref(p)))
        self.hdi = hdi; self.problem = problem
        n = C.c_longlong(); nnz = C.c_longlong(); nc = C.c_longlong(); nl = C.c_longlong()
        self._ck(lib().pb_system_sizes(self._h, C.byref(n), C.byref(nnz), C.byref(nc), C.byref(nl)))
        self.n_rows, self.nnz, self.ncells_local, self.n_local = n.value, nnz.value, nc.value, nl.value
        return self

    def assemble_pipelined(self, hdi, host_points_ptr=None, host_rhs_ptr=None, stab_kind=PB.STAB_FANCY, problem=PB.PROBLEM_HHO, fun=PB.FUN_SINSIN,
                           rhs_degree_increase=0, eta=5.0):
        """pb_assemble_pipelined: enqueue one step (upload of the vertices, stages 1+2, read-back of the rhs); page-locked
        host pointers as integers.  Call pipeline_sync() before using the results."""
        p = _AssemblyParams(hdi.cd, hdi.fd, stab_kind, problem, fun, rhs_degree_increase, eta, 0)
        self._ck(lib().pb_assemble_pipelined(self._h, C.byref(p), C.c_void_p(host_points_ptr), C.c_void_p(host_rhs_ptr)))
        self.hdi = hdi
        return self

    def strip_point_range(self):
        a = C.c_longlong(); b = C.c_longlong()
        self._ck(lib().pb_strip_point_range(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def pipeline_sync(self):
        self._ck(lib().pb_pipeline_sync(self._h))
        return self

    def csr(self, values=True):
        rp = np.zeros(self.n_rows + 1, np.int64); col = np.zeros(self.nnz, np.int32)
        val = np.zeros(self.nnz) if values else None; rhs = np.zeros(self.n_rows)
        self._ck(lib().pb_get_csr(self._h, _ptr(rp), _ptr(col), _ptr(val), _ptr(rhs)))
        return rp, col, val, rhs

    def local(self, what, cell_begin=0, count=None):
        count = self.ncells_local - cell_begin if count is None else count
        h = self.hdi; nf = 4 * h.fbs; n = h.cbs + nf
        shape = {PB.LOCAL_LC: (n, n), PB.LOCAL_OPER: (n, h.rbs), PB.LOCAL_RHS: (h.cbs,), PB.LOCAL_SCHUR: (nf, nf), PB.LOCAL_SCHUR_RHS: (nf,),
                 PB.LOCAL_REC_A: (nf, h.cbs), PB.LOCAL_REC_B: (h.cbs,), PB.LOCAL_DATA: (n, n)}[what]
        out = np.zeros((count,) + shape)
        self._ck(lib().pb_get_local(self._h, what, cell_begin, count, _ptr(out)))
        # device layout is column-major: out[c, col, row]; return [c, row, col]
        return out.transpose(0, 2, 1).copy() if len(shape) == 2 else out

    def timings(self):
        a = C.c_double(); b = C.c_double(); c = C.c_double()
        self._ck(lib().pb_last_timings(self._h, C.byref(a), C.byref(b), C.byref(c)))
        return dict(ms_local=a.value, ms_scatter=b.value, ms_pcg=c.value)

    # ---- stage 3
    def conjugated_gradient(self, cgp=None):
        cgp = cgp or cg_params()
        p = _CgParams(cgp.convergence_threshold, cgp.divergence_threshold, int(cgp.max_iter), int(cgp.apply_preconditioner))
        r = _CgResult()
        self._ck(lib().pb_pcg(self._h, C.byref(p), C.byref(r)))
        return dict(iterations=r.iterations, relative_residual=r.relative_residual, exit_reason=r.exit_reason, seconds=r.seconds)

    # ---- multigrid preconditioner (apply_preconditioner = 3)
    def mg_info(self):
        """builds the hierarchy for the current system if needed: (vertex levels, upper end of the smoothing range, setup ms)"""
        nl = C.c_int(); lm = C.c_double(); ms = C.c_double()
        self._ck(lib().pb_mg_info(self._h, C.byref(nl), C.byref(lm), C.byref(ms)))
        return nl.value, lm.value, ms.value

    def mg_level(self, level, stencil=True, elements=False):
        d = (C.c_longlong * 6)()
        self._ck(lib().pb_mg_level(self._h, int(level), d, None, None))
        Nx, Ny, row0, nrows, crow0, ncrows = [int(v) for v in d]
        st = np.zeros((9, nrows * (Nx + 1))) if stencil else None
        el = np.zeros((ncrows * Nx, 16)) if elements else None
        self._ck(lib().pb_mg_level(self._h, int(level), d, _ptr(st), _ptr(el)))
        return dict(Nx=Nx, Ny=Ny, row0=row0, nrows=nrows, crow0=crow0, ncrows=ncrows, stencil=None if st is None else st.T.copy(), elements=el)

    def mg_apply(self, r):
        r = np.ascontiguousarray(r, np.float64); assert r.shape == (self.n_rows,)
        z = np.zeros(self.n_rows)
        self._ck(lib().pb_mg_apply(self._h, _ptr(r), _ptr(z)))
        return z

    # ---- interface problem
    def interface_local(self, what):
        """per-cut-cell arrays of the interface problem, ascending cell order (see pb_interface_get_local); matrices are
        returned as [cell, row, col]"""
        h = self.hdi; nf = 4 * h.fbs; n2 = 2 * h.cbs + 2 * nf
        shape = {0: (2 * nf, 2 * nf), 1: (2 * nf,), 2: (n2, n2), 3: (2 * h.cbs,)}[what]
        ncut = self.cut_counts()[0]
        out = np.zeros((ncut,) + shape)
        self._ck(lib().pb_interface_get_local(self._h, what, _ptr(out)))
        return out.transpose(0, 2, 1).copy() if len(shape) == 2 else out

    def interface_face_table(self):
        t = np.zeros(self.nfaces, np.int64)
        self._ck(lib().pb_interface_face_table(self._h, _ptr(t)))
        return t

    def interface_cut_faces_owned(self):
        """cut faces among the faces this (strip) context owns, after the cut preprocessing"""
        k = C.c_longlong(0)
        self._ck(lib().pb_interface_cut_faces_owned(self._h, C.byref(k)))
        return k.value

    def interface_set_cut_offsets(self, before, total):
        """strip contexts without a communicator: cut faces owned by the strips below this one / by all strips"""
        self._ck(lib().pb_interface_set_cut_offsets(self._h, int(before), int(total))); return self

    def interface_strip_layout(self):
        o = (C.c_longlong * 7)()
        self._ck(lib().pb_interface_strip_layout(self._h, o))
        return dict(slot_begin=o[0], slot_end=o[1], slot_lo=o[2], slot_hi=o[3], slots=o[4], halo_prev=o[5], halo_next=o[6])

    # ---- driver outputs
    def pcg_history(self, capacity):
        """record the histfile lines of the following conjugated_gradient calls (0 switches it off)"""
        self._ck(lib().pb_pcg_history(self._h, C.c_longlong(int(capacity)))); return self

    def history(self):
        n = C.c_longlong()
        self._ck(lib().pb_pcg_get_history(self._h, None, C.c_longlong(0), C.byref(n)))
        out = np.zeros(n.value)
        if n.value:
            self._ck(lib().pb_pcg_get_history(self._h, _ptr(out), C.c_longlong(n.value), C.byref(n)))
        return out

    def sample_solution(self):
        """rows (x, y, u_T) of the driver's solution_N_*_k_*.dat for the owned cells"""
        n = C.c_longlong()
        self._ck(lib().pb_sample_solution(self._h, None, C.c_longlong(0), C.byref(n)))
        out = np.zeros((n.value, 3))
        self._ck(lib().pb_sample_solution(self._h, _ptr(out), C.c_longlong(n.value), C.byref(n)))
        return out

    def face_solution(self):
        x = np.zeros(self.n_rows); self._ck(lib().pb_get_face_solution(self._h, _ptr(x))); return x

    def set_face_solution(self, x):
        x = np.ascontiguousarray(x, np.float64); assert x.shape == (self.n_rows,)
        self._ck(lib().pb_set_face_solution(self._h, _ptr(x)))

    def recover(self):
        ncell_sets = self.ncells + (self.cut_counts()[0] if getattr(self, "problem", 0) == PB.PROBLEM_INTERFACE else 0)   # cut cells carry two sets
        full = np.zeros(ncell_sets * self.hdi.cbs + self.n_rows); self._ck(lib().pb_recover(self._h, _ptr(full))); return full

    def error_norms(self):
        out = np.zeros(3); self._ck(lib().pb_error_norms(self._h, _ptr(out))); return out

    def error_norms_static(self):
        """pb_error_norms_static: the same sums without any device allocation (scratch in a 16 MiB static device array)"""
        out = np.zeros(3); self._ck(lib().pb_error_norms_static(self._h, _ptr(out))); return out

    # ---- obstacle problem (apps/obstacle/obstacle.cpp)
    def obstacle_begin(self, face_degree, rhs_degree_increase=1):
        self._ck(lib().pb_obstacle_begin(self._h, int(face_degree), int(rhs_degree_increase)))
        self.hdi = hho_degree_info(0, face_degree)
        n = C.c_longlong(); nnz = C.c_longlong(); nc = C.c_longlong(); nl = C.c_longlong()
        self._ck(lib().pb_system_sizes(self._h, C.byref(n), C.byref(nnz), C.byref(nc), C.byref(nl)))
        self.n_rows, self.nnz, self.ncells_local, self.n_local = n.value, nnz.value, nc.value, nl.value

    def obstacle_iterate(self, cgp):
        p = _CgParams(cgp.convergence_threshold, cgp.divergence_threshold, int(cgp.max_iter), int(cgp.apply_preconditioner))
        d = C.c_double(); na = C.c_longlong(); it = C.c_longlong()
        self._ck(lib().pb_obstacle_iterate(self._h, C.byref(p), C.byref(d), C.byref(na), C.byref(it)))
        return d.value, na.value, it.value

    def obstacle_error(self):
        e = C.c_double(); self._ck(lib().pb_obstacle_error(self._h, C.byref(e))); return float(np.sqrt(e.value))

    def obstacle_state(self):
        a = np.zeros(self.ncells_local); b = np.zeros(self.ncells_local); ia = np.zeros(self.ncells_local, np.uint8)
        self._ck(lib().pb_obstacle_get(self._h, _ptr(a), _ptr(b), _ptr(ia))); return a, b, ia

    def obstacle_solve(self, face_degree, max_outer=50, tol=1e-7, cgp=None):
        """the driver's loop (obstacle.cpp:119-198): returns (error, outer iterations, log)"""
        cgp = cgp or cg_params(1e-13, 1e300, 100000, True)
        self.obstacle_begin(face_degree, 1)
        it, log = 0, []
        while it < max_outer:
            d, na, cgit = self.obstacle_iterate(cgp)
            log.append((d, na, cgit))
            if d < tol:
                break
            it += 1
        return self.obstacle_error(), it, log

    def probe_fp64_peak(self):
        t = C.c_double(); self._ck(lib().pb_probe_fp64_peak(self._h, C.byref(t))); return t.value
