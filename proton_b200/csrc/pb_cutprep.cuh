// cutHHO mesh preprocessing on the device: detect_node_position, detect_cut_faces, move_nodes,
// detect_cut_cells, refine_interface of src/methods/cuthho_bits/cuthho_geom.hpp:68-161, 275-340, 466-543, 609-673,
// called in the order of apps/cuthho/cuthho_square.cpp:2030-2052 (node-displacement branch).
// All decisions (sign tests, bisection paths) must follow the host arithmetic of the reference, which is compiled
// without FMA contraction: the level set and the bisection use explicit round-to-nearest add/mul intrinsics.
#pragma once
#include "pb_device.cuh"

namespace pb {

// circle_level_set / line_level_set (cuthho_square.cpp:56-124)
struct LevelSet {
    int kind;                       // 0 circle, 1 horizontal line
    double radius, cx, cy, cut_y;
};
PB_D double ls_eval(const LevelSet& L, double x, double y) {
    if (L.kind == 0) {
        const double dx = __dsub_rn(x, L.cx), dy = __dsub_rn(y, L.cy);
        return __dsub_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(L.radius, L.radius));
    }
    return __dsub_rn(y, L.cut_y);
}
// normal = gradient / |gradient| (cuthho_square.cpp:79-92)
PB_D void ls_normal(const LevelSet& L, double x, double y, double& nx, double& ny) {
    if (L.kind == 0) {
        const double gx = __dsub_rn(__dmul_rn(2.0, x), __dmul_rn(2.0, L.cx)), gy = __dsub_rn(__dmul_rn(2.0, y), __dmul_rn(2.0, L.cy));
        const double nrm = sqrt(__dadd_rn(__dmul_rn(gx, gx), __dmul_rn(gy, gy)));
        nx = gx / nrm; ny = gy / nrm;
    } else { nx = 0.0; ny = 1.0; }
}
PB_D double norm2d(double x, double y) { return sqrt(__dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y))); }

// find_zero_crossing (cuthho_geom.hpp:68-116): bisection with `while (dist > threshold && max_iter--)`
PB_D double2 find_zero_crossing(const LevelSet& L, double2 p0, double2 p1, double threshold) {
    double2 pa = p0, pb_ = p1;
    double2 pm = make_double2(__dadd_rn(pa.x, pb_.x) / 2.0, __dadd_rn(pa.y, pb_.y) / 2.0), pm_prev = pm;
    unsigned long long max_iter = 30;
    while (true) {
        const double lb = ls_eval(L, pb_.x, pb_.y), lm = ls_eval(L, pm.x, pm.y);
        pm_prev = pm;
        if ((lb >= 0 && lm >= 0) || (lb < 0 && lm < 0)) pb_ = pm; else pa = pm;
        pm = make_double2(__dadd_rn(pa.x, pb_.x) / 2.0, __dadd_rn(pa.y, pb_.y) / 2.0);
        const double ex = __dsub_rn(pm_prev.x, pm.x), ey = __dsub_rn(pm_prev.y, pm.y);
        const double dist = sqrt(__dadd_rn(__dmul_rn(ex, ex), __dmul_rn(ey, ey)));
        if (!(dist > threshold)) break;
        if (max_iter-- == 0) break;
    }
    return pm;
}

// face id -> its two points in sorted order (closed form of the sorted face array)
PB_HD void face_nodes(const Grid& g, long long f, long long& p0, long long& p1) {
    const long long per = 2 * g.Nx + 1, row = f / per, rem = f - row * per;
    if (row == g.Ny) { p0 = grid_point(g, rem, g.Ny); p1 = p0 + 1; return; }
    if (rem == 2 * g.Nx) { p0 = grid_point(g, g.Nx, row); p1 = p0 + (g.Nx + 1); return; }
    if ((rem & 1) == 0) { p0 = grid_point(g, rem / 2, row); p1 = p0 + 1; }
    else { p0 = grid_point(g, (rem - 1) / 2, row); p1 = p0 + (g.Nx + 1); }
}

enum { CUTERR_TOO_MANY_CUTS = 1, CUTERR_BAD_CUT_COUNT = 2, CUTERR_CONCAVE = 4, CUTERR_IFACE_NOT_FOUND = 8 };

// detect_node_position (cuthho_geom.hpp:118-130): strict < 0 is the negative side
// (all kernels of this file walk an index range [begin, end) of GLOBAL point / face / cell ids: the rows a strip context needs)
__global__ void cut_node_position_kernel(LevelSet L, long long begin, long long npoints, const double2* __restrict__ pts, int* __restrict__ node_loc) {
    for (long long p = begin + blockIdx.x * (long long)blockDim.x + threadIdx.x; p < npoints; p += (long long)gridDim.x * blockDim.x)
        node_loc[p] = ls_eval(L, pts[p].x, pts[p].y) < 0 ? PB_IN_NEGATIVE_SIDE : PB_IN_POSITIVE_SIDE;
}

// detect_cut_faces (cuthho_geom.hpp:132-161)
__global__ void cut_faces_kernel(LevelSet L, Grid g, long long begin, long long nfaces, const double2* __restrict__ pts, int* __restrict__ face_loc,
                                 double2* __restrict__ face_ip) {
    for (long long f = begin + blockIdx.x * (long long)blockDim.x + threadIdx.x; f < nfaces; f += (long long)gridDim.x * blockDim.x) {
        long long i0, i1; face_nodes(g, f, i0, i1);
        const double2 a = pts[i0], b = pts[i1];
        const double l0 = ls_eval(L, a.x, a.y), l1 = ls_eval(L, b.x, b.y);
        if (l0 >= 0 && l1 >= 0) { face_loc[f] = PB_IN_POSITIVE_SIDE; continue; }
        if (l0 < 0 && l1 < 0) { face_loc[f] = PB_IN_NEGATIVE_SIDE; continue; }
        const double threshold = norm2d(__dsub_rn(b.x, a.x), __dsub_rn(b.y, a.y)) / 1e4;
        face_ip[f] = find_zero_crossing(L, a, b, threshold);
        face_loc[f] = PB_ON_INTERFACE;
    }
}

// move_nodes, first loop (cuthho_geom.hpp:466-510): a cut closer than 0.4 |F| to an end node pushes that node away
// (pbegin, pend): only nodes of this range collect displacements (faces of the halo rows also touch nodes outside it)
__global__ void cut_displacements_kernel(Grid g, long long begin, long long nfaces, long long pbegin, long long pend, const double2* __restrict__ pts, const int* __restrict__ face_loc,
                                         const double2* __restrict__ face_ip, double* __restrict__ disp_xy, int* __restrict__ displaced) {
    for (long long f = begin + blockIdx.x * (long long)blockDim.x + threadIdx.x; f < nfaces; f += (long long)gridDim.x * blockDim.x) {
        if (face_loc[f] != PB_ON_INTERFACE) continue;
        long long i0, i1; face_nodes(g, f, i0, i1);
        const double2 a = pts[i0], b = pts[i1], ip = face_ip[f];
        const double barx = __dadd_rn(b.x, a.x) / 2.0, bary = __dadd_rn(b.y, a.y) / 2.0;
        const double lf = norm2d(__dsub_rn(b.x, a.x), __dsub_rn(b.y, a.y)), dp = norm2d(__dsub_rn(ip.x, a.x), __dsub_rn(ip.y, a.y));
        const double closeness = dp / lf, thresh = 0.4;
        long long ofs;
        if (closeness < thresh) ofs = i0;
        else if (closeness > 1.0 - thresh) ofs = i1;
        else continue;
        if (ofs < pbegin || ofs >= pend) continue;
        const double dx = __dsub_rn(barx, ip.x) / 2.0, dy = __dsub_rn(bary, ip.y) / 2.0;
        // displacement = displacement - delta; a node receives at most two contributions for a smooth interface, and
        // floating-point addition of two terms is commutative, so the atomic order does not change the bits
        atomicAdd(&disp_xy[2 * ofs], -dx); atomicAdd(&disp_xy[2 * ofs + 1], -dy);
        displaced[ofs] = 1;
    }
}
__global__ void cut_apply_displacements_kernel(long long begin, long long npoints, double2* __restrict__ pts, const double* __restrict__ disp_xy, const int* __restrict__ displaced) {
    for (long long p = begin + blockIdx.x * (long long)blockDim.x + threadIdx.x; p < npoints; p += (long long)gridDim.x * blockDim.x)
        if (displaced[p]) { double2 v = pts[p]; v.x = __dadd_rn(v.x, disp_xy[2 * p]); v.y = __dadd_rn(v.y, disp_xy[2 * p + 1]); pts[p] = v; }
}

// move_nodes, last loop (cuthho_geom.hpp:520-542): a displaced node must not make a cell concave;
// detect_cut_cells (cuthho_geom.hpp:275-340): 0 cut faces -> side by the vertices (POSITIVE only if all ls > 0),
// 2 cut faces -> ON_INTERFACE with the interface oriented so that the left normal points to the negative side.
__global__ void cut_cells_kernel(LevelSet L, Grid g, long long begin, long long ncells, const double2* __restrict__ pts, const int* __restrict__ displaced,
                                 const int* __restrict__ face_loc, const double2* __restrict__ face_ip, int* __restrict__ cell_loc,
                                 double2* __restrict__ cell_p01, int* __restrict__ is_cut, int* __restrict__ err) {
    for (long long c = begin + blockIdx.x * (long long)blockDim.x + threadIdx.x; c < ncells; c += (long long)gridDim.x * blockDim.x) {
        const long long ci = c % g.Nx, cj = c / g.Nx, p0 = grid_point(g, ci, cj);
        const long long pid[4] = {p0, p0 + 1, p0 + g.Nx + 2, p0 + g.Nx + 1};
        double2 v[4];
        bool distorted = false;
        for (int i = 0; i < 4; i++) { v[i] = pts[pid[i]]; distorted |= displaced[pid[i]] != 0; }
        if (distorted) {
            for (int i = 0; i < 4; i++) {
                const double2 a = v[i], b = v[(i + 1) & 3], d = v[(i + 2) & 3];
                const double v1x = __dsub_rn(b.x, a.x), v1y = __dsub_rn(b.y, a.y), v2x = __dsub_rn(d.x, b.x), v2y = __dsub_rn(d.y, b.y);
                if (__dsub_rn(__dmul_rn(v1x, v2y), __dmul_rn(v2x, v1y)) < 0) atomicOr(err, CUTERR_CONCAVE);
            }
        }
        long long fids[4]; cell_faces(g, ci, cj, fids);
        double2 cut[2]; int k = 0;
        for (int i = 0; i < 4; i++)
            if (face_loc[fids[i]] == PB_ON_INTERFACE) { if (k < 2) cut[k] = face_ip[fids[i]]; k++; }
        is_cut[c] = 0;
        if (k == 0) {
            bool allpos = true;
            for (int i = 0; i < 4; i++) if (!(ls_eval(L, v[i].x, v[i].y) > 0)) allpos = false;
            cell_loc[c] = allpos ? PB_IN_POSITIVE_SIDE : PB_IN_NEGATIVE_SIDE;
        } else if (k == 2) {
            cell_loc[c] = PB_ON_INTERFACE; is_cut[c] = 1;
            const double2 a = cut[0], b = cut[1];
            const double ptx = __dsub_rn(b.x, a.x), pty = __dsub_rn(b.y, a.y);
            const double pnx = __dadd_rn(a.x, -pty), pny = __dadd_rn(a.y, ptx);
            if (ls_eval(L, pnx, pny) >= 0) { cell_p01[2 * c] = b; cell_p01[2 * c + 1] = a; }
            else { cell_p01[2 * c] = a; cell_p01[2 * c + 1] = b; }
        } else {
            cell_loc[c] = PB_UNDEF;
            atomicOr(err, k > 2 ? CUTERR_TOO_MANY_CUTS : CUTERR_BAD_CUT_COUNT);
        }
    }
}

// cut list: slot -> cell from the exclusive scan of is_cut
__global__ void cut_compact_kernel(long long begin, long long ncells, const int* __restrict__ is_cut, const int* __restrict__ slot, int* __restrict__ cut_cells) {
    for (long long c = begin + blockIdx.x * (long long)blockDim.x + threadIdx.x; c < ncells; c += (long long)gridDim.x * blockDim.x)
        if (is_cut[c]) cut_cells[slot[c]] = (int)c;
}

// refine_interface (cuthho_geom.hpp:609-673): 2^levels segments per cut cell; every midpoint is projected onto the
// zero level set along the normal of its parent segment.  The recursion only couples a midpoint to its two parents,
// so it is evaluated level by level.  iface is [ncut][NI] with NI = 2^levels + 1.
__global__ void cut_refine_kernel(LevelSet L, Grid g, int ncut, int NI, const int* __restrict__ cut_cells, const double2* __restrict__ pts,
                                  const double2* __restrict__ cell_p01, double2* __restrict__ iface, int* __restrict__ err) {
    for (int s = blockIdx.x * blockDim.x + threadIdx.x; s < ncut; s += gridDim.x * blockDim.x) {
        const long long c = cut_cells[s], ci = c % g.Nx, cj = c / g.Nx, p0 = grid_point(g, ci, cj);
        const long long pid[4] = {p0, p0 + 1, p0 + g.Nx + 2, p0 + g.Nx + 1};
        double diam = 0.0;
        for (int i = 0; i < 4; i++) for (int j = i + 1; j < 4; j++) {
            const double2 a = pts[pid[i]], b = pts[pid[j]];
            diam = fmax(diam, norm2d(__dsub_rn(b.x, a.x), __dsub_rn(b.y, a.y)));
        }
        const double thr = diam / 10000.0;
        double2* I = iface + (long long)s * NI;
        I[0] = cell_p01[2 * c]; I[NI - 1] = cell_p01[2 * c + 1];
        for (int span = NI - 1; span >= 2; span >>= 1) {
            for (int mn = 0; mn + span <= NI - 1; mn += span) {
                const double2 a = I[mn], b = I[mn + span];
                const double pmx = __dadd_rn(a.x, b.x) / 2.0, pmy = __dadd_rn(a.y, b.y) / 2.0;
                const double ptx = __dsub_rn(b.x, a.x), pty = __dsub_rn(b.y, a.y);
                const double2 pm = make_double2(pmx, pmy);
                const double2 ps1 = make_double2(__dadd_rn(pmx, -pty), __dadd_rn(pmy, ptx)), ps2 = make_double2(__dsub_rn(pmx, -pty), __dsub_rn(pmy, ptx));
                const double lm = ls_eval(L, pmx, pmy), l1 = ls_eval(L, ps1.x, ps1.y), l2 = ls_eval(L, ps2.x, ps2.y);
                double2 ip = pm;
                if (!((lm >= 0 && l1 >= 0) || (lm < 0 && l1 < 0))) ip = find_zero_crossing(L, pm, ps1, thr);
                else if (!((lm >= 0 && l2 >= 0) || (lm < 0 && l2 < 0))) ip = find_zero_crossing(L, pm, ps2, thr);
                else atomicOr(err, CUTERR_IFACE_NOT_FOUND);
                I[mn + span / 2] = ip;
            }
        }
    }
}

}  // namespace pb
