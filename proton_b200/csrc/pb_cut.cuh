// K2 — local operator of one CUT cell of the fictitious-domain problem (where = negative side), followed by the same
// static condensation as K1.  Restates the driver-resident cut operators of apps/cuthho/cuthho_square.cpp:
//   make_hho_laplacian(msh, cl, level_set, hdi, where)   :308-388   (full R x R gr_lhs with the Nitsche terms, llt)
//   make_hho_cut_stabilization(msh, cl, di, where)       :566-621
//   make_rhs(msh, cl, degree, f, where, level_set, bcs)  :623-666   (interface quadrature of degree cd, :647)
// on the sub-cell quadrature of src/methods/cuthho_bits/cuthho_geom.hpp:675-895 (collect_triangulation_points, fan
// around the polygon centroid with Dunavant rules, cut faces integrate [p_where, ip] only, integrate_interface).
// Cut cells are O(perimeter) of the mesh (~2.8 N of N^2), so this kernel is written for clarity: one warp per cut
// cell, all matrices in shared memory, runtime loops.
#pragma once
#include "pb_cutprep.cuh"
#include "pb_local.cuh"

namespace pb {

struct CutArgs {
    Grid grid;                      // grid.row_begin = first local cell row
    LevelSet ls;
    const double2* points;
    const int* node_loc;            // per mesh point
    const int* face_loc;            // per mesh face
    const double2* face_ip;         // per mesh face
    const int* cut_cells;           // GLOBAL cell ids of the cut cells
    const double2* iface;           // [ncut][NI]
    int ncut, NI;
    long long cell_begin, cell_end; // global cell range stored by this context (local index = global - cell_begin)
    int fun; double eta;
    double *S, *g, *recA, *recb, *lc, *oper, *rhs;     // same layout as LocalArgs
    double* data;                   // gr.second alone, or nullptr
};

constexpr int CUT_MAX_TP = 4 + 65 + 1;      // polygon of the where-part: <= 3 vertices + interface polyline (refine <= 6)

// basis values and gradients of the scaled monomials at (bx, by); ih = 2/h
template <int RD>
PB_D void eval_cell_basis(double bx, double by, double ih, double* phi, double* dx, double* dy) {
    constexpr int R = cell_basis_size(RD);
    double xb[RD + 1], yb[RD + 1];
    powers<RD>(bx, xb); powers<RD>(by, yb);
#pragma unroll
    for (int r = 0; r < R; r++) {
        const int ex = mono_px(r), ey = mono_py(r);
        phi[r] = xb[ex] * yb[ey];
        dx[r] = ex == 0 ? 0.0 : ex * ih * xb[ex > 0 ? ex - 1 : 0] * yb[ey];
        dy[r] = ey == 0 ? 0.0 : ey * ih * xb[ex] * yb[ey > 0 ? ey - 1 : 0];
    }
}

// in-place LDL^T of the n x n SPD matrix A (column-major, leading dimension lda) by one warp.  On return the strict
// lower part holds unit-lower L and the diagonal holds d_i.
PB_D void warp_ldlt(double* A, int n, int lda, int lane) {
    for (int k = 0; k < n; k++) {
        __syncwarp();
        const double d = A[k + k * lda];
        const int m = n - k - 1;
        for (int idx = lane; idx < m * m; idx += 32) {
            const int jj = idx / m, ii = idx - jj * m, i = k + 1 + ii, j = k + 1 + jj;
            if (i >= j) A[i + j * lda] -= A[i + k * lda] * A[j + k * lda] / d;
        }
        __syncwarp();
        for (int i = k + 1 + lane; i < n; i += 32) A[i + k * lda] /= d;
    }
    __syncwarp();
}
// B(:, col) <- (L D L^T)^{-1} B(:, col), one lane per column
PB_D void warp_ldlt_solve(const double* A, int n, int lda, double* B, int ldb, int ncols, int lane) {
    for (int col = lane; col < ncols; col += 32) {
        double* x = B + (long long)col * ldb;
        for (int i = 0; i < n; i++) { double v = x[i]; for (int k = 0; k < i; k++) v -= A[i + k * lda] * x[k]; x[i] = v; }
        for (int i = 0; i < n; i++) x[i] /= A[i + i * lda];
        for (int i = n - 1; i >= 0; i--) { double v = x[i]; for (int k = i + 1; k < n; k++) v -= A[k + i * lda] * x[k]; x[i] = v; }
    }
    __syncwarp();
}

template <int CD, int FD>
struct CutSmem {
    using Z = Sizes<CD, FD>;
    static constexpr int R = Z::R, C = Z::C, F = Z::F, N = Z::N, NF = Z::NF;
    double tp[2 * CUT_MAX_TP];          // polygon of the where-part
    double tab[32 * (3 * R + F + 4)];   // per-point tables of one chunk of 32 quadrature points
    double stiff[R * R];                // gr_lhs (then its factor)
    double G[R * N];                    // gr_rhs, column-major R x N
    double X[R * N];                    // oper
    double LC[N * N];                   // data + stab, column-major
    double KT[C * C];                   // K_TT factor
    double AC[(NF + 1) * C];            // K_TT^{-1} [K_TF | f_T]
    double Sb[NF * (NF + 1)];           // Schur block + condensed rhs column
    double fT[C];
    double Mf[F * F], Tr[F * C], Pm[F * C];
    double seg[4 * 4];                  // where-part of each face: ax, ay, bx, by
    int seg_on[4];
    int ntp;
};

// Gauss nodes on a segment in the reference's node order (quadratures.hpp:404-432): point and weight*|seg|/2
PB_D void seg_qp(int ng, int q, double ax, double ay, double bx, double by, double& x, double& y, double& w) {
    const double t = c_qt.gx[ng][q];
    x = 0.5 * (1 - t) * ax + 0.5 * (1 + t) * bx; y = 0.5 * (1 - t) * ay + 0.5 * (1 + t) * by;
    const double dx = bx - ax, dy = by - ay;
    w = c_qt.gw[ng][q] * sqrt(dx * dx + dy * dy) * 0.5;
}

template <int CD, int FD>
__global__ void __launch_bounds__(32) local_cut_kernel(CutArgs A) {
    using Z = Sizes<CD, FD>;
    constexpr int C = Z::C, F = Z::F, R = Z::R, N = Z::N, NF = Z::NF, RD = Z::RD;
    constexpr int TW = 3 * R + F + 4;                 // doubles per point in the chunk tables
    constexpr int WHERE = PB_IN_NEGATIVE_SIDE;        // run_cuthho_fictdom integrates the negative side
    using SM = CutSmem<CD, FD>;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SM& s = *reinterpret_cast<SM*>(smem_raw);
    const int lane = threadIdx.x;
    const Grid& g = A.grid;

    for (int slot = blockIdx.x; slot < A.ncut; slot += gridDim.x) {
        const long long cell = A.cut_cells[slot];
        if (cell < A.cell_begin || cell >= A.cell_end) continue;          // another strip's cell (uniform per block)
        const long long lcell = cell - A.cell_begin;
        const long long ci = cell % g.Nx, cj = cell / g.Nx, p0 = grid_point(g, ci, cj);
        const long long pid[4] = {p0, p0 + 1, p0 + g.Nx + 2, p0 + g.Nx + 1};
        CellGeom geo; int nloc[4];
#pragma unroll
        for (int i = 0; i < 4; i++) { const double2 v = A.points[pid[i]]; geo.vx[i] = v.x; geo.vy[i] = v.y; nloc[i] = A.node_loc[pid[i]]; }
        cell_geometry(geo);                            // full-cell barycenter / diameter (the bases) and area (hT)
        const double ih = geo.ihalf, hT = geo.area;
        long long fids[4]; cell_faces(g, ci, cj, fids);
        const double2* I = A.iface + (long long)slot * A.NI;

        // ---- polygon of the where-part (collect_triangulation_points, cuthho_geom.hpp:675-728), where = negative:
        //      the interface is inserted in its stored orientation
        if (lane == 0) {
            int n = 0;
            auto push = [&](double x, double y) { s.tp[2 * n] = x; s.tp[2 * n + 1] = y; n++; };
            auto push_iface = [&]() { for (int i = 0; i < A.NI; i++) push(I[i].x, I[i].y); };
            const bool case1 = nloc[0] == WHERE && nloc[3] != WHERE, case2 = nloc[0] != WHERE && nloc[3] == WHERE, case3 = nloc[0] != WHERE && nloc[3] != WHERE;
            if (case1 || case2 || case3) {
                for (int i = 0; i < 4; i++) if (nloc[i] == WHERE) push(sel4(geo.vx, i), sel4(geo.vy, i));
                push_iface();
            } else {
                int i = 0;
                while (i < 4 && nloc[i] == WHERE) { push(sel4(geo.vx, i), sel4(geo.vy, i)); i++; }
                push_iface();
                while (i < 4 && nloc[i] != WHERE) i++;
                while (i < 4 && nloc[i] == WHERE) { push(sel4(geo.vx, i), sel4(geo.vy, i)); i++; }
            }
            s.ntp = n;
            // where-part of each face (points_where, cuthho_geom.hpp:547-569; integrate(fc, deg, where) :817-849)
            for (int li = 0; li < 4; li++) {
                const int fl = A.face_loc[fids[li]];
                double ax, ay, bx, by; face_points(geo, li, ax, ay, bx, by);
                int on = 0;
                if (fl == WHERE) on = 1;
                else if (fl == PB_ON_INTERFACE) {
                    on = 1;
                    int i0 = li, i1 = (li + 1) & 3; if (li >= 2) { int t = i0; i0 = i1; i1 = t; }     // sorted point order
                    const double2 ip = A.face_ip[fids[li]];
                    if (nloc[i0] == WHERE && nloc[i1] != WHERE) { bx = ip.x; by = ip.y; }
                    else if (nloc[i0] != WHERE && nloc[i1] == WHERE) { ax = ip.x; ay = ip.y; }
                    else on = 0;                        // "Invalid point configuration" cannot happen after detect_cut_cells
                }
                s.seg_on[li] = on; s.seg[4 * li] = ax; s.seg[4 * li + 1] = ay; s.seg[4 * li + 2] = bx; s.seg[4 * li + 3] = by;
            }
        }
        for (int i = lane; i < R * R; i += 32) s.stiff[i] = 0.0;
        for (int i = lane; i < R * N; i += 32) s.G[i] = 0.0;
        for (int i = lane; i < N * N; i += 32) s.LC[i] = 0.0;
        for (int i = lane; i < C; i += 32) s.fT[i] = 0.0;
        __syncwarp();
        const int ntp = s.ntp;
        // barycenter of the polygon (barycenter(msh, cl, where), cuthho_geom.hpp:588-607; basic_geom.hpp:247-286)
        double pbx, pby;
        {
            double rx = 0, ry = 0, den = 0;
            const double x0 = s.tp[0], y0 = s.tp[1];
            for (int i = 2; i < ntp; i++) {
                const double ax = s.tp[2 * (i - 1)] - x0, ay = s.tp[2 * (i - 1) + 1] - y0, bx = s.tp[2 * i] - x0, by = s.tp[2 * i + 1] - y0;
                const double d = (ax * by - ay * bx) / 2.0;
                rx += (ax + bx) * d; ry += (ay + by) * d; den += d;
            }
            pbx = x0 + rx / (den * 3); pby = y0 + ry / (den * 3);
        }
        // q-th point of integrate(msh, cl, degree, where) (cuthho_geom.hpp:754-815): triangles (bar, tp_i, tp_{i+1})
        auto cell_where_qp = [&](int ri, int np, int q, double& x, double& y, double& w) {
            const int t = q / np, p = q - t * np, t1 = (t + 1) % ntp;
            const double ax = s.tp[2 * t], ay = s.tp[2 * t + 1], bx = s.tp[2 * t1], by = s.tp[2 * t1 + 1];
            const double v0x = ax - pbx, v0y = ay - pby, v1x = bx - pbx, v1y = by - pby;
            const double area = fabs((v0x * v1y - v0y * v1x) / 2.0);
            const double* row = c_qt.dun[ri][p];
            x = pbx * row[0] + ax * row[1] + bx * row[2]; y = pby * row[0] + ay * row[1] + by * row[2];
            w = area * row[3];
        };
        // ---- stiffness over the where-part, degree 2*rd (cuthho_square.cpp:328-338)
        {
            const int ri = dunavant_rule_index(2 * RD), np = c_qt.dun_n[ri], nq = ntp * np;
            for (int base = 0; base < nq; base += 32) {
                const int q = base + lane;
                double* T = s.tab + lane * TW;
                if (q < nq) {
                    double x, y, w; cell_where_qp(ri, np, q, x, y, w);
                    double phi[R], dx[R], dy[R];
                    eval_cell_basis<RD>((x - geo.barx) * ih, (y - geo.bary) * ih, ih, phi, dx, dy);
#pragma unroll
                    for (int r = 0; r < R; r++) { T[r] = dx[r]; T[R + r] = dy[r]; }
                    T[3 * R] = w;
                }
                __syncwarp();
                const int cnt = min(32, nq - base);
                for (int idx = lane; idx < R * R; idx += 32) {
                    const int jj = idx / R, ii = idx - jj * R;
                    double acc = 0.0;
                    for (int p = 0; p < cnt; p++) {
                        const double* U = s.tab + p * TW;
                        acc += U[3 * R] * (U[ii] * U[jj] + U[R + ii] * U[R + jj]);
                    }
                    s.stiff[idx] += acc;
                }
                __syncwarp();
            }
        }
        // ---- interface: global sign of integrate_interface (cuthho_geom.hpp:851-895) and the Nitsche terms
        //      -w phi_i (grad phi_j . n) - w (grad phi_i . n) phi_j + w phi_i phi_j eta / hT   (cuthho_square.cpp:343-363)
        double int_sign;
        {
            const double vax = I[0].x - pbx, vay = I[0].y - pby, vbtx = I[1].x - I[0].x, vbty = I[1].y - I[0].y;
            int_sign = (vax * vbty + vay * (-vbtx)) < 0 ? -1.0 : 1.0;
        }
        {
            const int ng = gauss_nodes(2 * RD), nq = (A.NI - 1) * ng;
            for (int base = 0; base < nq; base += 32) {
                const int q = base + lane;
                double* T = s.tab + lane * TW;
                if (q < nq) {
                    const int sg = q / ng, p = q - sg * ng;
                    double x, y, w; seg_qp(ng, p, I[sg].x, I[sg].y, I[sg + 1].x, I[sg + 1].y, x, y, w);
                    w *= int_sign;
                    double nx, ny; ls_normal(A.ls, x, y, nx, ny);
                    double phi[R], dx[R], dy[R];
                    eval_cell_basis<RD>((x - geo.barx) * ih, (y - geo.bary) * ih, ih, phi, dx, dy);
#pragma unroll
                    for (int r = 0; r < R; r++) { T[r] = phi[r]; T[R + r] = dx[r] * nx + dy[r] * ny; }
                    T[3 * R] = w;
                }
                __syncwarp();
                const int cnt = min(32, nq - base);
                for (int idx = lane; idx < R * R; idx += 32) {
                    const int jj = idx / R, ii = idx - jj * R;
                    double acc = 0.0;
                    for (int p = 0; p < cnt; p++) {
                        const double* U = s.tab + p * TW;
                        acc += U[3 * R] * (-U[ii] * U[R + jj] - U[R + ii] * U[jj] + U[ii] * U[jj] * A.eta / hT);
                    }
                    s.stiff[idx] += acc;
                }
                __syncwarp();
            }
        }
        // gr_rhs(:, :C) = stiff(:, :C) (cuthho_square.cpp:366)
        for (int idx = lane; idx < R * C; idx += 32) s.G[idx] = s.stiff[idx];
        __syncwarp();
        // ---- face terms on the where-part of each face, degree 2*rd (cuthho_square.cpp:368-383); the face basis
        //      spans the FULL face (bases.hpp:241-291), the normal is the cell's outward normal of that face
        {
            const int ng = gauss_nodes(2 * RD);
            const int q = lane;                                   // 4 faces x ng <= 20 points
            double* T = s.tab + lane * TW;
            const int li = q / ng, p = q - li * ng;
            if (li < 4) {
                double w = 0.0, x = geo.barx, y = geo.bary;
                if (s.seg_on[li]) seg_qp(ng, p, s.seg[4 * li], s.seg[4 * li + 1], s.seg[4 * li + 2], s.seg[4 * li + 3], x, y, w);
                double ax, ay, bx, by; face_points(geo, li, ax, ay, bx, by);
                double nx, ny; face_normal(geo, li, nx, ny);
                double phi[R], dx[R], dy[R];
                eval_cell_basis<RD>((x - geo.barx) * ih, (y - geo.bary) * ih, ih, phi, dx, dy);
#pragma unroll
                for (int r = 0; r < R; r++) { T[r] = phi[r]; T[R + r] = dx[r] * nx + dy[r] * ny; }
                const double ex_ = bx - ax, ey_ = by - ay, len2 = ex_ * ex_ + ey_ * ey_;
                const double fbx = 0.5 * (ax + bx), fby = 0.5 * (ay + by);
                const double ep = 4.0 * ((fbx - ax) * (x - fbx) + (fby - ay) * (y - fby)) / len2;
                double fp = 1.0;
#pragma unroll
                for (int b = 0; b < F; b++) { T[2 * R + b] = fp; fp *= ep; }
                T[3 * R] = w;
            }
            __syncwarp();
            for (int idx = lane; idx < R * N; idx += 32) {
                const int col = idx / R, r = idx - col * R;
                double acc = 0.0;
                if (col < C) {
                    for (int pp = 0; pp < 4 * ng; pp++) { const double* U = s.tab + pp * TW; acc -= U[3 * R] * U[R + r] * U[col]; }
                } else {
                    const int fi = (col - C) / F, b = (col - C) - fi * F;
                    for (int pp = fi * ng; pp < (fi + 1) * ng; pp++) { const double* U = s.tab + pp * TW; acc += U[3 * R] * U[R + r] * U[2 * R + b]; }
                }
                s.G[idx] += acc;
            }
            __syncwarp();
        }
        // ---- oper = llt(gr_lhs).solve(gr_rhs), data = gr_rhs^T oper (cuthho_square.cpp:385-387)
        for (int idx = lane; idx < R * N; idx += 32) s.X[idx] = s.G[idx];
        warp_ldlt(s.stiff, R, R, lane);
        warp_ldlt_solve(s.stiff, R, R, s.X, R, N, lane);
        for (int idx = lane; idx < N * N; idx += 32) {
            const int j = idx / N, i = idx - j * N;
            double acc = 0.0;
            for (int r = 0; r < R; r++) acc += s.G[r + i * R] * s.X[r + j * R];
            s.LC[idx] = acc;
            if (A.data) A.data[lcell * (long long)(N * N) + idx] = acc;
        }
        __syncwarp();
        // ---- make_hho_cut_stabilization (cuthho_square.cpp:566-621): naive stabilisation on the where-part of each
        //      face, degree 2*fd, faces without points skipped, scaled by 1/area
        for (int li = 0; li < 4; li++) {
            if (!s.seg_on[li]) continue;                          // uniform per warp
            const int ng = gauss_nodes(2 * FD);
            double* T = s.tab + lane * TW;
            if (lane < ng) {
                double x, y, w; seg_qp(ng, lane, s.seg[4 * li], s.seg[4 * li + 1], s.seg[4 * li + 2], s.seg[4 * li + 3], x, y, w);
                double ax, ay, bx, by; face_points(geo, li, ax, ay, bx, by);
                double phi[R], dx[R], dy[R];
                eval_cell_basis<RD>((x - geo.barx) * ih, (y - geo.bary) * ih, ih, phi, dx, dy);
#pragma unroll
                for (int r = 0; r < R; r++) T[r] = phi[r];
                const double ex_ = bx - ax, ey_ = by - ay, len2 = ex_ * ex_ + ey_ * ey_;
                const double fbx = 0.5 * (ax + bx), fby = 0.5 * (ay + by);
                const double ep = 4.0 * ((fbx - ax) * (x - fbx) + (fby - ay) * (y - fby)) / len2;
                double fp = 1.0;
#pragma unroll
                for (int b = 0; b < F; b++) { T[2 * R + b] = fp; fp *= ep; }
                T[3 * R] = w;
            }
            __syncwarp();
            for (int idx = lane; idx < F * F; idx += 32) {
                const int b = idx / F, a = idx - b * F; double acc = 0.0;
                for (int p = 0; p < ng; p++) { const double* U = s.tab + p * TW; acc += U[3 * R] * U[2 * R + a] * U[2 * R + b]; }
                s.Mf[idx] = acc;
            }
            for (int idx = lane; idx < F * C; idx += 32) {
                const int c = idx / F, a = idx - c * F; double acc = 0.0;
                for (int p = 0; p < ng; p++) { const double* U = s.tab + p * TW; acc += U[3 * R] * U[2 * R + a] * U[c]; }
                s.Tr[idx] = acc; s.Pm[idx] = acc;
            }
            __syncwarp();
            // P = M^{-1} trace on a copy of M (the raw mass is needed below); the copy lives in the chunk table area
            double* Mfac = s.tab + 32 * TW - F * F;
            for (int idx = lane; idx < F * F; idx += 32) Mfac[idx] = s.Mf[idx];
            warp_ldlt(Mfac, F, F, lane);
            warp_ldlt_solve(Mfac, F, F, s.Pm, F, C, lane);
            // data += (1/hT) oper^T (M oper), oper = [P | ... -I_F ...]
            for (int idx = lane; idx < N * N; idx += 32) {
                const int j = idx / N, i = idx - j * N;
                const bool ic = i < C, jc = j < C, iface_i = !ic && (i - C) / F == li, iface_j = !jc && (j - C) / F == li;
                if (!(ic || iface_i) || !(jc || iface_j)) continue;
                double acc = 0.0;
                for (int a = 0; a < F; a++) {
                    const double oi = ic ? s.Pm[a + i * F] : ((i - C) - li * F == a ? -1.0 : 0.0);
                    if (oi == 0.0 && !ic) continue;
                    double mo = 0.0;                                   // (M oper)(a, j)
                    for (int b = 0; b < F; b++) {
                        const double oj = jc ? s.Pm[b + j * F] : ((j - C) - li * F == b ? -1.0 : 0.0);
                        mo += s.Mf[a + b * F] * oj;
                    }
                    acc += oi * mo;
                }
                s.LC[idx] += acc / hT;
            }
            __syncwarp();
        }
        // ---- fictitious-domain load (cuthho_square.cpp:623-666): int_{T cap Omega} f phi (degree 2*cd) plus the
        //      Nitsche datum  sum w g (phi eta/hT - grad phi . n)  with interface degree cd (not 2*cd, :647)
        {
            const int ri = dunavant_rule_index(2 * CD), np = c_qt.dun_n[ri], nq = ntp * np;
            for (int base = 0; base < nq; base += 32) {
                const int q = base + lane;
                double* T = s.tab + lane * TW;
                if (q < nq) {
                    double x, y, w; cell_where_qp(ri, np, q, x, y, w);
                    double phi[R], dx[R], dy[R];
                    eval_cell_basis<RD>((x - geo.barx) * ih, (y - geo.bary) * ih, ih, phi, dx, dy);
                    const double wf = w * fun_rhs(A.fun, x, y);
#pragma unroll
                    for (int r = 0; r < C; r++) T[r] = wf * phi[r];
                }
                __syncwarp();
                const int cnt = min(32, nq - base);
                for (int a = lane; a < C; a += 32) { double acc = 0.0; for (int p = 0; p < cnt; p++) acc += s.tab[p * TW + a]; s.fT[a] += acc; }
                __syncwarp();
            }
            const int ng = gauss_nodes(CD), nqi = (A.NI - 1) * ng;
            for (int base = 0; base < nqi; base += 32) {
                const int q = base + lane;
                double* T = s.tab + lane * TW;
                if (q < nqi) {
                    const int sg = q / ng, p = q - sg * ng;
                    double x, y, w; seg_qp(ng, p, I[sg].x, I[sg].y, I[sg + 1].x, I[sg + 1].y, x, y, w);
                    w *= int_sign;
                    double nx, ny; ls_normal(A.ls, x, y, nx, ny);
                    double phi[R], dx[R], dy[R];
                    eval_cell_basis<RD>((x - geo.barx) * ih, (y - geo.bary) * ih, ih, phi, dx, dy);
                    const double wg = w * fun_sol(A.fun, x, y);
#pragma unroll
                    for (int r = 0; r < C; r++) T[r] = wg * (phi[r] * A.eta / hT - (dx[r] * nx + dy[r] * ny));
                }
                __syncwarp();
                const int cnt = min(32, nqi - base);
                for (int a = lane; a < C; a += 32) { double acc = 0.0; for (int p = 0; p < cnt; p++) acc += s.tab[p * TW + a]; s.fT[a] += acc; }
                __syncwarp();
            }
        }
        if (A.lc) {          // test hook: lc = data + stab, oper (R rows), cell load
            for (int idx = lane; idx < N * N; idx += 32) A.lc[lcell * (long long)(N * N) + idx] = s.LC[idx];
            for (int idx = lane; idx < R * N; idx += 32) A.oper[lcell * (long long)(R * N) + idx] = s.X[idx];
            for (int a = lane; a < C; a += 32) A.rhs[lcell * (long long)C + a] = s.fT[a];
        }
        // ---- static condensation, as in K1 phase 6
        for (int idx = lane; idx < C * C; idx += 32) { const int j = idx / C, i = idx - j * C; s.KT[idx] = s.LC[i + j * N]; }
        for (int idx = lane; idx < (NF + 1) * C; idx += 32) {
            const int col = idx / C, a = idx - col * C;
            s.AC[idx] = col < NF ? s.LC[a + (C + col) * N] : s.fT[a];
        }
        warp_ldlt(s.KT, C, C, lane);
        warp_ldlt_solve(s.KT, C, C, s.AC, C, NF + 1, lane);
        for (int idx = lane; idx < NF * (NF + 1); idx += 32) {
            const int col = idx / NF, i = idx - col * NF;
            double v = col < NF ? s.LC[(C + i) + (C + col) * N] : 0.0;
            for (int a = 0; a < C; a++) v -= s.LC[a + (C + i) * N] * s.AC[a + col * C];       // K_FT(i, a) = K_TF(a, i)
            s.Sb[idx] = v;
        }
        __syncwarp();
        // Dirichlet lifting (hho.hpp:381-403) for boundary faces of the cut cell: full-face L2 projection of the datum
        for (int li = 0; li < 4; li++) {
            if (!cell_face_dirichlet(g, ci, cj, li)) continue;
            if (lane == 0) {
                constexpr int QF = Z::QF;
                double ax, ay, bx, by; face_points(geo, li, ax, ay, bx, by);
                double Mf[F * (F + 1) / 2], gd[F];
                for (int k = 0; k < F * (F + 1) / 2; k++) Mf[k] = 0.0;
                for (int k = 0; k < F; k++) gd[k] = 0.0;
                for (int q = 0; q < QF; q++) {
                    double x, y, w; seg_qp(QF, q, ax, ay, bx, by, x, y, w);
                    const double ex_ = bx - ax, ey_ = by - ay, len2 = ex_ * ex_ + ey_ * ey_;
                    const double fbx = 0.5 * (ax + bx), fby = 0.5 * (ay + by);
                    const double ep = 4.0 * ((fbx - ax) * (x - fbx) + (fby - ay) * (y - fby)) / len2;
                    double ph[F]; ph[0] = 1.0;
                    for (int k = 1; k < F; k++) ph[k] = ph[k - 1] * ep;
                    const double bc = fun_sol(A.fun, x, y);
                    for (int k = 0; k < F; k++) { gd[k] += w * ph[k] * bc; for (int m = 0; m <= k; m++) Mf[PB_TRI(k, m)] += w * ph[k] * ph[m]; }
                }
                ldlt_packed<F>(Mf); ldlt_solve<F>(Mf, gd);
                for (int i = 0; i < NF; i++) for (int a = 0; a < F; a++) s.Sb[i + NF * NF] -= s.Sb[i + (li * F + a) * NF] * gd[a];
            }
            __syncwarp();
        }
        for (int idx = lane; idx < NF * NF; idx += 32) A.S[lcell * (long long)(NF * NF) + idx] = s.Sb[idx];
        for (int i = lane; i < NF; i += 32) A.g[lcell * (long long)NF + i] = s.Sb[NF * NF + i];
        for (int idx = lane; idx < NF * C; idx += 32) A.recA[lcell * (long long)(NF * C) + idx] = s.AC[idx];
        for (int a = lane; a < C; a += 32) A.recb[lcell * (long long)C + a] = s.AC[NF * C + a];
        __syncwarp();
    }
}

}  // namespace pb
