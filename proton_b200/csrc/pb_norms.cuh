// K7 — error norms of the drivers' post-processing loops on the device:
//   PB_PROBLEM_HHO      convergence_test.cpp:254-306: sum of L2 errors^2 (degree 2*cd), the "mass-norm" term (added
//                       once per quadrature point, quirk kept) and the energy error^2 of the RECONSTRUCTION over the
//                       same degree-2*cd points (quirk kept)
//   PB_PROBLEM_FICTDOM  cuthho_square.cpp:1031-1050: H1 error^2 over the negative and cut cells, gradient of the
//                       CELL dofs in the rd basis, degree 2*cd on the negative part
// One warp per cell, matrices in shared memory, runtime loops (post-processing: clarity over speed).  The
// reconstruction is not stored by K1; it is recomputed here from  gr_lhs y = gr_rhs u  with one right-hand side:
//   (gr_rhs u)_r = a(u_T, phi_r) + sum_F int_F (grad phi_r . n)(u_F - u_T).
#pragma once
#include "pb_cut.cuh"

namespace pb {

struct NormArgs {
    Grid grid;                      // row_begin / row_end = owned cell rows
    const double2* points;
    const double* ucell;            // cell dofs, indexed by (global cell - row_begin*Nx) * C
    const double* xfull; long long xoff;   // face dofs for the columns [col_lo, col_hi), xoff = col_lo * F
    int fun, problem;
    // fictdom only
    const int* cell_loc_g; const int* node_loc; const int* cut_cells; const double2* iface; int ncut, NI;
    double* parts;                  // [gridDim.x][3]
};

template <int CD, int FD>
struct NormSmem {
    using Z = Sizes<CD, FD>;
    static constexpr int R = Z::R, C = Z::C, F = Z::F, N = Z::N;
    static constexpr int TW = 3 * R + F + 4;
    double tp[2 * CUT_MAX_TP];
    double tab[32 * TW];
    double stiff[R * R], mass[C * C], massf[C * C];
    double u[N], b[R], proj[C], rec[R];
    int ntp;
};

// QK selects integrate(msh, cl, deg) for uncut cells (tensor Gauss / Dunavant fan)
template <int CD, int FD, int QK>
__global__ void __launch_bounds__(32) error_norms_kernel(NormArgs A) {
    using Z = Sizes<CD, FD>;
    constexpr int C = Z::C, F = Z::F, R = Z::R, M = Z::M, RD = Z::RD, QF = Z::QF;
    using SM = NormSmem<CD, FD>;
    constexpr int TW = SM::TW;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SM& s = *reinterpret_cast<SM*>(smem_raw);
    const int lane = threadIdx.x;
    const Grid& g = A.grid;
    const long long ncells = (g.row_end - g.row_begin) * g.Nx;
    double sum0 = 0.0, sum1 = 0.0, sum2 = 0.0;       // lane 0 accumulates, cells in ascending order: deterministic

    for (long long lc = blockIdx.x; lc < ncells; lc += gridDim.x) {
        const long long ci = lc % g.Nx, cj = g.row_begin + lc / g.Nx, cell = cj * g.Nx + ci, p0 = grid_point(g, ci, cj);
        int loc = PB_IN_NEGATIVE_SIDE;
        if (A.problem == PB_PROBLEM_FICTDOM) { loc = A.cell_loc_g[cell]; if (loc == PB_IN_POSITIVE_SIDE) continue; }
        const long long pid[4] = {p0, p0 + 1, p0 + g.Nx + 2, p0 + g.Nx + 1};
        CellGeom geo;
#pragma unroll
        for (int i = 0; i < 4; i++) { const double2 v = A.points[pid[i]]; geo.vx[i] = v.x; geo.vy[i] = v.y; }
        cell_geometry(geo);
        const double ih = geo.ihalf;
        for (int i = lane; i < C; i += 32) s.u[i] = A.ucell[lc * C + i];
        __syncwarp();

        if (A.problem == PB_PROBLEM_FICTDOM) {
            // ---- polygon of the negative part: the whole cell, or collect_triangulation_points for a cut cell
            double pbx = geo.barx, pby = geo.bary;
            if (lane == 0) {
                int n = 0;
                auto push = [&](double x, double y) { s.tp[2 * n] = x; s.tp[2 * n + 1] = y; n++; };
                if (loc != PB_ON_INTERFACE) { for (int i = 0; i < 4; i++) push(sel4(geo.vx, i), sel4(geo.vy, i)); }
                else {
                    int lo = 0, hi = A.ncut - 1;                 // slot of this cell in the ascending cut list
                    while (lo < hi) { int mid = (lo + hi) / 2; if (A.cut_cells[mid] < cell) lo = mid + 1; else hi = mid; }
                    const double2* I = A.iface + (long long)lo * A.NI;
                    int nl[4]; for (int i = 0; i < 4; i++) nl[i] = A.node_loc[pid[i]];
                    const int W = PB_IN_NEGATIVE_SIDE;
                    auto push_iface = [&]() { for (int i = 0; i < A.NI; i++) push(I[i].x, I[i].y); };
                    const bool c1 = nl[0] == W && nl[3] != W, c2 = nl[0] != W && nl[3] == W, c3 = nl[0] != W && nl[3] != W;
                    if (c1 || c2 || c3) { for (int i = 0; i < 4; i++) if (nl[i] == W) push(sel4(geo.vx, i), sel4(geo.vy, i)); push_iface(); }
                    else {
                        int i = 0;
                        while (i < 4 && nl[i] == W) { push(sel4(geo.vx, i), sel4(geo.vy, i)); i++; }
                        push_iface();
                        while (i < 4 && nl[i] != W) i++;
                        while (i < 4 && nl[i] == W) { push(sel4(geo.vx, i), sel4(geo.vy, i)); i++; }
                    }
                }
                s.ntp = n;
            }
            __syncwarp();
            const int ntp = s.ntp;
            if (loc == PB_ON_INTERFACE) {
                double rx = 0, ry = 0, den = 0; const double x0 = s.tp[0], y0 = s.tp[1];
                for (int i = 2; i < ntp; i++) {
                    const double ax = s.tp[2 * (i - 1)] - x0, ay = s.tp[2 * (i - 1) + 1] - y0, bx = s.tp[2 * i] - x0, by = s.tp[2 * i + 1] - y0;
                    const double d = (ax * by - ay * bx) / 2.0; rx += (ax + bx) * d; ry += (ay + by) * d; den += d;
                }
                pbx = x0 + rx / (den * 3); pby = y0 + ry / (den * 3);
            }
            const int ri = dunavant_rule_index(2 * CD), np = c_qt.dun_n[ri], nq = ntp * np;
            double acc = 0.0;
            for (int q = lane; q < nq; q += 32) {
                const int t = q / np, p = q - t * np, t1 = (t + 1) % ntp;
                double ax = s.tp[2 * t], ay = s.tp[2 * t + 1], bx = s.tp[2 * t1], by = s.tp[2 * t1 + 1];
                const double* row = c_qt.dun[ri][p];
                double x, y, area;
                if (loc == PB_ON_INTERFACE) {      // triangles (bar, tp_i, tp_{i+1})
                    area = fabs(((ax - pbx) * (by - pby) - (ay - pby) * (bx - pbx)) / 2.0);
                    x = pbx * row[0] + ax * row[1] + bx * row[2]; y = pby * row[0] + ay * row[1] + by * row[2];
                } else {                            // integrate(poly cell): triangles (p_i, p_{i+1}, bar)
                    area = fabs(((bx - ax) * (pby - ay) - (by - ay) * (pbx - ax)) / 2.0);
                    x = ax * row[0] + bx * row[1] + pbx * row[2]; y = ay * row[0] + by * row[1] + pby * row[2];
                }
                double phi[R], dx[R], dy[R];
                eval_cell_basis<RD>((x - geo.barx) * ih, (y - geo.bary) * ih, ih, phi, dx, dy);
                double gx = 0, gy = 0;
#pragma unroll
                for (int i = 1; i < R; i++) if (i < C) { gx += s.u[i] * dx[i]; gy += s.u[i] * dy[i]; }
                double rgx, rgy; fun_grad(PB_FUN_SINSIN, x, y, rgx, rgy);
                acc += area * row[3] * ((rgx - gx) * (rgx - gx) + (rgy - gy) * (rgy - gy));
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
            sum0 += acc;
            __syncwarp();
            continue;
        }

        // ================= PB_PROBLEM_HHO (convergence_test.cpp:254-306)
        // face dofs of the cell: solution entries, or the L2 projection of the datum on Dirichlet faces (take_local_data,
        // hho.hpp:408-449)
        if (lane < 4) {
            const int li = lane;
            if (!cell_face_dirichlet(g, ci, cj, li)) {
                const long long cf = cell_face_compressed(g, ci, cj, li);
                for (int k = 0; k < F; k++) s.u[C + li * F + k] = A.xfull[cf * F + k - A.xoff];
            } else {
                double ax, ay, bx, by; face_points(geo, li, ax, ay, bx, by);
                double Mf[F * (F + 1) / 2], gd[F];
                for (int k = 0; k < F * (F + 1) / 2; k++) Mf[k] = 0.0;
                for (int k = 0; k < F; k++) gd[k] = 0.0;
                for (int q = 0; q < QF; q++) {
                    double x, y, w; seg_qp(QF, q, ax, ay, bx, by, x, y, w);
                    const double ex_ = bx - ax, ey_ = by - ay, len2 = ex_ * ex_ + ey_ * ey_, fbx = 0.5 * (ax + bx), fby = 0.5 * (ay + by);
                    const double ep = 4.0 * ((fbx - ax) * (x - fbx) + (fby - ay) * (y - fby)) / len2;
                    double ph[F]; ph[0] = 1.0;
                    for (int k = 1; k < F; k++) ph[k] = ph[k - 1] * ep;
                    const double bc = fun_sol(A.fun, x, y);
                    for (int k = 0; k < F; k++) { gd[k] += w * ph[k] * bc; for (int m = 0; m <= k; m++) Mf[PB_TRI(k, m)] += w * ph[k] * ph[m]; }
                }
                ldlt_packed<F>(Mf); ldlt_solve<F>(Mf, gd);
                for (int k = 0; k < F; k++) s.u[C + li * F + k] = gd[k];
            }
        }
        for (int i = lane; i < R * R; i += 32) s.stiff[i] = 0.0;
        for (int i = lane; i < C * C; i += 32) s.mass[i] = 0.0;
        for (int i = lane; i < R; i += 32) s.b[i] = 0.0;
        for (int i = lane; i < C; i += 32) s.proj[i] = 0.0;
        __syncwarp();
        // ---- stiffness, degree 2*rd (hho.hpp:52-59)
        {
            constexpr int nq = cell_qps(QK, 2 * RD);
            for (int base = 0; base < nq; base += 32) {
                const int q = base + lane;
                double* T = s.tab + lane * TW;
                if (q < nq) {
                    double x, y, w; cell_qp<QK, 2 * RD>(geo, q, x, y, w);
                    double phi[R], dx[R], dy[R];
                    eval_cell_basis<RD>((x - geo.barx) * ih, (y - geo.bary) * ih, ih, phi, dx, dy);
#pragma unroll
                    for (int r = 0; r < R; r++) { T[r] = dx[r]; T[R + r] = dy[r]; }
                    T[3 * R] = w;
                }
                __syncwarp();
                const int cnt = min(32, nq - base);
                for (int idx = lane; idx < R * R; idx += 32) {
                    const int jj = idx / R, ii = idx - jj * R; double acc = 0.0;
                    for (int p = 0; p < cnt; p++) { const double* U = s.tab + p * TW; acc += U[3 * R] * (U[ii] * U[jj] + U[R + ii] * U[R + jj]); }
                    s.stiff[idx] += acc;
                }
                __syncwarp();
            }
        }
        // (gr_rhs u)_r, r = 1..R-1: cell part, then the face terms with degree 2*fd (hho.hpp:64-85)
        for (int r = 1 + lane; r < R; r += 32) { double acc = 0.0; for (int c = 0; c < C; c++) acc += s.stiff[r + c * R] * s.u[c]; s.b[r] = acc; }
        __syncwarp();
        {
            double* T = s.tab + lane * TW;
            const int li = lane / QF, q = lane - li * QF;
            if (li < 4) {
                double ax, ay, bx, by; face_points(geo, li, ax, ay, bx, by);
                double nx, ny; face_normal(geo, li, nx, ny);
                double x, y, w; seg_qp(QF, q, ax, ay, bx, by, x, y, w);
                double phi[R], dx[R], dy[R];
                eval_cell_basis<RD>((x - geo.barx) * ih, (y - geo.bary) * ih, ih, phi, dx, dy);
                const double ex_ = bx - ax, ey_ = by - ay, len2 = ex_ * ex_ + ey_ * ey_, fbx = 0.5 * (ax + bx), fby = 0.5 * (ay + by);
                const double ep = 4.0 * ((fbx - ax) * (x - fbx) + (fby - ay) * (y - fby)) / len2;
                double uT = 0.0, uF = 0.0, fp = 1.0;
#pragma unroll
                for (int c = 0; c < C; c++) uT += s.u[c] * phi[c];
                for (int b = 0; b < F; b++) { uF += s.u[C + li * F + b] * fp; fp *= ep; }
#pragma unroll
                for (int r = 0; r < R; r++) T[r] = w * (dx[r] * nx + dy[r] * ny) * (uF - uT);
            }
            __syncwarp();
            for (int r = 1 + lane; r < R; r += 32) { double acc = 0.0; for (int p = 0; p < 4 * QF; p++) acc += s.tab[p * TW + r]; s.b[r] += acc; }
            __syncwarp();
        }
        // reconstruction dofs: gr_lhs = stiff[1:,1:] (hho.hpp:61, :92)
        warp_ldlt(s.stiff + 1 + R, M, R, lane);
        warp_ldlt_solve(s.stiff + 1 + R, M, R, s.b + 1, M, 1, lane);
        // ---- degree 2*cd points: mass matrix, projection of the exact solution, L2 and energy errors
        double l2 = 0.0, en = 0.0;
        constexpr int nq2 = cell_qps(QK, 2 * CD);
        for (int base = 0; base < nq2; base += 32) {
            const int q = base + lane;
            double* T = s.tab + lane * TW;
            if (q < nq2) {
                double x, y, w; cell_qp<QK, 2 * CD>(geo, q, x, y, w);
                double phi[R], dx[R], dy[R];
                eval_cell_basis<RD>((x - geo.barx) * ih, (y - geo.bary) * ih, ih, phi, dx, dy);
                double val = 0.0, gx = 0.0, gy = 0.0;
#pragma unroll
                for (int c = 0; c < C; c++) val += s.u[c] * phi[c];
#pragma unroll
                for (int i = 1; i < R; i++) { gx += s.b[i] * dx[i]; gy += s.b[i] * dy[i]; }
                const double rv = fun_sol(A.fun, x, y);
                double rgx, rgy; fun_grad(A.fun, x, y, rgx, rgy);
                l2 += w * (rv - val) * (rv - val);
                en += w * ((rgx - gx) * (rgx - gx) + (rgy - gy) * (rgy - gy));
#pragma unroll
                for (int c = 0; c < C; c++) T[c] = phi[c];
                T[3 * R] = w; T[3 * R + 1] = w * rv;
            }
            __syncwarp();
            const int cnt = min(32, nq2 - base);
            for (int idx = lane; idx < C * C; idx += 32) {
                const int jj = idx / C, ii = idx - jj * C; double acc = 0.0;
                for (int p = 0; p < cnt; p++) { const double* U = s.tab + p * TW; acc += U[3 * R] * U[ii] * U[jj]; }
                s.mass[idx] += acc;
            }
            for (int a = lane; a < C; a += 32) { double acc = 0.0; for (int p = 0; p < cnt; p++) { const double* U = s.tab + p * TW; acc += U[3 * R + 1] * U[a]; } s.proj[a] += acc; }
            __syncwarp();
        }
        for (int idx = lane; idx < C * C; idx += 32) s.massf[idx] = s.mass[idx];
        warp_ldlt(s.massf, C, C, lane);
        warp_ldlt_solve(s.massf, C, C, s.proj, C, 1, lane);
        double mm = 0.0;
        for (int i = lane; i < C; i += 32) {
            double acc = 0.0;
            for (int jx = 0; jx < C; jx++) acc += s.mass[i + jx * C] * (s.proj[jx] - s.u[jx]);
            mm += (s.proj[i] - s.u[i]) * acc;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { l2 += __shfl_down_sync(0xffffffffu, l2, o); en += __shfl_down_sync(0xffffffffu, en, o); mm += __shfl_down_sync(0xffffffffu, mm, o); }
        sum0 += l2; sum1 += mm * nq2; sum2 += en;
        __syncwarp();
    }
    if (lane == 0) { A.parts[3 * blockIdx.x] = sum0; A.parts[3 * blockIdx.x + 1] = sum1; A.parts[3 * blockIdx.x + 2] = sum2; }
}

// ---- solution_N_*_k_*.dat of the driver (convergence_test.cpp:250-252, 280-286): (x, y, u_T(x, y)) at the degree-2*cd
// points of every owned cell, cells ascending, points in integrate()'s order.  One thread per point.
inline long long sample_points_per_cell(int cd, int qk) { return qk == PB_QUAD_TENSOR_GAUSS ? cell_qps(PB_QUAD_TENSOR_GAUSS, 2 * cd) : cell_qps(PB_QUAD_POLY_DUNAVANT, 2 * cd); }
template <int CD, int QK>
__global__ void sample_solution_kernel(Grid g, const double2* __restrict__ points, const double* __restrict__ ucell, double* __restrict__ out) {
    constexpr int C = (CD + 1) * (CD + 2) / 2, nq = cell_qps(QK, 2 * CD);
    const long long total = (g.row_end - g.row_begin) * g.Nx * nq;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        const long long lc = t / nq; const int q = (int)(t - lc * nq);
        const long long ci = lc % g.Nx, cj = g.row_begin + lc / g.Nx, p0 = grid_point(g, ci, cj);
        const long long pid[4] = {p0, p0 + 1, p0 + g.Nx + 2, p0 + g.Nx + 1};
        CellGeom geo;
#pragma unroll
        for (int i = 0; i < 4; i++) { const double2 v = points[pid[i]]; geo.vx[i] = v.x; geo.vy[i] = v.y; }
        cell_geometry(geo);
        double x, y, w; cell_qp<QK, 2 * CD>(geo, q, x, y, w);
        double phi[C], dx[C], dy[C];
        eval_cell_basis<CD>((x - geo.barx) * geo.ihalf, (y - geo.bary) * geo.ihalf, geo.ihalf, phi, dx, dy);
        double val = 0.0;
#pragma unroll
        for (int c = 0; c < C; c++) val += ucell[lc * C + c] * phi[c];
        out[3 * t] = x; out[3 * t + 1] = y; out[3 * t + 2] = val;
    }
}

}  // namespace pb
