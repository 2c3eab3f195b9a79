// Generic local-operator kernel for ANY valid hho_degree_info(cd, fd) — in particular cd != rd, where the full HHO
// stabilisation (make_hho_fancy_stabilization, hho.hpp:155-237) does not collapse to the naive operator: the obstacle
// driver uses HHO(0, k) (apps/obstacle/obstacle.cpp:51, 249-253).  Literal restatement of
//   make_hho_laplacian            hho.hpp:32-96
//   make_hho_naive_stabilization  hho.hpp:99-148      (SK == 0)
//   make_hho_fancy_stabilization  hho.hpp:155-237     (SK == 1)
//   make_rhs(msh, cl, cd, f, di)  utils.hpp:153-174
// followed by the static condensation of K1.  One warp per cell, matrices in shared memory, runtime loops: this is the
// low-volume path (K1 is the tuned kernel for the HHO(k+1,k) drivers).
#pragma once
#include "pb_cut.cuh"

namespace pb {

template <int CD, int FD>
struct GenSmem {
    using Z = Sizes<CD, FD>;
    static constexpr int R = Z::R, C = Z::C, F = Z::F, N = Z::N, NF = Z::NF, M = Z::M;
    static constexpr int TW = 3 * R + F + 4;
    double tab[32 * TW];
    double stiff[R * R], mass[R * R];
    double G[M * N], X[M * N], LC[N * N];
    double P1[C * N], M1[C * C];
    double fm[F * F], fmf[F * F], ft[F * R], B2[F * N], B3[F * N];
    double KT[C * C], AC[(NF + 1) * C], Sb[NF * (NF + 1)], fT[C];
};

template <int CD, int FD, int QK, int SK>
__global__ void __launch_bounds__(32) local_generic_kernel(LocalArgs A, int rhs_di) {
    using Z = Sizes<CD, FD>;
    constexpr int C = Z::C, F = Z::F, R = Z::R, M = Z::M, N = Z::N, NF = Z::NF, RD = Z::RD, QF = Z::QF;
    using SM = GenSmem<CD, FD>;
    constexpr int TW = SM::TW;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SM& s = *reinterpret_cast<SM*>(smem_raw);
    const int lane = threadIdx.x;
    const Grid& g = A.grid;

    for (long long cell = blockIdx.x; cell < A.ncells; cell += gridDim.x) {
        int loc = PB_IN_NEGATIVE_SIDE;
        if (A.cell_loc) { loc = A.cell_loc[cell]; if (loc == PB_ON_INTERFACE) continue; }      // cut cells: K2 (uniform per warp)
        const bool has_load = !(A.problem == PB_PROBLEM_FICTDOM && loc != PB_IN_NEGATIVE_SIDE);   // cuthho_square.cpp:660-665
        const long long ci = cell % g.Nx, cj = g.row_begin + cell / g.Nx, p0 = grid_point(g, ci, cj);
        const long long pid[4] = {p0, p0 + 1, p0 + g.Nx + 2, p0 + g.Nx + 1};
        CellGeom geo;
#pragma unroll
        for (int i = 0; i < 4; i++) { const double2 v = A.points[pid[i]]; geo.vx[i] = v.x; geo.vy[i] = v.y; }
        cell_geometry(geo);
        const double ih = geo.ihalf;
        for (int i = lane; i < R * R; i += 32) { s.stiff[i] = 0.0; s.mass[i] = 0.0; }
        for (int i = lane; i < M * N; i += 32) s.G[i] = 0.0;
        for (int i = lane; i < C; i += 32) s.fT[i] = 0.0;
        __syncwarp();
        // ---- stiffness and mass of the rd basis, degree 2*rd (hho.hpp:52-59, 168-176)
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
                    for (int r = 0; r < R; r++) { T[r] = dx[r]; T[R + r] = dy[r]; T[2 * R + r] = phi[r]; }
                    T[3 * R] = w;
                }
                __syncwarp();
                const int cnt = min(32, nq - base);
                for (int idx = lane; idx < R * R; idx += 32) {
                    const int jj = idx / R, ii = idx - jj * R; double a1 = 0.0, a2 = 0.0;
                    for (int p = 0; p < cnt; p++) {
                        const double* U = s.tab + p * TW;
                        a1 += U[3 * R] * (U[ii] * U[jj] + U[R + ii] * U[R + jj]); a2 += U[3 * R] * U[2 * R + ii] * U[2 * R + jj];
                    }
                    s.stiff[idx] += a1; s.mass[idx] += a2;
                }
                __syncwarp();
            }
        }
        // ---- cell load, degree 2*(cd + di) (utils.hpp:153-174); the degree is a run-time value here
        {
            const int deg = 2 * (CD + rhs_di);
            const int nq = QK == PB_QUAD_TENSOR_GAUSS ? gauss_nodes(deg) * gauss_nodes(deg) : 4 * c_qt.dun_n[dunavant_rule_index(deg)];
            for (int base = 0; base < nq; base += 32) {
                const int q = base + lane;
                double* T = s.tab + lane * TW;
                if (q < nq) {
                    double x, y, w;
                    if (QK == PB_QUAD_TENSOR_GAUSS) {
                        const int n = gauss_nodes(deg), iy = q / n, ix = q - iy * n;
                        const double xi = c_qt.gx[n][ix], eta = c_qt.gx[n][iy];
                        const double a = (1 - xi) * (1 - eta), b = (1 + xi) * (1 - eta), c = (1 + xi) * (1 + eta), d = (1 - xi) * (1 + eta);
                        x = 0.25 * geo.vx[0] * a + 0.25 * geo.vx[1] * b + 0.25 * geo.vx[2] * c + 0.25 * geo.vx[3] * d;
                        y = 0.25 * geo.vy[0] * a + 0.25 * geo.vy[1] * b + 0.25 * geo.vy[2] * c + 0.25 * geo.vy[3] * d;
                        const double j11 = 0.25 * ((geo.vx[1] - geo.vx[0]) * (1 - eta) + (geo.vx[2] - geo.vx[3]) * (1 + eta));
                        const double j12 = 0.25 * ((geo.vy[1] - geo.vy[0]) * (1 - eta) + (geo.vy[2] - geo.vy[3]) * (1 + eta));
                        const double j21 = 0.25 * ((geo.vx[3] - geo.vx[0]) * (1 - xi) + (geo.vx[2] - geo.vx[1]) * (1 + xi));
                        const double j22 = 0.25 * ((geo.vy[3] - geo.vy[0]) * (1 - xi) + (geo.vy[2] - geo.vy[1]) * (1 + xi));
                        w = c_qt.gw[n][ix] * c_qt.gw[n][iy] * fabs(j11 * j22 - j12 * j21);
                    } else {
                        const int ri = dunavant_rule_index(deg), np = c_qt.dun_n[ri], t = q / np, p = q - t * np;
                        const double p0x = sel4(geo.vx, t), p0y = sel4(geo.vy, t), p1x = sel4(geo.vx, (t + 1) & 3), p1y = sel4(geo.vy, (t + 1) & 3);
                        const double area = fabs(((p1x - p0x) * (geo.bary - p0y) - (p1y - p0y) * (geo.barx - p0x)) / 2.0);
                        const double* row = c_qt.dun[ri][p];
                        x = p0x * row[0] + p1x * row[1] + geo.barx * row[2]; y = p0y * row[0] + p1y * row[1] + geo.bary * row[2]; w = area * row[3];
                    }
                    double phi[R], dx[R], dy[R];
                    eval_cell_basis<RD>((x - geo.barx) * ih, (y - geo.bary) * ih, ih, phi, dx, dy);
                    const double wf = has_load ? w * fun_rhs(A.fun, x, y) : 0.0;
#pragma unroll
                    for (int c = 0; c < C; c++) T[c] = wf * phi[c];
                }
                __syncwarp();
                const int cnt = min(32, nq - base);
                for (int a = lane; a < C; a += 32) { double acc = 0.0; for (int p = 0; p < cnt; p++) acc += s.tab[p * TW + a]; s.fT[a] += acc; }
                __syncwarp();
            }
        }
        // ---- face tables, degree 2*fd: phi (R), grad phi . n (R), face basis (F), w; lanes = (face, node)
        {
            double* T = s.tab + lane * TW;
            const int li = lane / QF, q = lane - li * QF;
            if (li < 4) {
                double ax, ay, bx, by; face_points(geo, li, ax, ay, bx, by);
                double nx, ny; face_normal(geo, li, nx, ny);
                double x, y, w; seg_qp(QF, q, ax, ay, bx, by, x, y, w);
                double phi[R], dx[R], dy[R];
                eval_cell_basis<RD>((x - geo.barx) * ih, (y - geo.bary) * ih, ih, phi, dx, dy);
#pragma unroll
                for (int r = 0; r < R; r++) { T[r] = phi[r]; T[R + r] = dx[r] * nx + dy[r] * ny; }
                const double ex_ = bx - ax, ey_ = by - ay, len2 = ex_ * ex_ + ey_ * ey_, fbx = 0.5 * (ax + bx), fby = 0.5 * (ay + by);
                const double ep = 4.0 * ((fbx - ax) * (x - fbx) + (fby - ay) * (y - fby)) / len2;
                double fp = 1.0;
#pragma unroll
                for (int b = 0; b < F; b++) { T[2 * R + b] = fp; fp *= ep; }
                T[3 * R] = w; T[3 * R + 1] = cell_face_dirichlet(g, ci, cj, li) ? w * fun_sol(A.fun, x, y) : 0.0;
            }
            __syncwarp();
        }
        // ---- gr_rhs (M x N, rows r = 1..R-1 of the rd basis), hho.hpp:64-85
        for (int idx = lane; idx < M * N; idx += 32) {
            const int col = idx / M, r = idx - col * M + 1;
            double acc = 0.0;
            if (col < C) {
                acc = s.stiff[r + col * R];
                for (int p = 0; p < 4 * QF; p++) { const double* U = s.tab + p * TW; acc -= U[3 * R] * U[R + r] * U[col]; }
            } else {
                const int fi = (col - C) / F, b = (col - C) - fi * F;
                for (int p = fi * QF; p < (fi + 1) * QF; p++) { const double* U = s.tab + p * TW; acc += U[3 * R] * U[R + r] * U[2 * R + b]; }
            }
            s.G[idx] = acc; s.X[idx] = acc;
        }
        __syncwarp();
        // oper = llt(gr_lhs).solve(gr_rhs), data = gr_rhs^T oper (hho.hpp:92-93); the factor overwrites stiff[1:,1:]
        warp_ldlt(s.stiff + 1 + R, M, R, lane);
        warp_ldlt_solve(s.stiff + 1 + R, M, R, s.X, M, N, lane);
        for (int idx = lane; idx < N * N; idx += 32) {
            const int j = idx / N, i = idx - j * N; double acc = 0.0;
            for (int r = 0; r < M; r++) acc += s.G[r + i * M] * s.X[r + j * M];
            s.LC[idx] = acc;
        }
        __syncwarp();
        if (A.data) for (int idx = lane; idx < N * N; idx += 32) A.data[cell * (long long)(N * N) + idx] = s.LC[idx];      // gr.second before the stabilisation
        // ---- stabilisation
        const double hdiam = geo.h, harea = geo.area;
        if (SK == 1) {
            // proj1 = I_C|0 - M1^{-1} M2 rec  (C x N), M1 = mass[:C,:C], M2 = mass[:C, 1:R]  (hho.hpp:178-189)
            for (int idx = lane; idx < C * C; idx += 32) { const int j = idx / C, i = idx - j * C; s.M1[idx] = s.mass[i + j * R]; }
            for (int idx = lane; idx < C * N; idx += 32) {
                const int col = idx / C, i = idx - col * C; double acc = 0.0;
                for (int r = 0; r < M; r++) acc += s.mass[i + (r + 1) * R] * s.X[r + col * M];
                s.P1[idx] = acc;
            }
            warp_ldlt(s.M1, C, C, lane);
            warp_ldlt_solve(s.M1, C, C, s.P1, C, N, lane);
            for (int idx = lane; idx < C * N; idx += 32) { const int col = idx / C, i = idx - col * C; s.P1[idx] = (i == col ? 1.0 : 0.0) - s.P1[idx]; }
            __syncwarp();
        }
        for (int fi = 0; fi < 4; fi++) {
            // face mass and trace on the rd basis, degree 2*fd (hho.hpp:127-135, 206-216)
            for (int idx = lane; idx < F * F; idx += 32) {
                const int b = idx / F, a = idx - b * F; double acc = 0.0;
                for (int p = fi * QF; p < (fi + 1) * QF; p++) { const double* U = s.tab + p * TW; acc += U[3 * R] * U[2 * R + a] * U[2 * R + b]; }
                s.fm[idx] = acc; s.fmf[idx] = acc;
            }
            for (int idx = lane; idx < F * R; idx += 32) {
                const int c = idx / F, a = idx - c * F; double acc = 0.0;
                for (int p = fi * QF; p < (fi + 1) * QF; p++) { const double* U = s.tab + p * TW; acc += U[3 * R] * U[2 * R + a] * U[c]; }
                s.ft[idx] = acc;
            }
            __syncwarp();
            warp_ldlt(s.fmf, F, F, lane);
            if (SK == 1) {
                // BRF = piKF(MR1 rec) - I_F + piKF(MR2 proj1)   (hho.hpp:218-231)
                for (int idx = lane; idx < F * N; idx += 32) {
                    const int col = idx / F, a = idx - col * F; double a2 = 0.0, a3 = 0.0;
                    for (int r = 0; r < M; r++) a2 += s.ft[a + (r + 1) * F] * s.X[r + col * M];
                    for (int c = 0; c < C; c++) a3 += s.ft[a + c * F] * s.P1[c + col * C];
                    s.B2[idx] = a2; s.B3[idx] = a3;
                }
                __syncwarp();
                warp_ldlt_solve(s.fmf, F, F, s.B2, F, N, lane);
                warp_ldlt_solve(s.fmf, F, F, s.B3, F, N, lane);
                for (int idx = lane; idx < F * N; idx += 32) {
                    const int col = idx / F, a = idx - col * F;
                    s.B2[idx] = s.B2[idx] - (col == C + fi * F + a ? 1.0 : 0.0) + s.B3[idx];
                }
            } else {
                // oper = [M^{-1} trace_C | ... -I_F ...]   (hho.hpp:137-143)
                for (int idx = lane; idx < F * N; idx += 32) { const int col = idx / F, a = idx - col * F; s.B2[idx] = col < C ? s.ft[a + col * F] : 0.0; }
                __syncwarp();
                warp_ldlt_solve(s.fmf, F, F, s.B2, F, C, lane);
                for (int idx = lane; idx < F * N; idx += 32) { const int col = idx / F, a = idx - col * F; if (col >= C) s.B2[idx] = (col == C + fi * F + a) ? -1.0 : 0.0; }
            }
            __syncwarp();
            // data += (1/h) BRF^T (fmass BRF): h = diam(T) (fancy) or area(T) (naive), SURVEY F8
            for (int idx = lane; idx < F * N; idx += 32) {
                const int col = idx / F, a = idx - col * F; double acc = 0.0;
                for (int b = 0; b < F; b++) acc += s.fm[a + b * F] * s.B2[b + col * F];
                s.B3[idx] = acc;
            }
            __syncwarp();
            const double sc = 1.0 / (SK == 1 ? hdiam : harea);
            for (int idx = lane; idx < N * N; idx += 32) {
                const int j = idx / N, i = idx - j * N; double acc = 0.0;
                for (int a = 0; a < F; a++) acc += s.B2[a + i * F] * s.B3[a + j * F];
                s.LC[idx] += sc * acc;
            }
            __syncwarp();
        }
        if (A.lc) {
            for (int idx = lane; idx < N * N; idx += 32) A.lc[cell * (long long)(N * N) + idx] = s.LC[idx];
            if (A.oper) for (int idx = lane; idx < N * R; idx += 32) { const int col = idx / R, r = idx - col * R; A.oper[cell * (long long)(N * R) + idx] = r < M ? s.X[r + col * M] : 0.0; }
            if (A.rhs) for (int a = lane; a < C; a += 32) A.rhs[cell * (long long)C + a] = s.fT[a];
        }
        // ---- static condensation (as K1 phase 6) and Dirichlet lifting (hho.hpp:381-403)
        for (int idx = lane; idx < C * C; idx += 32) { const int j = idx / C, i = idx - j * C; s.KT[idx] = s.LC[i + j * N]; }
        for (int idx = lane; idx < (NF + 1) * C; idx += 32) { const int col = idx / C, a = idx - col * C; s.AC[idx] = col < NF ? s.LC[a + (C + col) * N] : s.fT[a]; }
        warp_ldlt(s.KT, C, C, lane);
        warp_ldlt_solve(s.KT, C, C, s.AC, C, NF + 1, lane);
        for (int idx = lane; idx < NF * (NF + 1); idx += 32) {
            const int col = idx / NF, i = idx - col * NF;
            double v = col < NF ? s.LC[(C + i) + (C + col) * N] : 0.0;
            for (int a = 0; a < C; a++) v -= s.LC[a + (C + i) * N] * s.AC[a + col * C];
            s.Sb[idx] = v;
        }
        __syncwarp();
        for (int li = 0; li < 4; li++) {
            if (!cell_face_dirichlet(g, ci, cj, li)) continue;
            // face mass of this face again (fm holds face 3): projection of the datum with the tables still in `tab`
            if (lane == 0) {
                double Mf[F * (F + 1) / 2], gd[F];
                for (int k = 0; k < F * (F + 1) / 2; k++) Mf[k] = 0.0;
                for (int k = 0; k < F; k++) gd[k] = 0.0;
                for (int p = li * QF; p < (li + 1) * QF; p++) {
                    const double* U = s.tab + p * TW;
                    for (int k = 0; k < F; k++) { gd[k] += U[3 * R + 1] * U[2 * R + k]; for (int m = 0; m <= k; m++) Mf[PB_TRI(k, m)] += U[3 * R] * U[2 * R + k] * U[2 * R + m]; }
                }
                ldlt_packed<F>(Mf); ldlt_solve<F>(Mf, gd);
                for (int i = 0; i < NF; i++) for (int a = 0; a < F; a++) s.Sb[i + NF * NF] -= s.Sb[i + (li * F + a) * NF] * gd[a];
            }
            __syncwarp();
        }
        for (int idx = lane; idx < NF * NF; idx += 32) A.S[cell * (long long)(NF * NF) + idx] = s.Sb[idx];
        for (int i = lane; i < NF; i += 32) A.g[cell * (long long)NF + i] = s.Sb[NF * NF + i];
        for (int idx = lane; idx < NF * C; idx += 32) A.recA[cell * (long long)(NF * C) + idx] = s.AC[idx];
        for (int a = lane; a < C; a += 32) A.recb[cell * (long long)C + a] = s.AC[NF * C + a];
        __syncwarp();
    }
}

}  // namespace pb
