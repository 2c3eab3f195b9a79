// K2i — local operator of one CUT cell of the two-material INTERFACE problem (`cuthho_square -i`), followed
// by the static condensation of its 2*cbs cell unknowns.  Restates
//   make_hho_laplacian_interface(msh, cl, level_set, hdi, parms)   apps/cuthho/cuthho_square.cpp:390-502
//   make_hho_cut_stabilization(msh, cl, hdi, where) on both sides   :566-621, combined as in run_cuthho_interface :1689-1706
//   make_rhs(msh, cl, degree, where, f) on both sides               src/methods/cuthho_bits/cuthho_utils.hpp:65-86
// Local unknown order as in the driver: [cell-, cell+, faces-, faces+]  (2*cbs + 2*4*fbs).
// gr_lhs is singular along the common constant (1 on both sides): the reference hands it to Eigen's pivoted ldlt; every
// column of gr_rhs is orthogonal to that vector, so data = gr_rhs^T gr_lhs^+ gr_rhs is what any solution gives and the
// kernel factorises gr_lhs + v v^T, v = e_0 + e_rbs (positive definite, same data).
// Same style as K2 (pb_cut.cuh): one warp per cut cell, matrices in shared memory, runtime loops.
#pragma once
#include "pb_cut.cuh"

namespace pb {

struct IfcArgs {
    Grid grid;
    LevelSet ls;
    const double2* points;
    const int* node_loc; const int* face_loc; const double2* face_ip;
    const int* cut_cells; const double2* iface;
    int ncut, NI;
    int fun; double eta, kappa1, kappa2;
    double *Sx, *gx, *recAx, *recbx;     // per cut slot: (2NF)^2, 2NF, 2C x 2NF, 2C
    double* lcx;                         // test hook: per cut slot (2N)^2, or nullptr
    double* fx;                          // test hook: per cut slot 2C
};

template <int CD, int FD>
struct IfcSmem {
    using Z = Sizes<CD, FD>;
    static constexpr int R = Z::R, C = Z::C, F = Z::F, NF = Z::NF;
    static constexpr int R2 = 2 * R, C2 = 2 * C, NF2 = 2 * NF, N2 = C2 + NF2;
    double tp[2 * CUT_MAX_TP];
    double tab[32 * (3 * R + F + 4)];
    double stiff[R2 * R2];
    double G[R2 * N2];
    double X[R2 * N2];
    double LC[N2 * N2];
    double KT[C2 * C2];
    double AC[(NF2 + 1) * C2];
    double Sb[NF2 * (NF2 + 1)];
    double fT[C2];
    double Mf[F * F], Tr[F * C], Pm[F * C];
    double seg[4 * 4];
    int seg_on[4];
    int ntp;
};

template <int CD, int FD>
__global__ void __launch_bounds__(32) local_cut_interface_kernel(IfcArgs A) {
    using Z = Sizes<CD, FD>;
    constexpr int C = Z::C, F = Z::F, R = Z::R, NF = Z::NF, RD = Z::RD;
    using SM = IfcSmem<CD, FD>;
    constexpr int R2 = SM::R2, C2 = SM::C2, NF2 = SM::NF2, N2 = SM::N2;
    constexpr int TW = 3 * R + F + 4;
    static_assert(C == R, "the driver uses HHO(k+1,k): cell degree = reconstruction degree (cuthho_square.cpp:1662)");
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SM& s = *reinterpret_cast<SM*>(smem_raw);
    const int lane = threadIdx.x;
    const Grid& g = A.grid;

    for (int slot = blockIdx.x; slot < A.ncut; slot += gridDim.x) {
        const long long cell = A.cut_cells[slot];
        const long long ci = cell % g.Nx, cj = cell / g.Nx, p0 = grid_point(g, ci, cj);
        const long long pid[4] = {p0, p0 + 1, p0 + g.Nx + 2, p0 + g.Nx + 1};
        CellGeom geo; int nloc[4];
#pragma unroll
        for (int i = 0; i < 4; i++) { const double2 v = A.points[pid[i]]; geo.vx[i] = v.x; geo.vy[i] = v.y; nloc[i] = A.node_loc[pid[i]]; }
        cell_geometry(geo);
        const double ih = geo.ihalf, hT = geo.area;
        long long fids[4]; cell_faces(g, ci, cj, fids);
        const double2* I = A.iface + (long long)slot * A.NI;

        for (int i = lane; i < R2 * R2; i += 32) s.stiff[i] = 0.0;
        for (int i = lane; i < R2 * N2; i += 32) s.G[i] = 0.0;
        for (int i = lane; i < N2 * N2; i += 32) s.LC[i] = 0.0;
        for (int i = lane; i < C2; i += 32) s.fT[i] = 0.0;
        __syncwarp();

        double int_sign = 1.0;
        // ================= per side: polygon, stiffness block, load; the face terms and the stabilisation follow below
        for (int sd = 0; sd < 2; sd++) {
            const int WHERE = sd == 0 ? PB_IN_NEGATIVE_SIDE : PB_IN_POSITIVE_SIDE;
            const double kap = sd == 0 ? A.kappa1 : A.kappa2;
            __syncwarp();
            if (lane == 0) {       // collect_triangulation_points (cuthho_geom.hpp:675-728): the positive part takes the interface reversed
                int n = 0;
                auto push = [&](double x, double y) { s.tp[2 * n] = x; s.tp[2 * n + 1] = y; n++; };
                auto push_iface = [&]() {
                    if (sd == 0) for (int i = 0; i < A.NI; i++) push(I[i].x, I[i].y);
                    else for (int i = A.NI - 1; i >= 0; i--) push(I[i].x, I[i].y);
                };
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
            }
            __syncwarp();
            const int ntp = s.ntp;
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
            if (sd == 0) {        // sign of integrate_interface(msh, cl, deg, IN_NEGATIVE_SIDE) (cuthho_geom.hpp:851-895)
                const double vax = I[0].x - pbx, vay = I[0].y - pby, vbtx = I[1].x - I[0].x, vbty = I[1].y - I[0].y;
                int_sign = (vax * vbty + vay * (-vbtx)) < 0 ? -1.0 : 1.0;
            }
            auto cell_where_qp = [&](int ri, int np, int q, double& x, double& y, double& w) {
                const int t = q / np, p = q - t * np, t1 = (t + 1) % ntp;
                const double ax = s.tp[2 * t], ay = s.tp[2 * t + 1], bx = s.tp[2 * t1], by = s.tp[2 * t1 + 1];
                const double v0x = ax - pbx, v0y = ay - pby, v1x = bx - pbx, v1y = by - pby;
                const double area = fabs((v0x * v1y - v0y * v1x) / 2.0);
                const double* row = c_qt.dun[ri][p];
                x = pbx * row[0] + ax * row[1] + bx * row[2]; y = pby * row[0] + ay * row[1] + by * row[2];
                w = area * row[3];
            };
            // ---- kappa * stiffness over this part, degree 2*rd (cuthho_square.cpp:419-431) -> diagonal block sd
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
                        T[3 * R] = kap * w;
                    }
                    __syncwarp();
                    const int cnt = min(32, nq - base);
                    for (int idx = lane; idx < R * R; idx += 32) {
                        const int jj = idx / R, ii = idx - jj * R;
                        double acc = 0.0;
                        for (int p = 0; p < cnt; p++) { const double* U = s.tab + p * TW; acc += U[3 * R] * (U[ii] * U[jj] + U[R + ii] * U[R + jj]); }
                        s.stiff[(sd * R + ii) + (sd * R + jj) * R2] += acc;
                    }
                    __syncwarp();
                }
            }
            // ---- load of this part, degree 2*cd (cuthho_utils.hpp:65-86)
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
                    for (int a = lane; a < C; a += 32) { double acc = 0.0; for (int p = 0; p < cnt; p++) acc += s.tab[p * TW + a]; s.fT[sd * C + a] += acc; }
                    __syncwarp();
                }
            }
        }
        // ================= interface terms (cuthho_square.cpp:436-460), degree 2*rd, orientation of the negative side
        {
            const int ng = gauss_nodes(2 * RD), nq = (A.NI - 1) * ng;
            for (int base = 0; base < nq; base += 32) {
                const int q = base + lane;
                double* T = s.tab + lane * TW;
                if (q < nq) {
                    const int sg = q / ng, p = q - sg * ng;
                    double x, y, w; seg_qp(ng, p, I[sg].x, I[sg].y, I[sg + 1].x, I[sg + 1].y, x, y, w);
                    w *= int_sign * A.kappa1;
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
                    double a = 0.0, b = 0.0, c = 0.0;
                    for (int p = 0; p < cnt; p++) {
                        const double* U = s.tab + p * TW;
                        a += U[3 * R] * U[ii] * U[R + jj];
                        b += U[3 * R] * U[R + ii] * U[jj];
                        c += U[3 * R] * U[ii] * U[jj] * A.eta / hT;
                    }
                    s.stiff[ii + jj * R2] += -a - b + c;
                    s.stiff[(R + ii) + jj * R2] += a - c;
                    s.stiff[ii + (R + jj) * R2] += b - c;
                    s.stiff[(R + ii) + (R + jj) * R2] += c;
                }
                __syncwarp();
            }
        }
        // gr_rhs(:, 0:C) = stiff(:, 0:C), gr_rhs(:, C:2C) = stiff(:, R:R+C)   (:463-464)
        for (int idx = lane; idx < R2 * C2; idx += 32) {
            const int col = idx / R2, r = idx - col * R2;
            s.G[r + col * R2] = s.stiff[r + (col < C ? col : R + (col - C)) * R2];
        }
        __syncwarp();
        // ================= per side: face terms of gr_rhs (:466-497) and the cut stabilisation (:566-621, :1689-1706)
        for (int sd = 0; sd < 2; sd++) {
            const int WHERE = sd == 0 ? PB_IN_NEGATIVE_SIDE : PB_IN_POSITIVE_SIDE;
            const double kap = sd == 0 ? A.kappa1 : A.kappa2;
            __syncwarp();
            if (lane == 0) {       // points_where / integrate(fc, deg, where) (cuthho_geom.hpp:547-569, 817-849)
                for (int li = 0; li < 4; li++) {
                    const int fl = A.face_loc[fids[li]];
                    double ax, ay, bx, by; face_points(geo, li, ax, ay, bx, by);
                    int on = 0;
                    if (fl == WHERE) on = 1;
                    else if (fl == PB_ON_INTERFACE) {
                        on = 1;
                        int i0 = li, i1 = (li + 1) & 3; if (li >= 2) { int t = i0; i0 = i1; i1 = t; }
                        const double2 ip = A.face_ip[fids[li]];
                        if (nloc[i0] == WHERE && nloc[i1] != WHERE) { bx = ip.x; by = ip.y; }
                        else if (nloc[i0] != WHERE && nloc[i1] == WHERE) { ax = ip.x; ay = ip.y; }
                        else on = 0;
                    }
                    s.seg_on[li] = on; s.seg[4 * li] = ax; s.seg[4 * li + 1] = ay; s.seg[4 * li + 2] = bx; s.seg[4 * li + 3] = by;
                }
            }
            __syncwarp();
            {
                const int ng = gauss_nodes(2 * RD);
                const int q = lane;
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
                    T[3 * R] = kap * w;
                }
                __syncwarp();
                // rows sd*R + r; columns: cell block of this side, face block of this side
                for (int idx = lane; idx < R * (C + NF); idx += 32) {
                    const int col = idx / R, r = idx - col * R;
                    double acc = 0.0;
                    int gcol;
                    if (col < C) {
                        for (int pp = 0; pp < 4 * ng; pp++) { const double* U = s.tab + pp * TW; acc -= U[3 * R] * U[R + r] * U[col]; }
                        gcol = sd * C + col;
                    } else {
                        const int fi = (col - C) / F, b = (col - C) - fi * F;
                        for (int pp = fi * ng; pp < (fi + 1) * ng; pp++) { const double* U = s.tab + pp * TW; acc += U[3 * R] * U[R + r] * U[2 * R + b]; }
                        gcol = C2 + sd * NF + (col - C);
                    }
                    s.G[(sd * R + r) + gcol * R2] += acc;
                }
                __syncwarp();
            }
        }
        // ================= oper = (gr_lhs + v v^T)^{-1} gr_rhs, data = gr_rhs^T oper (:499-500)
        if (lane == 0) { s.stiff[0] += 1.0; s.stiff[R] += 1.0; s.stiff[R * R2] += 1.0; s.stiff[R + R * R2] += 1.0; }
        for (int idx = lane; idx < R2 * N2; idx += 32) s.X[idx] = s.G[idx];
        __syncwarp();
        warp_ldlt(s.stiff, R2, R2, lane);
        warp_ldlt_solve(s.stiff, R2, R2, s.X, R2, N2, lane);
        for (int idx = lane; idx < N2 * N2; idx += 32) {
            const int j = idx / N2, i = idx - j * N2;
            double acc = 0.0;
            for (int r = 0; r < R2; r++) acc += s.G[r + i * R2] * s.X[r + j * R2];
            s.LC[idx] = acc;
        }
        __syncwarp();
        // ================= stabilisation of both sides
        for (int sd = 0; sd < 2; sd++) {
            const int WHERE = sd == 0 ? PB_IN_NEGATIVE_SIDE : PB_IN_POSITIVE_SIDE;
            const double kap = sd == 0 ? A.kappa1 : A.kappa2;
            __syncwarp();
            if (lane == 0) {
                for (int li = 0; li < 4; li++) {
                    const int fl = A.face_loc[fids[li]];
                    double ax, ay, bx, by; face_points(geo, li, ax, ay, bx, by);
                    int on = 0;
                    if (fl == WHERE) on = 1;
                    else if (fl == PB_ON_INTERFACE) {
                        on = 1;
                        int i0 = li, i1 = (li + 1) & 3; if (li >= 2) { int t = i0; i0 = i1; i1 = t; }
                        const double2 ip = A.face_ip[fids[li]];
                        if (nloc[i0] == WHERE && nloc[i1] != WHERE) { bx = ip.x; by = ip.y; }
                        else if (nloc[i0] != WHERE && nloc[i1] == WHERE) { ax = ip.x; ay = ip.y; }
                        else on = 0;
                    }
                    s.seg_on[li] = on; s.seg[4 * li] = ax; s.seg[4 * li + 1] = ay; s.seg[4 * li + 2] = bx; s.seg[4 * li + 3] = by;
                }
            }
            __syncwarp();
            for (int li = 0; li < 4; li++) {
                if (!s.seg_on[li]) continue;
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
                double* Mfac = s.tab + 32 * TW - F * F;
                for (int idx = lane; idx < F * F; idx += 32) Mfac[idx] = s.Mf[idx];
                warp_ldlt(Mfac, F, F, lane);
                warp_ldlt_solve(Mfac, F, F, s.Pm, F, C, lane);
                // data += kappa/hT oper^T (M oper) on the one-sided unknowns [cell_sd | faces_sd]
                constexpr int N1 = C + NF;
                for (int idx = lane; idx < N1 * N1; idx += 32) {
                    const int j = idx / N1, i = idx - j * N1;
                    const bool ic = i < C, jc = j < C, iface_i = !ic && (i - C) / F == li, iface_j = !jc && (j - C) / F == li;
                    if (!(ic || iface_i) || !(jc || iface_j)) continue;
                    double acc = 0.0;
                    for (int a = 0; a < F; a++) {
                        const double oi = ic ? s.Pm[a + i * F] : ((i - C) - li * F == a ? -1.0 : 0.0);
                        if (oi == 0.0 && !ic) continue;
                        double mo = 0.0;
                        for (int b = 0; b < F; b++) {
                            const double oj = jc ? s.Pm[b + j * F] : ((j - C) - li * F == b ? -1.0 : 0.0);
                            mo += s.Mf[a + b * F] * oj;
                        }
                        acc += oi * mo;
                    }
                    const int gi = ic ? sd * C + i : C2 + sd * NF + (i - C), gj = jc ? sd * C + j : C2 + sd * NF + (j - C);
                    s.LC[gi + gj * N2] += kap * acc / hT;
                }
                __syncwarp();
            }
        }
        if (A.lcx) {
            for (int idx = lane; idx < N2 * N2; idx += 32) A.lcx[slot * (long long)(N2 * N2) + idx] = s.LC[idx];
            for (int a = lane; a < C2; a += 32) A.fx[slot * (long long)C2 + a] = s.fT[a];
        }
        // ================= static condensation of the 2*cbs cell unknowns
        for (int idx = lane; idx < C2 * C2; idx += 32) { const int j = idx / C2, i = idx - j * C2; s.KT[idx] = s.LC[i + j * N2]; }
        for (int idx = lane; idx < (NF2 + 1) * C2; idx += 32) {
            const int col = idx / C2, a = idx - col * C2;
            s.AC[idx] = col < NF2 ? s.LC[a + (C2 + col) * N2] : s.fT[a];
        }
        __syncwarp();
        warp_ldlt(s.KT, C2, C2, lane);
        warp_ldlt_solve(s.KT, C2, C2, s.AC, C2, NF2 + 1, lane);
        for (int idx = lane; idx < NF2 * (NF2 + 1); idx += 32) {
            const int col = idx / NF2, i = idx - col * NF2;
            double v = col < NF2 ? s.LC[(C2 + i) + (C2 + col) * N2] : 0.0;
            for (int a = 0; a < C2; a++) v -= s.LC[a + (C2 + i) * N2] * s.AC[a + col * C2];
            s.Sb[idx] = v;
        }
        __syncwarp();
        for (int idx = lane; idx < NF2 * NF2; idx += 32) A.Sx[slot * (long long)(NF2 * NF2) + idx] = s.Sb[idx];
        for (int i = lane; i < NF2; i += 32) A.gx[slot * (long long)NF2 + i] = s.Sb[NF2 * NF2 + i];
        for (int idx = lane; idx < NF2 * C2; idx += 32) A.recAx[slot * (long long)(NF2 * C2) + idx] = s.AC[idx];
        for (int a = lane; a < C2; a += 32) A.recbx[slot * (long long)C2 + a] = s.AC[NF2 * C2 + a];
        __syncwarp();
    }
}

}  // namespace pb
