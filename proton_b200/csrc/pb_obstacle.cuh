// Obstacle problem: primal-dual active-set loop of apps/obstacle/obstacle.cpp:119-198 on the CONDENSED
// SPD face system.  The reference assembles, per outer iteration, the saddle system of obstacle_assembler
// (src/methods/hho_bits/hho.hpp:471-751: all cell rows, then face rows; the cell column of an active cell is replaced
// by its multiplier) and solves it with SparseLU.  With cell degree 0 (one dof per cell, the only case the reference
// supports, hho.hpp:631) the cell rows eliminate in closed form:
//   inactive cell  u_T = (f_T - K_TF u_F) / K_TT,  lambda_T = 0   ->  S = K_FF - K_FT K_TF / K_TT,  g = -K_FT f_T / K_TT
//   active cell    u_T = gamma_T,  lambda_T = f_T - K_TT gamma_T - K_TF u_F   ->  S = K_FF,  g = -K_FT gamma_T
// so every outer iteration is: condense (this file) -> scatter (K3) -> PCG (K4/K5) -> update (this file).  The local
// matrices lc = gr.second + stab are computed once (pb_generic.cuh) and cached; the reference recomputes them in every
// outer iteration (obstacle.cpp:147-158).
#pragma once
#include "pb_generic.cuh"

namespace pb {

// obstacle_fun (obstacle.cpp:83-89): the obstacle is the zero function
PB_D double fun_obstacle_psi(double, double) { return 0.0; }

// face L2 projection of a function on local face li, quadrature degree 2*(fd+di) (make_mass_matrix / make_rhs of a
// face, utils.hpp:133-151, 176-199); returns the fd+1 coefficients
template <int FD>
PB_D void project_face(const CellGeom& geo, int li, int fun, int di, double* out) {
    constexpr int F = FD + 1;
    const int ng = gauss_nodes(2 * (FD + di));
    double ax, ay, bx, by; face_points(geo, li, ax, ay, bx, by);
    double Mf[F * (F + 1) / 2], gd[F];
    for (int k = 0; k < F * (F + 1) / 2; k++) Mf[k] = 0.0;
    for (int k = 0; k < F; k++) gd[k] = 0.0;
    for (int q = 0; q < ng; q++) {
        double x, y, w; seg_qp(ng, q, ax, ay, bx, by, x, y, w);
        const double ex_ = bx - ax, ey_ = by - ay, len2 = ex_ * ex_ + ey_ * ey_, fbx = 0.5 * (ax + bx), fby = 0.5 * (ay + by);
        const double ep = 4.0 * ((fbx - ax) * (x - fbx) + (fby - ay) * (y - fby)) / len2;
        double ph[F]; ph[0] = 1.0;
        for (int k = 1; k < F; k++) ph[k] = ph[k - 1] * ep;
        const double bc = fun_sol(fun, x, y);
        for (int k = 0; k < F; k++) { gd[k] += w * ph[k] * bc; for (int m = 0; m <= k; m++) Mf[PB_TRI(k, m)] += w * ph[k] * ph[m]; }
    }
    ldlt_packed<F>(Mf); ldlt_solve<F>(Mf, gd);
    for (int k = 0; k < F; k++) out[k] = gd[k];
}

PB_D void load_cell_geom(const Grid& g, const double2* points, long long ci, long long cj, CellGeom& geo) {
    const long long p0 = grid_point(g, ci, cj);
    const double2 a = points[p0], b = points[p0 + 1], c = points[p0 + g.Nx + 2], d = points[p0 + g.Nx + 1];
    geo.vx[0] = a.x; geo.vy[0] = a.y; geo.vx[1] = b.x; geo.vy[1] = b.y; geo.vx[2] = c.x; geo.vy[2] = c.y; geo.vx[3] = d.x; geo.vy[3] = d.y;
    cell_geometry(geo);
}

// alpha = 0, beta = 1, gamma = obstacle_fun(barycenter) (obstacle.cpp:98-113)
__global__ void obstacle_init_kernel(Grid g, long long ncells, const double2* __restrict__ points, double* alpha, double* beta, double* gamma) {
    for (long long c = blockIdx.x * (long long)blockDim.x + threadIdx.x; c < ncells; c += (long long)gridDim.x * blockDim.x) {
        CellGeom geo; load_cell_geom(g, points, c % g.Nx, g.row_begin + c / g.Nx, geo);
        alpha[c] = 0.0; beta[c] = 1.0; gamma[c] = fun_obstacle_psi(geo.barx, geo.bary);
    }
}

// active set (obstacle.cpp:133-145) and the condensed block of every cell; lc is [cell][N*N] column-major, N = 1 + 4F
template <int FD>
__global__ void obstacle_condense_kernel(Grid g, long long ncells, const double2* __restrict__ points, const double* __restrict__ lc,
                                         const double* __restrict__ fT, const double* __restrict__ alpha, const double* __restrict__ beta,
                                         const double* __restrict__ gamma, double cpar, int fun, unsigned char* __restrict__ in_A,
                                         double* __restrict__ S, double* __restrict__ gvec, int* __restrict__ n_active) {
    constexpr int F = FD + 1, NF = 4 * F, N = 1 + NF;
    int act = 0;
    for (long long c = blockIdx.x * (long long)blockDim.x + threadIdx.x; c < ncells; c += (long long)gridDim.x * blockDim.x) {
        const long long ci = c % g.Nx, cj = g.row_begin + c / g.Nx;
        const double* L = lc + c * (long long)(N * N);
        const bool A = (beta[c] + cpar * (alpha[c] - gamma[c])) < 0;
        in_A[c] = A ? 1 : 0; act += A ? 1 : 0;
        const double ktt = L[0], f = fT[c], gm = gamma[c];
        double gl[NF];
#pragma unroll
        for (int i = 0; i < NF; i++) {
            const double kft = L[(1 + i) * N];                 // K_TF(0, i) = K_FT(i, 0)
            gl[i] = A ? -kft * gm : -kft * f / ktt;
        }
        double* So = S + c * (long long)(NF * NF);
        for (int j = 0; j < NF; j++)
#pragma unroll
            for (int i = 0; i < NF; i++) {
                double v = L[(1 + i) + (1 + j) * N];
                if (!A) v -= L[(1 + i) * N] * L[(1 + j) * N] / ktt;
                So[i + j * NF] = v;
            }
        // Dirichlet faces: face projection of the datum moved to the rhs (hho.hpp:656-679)
        bool any = false;
        for (int li = 0; li < 4; li++) any |= cell_face_dirichlet(g, ci, cj, li);
        if (any) {
            CellGeom geo; load_cell_geom(g, points, ci, cj, geo);
            for (int li = 0; li < 4; li++) {
                if (!cell_face_dirichlet(g, ci, cj, li)) continue;
                double ud[F]; project_face<FD>(geo, li, fun, 0, ud);
                for (int i = 0; i < NF; i++) for (int a = 0; a < F; a++) gl[i] -= So[i + (li * F + a) * NF] * ud[a];
            }
        }
#pragma unroll
        for (int i = 0; i < NF; i++) gvec[c * NF + i] = gl[i];
    }
    if (act) atomicAdd(n_active, act);
}

// expand_solution (hho.hpp:698-744) for the cell part and the multiplier, plus ||alpha_prev - alpha||^2 over the
// cells; parts[blockIdx] receives the block's partial sum
template <int FD>
__global__ void obstacle_update_kernel(Grid g, long long ncells, const double2* __restrict__ points, const double* __restrict__ lc,
                                       const double* __restrict__ fT, const double* __restrict__ x, const unsigned char* __restrict__ in_A,
                                       const double* __restrict__ gamma, int fun, double* __restrict__ alpha, double* __restrict__ beta,
                                       double* __restrict__ parts) {
    constexpr int F = FD + 1, NF = 4 * F, N = 1 + NF;
    __shared__ double sh[32];
    double d2 = 0.0;
    for (long long c = blockIdx.x * (long long)blockDim.x + threadIdx.x; c < ncells; c += (long long)gridDim.x * blockDim.x) {
        const long long ci = c % g.Nx, cj = g.row_begin + c / g.Nx;
        const double* L = lc + c * (long long)(N * N);
        double ktf_u = 0.0;
        CellGeom geo; bool have_geo = false;
        for (int li = 0; li < 4; li++) {
            double uF[F];
            if (!cell_face_dirichlet(g, ci, cj, li)) {
                const long long cf = cell_face_compressed(g, ci, cj, li);
                for (int a = 0; a < F; a++) uF[a] = x[cf * F + a];
            } else {
                if (!have_geo) { load_cell_geom(g, points, ci, cj, geo); have_geo = true; }
                project_face<FD>(geo, li, fun, 0, uF);
            }
            for (int a = 0; a < F; a++) ktf_u += L[(1 + li * F + a) * N] * uF[a];
        }
        double an, bn;
        if (in_A[c]) { an = gamma[c]; bn = fT[c] - L[0] * an - ktf_u; }
        else { an = (fT[c] - ktf_u) / L[0]; bn = 0.0; }
        const double da = alpha[c] - an; d2 += da * da;
        alpha[c] = an; beta[c] = bn;
    }
    double t = block_reduce_sum(d2, sh); if (threadIdx.x == 0) parts[blockIdx.x] = t;
}

// ||x_prev - x||^2 of the face part of alpha, then x_prev = x
__global__ void obstacle_face_delta_kernel(long long n, const double* __restrict__ x, double* __restrict__ xprev, double* __restrict__ parts) {
    __shared__ double sh[32];
    double d2 = 0.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const double d = xprev[i] - x[i]; d2 += d * d; xprev[i] = x[i];
    }
    double t = block_reduce_sum(d2, sh); if (threadIdx.x == 0) parts[blockIdx.x] = t;
}

// final error (obstacle.cpp:201-218): sum over cells of (local - proj)^T lc (local - proj), proj = project_function with
// quadrature increase di = 1 (cell: mass / rhs of degree 2*(0+1); faces: degree 2*(fd+1))
template <int FD, int QK>
__global__ void obstacle_error_kernel(Grid g, long long ncells, const double2* __restrict__ points, const double* __restrict__ lc,
                                      const double* __restrict__ alpha, const double* __restrict__ x, int fun, double* __restrict__ parts) {
    constexpr int F = FD + 1, NF = 4 * F, N = 1 + NF;
    __shared__ double sh[32];
    double acc = 0.0;
    for (long long c = blockIdx.x * (long long)blockDim.x + threadIdx.x; c < ncells; c += (long long)gridDim.x * blockDim.x) {
        const long long ci = c % g.Nx, cj = g.row_begin + c / g.Nx;
        CellGeom geo; load_cell_geom(g, points, ci, cj, geo);
        double diff[N];
        {   // cell: constant basis, projection = int(sol) / int(1) with the degree-2 rule
            constexpr int nq = cell_qps(QK, 2);
            double m = 0.0, r = 0.0;
            for (int q = 0; q < nq; q++) { double px, py, w; cell_qp<QK, 2>(geo, q, px, py, w); m += w; r += w * fun_sol(fun, px, py); }
            diff[0] = alpha[c] - r / m;
        }
        for (int li = 0; li < 4; li++) {
            double uF[F], pr[F];
            if (!cell_face_dirichlet(g, ci, cj, li)) { const long long cf = cell_face_compressed(g, ci, cj, li); for (int a = 0; a < F; a++) uF[a] = x[cf * F + a]; }
            else project_face<FD>(geo, li, fun, 0, uF);
            project_face<FD>(geo, li, fun, 1, pr);
            for (int a = 0; a < F; a++) diff[1 + li * F + a] = uF[a] - pr[a];
        }
        const double* L = lc + c * (long long)(N * N);
        for (int j = 0; j < N; j++) { double v = 0.0; for (int i = 0; i < N; i++) v += L[i + j * N] * diff[i]; acc += v * diff[j]; }
    }
    double t = block_reduce_sum(acc, sh); if (threadIdx.x == 0) parts[blockIdx.x] = t;
}

}  // namespace pb
