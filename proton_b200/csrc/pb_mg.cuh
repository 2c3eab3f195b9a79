// Multigrid preconditioner for the condensed face system (apply_preconditioner = 3).  NOT reference
// behaviour: the reference's conjugated_gradient only knows the diagonal (solver_cg.hpp:78-81); the solution is the
// same to the tolerance, the iteration count no longer grows like N.
//
// Auxiliary-space two-grid on the face unknowns + geometric multigrid below it:
//   fine level   : the condensed face system A itself, smoothed by nu Chebyshev steps on D^{-1} A, D = the fbs x fbs
//                  diagonal blocks (pcg_bdiag_kernel), eigenvalue range [lmax * MG_CHEB_LO, lmax], lmax by power iteration
//   level 0      : continuous Q1 functions on the same Cartesian mesh (one unknown per interior vertex).  P maps a
//                  vertex function to the face monomials: a linear trace is exactly dof0 = mean, dof1 = half the
//                  difference (face_basis, bases.hpp:241-291: xi = -1 at the sorted first point), higher modes 0.
//                  On a cut face of the fictitious-domain problem the trace is the linear function through the value
//                  of the vertex inside the domain and 0 at the intersection point (the Nitsche condition), and vertices
//                  outside the domain carry no coarse unknown at all (mg_face_weights).  A_0 = P^T A P is formed cell by cell:
//                  E_T = P_T^T S_T P_T (4 x 4 per cell) from the condensed blocks K1/K2 leave in HBM.
//   levels 1...  : vertices with even indices (the last one kept when N is odd), bilinear interpolation, Galerkin
//                  operators again as 4 x 4 element matrices per coarse cell, 9-point stencils gathered from them,
//                  damped-Jacobi V(1,1) down to a single interior vertex.
// Multi-GPU: the face level and level 0 live on the strip of the rank (one halo row of vertices each side); levels
// >= 1 are replicated on every rank: the restriction 0 -> 1 (and the Galerkin product 0 -> 1) leaves partial sums of the
// owned rows that one all-reduce both completes and replicates.
// Everything index-related is a __host__ __device__ function so that tests/mg_host_check.cu can run the same code on
// host arrays against scipy's P^T A P (test infrastructure; the library has no host path).
#pragma once
#include "pb_device.cuh"
#include "pb_system.cuh"
#include "pb_pcg.cuh"

namespace pb {

constexpr double MG_CHEB_LO = 0.15;     // smoothing range [MG_CHEB_LO * lmax, lmax] of the Chebyshev smoother
constexpr double MG_OMEGA = 0.8;        // damping of the Jacobi sweeps on the vertex levels
constexpr int MG_MAX_LEVELS = 24;

struct MgGrid {             // one vertex level
    int Nx, Ny;             // cells per direction; vertices (Nx+1) x (Ny+1), boundary vertices carry 0
    int cx, cy;             // 1: the next coarser level halves this direction
    int row0, nrows;        // stored vertex rows [row0, row0 + nrows)
    int own0, own1;         // vertex rows this rank computes: [own0, own1)
    int crow0, ncrows;      // stored cell rows of the element matrices
};
PB_HD long long mg_node(const MgGrid& L, int i, int j) { return (long long)(j - L.row0) * (L.Nx + 1) + i; }
PB_HD bool mg_boundary(const MgGrid& L, int i, int j) { return i == 0 || j == 0 || i == L.Nx || j == L.Ny; }
PB_HD int mg_coarse_cells(int N) { return N > 2 ? (N + 1) / 2 : N; }
// fine index of coarse vertex I: 2I, the last fine vertex when N is odd
PB_HD int mg_cpos(int I, int N, int coarsen) { if (!coarsen) return I; const int p = 2 * I; return p < N ? p : N; }

// cut-cell data the prolongation needs (fictitious-domain problem only; node_loc == nullptr otherwise)
struct MgCut {
    const int* node_loc; const int* face_loc; const double2* face_ip; const double2* points;
};

// trace of a vertex function on a face with sorted end points a -> b: dof0 = wa[0] va + wb[0] vb, dof1 = wa[1] va + wb[1] vb.
// gf = global face id, pa / pb = point ids of the ends.
PB_HD void mg_face_weights(const MgCut& C, int F, long long gf, long long pa, long long pb, double wa[2], double wb[2]) {
    wa[0] = 0.5; wb[0] = 0.5; wa[1] = -0.5; wb[1] = 0.5;
    if (C.node_loc && C.face_loc[gf] == PB_ON_INTERFACE) {
        const double2 A = C.points[pa], B = C.points[pb], I = C.face_ip[gf];
        const double t = sqrt(((I.x - A.x) * (I.x - A.x) + (I.y - A.y) * (I.y - A.y)) / ((B.x - A.x) * (B.x - A.x) + (B.y - A.y) * (B.y - A.y)));
        const double xi = 2.0 * t - 1.0;
        wa[0] = wa[1] = wb[0] = wb[1] = 0.0;
        if (C.node_loc[pa] == PB_IN_NEGATIVE_SIDE) { wa[0] = F == 1 ? 0.5 : xi / (xi + 1.0); wa[1] = -1.0 / (xi + 1.0); }
        else { wb[0] = F == 1 ? 0.5 : -xi / (1.0 - xi); wb[1] = 1.0 / (1.0 - xi); }
    } else if (C.node_loc) {
        // vertices outside the domain carry no coarse unknown.  The fictitious-domain system decouples into the inside and the outside
        // of the interface (a cut cell integrates its inside part only); bilinear interpolation between vertex levels would couple the
        // two again, and the V-cycle would then excite the smooth modes of the outside problem, which CG has to remove one by one
        // (measured on the restatement: eigenvalues of M^-1 A down to 0.23 at 48^2, all of them outside modes; tools/scratch/
        // mg_slow_modes_cut.py).  With zero weights the outside rows and columns of every vertex-level operator vanish (those vertices
        // keep omega / diagonal = 0), so the outside faces see the smoother only; their right-hand side is zero (cuthho_square.cpp:626-665).
        if (C.node_loc[pa] != PB_IN_NEGATIVE_SIDE) { wa[0] = 0.0; wa[1] = 0.0; }
        if (C.node_loc[pb] != PB_IN_NEGATIVE_SIDE) { wb[0] = 0.0; wb[1] = 0.0; }
    }
    if (F == 1) { wa[1] = 0.0; wb[1] = 0.0; }
}

// E_T = P_T^T S_T P_T of cell (ci, cj).  S = the cell's condensed block, NF x NF column-major (symmetric).
// corners: 0 (i,j), 1 (i+1,j), 2 (i+1,j+1), 3 (i,j+1); faces(msh, cl) order bottom, right, top, left with sorted ends
// c0->c1, c1->c2, c3->c2, c0->c3.  Dirichlet faces are not unknowns: their rows of P_T are 0.
PB_HD void mg_element_from_S(const Grid& g, const MgCut& C, int F, long long ci, long long cj, const double* S, double E[16]) {
    const int NF = 4 * F, FA = F < 2 ? F : 2;
    const int ea[4] = {0, 1, 3, 0}, eb[4] = {1, 2, 2, 3};
    const long long p0 = grid_point(g, ci, cj);
    const long long pc[4] = {p0, p0 + 1, p0 + g.Nx + 2, p0 + g.Nx + 1};
    long long gf[4]; cell_faces(g, ci, cj, gf);
    double w[8][4];
    for (int r = 0; r < 8; r++) for (int c = 0; c < 4; c++) w[r][c] = 0.0;
    for (int li = 0; li < 4; li++) {
        if (cell_face_dirichlet(g, ci, cj, li)) continue;
        double wa[2], wb[2]; mg_face_weights(C, F, gf[li], pc[ea[li]], pc[eb[li]], wa, wb);
        for (int a = 0; a < FA; a++) { w[li * 2 + a][ea[li]] = wa[a]; w[li * 2 + a][eb[li]] = wb[a]; }
    }
    double T[8][4];
    for (int r = 0; r < 8; r++) for (int c = 0; c < 4; c++) T[r][c] = 0.0;
    for (int li = 0; li < 4; li++) for (int a = 0; a < FA; a++)
        for (int lj = 0; lj < 4; lj++) for (int b = 0; b < FA; b++) {
            const double s = S[(lj * F + b) * NF + li * F + a];
            for (int c = 0; c < 4; c++) T[li * 2 + a][c] += s * w[lj * 2 + b][c];
        }
    for (int c = 0; c < 4; c++) for (int d = 0; d < 4; d++) {
        double e = 0.0;
        for (int r = 0; r < 8; r++) e += w[r][c] * T[r][d];
        E[c * 4 + d] = e;
    }
}

// 9-point stencil of vertex (i, j) from the element matrices of the <= 4 cells around it; entry (dj+1)*3 + di+1
PB_HD void mg_node_stencil(const MgGrid& L, const double* E, int i, int j, double st[9]) {
    const int cdi[4] = {-1, 0, 0, -1}, cdj[4] = {-1, -1, 0, 0}, myc[4] = {2, 3, 0, 1};
    const int ox[4] = {0, 1, 1, 0}, oy[4] = {0, 0, 1, 1};
    for (int k = 0; k < 9; k++) st[k] = 0.0;
    for (int q = 0; q < 4; q++) {
        const int ci = i + cdi[q], cj = j + cdj[q];
        if (ci < 0 || ci >= L.Nx || cj < 0 || cj >= L.Ny) continue;
        const double* e = E + ((long long)(cj - L.crow0) * L.Nx + ci) * 16 + myc[q] * 4;
        for (int b = 0; b < 4; b++) st[(cj + oy[b] - j + 1) * 3 + (ci + ox[b] - i + 1)] += e[b];
    }
}

// Galerkin element matrix of coarse cell (I, J) of level Lc from the fine cells it covers (level Lf); fine cell rows
// outside [frow_lo, frow_hi) are left to another rank (their part arrives with the all-reduce)
PB_HD void mg_galerkin_cell(const MgGrid& Lf, const MgGrid& Lc, const double* Ef, int I, int J, int frow_lo, int frow_hi, double Ec[16]) {
    for (int k = 0; k < 16; k++) Ec[k] = 0.0;
    const int x0 = mg_cpos(I, Lf.Nx, Lf.cx), x1 = mg_cpos(I + 1, Lf.Nx, Lf.cx), y0 = mg_cpos(J, Lf.Ny, Lf.cy), y1 = mg_cpos(J + 1, Lf.Ny, Lf.cy);
    const double iwx = 1.0 / (x1 - x0), iwy = 1.0 / (y1 - y0);
    const int ox[4] = {0, 1, 1, 0}, oy[4] = {0, 0, 1, 1};
    for (int fj = y0; fj < y1; fj++) {
        if (fj < frow_lo || fj >= frow_hi) continue;
        for (int fi = x0; fi < x1; fi++) {
            double Q[4][4];          // fine corner a = sum_c Q[a][c] coarse corner c
            for (int a = 0; a < 4; a++) {
                const double tx = (fi + ox[a] - x0) * iwx, ty = (fj + oy[a] - y0) * iwy;
                Q[a][0] = (1 - tx) * (1 - ty); Q[a][1] = tx * (1 - ty); Q[a][2] = tx * ty; Q[a][3] = (1 - tx) * ty;
            }
            const double* e = Ef + ((long long)(fj - Lf.crow0) * Lf.Nx + fi) * 16;
            double T[4][4];
            for (int a = 0; a < 4; a++) for (int d = 0; d < 4; d++) {
                double t = 0.0;
                for (int b = 0; b < 4; b++) t += e[a * 4 + b] * Q[b][d];
                T[a][d] = t;
            }
            for (int c = 0; c < 4; c++) for (int d = 0; d < 4; d++) {
                double t = 0.0;
                for (int a = 0; a < 4; a++) t += Q[a][c] * T[a][d];
                Ec[c * 4 + d] += t;
            }
        }
    }
}

// full-weighting restriction (transpose of the bilinear interpolation): coarse vertex (I, J) of Lc gathers the fine
// residual; fine rows outside the owned range of Lf are left to their owner
PB_HD double mg_restrict_node(const MgGrid& Lf, const MgGrid& Lc, const double* rf, int I, int J) {
    const int fi = mg_cpos(I, Lf.Nx, Lf.cx), fj = mg_cpos(J, Lf.Ny, Lf.cy);
    double s = 0.0;
    for (int dj = (Lf.cy ? -1 : 0); dj <= (Lf.cy ? 1 : 0); dj++) {
        const int jj = fj + dj;
        if (jj <= 0 || jj >= Lf.Ny || jj < Lf.own0 || jj >= Lf.own1) continue;
        for (int di = (Lf.cx ? -1 : 0); di <= (Lf.cx ? 1 : 0); di++) {
            const int ii = fi + di;
            if (ii <= 0 || ii >= Lf.Nx) continue;
            s += (di ? 0.5 : 1.0) * (dj ? 0.5 : 1.0) * rf[mg_node(Lf, ii, jj)];
        }
    }
    return s;
}
// bilinear interpolation of the coarse correction at fine vertex (fi, fj)
PB_HD double mg_prolong_node(const MgGrid& Lf, const MgGrid& Lc, const double* xc, int fi, int fj) {
    int I0, I1, J0, J1; double wx1, wy1;
    if (!Lf.cx || (fi & 1) == 0 || fi == Lf.Nx) { I0 = I1 = Lf.cx ? (fi == Lf.Nx ? Lc.Nx : fi / 2) : fi; wx1 = 0.0; } else { I0 = (fi - 1) / 2; I1 = I0 + 1; wx1 = 0.5; }
    if (!Lf.cy || (fj & 1) == 0 || fj == Lf.Ny) { J0 = J1 = Lf.cy ? (fj == Lf.Ny ? Lc.Ny : fj / 2) : fj; wy1 = 0.0; } else { J0 = (fj - 1) / 2; J1 = J0 + 1; wy1 = 0.5; }
    const double a = xc[mg_node(Lc, I0, J0)], b = xc[mg_node(Lc, I1, J0)], c = xc[mg_node(Lc, I0, J1)], d = xc[mg_node(Lc, I1, J1)];
    return (1 - wy1) * ((1 - wx1) * a + wx1 * b) + wy1 * ((1 - wx1) * c + wx1 * d);
}

// face level <-> level 0.  Face vectors with halos start at compressed face col_lo.
// free face of the mesh by (horizontal?, i, j): horizontal h(i,j) joins vertices (i,j)-(i+1,j), vertical v(i,j) joins (i,j)-(i,j+1)
PB_HD void mg_face_ends(const Grid& g, bool horiz, long long i, long long j, long long& gf, long long& pa, long long& pb) {
    pa = grid_point(g, i, j);
    if (horiz) { gf = face_h(g, i, j); pb = pa + 1; } else { gf = face_v(g, i, j); pb = pa + g.Nx + 1; }
}
// P^T r at interior vertex (i, j): its four faces h(i-1,j) [end b], h(i,j) [end a], v(i,j-1) [end b], v(i,j) [end a]
PB_HD double mg_face_restrict_node(const Grid& g, const MgCut& C, int F, long long col_lo, const double* rfull, long long i, long long j) {
    double s = 0.0;
    for (int q = 0; q < 4; q++) {
        const bool horiz = q < 2, isb = (q & 1) == 0;
        const long long fi = horiz ? (isb ? i - 1 : i) : i, fj = horiz ? j : (isb ? j - 1 : j);
        long long gf, pa, pb; mg_face_ends(g, horiz, fi, fj, gf, pa, pb);
        double wa[2], wb[2]; mg_face_weights(C, F, gf, pa, pb, wa, wb);
        const long long cf = horiz ? compress_h(g, fi, fj) : compress_v(g, fi, fj);
        const double* r = rfull + (cf - col_lo) * F;
        s += (isb ? wb[0] : wa[0]) * r[0];
        if (F > 1) s += (isb ? wb[1] : wa[1]) * r[1];
    }
    return s;
}
// (P x)(face cf) for the first two modes; x0 = level-0 vector
PB_HD void mg_face_prolong(const Grid& g, const MgCut& C, int F, const MgGrid& L0, const double* x0, long long cf, double out[2]) {
    bool horiz; long long i, j; decompress_face(g, cf, horiz, i, j);
    long long gf, pa, pb; mg_face_ends(g, horiz, i, j, gf, pa, pb);
    double wa[2], wb[2]; mg_face_weights(C, F, gf, pa, pb, wa, wb);
    const double va = x0[mg_node(L0, (int)i, (int)j)], vb = x0[mg_node(L0, (int)(horiz ? i + 1 : i), (int)(horiz ? j : j + 1))];
    out[0] = wa[0] * va + wb[0] * vb; out[1] = wa[1] * va + wb[1] * vb;
}

#ifdef __CUDACC__
// ------------------------------------------------------------------------------------------------ setup kernels
template <int F>
__global__ void mg_elements_kernel(Grid g, MgCut C, long long cell_row0, long long ncells, const double* __restrict__ S, double* __restrict__ E) {
    constexpr int NF = 4 * F;
    for (long long c = blockIdx.x * (long long)blockDim.x + threadIdx.x; c < ncells; c += (long long)gridDim.x * blockDim.x) {
        const long long ci = c % g.Nx, cj = cell_row0 + c / g.Nx;
        double e[16]; mg_element_from_S(g, C, F, ci, cj, S + c * (NF * NF), e);
#pragma unroll
        for (int k = 0; k < 16; k++) E[c * 16 + k] = e[k];
    }
}
// stencils (SoA: st[k * stride + node]) and omega / diagonal of the owned vertex rows
__global__ void mg_stencil_kernel(MgGrid L, const double* __restrict__ E, double* __restrict__ st, double* __restrict__ idiag, double omega) {
    const long long stride = (long long)L.nrows * (L.Nx + 1), n = (long long)(L.own1 - L.own0) * (L.Nx + 1);
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) {
        const int j = L.own0 + (int)(t / (L.Nx + 1)), i = (int)(t % (L.Nx + 1));
        const long long nd = mg_node(L, i, j);
        double s[9];
        if (mg_boundary(L, i, j)) { for (int k = 0; k < 9; k++) s[k] = 0.0; s[4] = 1.0; } else mg_node_stencil(L, E, i, j, s);
#pragma unroll
        for (int k = 0; k < 9; k++) st[k * stride + nd] = s[k];
        idiag[nd] = s[4] > 0.0 ? omega / s[4] : 0.0;       // a vertex no cell matrix touches (none on these meshes) stays at 0
    }
}
__global__ void mg_galerkin_kernel(MgGrid Lf, MgGrid Lc, const double* __restrict__ Ef, double* __restrict__ Ec, int frow_lo, int frow_hi) {
    const long long n = (long long)Lc.Nx * Lc.Ny;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) {
        const int J = (int)(t / Lc.Nx), I = (int)(t % Lc.Nx);
        double e[16]; mg_galerkin_cell(Lf, Lc, Ef, I, J, frow_lo, frow_hi, e);
#pragma unroll
        for (int k = 0; k < 16; k++) Ec[t * 16 + k] = e[k];
    }
}

// ------------------------------------------------------------------------------------------------ vertex-level cycle
// x = omega D^{-1} b (first sweep from x = 0); boundary vertices 0
__global__ void mg_jacobi_first_kernel(MgGrid L, const double* __restrict__ idiag, const double* __restrict__ b, double* __restrict__ x) {
    const long long n = (long long)(L.own1 - L.own0) * (L.Nx + 1), base = mg_node(L, 0, L.own0);
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) {
        const int j = L.own0 + (int)(t / (L.Nx + 1)), i = (int)(t % (L.Nx + 1));
        x[base + t] = mg_boundary(L, i, j) ? 0.0 : idiag[base + t] * b[base + t];
    }
}
// MODE 0: out = b - A x (residual);  MODE 1: out = x + omega D^{-1} (b - A x) (one damped-Jacobi sweep)
template <int MODE>
__global__ void mg_stencil_apply_kernel(MgGrid L, const double* __restrict__ st, const double* __restrict__ idiag, const double* __restrict__ b,
                                        const double* __restrict__ x, double* __restrict__ out) {
    const int W = L.Nx + 1;
    const long long stride = (long long)L.nrows * W, n = (long long)(L.own1 - L.own0) * W, base = mg_node(L, 0, L.own0);
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) {
        const int j = L.own0 + (int)(t / W), i = (int)(t % W);
        const long long nd = base + t;
        double v = 0.0;
        if (!mg_boundary(L, i, j)) {
            double ax = 0.0;
#pragma unroll
            for (int dj = -1; dj <= 1; dj++)
#pragma unroll
                for (int di = -1; di <= 1; di++) ax += st[((dj + 1) * 3 + di + 1) * stride + nd] * x[nd + dj * W + di];
            const double r = b[nd] - ax;
            v = MODE == 0 ? r : x[nd] + idiag[nd] * r;
        }
        out[nd] = v;
    }
}
__global__ void mg_restrict_kernel(MgGrid Lf, MgGrid Lc, const double* __restrict__ rf, double* __restrict__ bc) {
    const long long n = (long long)(Lc.Nx + 1) * (Lc.Ny + 1);
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) {
        const int J = (int)(t / (Lc.Nx + 1)), I = (int)(t % (Lc.Nx + 1));
        bc[mg_node(Lc, I, J)] = mg_boundary(Lc, I, J) ? 0.0 : mg_restrict_node(Lf, Lc, rf, I, J);
    }
}
__global__ void mg_prolong_add_kernel(MgGrid Lf, MgGrid Lc, const double* __restrict__ xc, double* __restrict__ xf) {
    const long long n = (long long)(Lf.own1 - Lf.own0) * (Lf.Nx + 1), base = mg_node(Lf, 0, Lf.own0);
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) {
        const int j = Lf.own0 + (int)(t / (Lf.Nx + 1)), i = (int)(t % (Lf.Nx + 1));
        if (!mg_boundary(Lf, i, j)) xf[base + t] += mg_prolong_node(Lf, Lc, xc, i, j);
    }
}

// ------------------------------------------------------------------------------------------------ face level
// b0 = P^T r at the owned vertices of level 0 (r with halos, starting at compressed face col_lo)
__global__ void mg_face_restrict_kernel(Grid g, MgCut C, int F, MgGrid L0, long long col_lo, const double* __restrict__ rfull, double* __restrict__ b0) {
    const long long n = (long long)(L0.own1 - L0.own0) * (L0.Nx + 1), base = mg_node(L0, 0, L0.own0);
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) {
        const int j = L0.own0 + (int)(t / (L0.Nx + 1)), i = (int)(t % (L0.Nx + 1));
        b0[base + t] = mg_boundary(L0, i, j) ? 0.0 : mg_face_restrict_node(g, C, F, col_lo, rfull, i, j);
    }
}
// z += P x0 on the owned faces (z = owned part, block row br <-> compressed face cf_begin + br)
__global__ void mg_face_prolong_add_kernel(Grid g, MgCut C, int F, MgGrid L0, long long cf_begin, long long nbr, const double* __restrict__ x0, double* __restrict__ z) {
    for (long long br = blockIdx.x * (long long)blockDim.x + threadIdx.x; br < nbr; br += (long long)gridDim.x * blockDim.x) {
        double o[2]; mg_face_prolong(g, C, F, L0, x0, cf_begin + br, o);
        z[br * F] += o[0];
        if (F > 1) z[br * F + 1] += o[1];
    }
}

// Chebyshev smoother on D^{-1} A, one thread per face (F x F block of iAb, row-major).  With y = A d of the previous step:
//   FIRST = 1:  res = src - (HAVE_Y ? y : 0);   d = c2 * D^{-1} res;            z (+)= d      (ACC: z += d, else z = d)
//   FIRST = 0:  res = res - y;                  d = c1 * d + c2 * D^{-1} res;   z += d
// src = the right-hand side of the smoothing problem (FIRST only), res = running residual (written), d without halo
// offset (the caller passes the owned part).  LAST = 1: only the residual update (res -= y), no new direction.
template <int F>
__global__ void mg_cheb_kernel(long long nbr, int first, int have_y, int acc, int last, double c1, double c2, const double* __restrict__ iAb,
                               const double* __restrict__ src, const double* __restrict__ y, double* __restrict__ res, double* __restrict__ d,
                               double* __restrict__ z, const CgState* st) {
    if (st && st->done) return;
    for (long long br = blockIdx.x * (long long)blockDim.x + threadIdx.x; br < nbr; br += (long long)gridDim.x * blockDim.x) {
        double r[F];
#pragma unroll
        for (int a = 0; a < F; a++) {
            const long long e = br * F + a;
            r[a] = first ? src[e] - (have_y ? y[e] : 0.0) : res[e] - y[e];
            res[e] = r[a];
        }
        if (last) continue;
#pragma unroll
        for (int a = 0; a < F; a++) {
            const long long e = br * F + a;
            double t = 0.0;
#pragma unroll
            for (int c = 0; c < F; c++) t += iAb[e * F + c] * r[c];
            const double dn = first ? c2 * t : c1 * d[e] + c2 * t;
            d[e] = dn;
            z[e] = (first && !acc) ? dn : z[e] + dn;
        }
    }
}

// power iteration for lmax(D^{-1} A): w = D^{-1} y, partial sums of w.w
template <int F>
__global__ void mg_pow_apply_kernel(long long nbr, const double* __restrict__ iAb, const double* __restrict__ y, double* __restrict__ w, double* __restrict__ parts) {
    __shared__ double sh[32];
    double s = 0.0;
    for (long long br = blockIdx.x * (long long)blockDim.x + threadIdx.x; br < nbr; br += (long long)gridDim.x * blockDim.x) {
        double r[F];
#pragma unroll
        for (int a = 0; a < F; a++) r[a] = y[br * F + a];
#pragma unroll
        for (int a = 0; a < F; a++) {
            double t = 0.0;
#pragma unroll
            for (int c = 0; c < F; c++) t += iAb[(br * F + a) * F + c] * r[c];
            w[br * F + a] = t; s += t * t;
        }
    }
    const double t = block_reduce_sum(s, sh); if (threadIdx.x == 0) parts[blockIdx.x] = t;
}
// v = w / sqrt(*sum); lam = sqrt(*sum) (the iterate had norm 1)
__global__ void mg_pow_scale_kernel(long long n, const double* __restrict__ sum, const double* __restrict__ w, double* __restrict__ v, double* lam) {
    const double nrm = sqrt(*sum), inv = nrm > 0.0 ? 1.0 / nrm : 0.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) v[i] = w[i] * inv;
    if (blockIdx.x == 0 && threadIdx.x == 0) *lam = nrm;
}
// deterministic start vector of the power iteration: a hash of the GLOBAL unknown id (same vector for any partition)
__global__ void mg_pow_init_kernel(long long n, long long row0, double* __restrict__ v) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        unsigned long long h = (unsigned long long)(row0 + i) * 0x9E3779B97F4A7C15ull; h ^= h >> 29; h *= 0xBF58476D1CE4E5B9ull; h ^= h >> 32;
        v[i] = (double)(h & 0xFFFFF) / 1048576.0 + 0.25;
    }
}

// single-precision copy of the CSR values for the SpMVs of the V-cycle
__global__ void mg_val_to_float_kernel(long long nnz, const double* __restrict__ val, float* __restrict__ valf) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nnz; i += (long long)gridDim.x * blockDim.x) valf[i] = (float)val[i];
}

// ------------------------------------------------------------------------------------------------ PCG pieces
// x += alpha d, r -= alpha y with alpha = rho / (d.y); partial sums of r.r           (solver_cg.hpp:106-110)
__global__ void mg_xr_kernel(long long n, int nparts, const double* __restrict__ parts_dy, const double* __restrict__ d, const double* __restrict__ y,
                             double* __restrict__ x, double* __restrict__ r, double* __restrict__ parts_rr, const CgState* st) {
    __shared__ double sh[32];
    if (st->done) return;
    const double dy = reduce_partials(parts_dy, nparts, sh);
    const double alpha = st->rho / dy;
    double rr = 0.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        x[i] += alpha * d[i];
        const double ri = r[i] - alpha * y[i];
        r[i] = ri; rr += ri * ri;
    }
    const double t = block_reduce_sum(rr, sh); if (threadIdx.x == 0) parts_rr[blockIdx.x] = t;
}
// partial sums of a.b (and optionally a.a)
__global__ void mg_dot_kernel(long long n, const double* __restrict__ a, const double* __restrict__ b, double* __restrict__ parts_ab, double* __restrict__ parts_aa,
                              const CgState* st) {
    __shared__ double sh[32];
    if (st && st->done) return;
    double ab = 0.0, aa = 0.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) { ab += a[i] * b[i]; aa += a[i] * a[i]; }
    double t = block_reduce_sum(ab, sh); if (threadIdx.x == 0) parts_ab[blockIdx.x] = t;
    if (parts_aa) { t = block_reduce_sum(aa, sh); if (threadIdx.x == 0) parts_aa[blockIdx.x] = t; }
}
#endif  // __CUDACC__

}  // namespace pb

#ifdef __CUDACC__
namespace pb {
// exit tests of solver_cg.hpp:112-125 right after the residual update, so that a converged solve does not pay another
// V-cycle: in the exit case it records exactly what pcg_commit_kernel would (which then finds `done` set)
__global__ void mg_check_kernel(int nparts, const double* parts_rr, double tol, double divtol, long long max_iter, CgState* st, double* hist, long long hist_cap) {
    __shared__ double sh[32];
    if (st->done) return;
    const long long iter = st->iters;
    const double rr = reduce_partials(parts_rr, nparts, sh);
    if (threadIdx.x != 0) return;
    const double rel = sqrt(rr) / st->nr0;
    int reason = -1;
    if (rel < tol) reason = PB_CG_CONVERGED;
    else if (iter > max_iter) reason = PB_CG_MAX_ITER_REACHED;
    else if (rel > divtol) reason = PB_CG_DIVERGED;
    if (reason >= 0) {
        st->relres = rel;
        if (hist && iter + 1 < hist_cap) hist[iter + 1] = rel;
        st->done = 1; st->reason = reason; st->iters = iter;
    }
}

struct MgLevelData { MgGrid L; double omega; double *E, *st, *idiag, *x, *b, *r, *t; };
struct MgState {
    bool ready = false; long long gen = -1;      // assembly generation the operators were built from
    int nlev = 0, F = 0; long long key[4] = {0, 0, 0, 0};      // structure the arrays were allocated for
    MgLevelData lev[MG_MAX_LEVELS]{};
    double lmax = 0.0; double *resfull = nullptr, *zfull = nullptr, *wfull = nullptr, *lam = nullptr;
    long long nfull = 0;
    float* valf = nullptr; long long valf_n = 0;      // fp32 copy of the CSR values (V-cycle SpMVs), or nullptr
    // line-block Jacobi smoother (pb_mg_lines.cuh): factors of the two line sets (float unless PB_MG_FP64), line-solve output
    bool lines = false, lines_f64 = false; void *fac_h = nullptr, *fac_v = nullptr; double* tl = nullptr; long long fac_key[3] = {0, 0, 0};
};
}  // namespace pb
#endif
