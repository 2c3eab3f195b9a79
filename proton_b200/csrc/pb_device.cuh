// Device-side building blocks shared by the local-operator kernels: compile-time sizes, monomial index maps,
// quadrature tables in constant memory, problem functors, geometry of a quad cell.
// Reference behaviour cited as file:line relative to the reference root.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include "../../include/proton_b200.h"

#define PB_HD __host__ __device__ __forceinline__
#define PB_D __device__ __forceinline__

namespace pb {

// ---------------------------------------------------------------- compile-time sizes
PB_HD constexpr int cell_basis_size(int d) { return (d + 1) * (d + 2) / 2; }   // bases.hpp:85-91
// monomial a -> (px, py): graded order, within degree d: px = d-i, py = i (bases.hpp:114-127)
PB_HD constexpr int mono_deg(int a) { int d = 0; while (cell_basis_size(d) <= a) d++; return d; }
PB_HD constexpr int mono_py(int a) { return a - cell_basis_size(mono_deg(a) - 1); }
PB_HD constexpr int mono_px(int a) { return mono_deg(a) - mono_py(a); }
// moment index of x^s y^t in the same graded order
PB_HD constexpr int mom_idx(int s, int t) { return cell_basis_size(s + t - 1) + t; }

template <int CD_, int FD_>
struct Sizes {
    static constexpr int CD = CD_, FD = FD_, RD = FD_ + 1;
    static constexpr int C = cell_basis_size(CD), F = FD + 1, R = cell_basis_size(RD), M = R - 1;
    static constexpr int NF = 4 * F, N = C + NF;
    static constexpr int QF = FD + 1;                 // Gauss nodes for degree 2*FD (quadratures.hpp:82-88)
    static constexpr int QFR = RD + 1;                // Gauss nodes for degree 2*RD (cut faces, cuthho_square.cpp:373)
};

// ---------------------------------------------------------------- quadrature tables (constant memory)
// Gauss-Legendre closed forms for 1..5 nodes in the reference's node order (quadratures.hpp:96-150); filled
// from the host at context creation with the same double formulas.
struct QuadTables {
    double gx[6][5];
    double gw[6][5];
    // Dunavant rules 1..8 as rows {l0,l1,l2,w} (quadratures_dunavant.hpp:27-118), index 8 = the empty sentinel
    double dun[9][16][4];
    int dun_n[9];
    // graded monomial index -> exponents (bases.hpp:114-127), degrees 0..8
    int mono_x[45], mono_y[45];
};
static __constant__ QuadTables c_qt;   // one translation unit (pb_api.cu) includes this header

// number of Gauss nodes for a polynomial degree: even degrees bumped to odd (quadratures.hpp:82-88)
PB_HD constexpr int gauss_nodes(int degree) { return ((degree % 2 == 0 ? degree + 1 : degree) + 1) / 2; }
// triangle_quadrature(deg): `rules[deg]` on a 0-based table => rule deg+1; deg 0 -> 1; deg 8 -> 0 points
// (quadratures.hpp:242-268)
PB_HD constexpr int dunavant_rule_index(int deg) { return deg == 0 ? 1 : deg; }
PB_HD constexpr int dunavant_npts_of_rule(int ri) {
    return ri == 0 ? 1 : ri == 1 ? 3 : ri == 2 ? 4 : ri == 3 ? 6 : ri == 4 ? 7 : ri == 5 ? 12 : ri == 6 ? 13 : ri == 7 ? 16 : 0;
}
PB_HD constexpr int dunavant_npts(int deg) { return dunavant_npts_of_rule(dunavant_rule_index(deg)); }
// points of integrate(msh, cl, degree) on an uncut quad cell
PB_HD constexpr int cell_qps(int quad_kind, int degree) {
    return quad_kind == PB_QUAD_TENSOR_GAUSS ? gauss_nodes(degree) * gauss_nodes(degree) : 4 * dunavant_npts(degree);
}

// ---------------------------------------------------------------- problem functors
#define PB_PI 3.14159265358979323846
PB_D double fun_rhs(int kind, double x, double y) {
    // convergence_test.cpp:100-102.  sinpi(x) instead of sin(M_PI * x): exact argument reduction, no slow path; differs from
    // the host value by the rounding of M_PI * x (<= 3e-16 absolute)
    if (kind == PB_FUN_SINSIN) return 2.0 * PB_PI * PB_PI * sinpi(x) * sinpi(y);
    if (kind == PB_FUN_OBSTACLE) {                                                            // obstacle.cpp:67-74
        const double r0 = 0.7; double r = sqrt(x * x + y * y);
        return (r > r0) ? -16 * r * r + 8 * r0 * r0 : -8.0 * (r0 * r0 * (r0 * r0 + 1)) + 8 * r0 * r0 * r * r;
    }
    double th1 = tanh(30 - 60 * (x - y)), ch1 = cosh(30 - 60 * (x - y));                       // convergence_test.cpp:116-123
    double th2 = tanh(60 * y), ch2 = cosh(60 * y);
    return 14400 * th1 / (ch1 * ch1) + 7200 * th2 / (ch2 * ch2);
}
PB_D double fun_sol(int kind, double x, double y) {
    if (kind == PB_FUN_SINSIN) return sinpi(x) * sinpi(y);                                     // :104-106
    if (kind == PB_FUN_OBSTACLE) {                                                            // obstacle.cpp:76-81
        const double r0 = 0.7; double r = sqrt(x * x + y * y), s = r * r - r0 * r0, t = fmax(s, 0.0);
        return t * t;
    }
    return tanh(60 * y) - tanh(60 * (x - y) - 30);                                             // :125-127
}
PB_D void fun_grad(int kind, double x, double y, double& gx, double& gy) {
    if (kind == PB_FUN_SINSIN) {                                                              // :108-113
        gx = PB_PI * cospi(x) * sinpi(y); gy = PB_PI * sinpi(x) * cospi(y);
    } else { gx = 0; gy = 0; }
}

// ---------------------------------------------------------------- closed-form numbering
struct Grid {
    long long Nx, Ny;                  // global cell counts
    long long row_begin, row_end;      // strip of cell rows owned by this context
    double min_x, min_y, hx, hy;
};
PB_HD long long grid_point(const Grid& g, long long i, long long j) { return j * (g.Nx + 1) + i; }
PB_HD long long face_h(const Grid& g, long long i, long long j) { return j < g.Ny ? j * (2 * g.Nx + 1) + 2 * i : g.Ny * (2 * g.Nx + 1) + i; }
PB_HD long long face_v(const Grid& g, long long i, long long j) { return i < g.Nx ? j * (2 * g.Nx + 1) + 2 * i + 1 : j * (2 * g.Nx + 1) + 2 * g.Nx; }
// faces(msh, cl) order: bottom, right, top, left (basic_geom.hpp:183-212)
PB_HD void cell_faces(const Grid& g, long long i, long long j, long long f[4]) {
    f[0] = face_h(g, i, j); f[1] = face_v(g, i + 1, j); f[2] = face_h(g, i, j + 1); f[3] = face_v(g, i, j);
}
// is_boundary && DIRICHLET (basic_mesh.hpp:266-297): local face li of cell (i,j)
PB_HD bool cell_face_dirichlet(const Grid& g, long long i, long long j, int li) {
    return li == 0 ? j == 0 : li == 1 ? i == g.Nx - 1 : li == 2 ? j == g.Ny - 1 : i == 0;
}
// compress_table (hho.hpp:313-323) of local face li of cell (i,j); caller guarantees it is not Dirichlet
PB_HD long long compress_h(const Grid& g, long long i, long long j) {   // 1 <= j <= Ny-1
    return (g.Nx - 1) + (j - 1) * (2 * g.Nx - 1) + (i == 0 ? 0 : 2 * i - 1);
}
PB_HD long long compress_v(const Grid& g, long long i, long long j) {   // 1 <= i <= Nx-1
    return j == 0 ? i - 1 : (g.Nx - 1) + (j - 1) * (2 * g.Nx - 1) + 2 * i;
}
PB_HD long long cell_face_compressed(const Grid& g, long long i, long long j, int li) {
    return li == 0 ? compress_h(g, i, j) : li == 1 ? compress_v(g, i + 1, j) : li == 2 ? compress_h(g, i, j + 1) : compress_v(g, i, j);
}
PB_HD long long num_free_faces(const Grid& g) { return (g.Nx - 1) + (g.Ny - 1) * (2 * g.Nx - 1); }

}  // namespace pb
