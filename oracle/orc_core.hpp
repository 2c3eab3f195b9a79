/* ORACLE — TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement (plain C++17, no Eigen) of the HHO / cutHHO hot path of OmarDuran/ProtoN.
 * Nothing in the product (proton_b200/) may include, link or call this; only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs do.
 *
 * Parity status: the reference itself cannot be compiled here (Eigen3, SILO and Lua are un-vendored
 * system dependencies), so this restatement is pinned against the reference's own golden
 * tables (apps/obstacle/results/convergence.txt, apps/cuthho/cuthho.xlsx "F.D." block) to their 6 printed
 * digits by tests/test_oracle_golden.py.  Below 6 digits (per-cell matrices, numbering) the reference
 * ships no vectors: those are pinned by this restatement only.
 *
 * Each function cites the reference file:line (relative to /root/reference) it restates.
 * This file: dense helpers, points, mesh generator, geometry, bases, quadrature  (src/core/core_bits).
 */
#pragma once
#include <algorithm>
#include <array>
#include <cassert>
#include <cmath>
#include <cstddef>
#include <stdexcept>
#include <utility>
#include <vector>

namespace orc {

/* ---------------------------------------------------------------- dense helpers */
/* Column-major dynamic matrix, standing in for Eigen::Matrix<double,Dynamic,Dynamic>. */
struct Mat {
    int r = 0, c = 0;
    std::vector<double> a;
    Mat() {}
    Mat(int rr, int cc) : r(rr), c(cc), a((size_t)rr * cc, 0.0) {}
    double& operator()(int i, int j) { return a[(size_t)j * r + i]; }
    double operator()(int i, int j) const { return a[(size_t)j * r + i]; }
    static Mat identity(int n) { Mat m(n, n); for (int i = 0; i < n; i++) m(i, i) = 1.0; return m; }
    Mat block(int i0, int j0, int nr, int nc) const {
        Mat m(nr, nc);
        for (int j = 0; j < nc; j++) for (int i = 0; i < nr; i++) m(i, j) = (*this)(i0 + i, j0 + j);
        return m;
    }
    void set_block(int i0, int j0, const Mat& b) {
        for (int j = 0; j < b.c; j++) for (int i = 0; i < b.r; i++) (*this)(i0 + i, j0 + j) = b(i, j);
    }
    void add_block(int i0, int j0, const Mat& b, double s = 1.0) {
        for (int j = 0; j < b.c; j++) for (int i = 0; i < b.r; i++) (*this)(i0 + i, j0 + j) += s * b(i, j);
    }
    Mat transpose() const {
        Mat m(c, r);
        for (int j = 0; j < c; j++) for (int i = 0; i < r; i++) m(j, i) = (*this)(i, j);
        return m;
    }
};
typedef std::vector<double> Vec;

inline Mat operator*(const Mat& A, const Mat& B) {
    assert(A.c == B.r);
    Mat C(A.r, B.c);
    for (int j = 0; j < B.c; j++)
        for (int k = 0; k < A.c; k++) {
            double b = B(k, j);
            for (int i = 0; i < A.r; i++) C(i, j) += A(i, k) * b;
        }
    return C;
}
inline Mat operator+(const Mat& A, const Mat& B) {
    assert(A.r == B.r && A.c == B.c);
    Mat C(A.r, A.c);
    for (size_t i = 0; i < A.a.size(); i++) C.a[i] = A.a[i] + B.a[i];
    return C;
}
inline Mat operator-(const Mat& A, const Mat& B) {
    assert(A.r == B.r && A.c == B.c);
    Mat C(A.r, A.c);
    for (size_t i = 0; i < A.a.size(); i++) C.a[i] = A.a[i] - B.a[i];
    return C;
}
inline Mat operator*(double s, const Mat& A) {
    Mat C(A.r, A.c);
    for (size_t i = 0; i < A.a.size(); i++) C.a[i] = s * A.a[i];
    return C;
}
inline Vec matvec(const Mat& A, const Vec& x) {
    assert(A.c == (int)x.size());
    Vec y(A.r, 0.0);
    for (int j = 0; j < A.c; j++) for (int i = 0; i < A.r; i++) y[i] += A(i, j) * x[j];
    return y;
}
inline double dot(const Vec& a, const Vec& b) {
    double s = 0; for (size_t i = 0; i < a.size(); i++) s += a[i] * b[i]; return s;
}

/* Stand-in for Eigen's A.llt().solve(B): standard Cholesky A = L L^T, then two triangular solves.
 * Eigen does not check definiteness; neither do we (sqrt of a negative gives NaN). */
struct LLT {
    Mat L; int n;
    explicit LLT(const Mat& A) : L(A.r, A.r), n(A.r) {
        assert(A.r == A.c);
        for (int j = 0; j < n; j++) {
            double d = A(j, j);
            for (int k = 0; k < j; k++) d -= L(j, k) * L(j, k);
            d = std::sqrt(d);
            L(j, j) = d;
            for (int i = j + 1; i < n; i++) {
                double s = A(i, j);
                for (int k = 0; k < j; k++) s -= L(i, k) * L(j, k);
                L(i, j) = s / d;
            }
        }
    }
    Mat solve(const Mat& B) const {
        assert(B.r == n);
        Mat X = B;
        for (int c = 0; c < X.c; c++) {
            for (int i = 0; i < n; i++) {
                double s = X(i, c);
                for (int k = 0; k < i; k++) s -= L(i, k) * X(k, c);
                X(i, c) = s / L(i, i);
            }
            for (int i = n - 1; i >= 0; i--) {
                double s = X(i, c);
                for (int k = i + 1; k < n; k++) s -= L(k, i) * X(k, c);
                X(i, c) = s / L(i, i);
            }
        }
        return X;
    }
    Vec solve(const Vec& b) const {
        Mat B(n, 1); for (int i = 0; i < n; i++) B(i, 0) = b[i];
        Mat X = solve(B); Vec x(n); for (int i = 0; i < n; i++) x[i] = X(i, 0); return x;
    }
};

/* ---------------------------------------------------------------- points (point.hpp:35-112) */
struct Pt {
    double x = 0.0, y = 0.0;   /* point() zero-initialises, point.hpp:43-46 */
    Pt() {}
    Pt(double xx, double yy) : x(xx), y(yy) {}
};
inline Pt operator+(Pt a, Pt b) { return Pt(a.x + b.x, a.y + b.y); }
inline Pt operator-(Pt a, Pt b) { return Pt(a.x - b.x, a.y - b.y); }
inline Pt operator-(Pt a) { return Pt(-a.x, -a.y); }
inline Pt operator*(Pt a, double s) { return Pt(a.x * s, a.y * s); }
inline Pt operator*(double s, Pt a) { return Pt(a.x * s, a.y * s); }
inline Pt operator/(Pt a, double s) { return Pt(a.x / s, a.y / s); }
inline double det(Pt a, Pt b) { return a.x * b.y - a.y * b.x; }           /* point.hpp:106-111 */
inline double norm(Pt a) { return std::sqrt(a.x * a.x + a.y * a.y); }     /* to_vector().norm() */

/* ---------------------------------------------------------------- mesh (basic_mesh.hpp) */
enum Location { IN_NEGATIVE_SIDE = 0, IN_POSITIVE_SIDE = 1, ON_INTERFACE = 2, UNDEF = 3 }; /* cuthho_mesh.hpp:31-36 */

struct Node {                     /* basic_mesh.hpp:151-164 + node_cuthho_info cuthho_mesh.hpp:77-88 */
    size_t ptid = 0;
    int location = UNDEF; bool displaced = false; Pt displacement;
};
struct Face {                     /* basic_mesh.hpp:112-134 + face_cuthho_info cuthho_mesh.hpp:61-74 */
    std::array<size_t, 2> ptids{};
    bool is_boundary = false; bool dirichlet = false;
    int location = UNDEF; size_t node_inside = 0; Pt intersection_point;
    bool operator<(const Face& o) const { return ptids < o.ptids; }
    bool operator==(const Face& o) const { return ptids == o.ptids; }
};
struct Cell {                     /* basic_mesh.hpp:48-66 + cell_cuthho_info cuthho_mesh.hpp:45-58 */
    std::array<size_t, 4> ptids{};
    int location = UNDEF; Pt p0, p1; std::vector<Pt> interface; bool distorted = false;
    bool operator<(const Cell& o) const { return ptids < o.ptids; }
};

struct MeshParams { double min_x = 0, max_x = 1, min_y = 0, max_y = 1; size_t Nx = 4, Ny = 4; }; /* :178-197 */

/* Cartesian generator shared by quad_mesh (ET=4, basic_mesh.hpp:230-298) and the poly_mesh used by
 * cuthho_square (ET=0, :321-403; the random displacement there is commented out).  `poly` only selects
 * which integrate() overload a cell gets. */
struct Mesh {
    std::vector<Pt> points; std::vector<Node> nodes; std::vector<Face> faces; std::vector<Cell> cells;
    bool poly = false;
    Mesh() {}
    Mesh(const MeshParams& p, bool is_poly) : poly(is_poly) {
        double hx = (p.max_x - p.min_x) / p.Nx, hy = (p.max_y - p.min_y) / p.Ny;
        size_t pn = 0;
        for (size_t j = 0; j < p.Ny + 1; j++)
            for (size_t i = 0; i < p.Nx + 1; i++) {
                points.push_back(Pt(p.min_x + i * hx, p.min_y + j * hy));
                Node n; n.ptid = pn++; nodes.push_back(n);
            }
        for (size_t j = 0; j < p.Ny; j++)
            for (size_t i = 0; i < p.Nx; i++) {
                size_t p0 = j * (p.Nx + 1) + i, p1 = p0 + 1, p2 = p0 + p.Nx + 2, p3 = p0 + p.Nx + 1;
                Cell cl; cl.ptids = {p0, p1, p2, p3}; cells.push_back(cl);
                Face f0; f0.ptids = {p0, p1}; if (j == 0) f0.is_boundary = true; faces.push_back(f0);
                Face f1; f1.ptids = {p1, p2}; if (i == p.Nx - 1) f1.is_boundary = true; faces.push_back(f1);
                Face f2; f2.ptids = {p3, p2}; if (j == p.Ny - 1) f2.is_boundary = true; faces.push_back(f2);
                Face f3; f3.ptids = {p0, p3}; if (i == 0) f3.is_boundary = true; faces.push_back(f3);
            }
        std::sort(cells.begin(), cells.end());
        std::stable_sort(faces.begin(), faces.end());
        /* std::unique keeps the first of each run.  An interior face is pushed twice, never with the
         * boundary flag; a boundary face once.  So which duplicate survives is immaterial. */
        faces.erase(std::unique(faces.begin(), faces.end()), faces.end());
        for (auto& fc : faces) if (fc.is_boundary) fc.dirichlet = true;   /* :293-297 */
    }
};

/* basic_geom.hpp:30-61 — binary search in the sorted arrays */
inline size_t offset(const Mesh& m, const Cell& cl) {
    auto it = std::lower_bound(m.cells.begin(), m.cells.end(), cl);
    if (it == m.cells.end()) throw std::logic_error("Cell not found");
    return it - m.cells.begin();
}
inline size_t offset(const Mesh& m, const Face& fc) {
    auto it = std::lower_bound(m.faces.begin(), m.faces.end(), fc);
    if (it == m.faces.end()) throw std::logic_error("Face not found");
    return it - m.faces.begin();
}
/* basic_geom.hpp:122-178 */
inline std::array<Pt, 4> points(const Mesh& m, const Cell& cl) {
    return {m.points.at(cl.ptids[0]), m.points.at(cl.ptids[1]), m.points.at(cl.ptids[2]), m.points.at(cl.ptids[3])};
}
inline std::array<Pt, 2> points(const Mesh& m, const Face& fc) {
    return {m.points.at(fc.ptids[0]), m.points.at(fc.ptids[1])};
}
/* basic_geom.hpp:183-245 — local face i joins ptids[i], ptids[(i+1)%4], looked up with sorted ids */
inline std::array<size_t, 4> face_ids(const Mesh& m, const Cell& cl) {
    std::array<size_t, 4> ret{};
    for (size_t i = 0; i < 4; i++) {
        Face f; f.ptids = {cl.ptids[i], cl.ptids[(i + 1) % 4]};
        if (f.ptids[0] > f.ptids[1]) std::swap(f.ptids[0], f.ptids[1]);
        ret[i] = offset(m, f);
    }
    return ret;
}
/* basic_geom.hpp:247-286 */
template <typename It> inline Pt barycenter_range(It begin, It end) {
    Pt ret; double den = 0.0;
    size_t n = end - begin; Pt p0 = *begin;
    for (size_t i = 2; i < n; i++) {
        Pt pprev = *(begin + (i - 1)) - p0, pcur = *(begin + i) - p0;
        double d = det(pprev, pcur) / 2.0;
        ret = ret + (pprev + pcur) * d;
        den += d;
    }
    return p0 + ret / (den * 3);
}
inline Pt barycenter(const Mesh& m, const Cell& cl) { auto p = points(m, cl); return barycenter_range(p.begin(), p.end()); }
inline Pt barycenter(const Mesh& m, const Face& fc) { auto p = points(m, fc); return (p[0] + p[1]) / 2.0; }
/* basic_geom.hpp:288-315 */
inline double diameter(const Mesh& m, const Cell& cl) {
    double diam = 0.0;
    for (size_t i = 0; i < 4; i++) for (size_t j = i + 1; j < 4; j++)
        diam = std::max(diam, norm(m.points.at(cl.ptids[j]) - m.points.at(cl.ptids[i])));
    return diam;
}
inline double diameter(const Mesh& m, const Face& fc) { return norm(m.points.at(fc.ptids[1]) - m.points.at(fc.ptids[0])); }
/* basic_geom.hpp:317-345 */
inline double measure(const Mesh& m, const Cell& cl) {
    auto pts = points(m, cl); double acc = 0;
    for (size_t i = 1; i < 3; i++) {
        Pt u = pts[i] - pts[0], v = pts[i + 1] - pts[0];
        acc += std::abs(u.x * v.y - u.y * v.x) * 0.5;
    }
    return acc;
}
inline double measure(const Mesh& m, const Face& fc) { return diameter(m, fc); }
/* basic_geom.hpp:349-399 */
inline std::array<Pt, 4> normals(const Mesh& m, const Cell& cl) {
    auto pts = points(m, cl); std::array<Pt, 4> ret;
    for (size_t i = 0; i < 4; i++) {
        Pt v = pts[(i + 1) % 4] - pts[i];
        Pt n(v.y, -v.x);
        ret[i] = n / norm(n);
    }
    return ret;
}

/* ---------------------------------------------------------------- bases (bases.hpp) */
inline double iexp_pow(double x, size_t n) {   /* bases.hpp:27-50 */
    if (n == 0) return 1;
    double y = 1;
    while (n > 1) {
        if (n % 2 == 0) { x = x * x; n = n / 2; }
        else { y = x * y; x = x * x; n = (n - 1) / 2; }
    }
    return x * y;
}
inline size_t cell_basis_size(size_t d) { return (d + 2) * (d + 1) / 2; }
inline size_t face_basis_size(size_t d) { return d + 1; }

struct CellBasis {                 /* bases.hpp:70-195 */
    Pt bar; double h; size_t deg, size;
    CellBasis(const Mesh& m, const Cell& cl, size_t degree)
        : bar(barycenter(m, cl)), h(diameter(m, cl)), deg(degree), size(cell_basis_size(degree)) {}
    Vec eval_basis(Pt pt) const {
        Vec ret(size, 0.0);
        double bx = (pt.x - bar.x) / (0.5 * h), by = (pt.y - bar.y) / (0.5 * h);
        size_t pos = 0;
        for (size_t k = 0; k <= deg; k++)
            for (size_t i = 0; i <= k; i++)
                ret[pos++] = iexp_pow(bx, k - i) * iexp_pow(by, i);
        return ret;
    }
    Mat eval_gradients(Pt pt) const {
        Mat ret((int)size, 2);
        double bx = (pt.x - bar.x) / (0.5 * h), by = (pt.y - bar.y) / (0.5 * h), ih = 2.0 / h;
        size_t pos = 0;
        for (size_t k = 0; k <= deg; k++)
            for (size_t i = 0; i <= k; i++) {
                size_t pow_x = k - i, pow_y = i;
                double px = iexp_pow(bx, pow_x), py = iexp_pow(by, pow_y);
                double dx = (pow_x == 0) ? 0 : pow_x * ih * iexp_pow(bx, pow_x - 1);
                double dy = (pow_y == 0) ? 0 : pow_y * ih * iexp_pow(by, pow_y - 1);
                ret((int)pos, 0) = dx * py; ret((int)pos, 1) = px * dy; pos++;
            }
        return ret;
    }
};
struct FaceBasis {                 /* bases.hpp:241-291 */
    Pt bar, base; double h; size_t deg, size;
    FaceBasis(const Mesh& m, const Face& fc, size_t degree)
        : bar(barycenter(m, fc)), h(diameter(m, fc)), deg(degree), size(degree + 1) {
        base = bar - points(m, fc)[0];
    }
    Vec eval_basis(Pt pt) const {
        Vec ret(size, 0.0);
        Pt t = pt - bar;
        double d = base.x * t.x + base.y * t.y;
        double ep = 4.0 * d / (h * h);
        for (size_t i = 0; i <= deg; i++) ret[i] = iexp_pow(ep, i);
        return ret;
    }
};

} // namespace orc
