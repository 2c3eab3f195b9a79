/* ORACLE — TEST INFRASTRUCTURE ONLY (see orc_core.hpp header).
 * Quadrature: restates src/core/core_bits/quadratures.hpp and quadratures_dunavant.hpp. */
#pragma once
#include "orc_core.hpp"

namespace orc {

typedef std::pair<Pt, double> QP;          /* (point, weight) */
typedef std::pair<double, double> QP1;     /* 1-D (node, weight) */

/* quadratures.hpp:78-158 — closed forms for 1..5 nodes in the reference's node order.  The >=6-node
 * Golub–Welsch branch (:32-75) is only reached for k>=4 and is not restated (throws). */
inline std::vector<QP1> gauss_legendre(size_t degree) {
    size_t comp = degree; if (degree % 2 == 0) comp = degree + 1;
    size_t n = (comp + 1) / 2;
    std::vector<QP1> ret; double qp, qw, a1, a2;
    switch (n) {
    case 1: ret.push_back({0.0, 2.0}); return ret;
    case 2: qp = 1.0 / std::sqrt(3.0); ret.push_back({-qp, 1.0}); ret.push_back({qp, 1.0}); return ret;
    case 3:
        qp = std::sqrt(3.0 / 5.0); qw = 5.0 / 9.0;
        ret.push_back({-qp, qw}); ret.push_back({qp, qw}); ret.push_back({0.0, 8.0 / 9.0}); return ret;
    case 4:
        a1 = 3.0 / 7.0; a2 = 2.0 * std::sqrt(6.0 / 5.0) / 7.0;
        qp = std::sqrt(a1 - a2); qw = (18.0 + std::sqrt(30.0)) / 36.0;
        ret.push_back({-qp, qw}); ret.push_back({qp, qw});
        qp = std::sqrt(a1 + a2); qw = (18.0 - std::sqrt(30.0)) / 36.0;
        ret.push_back({-qp, qw}); ret.push_back({qp, qw}); return ret;
    case 5:
        ret.push_back({0.0, 128.0 / 225.0});
        a1 = 5.0; a2 = 2.0 * std::sqrt(10.0 / 7.0);
        qp = std::sqrt(a1 - a2) / 3.0; qw = (322 + 13.0 * std::sqrt(70.0)) / 900.0;
        ret.push_back({-qp, qw}); ret.push_back({qp, qw});
        qp = std::sqrt(a1 + a2) / 3.0; qw = (322 - 13.0 * std::sqrt(70.0)) / 900.0;
        ret.push_back({-qp, qw}); ret.push_back({qp, qw}); return ret;
    default: throw std::invalid_argument("oracle: Golub-Welsch (>5 Gauss nodes, k>=4) not restated");
    }
}

/* Dunavant rules 1..8 (quadratures_dunavant.hpp:27-118): published constants (Dunavant 1985) at the 15
 * digits the reference carries, kept as symmetry orbits and expanded in the reference's row order:
 *   S3   (a,a,a)                                  1 point
 *   S21  (a,b,b),(b,a,b),(b,b,a)                  3 points, a = the distinct coordinate
 *   S111 all 6 permutations of (a,b,c), lexicographic in position: abc acb bac bca cab cba */
struct DunOrbit { int kind; double a, b, c, w; };
inline const std::vector<std::vector<DunOrbit>>& dunavant_orbits() {
    static const std::vector<std::vector<DunOrbit>> R = {
        /* rule 1 */ {{1, 0.333333333333333, 0, 0, 1.000000000000000}},
        /* rule 2 */ {{3, 0.666666666666667, 0.166666666666667, 0, 0.333333333333333}},
        /* rule 3 */ {{1, 0.333333333333333, 0, 0, -0.562500000000000},
                      {3, 0.600000000000000, 0.200000000000000, 0, 0.520833333333333}},
        /* rule 4 */ {{3, 0.108103018168070, 0.445948490915965, 0, 0.223381589678011},
                      {3, 0.816847572980459, 0.091576213509771, 0, 0.109951743655322}},
        /* rule 5 */ {{1, 0.333333333333333, 0, 0, 0.225000000000000},
                      {3, 0.059715871789770, 0.470142064105115, 0, 0.132394152788506},
                      {3, 0.797426985353087, 0.101286507323456, 0, 0.125939180544827}},
        /* rule 6 */ {{3, 0.501426509658179, 0.249286745170910, 0, 0.116786275726379},
                      {3, 0.873821971016996, 0.063089014491502, 0, 0.050844906370207},
                      {6, 0.053145049844817, 0.310352451033784, 0.636502499121399, 0.082851075618374}},
        /* rule 7 */ {{1, 0.333333333333333, 0, 0, -0.149570044467682},
                      {3, 0.479308067841920, 0.260345966079040, 0, 0.175615257433208},
                      {3, 0.869739794195568, 0.065130102902216, 0, 0.053347235608838},
                      {6, 0.048690315425316, 0.312865496004874, 0.638444188569810, 0.077113760890257}},
        /* rule 8 */ {{1, 0.333333333333333, 0, 0, 0.144315607677787},
                      {3, 0.081414823414554, 0.459292588292723, 0, 0.095091634267285},
                      {3, 0.658861384496480, 0.170569307751760, 0, 0.103217370534718},
                      {3, 0.898905543365938, 0.050547228317031, 0, 0.032458497623198},
                      {6, 0.008394777409958, 0.263112829634638, 0.728492392955404, 0.027230314174435}},
    };
    return R;
}
/* rows {l0,l1,l2,w} of rule index ri (0-based: ri=0 is rule 1).  ri=8 is the reference's {0,NULL}
 * sentinel (quadratures_dunavant.hpp:120-130): zero points. */
inline std::vector<std::array<double, 4>> dunavant_rows(size_t ri) {
    std::vector<std::array<double, 4>> rows;
    if (ri >= 8) return rows;
    for (const auto& o : dunavant_orbits()[ri]) {
        if (o.kind == 1) rows.push_back({o.a, o.a, o.a, o.w});
        else if (o.kind == 3) {
            rows.push_back({o.a, o.b, o.b, o.w}); rows.push_back({o.b, o.a, o.b, o.w}); rows.push_back({o.b, o.b, o.a, o.w});
        } else {
            double a = o.a, b = o.b, c = o.c;
            rows.push_back({a, b, c, o.w}); rows.push_back({a, c, b, o.w}); rows.push_back({b, a, c, o.w});
            rows.push_back({b, c, a, o.w}); rows.push_back({c, a, b, o.w}); rows.push_back({c, b, a, o.w});
        }
    }
    return rows;
}

/* quadratures.hpp:238-271.  NOTE the off-by-one kept on purpose: rules[deg] on a 0-based
 * table, so degree d uses Dunavant rule d+1 and deg 8 hits the empty sentinel (0 points). */
inline std::vector<QP> triangle_quadrature(Pt p0, Pt p1, Pt p2, size_t deg) {
    if (deg == 0) deg = 1;
    if (deg > 8) throw std::invalid_argument("Quadrature order too high");
    Pt v0 = p1 - p0, v1 = p2 - p0;
    double area = std::abs((v0.x * v1.y - v0.y * v1.x) / 2.0);
    std::vector<QP> ret;
    for (const auto& row : dunavant_rows(deg)) {
        Pt qp = p0 * row[0] + p1 * row[1] + p2 * row[2];
        ret.push_back({qp, area * row[3]});
    }
    return ret;
}

/* quadratures.hpp:311-375 — tensor Gauss on the bilinear map, eta outer / xi inner */
inline std::vector<QP> integrate_quad_cell(const Mesh& m, const Cell& cl, size_t degree) {
    auto qps = gauss_legendre(degree);
    auto pts = points(m, cl);
    auto P = [&](double xi, double eta) {
        return 0.25 * pts[0].x * (1 - xi) * (1 - eta) + 0.25 * pts[1].x * (1 + xi) * (1 - eta) +
               0.25 * pts[2].x * (1 + xi) * (1 + eta) + 0.25 * pts[3].x * (1 - xi) * (1 + eta);
    };
    auto Q = [&](double xi, double eta) {
        return 0.25 * pts[0].y * (1 - xi) * (1 - eta) + 0.25 * pts[1].y * (1 + xi) * (1 - eta) +
               0.25 * pts[2].y * (1 + xi) * (1 + eta) + 0.25 * pts[3].y * (1 - xi) * (1 + eta);
    };
    auto J = [&](double xi, double eta) {
        double j11 = 0.25 * ((pts[1].x - pts[0].x) * (1 - eta) + (pts[2].x - pts[3].x) * (1 + eta));
        double j12 = 0.25 * ((pts[1].y - pts[0].y) * (1 - eta) + (pts[2].y - pts[3].y) * (1 + eta));
        double j21 = 0.25 * ((pts[3].x - pts[0].x) * (1 - xi) + (pts[2].x - pts[1].x) * (1 + xi));
        double j22 = 0.25 * ((pts[3].y - pts[0].y) * (1 - xi) + (pts[2].y - pts[1].y) * (1 + xi));
        return std::abs(j11 * j22 - j12 * j21);
    };
    std::vector<QP> ret;
    for (auto& qy : qps)
        for (auto& qx : qps) {
            double xi = qx.first, eta = qy.first;
            ret.push_back({Pt(P(xi, eta), Q(xi, eta)), qx.second * qy.second * J(xi, eta)});
        }
    return ret;
}
/* quadratures.hpp:377-402 — fan (p_i, p_{i+1}, barycenter) with Dunavant rules */
inline std::vector<QP> integrate_poly_cell(const Mesh& m, const Cell& cl, size_t degree) {
    Pt bar = barycenter(m, cl); auto pts = points(m, cl);
    std::vector<QP> ret;
    for (size_t i = 0; i < 4; i++) {
        auto q = triangle_quadrature(pts[i], pts[(i + 1) % 4], bar, degree);
        ret.insert(ret.end(), q.begin(), q.end());
    }
    return ret;
}
inline std::vector<QP> integrate(const Mesh& m, const Cell& cl, size_t degree) {
    return m.poly ? integrate_poly_cell(m, cl, degree) : integrate_quad_cell(m, cl, degree);
}
/* quadratures.hpp:404-432 — segment [p0,p1] in the face's sorted point order */
inline std::vector<QP> integrate_segment(Pt p0, Pt p1, size_t degree, double sign = 1.0) {
    auto qps = gauss_legendre(degree);
    double meas = norm(p1 - p0);
    std::vector<QP> ret;
    for (auto& q : qps) {
        double t = q.first;
        Pt p = 0.5 * (1 - t) * p0 + 0.5 * (1 + t) * p1;
        ret.push_back({p, sign * q.second * meas * 0.5});
    }
    return ret;
}
inline std::vector<QP> integrate(const Mesh& m, const Face& fc, size_t degree) {
    auto pts = points(m, fc);
    return integrate_segment(pts[0], pts[1], degree);
}

} // namespace orc
