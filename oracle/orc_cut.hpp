/* ORACLE — TEST INFRASTRUCTURE ONLY (see orc_core.hpp header).
 * cutHHO: restates src/methods/cuthho_bits/cuthho_geom.hpp:27-895, cuthho_utils.hpp and the
 * driver-resident cut operators of apps/cuthho/cuthho_square.cpp:56-124, 293-388, 566-666. */
#pragma once
#include "orc_hho.hpp"

namespace orc {

/* cuthho_square.cpp:56-124 */
struct LevelSet {
    int kind = 0;                 /* 0 circle, 1 horizontal line */
    double radius = 0.35, alpha = 0.5, beta = 0.5, cut_y = 0.5;
    double operator()(Pt p) const {
        if (kind == 0) return (p.x - alpha) * (p.x - alpha) + (p.y - beta) * (p.y - beta) - radius * radius;
        return p.y - cut_y;
    }
    Pt gradient(Pt p) const { return kind == 0 ? Pt(2 * p.x - 2 * alpha, 2 * p.y - 2 * beta) : Pt(0, 1); }
    Pt normal(Pt p) const { Pt g = gradient(p); return g / norm(g); }
};

/* cuthho_geom.hpp:68-116 */
inline Pt find_zero_crossing(Pt p0, Pt p1, const LevelSet& ls, double threshold) {
    Pt pa = p0, pb = p1, pm = (pa + pb) / 2.0, pm_prev = pm;
    double xd, yd; size_t max_iter = 30;
    do {
        double lb = ls(pb), lm = ls(pm);
        if ((lb >= 0 && lm >= 0) || (lb < 0 && lm < 0)) { pm_prev = pm; pb = pm; pm = (pa + pb) / 2.0; }
        else { pm_prev = pm; pa = pm; pm = (pa + pb) / 2.0; }
        xd = (pm_prev.x - pm.x) * (pm_prev.x - pm.x); yd = (pm_prev.y - pm.y) * (pm_prev.y - pm.y);
    } while ((std::sqrt(xd + yd) > threshold) && max_iter--);
    return pm;
}
/* :118-130 */
inline void detect_node_position(Mesh& m, const LevelSet& ls) {
    for (auto& n : m.nodes) n.location = (ls(m.points.at(n.ptid)) < 0) ? IN_NEGATIVE_SIDE : IN_POSITIVE_SIDE;
}
/* :132-161 */
inline void detect_cut_faces(Mesh& m, const LevelSet& ls) {
    for (auto& fc : m.faces) {
        auto pts = points(m, fc);
        double l0 = ls(pts[0]), l1 = ls(pts[1]);
        if (l0 >= 0 && l1 >= 0) { fc.location = IN_POSITIVE_SIDE; continue; }
        if (l0 < 0 && l1 < 0) { fc.location = IN_NEGATIVE_SIDE; continue; }
        double threshold = diameter(m, fc) / 1e4;
        Pt pm = find_zero_crossing(pts[0], pts[1], ls, threshold);
        fc.node_inside = (l0 < 0) ? 0 : 1;
        fc.location = ON_INTERFACE;
        fc.intersection_point = pm;
    }
}
/* :275-340 */
inline void detect_cut_cells(Mesh& m, const LevelSet& ls) {
    for (auto& cl : m.cells) {
        auto fids = face_ids(m, cl);
        std::array<Pt, 2> cut; size_t k = 0;
        for (size_t i = 0; i < 4; i++)
            if (m.faces[fids[i]].location == ON_INTERFACE) {
                if (k >= 2) throw std::out_of_range("more than two cut faces");   /* cut_faces.at(k++) */
                cut[k++] = m.faces[fids[i]].intersection_point;
            }
        if (k == 0) {
            auto pts = points(m, cl);
            bool allpos = true; for (auto& p : pts) if (!(ls(p) > 0)) allpos = false;
            cl.location = allpos ? IN_POSITIVE_SIDE : IN_NEGATIVE_SIDE;
        }
        if (k == 2) {
            cl.location = ON_INTERFACE;
            Pt p0 = cut[0], p1 = cut[1], pt = p1 - p0, pn = p0 + Pt(-pt.y, pt.x);
            if (ls(pn) >= 0) { cl.p0 = p1; cl.p1 = p0; } else { cl.p0 = p0; cl.p1 = p1; }
            cl.interface.push_back(cl.p0); cl.interface.push_back(cl.p1);
        }
        if (k != 0 && k != 2) throw std::logic_error("invalid number of cuts in cell");
    }
}
/* :466-543 (the non-USE_OLD_DISPLACEMENT branch).  Node `location`s are NOT recomputed afterwards. */
inline void move_nodes(Mesh& m, const LevelSet&) {
    double closeness_thresh = 0.4;
    for (auto& fc : m.faces) {
        if (fc.location != ON_INTERFACE) continue;
        auto pts = points(m, fc);
        Pt bar = (pts[1] + pts[0]) / 2.0;
        double lf = norm(pts[1] - pts[0]), dp = norm(fc.intersection_point - pts[0]);
        double closeness = dp / lf;
        size_t ofs;
        if (closeness < closeness_thresh) ofs = fc.ptids[0];
        else if (closeness > 1.0 - closeness_thresh) ofs = fc.ptids[1];
        else continue;
        Pt delta = (bar - fc.intersection_point) / 2;
        m.nodes.at(ofs).displacement = m.nodes.at(ofs).displacement - delta;
        m.nodes.at(ofs).displaced = true;
    }
    for (auto& nd : m.nodes) if (nd.displaced) m.points.at(nd.ptid) = m.points.at(nd.ptid) + nd.displacement;
    for (auto& cl : m.cells) for (auto id : cl.ptids) if (m.nodes.at(id).displaced) cl.distorted = true;
    for (auto& cl : m.cells) {
        if (!cl.distorted) continue;
        auto pts = points(m, cl);
        for (size_t i = 0; i < 4; i++) {
            Pt v1 = pts[(i + 1) % 4] - pts[i], v2 = pts[(i + 2) % 4] - pts[(i + 1) % 4];
            if (v1.x * v2.y - v2.x * v1.y < 0) throw std::logic_error("concave poly");
        }
    }
}
/* :609-673 */
inline void refine_interface_rec(Mesh& m, Cell& cl, const LevelSet& ls, size_t mn, size_t mx) {
    if ((mx - mn) < 2) return;
    size_t mid = (mx + mn) / 2;
    Pt p0 = cl.interface.at(mn), p1 = cl.interface.at(mx), pm = (p0 + p1) / 2.0, pt = p1 - p0, pn(-pt.y, pt.x);
    Pt ps1 = pm + pn, ps2 = pm - pn;
    double lm = ls(pm), ls1 = ls(ps1), ls2 = ls(ps2);
    Pt ip;
    if (!((lm >= 0 && ls1 >= 0) || (lm < 0 && ls1 < 0))) ip = find_zero_crossing(pm, ps1, ls, diameter(m, cl) / 10000.0);
    else if (!((lm >= 0 && ls2 >= 0) || (lm < 0 && ls2 < 0))) ip = find_zero_crossing(pm, ps2, ls, diameter(m, cl) / 10000.0);
    else throw std::logic_error("interface not found in search range");
    cl.interface.at(mid) = ip;
    refine_interface_rec(m, cl, ls, mn, mid);
    refine_interface_rec(m, cl, ls, mid, mx);
}
inline void refine_interface(Mesh& m, const LevelSet& ls, size_t levels) {
    if (levels == 0) return;
    size_t ip = (size_t)1 << levels;
    for (auto& cl : m.cells) {
        if (cl.location != ON_INTERFACE) continue;
        cl.interface.resize(ip + 1);
        cl.interface.at(0) = cl.p0; cl.interface.at(ip) = cl.p1;
        refine_interface_rec(m, cl, ls, 0, ip);
    }
}
/* cuthho_square.cpp:2036-2052 (node displacement branch) */
inline void cut_preprocess(Mesh& m, const LevelSet& ls, size_t refsteps) {
    detect_node_position(m, ls);
    detect_cut_faces(m, ls);
    move_nodes(m, ls);
    detect_cut_faces(m, ls);
    detect_cut_cells(m, ls);
    refine_interface(m, ls, refsteps);
}

/* cuthho_geom.hpp:547-569 */
inline std::array<Pt, 2> points_where(const Mesh& m, const Face& fc, int where) {
    if (fc.location != where && fc.location != ON_INTERFACE) throw std::invalid_argument("This face has no points where requested");
    auto pts = points(m, fc);
    if (fc.location != ON_INTERFACE) return pts;
    int l0 = m.nodes.at(fc.ptids[0]).location, l1 = m.nodes.at(fc.ptids[1]).location;
    if (l0 == where && l1 != where) pts[1] = fc.intersection_point;
    else if (l0 != where && l1 == where) pts[0] = fc.intersection_point;
    else throw std::logic_error("Invalid point configuration");
    return pts;
}
/* :675-728 */
inline std::vector<Pt> collect_triangulation_points(const Mesh& m, const Cell& cl, int where) {
    std::vector<Pt> ret;
    auto loc = [&](size_t i) { return m.nodes.at(cl.ptids[i]).location; };
    auto pt = [&](size_t i) { return m.points.at(cl.ptids[i]); };
    auto insert_interface = [&]() {
        if (where == IN_NEGATIVE_SIDE) ret.insert(ret.end(), cl.interface.begin(), cl.interface.end());
        else if (where == IN_POSITIVE_SIDE) ret.insert(ret.end(), cl.interface.rbegin(), cl.interface.rend());
        else throw std::logic_error("bad where");
    };
    bool case1 = loc(0) == where && loc(3) != where;
    bool case2 = loc(0) != where && loc(3) == where;
    bool case3 = loc(0) != where && loc(3) != where;
    if (case1 || case2 || case3) {
        for (size_t i = 0; i < 4; i++) if (loc(i) == where) ret.push_back(pt(i));
        insert_interface();
    } else {
        size_t i = 0;
        while (i < 4 && loc(i) == where) ret.push_back(pt(i++));
        insert_interface();
        while (i < 4 && loc(i) != where) i++;
        while (i < 4 && loc(i) == where) ret.push_back(pt(i++));
    }
    return ret;
}
/* :588-607 */
inline Pt barycenter_where(const Mesh& m, const Cell& cl, int where) {
    if (cl.location != ON_INTERFACE) return barycenter(m, cl);
    auto tp = collect_triangulation_points(m, cl, where);
    return barycenter_range(tp.begin(), tp.end());
}
/* :754-777 + :798-815 */
inline std::vector<QP> integrate_where(const Mesh& m, const Cell& cl, size_t degree, int where) {
    if (cl.location != ON_INTERFACE) return integrate(m, cl, degree);
    auto tp = collect_triangulation_points(m, cl, where);
    Pt bar = barycenter_range(tp.begin(), tp.end());
    std::vector<QP> ret;
    for (size_t i = 0; i < tp.size(); i++) {
        auto q = triangle_quadrature(bar, tp[i], tp[(i + 1) % tp.size()], degree);
        ret.insert(ret.end(), q.begin(), q.end());
    }
    return ret;
}
/* :817-849 */
inline std::vector<QP> integrate_where(const Mesh& m, const Face& fc, size_t degree, int where) {
    std::vector<QP> ret;
    if (fc.location != where && fc.location != ON_INTERFACE) return ret;
    if (fc.location != ON_INTERFACE) return integrate(m, fc, degree);
    auto pts = points_where(m, fc, where);
    return integrate_segment(pts[0], pts[1], degree);
}
/* :851-895 */
inline std::vector<QP> integrate_interface(const Mesh& m, const Cell& cl, size_t degree, int where) {
    std::vector<QP> ret;
    Pt pa = cl.interface.at(0), pb = cl.interface.at(1);
    Pt bar = barycenter_where(m, cl, where);
    Pt va = pa - bar, vbt = pb - pa, vb(vbt.y, -vbt.x);
    double int_sign = (va.x * vb.x + va.y * vb.y) < 0 ? -1.0 : +1.0;
    for (size_t i = 1; i < cl.interface.size(); i++) {
        auto q = integrate_segment(cl.interface.at(i - 1), cl.interface.at(i), degree, int_sign);
        ret.insert(ret.end(), q.begin(), q.end());
    }
    return ret;
}

const double CELL_ETA = 5.0;        /* cuthho_square.cpp:301-306 */

/* cuthho_square.cpp:308-388 */
inline std::pair<Mat, Mat> make_hho_laplacian_cut(const Mesh& m, const Cell& cl, const LevelSet& ls, DegreeInfo di, int where) {
    if (cl.location != ON_INTERFACE) return make_hho_laplacian(m, cl, di);
    CellBasis cb(m, cl, di.rd);
    int rbs = (int)cell_basis_size(di.rd), cbs = (int)cell_basis_size(di.cd), fbs = (int)face_basis_size(di.fd);
    auto fids = face_ids(m, cl);
    Mat stiff(rbs, rbs), gr_rhs(rbs, cbs + 4 * fbs);
    for (auto& qp : integrate_where(m, cl, 2 * di.rd, where)) {
        Mat dphi = cb.eval_gradients(qp.first);
        for (int j = 0; j < rbs; j++) for (int i = 0; i < rbs; i++)
            stiff(i, j) += qp.second * (dphi(i, 0) * dphi(j, 0) + dphi(i, 1) * dphi(j, 1));
    }
    double hT = measure(m, cl);
    for (auto& qp : integrate_interface(m, cl, 2 * di.rd, where)) {
        Vec phi = cb.eval_basis(qp.first); Mat dphi = cb.eval_gradients(qp.first); Pt n = ls.normal(qp.first);
        for (int j = 0; j < rbs; j++) {
            double dnj = dphi(j, 0) * n.x + dphi(j, 1) * n.y;
            for (int i = 0; i < rbs; i++) {
                double dni = dphi(i, 0) * n.x + dphi(i, 1) * n.y;
                stiff(i, j) -= qp.second * phi[i] * dnj;
                stiff(i, j) -= qp.second * dni * phi[j];
                stiff(i, j) += qp.second * phi[i] * phi[j] * CELL_ETA / hT;
            }
        }
    }
    Mat gr_lhs = stiff;
    gr_rhs.set_block(0, 0, stiff.block(0, 0, rbs, cbs));
    auto ns = normals(m, cl);
    for (size_t i = 0; i < 4; i++) {
        const Face& fc = m.faces[fids[i]]; Pt n = ns[i];
        FaceBasis fb(m, fc, di.fd);
        for (auto& qp : integrate_where(m, fc, 2 * di.rd, where)) {
            Vec c_phi = cb.eval_basis(qp.first), f_phi = fb.eval_basis(qp.first);
            Mat r_dphi = cb.eval_gradients(qp.first);
            for (int r = 0; r < rbs; r++) {
                double dn = r_dphi(r, 0) * n.x + r_dphi(r, 1) * n.y;
                for (int b = 0; b < fbs; b++) gr_rhs(r, cbs + (int)i * fbs + b) += qp.second * dn * f_phi[b];
                for (int c = 0; c < cbs; c++) gr_rhs(r, c) -= qp.second * dn * c_phi[c];
            }
        }
    }
    Mat oper = LLT(gr_lhs).solve(gr_rhs);
    Mat data = gr_rhs.transpose() * oper;
    return {oper, data};
}

/* cuthho_square.cpp:566-621 */
inline Mat make_hho_cut_stabilization(const Mesh& m, const Cell& cl, const DegreeInfo& di, int where) {
    if (cl.location != ON_INTERFACE) return make_hho_naive_stabilization(m, cl, di);
    int cbs = (int)cell_basis_size(di.cd), fbs = (int)face_basis_size(di.fd);
    auto fids = face_ids(m, cl);
    int msize = cbs + 4 * fbs; Mat data(msize, msize);
    CellBasis cb(m, cl, di.cd);
    double hT = measure(m, cl);
    for (size_t i = 0; i < 4; i++) {
        const Face& fc = m.faces[fids[i]]; FaceBasis fb(m, fc, di.fd);
        Mat oper(fbs, msize), mass(fbs, fbs), trace(fbs, cbs);
        for (int a = 0; a < fbs; a++) oper(a, cbs + (int)i * fbs + a) = -1.0;
        auto qps = integrate_where(m, fc, 2 * di.fd, where);
        for (auto& qp : qps) {
            Vec c_phi = cb.eval_basis(qp.first), f_phi = fb.eval_basis(qp.first);
            for (int b = 0; b < fbs; b++) for (int a = 0; a < fbs; a++) mass(a, b) += qp.second * f_phi[a] * f_phi[b];
            for (int c = 0; c < cbs; c++) for (int a = 0; a < fbs; a++) trace(a, c) += qp.second * f_phi[a] * c_phi[c];
        }
        if (qps.size() == 0) continue;
        oper.set_block(0, 0, LLT(mass).solve(trace));
        data = data + (1. / hT) * (oper.transpose() * (mass * oper));
    }
    return data;
}

/* cuthho_square.cpp:623-666 — interface quadrature degree is `degree`, not 2*degree (:647) */
inline Vec make_rhs_fictdom(const Mesh& m, const Cell& cl, size_t degree, const Fun& f, int where, const LevelSet& ls, const Fun& bcs) {
    if (cl.location == where) return make_rhs(m, cl, degree, f);
    size_t cbs = cell_basis_size(degree);
    Vec ret(cbs, 0.0);
    if (cl.location != ON_INTERFACE) return ret;
    CellBasis cb(m, cl, degree);
    double hT = measure(m, cl);
    for (auto& qp : integrate_where(m, cl, 2 * degree, where)) {
        Vec phi = cb.eval_basis(qp.first); double fv = f(qp.first);
        for (size_t i = 0; i < cbs; i++) ret[i] += qp.second * phi[i] * fv;
    }
    for (auto& qp : integrate_interface(m, cl, degree, where)) {
        Vec phi = cb.eval_basis(qp.first); Mat dphi = cb.eval_gradients(qp.first); Pt n = ls.normal(qp.first);
        double g = bcs(qp.first);
        for (size_t i = 0; i < cbs; i++)
            ret[i] += qp.second * g * (phi[i] * CELL_ETA / hT - (dphi((int)i, 0) * n.x + dphi((int)i, 1) * n.y));
    }
    return ret;
}

} // namespace orc
