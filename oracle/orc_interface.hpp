/* ORACLE — TEST INFRASTRUCTURE ONLY (see orc_core.hpp header).
 * CPU restatement of the two-material INTERFACE problem of the cut driver (`cuthho_square -i`):
 *   make_hho_laplacian_interface   apps/cuthho/cuthho_square.cpp:390-502
 *   interface_assembler            apps/cuthho/cuthho_square.cpp:1091-1436
 *   run_cuthho_interface           apps/cuthho/cuthho_square.cpp:1625-1846 (assembly loop :1667-1716, H1 error :1769-1835)
 *   make_rhs(msh, cl, deg, where, f)  src/methods/cuthho_bits/cuthho_utils.hpp:65-86
 * Pinned against the "Interface" table of apps/cuthho/cuthho.xlsx (tests/golden/reference_golden.json,
 * tests/test_oracle_golden.py).  No GPU path exists for this problem yet: this file is the checker the next
 * row of the scope table will be built against. */
#pragma once
#include "orc_cut.hpp"

namespace orc {

struct InterfaceParams { double kappa_1 = 1.0, kappa_2 = 1.0, eta = 5.0; };     /* cuthho_square.cpp:293-299 */

/* cuthho_square.cpp:390-502.  Returns `data` = gr_rhs^T gr_lhs^+ gr_rhs (2n x 2n, n = cbs + 4 fbs, unknown order
 * [cell-, cell+, faces-, faces+]).  gr_lhs is singular along (1, 1) (the same constant on both sides: the gradient,
 * consistency and penalty terms all vanish on it); the reference calls Eigen's pivoted ldlt().solve on it.  Every
 * column of gr_rhs is orthogonal to that vector, so `data` does not depend on which solution the factorisation
 * picks: the restatement solves with gr_lhs + v v^T, v = e_0 + e_rbs (positive definite, same `data`). */
inline Mat make_hho_laplacian_interface(const Mesh& m, const Cell& cl, const LevelSet& ls, DegreeInfo di, const InterfaceParams& parms) {
    if (cl.location != ON_INTERFACE) throw std::invalid_argument("The cell is not cut");
    CellBasis cb(m, cl, di.rd);
    const int rbs = (int)cell_basis_size(di.rd), cbs = (int)cell_basis_size(di.cd), fbs = (int)face_basis_size(di.fd);
    auto fids = face_ids(m, cl);
    const int nf = 4;
    Mat stiff(2 * rbs, 2 * rbs), gr_rhs(2 * rbs, 2 * (cbs + nf * fbs));
    const double kap[2] = {parms.kappa_1, parms.kappa_2};
    const int side[2] = {IN_NEGATIVE_SIDE, IN_POSITIVE_SIDE};
    for (int s = 0; s < 2; s++)
        for (auto& qp : integrate_where(m, cl, 2 * di.rd, side[s])) {
            Mat dphi = cb.eval_gradients(qp.first);
            for (int j = 0; j < rbs; j++) for (int i = 0; i < rbs; i++)
                stiff(s * rbs + i, s * rbs + j) += kap[s] * qp.second * (dphi(i, 0) * dphi(j, 0) + dphi(i, 1) * dphi(j, 1));
        }
    const double hT = measure(m, cl);
    for (auto& qp : integrate_interface(m, cl, 2 * di.rd, IN_NEGATIVE_SIDE)) {
        Vec phi = cb.eval_basis(qp.first); Mat dphi = cb.eval_gradients(qp.first); Pt n = ls.normal(qp.first);
        for (int j = 0; j < rbs; j++) {
            const double dnj = dphi(j, 0) * n.x + dphi(j, 1) * n.y;
            for (int i = 0; i < rbs; i++) {
                const double dni = dphi(i, 0) * n.x + dphi(i, 1) * n.y;
                const double a = parms.kappa_1 * qp.second * phi[i] * dnj;
                const double b = parms.kappa_1 * qp.second * dni * phi[j];
                const double c = parms.kappa_1 * qp.second * phi[i] * phi[j] * parms.eta / hT;
                stiff(i, j) -= a;        stiff(rbs + i, j) += a;
                stiff(i, j) -= b;        stiff(i, rbs + j) += b;
                stiff(i, j) += c;        stiff(i, rbs + j) -= c;
                stiff(rbs + i, j) -= c;  stiff(rbs + i, rbs + j) += c;
            }
        }
    }
    Mat gr_lhs = stiff;
    gr_rhs.set_block(0, 0, stiff.block(0, 0, 2 * rbs, cbs));
    gr_rhs.set_block(0, cbs, stiff.block(0, rbs, 2 * rbs, cbs));
    auto ns = normals(m, cl);
    for (int i = 0; i < nf; i++) {
        const Face& fc = m.faces[fids[i]]; Pt n = ns[i];
        FaceBasis fb(m, fc, di.fd);
        for (int s = 0; s < 2; s++)
            for (auto& qp : integrate_where(m, fc, 2 * di.rd, side[s])) {
                Vec c_phi = cb.eval_basis(qp.first), f_phi = fb.eval_basis(qp.first);
                Mat r_dphi = cb.eval_gradients(qp.first);
                const int col_ofs = 2 * cbs + s * fbs * nf + i * fbs;
                for (int r = 0; r < rbs; r++) {
                    const double dn = kap[s] * qp.second * (r_dphi(r, 0) * n.x + r_dphi(r, 1) * n.y);
                    for (int c = 0; c < cbs; c++) gr_rhs(s * rbs + r, s * cbs + c) -= dn * c_phi[c];
                    for (int b = 0; b < fbs; b++) gr_rhs(s * rbs + r, col_ofs + b) += dn * f_phi[b];
                }
            }
    }
    Mat reg = gr_lhs;
    reg(0, 0) += 1.0; reg(0, rbs) += 1.0; reg(rbs, 0) += 1.0; reg(rbs, rbs) += 1.0;
    Mat oper = LLT(reg).solve(gr_rhs);
    return gr_rhs.transpose() * oper;
}

/* cuthho_utils.hpp:65-86 */
inline Vec make_rhs_where(const Mesh& m, const Cell& cl, size_t degree, int where, const Fun& f) {
    CellBasis cb(m, cl, degree);
    Vec ret(cb.size, 0.0);
    for (auto& qp : integrate_where(m, cl, 2 * degree, where)) {
        Vec phi = cb.eval_basis(qp.first); const double fv = f(qp.first);
        for (size_t i = 0; i < cb.size; i++) ret[i] += qp.second * phi[i] * fv;
    }
    return ret;
}

/* cuthho_square.cpp:1091-1436 */
struct InterfaceAssembler {
    std::vector<size_t> cell_table, face_table;
    size_t num_all_cells = 0, num_all_faces = 0, system_size = 0;
    DegreeInfo di; std::vector<Triplet> triplets;
    Csr LHS; Vec RHS;
    InterfaceAssembler(const Mesh& m, DegreeInfo hdi) : di(hdi) {
        auto is_dirichlet = [](const Face& fc) { return fc.is_boundary && fc.dirichlet; };
        for (auto& cl : m.cells) { cell_table.push_back(num_all_cells); num_all_cells += cl.location == ON_INTERFACE ? 2 : 1; }
        size_t ndir = 0;
        for (auto& fc : m.faces) { num_all_faces += fc.location == ON_INTERFACE ? 2 : 1; if (is_dirichlet(fc)) ndir++; }
        face_table.assign(m.faces.size(), 0);
        size_t co = 0;
        for (size_t i = 0; i < m.faces.size(); i++)
            if (!is_dirichlet(m.faces[i])) { face_table[i] = co; co += m.faces[i].location == ON_INTERFACE ? 2 : 1; }
        system_size = cell_basis_size(di.cd) * num_all_cells + face_basis_size(di.fd) * (num_all_faces - ndir);
        RHS.assign(system_size, 0.0);
    }
    /* :1197-1263 (uncut cells) */
    void assemble(const Mesh& m, const Cell& cl, const Mat& lhs, const Vec& rhs, const Fun& dirichlet_bf) {
        if (cl.location == ON_INTERFACE) throw std::invalid_argument("UNcut cell expected.");
        const size_t cbs = cell_basis_size(di.cd), fbs = face_basis_size(di.fd);
        auto fids = face_ids(m, cl);
        std::vector<std::pair<size_t, bool>> asm_map;
        const size_t cell_LHS_offset = cell_table.at(offset(m, cl)) * cbs;
        for (size_t i = 0; i < cbs; i++) asm_map.push_back({cell_LHS_offset + i, true});
        Vec dirichlet_data(cbs + 4 * fbs, 0.0);
        for (size_t fi = 0; fi < 4; fi++) {
            const Face& fc = m.faces[fids[fi]];
            const size_t face_LHS_offset = num_all_cells * cbs + face_table.at(fids[fi]) * fbs;
            const bool dirichlet = fc.is_boundary && fc.dirichlet;
            for (size_t i = 0; i < fbs; i++) asm_map.push_back({face_LHS_offset + i, !dirichlet});
            if (dirichlet) {
                Vec v = LLT(make_mass_matrix(m, fc, di.fd)).solve(make_rhs(m, fc, di.fd, dirichlet_bf));
                std::copy(v.begin(), v.end(), dirichlet_data.begin() + cbs + fi * fbs);
            }
        }
        for (int i = 0; i < lhs.r; i++) {
            if (!asm_map[i].second) continue;
            for (int j = 0; j < lhs.c; j++) {
                if (asm_map[j].second) triplets.push_back({asm_map[i].first, asm_map[j].first, lhs(i, j)});
                else RHS[asm_map[i].first] -= lhs(i, j) * dirichlet_data[j];
            }
        }
        for (size_t i = 0; i < cbs; i++) RHS[cell_LHS_offset + i] += rhs[i];
    }
    /* :1265-1340 (cut cells: unknown order [cell-, cell+, faces-, faces+]; the second copy of a face that is not cut is
     * the same unknown as the first, `d = 0`) */
    void assemble_cut(const Mesh& m, const Cell& cl, const Mat& lhs, const Vec& rhs) {
        if (cl.location != ON_INTERFACE) throw std::invalid_argument("Cut cell expected.");
        const size_t cbs = cell_basis_size(di.cd), fbs = face_basis_size(di.fd);
        auto fids = face_ids(m, cl);
        std::vector<size_t> asm_map;
        const size_t cell_LHS_offset = cell_table.at(offset(m, cl)) * cbs;
        for (size_t i = 0; i < 2 * cbs; i++) asm_map.push_back(cell_LHS_offset + i);
        for (int pass = 0; pass < 2; pass++)
            for (size_t fi = 0; fi < 4; fi++) {
                const Face& fc = m.faces[fids[fi]];
                if (fc.is_boundary && fc.dirichlet) throw std::invalid_argument("Dirichlet boundary on cut cell not supported.");
                const size_t d = (pass == 1 && fc.location == ON_INTERFACE) ? fbs : 0;
                const size_t face_LHS_offset = num_all_cells * cbs + face_table.at(fids[fi]) * fbs + d;
                for (size_t i = 0; i < fbs; i++) asm_map.push_back(face_LHS_offset + i);
            }
        if ((int)asm_map.size() != lhs.r || lhs.r != lhs.c) throw std::logic_error("assemble_cut: size mismatch");
        for (int i = 0; i < lhs.r; i++) for (int j = 0; j < lhs.c; j++) triplets.push_back({asm_map[i], asm_map[j], lhs(i, j)});
        for (size_t i = 0; i < 2 * cbs; i++) RHS[cell_LHS_offset + i] += rhs[i];
    }
    /* cell part of take_local_data (:1343-1365): the H1 error of the driver only reads head(cbs) */
    Vec cell_dofs(const Mesh& m, const Cell& cl, const Vec& solution, int where) const {
        const size_t cbs = cell_basis_size(di.cd);
        size_t o = cell_table.at(offset(m, cl)) * cbs;
        if (cl.location == ON_INTERFACE && where == IN_POSITIVE_SIDE) o += cbs;
        return Vec(solution.begin() + o, solution.begin() + o + cbs);
    }
    void finalize() { LHS = csr_from_triplets(system_size, triplets); triplets.clear(); triplets.shrink_to_fit(); }
};

/* assembly loop of run_cuthho_interface, cuthho_square.cpp:1667-1716 (kappa_2 = 1, :1659) */
inline InterfaceAssembler assemble_interface_problem(const Mesh& m, const LevelSet& ls, size_t degree, const Fun& rhs_fun, const Fun& bcs_fun,
                                                     const InterfaceParams& parms = InterfaceParams()) {
    DegreeInfo hdi(degree + 1, degree);
    InterfaceAssembler as(m, hdi);
    const int cbs = (int)cell_basis_size(hdi.cd), fbs = (int)face_basis_size(hdi.fd), nfd = 4 * fbs;
    for (auto& cl : m.cells) {
        if (cl.location != ON_INTERFACE) {
            const double kappa = cl.location == IN_NEGATIVE_SIDE ? parms.kappa_1 : parms.kappa_2;
            auto gr = make_hho_laplacian(m, cl, hdi);
            Mat stab = make_hho_naive_stabilization(m, cl, hdi);
            Mat lc = stab; lc.add_block(0, 0, gr.second, kappa);
            as.assemble(m, cl, lc, make_rhs(m, cl, hdi.cd, rhs_fun), bcs_fun);
        } else {
            Mat lc = make_hho_laplacian_interface(m, cl, ls, hdi, parms);
            Mat sn = make_hho_cut_stabilization(m, cl, hdi, IN_NEGATIVE_SIDE), sp = make_hho_cut_stabilization(m, cl, hdi, IN_POSITIVE_SIDE);
            lc.add_block(0, 0, sn.block(0, 0, cbs, cbs), parms.kappa_1);
            lc.add_block(0, 2 * cbs, sn.block(0, cbs, cbs, nfd), parms.kappa_1);
            lc.add_block(2 * cbs, 0, sn.block(cbs, 0, nfd, cbs), parms.kappa_1);
            lc.add_block(2 * cbs, 2 * cbs, sn.block(cbs, cbs, nfd, nfd), parms.kappa_1);
            lc.add_block(cbs, cbs, sp.block(0, 0, cbs, cbs), parms.kappa_2);
            lc.add_block(cbs, 2 * cbs + nfd, sp.block(0, cbs, cbs, nfd), parms.kappa_2);
            lc.add_block(2 * cbs + nfd, cbs, sp.block(cbs, 0, nfd, cbs), parms.kappa_2);
            lc.add_block(2 * cbs + nfd, 2 * cbs + nfd, sp.block(cbs, cbs, nfd, nfd), parms.kappa_2);
            Vec f(2 * cbs, 0.0);
            Vec fn = make_rhs_where(m, cl, hdi.cd, IN_NEGATIVE_SIDE, rhs_fun), fp = make_rhs_where(m, cl, hdi.cd, IN_POSITIVE_SIDE, rhs_fun);
            std::copy(fn.begin(), fn.end(), f.begin()); std::copy(fp.begin(), fp.end(), f.begin() + cbs);
            as.assemble_cut(m, cl, lc, f);
        }
    }
    as.finalize();
    return as;
}

/* H1 error of run_cuthho_interface, cuthho_square.cpp:1769-1835: gradient of the CELL dofs (i >= 1), degree 2*cd, both
 * parts of a cut cell; grad(p, g) = exact gradient */
template <typename Grad>
inline double interface_h1_error_sq(const Mesh& m, const InterfaceAssembler& as, const Vec& sol, const Grad& grad) {
    const DegreeInfo& hdi = as.di; double H1 = 0.0;
    for (auto& cl : m.cells) {
        CellBasis cb(m, cl, hdi.cd); const size_t cbs = cb.size;
        auto accumulate = [&](const std::vector<QP>& qps, const Vec& dofs) {
            for (auto& qp : qps) {
                Mat dphi = cb.eval_gradients(qp.first); double gx = 0, gy = 0;
                for (size_t i = 1; i < cbs; i++) { gx += dofs[i] * dphi((int)i, 0); gy += dofs[i] * dphi((int)i, 1); }
                double rg[2]; grad(qp.first, rg);
                H1 += qp.second * ((rg[0] - gx) * (rg[0] - gx) + (rg[1] - gy) * (rg[1] - gy));
            }
        };
        if (cl.location == ON_INTERFACE) {
            accumulate(integrate_where(m, cl, 2 * hdi.cd, IN_NEGATIVE_SIDE), as.cell_dofs(m, cl, sol, IN_NEGATIVE_SIDE));
            accumulate(integrate_where(m, cl, 2 * hdi.cd, IN_POSITIVE_SIDE), as.cell_dofs(m, cl, sol, IN_POSITIVE_SIDE));
        } else accumulate(integrate(m, cl, 2 * hdi.cd), as.cell_dofs(m, cl, sol, IN_POSITIVE_SIDE));
    }
    return H1;
}

} // namespace orc
