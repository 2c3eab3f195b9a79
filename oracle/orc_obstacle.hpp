/* ORACLE — TEST INFRASTRUCTURE ONLY (see orc_core.hpp header).
 * Obstacle problem: restates obstacle_assembler (src/methods/hho_bits/hho.hpp:471-789) and the data of
 * apps/obstacle/obstacle.cpp:47-218.  The sparse direct solve (Eigen::SparseLU, obstacle.cpp:170-175)
 * is not restated: the saddle system is exported as CSR and solved by scipy's SuperLU in the tests. */
#pragma once
#include "orc_hho.hpp"

namespace orc {

struct ObstacleAssembler {
    std::vector<size_t> face_ct, A_ct, B_ct, face_et, A_et, B_et;
    std::vector<bool> is_in_set_A;
    DegreeInfo di; std::vector<Triplet> triplets;
    size_t num_all_faces, num_dirichlet_faces, num_other_faces, num_all_cells, num_A_cells, num_I_cells, system_size;
    Csr LHS; Vec RHS;

    ObstacleAssembler(const Mesh& m, const std::vector<bool>& in_A, DegreeInfo hdi) : is_in_set_A(in_A), di(hdi) {   /* :521-596 */
        num_all_faces = m.faces.size(); num_dirichlet_faces = 0;
        for (auto& f : m.faces) if (f.is_boundary && f.dirichlet) num_dirichlet_faces++;
        num_other_faces = num_all_faces - num_dirichlet_faces;
        num_all_cells = m.cells.size();
        num_A_cells = std::count(in_A.begin(), in_A.end(), true);
        num_I_cells = std::count(in_A.begin(), in_A.end(), false);
        A_ct.assign(num_all_cells, 0); A_et.assign(num_I_cells, 0);
        for (size_t i = 0, co = 0; i < num_all_cells; i++) if (!in_A[i]) { A_ct[i] = co; A_et[co] = i; co++; }
        face_ct.assign(num_all_faces, 0); face_et.assign(num_other_faces, 0);
        for (size_t i = 0, co = 0; i < num_all_faces; i++)
            if (!(m.faces[i].is_boundary && m.faces[i].dirichlet)) { face_ct[i] = co; face_et[co] = i; co++; }
        B_ct.assign(num_all_cells, 0); B_et.assign(num_A_cells, 0);
        for (size_t i = 0, co = 0; i < num_all_cells; i++) if (in_A[i]) { B_ct[i] = co; B_et[co] = i; co++; }
        size_t cbs = cell_basis_size(di.cd), fbs = face_basis_size(di.fd);
        system_size = cbs * (num_I_cells + num_A_cells) + fbs * num_other_faces;
        RHS.assign(system_size, 0.0);
    }

    void assemble(const Mesh& m, const Cell& cl, const Mat& lhs, const Vec& rhs, const Vec& gamma, const Fun& dirichlet_bf) {   /* :609-695 */
        size_t cbs = cell_basis_size(di.cd), fbs = face_basis_size(di.fd);
        std::vector<std::pair<size_t, bool>> row, col;
        size_t cell_offset = offset(m, cl);
        size_t cell_LHS_offset = A_ct.at(cell_offset) * cbs;
        bool needs_A = !is_in_set_A.at(cell_offset), needs_B = is_in_set_A.at(cell_offset);
        for (size_t i = 0; i < cbs; i++) { row.push_back({cell_offset + i, true}); col.push_back({cell_LHS_offset + i, needs_A}); }
        auto fids = face_ids(m, cl);
        Vec dirichlet_data(cbs + 4 * fbs, 0.0);
        for (size_t fi = 0; fi < 4; fi++) {
            const Face& fc = m.faces[fids[fi]];
            size_t rows_off = cbs * num_all_cells + face_ct.at(fids[fi]) * fbs;
            size_t cols_off = cbs * num_I_cells + face_ct.at(fids[fi]) * fbs;
            bool dirichlet = fc.is_boundary && fc.dirichlet;
            for (size_t i = 0; i < fbs; i++) { row.push_back({rows_off + i, !dirichlet}); col.push_back({cols_off + i, !dirichlet}); }
            if (dirichlet) {
                Vec v = LLT(make_mass_matrix(m, fc, di.fd)).solve(make_rhs(m, fc, di.fd, dirichlet_bf));
                std::copy(v.begin(), v.end(), dirichlet_data.begin() + cbs + fi * fbs);
            }
        }
        for (int i = 0; i < lhs.r; i++) {
            if (!row[i].second) continue;
            for (int j = 0; j < lhs.c; j++) {
                if (col[j].second) triplets.push_back({row[i].first, col[j].first, lhs(i, j)});
                else if ((size_t)j < cbs) RHS[row[i].first] -= lhs(i, j) * gamma[cell_offset];
                else RHS[row[i].first] -= lhs(i, j) * dirichlet_data[j];
            }
        }
        for (size_t i = 0; i < cbs; i++) RHS[cell_offset + i] += rhs[i];
        if (needs_B) {
            size_t orow = cell_offset * cbs;
            size_t ocol = num_I_cells * cbs + num_other_faces * fbs + B_ct.at(cell_offset);
            triplets.push_back({orow, ocol, 1.0});
        }
    }

    void expand_solution(const Mesh& m, const Vec& solution, const Fun& dirichlet_bf, const Vec& gamma, Vec& alpha, Vec& beta) {   /* :698-744 */
        size_t cbs = cell_basis_size(di.cd), fbs = face_basis_size(di.fd);
        for (size_t i = 0; i < num_all_cells * cbs; i++) alpha[i] = gamma[i];
        for (size_t i = 0; i < num_I_cells; i++) { size_t e = A_et.at(i); for (size_t k = 0; k < cbs; k++) alpha[e * cbs + k] = solution[i * cbs + k]; }
        beta.assign(m.cells.size(), 0.0);
        for (size_t i = 0; i < num_A_cells; i++) {
            size_t e = B_et.at(i);
            for (size_t k = 0; k < cbs; k++) beta[e * cbs + k] = solution[num_I_cells * cbs + num_other_faces * fbs + i * cbs + k];
        }
        for (size_t fo = 0; fo < m.faces.size(); fo++) {
            const Face& fc = m.faces[fo];
            if (fc.is_boundary && fc.dirichlet) {
                Vec v = LLT(make_mass_matrix(m, fc, di.fd)).solve(make_rhs(m, fc, di.fd, dirichlet_bf));
                for (size_t k = 0; k < fbs; k++) alpha[num_all_cells * cbs + fo * fbs + k] = v[k];
            } else {
                size_t so = cbs * num_I_cells + face_ct.at(fo) * fbs;
                for (size_t k = 0; k < fbs; k++) alpha[num_all_cells * cbs + fo * fbs + k] = solution[so + k];
            }
        }
    }
    void finalize() { LHS = csr_from_triplets(system_size, triplets); triplets.clear(); }
};

/* free take_local_data on the expanded solution, hho.hpp:753-781 */
inline Vec take_local_data_expanded(const Mesh& m, const Cell& cl, DegreeInfo di, const Vec& expanded) {
    size_t cbs = cell_basis_size(di.cd), fbs = face_basis_size(di.fd);
    size_t co = offset(m, cl) * cbs; auto fids = face_ids(m, cl);
    Vec ret(cbs + 4 * fbs, 0.0);
    for (size_t i = 0; i < cbs; i++) ret[i] = expanded[co + i];
    for (size_t fi = 0; fi < 4; fi++) {
        size_t so = cbs * m.cells.size() + fids[fi] * fbs;
        for (size_t i = 0; i < fbs; i++) ret[cbs + fi * fbs + i] = expanded[so + i];
    }
    return ret;
}

/* obstacle.cpp:65-89 problem data, r0 = 0.7 */
inline double obstacle_rhs(Pt pt) {
    double r0 = 0.7, r = std::sqrt(pt.x * pt.x + pt.y * pt.y);
    if (r > r0) return -16 * r * r + 8 * r0 * r0;
    return -8.0 * (r0 * r0 * (r0 * r0 + 1)) + 8 * r0 * r0 * r * r;
}
inline double obstacle_sol(Pt pt) {
    double r0 = 0.7, r = std::sqrt(pt.x * pt.x + pt.y * pt.y);
    double s = r * r - r0 * r0, t = std::max(s, 0.0);
    return t * t;
}

} // namespace orc
