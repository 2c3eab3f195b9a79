/* ORACLE — TEST INFRASTRUCTURE ONLY (see orc_core.hpp header).
 * HHO local operators and global assemblers: restates src/core/core_bits/utils.hpp,
 * src/methods/hho_bits/hho.hpp and src/core/core_bits/solver_cg.hpp. */
#pragma once
#include "orc_quad.hpp"
#include <functional>
#include <map>

namespace orc {

typedef std::function<double(Pt)> Fun;

struct DegreeInfo {                 /* hho_degree_info, utils.hpp:62-111 */
    size_t cd, fd, rd;
    DegreeInfo(size_t c, size_t f) {
        bool c1 = f > 0 && (c == f - 1 || c == f || c == f + 1);
        bool c2 = f == 0 && (c == f || c == f + 1);
        if (c1 || c2) { cd = c; fd = f; rd = f + 1; }
        else { cd = f; fd = f; rd = f + 1; }   /* "Invalid cell degree. Reverting to equal-order" */
    }
};

/* utils.hpp:113-151 */
inline Mat make_mass_matrix(const Mesh& m, const Cell& cl, size_t degree, size_t di = 0) {
    CellBasis cb(m, cl, degree); int n = (int)cb.size; Mat ret(n, n);
    for (auto& qp : integrate(m, cl, 2 * (degree + di))) {
        Vec phi = cb.eval_basis(qp.first);
        for (int j = 0; j < n; j++) for (int i = 0; i < n; i++) ret(i, j) += qp.second * phi[i] * phi[j];
    }
    return ret;
}
inline Mat make_mass_matrix(const Mesh& m, const Face& fc, size_t degree, size_t di = 0) {
    FaceBasis fb(m, fc, degree); int n = (int)fb.size; Mat ret(n, n);
    for (auto& qp : integrate(m, fc, 2 * (degree + di))) {
        Vec phi = fb.eval_basis(qp.first);
        for (int j = 0; j < n; j++) for (int i = 0; i < n; i++) ret(i, j) += qp.second * phi[i] * phi[j];
    }
    return ret;
}
/* utils.hpp:153-199 */
inline Vec make_rhs(const Mesh& m, const Cell& cl, size_t degree, const Fun& f, size_t di = 0) {
    CellBasis cb(m, cl, degree); Vec ret(cb.size, 0.0);
    for (auto& qp : integrate(m, cl, 2 * (degree + di))) {
        Vec phi = cb.eval_basis(qp.first); double fv = f(qp.first);
        for (size_t i = 0; i < cb.size; i++) ret[i] += qp.second * phi[i] * fv;
    }
    return ret;
}
inline Vec make_rhs(const Mesh& m, const Face& fc, size_t degree, const Fun& f, size_t di = 0) {
    FaceBasis fb(m, fc, degree); Vec ret(fb.size, 0.0);
    for (auto& qp : integrate(m, fc, 2 * (degree + di))) {
        Vec phi = fb.eval_basis(qp.first); double fv = f(qp.first);
        for (size_t i = 0; i < fb.size; i++) ret[i] += qp.second * phi[i] * fv;
    }
    return ret;
}
/* utils.hpp:201-231 */
inline Vec project_function(const Mesh& m, const Cell& cl, DegreeInfo hdi, const Fun& f, size_t di = 0) {
    size_t cbs = cell_basis_size(hdi.cd), fbs = face_basis_size(hdi.fd);
    auto fids = face_ids(m, cl);
    Vec ret(cbs + 4 * fbs, 0.0);
    Vec c = LLT(make_mass_matrix(m, cl, hdi.cd, di)).solve(make_rhs(m, cl, hdi.cd, f, di));
    std::copy(c.begin(), c.end(), ret.begin());
    for (size_t i = 0; i < 4; i++) {
        const Face& fc = m.faces[fids[i]];
        Vec v = LLT(make_mass_matrix(m, fc, hdi.fd, di)).solve(make_rhs(m, fc, hdi.fd, f, di));
        std::copy(v.begin(), v.end(), ret.begin() + cbs + i * fbs);
    }
    return ret;
}

/* hho.hpp:32-96 — gradient reconstruction.  Returns (oper, data). */
inline std::pair<Mat, Mat> make_hho_laplacian(const Mesh& m, const Cell& cl, const DegreeInfo& di) {
    CellBasis cb(m, cl, di.rd);
    int rbs = (int)cell_basis_size(di.rd), cbs = (int)cell_basis_size(di.cd), fbs = (int)face_basis_size(di.fd);
    auto fids = face_ids(m, cl);
    Mat stiff(rbs, rbs), gr_rhs(rbs - 1, cbs + 4 * fbs);
    for (auto& qp : integrate(m, cl, 2 * di.rd)) {
        Mat dphi = cb.eval_gradients(qp.first);
        for (int j = 0; j < rbs; j++) for (int i = 0; i < rbs; i++)
            stiff(i, j) += qp.second * (dphi(i, 0) * dphi(j, 0) + dphi(i, 1) * dphi(j, 1));
    }
    Mat gr_lhs = stiff.block(1, 1, rbs - 1, rbs - 1);
    gr_rhs.set_block(0, 0, stiff.block(1, 0, rbs - 1, cbs));
    auto ns = normals(m, cl);
    for (size_t i = 0; i < 4; i++) {
        const Face& fc = m.faces[fids[i]]; Pt n = ns[i];
        FaceBasis fb(m, fc, di.fd);
        for (auto& qp : integrate(m, fc, 2 * di.fd)) {
            Vec c_phi = cb.eval_basis(qp.first);
            Mat c_dphi = cb.eval_gradients(qp.first);
            Vec f_phi = fb.eval_basis(qp.first);
            for (int r = 0; r < rbs - 1; r++) {
                double dn = c_dphi(r + 1, 0) * n.x + c_dphi(r + 1, 1) * n.y;
                for (int b = 0; b < fbs; b++) gr_rhs(r, cbs + (int)i * fbs + b) += qp.second * dn * f_phi[b];
                for (int c = 0; c < cbs; c++) gr_rhs(r, c) -= qp.second * dn * c_phi[c];
            }
        }
    }
    Mat oper = LLT(gr_lhs).solve(gr_rhs);
    Mat data = gr_rhs.transpose() * oper;
    return {oper, data};
}

/* hho.hpp:99-148 — scaled by 1/area */
inline Mat make_hho_naive_stabilization(const Mesh& m, const Cell& cl, const DegreeInfo& di) {
    int cbs = (int)cell_basis_size(di.cd), fbs = (int)face_basis_size(di.fd);
    auto fids = face_ids(m, cl);
    int msize = cbs + 4 * fbs; Mat data(msize, msize);
    CellBasis cb(m, cl, di.cd);
    double h = measure(m, cl);
    for (size_t i = 0; i < 4; i++) {
        const Face& fc = m.faces[fids[i]]; FaceBasis fb(m, fc, di.fd);
        Mat oper(fbs, msize), mass(fbs, fbs), trace(fbs, cbs);
        for (int a = 0; a < fbs; a++) oper(a, cbs + (int)i * fbs + a) = -1.0;
        for (auto& qp : integrate(m, fc, 2 * di.fd)) {
            Vec c_phi = cb.eval_basis(qp.first), f_phi = fb.eval_basis(qp.first);
            for (int b = 0; b < fbs; b++) for (int a = 0; a < fbs; a++) mass(a, b) += qp.second * f_phi[a] * f_phi[b];
            for (int c = 0; c < cbs; c++) for (int a = 0; a < fbs; a++) trace(a, c) += qp.second * f_phi[a] * c_phi[c];
        }
        oper.set_block(0, 0, LLT(mass).solve(trace));
        data = data + (1. / h) * (oper.transpose() * (mass * oper));
    }
    return data;
}

/* hho.hpp:155-237 — full HHO stabilisation, scaled by 1/diam(T) on every face */
inline Mat make_hho_fancy_stabilization(const Mesh& m, const Cell& cl, const Mat& reconstruction, const DegreeInfo& di) {
    int rbs = (int)cell_basis_size(di.rd), cbs = (int)cell_basis_size(di.cd), fbs = (int)face_basis_size(di.fd);
    CellBasis cb(m, cl, di.rd);
    Mat mass_mat(rbs, rbs);
    for (auto& qp : integrate(m, cl, 2 * di.rd)) {
        Vec c_phi = cb.eval_basis(qp.first);
        for (int j = 0; j < rbs; j++) for (int i = 0; i < rbs; i++) mass_mat(i, j) += qp.second * c_phi[i] * c_phi[j];
    }
    Mat M1 = mass_mat.block(0, 0, cbs, cbs), M2 = mass_mat.block(0, 1, cbs, rbs - 1);
    Mat proj1 = -1.0 * LLT(M1).solve(M2 * reconstruction);
    for (int i = 0; i < cbs; i++) proj1(i, i) += 1.0;
    auto fids = face_ids(m, cl);
    int msize = cbs + 4 * fbs; Mat data(msize, msize);
    for (size_t fi = 0; fi < 4; fi++) {
        double h = diameter(m, cl);
        const Face& fc = m.faces[fids[fi]]; FaceBasis fb(m, fc, di.fd);
        Mat fmass(fbs, fbs), ftrace(fbs, rbs);
        for (auto& qp : integrate(m, fc, 2 * di.fd)) {
            Vec f_phi = fb.eval_basis(qp.first), c_phi = cb.eval_basis(qp.first);
            for (int b = 0; b < fbs; b++) for (int a = 0; a < fbs; a++) fmass(a, b) += (qp.second * f_phi[a]) * f_phi[b];
            for (int c = 0; c < rbs; c++) for (int a = 0; a < fbs; a++) ftrace(a, c) += (qp.second * f_phi[a]) * c_phi[c];
        }
        LLT piKF(fmass);
        Mat MR1 = ftrace.block(0, 1, fbs, rbs - 1);
        Mat proj2 = piKF.solve(MR1 * reconstruction);
        for (int a = 0; a < fbs; a++) proj2(a, cbs + (int)fi * fbs + a) -= 1.0;
        Mat MR2 = ftrace.block(0, 0, fbs, cbs);
        Mat proj3 = piKF.solve(MR2 * proj1);
        Mat BRF = proj2 + proj3;
        data = data + (1.0 / h) * (BRF.transpose() * (fmass * BRF));
    }
    return data;
}

/* ------------------------------------------------------------------ global system containers */
struct Triplet { size_t r, c; double v; };
struct Csr {                         /* what SparseMatrix::setFromTriplets produces, as sorted CSR */
    size_t n = 0; std::vector<long long> rowptr; std::vector<int> col; std::vector<double> val;
    Vec diagonal() const {
        Vec d(n, 0.0);
        for (size_t i = 0; i < n; i++) for (long long k = rowptr[i]; k < rowptr[i + 1]; k++) if ((size_t)col[k] == i) d[i] = val[k];
        return d;
    }
    void mul(const Vec& x, Vec& y) const {
        for (size_t i = 0; i < n; i++) { double s = 0; for (long long k = rowptr[i]; k < rowptr[i + 1]; k++) s += val[k] * x[col[k]]; y[i] = s; }
    }
};
/* hho.hpp:451-455 (setFromTriplets): duplicates summed, explicit zeros kept.
 * Eigen is not under /root/reference (un-vendored dependency, version unpinned: cmake/FindEigen3.cmake:18-30); this restates the
 * published algorithm of Eigen 3's internal::set_from_triplets (Eigen/src/SparseCore/SparseMatrix.h), which is LINEAR in the number
 * of triplets — no global sort:
 *   1. a temporary matrix of the OPPOSITE storage order: count the entries of every outer vector, reserve, append every triplet
 *      to its outer vector in input order (insertBackUncompressed);
 *   2. collapseDuplicates: per outer vector, a mask array over the inner index remembers where an inner index was first stored;
 *      later occurrences are added to it (so duplicates are summed in input order);
 *   3. the assignment temporary -> result transposes the storage order with a counting pass, which leaves every outer vector of
 *      the result sorted by inner index.
 * The reference's LHS is column-major, the oracle's Csr row-major: the roles of rows and columns are mirrored, the work is the same.
 * orc_finalize_kind = 1 selects the earlier std::sort-based build (same matrix up to the summation order of duplicates), kept so
 * that bench.py can report both timings. */
static int orc_finalize_kind = 0;
inline Csr csr_from_triplets_sort(size_t n, std::vector<Triplet>& t) {
    std::sort(t.begin(), t.end(), [](const Triplet& a, const Triplet& b) { return a.r != b.r ? a.r < b.r : a.c < b.c; });
    Csr A; A.n = n; A.rowptr.assign(n + 1, 0);
    for (size_t i = 0; i < t.size();) {
        size_t j = i; double s = 0;
        while (j < t.size() && t[j].r == t[i].r && t[j].c == t[i].c) s += t[j++].v;
        A.col.push_back((int)t[i].c); A.val.push_back(s); A.rowptr[t[i].r + 1]++;
        i = j;
    }
    for (size_t i = 0; i < n; i++) A.rowptr[i + 1] += A.rowptr[i];
    return A;
}
inline Csr csr_from_triplets(size_t n, std::vector<Triplet>& t) {
    if (orc_finalize_kind == 1) return csr_from_triplets_sort(n, t);
    /* 1. temporary, outer = column */
    std::vector<long long> cptr(n + 1, 0);
    for (const Triplet& e : t) cptr[e.c + 1]++;
    for (size_t j = 0; j < n; j++) cptr[j + 1] += cptr[j];
    std::vector<int> trow(t.size()); std::vector<double> tval(t.size());
    {
        std::vector<long long> fill(cptr.begin(), cptr.end() - 1);
        for (const Triplet& e : t) { const long long p = fill[e.c]++; trow[p] = (int)e.r; tval[p] = e.v; }
    }
    /* 2. collapse duplicates inside every column; cend[j] = one past the last kept entry of column j */
    std::vector<long long> cend(n, 0), mask(n, -1);
    for (size_t j = 0; j < n; j++) {
        const long long start = cptr[j]; long long w = start;
        for (long long k = start; k < cptr[j + 1]; k++) {
            const int i = trow[k];
            if (mask[i] >= start) tval[mask[i]] += tval[k];
            else { trow[w] = i; tval[w] = tval[k]; mask[i] = w; w++; }
        }
        cend[j] = w;
    }
    /* 3. transposing copy: rows of the result come out sorted by column */
    Csr A; A.n = n; A.rowptr.assign(n + 1, 0);
    for (size_t j = 0; j < n; j++) for (long long k = cptr[j]; k < cend[j]; k++) A.rowptr[trow[k] + 1]++;
    for (size_t i = 0; i < n; i++) A.rowptr[i + 1] += A.rowptr[i];
    A.col.resize((size_t)A.rowptr[n]); A.val.resize((size_t)A.rowptr[n]);
    std::vector<long long> fill(A.rowptr.begin(), A.rowptr.end() - 1);
    for (size_t j = 0; j < n; j++) for (long long k = cptr[j]; k < cend[j]; k++) { const long long p = fill[trow[k]]++; A.col[p] = (int)j; A.val[p] = tval[k]; }
    return A;
}

/* hho.hpp:252-463 */
struct Assembler {
    std::vector<size_t> compress_table, expand_table;
    DegreeInfo di; std::vector<Triplet> triplets;
    size_t system_size = 0, ncells = 0;
    Csr LHS; Vec RHS;
    Assembler(const Mesh& m, DegreeInfo hdi) : di(hdi) {
        size_t nall = m.faces.size(), ndir = 0;
        for (auto& f : m.faces) if (f.is_boundary && f.dirichlet) ndir++;
        compress_table.assign(nall, 0); expand_table.assign(nall - ndir, 0);
        size_t co = 0;
        for (size_t i = 0; i < nall; i++)
            if (!(m.faces[i].is_boundary && m.faces[i].dirichlet)) { compress_table[i] = co; expand_table[co] = i; co++; }
        ncells = m.cells.size();
        system_size = cell_basis_size(di.cd) * ncells + face_basis_size(di.fd) * (nall - ndir);
        RHS.assign(system_size, 0.0);
    }
    /* hho.hpp:344-406; also returns the face L2-projected Dirichlet data for reuse by tests */
    Vec assemble(const Mesh& m, const Cell& cl, const Mat& lhs, const Vec& rhs, const Fun& dirichlet_bf) {
        size_t cbs = cell_basis_size(di.cd), fbs = face_basis_size(di.fd);
        auto fids = face_ids(m, cl);
        std::vector<std::pair<size_t, bool>> asm_map;
        size_t cell_LHS_offset = offset(m, cl) * cbs;
        for (size_t i = 0; i < cbs; i++) asm_map.push_back({cell_LHS_offset + i, true});
        Vec dirichlet_data(cbs + 4 * fbs, 0.0);
        for (size_t fi = 0; fi < 4; fi++) {
            const Face& fc = m.faces[fids[fi]];
            size_t face_LHS_offset = cbs * m.cells.size() + compress_table.at(fids[fi]) * fbs;
            bool dirichlet = fc.is_boundary && fc.dirichlet;
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
        return dirichlet_data;
    }
    /* hho.hpp:408-449 */
    Vec take_local_data(const Mesh& m, const Cell& cl, const Vec& solution, const Fun& dirichlet_bf) const {
        size_t cbs = cell_basis_size(di.cd), fbs = face_basis_size(di.fd);
        size_t cell_SOL_offset = offset(m, cl) * cbs;
        auto fids = face_ids(m, cl);
        Vec ret(cbs + 4 * fbs, 0.0);
        for (size_t i = 0; i < cbs; i++) ret[i] = solution[cell_SOL_offset + i];
        for (size_t fi = 0; fi < 4; fi++) {
            const Face& fc = m.faces[fids[fi]];
            if (fc.is_boundary && fc.dirichlet) {
                Vec v = LLT(make_mass_matrix(m, fc, di.fd)).solve(make_rhs(m, fc, di.fd, dirichlet_bf));
                std::copy(v.begin(), v.end(), ret.begin() + cbs + fi * fbs);
            } else {
                size_t o = cbs * m.cells.size() + compress_table.at(fids[fi]) * fbs;
                for (size_t i = 0; i < fbs; i++) ret[cbs + fi * fbs + i] = solution[o + i];
            }
        }
        return ret;
    }
    void finalize() { LHS = csr_from_triplets(system_size, triplets); triplets.clear(); triplets.shrink_to_fit(); }
};

/* solver_cg.hpp:37-144 */
enum CgExit { CG_CONVERGED = 0, CG_DIVERGED = 1, CG_MAX_ITER_REACHED = 2 };
struct CgParams { double convergence_threshold = 1e-9, divergence_threshold = 100; size_t max_iter = 1000; bool apply_preconditioner = false; };
inline CgExit conjugated_gradient(const Csr& A, const Vec& b, Vec& x, const CgParams& p, size_t* iters_out = nullptr,
                                  double* relres_out = nullptr, std::vector<double>* history = nullptr) {
    size_t N = A.n, iter = 0; double nr, nr0, alpha, beta, rho;
    Vec d(N), r(N), y(N), iMr(N);
    x.assign(N, 0.0);
    Vec iA = A.diagonal();
    for (size_t i = 0; i < N; i++) iA[i] = 1. / iA[i];
    r = b;                                         /* r0 = r = b - A*x with x = 0 */
    for (size_t i = 0; i < N; i++) d[i] = p.apply_preconditioner ? iA[i] * r[i] : r[i];
    nr = nr0 = std::sqrt(dot(r, r));
    CgExit reason;
    while (1) {
        if (history) history->push_back(nr / nr0);
        A.mul(d, y);
        for (size_t i = 0; i < N; i++) iMr[i] = p.apply_preconditioner ? iA[i] * r[i] : r[i];
        rho = dot(r, iMr);
        alpha = rho / dot(d, y);
        for (size_t i = 0; i < N; i++) x[i] = x[i] + alpha * d[i];
        for (size_t i = 0; i < N; i++) r[i] = r[i] - alpha * y[i];
        nr = std::sqrt(dot(r, r));
        if (nr / nr0 < p.convergence_threshold) { reason = CG_CONVERGED; break; }
        if (iter > p.max_iter) { reason = CG_MAX_ITER_REACHED; break; }
        if (nr / nr0 > p.divergence_threshold) { reason = CG_DIVERGED; break; }
        for (size_t i = 0; i < N; i++) iMr[i] = p.apply_preconditioner ? iA[i] * r[i] : r[i];
        beta = dot(r, iMr) / rho;
        for (size_t i = 0; i < N; i++) d[i] = iMr[i] + beta * d[i];
        iter++;
    }
    if (history) history->push_back(nr / nr0);
    if (iters_out) *iters_out = iter;
    if (relres_out) *relres_out = nr / nr0;
    return reason;
}

} // namespace orc
