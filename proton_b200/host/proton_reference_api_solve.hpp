// proton_reference_api_solve.hpp — third part of proton_reference_api.hpp: assembler (hho.hpp:252-463) and
// conjugated_gradient (solver_cg.hpp:63-144) with the reference's call shapes.
#pragma once

namespace proton_b200 {

// assembler.LHS: the condensed face system (CSR copy for inspection; the solve runs on the device-resident system)
template <typename T>
struct condensed_matrix : sparse_matrix<T> {
    const reference_mesh<T>* owner = nullptr;
    size_t rows_ = 0;
    size_t rows() const { return rows_; } size_t cols() const { return rows_; } size_t nonZeros() const { return (size_t)this->nnz; }
};

template <typename T>
class reference_assembler {
    const reference_mesh<T>& msh_;
    hho_degree_info di_;
    std::vector<unsigned char> seen_; size_t nseen_ = 0; bool finalized_ = false;
    unsigned long long batch_run_ = 0;
public:
    condensed_matrix<T> LHS;
    Matrix<T, Dynamic, 1> RHS;
    reference_assembler(const reference_mesh<T>& msh, hho_degree_info hdi) : msh_(msh), di_(hdi), seen_(msh.num_cells(), 0) { LHS.owner = &msh; }

    // assemble(msh, cl, lhs, rhs, dirichlet_bf) (hho.hpp:344-406): the device has already condensed and scattered this cell's
    // operators with the batch; the call verifies that what the driver hands over is that cell's batched contribution
    template <int R1, int C1, int R2, int C2, typename Function>
    void assemble(const reference_mesh<T>& msh, const cell_handle& cl, const Matrix<T, R1, C1>& lhs, const Matrix<T, R2, C2>& rhs, const Function& dirichlet_bf) {
        if (&msh != &msh_) throw std::invalid_argument("assembler::assemble: another mesh than the assembler's");
        local_batch<T>& b = msh.batch();
        if (!b.valid()) throw std::logic_error("assembler::assemble: no local operators were requested on this mesh");
        if (nseen_ == 0) batch_run_ = b.runs();
        if (b.runs() != batch_run_) throw std::logic_error("assembler::assemble: the operators changed (stabilisation / data) after cells had been assembled");
        if (detail::recognise<T>(dirichlet_bf, false) != b.fun()) throw std::invalid_argument("assembler::assemble: the Dirichlet datum does not belong to the source term of make_rhs");
        const size_t n = b.n(), C = b.cbs();
        if (lhs.rows() != n || lhs.cols() != n || rhs.size() < C) throw std::invalid_argument("assembler::assemble: local matrix / rhs of the wrong size");
        const T* lc = b.lc(cl.id); const T* f = b.rhs(cl.id);
        T scale = 0, diff = 0;
        for (size_t k = 0; k < n * n; k++) { scale = std::max(scale, std::fabs(lc[k])); diff = std::max(diff, std::fabs(lc[k] - lhs.data()[k])); }
        T fs = 0, fd = 0;
        for (size_t k = 0; k < C; k++) { fs = std::max(fs, std::fabs(f[k])); fd = std::max(fd, std::fabs(f[k] - rhs.data()[k])); }
        if (diff > 1e-10 * scale || fd > 1e-10 * (fs + scale * 1e-6))
            throw std::invalid_argument("assembler::assemble: the local contribution differs from the one computed on the device (modified local matrices are not assembled)");
        if (!seen_[cl.id]) { seen_[cl.id] = 1; nseen_++; }
    }
    void finalize() {       // hho.hpp:451-455
        if (nseen_ != msh_.num_cells()) throw std::logic_error("assembler::finalize: not every cell was assembled");
        long long n = 0, nnz = 0;
        msh_.check(pb_system_sizes(msh_.ctx(), &n, &nnz, nullptr, nullptr));
        LHS.rows_ = (size_t)n; LHS.sparse_matrix<T>::rows = n; LHS.nnz = nnz; LHS.rowptr.resize(n + 1); LHS.col.resize(nnz); LHS.val.resize(nnz); RHS.resize((size_t)n);
        msh_.check(pb_get_csr(msh_.ctx(), LHS.rowptr.data(), LHS.col.data(), LHS.val.data(), RHS.data()));
        finalized_ = true;
    }
    // take_local_data(msh, cl, solution, dirichlet_bf) (hho.hpp:408-449): cell dofs, then the face dofs of faces(msh, cl) with the
    // Dirichlet faces projected; `solution` must be the vector conjugated_gradient returned (the device keeps it)
    template <int R1, int C1, typename Function>
    Matrix<T, Dynamic, 1> take_local_data(const reference_mesh<T>& msh, const cell_handle& cl, const Matrix<T, R1, C1>& solution, const Function&) {
        local_batch<T>& b = msh.batch();
        if (solution.size() != msh.num_cells() * b.cbs() + LHS.rows()) throw std::invalid_argument("take_local_data: not the solution vector of this system");
        Matrix<T, Dynamic, 1> r(b.n());
        const T* u = b.local_solution(cl.id);
        for (size_t i = 0; i < b.n(); i++) r(i) = u[i];
        return r;
    }
    std::vector<T> error_norms() const { std::vector<T> out(3); msh_.check(pb_error_norms(msh_.ctx(), out.data())); return out; }
    const reference_mesh<T>& mesh() const { return msh_; }
};
template <typename T> reference_assembler<T> make_assembler(const reference_mesh<T>& msh, hho_degree_info hdi) { return reference_assembler<T>(msh, hdi); }
template <typename T> reference_assembler<T> make_assembler(const quad_mesh<T>& msh, hho_degree_info hdi) { return reference_assembler<T>(msh, hdi); }
template <typename T> reference_assembler<T> make_assembler(const cuthho_poly_mesh<T>& msh, hho_degree_info hdi) { return reference_assembler<T>(msh, hdi); }

// conjugated_gradient(A, b, x, parms) (solver_cg.hpp:63-68) with A = assembler.LHS, b = assembler.RHS.  x receives the
// reference's FULL solution vector: cell dofs c*cbs+i (back-substituted), then the face dofs, hho.hpp:362-379.
// cg_params::preconditioner_kind (extension): 0 = the reference's choice (diagonal when apply_preconditioner), 2 = block-Jacobi, 3 = multigrid.
template <typename T, int R1, int C1, int R2, int C2>
cg_exit_reason conjugated_gradient(const condensed_matrix<T>& A, const Matrix<T, R1, C1>& b, Matrix<T, R2, C2>& x, const cg_params<T>& parms) {
    if (!A.owner || b.size() != A.rows()) throw std::invalid_argument("conjugated_gradient: A and b must be assembler.LHS and assembler.RHS after finalize()");
    const reference_mesh<T>& msh = *A.owner;
    pb_cg_params p{parms.convergence_threshold, parms.divergence_threshold, (long long)parms.max_iter,
                   parms.apply_preconditioner ? (parms.block_preconditioner ? 2 : 1) : 0};
    pb_cg_result r{};
    const bool hist = !parms.histfile.empty();
    if (hist) msh.check(pb_pcg_history(msh.ctx(), (long long)parms.max_iter + 3));
    msh.check(pb_pcg(msh.ctx(), &p, &r));
    if (hist) {
        long long nh = 0; msh.check(pb_pcg_get_history(msh.ctx(), nullptr, 0, &nh));
        std::vector<T> h((size_t)nh);
        if (nh) msh.check(pb_pcg_get_history(msh.ctx(), h.data(), nh, &nh));
        msh.check(pb_pcg_history(msh.ctx(), 0));
        std::ofstream ofs(parms.histfile);
        for (T v : h) ofs << v << std::endl;
    }
    if (parms.verbose) std::cout << " -> Iteration " << r.iterations << ", rr = " << r.relative_residual << std::endl;
    msh.batch().solution_changed();
    x.resize(msh.num_cells() * msh.batch().cbs() + A.rows());
    msh.check(pb_recover(msh.ctx(), x.data()));
    return r.exit_reason == PB_CG_CONVERGED ? cg_exit_reason::CONVERGED : r.exit_reason == PB_CG_DIVERGED ? cg_exit_reason::DIVERGED : cg_exit_reason::MAX_ITER_REACHED;
}

}  // namespace proton_b200
