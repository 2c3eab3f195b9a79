// proton_host.hpp — C++ host mirror of the reference's driver-facing names for the HHO hot path, forwarding to the
// C ABI of include/proton_b200.h.  Header-only, C++14, no dependencies beyond the C library.
//
// Names, argument meaning and error behaviour follow the reference (paths relative to the reference root):
//   mesh_init_params<T>        src/core/core_bits/basic_mesh.hpp:178-197
//   hho_degree_info            src/core/core_bits/utils.hpp:62-111 (invalid pairs revert to equal order)
//   cg_params<T>, cg_exit_reason, conjugated_gradient   src/core/core_bits/solver_cg.hpp:37-144
//   make_assembler(msh, hdi) -> .assemble() .finalize() .LHS .RHS .take_local_data()   src/methods/hho_bits/hho.hpp:252-463
// The reference calls make_hho_laplacian / make_hho_*_stabilization / make_rhs / assembler.assemble once per cell
// from a serial loop; here the whole loop is ONE batched call (assembler::assemble_all), and per-cell results are
// views into the batch (assembler::local_contrib).  Errors are exceptions, like the reference (std::logic_error /
// std::invalid_argument); there is no CPU fallback: without a CUDA device construction throws.
#pragma once
#include <array>
#include <cmath>
#include <cstddef>
#include <cstdio>
#include <fstream>
#include <iostream>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/proton_b200.h"

namespace proton_b200 {

template <typename T>
struct mesh_init_params {
    T min_x, max_x, min_y, max_y;
    size_t Nx, Ny;
    mesh_init_params() : min_x(0), max_x(1), min_y(0), max_y(1), Nx(4), Ny(4) {}
    T hx() const { return (max_x - min_x) / Nx; }
    T hy() const { return (max_y - min_y) / Ny; }
};

class hho_degree_info {
    size_t cell_deg, face_deg, reconstruction_deg;
public:
    hho_degree_info() : cell_deg(1), face_deg(1), reconstruction_deg(2) {}
    explicit hho_degree_info(size_t degree) : cell_deg(degree), face_deg(degree), reconstruction_deg(degree + 1) {}
    hho_degree_info(size_t cd, size_t fd) {
        bool c1 = fd > 0 && (cd == fd - 1 || cd == fd || cd == fd + 1);
        bool c2 = fd == 0 && (cd == fd || cd == fd + 1);
        if (c1 || c2) { cell_deg = cd; face_deg = fd; reconstruction_deg = fd + 1; }
        else { cell_deg = fd; face_deg = fd; reconstruction_deg = fd + 1; }     // "Reverting to equal-order"
    }
    size_t cell_degree() const { return cell_deg; }
    size_t face_degree() const { return face_deg; }
    size_t reconstruction_degree() const { return reconstruction_deg; }
};

enum class cg_exit_reason { CONVERGED, DIVERGED, MAX_ITER_REACHED };

template <typename T>
struct cg_params {
    T convergence_threshold, divergence_threshold;
    size_t max_iter;
    bool verbose, apply_preconditioner, save_reshist;
    std::string histfile;               // solver_cg.hpp:51: one line nr/nr0 per iteration + the closing line
    bool block_preconditioner;          // extension (no reference counterpart): fbs x fbs block-Jacobi instead of the diagonal
    cg_params() : convergence_threshold(1e-9), divergence_threshold(100), max_iter(1000), verbose(false),
                  apply_preconditioner(false), save_reshist(false), block_preconditioner(false) {}
};

// quad_mesh<T>(mip) / cuthho_poly_mesh<T>(mip): the Cartesian generator; cells, faces and the boundary flags are
// closed forms on the device, the points live in HBM.
template <typename T>
class cartesian_mesh {
    pb_ctx* ctx_ = nullptr;
    mesh_init_params<T> mip_;
    static void ck(pb_ctx* c, pb_status st) { if (st != PB_OK) throw std::logic_error(std::string("proton_b200: ") + pb_last_error(c)); }
public:
    enum kind { quad = PB_QUAD_TENSOR_GAUSS, cuthho_poly = PB_QUAD_POLY_DUNAVANT };
    cartesian_mesh(const mesh_init_params<T>& mip, kind k = quad, int device = 0, long long row_begin = 0, long long row_end = -1) : mip_(mip) {
        pb_status st = pb_ctx_create(device, &ctx_);
        if (st != PB_OK) throw std::logic_error(std::string("proton_b200: ") + pb_last_error(nullptr));
        ck(ctx_, pb_mesh_cartesian(ctx_, (long long)mip.Nx, (long long)mip.Ny, mip.min_x, mip.max_x, mip.min_y, mip.max_y, (int)k,
                                   row_begin, row_end < 0 ? (long long)mip.Ny : row_end));
    }
    ~cartesian_mesh() { pb_ctx_destroy(ctx_); }
    cartesian_mesh(const cartesian_mesh&) = delete;
    cartesian_mesh& operator=(const cartesian_mesh&) = delete;
    pb_ctx* ctx() const { return ctx_; }
    size_t num_cells() const { return mip_.Nx * mip_.Ny; }
    size_t num_faces() const { return 2 * mip_.Nx * mip_.Ny + mip_.Nx + mip_.Ny; }
    size_t num_points() const { return (mip_.Nx + 1) * (mip_.Ny + 1); }
    std::vector<T> points() const { std::vector<T> xy(2 * num_points()); ck(ctx_, pb_mesh_get_points(ctx_, xy.data())); return xy; }
    void check(pb_status st) const { ck(ctx_, st); }
    const mesh_init_params<T>& params() const { return mip_; }
    // detect_node_position, detect_cut_faces, move_nodes, detect_cut_faces, detect_cut_cells, refine_interface
    // (apps/cuthho/cuthho_square.cpp:2030-2052) for one of the reference's level sets, on the device
    template <typename LevelSet>
    void cut_preprocess(const LevelSet& ls, int refine_steps) {
        ck(ctx_, pb_cut_levelset(ctx_, ls.kind(), ls.radius, ls.alpha, ls.beta, ls.cut_y));
        ck(ctx_, pb_cut_preprocess(ctx_, refine_steps));
    }
};

// point<T,2> (src/core/core_bits/point.hpp): what the drivers' functors receive
template <typename T, size_t DIM> struct point;
template <typename T> struct point<T, 2> {
    T c[2];
    point() { c[0] = c[1] = T(0); }
    point(T x, T y) { c[0] = x; c[1] = y; }
    T x() const { return c[0]; } T y() const { return c[1]; }
    T& x() { return c[0]; } T& y() { return c[1]; }
};

// level sets of apps/cuthho/cuthho_square.cpp:56-124: host-evaluable like the reference's functors (operator(), gradient,
// normal on a point); the device uses the same formulas (pb_cut_levelset)
template <typename T>
struct circle_level_set {
    T radius, alpha, beta, cut_y = 0;
    circle_level_set(T r, T a, T b) : radius(r), alpha(a), beta(b) {}
    int kind() const { return 0; }
    T operator()(T x, T y) const { return (x - alpha) * (x - alpha) + (y - beta) * (y - beta) - radius * radius; }
    T operator()(const point<T, 2>& pt) const { return (*this)(pt.x(), pt.y()); }                      // :66-72
    std::array<T, 2> gradient(const point<T, 2>& pt) const { return {{2 * pt.x() - 2 * alpha, 2 * pt.y() - 2 * beta}}; }      // :74-81
    std::array<T, 2> normal(const point<T, 2>& pt) const {                                             // :83-90
        auto g = gradient(pt); const T n = std::sqrt(g[0] * g[0] + g[1] * g[1]); return {{g[0] / n, g[1] / n}};
    }
};
template <typename T>
struct line_level_set {
    T radius = 0, alpha = 0, beta = 0, cut_y;
    explicit line_level_set(T cy) : cut_y(cy) {}
    int kind() const { return 1; }
    T operator()(T, T y) const { return y - cut_y; }
    T operator()(const point<T, 2>& pt) const { return pt.y() - cut_y; }                               // :102-108
    std::array<T, 2> gradient(const point<T, 2>&) const { return {{T(0), T(1)}}; }                     // :110-117
    std::array<T, 2> normal(const point<T, 2>& pt) const { return gradient(pt); }
};

// CSR view of assembler.LHS (the condensed face system; the reference's full system has no counterpart here)
template <typename T>
struct sparse_matrix { long long rows = 0, nnz = 0; std::vector<long long> rowptr; std::vector<int> col; std::vector<T> val; };

template <typename T>
class assembler {
    const cartesian_mesh<T>& msh_;
    hho_degree_info di_;
    bool assembled_ = false; int problem_ = PB_PROBLEM_HHO;
public:
    sparse_matrix<T> LHS;
    std::vector<T> RHS;
    assembler(const cartesian_mesh<T>& msh, hho_degree_info hdi) : msh_(msh), di_(hdi) {}

    // the drivers' per-cell loop + assemble(): all cells in one batched launch sequence (K1/K2 + K3)
    void assemble_all(int stab_kind = PB_STAB_FANCY, int problem = PB_PROBLEM_HHO, int fun = PB_FUN_SINSIN, int rhs_degree_increase = 0,
                      double eta = 5.0, bool keep_local = false) {
        pb_assembly_params p{};
        p.cell_degree = (int)di_.cell_degree(); p.face_degree = (int)di_.face_degree(); p.stab_kind = stab_kind; p.problem = problem;
        p.fun = fun; p.rhs_degree_increase = rhs_degree_increase; p.eta = eta; p.keep_local = keep_local ? 1 : 0;
        msh_.check(pb_assemble(msh_.ctx(), &p));
        assembled_ = true; problem_ = problem;
    }
    // finalize(): hho.hpp:451-455 — here only the copy of the device CSR to the public members (optional: the solve
    // runs on the device-resident system)
    void finalize() {
        if (!assembled_) throw std::logic_error("assembler::finalize before assemble");
        long long n = 0, nnz = 0;
        msh_.check(pb_system_sizes(msh_.ctx(), &n, &nnz, nullptr, nullptr));
        LHS.rows = n; LHS.nnz = nnz; LHS.rowptr.resize(n + 1); LHS.col.resize(nnz); LHS.val.resize(nnz); RHS.resize(n);
        msh_.check(pb_get_csr(msh_.ctx(), LHS.rowptr.data(), LHS.col.data(), LHS.val.data(), RHS.data()));
    }
    // lc = gr.second + stab of one cell (column-major n x n), needs assemble_all(..., keep_local = true)
    std::vector<T> local_contrib(size_t cell_offset, size_t n_local) const {
        std::vector<T> out(n_local * n_local);
        msh_.check(pb_get_local(msh_.ctx(), 0, (long long)cell_offset, 1, out.data()));
        return out;
    }
    // take_local_data for all cells at once: the reference's full solution vector (cell dofs, then face dofs)
    std::vector<T> expand_solution(size_t cbs) const {
        long long n = 0; msh_.check(pb_system_sizes(msh_.ctx(), &n, nullptr, nullptr, nullptr));
        size_t cell_sets = msh_.num_cells();
        if (problem_ == PB_PROBLEM_INTERFACE) {          // cut cells carry two sets of cell unknowns (cell_table, cuthho_square.cpp:1143-1151)
            long long ncut = 0; msh_.check(pb_cut_download(msh_.ctx(), nullptr, nullptr, nullptr, nullptr, &ncut, nullptr, nullptr));
            cell_sets += (size_t)ncut;
        }
        std::vector<T> full(cell_sets * cbs + (size_t)n);
        msh_.check(pb_recover(msh_.ctx(), full.data()));
        return full;
    }
    const cartesian_mesh<T>& mesh() const { return msh_; }
    // the drivers' post-processing sums: {L2^2, mass-norm^2, energy^2} (convergence_test.cpp:254-306) or {H1^2, 0, 0}
    // for the fictitious-domain problem (cuthho_square.cpp:1031-1050)
    // rows "x y value" of the driver's solution_N_<N>_k_<k>.dat (convergence_test.cpp:250-252, 280-286)
    void write_solution(const std::string& filename) const {
        long long n = 0; msh_.check(pb_sample_solution(msh_.ctx(), nullptr, 0, &n));
        std::vector<T> xyv((size_t)3 * n);
        msh_.check(pb_sample_solution(msh_.ctx(), xyv.data(), n, &n));
        std::ofstream ofs(filename);
        for (long long i = 0; i < n; i++) ofs << xyv[3 * i] << " " << xyv[3 * i + 1] << " " << xyv[3 * i + 2] << "\n";
    }
    std::vector<T> error_norms() const { std::vector<T> out(3); msh_.check(pb_error_norms(msh_.ctx(), out.data())); return out; }
    size_t system_size() const { long long n = 0; msh_.check(pb_system_sizes(msh_.ctx(), &n, nullptr, nullptr, nullptr)); return (size_t)n; }
};

template <typename T>
assembler<T> make_assembler(const cartesian_mesh<T>& msh, hho_degree_info hdi) { return assembler<T>(msh, hdi); }

// conjugated_gradient(A, b, x, parms): A and b are the device-resident system of `assm`; x receives the face solution
template <typename T>
cg_exit_reason conjugated_gradient(const assembler<T>& assm, std::vector<T>& x, const cg_params<T>& parms, size_t* iterations = nullptr) {
    pb_cg_params p{parms.convergence_threshold, parms.divergence_threshold, (long long)parms.max_iter,
                   parms.apply_preconditioner ? (parms.block_preconditioner ? 2 : 1) : 0};
    pb_cg_result r{};
    const bool hist = !parms.histfile.empty();
    if (hist) assm.mesh().check(pb_pcg_history(assm.mesh().ctx(), (long long)parms.max_iter + 3));
    assm.mesh().check(pb_pcg(assm.mesh().ctx(), &p, &r));
    if (hist) {         // the same lines, in the same default stream format, as solver_cg.hpp:101-102, 133-137
        long long nh = 0; assm.mesh().check(pb_pcg_get_history(assm.mesh().ctx(), nullptr, 0, &nh));
        std::vector<T> h((size_t)nh);
        if (nh) assm.mesh().check(pb_pcg_get_history(assm.mesh().ctx(), h.data(), nh, &nh));
        assm.mesh().check(pb_pcg_history(assm.mesh().ctx(), 0));
        std::ofstream ofs(parms.histfile);
        for (T v : h) ofs << v << std::endl;
    }
    if (parms.verbose) std::cout << " -> Iteration " << r.iterations << ", rr = " << r.relative_residual << std::endl;
    long long n = 0; assm.mesh().check(pb_system_sizes(assm.mesh().ctx(), &n, nullptr, nullptr, nullptr));
    x.resize((size_t)n);
    assm.mesh().check(pb_get_face_solution(assm.mesh().ctx(), x.data()));
    if (iterations) *iterations = (size_t)r.iterations;
    return r.exit_reason == PB_CG_CONVERGED ? cg_exit_reason::CONVERGED : r.exit_reason == PB_CG_DIVERGED ? cg_exit_reason::DIVERGED : cg_exit_reason::MAX_ITER_REACHED;
}

// run_hho_obstacle (apps/obstacle/obstacle.cpp:47-218): the active-set loop; returns the driver's error
template <typename T>
T run_hho_obstacle(const cartesian_mesh<T>& msh, size_t degree, size_t* outer_iterations = nullptr, bool verbose = true) {
    msh.check(pb_obstacle_begin(msh.ctx(), (int)degree, 1));
    pb_cg_params cg{1e-13, 1e300, 1000000, 1};
    size_t iter = 0;
    while (iter < 50) {
        double delta = 0; long long na = 0, cgit = 0;
        msh.check(pb_obstacle_iterate(msh.ctx(), &cg, &delta, &na, &cgit));
        if (verbose) printf("Iteration %zu: active cells %lld, ||alpha_prev - alpha|| = %.6e, CG iterations %lld\n", iter + 1, na, delta, cgit);
        if (delta < 1e-7) break;
        iter++;
    }
    double e2 = 0; msh.check(pb_obstacle_error(msh.ctx(), &e2));
    if (outer_iterations) *outer_iterations = iter;
    return (T)std::sqrt(e2);
}

}  // namespace proton_b200
