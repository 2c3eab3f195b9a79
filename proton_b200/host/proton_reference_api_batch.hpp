// proton_reference_api_batch.hpp — second half of proton_reference_api.hpp (included by it): the batched evaluation
// behind the per-cell calls, the per-cell free functions with the reference's signatures, the assembler and the solver.
#pragma once

namespace proton_b200 {

namespace detail {
// the drivers' host functors, recognised by value at three probe points: rhs_fun / sol_fun of convergence_test.cpp:100-138
// (sin sin, tanh), cuthho_square.cpp:847-866 (sin sin), obstacle.cpp:65-89
inline double known_rhs(int kind, double x, double y) {
    const double pi = 3.14159265358979323846;
    if (kind == PB_FUN_SINSIN) return 2.0 * pi * pi * std::sin(pi * x) * std::sin(pi * y);
    if (kind == PB_FUN_OBSTACLE) { const double r0 = 0.7, r = std::sqrt(x * x + y * y); return r > r0 ? -16 * r * r + 8 * r0 * r0 : -8.0 * (r0 * r0 * (r0 * r0 + 1)) + 8 * r0 * r0 * r * r; }
    const double th1 = std::tanh(30 - 60 * (x - y)), ch1 = std::cosh(30 - 60 * (x - y)), th2 = std::tanh(60 * y), ch2 = std::cosh(60 * y);
    return 14400 * th1 / (ch1 * ch1) + 7200 * th2 / (ch2 * ch2);
}
inline double known_sol(int kind, double x, double y) {
    const double pi = 3.14159265358979323846;
    if (kind == PB_FUN_SINSIN) return std::sin(pi * x) * std::sin(pi * y);
    if (kind == PB_FUN_OBSTACLE) { const double r0 = 0.7, s = x * x + y * y - r0 * r0, t = s > 0 ? s : 0; return t * t; }
    return std::tanh(60 * y) - std::tanh(60 * (x - y) - 30);
}
template <typename T, typename Function>
int recognise(const Function& f, bool is_rhs) {
    const double px[3] = {0.3, 0.61, 0.11}, py[3] = {0.7, 0.23, 0.45};
    for (int kind = 0; kind < 3; kind++) {
        bool ok = true;
        for (int q = 0; q < 3 && ok; q++) {
            const double want = is_rhs ? known_rhs(kind, px[q], py[q]) : known_sol(kind, px[q], py[q]);
            const double got = (double)f(point<T, 2>((T)px[q], (T)py[q]));
            ok = std::fabs(got - want) <= 1e-12 * (1.0 + std::fabs(want));
        }
        if (ok) return kind;
    }
    throw std::invalid_argument("proton_b200: this functor is not one of the drivers' functors that exist on the device "
                                "(sin(pi x)sin(pi y) Poisson data, the tanh problem, the obstacle data)");
}
}  // namespace detail

// the batch behind the per-cell views: one pb_assemble(keep_local = 2) for all cells; chunks of the per-cell arrays are
// copied to the host on demand (a serial loop over msh.cells walks them in order)
template <typename T>
class local_batch {
    const reference_mesh<T>& msh_;
    bool valid_ = false; int cd_ = -1, fd_ = -1, stab_ = -1, problem_ = -1, fun_ = PB_FUN_SINSIN;
    size_t C_ = 0, F_ = 0, R_ = 0, n_ = 0;
    static const size_t CHUNK = 2048;
    size_t c0_ = 0, cnt_ = 0; std::vector<T> lc_, data_, oper_, rhs_;
    size_t s0_ = 0, scnt_ = 0; std::vector<T> uloc_;
    unsigned long long runs_ = 0;
    void fetch(size_t cell) {
        if (cnt_ && cell >= c0_ && cell < c0_ + cnt_) return;
        c0_ = cell; cnt_ = std::min(CHUNK, msh_.num_cells() - cell);
        lc_.resize(cnt_ * n_ * n_); data_.resize(cnt_ * n_ * n_); oper_.resize(cnt_ * n_ * R_); rhs_.resize(cnt_ * C_);
        msh_.check(pb_get_local(msh_.ctx(), 0, (long long)c0_, (long long)cnt_, lc_.data()));
        msh_.check(pb_get_local(msh_.ctx(), 7, (long long)c0_, (long long)cnt_, data_.data()));
        msh_.check(pb_get_local(msh_.ctx(), 1, (long long)c0_, (long long)cnt_, oper_.data()));
        msh_.check(pb_get_local(msh_.ctx(), 2, (long long)c0_, (long long)cnt_, rhs_.data()));
    }
public:
    explicit local_batch(const reference_mesh<T>& msh) : msh_(msh) {}
    // (re)run the batch when the request differs from what the device holds; stab / fun < 0 = keep the current choice
    void ensure(const hho_degree_info& hdi, int problem, int stab, int fun) {
        const int cd = (int)hdi.cell_degree(), fd = (int)hdi.face_degree();
        if (stab < 0) stab = (valid_ && problem == problem_) ? stab_ : (problem == PB_PROBLEM_FICTDOM ? PB_STAB_CUT : PB_STAB_FANCY);
        if (fun < 0) fun = fun_;
        if (valid_ && cd == cd_ && fd == fd_ && problem == problem_ && stab == stab_ && fun == fun_) return;
        pb_assembly_params p{};
        p.cell_degree = cd; p.face_degree = fd; p.stab_kind = stab; p.problem = problem; p.fun = fun; p.rhs_degree_increase = 0; p.eta = 5.0; p.keep_local = 2;
        msh_.check(pb_assemble(msh_.ctx(), &p));
        cd_ = cd; fd_ = fd; problem_ = problem; stab_ = stab; fun_ = fun; valid_ = true; cnt_ = 0; scnt_ = 0; runs_++;
        C_ = (cd + 1) * (cd + 2) / 2; F_ = fd + 1; R_ = (fd + 2) * (fd + 3) / 2; n_ = C_ + 4 * F_;
    }
    void invalidate() { valid_ = false; cnt_ = 0; scnt_ = 0; }
    bool valid() const { return valid_; }
    int problem() const { return problem_; } int stab() const { return stab_; } int fun() const { return fun_; }
    size_t n() const { return n_; } size_t cbs() const { return C_; } size_t fbs() const { return F_; } size_t rbs() const { return R_; }
    unsigned long long runs() const { return runs_; }
    const T* lc(size_t cell) { fetch(cell); return lc_.data() + (cell - c0_) * n_ * n_; }
    const T* data(size_t cell) { fetch(cell); return data_.data() + (cell - c0_) * n_ * n_; }
    const T* oper(size_t cell) { fetch(cell); return oper_.data() + (cell - c0_) * n_ * R_; }
    const T* rhs(size_t cell) { fetch(cell); return rhs_.data() + (cell - c0_) * C_; }
    // take_local_data of one cell from the device's face solution
    const T* local_solution(size_t cell) {
        if (!(scnt_ && cell >= s0_ && cell < s0_ + scnt_)) {
            s0_ = cell; scnt_ = std::min(CHUNK, msh_.num_cells() - cell); uloc_.resize(scnt_ * n_);
            msh_.check(pb_get_local_solution(msh_.ctx(), (long long)s0_, (long long)scnt_, uloc_.data()));
        }
        return uloc_.data() + (cell - s0_) * n_;
    }
    void solution_changed() { scnt_ = 0; }
};

namespace detail {
template <typename T> Matrix<T> copy_matrix(const T* src, size_t ld, size_t rows, size_t cols) {
    Matrix<T> m(rows, cols);
    for (size_t j = 0; j < cols; j++) for (size_t i = 0; i < rows; i++) m(i, j) = src[i + j * ld];
    return m;
}
template <typename T> std::pair<Matrix<T>, Matrix<T>> laplacian_view(local_batch<T>& b, size_t cell, bool cut_cell) {
    const size_t n = b.n(), R = b.rbs();
    // uncut cells: oper has R-1 rows (the constant is removed, hho.hpp:61-62), stored with leading dimension R; cut cells keep all R rows
    return std::make_pair(copy_matrix(b.oper(cell), R, cut_cell ? R : R - 1, n), copy_matrix(b.data(cell), n, n, n));
}
template <typename T> Matrix<T> stabilization_view(local_batch<T>& b, size_t cell) {
    const size_t n = b.n();
    Matrix<T> s = copy_matrix(b.lc(cell), n, n, n); s -= copy_matrix(b.data(cell), n, n, n);       // lc = gr.second + stab on the device
    return s;
}
}  // namespace detail

// ---------------------------------------------------------------------------------------------- per-cell calls, uncut (hho.hpp)
template <typename T>
std::pair<Matrix<T>, Matrix<T>> make_hho_laplacian(const reference_mesh<T>& msh, const cell_handle& cl, const hho_degree_info& di) {
    msh.batch().ensure(di, PB_PROBLEM_HHO, -1, -1);
    return detail::laplacian_view(msh.batch(), cl.id, false);
}
template <typename T>
Matrix<T> make_hho_naive_stabilization(const reference_mesh<T>& msh, const cell_handle& cl, const hho_degree_info& di) {
    msh.batch().ensure(di, PB_PROBLEM_HHO, PB_STAB_NAIVE, -1);
    return detail::stabilization_view(msh.batch(), cl.id);
}
template <typename T, int R1, int C1>
Matrix<T> make_hho_fancy_stabilization(const reference_mesh<T>& msh, const cell_handle& cl, const Matrix<T, R1, C1>& reconstruction, const hho_degree_info& di) {
    msh.batch().ensure(di, PB_PROBLEM_HHO, PB_STAB_FANCY, -1);
    if (reconstruction.rows() + 1 != msh.batch().rbs() || reconstruction.cols() != msh.batch().n())
        throw std::invalid_argument("make_hho_fancy_stabilization: the reconstruction operator is not gr.first of this cell");
    return detail::stabilization_view(msh.batch(), cl.id);
}
template <typename T, typename Function>
Matrix<T, Dynamic, 1> make_rhs(const reference_mesh<T>& msh, const cell_handle& cl, size_t degree, const Function& f) {
    local_batch<T>& b = msh.batch();
    if (!b.valid()) throw std::logic_error("make_rhs: call make_hho_laplacian on this mesh first (the degree pair comes from there)");
    if ((degree + 1) * (degree + 2) / 2 != b.cbs()) throw std::invalid_argument("make_rhs: degree is not the cell degree of the batch");
    b.ensure(hho_degree_info(degree, b.fbs() - 1), b.problem(), -1, detail::recognise<T>(f, true));
    Matrix<T, Dynamic, 1> r(b.cbs());
    const T* src = b.rhs(cl.id); for (size_t i = 0; i < b.cbs(); i++) r(i) = src[i];
    return r;
}

// ---------------------------------------------------------------------------------------------- cutHHO (cuthho_square.cpp)
template <typename T, typename LevelSet> void detect_node_position(reference_mesh<T>& msh, const LevelSet& ls) {
    msh.check(pb_cut_levelset(msh.ctx(), ls.kind(), ls.radius, ls.alpha, ls.beta, ls.cut_y)); msh.batch().invalidate();
}
template <typename T, typename LevelSet> void detect_cut_faces(reference_mesh<T>&, const LevelSet&) {}      // part of pb_cut_preprocess
template <typename T, typename LevelSet> void move_nodes(reference_mesh<T>&, const LevelSet&) {}
template <typename T, typename LevelSet> void detect_cut_cells(reference_mesh<T>&, const LevelSet&) {}
// the device runs detect_node_position .. refine_interface as ONE call (pb_cut_preprocess moves the mesh points, so it cannot be
// split); it runs here, where the reference's sequence ends (cuthho_square.cpp:2052), and fills location(msh, cl)
template <typename T, typename LevelSet> void refine_interface(reference_mesh<T>& msh, const LevelSet&, size_t refine_steps) {
    msh.check(pb_cut_preprocess(msh.ctx(), (int)refine_steps));
    std::vector<int> loc(msh.num_cells());
    msh.check(pb_cut_download(msh.ctx(), nullptr, nullptr, nullptr, loc.data(), nullptr, nullptr, nullptr));
    for (size_t i = 0; i < loc.size(); i++)
        msh.cells[i].user_data.location = loc[i] == PB_IN_NEGATIVE_SIDE ? element_location::IN_NEGATIVE_SIDE : loc[i] == PB_IN_POSITIVE_SIDE ? element_location::IN_POSITIVE_SIDE
                                          : loc[i] == PB_ON_INTERFACE ? element_location::ON_INTERFACE : element_location::UNDEF;
    msh.batch().invalidate();
}
namespace detail {
inline void need_negative_side(element_location where) {
    if (where != element_location::IN_NEGATIVE_SIDE) throw std::invalid_argument("proton_b200: the fictitious-domain operators are built for where = IN_NEGATIVE_SIDE (run_cuthho_fictdom, cuthho_square.cpp:873)");
}
}
template <typename T, typename LevelSet>
std::pair<Matrix<T>, Matrix<T>> make_hho_laplacian(const reference_mesh<T>& msh, const cell_handle& cl, const LevelSet&, const hho_degree_info& di, element_location where) {
    detail::need_negative_side(where);
    msh.batch().ensure(di, PB_PROBLEM_FICTDOM, PB_STAB_CUT, -1);
    return detail::laplacian_view(msh.batch(), cl.id, cl.user_data.location == element_location::ON_INTERFACE);
}
template <typename T>
Matrix<T> make_hho_cut_stabilization(const reference_mesh<T>& msh, const cell_handle& cl, const hho_degree_info& di, element_location where) {
    detail::need_negative_side(where);
    msh.batch().ensure(di, PB_PROBLEM_FICTDOM, PB_STAB_CUT, -1);
    return detail::stabilization_view(msh.batch(), cl.id);
}
template <typename T, typename F1, typename LevelSet, typename F2>
Matrix<T, Dynamic, 1> make_rhs(const reference_mesh<T>& msh, const cell_handle& cl, size_t degree, const F1& f, element_location where, const LevelSet&, const F2& bcs) {
    detail::need_negative_side(where);
    local_batch<T>& b = msh.batch();
    if (!b.valid()) throw std::logic_error("make_rhs: call make_hho_laplacian on this mesh first");
    const int kind = detail::recognise<T>(f, true);
    if (detail::recognise<T>(bcs, false) != kind) throw std::invalid_argument("make_rhs: source and Nitsche datum belong to different problems");
    b.ensure(hho_degree_info(degree, b.fbs() - 1), PB_PROBLEM_FICTDOM, PB_STAB_CUT, kind);
    Matrix<T, Dynamic, 1> r(b.cbs());
    const T* src = b.rhs(cl.id); for (size_t i = 0; i < b.cbs(); i++) r(i) = src[i];
    return r;
}

}  // namespace proton_b200
#include "proton_reference_api_solve.hpp"
