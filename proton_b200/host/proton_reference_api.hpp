// proton_reference_api.hpp — the reference's per-cell driver API on top of the batched C ABI, so that a ProtoN driver loop
// compiles against this header with only the include and the namespace changed (SURVEY 8b "signatures to keep").
//
//   quad_mesh<T>(mip), cuthho_poly_mesh<T>(mip)       src/core/core_bits/basic_mesh.hpp:178-197, 479-486; cuthho_bits/cuthho_mesh.hpp:93-100
//   for (auto& cl : msh.cells) / for (auto& fc : msh.faces), fc.is_boundary, fc.bndtype, location(msh, cl)
//   make_hho_laplacian(msh, cl, hdi) -> pair(oper, data)                 src/methods/hho_bits/hho.hpp:32-96
//   make_hho_naive_stabilization / make_hho_fancy_stabilization          hho.hpp:99-148, 155-237
//   make_rhs(msh, cl, degree, f)                                         src/core/core_bits/utils.hpp:153-174
//   cut overloads (msh, cl, level_set, hdi, where) / make_hho_cut_stabilization / make_rhs(.., where, level_set, bcs)
//                                                                        apps/cuthho/cuthho_square.cpp:308-388, 566-666
//   detect_node_position, detect_cut_faces, move_nodes, detect_cut_cells, refine_interface   cuthho_square.cpp:2030-2052
//   make_assembler(msh, hdi): assemble(msh, cl, lc, f, bf), finalize(), LHS, RHS, take_local_data(msh, cl, sol, bf)   hho.hpp:252-463
//   conjugated_gradient(A, b, x, parms)                                  src/core/core_bits/solver_cg.hpp:63-144
//   circle_level_set / line_level_set with operator()(pt), gradient(pt), normal(pt)          cuthho_square.cpp:56-124
//
// How a serial per-cell loop meets a batched device path: the FIRST per-cell call on a mesh runs the whole batch (all cells,
// pb_assemble with keep_local = 2) and every per-cell call returns a view (a copy of that cell's block) of the batched
// result; assembler.assemble() checks that the local matrix it is handed IS the batched one (a driver that modifies lc
// before assembling gets std::invalid_argument: the device assembles what the device computed) and finalize() publishes
// the condensed system.  Host functors (rhs_fun, sol_fun, the level set) cannot run on the device: they are recognised
// by sampling them against the device functors of the drivers (convergence_test.cpp:100-138, cuthho_square.cpp:847-866,
// obstacle.cpp:65-89); an unknown functor throws std::invalid_argument.  There is no host arithmetic for the operators.
// What differs from the reference by construction: LHS / RHS are the CONDENSED face system (the reference has no static
// condensation); conjugated_gradient returns the reference's FULL vector (cell dofs, then face dofs) after
// back-substitution, so post-processing code that slices `sol` like the reference keeps working.
#pragma once
#include <array>
#include <utility>

#include "proton_host.hpp"

namespace proton_b200 {

// ---------------------------------------------------------------------------------------------- dense value types
const int Dynamic = -1;
// the little of Eigen::Matrix<T, Rows, Cols> the driver loops use: column-major storage, value semantics
template <typename T, int Rows = Dynamic, int Cols = Dynamic>
class Matrix {
    size_t r_ = Rows > 0 ? Rows : 0, c_ = Cols > 0 ? Cols : 0;
    std::vector<T> v_ = std::vector<T>((Rows > 0 ? Rows : 0) * (Cols > 0 ? Cols : 0), T(0));
public:
    Matrix() {}
    Matrix(size_t r, size_t c) : r_(r), c_(c), v_(r * c, T(0)) {}
    explicit Matrix(size_t r) : r_(r), c_(Cols > 0 ? Cols : 1), v_(r * (Cols > 0 ? Cols : 1), T(0)) {}
    template <int R2, int C2> Matrix(const Matrix<T, R2, C2>& o) : r_(o.rows()), c_(o.cols()), v_(o.data(), o.data() + o.size()) {}
    static Matrix Zero(size_t r, size_t c = (Cols > 0 ? Cols : 1)) { return Matrix(r, c); }
    static Matrix Zero() { return Matrix(Rows > 0 ? Rows : 0, Cols > 0 ? Cols : 0); }
    size_t rows() const { return r_; } size_t cols() const { return c_; } size_t size() const { return v_.size(); }
    T* data() { return v_.data(); } const T* data() const { return v_.data(); }
    T& operator()(size_t i, size_t j) { return v_[i + j * r_]; } const T& operator()(size_t i, size_t j) const { return v_[i + j * r_]; }
    T& operator()(size_t i) { return v_[i]; } const T& operator()(size_t i) const { return v_[i]; }
    void resize(size_t r, size_t c = 1) { r_ = r; c_ = c; v_.assign(r * c, T(0)); }
    Matrix<T> block(size_t i0, size_t j0, size_t nr, size_t nc) const {
        Matrix<T> b(nr, nc);
        for (size_t j = 0; j < nc; j++) for (size_t i = 0; i < nr; i++) b(i, j) = (*this)(i0 + i, j0 + j);
        return b;
    }
    Matrix<T> head(size_t n) const { return block(0, 0, n, 1); }
    Matrix<T> tail(size_t n) const { return block(r_ - n, 0, n, 1); }
    Matrix<T> transpose() const { Matrix<T> t(c_, r_); for (size_t j = 0; j < c_; j++) for (size_t i = 0; i < r_; i++) t(j, i) = (*this)(i, j); return t; }
    template <int R2, int C2> T dot(const Matrix<T, R2, C2>& o) const {
        if (o.size() != size()) throw std::invalid_argument("Matrix::dot: size mismatch");
        T s = 0; for (size_t k = 0; k < v_.size(); k++) s += v_[k] * o.data()[k]; return s;
    }
    T norm() const { T s = 0; for (T x : v_) s += x * x; return std::sqrt(s); }
    template <int R2, int C2> Matrix& operator+=(const Matrix<T, R2, C2>& o) { same(o); for (size_t k = 0; k < v_.size(); k++) v_[k] += o.data()[k]; return *this; }
    template <int R2, int C2> Matrix& operator-=(const Matrix<T, R2, C2>& o) { same(o); for (size_t k = 0; k < v_.size(); k++) v_[k] -= o.data()[k]; return *this; }
    Matrix& operator*=(T a) { for (T& x : v_) x *= a; return *this; }
private:
    template <int R2, int C2> void same(const Matrix<T, R2, C2>& o) const { if (o.rows() != r_ || o.cols() != c_) throw std::invalid_argument("Matrix: shape mismatch"); }
};
template <typename T, int R1, int C1, int R2, int C2> Matrix<T> operator+(const Matrix<T, R1, C1>& a, const Matrix<T, R2, C2>& b) { Matrix<T> r(a); r += b; return r; }
template <typename T, int R1, int C1, int R2, int C2> Matrix<T> operator-(const Matrix<T, R1, C1>& a, const Matrix<T, R2, C2>& b) { Matrix<T> r(a); r -= b; return r; }
template <typename T, int R1, int C1> Matrix<T> operator*(T s, const Matrix<T, R1, C1>& a) { Matrix<T> r(a); r *= s; return r; }
template <typename T, int R1, int C1> Matrix<T> operator*(const Matrix<T, R1, C1>& a, T s) { Matrix<T> r(a); r *= s; return r; }
template <typename T, int R1, int C1, int R2, int C2> Matrix<T> operator*(const Matrix<T, R1, C1>& a, const Matrix<T, R2, C2>& b) {
    if (a.cols() != b.rows()) throw std::invalid_argument("Matrix product: shape mismatch");
    Matrix<T> r(a.rows(), b.cols());
    for (size_t j = 0; j < b.cols(); j++) for (size_t k = 0; k < a.cols(); k++) { const T bkj = b(k, j); for (size_t i = 0; i < a.rows(); i++) r(i, j) += a(i, k) * bkj; }
    return r;
}

// ---------------------------------------------------------------------------------------------- mesh handles
enum class element_location { IN_NEGATIVE_SIDE, IN_POSITIVE_SIDE, ON_INTERFACE, UNDEF };     // cuthho_mesh.hpp:31-36
enum class boundary { NONE, DIRICHLET, NEUMANN, ROBIN };                                     // basic_mesh.hpp:37-43

struct cell_user_data { element_location location = element_location::UNDEF; bool distorted = false; };
struct cell_handle { size_t id = 0; cell_user_data user_data; };
struct face_handle { size_t id = 0; bool is_boundary = false; boundary bndtype = boundary::NONE; };

template <typename T> class local_batch;

// common part of quad_mesh / cuthho_poly_mesh: the Cartesian generator of basic_mesh.hpp:230-298 plus iterable handles
template <typename T>
class reference_mesh : public cartesian_mesh<T> {
public:
    typedef T coordinate_type;
    typedef point<T, 2> point_type;
    typedef cell_handle cell_type;
    typedef face_handle face_type;
    std::vector<cell_handle> cells;
    std::vector<face_handle> faces;
    reference_mesh(const mesh_init_params<T>& mip, typename cartesian_mesh<T>::kind k) : cartesian_mesh<T>(mip, k), batch_(new local_batch<T>(*this)) {
        cells.resize(this->num_cells()); for (size_t i = 0; i < cells.size(); i++) cells[i].id = i;
        const size_t Nx = mip.Nx, Ny = mip.Ny;
        faces.resize(this->num_faces());
        for (size_t f = 0; f < faces.size(); f++) faces[f].id = f;
        // boundary faces of the generator are DIRICHLET (basic_mesh.hpp:293-297); closed-form ids: SURVEY App. B
        auto mark = [&](size_t f) { faces[f].is_boundary = true; faces[f].bndtype = boundary::DIRICHLET; };
        for (size_t i = 0; i < Nx; i++) { mark(2 * i); mark(Ny * (2 * Nx + 1) + i); }
        for (size_t j = 0; j < Ny; j++) { mark(j * (2 * Nx + 1) + 1); mark(j * (2 * Nx + 1) + 2 * Nx); }
    }
    ~reference_mesh() { delete batch_; }
    local_batch<T>& batch() const { return *batch_; }
private:
    local_batch<T>* batch_;
};
template <typename T> struct quad_mesh : reference_mesh<T> {
    explicit quad_mesh(const mesh_init_params<T>& mip) : reference_mesh<T>(mip, cartesian_mesh<T>::quad) {}
};
template <typename T> struct cuthho_poly_mesh : reference_mesh<T> {
    explicit cuthho_poly_mesh(const mesh_init_params<T>& mip) : reference_mesh<T>(mip, cartesian_mesh<T>::cuthho_poly) {}
};
template <typename T> size_t offset(const reference_mesh<T>&, const cell_handle& cl) { return cl.id; }
template <typename T> size_t offset(const reference_mesh<T>&, const face_handle& fc) { return fc.id; }
template <typename T> element_location location(const reference_mesh<T>&, const cell_handle& cl) { return cl.user_data.location; }

}  // namespace proton_b200
#include "proton_reference_api_batch.hpp"
