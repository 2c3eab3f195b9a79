/* ORACLE — TEST INFRASTRUCTURE ONLY (see orc_core.hpp header).
 * extern "C" surface over the restatement so tests/ and bench.py's cpu_baseline leg can drive it through
 * ctypes.  Drivers restated here: the assembly loops of apps/convergence_test/convergence_test.cpp:199-217,
 * apps/cuthho/cuthho_square.cpp:883-902, apps/obstacle/obstacle.cpp:147-158 and their error norms. */
#include "orc_cut.hpp"
#include "orc_obstacle.hpp"
#include "orc_interface.hpp"
#include <chrono>
#include <cstring>
#include <string>

using namespace orc;

namespace {
thread_local std::string g_err;

struct MeshH { Mesh m; LevelSet ls; bool cut_done = false; };

Fun rhs_fun(int kind) {
    switch (kind) {
    case 0: return [](Pt p) { return 2.0 * M_PI * M_PI * std::sin(M_PI * p.x) * std::sin(M_PI * p.y); };  /* convergence_test.cpp:100-102 */
    case 1: return obstacle_rhs;
    default: return [](Pt p) {                                                                            /* convergence_test.cpp:116-123 (commented-out tanh problem) */
        double th1 = std::tanh(30 - 60 * (p.x - p.y)), ch1 = std::cosh(30 - 60 * (p.x - p.y));
        double th2 = std::tanh(60 * p.y), ch2 = std::cosh(60 * p.y);
        return 14400 * th1 / (ch1 * ch1) + 7200 * th2 / (ch2 * ch2); };
    }
}
Fun sol_fun(int kind) {
    switch (kind) {
    case 0: return [](Pt p) { return std::sin(M_PI * p.x) * std::sin(M_PI * p.y); };                      /* :104-106 */
    case 1: return obstacle_sol;
    default: return [](Pt p) { return std::tanh(60 * p.y) - std::tanh(60 * (p.x - p.y) - 30); };          /* :125-127 */
    }
}
void sol_grad(int kind, Pt p, double g[2]) {
    if (kind == 0) { g[0] = M_PI * std::cos(M_PI * p.x) * std::sin(M_PI * p.y); g[1] = M_PI * std::sin(M_PI * p.x) * std::cos(M_PI * p.y); }   /* :108-113 */
    else { g[0] = g[1] = 0; }
}

struct Problem { DegreeInfo di; int stab; int problem; int fun; size_t rhs_di; Problem(int cd, int fd) : di(cd, fd) {} };

/* stab: 0 naive (hho.hpp:99), 1 fancy (hho.hpp:155), 2 cut (cuthho_square.cpp:566).  problem: 0 plain HHO, 1 fictitious domain. */
void local_ops(const MeshH& H, const Cell& cl, const Problem& P, Mat& oper, Mat& data, Mat& stab, Vec& rhs) {
    const int where = IN_NEGATIVE_SIDE;
    if (P.problem == 1) {
        auto gr = make_hho_laplacian_cut(H.m, cl, H.ls, P.di, where);
        oper = gr.first; data = gr.second;
        stab = make_hho_cut_stabilization(H.m, cl, P.di, where);
        rhs = make_rhs_fictdom(H.m, cl, P.di.cd, rhs_fun(P.fun), where, H.ls, sol_fun(P.fun));
    } else {
        auto gr = make_hho_laplacian(H.m, cl, P.di);
        oper = gr.first; data = gr.second;
        stab = (P.stab == 1) ? make_hho_fancy_stabilization(H.m, cl, oper, P.di) : make_hho_naive_stabilization(H.m, cl, P.di);
        rhs = make_rhs(H.m, cl, P.di.cd, rhs_fun(P.fun), P.rhs_di);
    }
}

struct SysH { Assembler as; const MeshH* mh; Problem P; SysH(const MeshH* h, const Problem& p) : as(h->m, p.di), mh(h), P(p) {} };
struct IfcH { InterfaceAssembler as; const MeshH* mh; int fun; IfcH(const MeshH* h, InterfaceAssembler&& a, int f) : as(std::move(a)), mh(h), fun(f) {} };
struct ObsH { ObstacleAssembler as; const MeshH* mh; ObsH(const MeshH* h, const std::vector<bool>& inA, DegreeInfo di) : as(h->m, inA, di), mh(h) {} };
}  // namespace

#define ORC_TRY try {
#define ORC_CATCH(rv) } catch (const std::exception& e) { g_err = e.what(); return rv; }

extern "C" {

const char* orc_last_error() { return g_err.c_str(); }

void* orc_mesh_new(long long Nx, long long Ny, double minx, double maxx, double miny, double maxy, int poly) {
    MeshParams p; p.Nx = Nx; p.Ny = Ny; p.min_x = minx; p.max_x = maxx; p.min_y = miny; p.max_y = maxy;
    auto* h = new MeshH; h->m = Mesh(p, poly != 0); return h;
}
void orc_mesh_free(void* h) { delete (MeshH*)h; }
void orc_mesh_sizes(void* h, long long out[3]) { auto& m = ((MeshH*)h)->m; out[0] = m.cells.size(); out[1] = m.faces.size(); out[2] = m.points.size(); }
void orc_mesh_points(void* h, double* xy) { auto& m = ((MeshH*)h)->m; for (size_t i = 0; i < m.points.size(); i++) { xy[2 * i] = m.points[i].x; xy[2 * i + 1] = m.points[i].y; } }
void orc_mesh_cells(void* h, long long* ptids, long long* fids) {
    auto& m = ((MeshH*)h)->m;
    for (size_t c = 0; c < m.cells.size(); c++) {
        auto f = face_ids(m, m.cells[c]);
        for (int i = 0; i < 4; i++) { ptids[4 * c + i] = m.cells[c].ptids[i]; fids[4 * c + i] = f[i]; }
    }
}
void orc_mesh_faces(void* h, long long* ptids, int* dirichlet) {
    auto& m = ((MeshH*)h)->m;
    for (size_t f = 0; f < m.faces.size(); f++) { ptids[2 * f] = m.faces[f].ptids[0]; ptids[2 * f + 1] = m.faces[f].ptids[1]; dirichlet[f] = m.faces[f].is_boundary && m.faces[f].dirichlet; }
}

/* cuthho_square.cpp:2030-2052 */
int orc_cut_preprocess(void* h, int ls_kind, double radius, double cx, double cy, double cut_y, int refsteps) {
    ORC_TRY
    auto* H = (MeshH*)h;
    H->ls.kind = ls_kind; H->ls.radius = radius; H->ls.alpha = cx; H->ls.beta = cy; H->ls.cut_y = cut_y;
    cut_preprocess(H->m, H->ls, refsteps); H->cut_done = true; return 0;
    ORC_CATCH(1)
}
/* iface_off has ncells+1 entries; call with iface_pts == NULL first to size it */
void orc_cut_info(void* h, int* node_loc, int* node_displaced, int* face_loc, double* face_ip, int* cell_loc, long long* iface_off, double* iface_pts) {
    auto& m = ((MeshH*)h)->m;
    for (size_t i = 0; i < m.nodes.size(); i++) { node_loc[i] = m.nodes[i].location; node_displaced[i] = m.nodes[i].displaced; }
    for (size_t f = 0; f < m.faces.size(); f++) { face_loc[f] = m.faces[f].location; face_ip[2 * f] = m.faces[f].intersection_point.x; face_ip[2 * f + 1] = m.faces[f].intersection_point.y; }
    long long o = 0;
    for (size_t c = 0; c < m.cells.size(); c++) {
        cell_loc[c] = m.cells[c].location; iface_off[c] = o;
        if (iface_pts) for (auto& p : m.cells[c].interface) { iface_pts[2 * o] = p.x; iface_pts[2 * o + 1] = p.y; o++; }
        else o += m.cells[c].interface.size();
    }
    iface_off[m.cells.size()] = o;
}

void orc_dims(int cd, int fd, int out[4]) { DegreeInfo di(cd, fd); out[0] = (int)cell_basis_size(di.cd); out[1] = (int)face_basis_size(di.fd); out[2] = (int)cell_basis_size(di.rd); out[3] = out[0] + 4 * out[1]; }

/* One cell.  oper is (rows x n) column-major with rows = rbs-1 (uncut) or rbs (cut cell of the fictdom problem);
 * *oper_rows receives it.  data, stab: n x n column-major.  rhs: cbs. */
int orc_local(void* h, long long cell, int cd, int fd, int stab_kind, int problem, int fun, int rhs_di,
              double* oper, int* oper_rows, double* data, double* stab, double* rhs) {
    ORC_TRY
    auto* H = (MeshH*)h; Problem P(cd, fd); P.stab = stab_kind; P.problem = problem; P.fun = fun; P.rhs_di = rhs_di;
    Mat o, d, s; Vec r;
    local_ops(*H, H->m.cells.at(cell), P, o, d, s, r);
    if (oper_rows) *oper_rows = o.r;
    if (oper) std::memcpy(oper, o.a.data(), o.a.size() * 8);
    if (data) std::memcpy(data, d.a.data(), d.a.size() * 8);
    if (stab) std::memcpy(stab, s.a.data(), s.a.size() * 8);
    if (rhs) std::memcpy(rhs, r.data(), r.size() * 8);
    return 0;
    ORC_CATCH(1)
}

/* Per-cell quadrature dump for unit tests: kind 0 = integrate(cell,deg) [where-aware when the mesh is cut and
 * problem==1], 1 = integrate_interface.  Returns the count; pts/w may be NULL to size. */
long long orc_cell_quadrature(void* h, long long cell, int degree, int kind, int where_aware, double* pts, double* w) {
    ORC_TRY
    auto* H = (MeshH*)h; const Cell& cl = H->m.cells.at(cell);
    std::vector<QP> q = (kind == 1) ? integrate_interface(H->m, cl, degree, IN_NEGATIVE_SIDE)
                        : (where_aware ? integrate_where(H->m, cl, degree, where_aware == 2 ? IN_POSITIVE_SIDE : IN_NEGATIVE_SIDE) : integrate(H->m, cl, degree));   /* where_aware: 1 negative part, 2 positive part */
    if (pts) for (size_t i = 0; i < q.size(); i++) { pts[2 * i] = q[i].first.x; pts[2 * i + 1] = q[i].first.y; w[i] = q[i].second; }
    return (long long)q.size();
    ORC_CATCH(-1)
}

/* Assembly loop of the drivers: local operators -> assembler.assemble -> finalize. */
void* orc_assemble(void* h, int cd, int fd, int stab_kind, int problem, int fun, int rhs_di, double* seconds_local, double* seconds_global) {
    ORC_TRY
    auto* H = (MeshH*)h; Problem P(cd, fd); P.stab = stab_kind; P.problem = problem; P.fun = fun; P.rhs_di = rhs_di;
    auto* S = new SysH(H, P);
    Fun bc = sol_fun(fun);
    double tl = 0, tg = 0;
    for (auto& cl : H->m.cells) {
        auto t0 = std::chrono::steady_clock::now();
        Mat o, d, s; Vec r; local_ops(*H, cl, P, o, d, s, r);
        Mat lc = d + s;
        auto t1 = std::chrono::steady_clock::now();
        S->as.assemble(H->m, cl, lc, r, bc);
        auto t2 = std::chrono::steady_clock::now();
        tl += std::chrono::duration<double>(t1 - t0).count(); tg += std::chrono::duration<double>(t2 - t1).count();
    }
    auto t0 = std::chrono::steady_clock::now();
    S->as.finalize();
    tg += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (seconds_local) *seconds_local = tl;
    if (seconds_global) *seconds_global = tg;
    return S;
    ORC_CATCH(nullptr)
}
/* The same loop with the local operators computed by `nthreads` OpenMP threads over the cells (the reference is serial; this
   is the "all host threads" variant of the CPU baseline): assembler.assemble and finalize stay serial and in cell order, so the
   system is identical to orc_assemble's. */
void* orc_assemble_mt(void* h, int cd, int fd, int stab_kind, int problem, int fun, int rhs_di, int nthreads, double* seconds_local, double* seconds_global) {
    ORC_TRY
    auto* H = (MeshH*)h; Problem P(cd, fd); P.stab = stab_kind; P.problem = problem; P.fun = fun; P.rhs_di = rhs_di;
    auto* S = new SysH(H, P);
    Fun bc = sol_fun(fun);
    const long long nc = (long long)H->m.cells.size();
    std::vector<Mat> lcs(nc); std::vector<Vec> rs(nc);
    auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for schedule(static) num_threads(nthreads)
    for (long long c = 0; c < nc; c++) {
        Mat o, d, s; local_ops(*H, H->m.cells[c], P, o, d, s, rs[c]);
        lcs[c] = d + s;
    }
    auto t1 = std::chrono::steady_clock::now();
    for (long long c = 0; c < nc; c++) S->as.assemble(H->m, H->m.cells[c], lcs[c], rs[c], bc);
    S->as.finalize();
    auto t2 = std::chrono::steady_clock::now();
    if (seconds_local) *seconds_local = std::chrono::duration<double>(t1 - t0).count();
    if (seconds_global) *seconds_global = std::chrono::duration<double>(t2 - t1).count();
    return S;
    ORC_CATCH(nullptr)
}
int orc_omp_enabled() {
#ifdef _OPENMP
    return 1;
#else
    return 0;
#endif
}
void orc_sys_free(void* s) { delete (SysH*)s; }
void orc_sys_sizes(void* s, long long out[2]) { auto* S = (SysH*)s; out[0] = S->as.LHS.n; out[1] = S->as.LHS.val.size(); }
void orc_sys_get(void* s, long long* rowptr, int* col, double* val, double* rhs, long long* compress) {
    auto* S = (SysH*)s; auto& A = S->as.LHS;
    std::memcpy(rowptr, A.rowptr.data(), A.rowptr.size() * 8); std::memcpy(col, A.col.data(), A.col.size() * 4);
    std::memcpy(val, A.val.data(), A.val.size() * 8); std::memcpy(rhs, S->as.RHS.data(), S->as.RHS.size() * 8);
    if (compress) for (size_t i = 0; i < S->as.compress_table.size(); i++) compress[i] = S->as.compress_table[i];
}

/* solver_cg.hpp:63-144 on caller-supplied CSR arrays (symmetric matrices: CSR == the reference's CSC) */
int orc_cg(long long n, const long long* rowptr, const int* col, const double* val, const double* b, double tol, long long max_iter,
           int precond, double* x, long long* iters, double* relres, double* seconds) {
    Csr A; A.n = n; A.rowptr.assign(rowptr, rowptr + n + 1); A.col.assign(col, col + rowptr[n]); A.val.assign(val, val + rowptr[n]);
    Vec bb(b, b + n), xx; CgParams p; p.convergence_threshold = tol; p.max_iter = max_iter; p.apply_preconditioner = precond;
    size_t it; double rr;
    auto t0 = std::chrono::steady_clock::now();
    CgExit e = conjugated_gradient(A, bb, xx, p, &it, &rr);
    if (seconds) *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    std::memcpy(x, xx.data(), n * 8); if (iters) *iters = it; if (relres) *relres = rr;
    return (int)e;
}

/* convergence_test.cpp:249-306: out = {sum L2^2, sum mass-norm^2 (once per quadrature point, quirk kept),
 * sum energy^2 over qps of degree 2*cd (quirk kept)}; sol is the FULL system vector (cells then faces). */
int orc_errors_convergence(void* s, const double* sol, double out[3]) {
    ORC_TRY
    auto* S = (SysH*)s; const Mesh& m = S->mh->m; DegreeInfo hdi = S->P.di; Fun sf = sol_fun(S->P.fun);
    Vec solv(sol, sol + S->as.system_size);
    out[0] = out[1] = out[2] = 0; size_t cell_i = 0;
    for (auto& cl : m.cells) {
        CellBasis cb(m, cl, hdi.cd); size_t cbs = cb.size;
        Vec cdofs(solv.begin() + cell_i * cbs, solv.begin() + (cell_i + 1) * cbs);
        auto qps = integrate(m, cl, 2 * hdi.cd);
        Mat mass = make_mass_matrix(m, cl, hdi.cd);
        Vec real_dofs = LLT(mass).solve(make_rhs(m, cl, hdi.cd, sf));
        Vec diff(cbs); for (size_t i = 0; i < cbs; i++) diff[i] = real_dofs[i] - cdofs[i];
        double mm = dot(diff, matvec(mass, diff));
        for (auto& qp : qps) {
            double val = dot(cdofs, cb.eval_basis(qp.first)), rv = sf(qp.first);
            out[0] += qp.second * (rv - val) * (rv - val);
            out[1] += mm;
        }
        CellBasis rb(m, cl, hdi.rd);
        Vec alldofs = S->as.take_local_data(m, cl, solv, sf);
        auto gr = make_hho_laplacian(m, cl, hdi);
        Vec recdofs = matvec(gr.first, alldofs);
        for (auto& qp : qps) {
            Mat dphi = rb.eval_gradients(qp.first); double gx = 0, gy = 0;
            for (int i = 1; i < dphi.r; i++) { gx += recdofs[i - 1] * dphi(i, 0); gy += recdofs[i - 1] * dphi(i, 1); }
            double rg[2]; sol_grad(S->P.fun, qp.first, rg);
            out[2] += qp.second * ((rg[0] - gx) * (rg[0] - gx) + (rg[1] - gy) * (rg[1] - gy));
        }
        cell_i++;
    }
    return 0;
    ORC_CATCH(1)
}

/* cuthho_square.cpp:1031-1064: H1 error over NEG and CUT cells, gradient of the *cell* dofs in the rd basis */
int orc_error_fictdom(void* s, const double* sol, double* h1_sq) {
    ORC_TRY
    auto* S = (SysH*)s; const Mesh& m = S->mh->m; DegreeInfo hdi = S->P.di;
    *h1_sq = 0; size_t cell_i = 0; size_t cbs = cell_basis_size(hdi.cd), rbs = cell_basis_size(hdi.rd);
    for (auto& cl : m.cells) {
        if (cl.location == IN_NEGATIVE_SIDE || cl.location == ON_INTERFACE) {
            CellBasis rb(m, cl, hdi.rd);
            for (auto& qp : integrate_where(m, cl, 2 * hdi.cd, IN_NEGATIVE_SIDE)) {
                Mat d = rb.eval_gradients(qp.first); double gx = 0, gy = 0;
                for (size_t i = 1; i < rbs; i++) { gx += sol[cell_i * cbs + i] * d((int)i, 0); gy += sol[cell_i * cbs + i] * d((int)i, 1); }
                double rg[2]; sol_grad(0, qp.first, rg);
                *h1_sq += qp.second * ((rg[0] - gx) * (rg[0] - gx) + (rg[1] - gy) * (rg[1] - gy));
            }
        }
        cell_i++;
    }
    return 0;
    ORC_CATCH(1)
}

/* obstacle.cpp:147-158: one active-set assembly with HHO(0,fd), fancy stabilisation, rhs with di=1 */
void* orc_obstacle_assemble(void* h, int fd, const unsigned char* in_A, const double* gamma) {
    ORC_TRY
    auto* H = (MeshH*)h; DegreeInfo hdi(0, fd); size_t nc = H->m.cells.size();
    std::vector<bool> inA(nc); for (size_t i = 0; i < nc; i++) inA[i] = in_A[i] != 0;
    Vec g(gamma, gamma + nc);
    auto* O = new ObsH(H, inA, hdi);
    for (auto& cl : H->m.cells) {
        auto gr = make_hho_laplacian(H->m, cl, hdi);
        Mat lc = gr.second + make_hho_fancy_stabilization(H->m, cl, gr.first, hdi);
        Vec f = make_rhs(H->m, cl, hdi.cd, obstacle_rhs, 1);
        O->as.assemble(H->m, cl, lc, f, g, obstacle_sol);
    }
    O->as.finalize();
    return O;
    ORC_CATCH(nullptr)
}
void orc_obstacle_free(void* o) { delete (ObsH*)o; }
void orc_obstacle_sizes(void* o, long long out[2]) { auto* O = (ObsH*)o; out[0] = O->as.LHS.n; out[1] = O->as.LHS.val.size(); }
void orc_obstacle_get(void* o, long long* rowptr, int* col, double* val, double* rhs) {
    auto* O = (ObsH*)o; auto& A = O->as.LHS;
    std::memcpy(rowptr, A.rowptr.data(), A.rowptr.size() * 8); std::memcpy(col, A.col.data(), A.col.size() * 4);
    std::memcpy(val, A.val.data(), A.val.size() * 8); std::memcpy(rhs, O->as.RHS.data(), O->as.RHS.size() * 8);
}
/* alpha has ncells + fbs*nfaces entries (in/out), beta ncells (out) */
void orc_obstacle_expand(void* o, const double* sol, const double* gamma, double* alpha, double* beta) {
    auto* O = (ObsH*)o; const Mesh& m = O->mh->m; size_t nc = m.cells.size(), na = nc + face_basis_size(O->as.di.fd) * m.faces.size();
    Vec s(sol, sol + O->as.system_size), g(gamma, gamma + nc), a(alpha, alpha + na), b;
    O->as.expand_solution(m, s, obstacle_sol, g, a, b);
    std::memcpy(alpha, a.data(), na * 8); std::memcpy(beta, b.data(), nc * 8);
}
/* obstacle.cpp:199-218 */
double orc_obstacle_error(void* h, int fd, const double* alpha) {
    auto* H = (MeshH*)h; const Mesh& m = H->m; DegreeInfo hdi(0, fd);
    Vec a(alpha, alpha + m.cells.size() + face_basis_size(fd) * m.faces.size());
    double error = 0;
    for (auto& cl : m.cells) {
        Vec local = take_local_data_expanded(m, cl, hdi, a);
        Vec proj = project_function(m, cl, hdi, obstacle_sol, 1);
        auto gr = make_hho_laplacian(m, cl, hdi);
        Mat lc = gr.second + make_hho_fancy_stabilization(m, cl, gr.first, hdi);
        Vec diff(local.size()); for (size_t i = 0; i < diff.size(); i++) diff[i] = local[i] - proj[i];
        error += dot(diff, matvec(lc, diff));
    }
    return std::sqrt(error);
}
/* cell barycenters (obstacle.cpp:108-114 evaluates gamma there) */
void orc_cell_barycenters(void* h, double* xy) { auto& m = ((MeshH*)h)->m; for (size_t c = 0; c < m.cells.size(); c++) { Pt b = barycenter(m, m.cells[c]); xy[2 * c] = b.x; xy[2 * c + 1] = b.y; } }

}  // extern "C"

extern "C" void orc_mesh_set_points(void* h, const double* xy) {
    auto& m = ((MeshH*)h)->m;
    for (size_t i = 0; i < m.points.size(); i++) m.points[i] = Pt(xy[2 * i], xy[2 * i + 1]);
}

/* ---- two-material interface problem (`cuthho_square -i`, cuthho_square.cpp:1625-1846); needs orc_cut_preprocess first ---- */
extern "C" {
void* orc_interface_assemble(void* h, int degree, int fun) {
    ORC_TRY
    auto* H = (MeshH*)h;
    if (!H->cut_done) throw std::logic_error("orc_interface_assemble: run orc_cut_preprocess first");
    return new IfcH(H, assemble_interface_problem(H->m, H->ls, (size_t)degree, rhs_fun(fun), sol_fun(fun)), fun);
    ORC_CATCH(nullptr)
}
void orc_interface_free(void* s) { delete (IfcH*)s; }
void orc_interface_sizes(void* s, long long out[4]) {
    auto* S = (IfcH*)s; out[0] = S->as.LHS.n; out[1] = S->as.LHS.val.size(); out[2] = S->as.num_all_cells; out[3] = S->as.num_all_faces;
}
void orc_interface_get(void* s, long long* rowptr, int* col, double* val, double* rhs, long long* cell_table, long long* face_table) {
    auto* S = (IfcH*)s; auto& A = S->as.LHS;
    std::memcpy(rowptr, A.rowptr.data(), A.rowptr.size() * 8); std::memcpy(col, A.col.data(), A.col.size() * 4);
    std::memcpy(val, A.val.data(), A.val.size() * 8); std::memcpy(rhs, S->as.RHS.data(), S->as.RHS.size() * 8);
    if (cell_table) for (size_t i = 0; i < S->as.cell_table.size(); i++) cell_table[i] = S->as.cell_table[i];
    if (face_table) for (size_t i = 0; i < S->as.face_table.size(); i++) face_table[i] = S->as.face_table[i];
}
/* local matrix of a CUT cell in the driver's unknown order [cell-, cell+, faces-, faces+] (2n x 2n column-major) and its load (2 cbs) */
int orc_interface_local(void* h, long long cell, int degree, double* lc_out, double* rhs_out) {
    ORC_TRY
    auto* H = (MeshH*)h; const Cell& cl = H->m.cells.at(cell); DegreeInfo hdi(degree + 1, degree); InterfaceParams parms;
    const int cbs = (int)cell_basis_size(hdi.cd), nfd = 4 * (int)face_basis_size(hdi.fd);
    Mat lc = make_hho_laplacian_interface(H->m, cl, H->ls, hdi, parms);
    Mat sn = make_hho_cut_stabilization(H->m, cl, hdi, IN_NEGATIVE_SIDE), sp = make_hho_cut_stabilization(H->m, cl, hdi, IN_POSITIVE_SIDE);
    lc.add_block(0, 0, sn.block(0, 0, cbs, cbs)); lc.add_block(0, 2 * cbs, sn.block(0, cbs, cbs, nfd));
    lc.add_block(2 * cbs, 0, sn.block(cbs, 0, nfd, cbs)); lc.add_block(2 * cbs, 2 * cbs, sn.block(cbs, cbs, nfd, nfd));
    lc.add_block(cbs, cbs, sp.block(0, 0, cbs, cbs)); lc.add_block(cbs, 2 * cbs + nfd, sp.block(0, cbs, cbs, nfd));
    lc.add_block(2 * cbs + nfd, cbs, sp.block(cbs, 0, nfd, cbs)); lc.add_block(2 * cbs + nfd, 2 * cbs + nfd, sp.block(cbs, cbs, nfd, nfd));
    std::memcpy(lc_out, lc.a.data(), lc.a.size() * 8);
    if (rhs_out) {
        Vec fn = make_rhs_where(H->m, cl, hdi.cd, IN_NEGATIVE_SIDE, rhs_fun(0)), fp = make_rhs_where(H->m, cl, hdi.cd, IN_POSITIVE_SIDE, rhs_fun(0));
        std::memcpy(rhs_out, fn.data(), cbs * 8); std::memcpy(rhs_out + cbs, fp.data(), cbs * 8);
    }
    return 0;
    ORC_CATCH(1)
}
/* cuthho_square.cpp:1769-1837: H1 error^2 of the cell unknowns, both parts of the cut cells */
int orc_interface_error(void* s, const double* sol, double* h1_sq) {
    ORC_TRY
    auto* S = (IfcH*)s; Vec solv(sol, sol + S->as.system_size); const int fun = S->fun;
    *h1_sq = interface_h1_error_sq(S->mh->m, S->as, solv, [fun](Pt p, double g[2]) { sol_grad(fun, p, g); });
    return 0;
    ORC_CATCH(1)
}
}
/* 0 = the restatement of Eigen's linear setFromTriplets (default), 1 = the earlier std::sort build (timing comparison only) */
extern "C" void orc_set_finalize(int kind) { orc_finalize_kind = kind; }
