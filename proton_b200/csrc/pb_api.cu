// C ABI of proton_b200 (include/proton_b200.h): context, mesh, dispatch of the hand-written sm_100a kernels.
#include <cuda_runtime.h>
#include <cub/device/device_scan.cuh>
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_reduce.cuh>
#include <cub/device/device_select.cuh>
#include <cub/iterator/counting_input_iterator.cuh>
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "pb_local.cuh"
#include "pb_cut.cuh"
#include "pb_interface.cuh"
#include "pb_norms.cuh"
#include "pb_pcg.cuh"
#include "pb_mg.cuh"
#include "pb_mg_lines.cuh"
#include "pb_system.cuh"
#include "pb_generic.cuh"
#include "pb_obstacle.cuh"

using namespace pb;

static std::string g_create_error;

struct pb_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    std::string err;
    // mesh
    bool have_mesh = false;
    Grid grid{}; Strip strip{}; int quad_kind = 0;
    long long npoints = 0, ncells_local = 0, ncells_global = 0;
    double2* points = nullptr;
    StiffEnt* stiff_tab = nullptr; int stiff_key = -1;      // K1: stiff_table of the last reconstruction degree
    // cutHHO user data (whole mesh)
    LevelSet ls{}; bool have_ls = false, have_cut = false;
    // node_loc / face_loc / face_ip / cell_loc_g are indexed by GLOBAL point / face / cell id but only hold the window of rows this
    // context needs (cut_window, pb_api_cut.inl): the pointers are the allocations (*_a) shifted back by the window start
    int *node_loc = nullptr, *face_loc = nullptr, *cell_loc_g = nullptr, *cut_cells = nullptr;
    double2 *face_ip = nullptr, *iface = nullptr;
    int *node_loc_a = nullptr, *face_loc_a = nullptr, *cell_loc_a = nullptr; double2* face_ip_a = nullptr;
    long long pw0 = 0, pw1 = 0, fw0 = 0, fw1 = 0, cw0 = 0, cw1 = 0;      // windows [w0, w1) of point, face and cell ids
    int ncut = 0, NI = 0, ncut_owned = 0;
    long long nfaces = 0;
    // assembly
    bool assembled = false;
    pb_assembly_params prm{};
    int CD = 0, FD = 0, C = 0, F = 0, R = 0, N = 0, NF = 0;
    double *S = nullptr, *g = nullptr, *recA = nullptr, *recb = nullptr, *lc = nullptr, *oper = nullptr, *rhsloc = nullptr, *data = nullptr;
    long long nrows_blk = 0, nrows = 0, nnz = 0;
    BlockRow* brows = nullptr; long long* brp = nullptr;
    long long* rowptr = nullptr; int* col = nullptr; double *val = nullptr, *rhs = nullptr;
    // interface problem (pb_api_interface.inl): per-cut-cell blocks, face_table, general-CSR flag
    double *Sx = nullptr, *gx = nullptr, *recAx = nullptr, *recbx = nullptr, *lcx = nullptr, *fx = nullptr;
    int* ftab = nullptr; int* ftab_a = nullptr; bool general_csr = false;      // ftab: indexed by GLOBAL face id over the face window [fw0, fw1] (ftab_a shifted back)
    long long ifc_cut_before = -1, ifc_cut_total = -1;      // strips: cut faces owned by the strips below / by all strips (pb_interface_set_cut_offsets, or all-reduced)
    long long ifc_nglob = 0;                                 // unknowns of the whole interface system (modulus of the sort keys)
    // block rows of the system this context owns [rb_begin, rb_end) and references [rb_lo, rb_hi), in the numbering of the assembled
    // problem (plain / fictitious domain: compressed faces = strip.cf_*, col_*; interface: face_table slots), and the block rows its
    // neighbours need (halo_prev: my first face row, the upper halo of rank-1; halo_next: my last face row, the lower halo of rank+1)
    long long rb_begin = 0, rb_end = 0, rb_lo = 0, rb_hi = 0, halo_prev = 0, halo_next = 0;
    int* irr_rows = nullptr; int n_irr_rows = 0; int* irr_faces = nullptr; int n_irr_faces = 0; bool ifc_struct_ok = false;
    unsigned long long *ifc_k0 = nullptr, *ifc_k1 = nullptr; double *ifc_v0 = nullptr, *ifc_v1 = nullptr; long long* ifc_dnum = nullptr;
    void* ifc_tmp = nullptr; size_t ifc_tmp_bytes = 0; long long ifc_T = 0, ifc_ncut_alloc = 0;       // block rows of the cut cells' faces (scalar SpMV pass)
    // solver
    double *x = nullptr, *r = nullptr, *z = nullptr, *d = nullptr, *y = nullptr, *iA = nullptr, *iAb = nullptr, *parts = nullptr;
    CgState* cg = nullptr;
    double *dfull = nullptr, *scal = nullptr;      // d with halos (d = dfull + halo_lo); all-reduced scalars
    long long solver_n = -1, solver_nfull = -1, iAb_n = -1;      // sizes the solver vectors / iAb were allocated for
    long long asm_gen = 0;                          // bumped whenever the system values change (multigrid setup cache)
    MgState mg;                                     // multigrid preconditioner (pb_api_mg.inl)
    // obstacle active-set state (per cell)
    double *ob_alpha = nullptr, *ob_beta = nullptr, *ob_gamma = nullptr, *ob_xprev = nullptr; unsigned char* ob_inA = nullptr; bool ob_ready = false;
    // multi-GPU
    int rank = 0, world = 1; void* comm = nullptr;
    cudaStream_t s_halo = nullptr; cudaEvent_t ev_fork = nullptr, ev_join = nullptr;      // halo exchange beside the interior rows of the SpMV
    // timings
    cudaEvent_t ev[4]{};
    double ms_local = 0, ms_scatter = 0, ms_pcg = 0;
    // pipelined steps (pb_assemble_pipelined): copy streams and their events
    cudaStream_t s_in = nullptr, s_out = nullptr;
    cudaEvent_t ev_in = nullptr, ev_out = nullptr;
    bool pipe_in_flight = false, out_pending = false, step_timed = false;
    // residual history of the last pb_pcg (pb_pcg_history)
    double* hist = nullptr; long long hist_cap = 0, hist_n = 0;
    size_t alloc_bytes = 0;
};

static pb_status fail(pb_ctx* c, pb_status st, const char* fmt, ...) {
    char buf[512]; va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    if (c) c->err = buf; else g_create_error = buf;
    return st;
}
#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(ctx, PB_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); } while (0)
#define NEED(cond, st, msg) do { if (!(cond)) return fail(ctx, st, "%s", msg); } while (0)
// entry points that read or overwrite buffers of a pipelined step first wait for the steps still in flight
#define DRAIN() do { if (ctx->pipe_in_flight) { pb_status st_ = pb_pipeline_sync(ctx); if (st_ != PB_OK) return st_; } } while (0)

template <typename T> static cudaError_t dalloc(pb_ctx* c, T** p, size_t n) {
    if (*p) { cudaFree(*p); *p = nullptr; }
    if (n == 0) n = 1;
    cudaError_t e = cudaMalloc((void**)p, n * sizeof(T));
    if (e == cudaSuccess) c->alloc_bytes += n * sizeof(T);
    return e;
}
template <typename T> static void dfree(T** p) { if (*p) cudaFree(*p); *p = nullptr; }

// Gauss-Legendre closed forms in the reference's node order (quadratures.hpp:96-150) and the Dunavant rows
// (quadratures_dunavant.hpp:27-118; published constants, stored as symmetry orbits and expanded in the reference order)
static void fill_quad_tables(QuadTables& t) {
    memset(&t, 0, sizeof t);
    t.gx[1][0] = 0.0; t.gw[1][0] = 2.0;
    { double q = 1.0 / std::sqrt(3.0); t.gx[2][0] = -q; t.gx[2][1] = q; t.gw[2][0] = t.gw[2][1] = 1.0; }
    { double q = std::sqrt(3.0 / 5.0); t.gx[3][0] = -q; t.gx[3][1] = q; t.gx[3][2] = 0.0; t.gw[3][0] = t.gw[3][1] = 5.0 / 9.0; t.gw[3][2] = 8.0 / 9.0; }
    { double a1 = 3.0 / 7.0, a2 = 2.0 * std::sqrt(6.0 / 5.0) / 7.0;
      double q = std::sqrt(a1 - a2), w = (18.0 + std::sqrt(30.0)) / 36.0; t.gx[4][0] = -q; t.gx[4][1] = q; t.gw[4][0] = t.gw[4][1] = w;
      q = std::sqrt(a1 + a2); w = (18.0 - std::sqrt(30.0)) / 36.0; t.gx[4][2] = -q; t.gx[4][3] = q; t.gw[4][2] = t.gw[4][3] = w; }
    { t.gx[5][0] = 0.0; t.gw[5][0] = 128.0 / 225.0;
      double a1 = 5.0, a2 = 2.0 * std::sqrt(10.0 / 7.0);
      double q = std::sqrt(a1 - a2) / 3.0, w = (322 + 13.0 * std::sqrt(70.0)) / 900.0; t.gx[5][1] = -q; t.gx[5][2] = q; t.gw[5][1] = t.gw[5][2] = w;
      q = std::sqrt(a1 + a2) / 3.0; w = (322 - 13.0 * std::sqrt(70.0)) / 900.0; t.gx[5][3] = -q; t.gx[5][4] = q; t.gw[5][3] = t.gw[5][4] = w; }
    for (int a = 0; a < 45; a++) { t.mono_x[a] = mono_px(a); t.mono_y[a] = mono_py(a); }
    struct Orb { int rule, kind; double a, b, c, w; };
    static const Orb orbs[] = {
        {0, 1, 0.333333333333333, 0, 0, 1.000000000000000},
        {1, 3, 0.666666666666667, 0.166666666666667, 0, 0.333333333333333},
        {2, 1, 0.333333333333333, 0, 0, -0.562500000000000}, {2, 3, 0.600000000000000, 0.200000000000000, 0, 0.520833333333333},
        {3, 3, 0.108103018168070, 0.445948490915965, 0, 0.223381589678011}, {3, 3, 0.816847572980459, 0.091576213509771, 0, 0.109951743655322},
        {4, 1, 0.333333333333333, 0, 0, 0.225000000000000}, {4, 3, 0.059715871789770, 0.470142064105115, 0, 0.132394152788506},
        {4, 3, 0.797426985353087, 0.101286507323456, 0, 0.125939180544827},
        {5, 3, 0.501426509658179, 0.249286745170910, 0, 0.116786275726379}, {5, 3, 0.873821971016996, 0.063089014491502, 0, 0.050844906370207},
        {5, 6, 0.053145049844817, 0.310352451033784, 0.636502499121399, 0.082851075618374},
        {6, 1, 0.333333333333333, 0, 0, -0.149570044467682}, {6, 3, 0.479308067841920, 0.260345966079040, 0, 0.175615257433208},
        {6, 3, 0.869739794195568, 0.065130102902216, 0, 0.053347235608838},
        {6, 6, 0.048690315425316, 0.312865496004874, 0.638444188569810, 0.077113760890257},
        {7, 1, 0.333333333333333, 0, 0, 0.144315607677787}, {7, 3, 0.081414823414554, 0.459292588292723, 0, 0.095091634267285},
        {7, 3, 0.658861384496480, 0.170569307751760, 0, 0.103217370534718}, {7, 3, 0.898905543365938, 0.050547228317031, 0, 0.032458497623198},
        {7, 6, 0.008394777409958, 0.263112829634638, 0.728492392955404, 0.027230314174435},
    };
    for (const Orb& o : orbs) {
        int& n = t.dun_n[o.rule];
        auto put = [&](double l0, double l1, double l2) { double* r = t.dun[o.rule][n++]; r[0] = l0; r[1] = l1; r[2] = l2; r[3] = o.w; };
        if (o.kind == 1) put(o.a, o.a, o.a);
        else if (o.kind == 3) { put(o.a, o.b, o.b); put(o.b, o.a, o.b); put(o.b, o.b, o.a); }
        else { put(o.a, o.b, o.c); put(o.a, o.c, o.b); put(o.b, o.a, o.c); put(o.b, o.c, o.a); put(o.c, o.a, o.b); put(o.c, o.b, o.a); }
    }
}

extern "C" int pb_abi_version(void) { return 1; }

extern "C" pb_status pb_ctx_create(int device, pb_ctx** out) {
    pb_ctx* ctx = nullptr;
    if (!out) return fail(nullptr, PB_ERR_INVALID, "pb_ctx_create: null output pointer");
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(nullptr, PB_ERR_CUDA, "pb_ctx_create: no CUDA device (%s); proton_b200 has no CPU fallback", e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
    if (device < 0 || device >= ndev) return fail(nullptr, PB_ERR_INVALID, "pb_ctx_create: device %d out of range [0,%d)", device, ndev);
    if ((e = cudaSetDevice(device)) != cudaSuccess) return fail(nullptr, PB_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));
    ctx = new pb_ctx; ctx->device = device;
    QuadTables t; fill_quad_tables(t);
    if ((e = cudaMemcpyToSymbol(c_qt, &t, sizeof t)) != cudaSuccess || (e = cudaStreamCreate(&ctx->stream)) != cudaSuccess) {
        pb_status st = fail(nullptr, PB_ERR_CUDA, "pb_ctx_create: %s", cudaGetErrorString(e)); delete ctx; return st;
    }
    for (auto& ev : ctx->ev) cudaEventCreate(&ev);
    cudaMalloc((void**)&ctx->cg, sizeof(CgState));
    *out = ctx; return PB_OK;
}

static void mg_free(pb_ctx* c);      // pb_api_mg.inl
static void free_solver(pb_ctx* c) {
    dfree(&c->x); dfree(&c->r); dfree(&c->z); dfree(&c->dfull); c->d = nullptr; dfree(&c->y); dfree(&c->iA); dfree(&c->iAb); dfree(&c->parts); dfree(&c->scal);
    c->solver_n = c->solver_nfull = c->iAb_n = -1;
    mg_free(c);
}
static void free_system(pb_ctx* c) {
    dfree(&c->S); dfree(&c->g); dfree(&c->recA); dfree(&c->recb); dfree(&c->lc); dfree(&c->oper); dfree(&c->rhsloc); dfree(&c->data);
    dfree(&c->Sx); dfree(&c->gx); dfree(&c->recAx); dfree(&c->recbx); dfree(&c->lcx); dfree(&c->fx); dfree(&c->ftab_a); c->ftab = nullptr; dfree(&c->irr_rows); c->n_irr_rows = 0; dfree(&c->irr_faces); c->n_irr_faces = 0; c->ifc_struct_ok = false;
    dfree(&c->ifc_k0); dfree(&c->ifc_k1); dfree(&c->ifc_v0); dfree(&c->ifc_v1); dfree(&c->ifc_dnum); if (c->ifc_tmp) { cudaFree(c->ifc_tmp); c->ifc_tmp = nullptr; } c->ifc_T = 0; c->ifc_ncut_alloc = 0; c->general_csr = false;
    dfree(&c->brows); dfree(&c->brp); dfree(&c->rowptr); dfree(&c->col); dfree(&c->val); dfree(&c->rhs);
    free_solver(c);
    dfree(&c->ob_alpha); dfree(&c->ob_beta); dfree(&c->ob_gamma); dfree(&c->ob_xprev); dfree(&c->ob_inA); c->ob_ready = false;
    c->assembled = false;
}

static void dist_release(pb_ctx* c);      // pb_api_dist.inl
static void free_cut(pb_ctx* c) {
    dfree(&c->node_loc_a); dfree(&c->face_loc_a); dfree(&c->cell_loc_a); dfree(&c->face_ip_a); dfree(&c->cut_cells); dfree(&c->iface);
    c->node_loc = c->face_loc = c->cell_loc_g = nullptr; c->face_ip = nullptr;
    c->have_cut = false; c->ncut = 0; c->NI = 0; c->ncut_owned = 0;
}

extern "C" pb_status pb_ctx_destroy(pb_ctx* ctx) {
    if (!ctx) return PB_OK;
    cudaSetDevice(ctx->device);
    free_system(ctx); free_cut(ctx); dfree(&ctx->hist); dfree(&ctx->points); dfree(&ctx->stiff_tab); dfree(&ctx->cg);
    dist_release(ctx);
    for (auto& ev : ctx->ev) if (ev) cudaEventDestroy(ev);
    if (ctx->ev_in) cudaEventDestroy(ctx->ev_in); if (ctx->ev_out) cudaEventDestroy(ctx->ev_out);
    if (ctx->s_in) cudaStreamDestroy(ctx->s_in); if (ctx->s_out) cudaStreamDestroy(ctx->s_out);
    if (ctx->s_halo) cudaStreamDestroy(ctx->s_halo); if (ctx->ev_fork) cudaEventDestroy(ctx->ev_fork); if (ctx->ev_join) cudaEventDestroy(ctx->ev_join);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx; return PB_OK;
}

extern "C" const char* pb_last_error(const pb_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

// ------------------------------------------------------------------------------------------------ mesh
// points p(i,j) = (min_x + i*hx, min_y + j*hy) exactly as the host generator rounds them (basic_mesh.hpp:240-249):
// product and sum rounded separately, no FMA contraction.
__global__ void gen_points_kernel(Grid g, double2* pts) {
    long long n = (g.Nx + 1) * (g.Ny + 1);
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) {
        long long j = t / (g.Nx + 1), i = t - j * (g.Nx + 1);
        pts[t] = make_double2(__dadd_rn(g.min_x, __dmul_rn((double)i, g.hx)), __dadd_rn(g.min_y, __dmul_rn((double)j, g.hy)));
    }
}

extern "C" pb_status pb_mesh_cartesian(pb_ctx* ctx, long long Nx, long long Ny, double min_x, double max_x, double min_y, double max_y,
                                       int quad_kind, long long row_begin, long long row_end) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(Nx >= 2 && Ny >= 2, PB_ERR_INVALID, "pb_mesh_cartesian: need Nx, Ny >= 2");
    NEED(quad_kind == PB_QUAD_TENSOR_GAUSS || quad_kind == PB_QUAD_POLY_DUNAVANT, PB_ERR_INVALID, "pb_mesh_cartesian: bad quad_kind");
    NEED(0 <= row_begin && row_begin < row_end && row_end <= Ny, PB_ERR_INVALID, "pb_mesh_cartesian: bad row range");
    NEED((Nx + 1) * (Ny + 1) < (1LL << 31) && 4 * (2 * Nx * Ny) < (1LL << 31), PB_ERR_UNSUPPORTED, "pb_mesh_cartesian: mesh too large for 32-bit column indices");
    CK(cudaSetDevice(ctx->device));
    free_system(ctx);
    Grid& g = ctx->grid;
    g.Nx = Nx; g.Ny = Ny; g.row_begin = row_begin; g.row_end = row_end; g.min_x = min_x; g.min_y = min_y;
    g.hx = (max_x - min_x) / Nx; g.hy = (max_y - min_y) / Ny;      // mesh_init_params::hx/hy, basic_mesh.hpp:190-196
    ctx->quad_kind = quad_kind;
    ctx->npoints = (Nx + 1) * (Ny + 1);
    ctx->ncells_global = Nx * Ny;
    // owned block rows: faces whose id lies in face-rows [row_begin, row_end): bottom and vertical faces of the owned
    // cell rows.  The rows of the strip's bottom faces need the cell row below => one ghost row of cells.
    Strip& s = ctx->strip;
    s.cf_begin = compress_row_start(g, row_begin);
    s.cf_end = row_end == Ny ? num_free_faces(g) : compress_row_start(g, row_end);
    s.cell_row0 = row_begin > 0 ? row_begin - 1 : 0;
    s.col_lo = compress_row_start(g, s.cell_row0);
    s.col_hi = row_end == Ny ? num_free_faces(g) : compress_row_start(g, row_end) + (2 * Nx - 1);   // through the top faces h(.,row_end)
    if (s.col_hi > num_free_faces(g)) s.col_hi = num_free_faces(g);
    ctx->ncells_local = (row_end - s.cell_row0) * Nx;
    CK(dalloc(ctx, &ctx->points, (size_t)ctx->npoints));
    free_cut(ctx);
    ctx->nfaces = 2 * Nx * Ny + Nx + Ny;
    gen_points_kernel<<<148 * 4, 256, 0, ctx->stream>>>(g, ctx->points);
    CK(cudaGetLastError()); CK(cudaStreamSynchronize(ctx->stream));
    ctx->have_mesh = true; return PB_OK;
}

extern "C" pb_status pb_mesh_get_points(pb_ctx* ctx, double* xy) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->have_mesh && xy, PB_ERR_STATE, "pb_mesh_get_points: no mesh");
    CK(cudaSetDevice(ctx->device));
    CK(cudaMemcpy(xy, ctx->points, (size_t)ctx->npoints * 16, cudaMemcpyDeviceToHost)); return PB_OK;
}
extern "C" pb_status pb_mesh_set_points(pb_ctx* ctx, const double* xy) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->have_mesh && xy, PB_ERR_STATE, "pb_mesh_set_points: no mesh");
    DRAIN();
    CK(cudaSetDevice(ctx->device));
    CK(cudaMemcpy(ctx->points, xy, (size_t)ctx->npoints * 16, cudaMemcpyHostToDevice)); ctx->assembled = false; return PB_OK;
}
extern "C" pb_status pb_mesh_cell_faces(pb_ctx* ctx, long long* out) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->have_mesh && out, PB_ERR_STATE, "pb_mesh_cell_faces: no mesh");
    const Grid& g = ctx->grid;
    for (long long j = 0; j < g.Ny; j++) for (long long i = 0; i < g.Nx; i++) cell_faces(g, i, j, out + 4 * (j * g.Nx + i));
    return PB_OK;
}
extern "C" pb_status pb_mesh_compress_table(pb_ctx* ctx, long long* table) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->have_mesh && table, PB_ERR_STATE, "pb_mesh_compress_table: no mesh");
    const Grid& g = ctx->grid;
    long long nf = 2 * g.Nx * g.Ny + g.Nx + g.Ny;
    for (long long f = 0; f < nf; f++) table[f] = -1;
    for (long long j = 0; j < g.Ny; j++) for (long long i = 0; i < g.Nx; i++) {
        long long f[4]; cell_faces(g, i, j, f);
        for (int li = 0; li < 4; li++) if (!cell_face_dirichlet(g, i, j, li)) table[f[li]] = cell_face_compressed(g, i, j, li);
    }
    return PB_OK;
}
#include "pb_api_cut.inl"
static pb_status scatter_system(pb_ctx* ctx);
#include "pb_api_interface.inl"

// ------------------------------------------------------------------------------------------------ stage 1+2
template <int CD, int FD, int QK, int SK, int RQD, int GS, int TPB, int MINB, int MQ = QK>
static cudaError_t launch_local(pb_ctx* c, const LocalArgs& A_) {
    constexpr int QC = cell_qps(MQ, 2 * (FD + 1)), QR = cell_qps(QK, RQD), QCMAX = QC > QR ? QC : QR;
    using SM = LocalSmem<CD, FD, GS, QCMAX>;
    LocalArgs A = A_;
    if (A.ncells >= (1ll << 31)) return cudaErrorInvalidValue;      // K1 indexes local cells with 32 bits
    {   // moment -> stiffness table of this reconstruction degree
        constexpr int RD = FD + 1, M = cell_basis_size(RD) - 1;
        if (c->stiff_key != RD * 1000 + SM::MP) {
            std::vector<StiffEnt> tab((size_t)M * (M + 1) / 2);
            stiff_table<RD, SM::MP>(tab.data());
            cudaError_t e0 = dalloc(c, &c->stiff_tab, tab.size());
            if (e0 == cudaSuccess) e0 = cudaMemcpyAsync(c->stiff_tab, tab.data(), tab.size() * sizeof(StiffEnt), cudaMemcpyHostToDevice, c->stream);
            if (e0 == cudaSuccess) e0 = cudaStreamSynchronize(c->stream);
            if (e0 != cudaSuccess) return e0;
            c->stiff_key = RD * 1000 + SM::MP;
        }
        A.stiff = c->stiff_tab;
    }
    const size_t smem = local_ref_bytes<CD, FD, QK, RQD, MQ>() + sizeof(SM) * (TPB / GS);
    auto kern = local_uncut_kernel<CD, FD, QK, SK, RQD, GS, TPB, MINB, MQ>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int bps = 0, nsm = 0;
    if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, kern, TPB, smem)) != cudaSuccess) return e;
    cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, c->device);
    if (bps < 1) return cudaErrorLaunchOutOfResources;
    long long need = (A.ncells + (TPB / GS) - 1) / (TPB / GS), cap = (long long)nsm * bps;
    int grid = (int)(need < cap ? need : cap);
    kern<<<grid, TPB, smem, c->stream>>>(A);
    return cudaGetLastError();
}

// (cell degree, face degree, quadrature kind, stabilisation, load quadrature degree) -> kernel instance
static cudaError_t dispatch_local(pb_ctx* c, const LocalArgs& A, int cd, int fd, int qk, int sk, int rqd, bool* found) {
    *found = true;
#define PB_CASE(CD_, FD_, QK_, SK_, RQD_, GS_, TPB_, MINB_) \
    if (cd == CD_ && fd == FD_ && qk == QK_ && sk == SK_ && rqd == RQD_) return launch_local<CD_, FD_, QK_, SK_, RQD_, GS_, TPB_, MINB_>(c, A);
    {   // tuning hook: alternative lane mappings (developer experiments, PB_K1_GS=<id>)
        const char* e = getenv("PB_K1_GS");
        const int v = e ? atoi(e) : 0;
        if (v && qk == 0 && sk == 1) {
#ifndef PB_FAST_BUILD
            if (cd == 2 && fd == 1 && rqd == 4) {
                if (v == 1) return launch_local<2, 1, 0, 1, 4, 4, 32, 10>(c, A);
                if (v == 2) return launch_local<2, 1, 0, 1, 4, 4, 64, 6>(c, A);
                if (v == 3) return launch_local<2, 1, 0, 1, 4, 4, 32, 14>(c, A);
            }
            if (cd == 1 && fd == 0 && rqd == 2) {
                if (v == 1) return launch_local<1, 0, 0, 1, 2, 2, 32, 16>(c, A);
                if (v == 2) return launch_local<1, 0, 0, 1, 2, 2, 64, 8>(c, A);
                if (v == 3) return launch_local<1, 0, 0, 1, 2, 4, 32, 24>(c, A);
            }
#endif
        }
    }
#ifdef PB_FAST_BUILD   // developer builds: only the headline instance
    PB_CASE(3, 2, 0, 1, 6, 8, 32, 8)
#else
    // HHO(k+1,k), tensor Gauss (convergence_test): fancy and naive
    PB_CASE(1, 0, 0, 1, 2, 4, 32, 16)  PB_CASE(1, 0, 0, 0, 2, 4, 32, 16)
    PB_CASE(2, 1, 0, 1, 4, 4, 32, 12) PB_CASE(2, 1, 0, 0, 4, 4, 32, 12)
    PB_CASE(3, 2, 0, 1, 6, 8, 32, 8) PB_CASE(3, 2, 0, 0, 6, 8, 32, 8)
    PB_CASE(4, 3, 0, 1, 8, 16, 32, 5) PB_CASE(4, 3, 0, 0, 8, 16, 32, 5)
    // HHO(k+1,k), Dunavant fan (uncut cells of cuthho_square): naive.  (Tried in round 2: moments from the 16 tensor Gauss points, which are
    // exact for them, and the fan only for the load of the cells inside the domain — template parameter MQ of local_uncut_kernel.  Same
    // matrices to rounding, all tests green, but SLOWER: 73.1 against 63.4 ms at 4096^2, k = 2 (profiles/r2p_fan_moments.txt): the second
    // quadrature loop lengthens an instruction stream that is already the bottleneck.  The instances are not built.)
    PB_CASE(1, 0, 1, 0, 2, 4, 32, 16)
    PB_CASE(2, 1, 1, 0, 4, 4, 32, 12)
    PB_CASE(3, 2, 1, 0, 6, 8, 32, 8)
#endif
#undef PB_CASE
    *found = false; return cudaSuccess;
}

static pb_status build_structure(pb_ctx* ctx) {
    const long long nbr = ctx->strip.cf_end - ctx->strip.cf_begin;
    const int F = ctx->F;
    ctx->nrows_blk = nbr; ctx->nrows = nbr * F;
    {   // owned / referenced block rows and what the neighbours need of them (see halo_exchange)
        const Grid& g = ctx->grid; const Strip& s = ctx->strip;
        ctx->rb_begin = s.cf_begin; ctx->rb_end = s.cf_end; ctx->rb_lo = s.col_lo; ctx->rb_hi = s.col_hi;
        ctx->halo_prev = std::min(s.cf_begin + (2 * g.Nx - 1), num_free_faces(g)) - s.cf_begin;
        ctx->halo_next = s.cf_end - compress_row_start(g, g.row_end - 1);
    }
    int* nb = nullptr;
    CK(dalloc(ctx, &ctx->brows, (size_t)nbr)); CK(dalloc(ctx, &ctx->brp, (size_t)nbr + 1));
    CK(cudaMalloc((void**)&nb, (size_t)(nbr + 1) * 4));
    CK(cudaMemsetAsync(nb, 0, (size_t)(nbr + 1) * 4, ctx->stream));
    build_block_rows_kernel<<<(unsigned)((nbr + 255) / 256), 256, 0, ctx->stream>>>(ctx->grid, ctx->strip, ctx->brows, nb);
    CK(cudaGetLastError());
    void* tmp = nullptr; size_t tmpb = 0;
    // exclusive scan of nb (int) into brp (long long) over nbr+1 items (the extra 0 yields the total)
    CK(cub::DeviceScan::ExclusiveScan(nullptr, tmpb, nb, ctx->brp, cuda::std::plus<>(), 0LL, (int)(nbr + 1), ctx->stream));
    CK(cudaMalloc(&tmp, tmpb));
    CK(cub::DeviceScan::ExclusiveScan(tmp, tmpb, nb, ctx->brp, cuda::std::plus<>(), 0LL, (int)(nbr + 1), ctx->stream));
    long long nblk = 0;
    CK(cudaMemcpyAsync(&nblk, ctx->brp + nbr, 8, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    cudaFree(tmp); cudaFree(nb);
    ctx->nnz = nblk * F * F;
    CK(dalloc(ctx, &ctx->rowptr, (size_t)ctx->nrows + 1)); CK(dalloc(ctx, &ctx->col, (size_t)ctx->nnz));
    CK(dalloc(ctx, &ctx->val, (size_t)ctx->nnz)); CK(dalloc(ctx, &ctx->rhs, (size_t)ctx->nrows));
    fill_csr_pattern_kernel<<<(unsigned)((ctx->nrows + 1 + 255) / 256), 256, 0, ctx->stream>>>(nbr, F, ctx->brows, ctx->brp, ctx->rowptr, ctx->col);
    CK(cudaGetLastError());
    return PB_OK;
}

template <int CD, int FD, int QK, int SK>
static cudaError_t launch_generic_t(pb_ctx* c, const LocalArgs& A, int rhs_di) {
    const size_t smem = sizeof(GenSmem<CD, FD>);
    auto kern = local_generic_kernel<CD, FD, QK, SK>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    long long grid = A.ncells < 148 * 32 ? A.ncells : 148 * 32;
    kern<<<(unsigned)grid, 32, smem, c->stream>>>(A, rhs_di);
    return cudaGetLastError();
}
// every valid hho_degree_info that the tuned K1 has no instance for: equal order (k,k) and the obstacle's HHO(0,k)
static pb_status launch_generic(pb_ctx* ctx, const LocalArgs& A, int cd, int fd, int qk, int sk, int rhs_di, bool* found) {
    *found = true; cudaError_t e = cudaSuccess;
#define PB_GEN(CD_, FD_) \
    if (cd == CD_ && fd == FD_) { \
        if (qk == 0 && sk == 1) e = launch_generic_t<CD_, FD_, 0, 1>(ctx, A, rhs_di); else if (qk == 0) e = launch_generic_t<CD_, FD_, 0, 0>(ctx, A, rhs_di); \
        else if (sk == 1) e = launch_generic_t<CD_, FD_, 1, 1>(ctx, A, rhs_di); else e = launch_generic_t<CD_, FD_, 1, 0>(ctx, A, rhs_di); \
        if (e != cudaSuccess) return fail(ctx, PB_ERR_CUDA, "local_generic_kernel launch failed: %s", cudaGetErrorString(e)); \
        return PB_OK; }
    PB_GEN(0, 0) PB_GEN(0, 1) PB_GEN(1, 1) PB_GEN(2, 2) PB_GEN(1, 2)
#ifndef PB_FAST_BUILD
    PB_GEN(1, 0) PB_GEN(2, 1) PB_GEN(3, 2) PB_GEN(4, 3)      // HHO(k+1,k) through the literal kernel: keep_local = 2 (per-cell views of the C++ shim)
#endif
#undef PB_GEN
    *found = false; return PB_OK;
}

template <int F>
static pb_status launch_scatter_tile(pb_ctx* ctx) {
    constexpr int T = 32;
    const size_t smem = sizeof(ScatterTileSmem<F, T>);
    auto kern = scatter_tile_kernel<F, T>;
    static int bps = 0;
    if (!bps) {
        CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, kern, 256, smem));
        if (bps < 1) return fail(ctx, PB_ERR_CUDA, "scatter_tile_kernel does not fit on an SM");
    }
    int nsm = 148; cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, ctx->device);
    const long long ntiles = (ctx->nrows_blk + T - 1) / T;
    // several waves: tiles are short.  The wave count hardly matters (2.05 ms with one wave, 2.02 with four, 1.99 with eight at 2048^2, k = 2:
    // profiles/r3h_k3_waves.txt) — in particular a single persistent wave, which would keep the two uses of a cell's block (the face rows below
    // and above it) close in time, does not lower the time: the L2 miss rate is not what limits K3.  PB_K3_WAVES overrides.
    static const int waves = getenv("PB_K3_WAVES") ? std::max(1, atoi(getenv("PB_K3_WAVES"))) : 8;
    const long long grid = std::min<long long>(ntiles, (long long)nsm * bps * waves);
    kern<<<(unsigned)(grid < 1 ? 1 : grid), 256, smem, ctx->stream>>>(ctx->nrows_blk, ctx->brows, ctx->brp, ctx->S, ctx->g, ctx->val, ctx->rhs);
    return PB_OK;
}

// K3: condensed blocks -> CSR values and rhs
static pb_status scatter_system(pb_ctx* ctx) {
    const int F = ctx->F;
#define PB_SC(F_) scatter_kernel_t<F_, 1><<<148 * 32, 256, 0, ctx->stream>>>(ctx->nrows_blk, ctx->brows, ctx->brp, ctx->S, ctx->g, ctx->val, ctx->rhs)
    if (F == 1) scatter_packed_kernel_t<1, 1><<<148 * 32, 256, 0, ctx->stream>>>(ctx->nrows_blk, ctx->brows, ctx->brp, ctx->S, ctx->g, ctx->val, ctx->rhs);
    else if (F == 2) scatter_packed_kernel_t<2, 1><<<148 * 32, 256, 0, ctx->stream>>>(ctx->nrows_blk, ctx->brows, ctx->brp, ctx->S, ctx->g, ctx->val, ctx->rhs);
    // F = 3: tile kernel (2.02 against 2.58 ms at 2048^2, bit-identical values); F = 4 keeps the one-warp-per-row kernel, whose 28-lane
    // rows already fill a warp (0.82 against 0.87 ms at 1024^2).  PB_SCATTER_ROWS=1 / PB_SCATTER_TILE=1 force one or the other.
    else if (F == 3 && !getenv("PB_SCATTER_ROWS")) { pb_status st = launch_scatter_tile<3>(ctx); if (st != PB_OK) return st; }
    else if (F == 4 && getenv("PB_SCATTER_TILE")) { pb_status st = launch_scatter_tile<4>(ctx); if (st != PB_OK) return st; }
    else if (F == 3) PB_SC(3); else if (F == 4) PB_SC(4);
    else {      // face degree > 3 (generic local kernel)
        scatter_values_kernel<<<148 * 16, 256, 0, ctx->stream>>>(ctx->nrows_blk, F, ctx->brows, ctx->brp, ctx->S, ctx->val);
        CK(cudaGetLastError());
        scatter_rhs_kernel<<<(unsigned)((ctx->nrows + 255) / 256), 256, 0, ctx->stream>>>(ctx->nrows_blk, F, ctx->brows, ctx->g, ctx->rhs);
    }
#undef PB_SC
    CK(cudaGetLastError());
    return PB_OK;
}

// stages 1+2 of one step.  host_points / host_rhs (pinned, may be null) turn the step into a pipelined one: the
// vertex upload runs on its own stream as soon as the previous step's local-operator kernel has released the vertex
// array, the read-back of the condensed rhs on another one after K3, and the next step's K3 waits for that read-back.
static pb_status assemble_step(pb_ctx* ctx, const pb_assembly_params* p, const double* host_points, double* host_rhs, bool sync) {
    NEED(ctx->have_mesh && p, PB_ERR_STATE, "pb_assemble: no mesh");
    CK(cudaSetDevice(ctx->device));
    int cd = p->cell_degree, fd = p->face_degree;
    NEED(cd >= 0 && fd >= 0, PB_ERR_INVALID, "pb_assemble: negative degree");
    {   // hho_degree_info(cd, fd): invalid pairs revert to equal order (utils.hpp:75-95)
        bool c1 = fd > 0 && (cd == fd - 1 || cd == fd || cd == fd + 1), c2 = fd == 0 && (cd == fd || cd == fd + 1);
        if (!(c1 || c2)) cd = fd;
    }
    NEED(p->problem == PB_PROBLEM_HHO || p->problem == PB_PROBLEM_FICTDOM || p->problem == PB_PROBLEM_INTERFACE, PB_ERR_INVALID, "pb_assemble: bad problem kind");
    const bool ifc = p->problem == PB_PROBLEM_INTERFACE;
    if (ifc) {
        NEED(ctx->have_cut && ctx->have_ls, PB_ERR_STATE, "pb_assemble: the interface problem needs pb_cut_levelset and pb_cut_preprocess / pb_cut_upload first");
        NEED(ctx->quad_kind == PB_QUAD_POLY_DUNAVANT, PB_ERR_INVALID, "pb_assemble: the interface problem runs on the cuthho_poly mesh (PB_QUAD_POLY_DUNAVANT)");
        NEED(p->stab_kind != PB_STAB_FANCY, PB_ERR_INVALID, "pb_assemble: the interface driver uses the naive / cut stabilisation");
    }
    if (p->problem == PB_PROBLEM_FICTDOM) NEED(ctx->have_cut && ctx->have_ls, PB_ERR_STATE, "pb_assemble: fictitious-domain problem needs pb_cut_levelset and pb_cut_preprocess / pb_cut_upload first");
    int sk = p->stab_kind == PB_STAB_FANCY ? 1 : 0;
    const int rqd = 2 * (cd + p->rhs_degree_increase);
    bool relayout = !ctx->S || ctx->CD != cd || ctx->FD != fd || (p->keep_local && !ctx->lc) || (p->keep_local == 2 && !ctx->data) || ctx->general_csr != ifc;
    ctx->prm = *p; ctx->CD = cd; ctx->FD = fd;
    ctx->C = cell_basis_size(cd); ctx->F = fd + 1; ctx->R = cell_basis_size(fd + 1); ctx->NF = 4 * ctx->F; ctx->N = ctx->C + ctx->NF;
    const long long nc = ctx->ncells_local;
    if (relayout) {
        free_system(ctx);
        CK(dalloc(ctx, &ctx->S, (size_t)nc * ctx->NF * ctx->NF)); CK(dalloc(ctx, &ctx->g, (size_t)nc * ctx->NF));
        CK(dalloc(ctx, &ctx->recA, (size_t)nc * ctx->NF * ctx->C)); CK(dalloc(ctx, &ctx->recb, (size_t)nc * ctx->C));
        if (p->keep_local) {
            CK(dalloc(ctx, &ctx->lc, (size_t)nc * ctx->N * ctx->N)); CK(dalloc(ctx, &ctx->oper, (size_t)nc * ctx->N * ctx->R));
            CK(dalloc(ctx, &ctx->rhsloc, (size_t)nc * ctx->C));
            if (p->keep_local == 2) CK(dalloc(ctx, &ctx->data, (size_t)nc * ctx->N * ctx->N));
        }
        if (!ifc) { pb_status st = build_structure(ctx); if (st != PB_OK) return st; }
        ctx->general_csr = ifc;
    }
    LocalArgs A{};
    A.grid = ctx->grid; A.grid.row_begin = ctx->strip.cell_row0;      // local cell 0 sits in row cell_row0
    A.points = ctx->points; A.cell_loc = (p->problem == PB_PROBLEM_FICTDOM || ifc) ? ctx->cell_loc_g + ctx->strip.cell_row0 * ctx->grid.Nx : nullptr;
    A.ncells = nc; A.fun = p->fun; A.problem = p->problem; A.eta = p->eta;
    A.S = ctx->S; A.g = ctx->g; A.recA = ctx->recA; A.recb = ctx->recb;
    A.lc = p->keep_local ? ctx->lc : nullptr; A.oper = p->keep_local ? ctx->oper : nullptr; A.rhs = p->keep_local ? ctx->rhsloc : nullptr;
    A.data = p->keep_local == 2 ? ctx->data : nullptr;
    NEED(!(ifc && p->keep_local == 2), PB_ERR_UNSUPPORTED, "pb_assemble: keep_local = 2 (gr.second apart from the stabilisation) is built for the plain and fictitious-domain problems");
    if (host_points) {
        if (!ctx->s_in) { CK(cudaStreamCreateWithFlags(&ctx->s_in, cudaStreamNonBlocking)); CK(cudaEventCreateWithFlags(&ctx->ev_in, cudaEventDisableTiming)); }
        if (ctx->pipe_in_flight) CK(cudaStreamWaitEvent(ctx->s_in, ctx->ev[1], 0));      // previous K1 has read the vertices
        // host_points is the whole-mesh array; a strip context reads only the vertex rows of its own cells
        const long long p0 = ctx->strip.cell_row0 * (ctx->grid.Nx + 1), pn = (ctx->grid.row_end - ctx->strip.cell_row0 + 1) * (ctx->grid.Nx + 1);
        CK(cudaMemcpyAsync(ctx->points + p0, host_points + 2 * p0, (size_t)pn * 16, cudaMemcpyHostToDevice, ctx->s_in));
        CK(cudaEventRecord(ctx->ev_in, ctx->s_in));
        CK(cudaStreamWaitEvent(ctx->stream, ctx->ev_in, 0));
    }
    CK(cudaEventRecord(ctx->ev[0], ctx->stream));
    bool found = false;
    // keep_local = 2 asks for gr.second apart from the stabilisation: only the literal kernels hold the two separately
    if (p->keep_local != 2) CK(dispatch_local(ctx, A, cd, fd, ctx->quad_kind, sk, rqd, &found));
    if (!found) { pb_status st = launch_generic(ctx, A, cd, fd, ctx->quad_kind, sk, p->rhs_degree_increase, &found); if (st != PB_OK) return st; }
    if (!found) return fail(ctx, PB_ERR_UNSUPPORTED, "pb_assemble: no kernel instance for HHO(%d,%d) quad_kind=%d stab=%d load degree %d", cd, fd, ctx->quad_kind, p->stab_kind, rqd);
    if (p->problem == PB_PROBLEM_FICTDOM && ctx->ncut > 0) {
        pb_status st = launch_cut(ctx, A, cd, fd); if (st != PB_OK) return st;
    }
    if (ifc) { pb_status st = launch_interface_cut(ctx, p, cd, fd); if (st != PB_OK) return st; }
    CK(cudaEventRecord(ctx->ev[1], ctx->stream));
    if (ctx->out_pending) CK(cudaStreamWaitEvent(ctx->stream, ctx->ev_out, 0));         // previous rhs read-back done before K3 overwrites rhs
    if (ifc) {
        if (ctx->out_pending) { CK(cudaStreamSynchronize(ctx->s_out)); ctx->out_pending = false; }   // the system arrays are reallocated below
        pb_status st = build_interface_system(ctx); if (st != PB_OK) return st;
    } else { pb_status st = scatter_system(ctx); if (st != PB_OK) return st; }
    CK(cudaEventRecord(ctx->ev[2], ctx->stream));
    if (host_rhs) {
        if (!ctx->s_out) { CK(cudaStreamCreateWithFlags(&ctx->s_out, cudaStreamNonBlocking)); CK(cudaEventCreateWithFlags(&ctx->ev_out, cudaEventDisableTiming)); }
        CK(cudaStreamWaitEvent(ctx->s_out, ctx->ev[2], 0));
        CK(cudaMemcpyAsync(host_rhs, ctx->rhs, (size_t)ctx->nrows * 8, cudaMemcpyDeviceToHost, ctx->s_out));
        CK(cudaEventRecord(ctx->ev_out, ctx->s_out));
        ctx->out_pending = true;
    }
    ctx->pipe_in_flight = !sync; ctx->step_timed = true;
    ctx->assembled = true; ctx->asm_gen++;
    if (sync) return pb_pipeline_sync(ctx);
    return PB_OK;
}
extern "C" pb_status pb_strip_point_range(pb_ctx* ctx, long long* first, long long* count) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->have_mesh && first && count, PB_ERR_STATE, "pb_strip_point_range: no mesh");
    *first = ctx->strip.cell_row0 * (ctx->grid.Nx + 1); *count = (ctx->grid.row_end - ctx->strip.cell_row0 + 1) * (ctx->grid.Nx + 1);
    return PB_OK;
}
extern "C" pb_status pb_pipeline_sync(pb_ctx* ctx) {
    if (!ctx) return PB_ERR_INVALID;
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    if (ctx->s_out) CK(cudaStreamSynchronize(ctx->s_out));
    ctx->pipe_in_flight = false; ctx->out_pending = false;
    if (ctx->step_timed) {
        ctx->step_timed = false;
        float ms = 0; cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]); ctx->ms_local = ms;
        cudaEventElapsedTime(&ms, ctx->ev[1], ctx->ev[2]); ctx->ms_scatter = ms;
    }
    return PB_OK;
}
extern "C" pb_status pb_assemble(pb_ctx* ctx, const pb_assembly_params* p) {
    if (!ctx) return PB_ERR_INVALID;
    if (ctx->pipe_in_flight) { pb_status st = pb_pipeline_sync(ctx); if (st != PB_OK) return st; }
    return assemble_step(ctx, p, nullptr, nullptr, true);
}
extern "C" pb_status pb_assemble_pipelined(pb_ctx* ctx, const pb_assembly_params* p, const double* host_points, double* host_rhs) {
    if (!ctx) return PB_ERR_INVALID;
    return assemble_step(ctx, p, host_points, host_rhs, false);
}

extern "C" pb_status pb_system_sizes(pb_ctx* ctx, long long* n_rows, long long* nnz, long long* n_cells, long long* n_local) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->assembled, PB_ERR_STATE, "pb_system_sizes: nothing assembled");
    if (n_rows) *n_rows = ctx->nrows; if (nnz) *nnz = ctx->nnz; if (n_cells) *n_cells = ctx->ncells_local; if (n_local) *n_local = ctx->N;
    return PB_OK;
}
extern "C" pb_status pb_get_csr(pb_ctx* ctx, long long* rowptr, int* col, double* val, double* rhs) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->assembled, PB_ERR_STATE, "pb_get_csr: nothing assembled");
    DRAIN();
    CK(cudaSetDevice(ctx->device));
    if (rowptr) CK(cudaMemcpy(rowptr, ctx->rowptr, (size_t)(ctx->nrows + 1) * 8, cudaMemcpyDeviceToHost));
    if (col) CK(cudaMemcpy(col, ctx->col, (size_t)ctx->nnz * 4, cudaMemcpyDeviceToHost));
    if (val) CK(cudaMemcpy(val, ctx->val, (size_t)ctx->nnz * 8, cudaMemcpyDeviceToHost));
    if (rhs) CK(cudaMemcpy(rhs, ctx->rhs, (size_t)ctx->nrows * 8, cudaMemcpyDeviceToHost));
    return PB_OK;
}
extern "C" pb_status pb_get_local(pb_ctx* ctx, int what, long long cell_begin, long long count, double* out) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->assembled && out, PB_ERR_STATE, "pb_get_local: nothing assembled");
    DRAIN();
    NEED(cell_begin >= 0 && count >= 0 && cell_begin + count <= ctx->ncells_local, PB_ERR_INVALID, "pb_get_local: cell range out of bounds");
    const double* src = nullptr; long long per = 0;
    switch (what) {
    case 0: src = ctx->lc; per = (long long)ctx->N * ctx->N; break;
    case 1: src = ctx->oper; per = (long long)ctx->N * ctx->R; break;
    case 2: src = ctx->rhsloc; per = ctx->C; break;
    case 3: src = ctx->S; per = (long long)ctx->NF * ctx->NF; break;
    case 4: src = ctx->g; per = ctx->NF; break;
    case 5: src = ctx->recA; per = (long long)ctx->NF * ctx->C; break;
    case 6: src = ctx->recb; per = ctx->C; break;
    case 7: src = ctx->data; per = (long long)ctx->N * ctx->N; break;
    default: return fail(ctx, PB_ERR_INVALID, "pb_get_local: unknown selector %d", what);
    }
    NEED(src != nullptr, PB_ERR_STATE, "pb_get_local: assemble with keep_local = 1 to read lc / oper / rhs, keep_local = 2 to read gr.second (selector 7)");
    CK(cudaSetDevice(ctx->device));
    CK(cudaMemcpy(out, src + cell_begin * per, (size_t)(count * per) * 8, cudaMemcpyDeviceToHost));
    return PB_OK;
}

// ------------------------------------------------------------------------------------------------ stage 3
static pb_status ensure_solver(pb_ctx* ctx);      // pb_api_dist.inl
extern "C" pb_status pb_get_face_solution(pb_ctx* ctx, double* x) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->assembled && ctx->x && x && ctx->solver_n == ctx->nrows, PB_ERR_STATE, "pb_get_face_solution: no solution");
    CK(cudaSetDevice(ctx->device));
    CK(cudaMemcpy(x, ctx->x, (size_t)ctx->nrows * 8, cudaMemcpyDeviceToHost)); return PB_OK;
}
extern "C" pb_status pb_set_face_solution(pb_ctx* ctx, const double* x) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->assembled && x, PB_ERR_STATE, "pb_set_face_solution: nothing assembled");
    CK(cudaSetDevice(ctx->device));
    { pb_status st = ensure_solver(ctx); if (st != PB_OK) return st; }
    CK(cudaMemcpy(ctx->x, x, (size_t)ctx->nrows * 8, cudaMemcpyHostToDevice)); return PB_OK;
}

// K6 — back-substitution u_T = K_TT^{-1} f_T - (K_TT^{-1} K_TF) u_F with the Dirichlet face values re-projected
// (take_local_data, hho.hpp:408-449).  One thread per cell.
template <int FD>
__global__ void recover_kernel(Grid g, long long cell_row0, long long ncells, long long row_first_owned, int C, int fun, const double2* __restrict__ points,
                               const double* __restrict__ recA, const double* __restrict__ recb, const double* __restrict__ xf, long long xoff,
                               double* __restrict__ ucell, double* __restrict__ ulocal) {
    constexpr int F = FD + 1, NF = 4 * F, QF = FD + 1;
    for (long long c = blockIdx.x * (long long)blockDim.x + threadIdx.x; c < ncells; c += (long long)gridDim.x * blockDim.x) {
        long long ci = c % g.Nx, cj = cell_row0 + c / g.Nx;
        if (cj < row_first_owned) continue;                 // ghost row: recovered by its owner
        double uF[NF];
        CellGeom geo;
        long long p0 = grid_point(g, ci, cj);
        double2 a = points[p0], b = points[p0 + 1], cc = points[p0 + g.Nx + 2], d = points[p0 + g.Nx + 1];
        geo.vx[0] = a.x; geo.vy[0] = a.y; geo.vx[1] = b.x; geo.vy[1] = b.y; geo.vx[2] = cc.x; geo.vy[2] = cc.y; geo.vx[3] = d.x; geo.vy[3] = d.y;
#pragma unroll
        for (int li = 0; li < 4; li++) {
            if (!cell_face_dirichlet(g, ci, cj, li)) {
                long long cf = cell_face_compressed(g, ci, cj, li);
#pragma unroll
                for (int k = 0; k < F; k++) uF[li * F + k] = xf[cf * F + k - xoff];
            } else {
                double ax, ay, bx, by; face_points(geo, li, ax, ay, bx, by);
                double len = sqrt((bx - ax) * (bx - ax) + (by - ay) * (by - ay)), fbx = 0.5 * (ax + bx), fby = 0.5 * (ay + by);
                double Mf[F * (F + 1) / 2], gd[F];
#pragma unroll
                for (int k = 0; k < F * (F + 1) / 2; k++) Mf[k] = 0.0;
#pragma unroll
                for (int k = 0; k < F; k++) gd[k] = 0.0;
#pragma unroll
                for (int q = 0; q < QF; q++) {
                    double t = c_qt.gx[QF][q], px = 0.5 * (1 - t) * ax + 0.5 * (1 + t) * bx, py = 0.5 * (1 - t) * ay + 0.5 * (1 + t) * by;
                    double w = c_qt.gw[QF][q] * len * 0.5;
                    double ep = 4.0 * ((fbx - ax) * (px - fbx) + (fby - ay) * (py - fby)) / (len * len);
                    double ph[F]; ph[0] = 1.0;
#pragma unroll
                    for (int k = 1; k < F; k++) ph[k] = ph[k - 1] * ep;
                    double bc = fun_sol(fun, px, py);
#pragma unroll
                    for (int k = 0; k < F; k++) {
                        gd[k] += w * ph[k] * bc;
#pragma unroll
                        for (int m = 0; m <= k; m++) Mf[PB_TRI(k, m)] += w * ph[k] * ph[m];
                    }
                }
                ldlt_packed<F>(Mf); ldlt_solve<F>(Mf, gd);
#pragma unroll
                for (int k = 0; k < F; k++) uF[li * F + k] = gd[k];
            }
        }
        long long gc = cj * g.Nx + ci;
        for (int k = 0; k < C; k++) {
            double v = recb[c * C + k];
#pragma unroll
            for (int i = 0; i < NF; i++) v -= recA[(c * NF + i) * C + k] * uF[i];
            ucell[gc * C + k] = v;
            if (ulocal) ulocal[(gc - row_first_owned * g.Nx) * (C + NF) + k] = v;
        }
        if (ulocal)       // take_local_data (hho.hpp:408-449): cell dofs, then the face dofs with the Dirichlet faces projected
            for (int i = 0; i < NF; i++) ulocal[(gc - row_first_owned * g.Nx) * (C + NF) + C + i] = uF[i];
    }
}

static pb_status interface_recover(pb_ctx* ctx, double* full);
static pb_status interface_error(pb_ctx* ctx, double out[3]);
static pb_status interface_sample(pb_ctx* ctx, double* xyv, long long capacity_points, long long* n_points);
#include "pb_api_dist.inl"
#include "pb_api_mg.inl"
#include "pb_api_interface_post.inl"
#include "pb_api_obstacle.inl"
#include "pb_api_static.inl"

extern "C" pb_status pb_last_timings(pb_ctx* ctx, double* ms_local, double* ms_scatter, double* ms_pcg) {
    if (!ctx) return PB_ERR_INVALID;
    if (ms_local) *ms_local = ctx->ms_local; if (ms_scatter) *ms_scatter = ctx->ms_scatter; if (ms_pcg) *ms_pcg = ctx->ms_pcg;
    return PB_OK;
}

// register-resident DFMA chains: 8 independent accumulators per thread, 2 flop per FMA
__global__ void fp64_probe_kernel(double* out, int iters, double a, double b) {
    double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    for (int i = 0; i < iters; i++) {
        x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
        x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
    }
    out[blockIdx.x * (size_t)blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
}
extern "C" pb_status pb_probe_fp64_peak(pb_ctx* ctx, double* tflops) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(tflops, PB_ERR_INVALID, "pb_probe_fp64_peak: null output");
    CK(cudaSetDevice(ctx->device));
    int nsm = 148; cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, ctx->device);
    const int blocks = nsm * 8, threads = 256, iters = 1 << 16;
    double* out = nullptr; CK(cudaMalloc((void**)&out, (size_t)blocks * threads * 8));
    double best = 0;
    for (int rep = 0; rep < 4; rep++) {
        CK(cudaEventRecord(ctx->ev[0], ctx->stream));
        fp64_probe_kernel<<<blocks, threads, 0, ctx->stream>>>(out, iters, 0.999999, 1e-7);
        CK(cudaEventRecord(ctx->ev[1], ctx->stream)); CK(cudaStreamSynchronize(ctx->stream));
        float ms = 0; cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]);
        double tf = 2.0 * 8 * (double)iters * blocks * threads / (ms * 1e-3) / 1e12;
        if (rep > 0 && tf > best) best = tf;
    }
    cudaFree(out); *tflops = best; return PB_OK;
}
