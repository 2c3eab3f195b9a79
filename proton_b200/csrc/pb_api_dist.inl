// C ABI, solver part (included by pb_api.cu): strip layout, NCCL plumbing (dlopen), the PCG driver for one rank or a
// row of ranks, back-substitution.
#include <dlfcn.h>

// ---- minimal NCCL surface, resolved at run time: the process that drives several ranks has already loaded NCCL
// (torch.distributed), a single-rank user never needs it
namespace nccl_dl {
typedef struct { char internal[128]; } UniqueId;
typedef void* Comm;
enum { Sum = 0, Float64 = 8 };
struct Api {
    void* lib = nullptr;
    int (*GetUniqueId)(UniqueId*) = nullptr;
    int (*CommInitRank)(Comm*, int, UniqueId, int) = nullptr;
    int (*CommDestroy)(Comm) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, Comm, cudaStream_t) = nullptr;
    int (*Send)(const void*, size_t, int, int, Comm, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, Comm, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};
static Api g;
static bool load(std::string& why) {
    if (g.lib) return true;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) { g.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (g.lib) break; }
    if (!g.lib) { why = std::string("cannot load NCCL: ") + dlerror(); return false; }
#define PB_SYM(field, name) *(void**)(&g.field) = dlsym(g.lib, name); if (!g.field) { why = std::string("NCCL symbol missing: ") + name; g.lib = nullptr; return false; }
    PB_SYM(GetUniqueId, "ncclGetUniqueId") PB_SYM(CommInitRank, "ncclCommInitRank") PB_SYM(CommDestroy, "ncclCommDestroy")
    PB_SYM(AllReduce, "ncclAllReduce") PB_SYM(Send, "ncclSend") PB_SYM(Recv, "ncclRecv") PB_SYM(GroupStart, "ncclGroupStart")
    PB_SYM(GroupEnd, "ncclGroupEnd") PB_SYM(GetErrorString, "ncclGetErrorString")
#undef PB_SYM
    return true;
}
}  // namespace nccl_dl
#define NCK(call) do { int r_ = (call); if (r_ != 0) return fail(ctx, PB_ERR_CUDA, "%s failed: %s", #call, nccl_dl::g.GetErrorString(r_)); } while (0)

static void dist_release(pb_ctx* c) { if (c->comm && nccl_dl::g.CommDestroy) nccl_dl::g.CommDestroy(c->comm); c->comm = nullptr; }

extern "C" pb_status pb_strip_layout(long long Nx, long long Ny, long long row_begin, long long row_end, int fbs, long long out[8]) {
    if (!out || Nx < 2 || Ny < 2 || fbs < 1 || row_begin < 0 || row_begin >= row_end || row_end > Ny) return PB_ERR_INVALID;
    Grid g{}; g.Nx = Nx; g.Ny = Ny;
    const long long nfree = num_free_faces(g);
    const long long cf_begin = compress_row_start(g, row_begin), cf_end = row_end == Ny ? nfree : compress_row_start(g, row_end);
    const long long cell_row0 = row_begin > 0 ? row_begin - 1 : 0;
    const long long col_lo = compress_row_start(g, cell_row0);
    long long col_hi = row_end == Ny ? nfree : compress_row_start(g, row_end) + (2 * Nx - 1);
    if (col_hi > nfree) col_hi = nfree;
    out[0] = cf_begin; out[1] = cf_end; out[2] = col_lo; out[3] = col_hi; out[4] = cell_row0;
    out[5] = (cf_end - cf_begin) * fbs; out[6] = (cf_begin - col_lo) * fbs; out[7] = (col_hi - cf_end) * fbs;
    return PB_OK;
}

extern "C" pb_status pb_dist_unique_id(unsigned char id[128]) {
    pb_ctx* ctx = nullptr;
    std::string why;
    if (!id) return fail(nullptr, PB_ERR_INVALID, "pb_dist_unique_id: null output");
    if (!nccl_dl::load(why)) return fail(nullptr, PB_ERR_CUDA, "pb_dist_unique_id: %s", why.c_str());
    nccl_dl::UniqueId u;
    NCK(nccl_dl::g.GetUniqueId(&u));
    memcpy(id, u.internal, 128);
    return PB_OK;
}

extern "C" pb_status pb_dist_init(pb_ctx* ctx, int rank, int world, const unsigned char id[128]) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(world >= 1 && rank >= 0 && rank < world && id, PB_ERR_INVALID, "pb_dist_init: bad rank / world / id");
    CK(cudaSetDevice(ctx->device));
    ctx->rank = rank; ctx->world = world;
    if (world == 1) return PB_OK;
    std::string why;
    if (!nccl_dl::load(why)) return fail(ctx, PB_ERR_CUDA, "pb_dist_init: %s", why.c_str());
    nccl_dl::UniqueId u; memcpy(u.internal, id, 128);
    NCK(nccl_dl::g.CommInitRank(&ctx->comm, world, u, rank));
    return PB_OK;
}

// ---- halo exchange of a face vector stored for the columns [col_lo, col_hi): `full` points at col_lo, the owned part
// starts `down` doubles in and has n entries.  Rank r-1 needs my first face-row part (its upper halo: the faces
// through h(., row_begin)), rank r+1 needs my last face row (its lower halo).
static pb_status halo_exchange(pb_ctx* ctx, double* full, cudaStream_t stream = nullptr) {
    if (ctx->world == 1) return PB_OK;
    if (!stream) stream = ctx->stream;
    const int F = ctx->F;
    const long long n = ctx->nrows, down = (ctx->rb_begin - ctx->rb_lo) * F, up = (ctx->rb_hi - ctx->rb_end) * F;
    const long long to_prev = ctx->halo_prev * F;        // upper halo of rank r-1: my first face row
    const long long to_next = ctx->halo_next * F;        // lower halo of rank r+1: my last face row
    double* own = full + down;
    NCK(nccl_dl::g.GroupStart());
    if (ctx->rank > 0) {
        NCK(nccl_dl::g.Send(own, (size_t)to_prev, nccl_dl::Float64, ctx->rank - 1, ctx->comm, stream));
        NCK(nccl_dl::g.Recv(full, (size_t)down, nccl_dl::Float64, ctx->rank - 1, ctx->comm, stream));
    }
    if (ctx->rank < ctx->world - 1) {
        NCK(nccl_dl::g.Send(own + n - to_next, (size_t)to_next, nccl_dl::Float64, ctx->rank + 1, ctx->comm, stream));
        NCK(nccl_dl::g.Recv(own + n, (size_t)up, nccl_dl::Float64, ctx->rank + 1, ctx->comm, stream));
    }
    NCK(nccl_dl::g.GroupEnd());
    return PB_OK;
}

// every rank contributes one count (interface problem: the cut faces it owns): sum over the ranks below this one, and over all
static pb_status dist_gather_counts(pb_ctx* ctx, long long mine, long long* before, long long* total) {
    const int W = ctx->world;
    double* d = nullptr;
    struct Cleanup { double*& d; ~Cleanup() { cudaFree(d); } } cleanup{d};
    CK(cudaMalloc((void**)&d, (size_t)W * 8));
    std::vector<double> h((size_t)W, 0.0); h[(size_t)ctx->rank] = (double)mine;      // counts are far below 2^53: exact in double
    CK(cudaMemcpyAsync(d, h.data(), (size_t)W * 8, cudaMemcpyHostToDevice, ctx->stream));
    NCK(nccl_dl::g.AllReduce(d, d, (size_t)W, nccl_dl::Float64, nccl_dl::Sum, ctx->comm, ctx->stream));
    CK(cudaMemcpyAsync(h.data(), d, (size_t)W * 8, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    *before = 0; *total = 0;
    for (int q = 0; q < W; q++) { if (q < ctx->rank) *before += (long long)h[(size_t)q]; *total += (long long)h[(size_t)q]; }
    return PB_OK;
}

// resident blocks per SM of the block SpMV instance for face basis size F: its persistent grid must be exactly one wave
// (a second, partial wave costs a third of the kernel: measured)
template <int F> static int spmv_blk_blocks_per_sm() {
    static int cached = 0;
    if (!cached) { if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&cached, pcg_spmv_blk_kernel<8, F, 0>, PCG_THREADS, 0) != cudaSuccess || cached < 1) cached = 4; }
    return cached;
}
static int spmv_lanes(int F) { int len = 7 * F; return len <= 8 ? 4 : len <= 16 ? 8 : len <= 24 ? 8 : 16; }

// ---- solver workspace: every vector of the Krylov solvers is (re)allocated together whenever the system size differs
// from the size they were made for (pb_set_face_solution on a fresh assembly, a re-assembled interface system with
// another cut, ...), so no entry point can run on a partly allocated or stale-sized set
static pb_status ensure_solver(pb_ctx* ctx) {
    const long long n = ctx->nrows; const int F = ctx->F;
    const long long nfull = (ctx->rb_hi - ctx->rb_lo) * F;
    const bool ok = ctx->solver_n == n && ctx->solver_nfull == nfull && ctx->x && ctx->r && ctx->z && ctx->dfull && ctx->y && ctx->iA && ctx->parts && ctx->scal;
    if (!ok) {
        free_solver(ctx);
        CK(dalloc(ctx, &ctx->x, (size_t)n)); CK(dalloc(ctx, &ctx->r, (size_t)n)); CK(dalloc(ctx, &ctx->z, (size_t)n));
        CK(dalloc(ctx, &ctx->dfull, (size_t)nfull)); CK(dalloc(ctx, &ctx->y, (size_t)n)); CK(dalloc(ctx, &ctx->iA, (size_t)n));
        CK(dalloc(ctx, &ctx->parts, (size_t)6 * PCG_MAX_BLOCKS)); CK(dalloc(ctx, &ctx->scal, 4));      // [d.y: 2 | r.r | r.z | scratch of the multigrid: 2] x PCG_MAX_BLOCKS
        CK(cudaMemsetAsync(ctx->dfull, 0, (size_t)nfull * 8, ctx->stream));
        CK(cudaMemsetAsync(ctx->x, 0, (size_t)n * 8, ctx->stream));
        ctx->solver_n = n; ctx->solver_nfull = nfull;
    }
    ctx->d = ctx->dfull + (ctx->rb_begin - ctx->rb_lo) * F;
    return PB_OK;
}
static pb_status ensure_block_inverse(pb_ctx* ctx) {
    const long long want = ctx->nrows * ctx->F;
    if (!ctx->iAb || ctx->iAb_n != want) { CK(dalloc(ctx, &ctx->iAb, (size_t)want)); ctx->iAb_n = want; }
    return PB_OK;
}

// ---- y = A v on the owned rows (v with halos), with the partial sums of v.y: which kernel, which grid
struct SpmvPlan {
    long long n, xoff, down; int F, lpr, nblk, nblk_sp, nblk_irr; bool use_csr;
    // several ranks: the block rows [0, lo_rows) (first face row of the strip) and [nbr - hi_rows, nbr) (last face row) are the only
    // ones that read halo columns; the interior is launched while the halo travels (nblk_bnd blocks for each boundary launch)
    bool overlap = false; long long lo_rows = 0, hi_rows = 0; int nblk_lo = 0, nblk_hi = 0;
    const float* valf = nullptr;      // set by the multigrid preconditioner: run the block product on the single-precision copy of the values
    int ndy() const { return nblk_sp + nblk_irr + nblk_lo + nblk_hi; }      // number of d.y partials
};
static SpmvPlan make_spmv_plan(pb_ctx* ctx) {
    SpmvPlan P{};
    const int F = ctx->F; const long long n = ctx->nrows;
    P.n = n; P.F = F; P.xoff = ctx->rb_lo * F; P.down = (ctx->rb_begin - ctx->rb_lo) * F;
    int nsm = 148; cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, ctx->device);
    int nblk = nsm * 8; if (nblk > PCG_MAX_BLOCKS) nblk = PCG_MAX_BLOCKS;
    const long long want = (n + PCG_THREADS - 1) / PCG_THREADS; if (want < nblk) nblk = (int)(want < 1 ? 1 : want);
    P.nblk = nblk; P.lpr = spmv_lanes(F);
    // scalar CSR: comparison runs, F = 1 (a 64-byte face record per row costs more than 7 column indices) and F > 4
    P.use_csr = getenv("PB_SPMV_CSR") != nullptr || F > 4 || F == 1 || (ctx->general_csr && !ctx->ifc_struct_ok);
    // interface problem: block kernel on the block rows in interface order + a scalar pass over the rows of the cut cells' faces
    P.nblk_irr = (!P.use_csr && ctx->general_csr && ctx->n_irr_rows > 0) ? (int)std::min<long long>(PCG_MAX_BLOCKS, ((long long)ctx->n_irr_rows * F + 7) / 8) : 0;
    P.nblk_sp = nblk;                                                  // grid of the SpMV = number of d.y partials
    if (!P.use_csr) {
        const int occ = F == 1 ? spmv_blk_blocks_per_sm<1>() : F == 2 ? spmv_blk_blocks_per_sm<2>() : F == 3 ? spmv_blk_blocks_per_sm<3>() : spmv_blk_blocks_per_sm<4>();
        P.nblk_sp = nsm * occ; if (P.nblk_sp > PCG_MAX_BLOCKS) P.nblk_sp = PCG_MAX_BLOCKS;
        const long long wantg = (n / F * 8 + PCG_THREADS - 1) / PCG_THREADS; if (wantg < P.nblk_sp) P.nblk_sp = (int)(wantg < 1 ? 1 : wantg);
        // opt-in (PB_HALO_OVERLAP=1): measured on 2 GPUs at 4096^2 the two extra boundary launches and the fork / join cost what the hidden
        // exchange (98 KB, ~25 us) saves: Jacobi-PCG 3.007 against 2.970 ms per iteration, multigrid-PCG 10.19 against 10.11 (profiles/r2m_overlap.txt)
        if (ctx->world > 1 && ctx->comm && !ctx->general_csr && getenv("PB_HALO_OVERLAP")) {
            const Grid& g = ctx->grid; const long long nbr = n / F, per = 2 * g.Nx - 1;
            const long long lo = ctx->rank > 0 ? std::min(per, nbr) : 0;
            const long long hi = ctx->rank < ctx->world - 1 ? ctx->strip.cf_end - compress_row_start(g, g.row_end - 1) : 0;
            if (lo + hi > 0 && nbr - lo - hi >= 4 * per) {      // enough interior rows to hide the exchange behind
                P.overlap = true; P.lo_rows = lo; P.hi_rows = hi;
                auto blocks_for = [](long long rows) { return rows ? (int)std::max<long long>(1, std::min<long long>(PCG_MAX_BLOCKS / 8, (rows * 8 + PCG_THREADS - 1) / PCG_THREADS)) : 0; };
                P.nblk_lo = blocks_for(lo); P.nblk_hi = blocks_for(hi);
                const long long wanti = ((nbr - lo - hi) * 8 + PCG_THREADS - 1) / PCG_THREADS; if (wanti < P.nblk_sp) P.nblk_sp = (int)(wanti < 1 ? 1 : wanti);
            }
        }
    }
    return P;
}
// block SpMV on the block rows [r0, r0 + cnt) of the owned rows, partial sums of v.y into `parts` (one per block)
static void launch_spmv_blk_range(pb_ctx* ctx, const SpmvPlan& P, long long r0, long long cnt, int blocks, const double* vfull, double* y, double* parts) {
    const int F = P.F; cudaStream_t sm = ctx->stream;
#define PB_SPMV_R(F_) do { if (P.valf) pcg_spmv_blk_kernel<8, F_, 0, float><<<blocks, PCG_THREADS, 0, sm>>>(cnt * F, P.xoff, ctx->brows + r0, ctx->brp + r0, P.valf, vfull, P.down + r0 * F, y + r0 * F, parts, ctx->cg); \
    else pcg_spmv_blk_kernel<8, F_, 0><<<blocks, PCG_THREADS, 0, sm>>>(cnt * F, P.xoff, ctx->brows + r0, ctx->brp + r0, ctx->val, vfull, P.down + r0 * F, y + r0 * F, parts, ctx->cg); } while (0)
    switch (F) { case 1: PB_SPMV_R(1); break; case 2: PB_SPMV_R(2); break; case 3: PB_SPMV_R(3); break; default: PB_SPMV_R(4); break; }
#undef PB_SPMV_R
}
// halo exchange of vfull + y = A v.  Several ranks: the exchange runs on its own stream beside the interior block rows, the rows of
// the strip's first and last face row follow once the halo has arrived (fork / join by events, also valid inside a graph capture).
static pb_status launch_spmv(pb_ctx* ctx, const SpmvPlan& P, const double* vfull, double* y, double* p_dy);
static pb_status spmv_with_halo(pb_ctx* ctx, const SpmvPlan& P, double* vfull, double* y, double* p_dy) {
    if (!P.overlap) {
        pb_status st = halo_exchange(ctx, vfull); if (st != PB_OK) return st;
        return launch_spmv(ctx, P, vfull, y, p_dy);
    }
    if (!ctx->s_halo) {
        CK(cudaStreamCreateWithFlags(&ctx->s_halo, cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming)); CK(cudaEventCreateWithFlags(&ctx->ev_join, cudaEventDisableTiming));
    }
    const long long nbr = P.n / P.F;
    CK(cudaEventRecord(ctx->ev_fork, ctx->stream));
    CK(cudaStreamWaitEvent(ctx->s_halo, ctx->ev_fork, 0));
    { pb_status st = halo_exchange(ctx, vfull, ctx->s_halo); if (st != PB_OK) return st; }
    CK(cudaEventRecord(ctx->ev_join, ctx->s_halo));
    launch_spmv_blk_range(ctx, P, P.lo_rows, nbr - P.lo_rows - P.hi_rows, P.nblk_sp, vfull, y, p_dy);
    CK(cudaStreamWaitEvent(ctx->stream, ctx->ev_join, 0));
    if (P.lo_rows) launch_spmv_blk_range(ctx, P, 0, P.lo_rows, P.nblk_lo, vfull, y, p_dy + P.nblk_sp + P.nblk_irr);
    if (P.hi_rows) launch_spmv_blk_range(ctx, P, nbr - P.hi_rows, P.hi_rows, P.nblk_hi, vfull, y, p_dy + P.nblk_sp + P.nblk_irr + P.nblk_lo);
    CK(cudaGetLastError());
    return PB_OK;
}
static pb_status launch_spmv(pb_ctx* ctx, const SpmvPlan& P, const double* vfull, double* y, double* p_dy) {
    cudaStream_t sm = ctx->stream; const long long n = P.n; const int F = P.F;
    if (P.overlap) { launch_spmv_blk_range(ctx, P, 0, n / F, P.nblk_sp, vfull, y, p_dy);      // (callers that exchanged the halo themselves)
        if (P.nblk_lo + P.nblk_hi) cudaMemsetAsync(p_dy + P.nblk_sp + P.nblk_irr, 0, (size_t)(P.nblk_lo + P.nblk_hi) * 8, sm);
        CK(cudaGetLastError()); return PB_OK; }
#define PB_SPMV(LPR_, F_, CS_) pcg_spmv_blk_kernel<LPR_, F_, CS_><<<P.nblk_sp, PCG_THREADS, 0, sm>>>(n, P.xoff, ctx->brows, ctx->brp, ctx->val, vfull, P.down, y, p_dy, ctx->cg)
    if (P.use_csr) {
        switch (P.lpr) {
        case 4: pcg_spmv_kernel<4><<<P.nblk, PCG_THREADS, 0, sm>>>(n, P.xoff, ctx->rowptr, ctx->col, ctx->val, vfull, P.down, y, p_dy, ctx->cg); break;
        case 8: pcg_spmv_kernel<8><<<P.nblk, PCG_THREADS, 0, sm>>>(n, P.xoff, ctx->rowptr, ctx->col, ctx->val, vfull, P.down, y, p_dy, ctx->cg); break;
        default: pcg_spmv_kernel<16><<<P.nblk, PCG_THREADS, 0, sm>>>(n, P.xoff, ctx->rowptr, ctx->col, ctx->val, vfull, P.down, y, p_dy, ctx->cg); break;
        }
    } else if (P.valf) {
        launch_spmv_blk_range(ctx, P, 0, n / F, P.nblk_sp, vfull, y, p_dy);
    } else {
        switch (F) { case 1: PB_SPMV(8, 1, 0); break; case 2: PB_SPMV(8, 2, 0); break; case 3: PB_SPMV(8, 3, 0); break; default: PB_SPMV(8, 4, 0); break; }
    }
#undef PB_SPMV
    if (P.nblk_irr) pcg_spmv_rows_kernel<<<P.nblk_irr, PCG_THREADS, 0, sm>>>(ctx->n_irr_rows, F, ctx->irr_rows, ctx->rowptr, ctx->col, ctx->val, vfull - P.xoff, P.xoff + P.down, y, p_dy + P.nblk_sp, ctx->cg);
    CK(cudaGetLastError());
    return PB_OK;
}

static pb_status pcg_multigrid(pb_ctx* ctx, const pb_cg_params* p, pb_cg_result* res);      // pb_api_mg.inl
static pb_status interface_recover_owned(pb_ctx* ctx, double* cell_dofs, double* face_dofs);      // pb_api_interface_post.inl

// conjugated_gradient (solver_cg.hpp:63-144) on the condensed system of this context, or of the row of contexts that
// share a communicator: same update order, exit tests and iteration count; dot products are all-reduced.
extern "C" pb_status pb_pcg(pb_ctx* ctx, const pb_cg_params* p, pb_cg_result* res) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->assembled && p, PB_ERR_STATE, "pb_pcg: nothing assembled");
    DRAIN();
    const bool whole = ctx->strip.cf_begin == 0 && ctx->strip.cf_end == num_free_faces(ctx->grid);
    NEED(whole || (ctx->world > 1 && ctx->comm), PB_ERR_STATE, "pb_pcg: this context owns a strip: call pb_dist_init on every rank first");
    NEED(p->apply_preconditioner >= 0 && p->apply_preconditioner <= 3, PB_ERR_INVALID, "pb_pcg: apply_preconditioner must be 0 (none), 1 (diagonal), 2 (block-Jacobi) or 3 (multigrid)");
    CK(cudaSetDevice(ctx->device));
    { pb_status st = ensure_solver(ctx); if (st != PB_OK) return st; }
    if (p->apply_preconditioner == 3) return pcg_multigrid(ctx, p, res);
    const long long n = ctx->nrows; const int F = ctx->F;
    const bool multi = ctx->world > 1;
    const SpmvPlan plan = make_spmv_plan(ctx);
    const int nblk = plan.nblk, nblk_sp = plan.nblk_sp;
    double *p_dy = ctx->parts, *p_rr = ctx->parts + 2 * PCG_MAX_BLOCKS, *p_rz = ctx->parts + 3 * PCG_MAX_BLOCKS;      // d.y: block SpMV partials, then the irregular-row pass
    double *s_dy = ctx->scal, *s_rr = ctx->scal + 1, *s_rz = ctx->scal + 2;
    // what the consuming kernels read: the per-block partials (one rank) or the all-reduced scalars (several)
    const double *c_dy = multi ? s_dy : p_dy, *c_rr = multi ? s_rr : p_rr, *c_rz = multi ? s_rz : p_rz;
    const int c_n = multi ? 1 : nblk;
    const int pre = p->apply_preconditioner ? 1 : 0;
    const bool blockpre = p->apply_preconditioner == 2 && F >= 2 && F <= 4;      // F = 1: the block is the diagonal
    if (blockpre) { pb_status st = ensure_block_inverse(ctx); if (st != PB_OK) return st; }
    const long long nbr_own = n / F;
    const int c_ndy = multi ? 1 : plan.ndy();
    cudaStream_t sm = ctx->stream;
    auto reduce_rr_rz = [&]() -> pb_status {
        if (!multi) return PB_OK;
        pcg_sum_partials_kernel<<<1, PCG_THREADS, 0, sm>>>(nblk, p_rr, s_rr, p_rz, s_rz);
        NCK(nccl_dl::g.AllReduce(s_rr, s_rr, 2, nccl_dl::Float64, nccl_dl::Sum, ctx->comm, sm));
        return PB_OK;
    };
    auto t0 = std::chrono::steady_clock::now();
    CK(cudaEventRecord(ctx->ev[0], sm));
    if (blockpre) {
#define PB_BJ(F_) do { pcg_bdiag_kernel<F_><<<nblk, PCG_THREADS, 0, sm>>>(nbr_own, ctx->rb_begin * F, ctx->rowptr, ctx->col, ctx->val, ctx->iAb); \
        pcg_init_blk_kernel<F_><<<nblk, PCG_THREADS, 0, sm>>>(nbr_own, ctx->rhs, ctx->iAb, ctx->x, ctx->r, ctx->z, ctx->d, p_rr, p_rz); } while (0)
        if (F == 2) PB_BJ(2); else if (F == 3) PB_BJ(3); else PB_BJ(4);
#undef PB_BJ
    } else {
        pcg_diag_kernel<<<nblk, PCG_THREADS, 0, sm>>>(n, ctx->rb_begin * F, ctx->rowptr, ctx->col, ctx->val, ctx->iA);
        pcg_init_kernel<<<nblk, PCG_THREADS, 0, sm>>>(n, pre, ctx->rhs, ctx->iA, ctx->x, ctx->r, ctx->z, ctx->d, p_rr, p_rz);
    }
    { pb_status st = reduce_rr_rz(); if (st != PB_OK) return st; }
    pcg_init_state_kernel<<<1, PCG_THREADS, 0, sm>>>(c_n, c_rr, c_rz, ctx->cg, ctx->hist);
    CK(cudaGetLastError());
    // one iteration = [dir, commit] (skipped in the very first one), halo exchange, spmv, xrz (+ all-reduces)
    auto spmv_xrz = [&]() -> pb_status {
        { pb_status st = spmv_with_halo(ctx, plan, ctx->dfull, ctx->y, p_dy); if (st != PB_OK) return st; }
        if (multi) {
            pcg_sum_partials_kernel<<<1, PCG_THREADS, 0, sm>>>(plan.ndy(), p_dy, s_dy, nullptr, nullptr);
            NCK(nccl_dl::g.AllReduce(s_dy, s_dy, 1, nccl_dl::Float64, nccl_dl::Sum, ctx->comm, sm));
        }
        if (blockpre) {
#define PB_BJ(F_) pcg_xrz_blk_kernel<F_><<<nblk, PCG_THREADS, 0, sm>>>(nbr_own, c_ndy, c_dy, ctx->iAb, ctx->d, ctx->y, ctx->x, ctx->r, ctx->z, p_rr, p_rz, ctx->cg)
            if (F == 2) PB_BJ(2); else if (F == 3) PB_BJ(3); else PB_BJ(4);
#undef PB_BJ
        } else
            pcg_xrz_kernel<<<nblk, PCG_THREADS, 0, sm>>>(n, pre, c_ndy, c_dy, ctx->iA, ctx->d, ctx->y, ctx->x, ctx->r, ctx->z, p_rr, p_rz, ctx->cg);
        return reduce_rr_rz();
    };
    auto full_iteration = [&]() -> pb_status {
        pcg_dir_kernel<<<nblk, PCG_THREADS, 0, sm>>>(n, c_n, c_rr, c_rz, ctx->z, ctx->d, p->convergence_threshold, p->divergence_threshold, p->max_iter, ctx->cg);
        pcg_commit_kernel<<<1, PCG_THREADS, 0, sm>>>(c_n, c_rr, c_rz, p->convergence_threshold, p->divergence_threshold, p->max_iter, ctx->cg, ctx->hist, ctx->hist_cap);
        return spmv_xrz();
    };
    CgState hs{}; const int check_every = 32;
    { pb_status st = spmv_xrz(); if (st != PB_OK) return st; }
    // the iteration carries no host-side state (the kernels read the iteration number and the exit flag from device
    // memory), so a batch of `check_every` iterations is captured once into a CUDA graph and replayed; the host only
    // polls the state between batches.  Multi-rank runs keep plain launches (NCCL calls between the kernels).
    cudaGraph_t graph = nullptr; cudaGraphExec_t gexec = nullptr;
    // (PB_PCG_GRAPH_MULTI=1: capture the multi-rank iteration too — NCCL's send/recv and all-reduces are capturable; opt-in)
    if ((!multi || getenv("PB_PCG_GRAPH_MULTI")) && !getenv("PB_PCG_NO_GRAPH")) {
        CK(cudaStreamBeginCapture(sm, cudaStreamCaptureModeThreadLocal));
        pb_status cst = PB_OK;
        for (int k = 0; k < check_every && cst == PB_OK; k++) cst = full_iteration();
        cudaError_t ce = cudaStreamEndCapture(sm, &graph);      // always closes the capture, also on the error path
        if (cst != PB_OK) { if (graph) cudaGraphDestroy(graph); return cst; }
        CK(ce);
        CK(cudaGraphInstantiate(&gexec, graph, 0));
    }
    while (true) {
        if (gexec) CK(cudaGraphLaunch(gexec, sm));
        else for (int k = 0; k < check_every; k++) { pb_status st = full_iteration(); if (st != PB_OK) return st; }
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(&hs, ctx->cg, sizeof hs, cudaMemcpyDeviceToHost, sm));
        CK(cudaStreamSynchronize(sm));
        if (hs.done) break;
    }
    if (gexec) { cudaGraphExecDestroy(gexec); cudaGraphDestroy(graph); }
    CK(cudaEventRecord(ctx->ev[1], sm)); CK(cudaStreamSynchronize(sm));
    float ms = 0; cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]); ctx->ms_pcg = ms;
    ctx->hist_n = ctx->hist ? std::min<long long>(hs.iters + 2, ctx->hist_cap) : 0;
    if (res) {
        res->iterations = hs.iters; res->relative_residual = hs.relres; res->exit_reason = hs.reason;
        res->seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    }
    return PB_OK;
}

// residual history = the lines conjugated_gradient writes to cg_params::histfile (solver_cg.hpp:92-93, 101-102, 133-137)
extern "C" pb_status pb_pcg_history(pb_ctx* ctx, long long capacity) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(capacity >= 0, PB_ERR_INVALID, "pb_pcg_history: negative capacity");
    CK(cudaSetDevice(ctx->device));
    dfree(&ctx->hist); ctx->hist_cap = 0; ctx->hist_n = 0;
    if (capacity > 0) { CK(dalloc(ctx, &ctx->hist, (size_t)capacity)); ctx->hist_cap = capacity; }
    return PB_OK;
}
extern "C" pb_status pb_pcg_get_history(pb_ctx* ctx, double* out, long long capacity, long long* n) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(n, PB_ERR_INVALID, "pb_pcg_get_history: null count");
    CK(cudaSetDevice(ctx->device));
    *n = ctx->hist_n;
    if (out && ctx->hist_n > 0) CK(cudaMemcpy(out, ctx->hist, (size_t)std::min(capacity, ctx->hist_n) * 8, cudaMemcpyDeviceToHost));
    return PB_OK;
}

// back-substitution (K6) for the owned cell rows; xfull = face solution for the columns [col_lo, col_hi)
static pb_status recover_owned(pb_ctx* ctx, double* cell_dofs_host, double** keep_uc = nullptr, double** keep_xfull = nullptr, double* ulocal_dev = nullptr) {
    const Grid& g = ctx->grid; const int F = ctx->F, C = ctx->C;
    const long long n = ctx->nrows, down = (ctx->strip.cf_begin - ctx->strip.col_lo) * F, nfull = (ctx->strip.col_hi - ctx->strip.col_lo) * F;
    const long long owned_cells = (g.row_end - g.row_begin) * g.Nx;
    double *xfull = nullptr, *uc = nullptr;
    CK(cudaMalloc((void**)&xfull, (size_t)nfull * 8)); CK(cudaMalloc((void**)&uc, (size_t)owned_cells * C * 8));
    CK(cudaMemsetAsync(xfull, 0, (size_t)nfull * 8, ctx->stream));
    CK(cudaMemcpyAsync(xfull + down, ctx->x, (size_t)n * 8, cudaMemcpyDeviceToDevice, ctx->stream));
    { pb_status st = halo_exchange(ctx, xfull); if (st != PB_OK) return st; }
    double* uc_shift = uc - g.row_begin * g.Nx * C;          // the kernel indexes cell dofs by GLOBAL cell id
#define PB_REC(FD_) recover_kernel<FD_><<<148 * 4, 128, 0, ctx->stream>>>(g, ctx->strip.cell_row0, ctx->ncells_local, g.row_begin, C, ctx->prm.fun, \
        ctx->points, ctx->recA, ctx->recb, xfull, ctx->strip.col_lo * F, uc_shift, ulocal_dev)
    switch (ctx->FD) { case 0: PB_REC(0); break; case 1: PB_REC(1); break; case 2: PB_REC(2); break; case 3: PB_REC(3); break;
    default: cudaFree(uc); cudaFree(xfull); return fail(ctx, PB_ERR_UNSUPPORTED, "pb_recover: face degree %d", ctx->FD); }
#undef PB_REC
    CK(cudaGetLastError());
    if (cell_dofs_host) CK(cudaMemcpyAsync(cell_dofs_host, uc, (size_t)owned_cells * C * 8, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (keep_uc) *keep_uc = uc; else cudaFree(uc);
    if (keep_xfull) *keep_xfull = xfull; else cudaFree(xfull);
    return PB_OK;
}

extern "C" pb_status pb_recover_strip(pb_ctx* ctx, double* cell_dofs, double* face_dofs) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->assembled && ctx->x && cell_dofs, PB_ERR_STATE, "pb_recover_strip: no face solution");
    CK(cudaSetDevice(ctx->device));
    if (ctx->general_csr) return interface_recover_owned(ctx, cell_dofs, face_dofs);      // (owned cells + owned cut cells) * cbs cell dofs
    pb_status st = recover_owned(ctx, cell_dofs); if (st != PB_OK) return st;
    if (face_dofs) CK(cudaMemcpy(face_dofs, ctx->x, (size_t)ctx->nrows * 8, cudaMemcpyDeviceToHost));
    return PB_OK;
}

// take_local_data (hho.hpp:408-449) of every owned cell: cbs cell dofs, then 4 fbs face dofs, Dirichlet faces projected
extern "C" pb_status pb_get_local_solution(pb_ctx* ctx, long long cell_begin, long long count, double* out) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->assembled && ctx->x && out && ctx->solver_n == ctx->nrows, PB_ERR_STATE, "pb_get_local_solution: no face solution");
    NEED(!ctx->general_csr, PB_ERR_UNSUPPORTED, "pb_get_local_solution: plain and fictitious-domain problems");
    const long long owned = (ctx->grid.row_end - ctx->grid.row_begin) * ctx->grid.Nx;
    NEED(cell_begin >= 0 && count >= 0 && cell_begin + count <= owned, PB_ERR_INVALID, "pb_get_local_solution: cell range out of bounds (owned cells, 0-based)");
    CK(cudaSetDevice(ctx->device));
    double* ul = nullptr; CK(cudaMalloc((void**)&ul, (size_t)owned * ctx->N * 8));
    pb_status st = recover_owned(ctx, nullptr, nullptr, nullptr, ul);
    if (st == PB_OK) { cudaError_t e = cudaMemcpy(out, ul + cell_begin * ctx->N, (size_t)count * ctx->N * 8, cudaMemcpyDeviceToHost); if (e != cudaSuccess) st = fail(ctx, PB_ERR_CUDA, "pb_get_local_solution: %s", cudaGetErrorString(e)); }
    cudaFree(ul);
    return st;
}

// the reference's FULL solution vector (cell dofs c*cbs+i, then face dofs cbs*ncells + compress[f]*fbs + a)
extern "C" pb_status pb_recover(pb_ctx* ctx, double* full) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->assembled && ctx->x && full, PB_ERR_STATE, "pb_recover: no face solution");
    if (ctx->general_csr) { CK(cudaSetDevice(ctx->device)); return interface_recover(ctx, full); }
    NEED(ctx->strip.cf_begin == 0 && ctx->strip.cf_end == num_free_faces(ctx->grid), PB_ERR_UNSUPPORTED, "pb_recover: whole-mesh contexts only (strips: pb_recover_strip)");
    return pb_recover_strip(ctx, full, full + ctx->ncells_global * ctx->C);
}

// ---- K7: error norms of the drivers (pb_norms.cuh) over the owned cells; strips return partial sums
template <int CD, int FD, int QK>
static cudaError_t launch_norms_t(pb_ctx* c, NormArgs& A, int grid) {
    const size_t smem = sizeof(NormSmem<CD, FD>);
    auto kern = error_norms_kernel<CD, FD, QK>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    kern<<<grid, 32, smem, c->stream>>>(A);
    return cudaGetLastError();
}

// the driver's solution_N_*_k_*.dat (convergence_test.cpp:250-252, 280-286): cell solution at the degree-2*cd points
extern "C" pb_status pb_sample_solution(pb_ctx* ctx, double* xyv, long long capacity_points, long long* n_points) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->assembled && ctx->x && n_points, PB_ERR_STATE, "pb_sample_solution: no face solution");
    if (ctx->general_csr) { CK(cudaSetDevice(ctx->device)); return interface_sample(ctx, xyv, capacity_points, n_points); }      // interface_uT.dat
    NEED(ctx->prm.problem == PB_PROBLEM_HHO, PB_ERR_UNSUPPORTED, "pb_sample_solution: uncut HHO and interface problems only");
    CK(cudaSetDevice(ctx->device));
    const int cd = ctx->CD, qk = ctx->quad_kind;
    NEED(cd >= 0 && cd <= 4, PB_ERR_UNSUPPORTED, "pb_sample_solution: cell degree > 4");
    const Grid& g = ctx->grid;
    const long long owned = (g.row_end - g.row_begin) * g.Nx;
    const long long nq = sample_points_per_cell(cd, qk);
    *n_points = owned * nq;
    if (!xyv) return PB_OK;
    NEED(capacity_points >= owned * nq, PB_ERR_INVALID, "pb_sample_solution: buffer too small");
    double *uc = nullptr, *xfull = nullptr;
    pb_status st = recover_owned(ctx, nullptr, &uc, &xfull); if (st != PB_OK) return st;
    cudaFree(xfull);
    double* out = nullptr; CK(cudaMalloc((void**)&out, (size_t)owned * nq * 24));
    const unsigned grid = (unsigned)std::min<long long>((owned * nq + 255) / 256, 148 * 64);
    bool ok = true;
#define PB_SMP(CD_) case CD_: if (qk == 0) sample_solution_kernel<CD_, 0><<<grid, 256, 0, ctx->stream>>>(g, ctx->points, uc, out); \
                             else sample_solution_kernel<CD_, 1><<<grid, 256, 0, ctx->stream>>>(g, ctx->points, uc, out); break;
    switch (cd) { PB_SMP(0) PB_SMP(1) PB_SMP(2) PB_SMP(3) PB_SMP(4) default: ok = false; }
#undef PB_SMP
    cudaError_t e = ok ? cudaGetLastError() : cudaErrorInvalidValue;
    if (e == cudaSuccess) e = cudaMemcpyAsync(xyv, out, (size_t)owned * nq * 24, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(out); cudaFree(uc);
    if (e != cudaSuccess) return fail(ctx, PB_ERR_CUDA, "pb_sample_solution: %s", cudaGetErrorString(e));
    return PB_OK;
}

extern "C" pb_status pb_error_norms(pb_ctx* ctx, double out[3]) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->assembled && ctx->x && out, PB_ERR_STATE, "pb_error_norms: no face solution");
    if (ctx->general_csr) { CK(cudaSetDevice(ctx->device)); return interface_error(ctx, out); }
    CK(cudaSetDevice(ctx->device));
    double *uc = nullptr, *xfull = nullptr;
    pb_status st = recover_owned(ctx, nullptr, &uc, &xfull); if (st != PB_OK) return st;
    const int grid = 148 * 16;
    double* parts = nullptr; CK(cudaMalloc((void**)&parts, (size_t)grid * 3 * 8));
    NormArgs A{};
    A.grid = ctx->grid; A.points = ctx->points; A.ucell = uc; A.xfull = xfull; A.xoff = ctx->strip.col_lo * ctx->F;
    A.fun = ctx->prm.fun; A.problem = ctx->prm.problem;
    A.cell_loc_g = ctx->cell_loc_g; A.node_loc = ctx->node_loc; A.cut_cells = ctx->cut_cells; A.iface = ctx->iface; A.ncut = ctx->ncut; A.NI = ctx->NI;
    A.parts = parts;
    cudaError_t e = cudaErrorInvalidValue; const int cd = ctx->CD, fd = ctx->FD, qk = ctx->quad_kind;
#define PB_NORM(CD_, FD_, QK_) if (cd == CD_ && fd == FD_ && qk == QK_) e = launch_norms_t<CD_, FD_, QK_>(ctx, A, grid);
    PB_NORM(1, 0, 0) PB_NORM(2, 1, 0) PB_NORM(3, 2, 0) PB_NORM(4, 3, 0) PB_NORM(1, 0, 1) PB_NORM(2, 1, 1) PB_NORM(3, 2, 1)
#undef PB_NORM
    if (e != cudaSuccess) { cudaFree(parts); cudaFree(uc); cudaFree(xfull); return fail(ctx, PB_ERR_UNSUPPORTED, "pb_error_norms: no kernel instance for HHO(%d,%d) quad_kind=%d (%s)", cd, fd, qk, cudaGetErrorString(e)); }
    std::vector<double> h((size_t)grid * 3);
    CK(cudaMemcpyAsync(h.data(), parts, h.size() * 8, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    out[0] = out[1] = out[2] = 0.0;
    for (int b = 0; b < grid; b++) for (int k = 0; k < 3; k++) out[k] += h[3 * b + k];
    cudaFree(parts); cudaFree(uc); cudaFree(xfull);
    return PB_OK;
}
