// C ABI, allocation-free variant of pb_error_norms (included by pb_api.cu).  No counterpart in the reference: added
// for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
// loops, harnesses that forbid it).
//
// pb_error_norms (pb_api_dist.inl) runs K6 (recover_kernel) and K7 (error_norms_kernel) on three scratch arrays it takes
// from cudaMalloc on every call: the face solution padded to the referenced columns, the recovered cell dofs, and the
// per-block partial sums of K7.  Here the same kernels, launched the same way, take that scratch from ONE file-scope device
// array of PB_STATIC_SCRATCH_BYTES (16 MiB; every GPU that loads the library holds it until exit).  The entry point calls
// no cudaMalloc / cudaFree / allocating library routine; a system that needs more scratch is refused with PB_ERR_UNSUPPORTED.
// All calls (all contexts of a device) share the array and it keeps the last call's contents: the padded face vector,
// which must start at zero, is cleared on every call, the other two parts are fully overwritten before they are read.
// Two calls that overlap on the GPU (different contexts = different streams, or different host threads) RACE on it.
//
// In host code `g_pb_static_scratch` is only a stand-in for the symbol, not a device address: the address is taken with
// cudaGetSymbolAddress once per device, with that device current, and that pointer is what the kernels receive.

constexpr size_t PB_STATIC_SCRATCH_BYTES = 16u << 20;
static __device__ __align__(256) unsigned char g_pb_static_scratch[PB_STATIC_SCRATCH_BYTES];

constexpr int PB_STATIC_MAX_DEVICES = 64;
static unsigned char* g_static_scratch_addr[PB_STATIC_MAX_DEVICES] = {};      // device address of the array, per device

// set-up, once per device: `device` must be the current device
static cudaError_t static_scratch_address(int device, unsigned char** out) {
    *out = nullptr;
    if (device < 0 || device >= PB_STATIC_MAX_DEVICES) return cudaErrorInvalidDevice;
    if (!g_static_scratch_addr[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_pb_static_scratch);
        if (e != cudaSuccess) return e;
        g_static_scratch_addr[device] = (unsigned char*)p;
    }
    *out = g_static_scratch_addr[device];
    return cudaSuccess;
}

static size_t static_round256(size_t b) { return (b + 255) & ~(size_t)255; }

extern "C" pb_status pb_error_norms_static(pb_ctx* ctx, double out[3]) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->assembled && ctx->x && out && ctx->solver_n == ctx->nrows, PB_ERR_STATE, "pb_error_norms_static: no face solution");
    NEED(!ctx->general_csr, PB_ERR_UNSUPPORTED, "pb_error_norms_static: plain and fictitious-domain problems (the interface problem has its own post-processing)");
    NEED(ctx->world == 1 && ctx->strip.cf_begin == 0 && ctx->strip.cf_end == num_free_faces(ctx->grid), PB_ERR_UNSUPPORTED,
         "pb_error_norms_static: whole-mesh contexts on one GPU");
    DRAIN();
    CK(cudaSetDevice(ctx->device));
    const Grid& g = ctx->grid; const int F = ctx->F, C = ctx->C;
    const long long n = ctx->nrows, nfull = (ctx->strip.col_hi - ctx->strip.col_lo) * F, down = (ctx->strip.cf_begin - ctx->strip.col_lo) * F;
    const long long owned_cells = (g.row_end - g.row_begin) * g.Nx;
    constexpr int grid = 148 * 16;                                   // the launch of pb_error_norms
    const size_t b_x = static_round256((size_t)nfull * 8), b_uc = static_round256((size_t)owned_cells * C * 8), b_parts = static_round256((size_t)grid * 3 * 8);
    if (b_x + b_uc + b_parts > PB_STATIC_SCRATCH_BYTES)
        return fail(ctx, PB_ERR_UNSUPPORTED, "pb_error_norms_static: this system needs %zu bytes of scratch (%lld face unknowns, %lld cells x %d cell dofs), the static "
                                             "array holds %zu: use pb_error_norms", b_x + b_uc + b_parts, n, owned_cells, C, PB_STATIC_SCRATCH_BYTES);
    unsigned char* base = nullptr;
    cudaError_t se = static_scratch_address(ctx->device, &base);
    if (se != cudaSuccess || !base) return fail(ctx, PB_ERR_CUDA, "pb_error_norms_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(se));
    double* xfull = (double*)base; double* uc = (double*)(base + b_x); double* parts = (double*)(base + b_x + b_uc);
    cudaStream_t sm = ctx->stream;
    // the padded face vector must start at zero (columns outside the owned rows: none on a whole mesh, kept for the layout)
    CK(cudaMemsetAsync(xfull, 0, (size_t)nfull * 8, sm));
    CK(cudaMemcpyAsync(xfull + down, ctx->x, (size_t)n * 8, cudaMemcpyDeviceToDevice, sm));
    double* uc_shift = uc - g.row_begin * g.Nx * C;                  // K6 indexes cell dofs by GLOBAL cell id
#define PB_REC(FD_) recover_kernel<FD_><<<148 * 4, 128, 0, sm>>>(g, ctx->strip.cell_row0, ctx->ncells_local, g.row_begin, C, ctx->prm.fun, \
        ctx->points, ctx->recA, ctx->recb, xfull, ctx->strip.col_lo * F, uc_shift, nullptr)
    switch (ctx->FD) { case 0: PB_REC(0); break; case 1: PB_REC(1); break; case 2: PB_REC(2); break; case 3: PB_REC(3); break;
    default: return fail(ctx, PB_ERR_UNSUPPORTED, "pb_error_norms_static: face degree %d", ctx->FD); }
#undef PB_REC
    CK(cudaGetLastError());
    NormArgs A{};
    A.grid = ctx->grid; A.points = ctx->points; A.ucell = uc; A.xfull = xfull; A.xoff = ctx->strip.col_lo * F;
    A.fun = ctx->prm.fun; A.problem = ctx->prm.problem;
    A.cell_loc_g = ctx->cell_loc_g; A.node_loc = ctx->node_loc; A.cut_cells = ctx->cut_cells; A.iface = ctx->iface; A.ncut = ctx->ncut; A.NI = ctx->NI;
    A.parts = parts;
    cudaError_t e = cudaErrorInvalidValue; const int cd = ctx->CD, fd = ctx->FD, qk = ctx->quad_kind;
#define PB_NORM(CD_, FD_, QK_) if (cd == CD_ && fd == FD_ && qk == QK_) e = launch_norms_t<CD_, FD_, QK_>(ctx, A, grid);
    PB_NORM(1, 0, 0) PB_NORM(2, 1, 0) PB_NORM(3, 2, 0) PB_NORM(4, 3, 0) PB_NORM(1, 0, 1) PB_NORM(2, 1, 1) PB_NORM(3, 2, 1)
#undef PB_NORM
    if (e != cudaSuccess) return fail(ctx, PB_ERR_UNSUPPORTED, "pb_error_norms_static: no kernel instance for HHO(%d,%d) quad_kind=%d (%s)", cd, fd, qk, cudaGetErrorString(e));
    static thread_local double h[grid * 3];                          // host side of the final sum: no allocation either
    CK(cudaMemcpyAsync(h, parts, sizeof h, cudaMemcpyDeviceToHost, sm));
    CK(cudaStreamSynchronize(sm));
    out[0] = out[1] = out[2] = 0.0;
    for (int b = 0; b < grid; b++) for (int k = 0; k < 3; k++) out[k] += h[3 * b + k];      // the summation order of pb_error_norms
    return PB_OK;
}
extern "C" long long pb_static_scratch_bytes(void) { return (long long)PB_STATIC_SCRATCH_BYTES; }
