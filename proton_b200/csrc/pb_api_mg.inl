// C ABI, multigrid-preconditioned CG (included by pb_api.cu): apply_preconditioner = 3 of pb_pcg.  Kernels and the
// description of the method: pb_mg.cuh.  Not reference behaviour: conjugated_gradient's update order, exit
// tests and iteration counting are kept (solver_cg.hpp:63-144), the preconditioner is the V-cycle instead of the diagonal.

static void mg_free(pb_ctx* c) {
    MgState& M = c->mg;
    for (int l = 0; l < MG_MAX_LEVELS; l++) {
        MgLevelData& D = M.lev[l];
        dfree(&D.E); dfree(&D.st); dfree(&D.idiag); dfree(&D.x); dfree(&D.b); dfree(&D.r); dfree(&D.t);
    }
    dfree(&M.resfull); dfree(&M.zfull); dfree(&M.wfull); dfree(&M.lam); dfree(&M.valf); M.valf_n = 0;
    if (M.fac_h) { cudaFree(M.fac_h); M.fac_h = nullptr; } if (M.fac_v) { cudaFree(M.fac_v); M.fac_v = nullptr; } dfree(&M.tl); M.fac_key[0] = M.fac_key[1] = M.fac_key[2] = 0; M.lines = false;
    M.ready = false; M.gen = -1; M.nlev = 0; M.nfull = 0; M.key[0] = M.key[1] = M.key[2] = M.key[3] = 0;
}

static int mg_blocks(long long n) { long long b = (n + 255) / 256; return (int)(b < 1 ? 1 : b > 148 * 8 ? 148 * 8 : b); }

static MgCut mg_cut_of(pb_ctx* ctx) {
    MgCut C{};
    if (ctx->prm.problem == PB_PROBLEM_FICTDOM) { C.node_loc = ctx->node_loc; C.face_loc = ctx->face_loc; C.face_ip = ctx->face_ip; C.points = ctx->points; }
    return C;
}

// halo rows of a level-0 vertex vector: my first owned row goes to the rank below, my last one to the rank above
static pb_status mg_halo_nodes(pb_ctx* ctx, double* v) {
    if (ctx->world == 1) return PB_OK;
    const MgGrid& L = ctx->mg.lev[0].L; const long long W = L.Nx + 1;
    const long long rb = ctx->grid.row_begin, re = ctx->grid.row_end;
    NCK(nccl_dl::g.GroupStart());
    if (ctx->rank > 0) {
        NCK(nccl_dl::g.Send(v + (rb - L.row0) * W, (size_t)W, nccl_dl::Float64, ctx->rank - 1, ctx->comm, ctx->stream));
        NCK(nccl_dl::g.Recv(v + (rb - 1 - L.row0) * W, (size_t)W, nccl_dl::Float64, ctx->rank - 1, ctx->comm, ctx->stream));
    }
    if (ctx->rank < ctx->world - 1) {
        NCK(nccl_dl::g.Send(v + (re - 1 - L.row0) * W, (size_t)W, nccl_dl::Float64, ctx->rank + 1, ctx->comm, ctx->stream));
        NCK(nccl_dl::g.Recv(v + (re - L.row0) * W, (size_t)W, nccl_dl::Float64, ctx->rank + 1, ctx->comm, ctx->stream));
    }
    NCK(nccl_dl::g.GroupEnd());
    return PB_OK;
}

// hierarchy of vertex levels for this context's mesh / strip; (re)allocates when the structure changed
static pb_status mg_layout(pb_ctx* ctx) {
    MgState& M = ctx->mg;
    const Grid& g = ctx->grid; const int F = ctx->F;
    const long long nfull = (ctx->strip.col_hi - ctx->strip.col_lo) * F;
    const long long key[4] = {g.Nx * 100 + F, g.Ny, g.row_begin, g.row_end * 8 + ctx->world};
    if (M.nlev > 0 && M.key[0] == key[0] && M.key[1] == key[1] && M.key[2] == key[2] && M.key[3] == key[3] && M.nfull == nfull) return PB_OK;
    mg_free(ctx);
    NEED(g.Nx < (1 << 30) && g.Ny < (1 << 30), PB_ERR_UNSUPPORTED, "pb_pcg: mesh too large for the multigrid levels");
    int nlev = 0; int Nx = (int)g.Nx, Ny = (int)g.Ny;
    while (true) {
        NEED(nlev < MG_MAX_LEVELS, PB_ERR_UNSUPPORTED, "pb_pcg: too many multigrid levels");
        MgLevelData& D = M.lev[nlev]; MgGrid& L = D.L;
        L.Nx = Nx; L.Ny = Ny; L.cx = Nx > 2; L.cy = Ny > 2;
        if (nlev == 0) {      // the strip of this rank: owned vertex rows [row_begin, row_end), one halo row each side
            L.row0 = (int)(g.row_begin > 0 ? g.row_begin - 1 : 0);
            L.nrows = (int)((g.row_end + 1 <= g.Ny ? g.row_end + 1 : g.Ny + 1) - L.row0);
            L.own0 = (int)g.row_begin; L.own1 = (int)g.row_end;
            L.crow0 = (int)ctx->strip.cell_row0; L.ncrows = (int)(g.row_end - ctx->strip.cell_row0);
        } else { L.row0 = 0; L.nrows = Ny + 1; L.own0 = 0; L.own1 = Ny + 1; L.crow0 = 0; L.ncrows = Ny; }
        const bool last = !(L.cx || L.cy);
        D.omega = last ? 1.0 : MG_OMEGA;      // the coarsest level has a single interior vertex: one undamped sweep solves it
        const size_t nn = (size_t)L.nrows * (L.Nx + 1), ncell = (size_t)L.ncrows * L.Nx;
        CK(dalloc(ctx, &D.E, ncell * 16)); CK(dalloc(ctx, &D.st, nn * 9)); CK(dalloc(ctx, &D.idiag, nn));
        CK(dalloc(ctx, &D.x, nn)); CK(dalloc(ctx, &D.b, nn)); CK(dalloc(ctx, &D.r, nn)); CK(dalloc(ctx, &D.t, nn));
        for (double* v : {D.idiag, D.x, D.b, D.r, D.t}) CK(cudaMemsetAsync(v, 0, nn * 8, ctx->stream));
        CK(cudaMemsetAsync(D.st, 0, nn * 72, ctx->stream));
        nlev++;
        if (last) break;
        Nx = mg_coarse_cells(Nx); Ny = mg_coarse_cells(Ny);
    }
    M.nlev = nlev; M.F = F; M.nfull = nfull;
    CK(dalloc(ctx, &M.resfull, (size_t)nfull)); CK(dalloc(ctx, &M.zfull, (size_t)nfull)); CK(dalloc(ctx, &M.wfull, (size_t)nfull)); CK(dalloc(ctx, &M.lam, 2));
    for (double* v : {M.resfull, M.zfull, M.wfull}) CK(cudaMemsetAsync(v, 0, (size_t)nfull * 8, ctx->stream));
    for (int k = 0; k < 4; k++) M.key[k] = key[k];
    return PB_OK;
}

template <int F> static void mg_launch_bdiag(pb_ctx* ctx, int nblk, long long nbr) {
    pcg_bdiag_kernel<F><<<nblk, PCG_THREADS, 0, ctx->stream>>>(nbr, ctx->strip.cf_begin * F, ctx->rowptr, ctx->col, ctx->val, ctx->iAb);
}
template <int F> static void mg_launch_cheb(pb_ctx* ctx, int nblk, long long nbr, int first, int have_y, int acc, int last, double c1, double c2,
                                            const double* src, double* res, double* d, double* z) {
    mg_cheb_kernel<F><<<nblk, PCG_THREADS, 0, ctx->stream>>>(nbr, first, have_y, acc, last, c1, c2, ctx->iAb, src, ctx->y, res, d, z, ctx->cg);
}
// line-block Jacobi (pb_mg_lines.cuh): the problems whose stabilisation is scaled by 1/area (naive and cut stabilisation), on
// single-rank contexts.  On several ranks the face-block smoother is the default: the rows of vertical faces of a strip keep their full
// length, so the line solve does not get cheaper with the number of ranks while everything else does (BASELINE target on 8 GPUs: 41 iterations
// x 5.6 ms with the lines against 60 x 3.1 ms without, profiles/r3c_mg_dist8*.txt; one GPU: 36 x 24.3 against 60 x 16.8).  PB_MG_LINES=0 / 1 overrides.
static bool mg_wants_lines(const pb_ctx* ctx) {
    const char* e = getenv("PB_MG_LINES");
    if (e) return atoi(e) != 0;
    return ctx->world == 1 && (ctx->prm.problem == PB_PROBLEM_FICTDOM || ctx->prm.stab_kind != PB_STAB_FANCY);
}
template <int F, typename VT> static void mg_launch_line_factor(pb_ctx* ctx, const MgLineSet& H, const MgLineSet& V) {
    MgState& M = ctx->mg; const long long row0 = ctx->strip.cf_begin * F;
    for (int dir = 0; dir < 2; dir++) {
        const MgLineSet& S = dir ? V : H; VT* fac = (VT*)(dir ? M.fac_v : M.fac_h);
        if (S.nlines <= 0 || S.nsteps <= 0) continue;
        mg_line_gather_kernel<F, VT><<<mg_blocks((long long)S.nlines * S.nsteps), 256, 0, ctx->stream>>>(S, row0, ctx->rowptr, ctx->col, ctx->val, fac);
        mg_line_factor_kernel<F, VT><<<(S.nlines + 31) / 32, 32, 0, ctx->stream>>>(S, fac);
    }
}
// tl = L^{-1} r (owned rows)
template <int F, typename VT> static void mg_launch_line_solve(pb_ctx* ctx, const MgLineSet& H, const MgLineSet& V, const double* r, double* t, const CgState* st) {
    MgState& M = ctx->mg;
    static const int lanes = [] { const char* e = getenv("PB_MG_LINE_LANES"); const int v = e ? atoi(e) : 16; return v >= 32 ? 32 : v >= 16 ? 16 : v >= 8 ? 8 : 4; }();
    const int nbA = mg_line_blocks(H, lanes), nbB = mg_line_blocks(V, lanes);
    if (nbA + nbB > 0) mg_line_solve_kernel<F, VT><<<nbA + nbB, 32, 0, ctx->stream>>>(H, (const VT*)M.fac_h, nbA, V, (const VT*)M.fac_v, lanes, r, t, st);
}
static void mg_line_solve(pb_ctx* ctx, const double* r, double* t, const CgState* st) {
    MgLineSet H, V; mg_line_sets(ctx->grid, ctx->grid.row_begin, ctx->grid.row_end, ctx->strip.cf_begin, H, V);
    const bool f64 = ctx->mg.lines_f64;
    switch (ctx->F) {
    case 1: if (f64) mg_launch_line_solve<1, double>(ctx, H, V, r, t, st); else mg_launch_line_solve<1, float>(ctx, H, V, r, t, st); break;
    case 2: if (f64) mg_launch_line_solve<2, double>(ctx, H, V, r, t, st); else mg_launch_line_solve<2, float>(ctx, H, V, r, t, st); break;
    case 3: if (f64) mg_launch_line_solve<3, double>(ctx, H, V, r, t, st); else mg_launch_line_solve<3, float>(ctx, H, V, r, t, st); break;
    default: if (f64) mg_launch_line_solve<4, double>(ctx, H, V, r, t, st); else mg_launch_line_solve<4, float>(ctx, H, V, r, t, st); break;
    }
}
// factors of the current system values
static pb_status mg_line_setup(pb_ctx* ctx) {
    MgState& M = ctx->mg; const int F = ctx->F;
    MgLineSet H, V; mg_line_sets(ctx->grid, ctx->grid.row_begin, ctx->grid.row_end, ctx->strip.cf_begin, H, V);
    const bool f64 = getenv("PB_MG_FP64") != nullptr;
    const long long nh = (long long)H.nlines * H.nsteps, nv = (long long)V.nlines * V.nsteps;
    NEED((nh + nv) * F == ctx->nrows, PB_ERR_STATE, "pb_pcg: the mesh lines do not cover the owned faces");
    const long long key[3] = {nh * F * F, nv * F * F, f64 ? 8 : 4};
    if (key[0] != M.fac_key[0] || key[1] != M.fac_key[1] || key[2] != M.fac_key[2] || !M.tl) {
        if (M.fac_h) { cudaFree(M.fac_h); M.fac_h = nullptr; } if (M.fac_v) { cudaFree(M.fac_v); M.fac_v = nullptr; }
        CK(cudaMalloc(&M.fac_h, (size_t)std::max<long long>(1, nh) * 2 * F * F * key[2])); CK(cudaMalloc(&M.fac_v, (size_t)std::max<long long>(1, nv) * 2 * F * F * key[2]));
        CK(dalloc(ctx, &M.tl, (size_t)ctx->nrows));
        for (int k = 0; k < 3; k++) M.fac_key[k] = key[k];
    }
    M.lines_f64 = f64;
    switch (F) {
    case 1: if (f64) mg_launch_line_factor<1, double>(ctx, H, V); else mg_launch_line_factor<1, float>(ctx, H, V); break;
    case 2: if (f64) mg_launch_line_factor<2, double>(ctx, H, V); else mg_launch_line_factor<2, float>(ctx, H, V); break;
    case 3: if (f64) mg_launch_line_factor<3, double>(ctx, H, V); else mg_launch_line_factor<3, float>(ctx, H, V); break;
    default: if (f64) mg_launch_line_factor<4, double>(ctx, H, V); else mg_launch_line_factor<4, float>(ctx, H, V); break;
    }
    CK(cudaGetLastError());
    return PB_OK;
}
#define PB_MG_F(call_) do { switch (ctx->F) { case 1: call_(1); break; case 2: call_(2); break; case 3: call_(3); break; default: call_(4); break; } } while (0)

// operators of all levels from the current system values: block inverses, lmax, element matrices, stencils
static pb_status mg_setup(pb_ctx* ctx) {
    MgState& M = ctx->mg;
    const int F = ctx->F;
    NEED(!ctx->general_csr, PB_ERR_UNSUPPORTED, "pb_pcg: the multigrid preconditioner is built for the plain and fictitious-domain problems (the interface problem keeps apply_preconditioner <= 2)");
    NEED(F >= 1 && F <= 4 && ctx->S, PB_ERR_UNSUPPORTED, "pb_pcg: the multigrid preconditioner needs face degree <= 3");
    { pb_status st = mg_layout(ctx); if (st != PB_OK) return st; }
    { pb_status st = ensure_block_inverse(ctx); if (st != PB_OK) return st; }
    if (M.ready && M.gen == ctx->asm_gen) return PB_OK;
    cudaStream_t sm = ctx->stream;
    const bool multi = ctx->world > 1;
    pcg_init_state_kernel<<<1, 32, 0, sm>>>(0, ctx->parts, ctx->parts, ctx->cg, nullptr);      // the SpMV kernels return early on a finished solve's `done`
    const SpmvPlan plan = make_spmv_plan(ctx);
    const long long n = ctx->nrows, nbr = n / F, down = plan.down;
    // fine level: D^{-1}
#define PB_BD(F_) mg_launch_bdiag<F_>(ctx, plan.nblk, nbr)
    PB_MG_F(PB_BD);
#undef PB_BD
    CK(cudaGetLastError());
    M.lines = mg_wants_lines(ctx);
    if (M.lines) { pb_status st = mg_line_setup(ctx); if (st != PB_OK) return st; }
    // level 0 element matrices E_T = P_T^T S_T P_T of the local cells, Galerkin products below, stencils
    const MgCut C = mg_cut_of(ctx);
    Grid g = ctx->grid;
#define PB_EL(F_) mg_elements_kernel<F_><<<mg_blocks(ctx->ncells_local), 256, 0, sm>>>(g, C, ctx->strip.cell_row0, ctx->ncells_local, ctx->S, M.lev[0].E)
    PB_MG_F(PB_EL);
#undef PB_EL
    CK(cudaGetLastError());
    for (int l = 0; l + 1 < M.nlev; l++) {
        const MgGrid &Lf = M.lev[l].L, &Lc = M.lev[l + 1].L;
        const long long nc = (long long)Lc.Nx * Lc.Ny;
        const int lo = l == 0 ? (int)g.row_begin : 0, hi = l == 0 ? (int)g.row_end : Lf.Ny;      // every fine cell is counted by exactly one rank
        mg_galerkin_kernel<<<mg_blocks(nc), 256, 0, sm>>>(Lf, Lc, M.lev[l].E, M.lev[l + 1].E, lo, hi);
        CK(cudaGetLastError());
        if (l == 0 && multi) NCK(nccl_dl::g.AllReduce(M.lev[1].E, M.lev[1].E, (size_t)nc * 16, nccl_dl::Float64, nccl_dl::Sum, ctx->comm, sm));
    }
    for (int l = 0; l < M.nlev; l++) {
        const MgGrid& L = M.lev[l].L;
        mg_stencil_kernel<<<mg_blocks((long long)(L.own1 - L.own0) * (L.Nx + 1)), 256, 0, sm>>>(L, M.lev[l].E, M.lev[l].st, M.lev[l].idiag, M.lev[l].omega);
        CK(cudaGetLastError());
    }
    // lmax(D^{-1} A) by power iteration from a fixed start vector (the same for any partition of the rows)
    double* v = M.wfull + down; double* w = M.resfull + down;
    double *parts = ctx->parts, *ssum = ctx->scal + 3;
    mg_pow_init_kernel<<<plan.nblk, PCG_THREADS, 0, sm>>>(n, ctx->strip.cf_begin * F, v);
    const int iters = 14;
    for (int it = 0; it <= iters; it++) {
        if (it > 0) {
            { pb_status st = spmv_with_halo(ctx, plan, M.wfull, ctx->y, parts + 4 * PCG_MAX_BLOCKS); if (st != PB_OK) return st; }
#define PB_PW(F_) mg_pow_apply_kernel<F_><<<plan.nblk, PCG_THREADS, 0, sm>>>(nbr, ctx->iAb, ctx->y, w, parts)
            if (M.lines) { mg_line_solve(ctx, ctx->y, w, nullptr); mg_norm2_kernel<<<plan.nblk, PCG_THREADS, 0, sm>>>(n, w, parts); }
            else PB_MG_F(PB_PW);
#undef PB_PW
        } else {
            mg_dot_kernel<<<plan.nblk, PCG_THREADS, 0, sm>>>(n, v, v, parts, nullptr, nullptr);      // normalise the start vector
            CK(cudaMemcpyAsync(w, v, (size_t)n * 8, cudaMemcpyDeviceToDevice, sm));
        }
        pcg_sum_partials_kernel<<<1, PCG_THREADS, 0, sm>>>(plan.nblk, parts, ssum, nullptr, nullptr);
        if (multi) NCK(nccl_dl::g.AllReduce(ssum, ssum, 1, nccl_dl::Float64, nccl_dl::Sum, ctx->comm, sm));
        mg_pow_scale_kernel<<<plan.nblk, PCG_THREADS, 0, sm>>>(n, ssum, w, v, M.lam);
        CK(cudaGetLastError());
    }
    double lam = 0.0;
    CK(cudaMemcpyAsync(&lam, M.lam, 8, cudaMemcpyDeviceToHost, sm));
    CK(cudaStreamSynchronize(sm));
    NEED(lam > 0.0 && std::isfinite(lam), PB_ERR_STATE, "pb_pcg: power iteration for the smoother failed (singular diagonal block?)");
    M.lmax = 1.1 * lam;      // the iteration approaches lmax from below
    // single-precision copy of the values for the SpMVs INSIDE the V-cycle (4 of the 5 products of an iteration): a V-cycle on a
    // slightly perturbed symmetric matrix is still a fixed SPD operator, so the CG recurrence, which keeps the double matrix, double
    // vectors and double dot products, converges to the same solution in the same number of iterations, at half the matrix traffic
    // of those products.  PB_MG_FP64=1 keeps the V-cycle on the double values (the tests compare that variant with the restatement).
    if (!plan.use_csr && !getenv("PB_MG_FP64")) {
        if (M.valf_n != ctx->nnz) { CK(dalloc(ctx, &M.valf, (size_t)ctx->nnz)); M.valf_n = ctx->nnz; }
        mg_val_to_float_kernel<<<148 * 8, 256, 0, sm>>>(ctx->nnz, ctx->val, M.valf);
        CK(cudaGetLastError());
    } else { dfree(&M.valf); M.valf_n = 0; }
    CK(cudaMemsetAsync(M.wfull, 0, (size_t)M.nfull * 8, sm)); CK(cudaMemsetAsync(M.resfull, 0, (size_t)M.nfull * 8, sm));
    M.ready = true; M.gen = ctx->asm_gen;
    return PB_OK;
}

// V(1,1) cycle on the vertex levels: lev[0].b -> lev[0].x (halo rows of x valid on return)
static pb_status mg_vcycle_nodes(pb_ctx* ctx) {
    MgState& M = ctx->mg; cudaStream_t sm = ctx->stream;
    const bool multi = ctx->world > 1;
    for (int l = 0; l < M.nlev; l++) {
        MgLevelData& D = M.lev[l]; const MgGrid& L = D.L;
        const int nb = mg_blocks((long long)(L.own1 - L.own0) * (L.Nx + 1));
        mg_jacobi_first_kernel<<<nb, 256, 0, sm>>>(L, D.idiag, D.b, D.x);
        if (l == M.nlev - 1) break;
        if (l == 0 && multi) { pb_status st = mg_halo_nodes(ctx, D.x); if (st != PB_OK) return st; }
        mg_stencil_apply_kernel<0><<<nb, 256, 0, sm>>>(L, D.st, D.idiag, D.b, D.x, D.r);
        const MgGrid& Lc = M.lev[l + 1].L;
        const long long ncn = (long long)(Lc.Nx + 1) * (Lc.Ny + 1);
        mg_restrict_kernel<<<mg_blocks(ncn), 256, 0, sm>>>(L, Lc, D.r, M.lev[l + 1].b);
        CK(cudaGetLastError());
        if (l == 0 && multi) NCK(nccl_dl::g.AllReduce(M.lev[1].b, M.lev[1].b, (size_t)ncn, nccl_dl::Float64, nccl_dl::Sum, ctx->comm, sm));
    }
    for (int l = M.nlev - 2; l >= 0; l--) {
        MgLevelData& D = M.lev[l]; const MgGrid& L = D.L;
        const int nb = mg_blocks((long long)(L.own1 - L.own0) * (L.Nx + 1));
        mg_prolong_add_kernel<<<nb, 256, 0, sm>>>(L, M.lev[l + 1].L, M.lev[l + 1].x, D.x);
        if (l == 0 && multi) { pb_status st = mg_halo_nodes(ctx, D.x); if (st != PB_OK) return st; }
        mg_stencil_apply_kernel<1><<<nb, 256, 0, sm>>>(L, D.st, D.idiag, D.b, D.x, D.t);
        std::swap(D.x, D.t);
        CK(cudaGetLastError());
    }
    if (multi) { pb_status st = mg_halo_nodes(ctx, M.lev[0].x); if (st != PB_OK) return st; }
    CK(cudaGetLastError());
    return PB_OK;
}

// z = M^{-1} r: nu Chebyshev steps, coarse correction through the vertex levels, nu Chebyshev steps (symmetric).
// r = owned rows; the result is the owned part of mg.zfull.
static pb_status mg_apply(pb_ctx* ctx, const SpmvPlan& plan_d, const double* r, int nu) {
    MgState& M = ctx->mg; cudaStream_t sm = ctx->stream;
    SpmvPlan plan = plan_d; plan.valf = M.valf;      // the products of the V-cycle run on the single-precision copy when there is one
    const int F = ctx->F; const long long n = ctx->nrows, nbr = n / F, down = plan.down;
    double *z = M.zfull + down, *w = M.wfull + down, *rs = M.resfull + down;
    double* pdummy = ctx->parts + 4 * PCG_MAX_BLOCKS;      // the SpMV's v.y partials are not used here (scratch area of the partials buffer)
    const double lmax = M.lmax, lmin = MG_CHEB_LO * lmax, theta = 0.5 * (lmax + lmin), delta = 0.5 * (lmax - lmin), sigma = theta / delta;
    const MgCut C = mg_cut_of(ctx);
    auto cheb = [&](int first, int have_y, int acc, int last, double c1, double c2) {
        if (M.lines) {      // residual update, t = L^{-1} res along the mesh lines, direction update
#define PB_CH(F_) mg_launch_cheb<F_>(ctx, plan.nblk, nbr, first, have_y, acc, 1, c1, c2, r, rs, w, z)
            PB_MG_F(PB_CH);
#undef PB_CH
            if (last) return;
            mg_line_solve(ctx, rs, M.tl, ctx->cg);
            mg_cheb_combine_kernel<<<plan.nblk, PCG_THREADS, 0, sm>>>(n, first, acc, c1, c2, M.tl, w, z, ctx->cg);
            return;
        }
#define PB_CH(F_) mg_launch_cheb<F_>(ctx, plan.nblk, nbr, first, have_y, acc, last, c1, c2, r, rs, w, z)
        PB_MG_F(PB_CH);
#undef PB_CH
    };
    auto spmv = [&](double* vfull) -> pb_status { return spmv_with_halo(ctx, plan, vfull, ctx->y, pdummy); };
    auto steps = [&](bool from_zero) -> pb_status {      // the Chebyshev recurrence after its first step
        double rho = 1.0 / sigma;
        for (int k = 1; k < nu; k++) {
            { pb_status st = spmv(M.wfull); if (st != PB_OK) return st; }
            const double rho_new = 1.0 / (2.0 * sigma - rho);
            cheb(0, 1, 1, 0, rho_new * rho, 2.0 * rho_new / delta);
            rho = rho_new;
        }
        if (from_zero) {      // residual after the last step, for the restriction
            { pb_status st = spmv(M.wfull); if (st != PB_OK) return st; }
            cheb(0, 1, 1, 1, 0.0, 0.0);
        }
        return PB_OK;
    };
    // pre-smoothing from z = 0
    cheb(1, 0, 0, 0, 0.0, 1.0 / theta);
    { pb_status st = steps(true); if (st != PB_OK) return st; }
    // coarse correction
    { pb_status st = halo_exchange(ctx, M.resfull); if (st != PB_OK) return st; }
    const MgGrid& L0 = M.lev[0].L;
    mg_face_restrict_kernel<<<mg_blocks((long long)(L0.own1 - L0.own0) * (L0.Nx + 1)), 256, 0, sm>>>(ctx->grid, C, F, L0, ctx->strip.col_lo, M.resfull, M.lev[0].b);
    CK(cudaGetLastError());
    { pb_status st = mg_vcycle_nodes(ctx); if (st != PB_OK) return st; }
    mg_face_prolong_add_kernel<<<plan.nblk, PCG_THREADS, 0, sm>>>(ctx->grid, C, F, L0, ctx->strip.cf_begin, nbr, M.lev[0].x, z);
    CK(cudaGetLastError());
    // post-smoothing
    { pb_status st = spmv(M.zfull); if (st != PB_OK) return st; }
    cheb(1, 1, 1, 0, 0.0, 1.0 / theta);
    { pb_status st = steps(false); if (st != PB_OK) return st; }
    CK(cudaGetLastError());
    return PB_OK;
}

// Chebyshev steps per smoothing sweep: 1.  More steps buy fewer iterations than they cost: 2048^2, k = 2, plain problem 14 x 4.2 ms (nu = 1) against
// 11 x 6.6 (nu = 2) and 10 x 8.8 (nu = 3); 4096^2 fictitious-domain problem with the line smoother 36 x 23.9 ms against 21 x 40.1 and 19 x 56.2
// (profiles/r2y_mg_lines.txt).  PB_MG_NU overrides.
static int mg_smoothing_steps(const pb_ctx* ctx) {
    const char* e = getenv("PB_MG_NU");
    const int v = e ? atoi(e) : 1;
    (void)ctx;
    return v < 1 ? 1 : v > 8 ? 8 : v;
}

static pb_status pcg_multigrid(pb_ctx* ctx, const pb_cg_params* p, pb_cg_result* res) {
    cudaStream_t sm = ctx->stream;
    auto t0 = std::chrono::steady_clock::now();
    CK(cudaEventRecord(ctx->ev[0], sm));
    { pb_status st = mg_setup(ctx); if (st != PB_OK) return st; }
    MgState& M = ctx->mg;
    const SpmvPlan plan = make_spmv_plan(ctx);
    const long long n = ctx->nrows; const int nblk = plan.nblk, nu = mg_smoothing_steps(ctx);
    const bool multi = ctx->world > 1;
    double *p_dy = ctx->parts, *p_rr = ctx->parts + 2 * PCG_MAX_BLOCKS, *p_rz = ctx->parts + 3 * PCG_MAX_BLOCKS;
    double *s_dy = ctx->scal, *s_rr = ctx->scal + 1, *s_rz = ctx->scal + 2;
    const double *c_dy = multi ? s_dy : p_dy, *c_rr = multi ? s_rr : p_rr, *c_rz = multi ? s_rz : p_rz;
    const int c_n = multi ? 1 : nblk, c_ndy = multi ? 1 : plan.ndy();
    double* z = M.zfull + plan.down;
    auto reduce1 = [&](int np, double* parts, double* s) -> pb_status {
        if (!multi) return PB_OK;
        pcg_sum_partials_kernel<<<1, PCG_THREADS, 0, sm>>>(np, parts, s, nullptr, nullptr);
        NCK(nccl_dl::g.AllReduce(s, s, 1, nccl_dl::Float64, nccl_dl::Sum, ctx->comm, sm));
        return PB_OK;
    };
    // x = 0, r = b, z = M^{-1} r, d = z                                              (solver_cg.hpp:74-85)
    CK(cudaMemsetAsync(ctx->x, 0, (size_t)n * 8, sm));
    CK(cudaMemcpyAsync(ctx->r, ctx->rhs, (size_t)n * 8, cudaMemcpyDeviceToDevice, sm));
    pcg_init_state_kernel<<<1, 32, 0, sm>>>(0, p_rr, p_rz, ctx->cg, nullptr);      // clears `done` before the first V-cycle reads it
    { pb_status st = mg_apply(ctx, plan, ctx->r, nu); if (st != PB_OK) return st; }
    mg_dot_kernel<<<nblk, PCG_THREADS, 0, sm>>>(n, ctx->r, z, p_rz, p_rr, nullptr);
    { pb_status st = reduce1(nblk, p_rr, s_rr); if (st != PB_OK) return st; }
    { pb_status st = reduce1(nblk, p_rz, s_rz); if (st != PB_OK) return st; }
    pcg_init_state_kernel<<<1, PCG_THREADS, 0, sm>>>(c_n, c_rr, c_rz, ctx->cg, ctx->hist);
    CK(cudaMemcpyAsync(ctx->d, z, (size_t)n * 8, cudaMemcpyDeviceToDevice, sm));
    CK(cudaGetLastError());
    CgState hs{};
    // one iteration; `first` skips the direction update (d = z already)
    auto iteration = [&](bool first) -> pb_status {
        if (!first) {      // exit tests, beta, d = z + beta d, iter++                            (solver_cg.hpp:112-131)
            pcg_dir_kernel<<<nblk, PCG_THREADS, 0, sm>>>(n, c_n, c_rr, c_rz, z, ctx->d, p->convergence_threshold, p->divergence_threshold, p->max_iter, ctx->cg);
            pcg_commit_kernel<<<1, PCG_THREADS, 0, sm>>>(c_n, c_rr, c_rz, p->convergence_threshold, p->divergence_threshold, p->max_iter, ctx->cg, ctx->hist, ctx->hist_cap);
        }
        { pb_status st = spmv_with_halo(ctx, plan, ctx->dfull, ctx->y, p_dy); if (st != PB_OK) return st; }
        { pb_status st = reduce1(plan.ndy(), p_dy, s_dy); if (st != PB_OK) return st; }
        mg_xr_kernel<<<nblk, PCG_THREADS, 0, sm>>>(n, c_ndy, c_dy, ctx->d, ctx->y, ctx->x, ctx->r, p_rr, ctx->cg);
        { pb_status st = reduce1(nblk, p_rr, s_rr); if (st != PB_OK) return st; }
        mg_check_kernel<<<1, PCG_THREADS, 0, sm>>>(c_n, c_rr, p->convergence_threshold, p->divergence_threshold, p->max_iter, ctx->cg, ctx->hist, ctx->hist_cap);
        { pb_status st = mg_apply(ctx, plan, ctx->r, nu); if (st != PB_OK) return st; }
        mg_dot_kernel<<<nblk, PCG_THREADS, 0, sm>>>(n, ctx->r, z, p_rz, nullptr, ctx->cg);
        { pb_status st = reduce1(nblk, p_rz, s_rz); if (st != PB_OK) return st; }
        CK(cudaGetLastError());
        return PB_OK;
    };
    { pb_status st = iteration(true); if (st != PB_OK) return st; }
    // An iteration is ~100 launches (12 vertex levels, most of them tiny) and carries no host state: batches of `check_every`
    // iterations are captured once into a CUDA graph (NCCL's send/recv and all-reduces included on several ranks) and
    // replayed; the host polls the exit flag between batches, the kernels of an overshooting batch return on `done`.
    // check_every must be EVEN: a V-cycle swaps the two sweep buffers of every vertex level (host pointers), and a graph
    // bakes pointers in — after an even number of cycles the host state is back where the capture started.
    const int check_every = 2;
    cudaGraph_t graph = nullptr; cudaGraphExec_t gexec = nullptr;
    if (!getenv("PB_PCG_NO_GRAPH")) {
        CK(cudaStreamBeginCapture(sm, cudaStreamCaptureModeThreadLocal));
        pb_status cst = PB_OK;
        for (int k = 0; k < check_every && cst == PB_OK; k++) cst = iteration(false);
        cudaError_t ce = cudaStreamEndCapture(sm, &graph);      // always closes the capture, also on the error path
        if (cst != PB_OK) { if (graph) cudaGraphDestroy(graph); return cst; }
        CK(ce);
        cudaError_t ie = cudaGraphInstantiate(&gexec, graph, 0);
        if (ie != cudaSuccess) { cudaGraphDestroy(graph); return fail(ctx, PB_ERR_CUDA, "pb_pcg: cudaGraphInstantiate failed: %s", cudaGetErrorString(ie)); }
    }
    pb_status lst = PB_OK;
    while (lst == PB_OK) {
        if (gexec) { cudaError_t le = cudaGraphLaunch(gexec, sm); if (le != cudaSuccess) { lst = fail(ctx, PB_ERR_CUDA, "pb_pcg: cudaGraphLaunch failed: %s", cudaGetErrorString(le)); break; } }
        else for (int k = 0; k < check_every && lst == PB_OK; k++) lst = iteration(false);
        if (lst != PB_OK) break;
        cudaError_t me = cudaMemcpyAsync(&hs, ctx->cg, sizeof hs, cudaMemcpyDeviceToHost, sm);
        if (me == cudaSuccess) me = cudaStreamSynchronize(sm);
        if (me != cudaSuccess) { lst = fail(ctx, PB_ERR_CUDA, "pb_pcg: %s", cudaGetErrorString(me)); break; }
        if (hs.done) break;
    }
    if (gexec) { cudaGraphExecDestroy(gexec); cudaGraphDestroy(graph); }
    if (lst != PB_OK) return lst;
    CK(cudaEventRecord(ctx->ev[1], sm)); CK(cudaStreamSynchronize(sm));
    float ms = 0; cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]); ctx->ms_pcg = ms;
    ctx->hist_n = ctx->hist ? std::min<long long>(hs.iters + 2, ctx->hist_cap) : 0;
    if (res) {
        res->iterations = hs.iters; res->relative_residual = hs.relres; res->exit_reason = hs.reason;
        res->seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    }
    return PB_OK;
}

// ---- inspection of the preconditioner (tests, bench): levels, smoother range, one application
extern "C" pb_status pb_mg_info(pb_ctx* ctx, int* n_levels, double* lmax, double* setup_ms) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->assembled, PB_ERR_STATE, "pb_mg_info: nothing assembled");
    CK(cudaSetDevice(ctx->device));
    { pb_status st = ensure_solver(ctx); if (st != PB_OK) return st; }
    CK(cudaEventRecord(ctx->ev[0], ctx->stream));
    { pb_status st = mg_setup(ctx); if (st != PB_OK) return st; }
    CK(cudaEventRecord(ctx->ev[1], ctx->stream)); CK(cudaStreamSynchronize(ctx->stream));
    float ms = 0; cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]);
    if (n_levels) *n_levels = ctx->mg.nlev; if (lmax) *lmax = ctx->mg.lmax; if (setup_ms) *setup_ms = ms;
    return PB_OK;
}
extern "C" pb_status pb_mg_level(pb_ctx* ctx, int level, long long dims[6], double* stencil9, double* elements16) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->mg.ready && level >= 0 && level < ctx->mg.nlev && dims, PB_ERR_STATE, "pb_mg_level: no such level (call pb_mg_info first)");
    CK(cudaSetDevice(ctx->device));
    const MgLevelData& D = ctx->mg.lev[level]; const MgGrid& L = D.L;
    dims[0] = L.Nx; dims[1] = L.Ny; dims[2] = L.row0; dims[3] = L.nrows; dims[4] = L.crow0; dims[5] = L.ncrows;
    if (stencil9) CK(cudaMemcpy(stencil9, D.st, (size_t)L.nrows * (L.Nx + 1) * 72, cudaMemcpyDeviceToHost));
    if (elements16) CK(cudaMemcpy(elements16, D.E, (size_t)L.ncrows * L.Nx * 128, cudaMemcpyDeviceToHost));
    return PB_OK;
}
extern "C" pb_status pb_mg_apply(pb_ctx* ctx, const double* r, double* z) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->assembled && r && z, PB_ERR_STATE, "pb_mg_apply: nothing assembled");
    DRAIN();
    CK(cudaSetDevice(ctx->device));
    { pb_status st = ensure_solver(ctx); if (st != PB_OK) return st; }
    { pb_status st = mg_setup(ctx); if (st != PB_OK) return st; }
    const SpmvPlan plan = make_spmv_plan(ctx);
    CK(cudaMemcpyAsync(ctx->r, r, (size_t)ctx->nrows * 8, cudaMemcpyHostToDevice, ctx->stream));
    pcg_init_state_kernel<<<1, 32, 0, ctx->stream>>>(0, ctx->parts, ctx->parts, ctx->cg, nullptr);
    { pb_status st = mg_apply(ctx, plan, ctx->r, mg_smoothing_steps(ctx)); if (st != PB_OK) return st; }
    CK(cudaMemcpyAsync(z, ctx->mg.zfull + plan.down, (size_t)ctx->nrows * 8, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return PB_OK;
}
