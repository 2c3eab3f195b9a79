// C ABI, obstacle problem (included by pb_api.cu): the active-set loop of apps/obstacle/obstacle.cpp:119-198, one
// entry point per outer iteration so that the driver keeps the reference's loop, stopping test and prints.

static double sum_parts(pb_ctx* ctx, const double* dparts, int n, cudaError_t* err) {
    std::vector<double> h(n);
    *err = cudaMemcpyAsync(h.data(), dparts, (size_t)n * 8, cudaMemcpyDeviceToHost, ctx->stream);
    if (*err == cudaSuccess) *err = cudaStreamSynchronize(ctx->stream);
    double t = 0.0; for (double v : h) t += v;
    return t;
}

extern "C" pb_status pb_obstacle_begin(pb_ctx* ctx, int face_degree, int rhs_degree_increase) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->have_mesh, PB_ERR_STATE, "pb_obstacle_begin: no mesh");
    NEED(face_degree == 0 || face_degree == 1, PB_ERR_INVALID, "pb_obstacle_begin: the reference supports face degree 0 or 1 (obstacle.cpp:249-253)");
    NEED(ctx->strip.cf_begin == 0 && ctx->strip.cf_end == num_free_faces(ctx->grid), PB_ERR_UNSUPPORTED, "pb_obstacle_begin: whole-mesh contexts only");
    pb_assembly_params p{};
    p.cell_degree = 0; p.face_degree = face_degree; p.stab_kind = PB_STAB_FANCY; p.problem = PB_PROBLEM_HHO; p.fun = PB_FUN_OBSTACLE;
    p.rhs_degree_increase = rhs_degree_increase; p.eta = 0.0; p.keep_local = 1;
    pb_status st = pb_assemble(ctx, &p); if (st != PB_OK) return st;      // local operators once; lc and f_T stay resident
    const long long nc = ctx->ncells_local;
    CK(dalloc(ctx, &ctx->ob_alpha, (size_t)nc)); CK(dalloc(ctx, &ctx->ob_beta, (size_t)nc)); CK(dalloc(ctx, &ctx->ob_gamma, (size_t)nc));
    CK(dalloc(ctx, &ctx->ob_inA, (size_t)nc)); CK(dalloc(ctx, &ctx->ob_xprev, (size_t)ctx->nrows));
    CK(cudaMemsetAsync(ctx->ob_xprev, 0, (size_t)ctx->nrows * 8, ctx->stream));
    Grid g = ctx->grid; g.row_begin = ctx->strip.cell_row0;
    obstacle_init_kernel<<<148 * 4, 256, 0, ctx->stream>>>(g, nc, ctx->points, ctx->ob_alpha, ctx->ob_beta, ctx->ob_gamma);
    CK(cudaGetLastError()); CK(cudaStreamSynchronize(ctx->stream));
    ctx->ob_ready = true;
    return PB_OK;
}

extern "C" pb_status pb_obstacle_iterate(pb_ctx* ctx, const pb_cg_params* cgp, double* delta_alpha_norm, long long* n_active, long long* cg_iterations) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->ob_ready && cgp, PB_ERR_STATE, "pb_obstacle_iterate: call pb_obstacle_begin first");
    CK(cudaSetDevice(ctx->device));
    const long long nc = ctx->ncells_local; const int fd = ctx->FD;
    Grid g = ctx->grid; g.row_begin = ctx->strip.cell_row0;
    const int B = 128, G = 148 * 8;
    int* d_act = nullptr; double* parts = nullptr;
    struct Cleanup { int** a; double** b; ~Cleanup() { if (*a) cudaFree(*a); if (*b) cudaFree(*b); } } cleanup{&d_act, &parts};      // also on the error paths
    CK(cudaMalloc((void**)&d_act, 4)); CK(cudaMalloc((void**)&parts, (size_t)G * 8));
    CK(cudaMemsetAsync(d_act, 0, 4, ctx->stream));
    if (fd == 0) obstacle_condense_kernel<0><<<G, B, 0, ctx->stream>>>(g, nc, ctx->points, ctx->lc, ctx->rhsloc, ctx->ob_alpha, ctx->ob_beta, ctx->ob_gamma, 1.0, ctx->prm.fun, ctx->ob_inA, ctx->S, ctx->g, d_act);
    else obstacle_condense_kernel<1><<<G, B, 0, ctx->stream>>>(g, nc, ctx->points, ctx->lc, ctx->rhsloc, ctx->ob_alpha, ctx->ob_beta, ctx->ob_gamma, 1.0, ctx->prm.fun, ctx->ob_inA, ctx->S, ctx->g, d_act);
    CK(cudaGetLastError());
    { pb_status st = scatter_system(ctx); if (st != PB_OK) return st; }
    ctx->asm_gen++;      // new system values: cached preconditioner data is stale
    pb_cg_result res{};
    { pb_status st = pb_pcg(ctx, cgp, &res); if (st != PB_OK) return st; }
    if (fd == 0) obstacle_update_kernel<0><<<G, B, 0, ctx->stream>>>(g, nc, ctx->points, ctx->lc, ctx->rhsloc, ctx->x, ctx->ob_inA, ctx->ob_gamma, ctx->prm.fun, ctx->ob_alpha, ctx->ob_beta, parts);
    else obstacle_update_kernel<1><<<G, B, 0, ctx->stream>>>(g, nc, ctx->points, ctx->lc, ctx->rhsloc, ctx->x, ctx->ob_inA, ctx->ob_gamma, ctx->prm.fun, ctx->ob_alpha, ctx->ob_beta, parts);
    CK(cudaGetLastError());
    cudaError_t e; double d2 = sum_parts(ctx, parts, G, &e); CK(e);
    obstacle_face_delta_kernel<<<G, B, 0, ctx->stream>>>(ctx->nrows, ctx->x, ctx->ob_xprev, parts);
    CK(cudaGetLastError());
    d2 += sum_parts(ctx, parts, G, &e); CK(e);
    int act = 0; CK(cudaMemcpy(&act, d_act, 4, cudaMemcpyDeviceToHost));
    // the Dirichlet-face entries of alpha change only in the first outer iteration (0 -> projection), where the
    // reference cannot stop either; they are left out of the norm
    if (delta_alpha_norm) *delta_alpha_norm = std::sqrt(d2);
    if (n_active) *n_active = act;
    if (cg_iterations) *cg_iterations = res.iterations;
    NEED(res.exit_reason == PB_CG_CONVERGED, PB_ERR_STATE, "pb_obstacle_iterate: the inner PCG did not converge");
    return PB_OK;
}

extern "C" pb_status pb_obstacle_error(pb_ctx* ctx, double* error_squared) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->ob_ready && ctx->x && error_squared, PB_ERR_STATE, "pb_obstacle_error: no solution");
    CK(cudaSetDevice(ctx->device));
    const long long nc = ctx->ncells_local; const int fd = ctx->FD, qk = ctx->quad_kind;
    Grid g = ctx->grid; g.row_begin = ctx->strip.cell_row0;
    const int B = 128, G = 148 * 8;
    double* parts = nullptr; CK(cudaMalloc((void**)&parts, (size_t)G * 8));
#define PB_OE(FD_, QK_) obstacle_error_kernel<FD_, QK_><<<G, B, 0, ctx->stream>>>(g, nc, ctx->points, ctx->lc, ctx->ob_alpha, ctx->x, ctx->prm.fun, parts)
    if (fd == 0 && qk == 0) PB_OE(0, 0); else if (fd == 1 && qk == 0) PB_OE(1, 0); else if (fd == 0) PB_OE(0, 1); else PB_OE(1, 1);
#undef PB_OE
    CK(cudaGetLastError());
    cudaError_t e; *error_squared = sum_parts(ctx, parts, G, &e); CK(e);
    cudaFree(parts);
    return PB_OK;
}

extern "C" pb_status pb_obstacle_get(pb_ctx* ctx, double* alpha_cells, double* beta, unsigned char* in_A) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->ob_ready, PB_ERR_STATE, "pb_obstacle_get: call pb_obstacle_begin first");
    CK(cudaSetDevice(ctx->device));
    const long long nc = ctx->ncells_local;
    if (alpha_cells) CK(cudaMemcpy(alpha_cells, ctx->ob_alpha, (size_t)nc * 8, cudaMemcpyDeviceToHost));
    if (beta) CK(cudaMemcpy(beta, ctx->ob_beta, (size_t)nc * 8, cudaMemcpyDeviceToHost));
    if (in_A) CK(cudaMemcpy(in_A, ctx->ob_inA, (size_t)nc, cudaMemcpyDeviceToHost));
    return PB_OK;
}
