// C ABI, cutHHO part (included by pb_api.cu): level set, mesh preprocessing on the device, upload / download of the
// cut user data, launch of K2.

extern "C" pb_status pb_cut_levelset(pb_ctx* ctx, int kind, double radius, double cx, double cy, double cut_y) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(kind == 0 || kind == 1, PB_ERR_INVALID, "pb_cut_levelset: kind must be 0 (circle) or 1 (line)");
    ctx->ls.kind = kind; ctx->ls.radius = radius; ctx->ls.cx = cx; ctx->ls.cy = cy; ctx->ls.cut_y = cut_y;
    ctx->have_ls = true; ctx->assembled = false; ctx->ifc_struct_ok = false;
    return PB_OK;
}

static pb_status alloc_cut_arrays(pb_ctx* ctx) {
    CK(dalloc(ctx, &ctx->node_loc, (size_t)ctx->npoints)); CK(dalloc(ctx, &ctx->face_loc, (size_t)ctx->nfaces));
    CK(dalloc(ctx, &ctx->face_ip, (size_t)ctx->nfaces)); CK(dalloc(ctx, &ctx->cell_loc_g, (size_t)ctx->ncells_global));
    return PB_OK;
}

static const char* cut_error_text(int e) {
    if (e & CUTERR_TOO_MANY_CUTS) return "more than two cut faces in a cell (cuthho_geom.hpp:290, cut_faces.at(k++))";
    if (e & CUTERR_BAD_CUT_COUNT) return "invalid number of cuts in cell (cuthho_geom.hpp:336)";
    if (e & CUTERR_CONCAVE) return "move_nodes produced a concave cell (cuthho_geom.hpp:538)";
    return "interface not found in search range (cuthho_geom.hpp:648)";
}

extern "C" pb_status pb_cut_preprocess(pb_ctx* ctx, int refine_steps) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->have_mesh && ctx->have_ls, PB_ERR_STATE, "pb_cut_preprocess: needs a mesh and pb_cut_levelset");
    NEED(refine_steps >= 0 && refine_steps <= 6, PB_ERR_INVALID, "pb_cut_preprocess: refine_steps must be in [0, 6]");
    CK(cudaSetDevice(ctx->device));
    free_cut(ctx);
    pb_status st = alloc_cut_arrays(ctx); if (st != PB_OK) return st;
    const Grid& g = ctx->grid; const LevelSet L = ctx->ls;
    const long long np = ctx->npoints, nf = ctx->nfaces, nc = ctx->ncells_global;
    cudaStream_t sm = ctx->stream;
    double* disp = nullptr; int *displaced = nullptr, *is_cut = nullptr, *slot = nullptr, *err = nullptr; double2* p01 = nullptr;
    CK(cudaMalloc((void**)&disp, (size_t)np * 16)); CK(cudaMalloc((void**)&displaced, (size_t)np * 4));
    CK(cudaMalloc((void**)&is_cut, (size_t)(nc + 1) * 4)); CK(cudaMalloc((void**)&slot, (size_t)(nc + 1) * 4));
    CK(cudaMalloc((void**)&err, 4)); CK(cudaMalloc((void**)&p01, (size_t)nc * 32));
    CK(cudaMemsetAsync(disp, 0, (size_t)np * 16, sm)); CK(cudaMemsetAsync(displaced, 0, (size_t)np * 4, sm));
    CK(cudaMemsetAsync(err, 0, 4, sm)); CK(cudaMemsetAsync(is_cut, 0, (size_t)(nc + 1) * 4, sm));
    const int B = 256, G = 148 * 8;
    cut_node_position_kernel<<<G, B, 0, sm>>>(L, np, ctx->points, ctx->node_loc);
    cut_faces_kernel<<<G, B, 0, sm>>>(L, g, nf, ctx->points, ctx->face_loc, ctx->face_ip);
    cut_displacements_kernel<<<G, B, 0, sm>>>(g, nf, ctx->points, ctx->face_loc, ctx->face_ip, disp, displaced);
    cut_apply_displacements_kernel<<<G, B, 0, sm>>>(np, ctx->points, disp, displaced);
    cut_faces_kernel<<<G, B, 0, sm>>>(L, g, nf, ctx->points, ctx->face_loc, ctx->face_ip);          // cuthho_square.cpp:2048
    cut_cells_kernel<<<G, B, 0, sm>>>(L, g, nc, ctx->points, displaced, ctx->face_loc, ctx->face_ip, ctx->cell_loc_g, p01, is_cut, err);
    CK(cudaGetLastError());
    void* tmp = nullptr; size_t tmpb = 0;
    CK(cub::DeviceScan::ExclusiveSum(nullptr, tmpb, is_cut, slot, (int)(nc + 1), sm));
    CK(cudaMalloc(&tmp, tmpb));
    CK(cub::DeviceScan::ExclusiveSum(tmp, tmpb, is_cut, slot, (int)(nc + 1), sm));
    int ncut = 0, herr = 0;
    CK(cudaMemcpyAsync(&ncut, slot + nc, 4, cudaMemcpyDeviceToHost, sm));
    CK(cudaMemcpyAsync(&herr, err, 4, cudaMemcpyDeviceToHost, sm));
    CK(cudaStreamSynchronize(sm));
    pb_status rc = PB_OK;
    if (herr) rc = fail(ctx, PB_ERR_INVALID, "pb_cut_preprocess: %s", cut_error_text(herr));
    if (rc == PB_OK) {
        ctx->ncut = ncut; ctx->NI = (1 << refine_steps) + 1;
        CK(dalloc(ctx, &ctx->cut_cells, (size_t)ncut)); CK(dalloc(ctx, &ctx->iface, (size_t)ncut * ctx->NI));
        if (ncut > 0) {
            cut_compact_kernel<<<G, B, 0, sm>>>(nc, is_cut, slot, ctx->cut_cells);
            cut_refine_kernel<<<(ncut + 127) / 128, 128, 0, sm>>>(L, g, ncut, ctx->NI, ctx->cut_cells, ctx->points, p01, ctx->iface, err);
            CK(cudaGetLastError());
            CK(cudaMemcpyAsync(&herr, err, 4, cudaMemcpyDeviceToHost, sm));
            CK(cudaStreamSynchronize(sm));
            if (herr) rc = fail(ctx, PB_ERR_INVALID, "pb_cut_preprocess: %s", cut_error_text(herr));
        }
    }
    cudaFree(tmp); cudaFree(disp); cudaFree(displaced); cudaFree(is_cut); cudaFree(slot); cudaFree(err); cudaFree(p01);
    ctx->have_cut = rc == PB_OK; ctx->assembled = false; ctx->ifc_struct_ok = false;
    return rc;
}

extern "C" pb_status pb_cut_upload(pb_ctx* ctx, const int* node_location, const int* face_location, const double* face_ip_xy,
                                   const int* cell_location, const long long* iface_off, const double* iface_xy) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->have_mesh, PB_ERR_STATE, "pb_cut_upload: no mesh");
    NEED(node_location && face_location && face_ip_xy && cell_location && iface_off && iface_xy, PB_ERR_INVALID, "pb_cut_upload: null array");
    CK(cudaSetDevice(ctx->device));
    free_cut(ctx);
    pb_status st = alloc_cut_arrays(ctx); if (st != PB_OK) return st;
    const long long nc = ctx->ncells_global;
    std::vector<int> cells; std::vector<double> pts; int NI = 0;
    for (long long c = 0; c < nc; c++) {
        const long long n = iface_off[c + 1] - iface_off[c];
        if (cell_location[c] != PB_ON_INTERFACE) continue;
        if (NI == 0) NI = (int)n;
        NEED(n == NI && n >= 2, PB_ERR_INVALID, "pb_cut_upload: every cut cell must carry the same number (>= 2) of interface points");
        cells.push_back((int)c);
        pts.insert(pts.end(), iface_xy + 2 * iface_off[c], iface_xy + 2 * iface_off[c + 1]);
    }
    ctx->ncut = (int)cells.size(); ctx->NI = NI;
    CK(cudaMemcpy(ctx->node_loc, node_location, (size_t)ctx->npoints * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(ctx->face_loc, face_location, (size_t)ctx->nfaces * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(ctx->face_ip, face_ip_xy, (size_t)ctx->nfaces * 16, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(ctx->cell_loc_g, cell_location, (size_t)nc * 4, cudaMemcpyHostToDevice));
    CK(dalloc(ctx, &ctx->cut_cells, cells.size())); CK(dalloc(ctx, &ctx->iface, pts.size() / 2));
    if (!cells.empty()) {
        CK(cudaMemcpy(ctx->cut_cells, cells.data(), cells.size() * 4, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(ctx->iface, pts.data(), pts.size() * 8, cudaMemcpyHostToDevice));
    }
    ctx->have_cut = true; ctx->assembled = false; ctx->ifc_struct_ok = false;
    return PB_OK;
}

extern "C" pb_status pb_cut_download(pb_ctx* ctx, int* node_location, int* face_location, double* face_ip_xy, int* cell_location,
                                     long long* n_cut_cells, int* n_iface_pts, double* iface_xy) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->have_cut, PB_ERR_STATE, "pb_cut_download: no cut data");
    CK(cudaSetDevice(ctx->device));
    if (node_location) CK(cudaMemcpy(node_location, ctx->node_loc, (size_t)ctx->npoints * 4, cudaMemcpyDeviceToHost));
    if (face_location) CK(cudaMemcpy(face_location, ctx->face_loc, (size_t)ctx->nfaces * 4, cudaMemcpyDeviceToHost));
    if (face_ip_xy) CK(cudaMemcpy(face_ip_xy, ctx->face_ip, (size_t)ctx->nfaces * 16, cudaMemcpyDeviceToHost));
    if (cell_location) CK(cudaMemcpy(cell_location, ctx->cell_loc_g, (size_t)ctx->ncells_global * 4, cudaMemcpyDeviceToHost));
    if (n_cut_cells) *n_cut_cells = ctx->ncut;
    if (n_iface_pts) *n_iface_pts = ctx->NI;
    if (iface_xy && ctx->ncut > 0) CK(cudaMemcpy(iface_xy, ctx->iface, (size_t)ctx->ncut * ctx->NI * 16, cudaMemcpyDeviceToHost));   // cut list is in ascending cell order
    return PB_OK;
}

template <int CD, int FD>
static cudaError_t launch_cut_t(pb_ctx* c, const CutArgs& A) {
    const size_t smem = sizeof(CutSmem<CD, FD>);
    auto kern = local_cut_kernel<CD, FD>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int grid = A.ncut < 148 * 8 ? A.ncut : 148 * 8;
    kern<<<grid, 32, smem, c->stream>>>(A);
    return cudaGetLastError();
}

static pb_status launch_cut(pb_ctx* ctx, const LocalArgs& L, int cd, int fd) {
    CutArgs A{};
    A.grid = ctx->grid; A.ls = ctx->ls; A.points = ctx->points; A.node_loc = ctx->node_loc; A.face_loc = ctx->face_loc; A.face_ip = ctx->face_ip;
    A.cut_cells = ctx->cut_cells; A.iface = ctx->iface; A.ncut = ctx->ncut; A.NI = ctx->NI;
    A.cell_begin = ctx->strip.cell_row0 * ctx->grid.Nx; A.cell_end = A.cell_begin + ctx->ncells_local;
    A.fun = L.fun; A.eta = L.eta;
    A.S = L.S; A.g = L.g; A.recA = L.recA; A.recb = L.recb; A.lc = L.lc; A.oper = L.oper; A.rhs = L.rhs; A.data = L.data;
    cudaError_t e;
    if (cd == 1 && fd == 0) e = launch_cut_t<1, 0>(ctx, A);
    else if (cd == 2 && fd == 1) e = launch_cut_t<2, 1>(ctx, A);
    else if (cd == 3 && fd == 2) e = launch_cut_t<3, 2>(ctx, A);
    else return fail(ctx, PB_ERR_UNSUPPORTED, "pb_assemble: cut cells are built for HHO(k+1,k), k <= 2 (k = 3 has no valid reference answer: the "
                                              "Dunavant table has no rule for degree 8, SURVEY F4)");
    if (e != cudaSuccess) return fail(ctx, PB_ERR_CUDA, "local_cut_kernel launch failed: %s", cudaGetErrorString(e));
    return PB_OK;
}
