// C ABI, cutHHO part (included by pb_api.cu): level set, mesh preprocessing on the device, upload / download of the
// cut user data, launch of K2.

extern "C" pb_status pb_cut_levelset(pb_ctx* ctx, int kind, double radius, double cx, double cy, double cut_y) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(kind == 0 || kind == 1, PB_ERR_INVALID, "pb_cut_levelset: kind must be 0 (circle) or 1 (line)");
    ctx->ls.kind = kind; ctx->ls.radius = radius; ctx->ls.cx = cx; ctx->ls.cy = cy; ctx->ls.cut_y = cut_y;
    ctx->have_ls = true; ctx->assembled = false; ctx->ifc_struct_ok = false;
    return PB_OK;
}

// rows of the mesh this context needs cut data for.  Cells: its local cells (own rows + the ghost row below).  Faces: the face
// rows of those cells and the row above them (top faces of the last cell row; a face row holds the bottom and vertical faces of a
// cell row).  Points: the vertex rows of those faces.  A whole-mesh context gets the whole mesh.
static void cut_window(pb_ctx* ctx) {
    const Grid& g = ctx->grid;
    const long long r0 = ctx->strip.cell_row0, r1 = g.row_end;
    ctx->cw0 = r0 * g.Nx; ctx->cw1 = r1 * g.Nx;
    ctx->fw0 = r0 * (2 * g.Nx + 1); ctx->fw1 = r1 < g.Ny ? (r1 + 1) * (2 * g.Nx + 1) : ctx->nfaces;
    ctx->pw0 = r0 * (g.Nx + 1); ctx->pw1 = std::min(r1 + 2, g.Ny + 1) * (g.Nx + 1);
}
static pb_status alloc_cut_arrays(pb_ctx* ctx) {
    cut_window(ctx);
    CK(dalloc(ctx, &ctx->node_loc_a, (size_t)(ctx->pw1 - ctx->pw0))); CK(dalloc(ctx, &ctx->face_loc_a, (size_t)(ctx->fw1 - ctx->fw0)));
    CK(dalloc(ctx, &ctx->face_ip_a, (size_t)(ctx->fw1 - ctx->fw0))); CK(dalloc(ctx, &ctx->cell_loc_a, (size_t)(ctx->cw1 - ctx->cw0)));
    ctx->node_loc = ctx->node_loc_a - ctx->pw0; ctx->face_loc = ctx->face_loc_a - ctx->fw0;
    ctx->face_ip = ctx->face_ip_a - ctx->fw0; ctx->cell_loc_g = ctx->cell_loc_a - ctx->cw0;
    return PB_OK;
}
// cut cells of the OWNED rows among the (ascending) list of the window: everything from the first cell of row_begin on
static pb_status count_owned_cut(pb_ctx* ctx) {
    ctx->ncut_owned = 0;
    if (ctx->ncut == 0) return PB_OK;
    std::vector<int> h((size_t)ctx->ncut);
    CK(cudaMemcpy(h.data(), ctx->cut_cells, (size_t)ctx->ncut * 4, cudaMemcpyDeviceToHost));
    const long long first = ctx->grid.row_begin * ctx->grid.Nx;
    ctx->ncut_owned = (int)(h.end() - std::lower_bound(h.begin(), h.end(), (int)first));
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
    NEED(refine_steps >= 0, PB_ERR_INVALID, "pb_cut_preprocess: negative refine_steps");
    NEED(refine_steps <= 6, PB_ERR_UNSUPPORTED, "pb_cut_preprocess: refine_steps > 6 is not built (the cut-cell kernels hold the 2^r + 1 interface points of a cell in shared memory; the reference's -r is uncapped)");
    CK(cudaSetDevice(ctx->device));
    free_cut(ctx);
    pb_status st = alloc_cut_arrays(ctx); if (st != PB_OK) return st;
    const Grid& g = ctx->grid; const LevelSet L = ctx->ls;
    cudaStream_t sm = ctx->stream;
    // windows (global ids).  The node displacement of move_nodes gathers from the faces around a node, so the FIRST face pass and
    // the displacement pass run one face row further out on both sides; everything else runs on the windows of cut_window().
    const long long pw0 = ctx->pw0, pw1 = ctx->pw1, fw0 = ctx->fw0, fw1 = ctx->fw1, cw0 = ctx->cw0, cw1 = ctx->cw1;
    const long long r0 = ctx->strip.cell_row0, r1 = g.row_end;
    const long long fh0 = (r0 > 0 ? r0 - 1 : 0) * (2 * g.Nx + 1), fh1 = r1 + 2 <= g.Ny ? (r1 + 2) * (2 * g.Nx + 1) : ctx->nfaces;      // first-pass faces
    const long long np = pw1 - pw0, nc = cw1 - cw0, nfh = fh1 - fh0;
    // all temporaries in ONE allocation (eight ranks allocating and freeing a dozen arrays each contend in the driver: the call took
    // 16 ms on 8 GPUs against 7 ms on one), released on every exit, error paths included
    size_t tmpb = 0;
    CK(cub::DeviceScan::ExclusiveSum(nullptr, tmpb, (int*)nullptr, (int*)nullptr, (int)(nc + 1), sm));
    auto up = [](size_t b) { return (b + 255) & ~(size_t)255; };
    const size_t o_disp = 0, o_displaced = o_disp + up((size_t)np * 16), o_iscut = o_displaced + up((size_t)np * 4), o_slot = o_iscut + up((size_t)(nc + 1) * 4),
                 o_err = o_slot + up((size_t)(nc + 1) * 4), o_p01 = o_err + 256, o_floc = o_p01 + up((size_t)nc * 32), o_fip = o_floc + up((size_t)nfh * 4),
                 o_tmp = o_fip + up((size_t)nfh * 16), total = o_tmp + up(tmpb);
    unsigned char* pool = nullptr;
    struct Cleanup { unsigned char** q; ~Cleanup() { if (*q) { cudaFree(*q); *q = nullptr; } } } cleanup{&pool};
    CK(cudaMalloc((void**)&pool, total));
    double* disp = (double*)(pool + o_disp); int* displaced = (int*)(pool + o_displaced); int* is_cut = (int*)(pool + o_iscut); int* slot = (int*)(pool + o_slot);
    int* err = (int*)(pool + o_err); double2* p01 = (double2*)(pool + o_p01); int* floc1 = (int*)(pool + o_floc); double2* fip1 = (double2*)(pool + o_fip);
    void* tmp = pool + o_tmp;
    CK(cudaMemsetAsync(pool, 0, o_p01, sm));      // disp, displaced, is_cut, slot, err
    // the temporaries are indexed by global ids too: shift them back by their window start
    double* disp_g = disp - 2 * pw0; int* displaced_g = displaced - pw0; int* is_cut_g = is_cut - cw0; int* slot_g = slot - cw0; double2* p01_g = p01 - 2 * cw0;
    int* floc1_g = floc1 - fh0; double2* fip1_g = fip1 - fh0;
    const int B = 256, G = 148 * 8;
    cut_node_position_kernel<<<G, B, 0, sm>>>(L, pw0, pw1, ctx->points, ctx->node_loc);
    cut_faces_kernel<<<G, B, 0, sm>>>(L, g, fh0, fh1, ctx->points, floc1_g, fip1_g);
    cut_displacements_kernel<<<G, B, 0, sm>>>(g, fh0, fh1, pw0, pw1, ctx->points, floc1_g, fip1_g, disp_g, displaced_g);
    cut_apply_displacements_kernel<<<G, B, 0, sm>>>(pw0, pw1, ctx->points, disp_g, displaced_g);
    cut_faces_kernel<<<G, B, 0, sm>>>(L, g, fw0, fw1, ctx->points, ctx->face_loc, ctx->face_ip);          // cuthho_square.cpp:2048
    cut_cells_kernel<<<G, B, 0, sm>>>(L, g, cw0, cw1, ctx->points, displaced_g, ctx->face_loc, ctx->face_ip, ctx->cell_loc_g, p01_g, is_cut_g, err);
    CK(cudaGetLastError());
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
            cut_compact_kernel<<<G, B, 0, sm>>>(cw0, cw1, is_cut_g, slot_g, ctx->cut_cells);
            cut_refine_kernel<<<(ncut + 127) / 128, 128, 0, sm>>>(L, g, ncut, ctx->NI, ctx->cut_cells, ctx->points, p01_g, ctx->iface, err);
            CK(cudaGetLastError());
            CK(cudaMemcpyAsync(&herr, err, 4, cudaMemcpyDeviceToHost, sm));
            CK(cudaStreamSynchronize(sm));
            if (herr) rc = fail(ctx, PB_ERR_INVALID, "pb_cut_preprocess: %s", cut_error_text(herr));
        }
        if (rc == PB_OK) rc = count_owned_cut(ctx);
    }
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
    std::vector<int> cells; std::vector<double> pts; int NI = 0;
    for (long long c = ctx->cw0; c < ctx->cw1; c++) {      // the caller passes whole-mesh arrays; a strip context keeps its window
        const long long n = iface_off[c + 1] - iface_off[c];
        if (cell_location[c] != PB_ON_INTERFACE) continue;
        if (NI == 0) NI = (int)n;
        NEED(n == NI && n >= 2, PB_ERR_INVALID, "pb_cut_upload: every cut cell must carry the same number (>= 2) of interface points");
        NEED(n <= 65, PB_ERR_UNSUPPORTED, "pb_cut_upload: more than 65 interface points per cell (refine_steps > 6) are not built");
        cells.push_back((int)c);
        pts.insert(pts.end(), iface_xy + 2 * iface_off[c], iface_xy + 2 * iface_off[c + 1]);
    }
    ctx->ncut = (int)cells.size(); ctx->NI = NI;
    CK(cudaMemcpy(ctx->node_loc_a, node_location + ctx->pw0, (size_t)(ctx->pw1 - ctx->pw0) * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(ctx->face_loc_a, face_location + ctx->fw0, (size_t)(ctx->fw1 - ctx->fw0) * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(ctx->face_ip_a, face_ip_xy + 2 * ctx->fw0, (size_t)(ctx->fw1 - ctx->fw0) * 16, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(ctx->cell_loc_a, cell_location + ctx->cw0, (size_t)(ctx->cw1 - ctx->cw0) * 4, cudaMemcpyHostToDevice));
    CK(dalloc(ctx, &ctx->cut_cells, cells.size())); CK(dalloc(ctx, &ctx->iface, pts.size() / 2));
    if (!cells.empty()) {
        CK(cudaMemcpy(ctx->cut_cells, cells.data(), cells.size() * 4, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(ctx->iface, pts.data(), pts.size() * 8, cudaMemcpyHostToDevice));
    }
    { pb_status st = count_owned_cut(ctx); if (st != PB_OK) return st; }
    ctx->have_cut = true; ctx->assembled = false; ctx->ifc_struct_ok = false;
    return PB_OK;
}

extern "C" pb_status pb_cut_download(pb_ctx* ctx, int* node_location, int* face_location, double* face_ip_xy, int* cell_location,
                                     long long* n_cut_cells, int* n_iface_pts, double* iface_xy) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->have_cut, PB_ERR_STATE, "pb_cut_download: no cut data");
    CK(cudaSetDevice(ctx->device));
    // whole-mesh host arrays; outside the rows this context holds: PB_UNDEF / zero points
    if (node_location) { std::fill(node_location, node_location + ctx->npoints, (int)PB_UNDEF); CK(cudaMemcpy(node_location + ctx->pw0, ctx->node_loc_a, (size_t)(ctx->pw1 - ctx->pw0) * 4, cudaMemcpyDeviceToHost)); }
    if (face_location) { std::fill(face_location, face_location + ctx->nfaces, (int)PB_UNDEF); CK(cudaMemcpy(face_location + ctx->fw0, ctx->face_loc_a, (size_t)(ctx->fw1 - ctx->fw0) * 4, cudaMemcpyDeviceToHost)); }
    if (face_ip_xy) { std::fill(face_ip_xy, face_ip_xy + 2 * ctx->nfaces, 0.0); CK(cudaMemcpy(face_ip_xy + 2 * ctx->fw0, ctx->face_ip_a, (size_t)(ctx->fw1 - ctx->fw0) * 16, cudaMemcpyDeviceToHost)); }
    if (cell_location) { std::fill(cell_location, cell_location + ctx->ncells_global, (int)PB_UNDEF); CK(cudaMemcpy(cell_location + ctx->cw0, ctx->cell_loc_a, (size_t)(ctx->cw1 - ctx->cw0) * 4, cudaMemcpyDeviceToHost)); }
    if (n_cut_cells) *n_cut_cells = ctx->ncut;
    if (n_iface_pts) *n_iface_pts = ctx->NI;
    if (iface_xy && ctx->ncut > 0) CK(cudaMemcpy(iface_xy, ctx->iface, (size_t)ctx->ncut * ctx->NI * 16, cudaMemcpyDeviceToHost));   // cut list is in ascending cell order
    return PB_OK;
}

extern "C" pb_status pb_cut_counts(pb_ctx* ctx, long long* n_window, long long* n_owned) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->have_cut, PB_ERR_STATE, "pb_cut_counts: no cut data");
    if (n_window) *n_window = ctx->ncut; if (n_owned) *n_owned = ctx->ncut_owned;
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
