// C ABI — interface problem, after the solve: back-substitution in the layout of the reference's solution vector
// (cell unknowns with cell_table, cut cells twice: negative then positive part, cuthho_square.cpp:1143-1151, 1343-1365;
// then the face unknowns) and the driver's H1 error (run_cuthho_interface, :1769-1837).

// (cells [c_begin, c_end) = the owned cells; x is addressed by global unknown of the interface system, xreg by global compressed
//  face unknown: both through pointers shifted back by the first column this context holds)
__global__ void ifc_gather_regular_kernel(Grid g, int F, long long c_begin, long long c_end, const int* __restrict__ ftab, const double* __restrict__ x, double* __restrict__ xreg) {
    for (long long c = c_begin + blockIdx.x * (long long)blockDim.x + threadIdx.x; c < c_end; c += (long long)gridDim.x * blockDim.x) {
        const long long ci = c % g.Nx, cj = c / g.Nx;
        long long f[4]; cell_faces(g, ci, cj, f);
        for (int li = 0; li < 4; li++) {
            if (cell_face_dirichlet(g, ci, cj, li)) continue;
            const long long cf = cell_face_compressed(g, ci, cj, li);
            for (int a = 0; a < F; a++) xreg[cf * F + a] = x[(long long)ftab[f[li]] * F + a];     // first set (both cells write the same value)
        }
    }
}
// cut slots [slot0, ncut) = the owned cut cells of the window list
__global__ void ifc_recover_cut_kernel(Grid g, int F, int C, int slot0, int ncut, const int* __restrict__ cut_cells, const int* __restrict__ face_loc,
                                       const int* __restrict__ ftab, const double* __restrict__ recAx, const double* __restrict__ recbx,
                                       const double* __restrict__ x, double* __restrict__ ucut) {
    const int C2 = 2 * C, NF2 = 8 * F;
    for (long long t = (long long)slot0 * C2 + blockIdx.x * (long long)blockDim.x + threadIdx.x; t < (long long)ncut * C2; t += (long long)gridDim.x * blockDim.x) {
        const long long slot = t / C2; const int a = (int)(t - slot * C2);
        const long long c = cut_cells[slot], ci = c % g.Nx, cj = c / g.Nx;
        long long f[4]; cell_faces(g, ci, cj, f);
        double v = recbx[t];
        for (int col = 0; col < NF2; col++) v -= recAx[slot * (long long)(NF2 * C2) + a + col * C2] * x[ifc_slot_dof(g, F, face_loc, ftab, f, col)];
        ucut[t] = v;
    }
}
// cells [c_begin, c_end) = the owned cells, slot0 = first owned slot of the cut-cell list: the packed layout starts at the first owned cell
// (uc by global cell id through a shifted pointer, ucut by slot of the window list)
__global__ void ifc_pack_cells_kernel(Grid g, int C, long long c_begin, long long c_end, int slot0, int ncut, const int* __restrict__ cut_cells, const int* __restrict__ cell_loc,
                                      const double* __restrict__ uc, const double* __restrict__ ucut, double* __restrict__ out) {
    for (long long c = c_begin + blockIdx.x * (long long)blockDim.x + threadIdx.x; c < c_end; c += (long long)gridDim.x * blockDim.x) {
        int lo = slot0, hi = ncut;                  // cut cells before c = cell_table[c] - c
        while (lo < hi) { const int mid = (lo + hi) / 2; if (cut_cells[mid] < c) lo = mid + 1; else hi = mid; }
        double* o = out + ((c - c_begin) + (lo - slot0)) * C;
        if (cell_loc[c] == PB_ON_INTERFACE) for (int a = 0; a < 2 * C; a++) o[a] = ucut[(long long)lo * 2 * C + a];
        else for (int a = 0; a < C; a++) o[a] = uc[c * C + a];
    }
}

// H1 error^2 of the cell unknowns: uncut cells over integrate(msh, cl, 2 cd), cut cells over both parts.  One warp per
// cell, lane 0 accumulates in cell order: deterministic.
template <int CD>
__global__ void __launch_bounds__(32) interface_h1_kernel(Grid g, const double2* __restrict__ points, const int* __restrict__ node_loc,
                                                           const int* __restrict__ cell_loc, const int* __restrict__ cut_cells, const double2* __restrict__ iface,
                                                           int ncut, int NI, int fun, long long c_begin, long long c_end, const double* __restrict__ uc,
                                                           const double* __restrict__ ucut, double* __restrict__ parts) {
    // (cells [c_begin, c_end) = the owned cells; uc is addressed by global cell id through a shifted pointer)
    constexpr int C = cell_basis_size(CD);
    __shared__ double tp[2 * CUT_MAX_TP];
    __shared__ int s_ntp;
    const int lane = threadIdx.x;
    double sum = 0.0;
    for (long long c = c_begin + blockIdx.x; c < c_end; c += gridDim.x) {
        const long long ci = c % g.Nx, cj = c / g.Nx, p0 = grid_point(g, ci, cj);
        const long long pid[4] = {p0, p0 + 1, p0 + g.Nx + 2, p0 + g.Nx + 1};
        CellGeom geo; int nloc[4];
#pragma unroll
        for (int i = 0; i < 4; i++) { const double2 v = points[pid[i]]; geo.vx[i] = v.x; geo.vy[i] = v.y; nloc[i] = node_loc[pid[i]]; }
        cell_geometry(geo);
        const double ih = geo.ihalf;
        double acc = 0.0;
        auto point_term = [&](const double* u, double x, double y, double w) {
            double phi[C], dx[C], dy[C];
            eval_cell_basis<CD>((x - geo.barx) * ih, (y - geo.bary) * ih, ih, phi, dx, dy);
            double gx = 0, gy = 0;
#pragma unroll
            for (int i = 1; i < C; i++) { gx += u[i] * dx[i]; gy += u[i] * dy[i]; }
            double rgx, rgy; fun_grad(fun, x, y, rgx, rgy);
            acc += w * ((rgx - gx) * (rgx - gx) + (rgy - gy) * (rgy - gy));
        };
        if (cell_loc[c] != PB_ON_INTERFACE) {
            constexpr int nq = cell_qps(PB_QUAD_POLY_DUNAVANT, 2 * CD);
            for (int q = lane; q < nq; q += 32) { double x, y, w; cell_qp<PB_QUAD_POLY_DUNAVANT, 2 * CD>(geo, q, x, y, w); point_term(uc + c * C, x, y, w); }
        } else {
            int lo = 0, hi = ncut - 1;
            while (lo < hi) { const int mid = (lo + hi) / 2; if (cut_cells[mid] < c) lo = mid + 1; else hi = mid; }
            const double2* I = iface + (long long)lo * NI;
            for (int sd = 0; sd < 2; sd++) {
                const int WHERE = sd == 0 ? PB_IN_NEGATIVE_SIDE : PB_IN_POSITIVE_SIDE;
                __syncwarp();
                if (lane == 0) {
                    int n = 0;
                    auto push = [&](double x, double y) { tp[2 * n] = x; tp[2 * n + 1] = y; n++; };
                    auto push_iface = [&]() {
                        if (sd == 0) for (int i = 0; i < NI; i++) push(I[i].x, I[i].y);
                        else for (int i = NI - 1; i >= 0; i--) push(I[i].x, I[i].y);
                    };
                    const bool c1 = nloc[0] == WHERE && nloc[3] != WHERE, c2 = nloc[0] != WHERE && nloc[3] == WHERE, c3 = nloc[0] != WHERE && nloc[3] != WHERE;
                    if (c1 || c2 || c3) { for (int i = 0; i < 4; i++) if (nloc[i] == WHERE) push(sel4(geo.vx, i), sel4(geo.vy, i)); push_iface(); }
                    else {
                        int i = 0;
                        while (i < 4 && nloc[i] == WHERE) { push(sel4(geo.vx, i), sel4(geo.vy, i)); i++; }
                        push_iface();
                        while (i < 4 && nloc[i] != WHERE) i++;
                        while (i < 4 && nloc[i] == WHERE) { push(sel4(geo.vx, i), sel4(geo.vy, i)); i++; }
                    }
                    s_ntp = n;
                }
                __syncwarp();
                const int ntp = s_ntp;
                double rx = 0, ry = 0, den = 0; const double x0 = tp[0], y0 = tp[1];
                for (int i = 2; i < ntp; i++) {
                    const double ax = tp[2 * (i - 1)] - x0, ay = tp[2 * (i - 1) + 1] - y0, bx = tp[2 * i] - x0, by = tp[2 * i + 1] - y0;
                    const double d = (ax * by - ay * bx) / 2.0; rx += (ax + bx) * d; ry += (ay + by) * d; den += d;
                }
                const double pbx = x0 + rx / (den * 3), pby = y0 + ry / (den * 3);
                const int ri = dunavant_rule_index(2 * CD), np = c_qt.dun_n[ri], nq = ntp * np;
                for (int q = lane; q < nq; q += 32) {
                    const int t = q / np, p = q - t * np, t1 = (t + 1) % ntp;
                    const double ax = tp[2 * t], ay = tp[2 * t + 1], bx = tp[2 * t1], by = tp[2 * t1 + 1];
                    const double area = fabs(((ax - pbx) * (by - pby) - (ay - pby) * (bx - pbx)) / 2.0);
                    const double* row = c_qt.dun[ri][p];
                    point_term(ucut + (long long)lo * 2 * C + sd * C, pbx * row[0] + ax * row[1] + bx * row[2], pby * row[0] + ay * row[1] + by * row[2], area * row[3]);
                }
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
        sum += acc;
        __syncwarp();
    }
    if (lane == 0) parts[blockIdx.x] = sum;
}

static pb_status interface_cell_unknowns(pb_ctx* ctx, double** uc_out, double** ucut_out);
// rows "x y u_T" of the driver's interface_uT.dat (cuthho_square.cpp:1775-1833, gnuplot_output_object :754-780): cells in
// order, uncut cells over integrate(msh, cl, 2 cd), cut cells over the negative then the positive part.  A cut cell always has
// (4 + 2 NI) * np points (every vertex lies on exactly one side), so the row offsets are closed forms.  One warp per cell.
template <int CD>
__global__ void __launch_bounds__(32) interface_sample_kernel(Grid g, const double2* __restrict__ points, const int* __restrict__ node_loc,
                                                               const int* __restrict__ cell_loc, const int* __restrict__ cut_cells, const double2* __restrict__ iface,
                                                               int ncut, int NI, long long c_begin, long long c_end, int slot0, const double* __restrict__ uc,
                                                               const double* __restrict__ ucut, double* __restrict__ out) {
    // (cells [c_begin, c_end) = the owned cells, slot0 = the first owned slot of the cut-cell list: the rows start at the first owned cell;
    //  uc by global cell id through a shifted pointer)
    constexpr int C = cell_basis_size(CD);
    constexpr int nq0 = cell_qps(PB_QUAD_POLY_DUNAVANT, 2 * CD);
    __shared__ double tp[2 * CUT_MAX_TP];
    __shared__ int s_ntp;
    const int lane = threadIdx.x;
    const int ri = dunavant_rule_index(2 * CD), np = c_qt.dun_n[ri];
    const long long ncutq = (long long)(4 + 2 * NI) * np;
    for (long long c = c_begin + blockIdx.x; c < c_end; c += gridDim.x) {
        const long long ci = c % g.Nx, cj = c / g.Nx, p0 = grid_point(g, ci, cj);
        const long long pid[4] = {p0, p0 + 1, p0 + g.Nx + 2, p0 + g.Nx + 1};
        CellGeom geo; int nloc[4];
#pragma unroll
        for (int i = 0; i < 4; i++) { const double2 v = points[pid[i]]; geo.vx[i] = v.x; geo.vy[i] = v.y; nloc[i] = node_loc[pid[i]]; }
        cell_geometry(geo);
        const double ih = geo.ihalf;
        int lo = slot0, hi = ncut;                  // cut cells before c
        while (lo < hi) { const int mid = (lo + hi) / 2; if (cut_cells[mid] < c) lo = mid + 1; else hi = mid; }
        double* o = out + 3 * (((c - c_begin) - (lo - slot0)) * nq0 + (lo - slot0) * ncutq);
        auto emit = [&](const double* u, long long idx, double x, double y) {
            double phi[C], dx[C], dy[C];
            eval_cell_basis<CD>((x - geo.barx) * ih, (y - geo.bary) * ih, ih, phi, dx, dy);
            double v = 0.0;
#pragma unroll
            for (int i = 0; i < C; i++) v += u[i] * phi[i];
            o[3 * idx] = x; o[3 * idx + 1] = y; o[3 * idx + 2] = v;
        };
        if (cell_loc[c] != PB_ON_INTERFACE) {
            for (int q = lane; q < nq0; q += 32) { double x, y, w; cell_qp<PB_QUAD_POLY_DUNAVANT, 2 * CD>(geo, q, x, y, w); emit(uc + c * C, q, x, y); }
        } else {
            const double2* I = iface + (long long)lo * NI;
            long long done = 0;
            for (int sd = 0; sd < 2; sd++) {
                const int WHERE = sd == 0 ? PB_IN_NEGATIVE_SIDE : PB_IN_POSITIVE_SIDE;
                __syncwarp();
                if (lane == 0) {
                    int n = 0;
                    auto push = [&](double x, double y) { tp[2 * n] = x; tp[2 * n + 1] = y; n++; };
                    auto push_iface = [&]() {
                        if (sd == 0) for (int i = 0; i < NI; i++) push(I[i].x, I[i].y);
                        else for (int i = NI - 1; i >= 0; i--) push(I[i].x, I[i].y);
                    };
                    const bool c1 = nloc[0] == WHERE && nloc[3] != WHERE, c2 = nloc[0] != WHERE && nloc[3] == WHERE, c3 = nloc[0] != WHERE && nloc[3] != WHERE;
                    if (c1 || c2 || c3) { for (int i = 0; i < 4; i++) if (nloc[i] == WHERE) push(sel4(geo.vx, i), sel4(geo.vy, i)); push_iface(); }
                    else {
                        int i = 0;
                        while (i < 4 && nloc[i] == WHERE) { push(sel4(geo.vx, i), sel4(geo.vy, i)); i++; }
                        push_iface();
                        while (i < 4 && nloc[i] != WHERE) i++;
                        while (i < 4 && nloc[i] == WHERE) { push(sel4(geo.vx, i), sel4(geo.vy, i)); i++; }
                    }
                    s_ntp = n;
                }
                __syncwarp();
                const int ntp = s_ntp;
                double rx = 0, ry = 0, den = 0; const double x0 = tp[0], y0 = tp[1];
                for (int i = 2; i < ntp; i++) {
                    const double ax = tp[2 * (i - 1)] - x0, ay = tp[2 * (i - 1) + 1] - y0, bx = tp[2 * i] - x0, by = tp[2 * i + 1] - y0;
                    const double d = (ax * by - ay * bx) / 2.0; rx += (ax + bx) * d; ry += (ay + by) * d; den += d;
                }
                const double pbx = x0 + rx / (den * 3), pby = y0 + ry / (den * 3);
                const int nq = ntp * np;
                for (int q = lane; q < nq; q += 32) {
                    const int t = q / np, p = q - t * np, t1 = (t + 1) % ntp;
                    const double ax = tp[2 * t], ay = tp[2 * t + 1], bx = tp[2 * t1], by = tp[2 * t1 + 1];
                    const double* row = c_qt.dun[ri][p];
                    emit(ucut + (long long)lo * 2 * C + sd * C, done + q, pbx * row[0] + ax * row[1] + bx * row[2], pby * row[0] + ay * row[1] + by * row[2]);
                }
                done += nq;
            }
        }
        __syncwarp();
    }
}

static pb_status interface_sample(pb_ctx* ctx, double* xyv, long long capacity_points, long long* n_points) {      // rows of the OWNED cells
    const Grid& g = ctx->grid; const int CD = ctx->CD;
    const long long c_begin = g.row_begin * g.Nx, c_end = g.row_end * g.Nx, nc = c_end - c_begin;
    const int np = dunavant_npts(2 * CD), slot0 = ctx->ncut - ctx->ncut_owned;
    const long long nq0 = cell_qps(PB_QUAD_POLY_DUNAVANT, 2 * CD), ncutq = (long long)(4 + 2 * ctx->NI) * np;
    const long long total = (nc - ctx->ncut_owned) * nq0 + (long long)ctx->ncut_owned * ncutq;
    *n_points = total;
    if (!xyv) return PB_OK;
    NEED(capacity_points >= total, PB_ERR_INVALID, "pb_sample_solution: buffer too small");
    double *uc = nullptr, *ucut = nullptr, *out = nullptr;
    pb_status st = interface_cell_unknowns(ctx, &uc, &ucut); if (st != PB_OK) return st;
    CK(cudaMalloc((void**)&out, (size_t)total * 24));
    const int grid = (int)std::min<long long>(nc, 148 * 32);
#define PB_SMP(CD_) interface_sample_kernel<CD_><<<grid, 32, 0, ctx->stream>>>(g, ctx->points, ctx->node_loc, ctx->cell_loc_g, ctx->cut_cells, ctx->iface, ctx->ncut, ctx->NI, \
                                                                              c_begin, c_end, slot0, uc - c_begin * ctx->C, ucut, out)
    switch (CD) { case 1: PB_SMP(1); break; case 2: PB_SMP(2); break; default: PB_SMP(3); break; }
#undef PB_SMP
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(xyv, out, (size_t)total * 24, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(uc); cudaFree(ucut); cudaFree(out);
    if (e != cudaSuccess) return fail(ctx, PB_ERR_CUDA, "pb_sample_solution (interface): %s", cudaGetErrorString(e));
    return PB_OK;
}

// cell unknowns of the owned cells from the face solution ctx->x (uc: owned cells in order, *uc_out is the allocation; ucut:
// [u-, u+] per cut slot of the window list, filled for the owned cut cells).  Strips fetch the halo part of x first.
static pb_status interface_cell_unknowns(pb_ctx* ctx, double** uc_out, double** ucut_out) {
    const Grid& g = ctx->grid; const int F = ctx->F, C = ctx->C; const Strip& st = ctx->strip;
    const long long c_begin = g.row_begin * g.Nx, c_end = g.row_end * g.Nx, owned = c_end - c_begin;
    const long long nreg = (st.col_hi - st.col_lo) * F, nfull = (ctx->rb_hi - ctx->rb_lo) * F, down = (ctx->rb_begin - ctx->rb_lo) * F;
    double *xs = nullptr, *xreg = nullptr, *uc = nullptr, *ucut = nullptr;
    auto release = [&]() { cudaFree(xs); cudaFree(xreg); cudaFree(uc); cudaFree(ucut); };
#define IFP(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { release(); return fail(ctx, PB_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); } } while (0)
    IFP(cudaMalloc((void**)&xs, (size_t)nfull * 8)); IFP(cudaMalloc((void**)&xreg, (size_t)nreg * 8)); IFP(cudaMalloc((void**)&uc, (size_t)owned * C * 8));
    IFP(cudaMalloc((void**)&ucut, (size_t)(ctx->ncut > 0 ? ctx->ncut : 1) * 2 * C * 8));
    IFP(cudaMemsetAsync(xs, 0, (size_t)nfull * 8, ctx->stream)); IFP(cudaMemsetAsync(xreg, 0, (size_t)nreg * 8, ctx->stream));
    IFP(cudaMemsetAsync(ucut, 0, (size_t)(ctx->ncut > 0 ? ctx->ncut : 1) * 2 * C * 8, ctx->stream));
    IFP(cudaMemcpyAsync(xs + down, ctx->x, (size_t)ctx->nrows * 8, cudaMemcpyDeviceToDevice, ctx->stream));
    { pb_status s2 = halo_exchange(ctx, xs); if (s2 != PB_OK) { release(); return s2; } }
    const double* xsg = xs - ctx->rb_lo * F;                 // by global unknown of the interface system
    ifc_gather_regular_kernel<<<148 * 8, 256, 0, ctx->stream>>>(g, F, c_begin, c_end, ctx->ftab, xsg, xreg - st.col_lo * F);
    double* uc_shift = uc - c_begin * C;                     // K6 indexes cell dofs by GLOBAL cell id
#define PB_REC(FD_) recover_kernel<FD_><<<148 * 4, 128, 0, ctx->stream>>>(g, st.cell_row0, ctx->ncells_local, g.row_begin, C, ctx->prm.fun, ctx->points, ctx->recA, ctx->recb, \
                                                                         xreg, st.col_lo * F, uc_shift, nullptr)
    switch (ctx->FD) { case 0: PB_REC(0); break; case 1: PB_REC(1); break; default: PB_REC(2); break; }
#undef PB_REC
    if (ctx->ncut_owned > 0)
        ifc_recover_cut_kernel<<<148 * 2, 256, 0, ctx->stream>>>(g, F, C, ctx->ncut - ctx->ncut_owned, ctx->ncut, ctx->cut_cells, ctx->face_loc, ctx->ftab, ctx->recAx, ctx->recbx, xsg, ucut);
    IFP(cudaGetLastError());
    IFP(cudaStreamSynchronize(ctx->stream));
#undef IFP
    cudaFree(xs); cudaFree(xreg);
    *uc_out = uc; *ucut_out = ucut;
    return PB_OK;
}

// the owned part of the reference's solution vector: cell blocks of the owned cells in cell_table order (a cut cell carries its negative, then
// its positive set), (owned cells + owned cut cells) * cbs doubles, and the owned face unknowns.  The ranks' parts, concatenated in rank
// order (cells, then faces), are the whole-mesh vector.
static pb_status interface_recover_owned(pb_ctx* ctx, double* cell_dofs, double* face_dofs) {
    const Grid& g = ctx->grid; const int C = ctx->C;
    const long long c_begin = g.row_begin * g.Nx, c_end = g.row_end * g.Nx, ncell_dofs = ((c_end - c_begin) + ctx->ncut_owned) * C;
    double *uc = nullptr, *ucut = nullptr, *packed = nullptr;
    pb_status st = interface_cell_unknowns(ctx, &uc, &ucut); if (st != PB_OK) return st;
    cudaError_t e = cudaMalloc((void**)&packed, (size_t)ncell_dofs * 8);
    if (e == cudaSuccess) {
        ifc_pack_cells_kernel<<<148 * 8, 256, 0, ctx->stream>>>(g, C, c_begin, c_end, ctx->ncut - ctx->ncut_owned, ctx->ncut, ctx->cut_cells, ctx->cell_loc_g, uc - c_begin * C, ucut, packed);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(cell_dofs, packed, (size_t)ncell_dofs * 8, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess && face_dofs) e = cudaMemcpyAsync(face_dofs, ctx->x, (size_t)ctx->nrows * 8, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(uc); cudaFree(ucut); cudaFree(packed);
    if (e != cudaSuccess) return fail(ctx, PB_ERR_CUDA, "pb_recover (interface): %s", cudaGetErrorString(e));
    return PB_OK;
}
static pb_status interface_recover(pb_ctx* ctx, double* full) {
    NEED(ctx->world == 1 && ctx->strip.cf_begin == 0 && ctx->strip.cf_end == num_free_faces(ctx->grid), PB_ERR_UNSUPPORTED,
         "pb_recover (interface): whole-mesh contexts (strips: pb_recover_strip, whose parts concatenate to this vector)");
    const Grid& g = ctx->grid;
    return interface_recover_owned(ctx, full, full + (g.Nx * g.Ny + ctx->ncut) * ctx->C);
}

static pb_status interface_error(pb_ctx* ctx, double out[3]) {
    const Grid& g = ctx->grid;
    double *uc = nullptr, *ucut = nullptr, *parts = nullptr;
    pb_status st = interface_cell_unknowns(ctx, &uc, &ucut); if (st != PB_OK) return st;
    const int grid = 148 * 16;
    CK(cudaMalloc((void**)&parts, (size_t)grid * 8));
#define PB_H1(CD_) interface_h1_kernel<CD_><<<grid, 32, 0, ctx->stream>>>(g, ctx->points, ctx->node_loc, ctx->cell_loc_g, ctx->cut_cells, ctx->iface, ctx->ncut, ctx->NI, \
                                                                         ctx->prm.fun, g.row_begin * g.Nx, g.row_end * g.Nx, uc - g.row_begin * g.Nx * ctx->C, ucut, parts)
    switch (ctx->CD) { case 1: PB_H1(1); break; case 2: PB_H1(2); break; default: PB_H1(3); break; }
#undef PB_H1
    std::vector<double> h((size_t)grid);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(h.data(), parts, h.size() * 8, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(uc); cudaFree(ucut); cudaFree(parts);
    if (e != cudaSuccess) return fail(ctx, PB_ERR_CUDA, "pb_error_norms (interface): %s", cudaGetErrorString(e));
    out[0] = out[1] = out[2] = 0.0;
    for (double v : h) out[0] += v;
    return PB_OK;
}
