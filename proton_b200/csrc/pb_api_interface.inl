// C ABI — the two-material interface problem (`cuthho_square -i`): cut cells by K2i (pb_interface.cuh), uncut
// cells by K1, and the condensed face system in the numbering of interface_assembler (apps/cuthho/cuthho_square.cpp:1091-
// 1436): every cut face carries two sets of unknowns, face_table counts them in face order (:1161-1175), the second set
// of face f sits fbs after the first (:1301).  The rows of faces between two uncut cells keep the closed-form block rows of
// the plain problems (renumbered by face_table) and are served by K3 and the block SpMV; the rows of the cut cells' faces
// are built the way the reference builds everything (one triplet per local entry, sorted and summed — setFromTriplets,
// :1430-1434) with cub on the device.  build_interface_system_sorted is the first version, which sorted every local entry;
// PB_IFC_SORT_ALL=1 still selects it on whole-mesh contexts (comparison runs, tests).
//
// Strips (SURVEY 8e on 8f.3): face_table is monotone in the face id, so a strip of cell rows owns a contiguous range of slots
// [rb_begin, rb_end) and references the contiguous range [rb_lo, rb_hi) of the face rows of its local cells (ghost row below,
// face row above).  slot(f) = compressed id of f + number of cut faces before f; the cut faces of the strips below are the one
// piece of global information: ranks that share a communicator all-reduce their counts, strip contexts without one are told by
// the host (pb_interface_cut_faces_owned / pb_interface_set_cut_offsets).  Every index array (face weights, face_table,
// irregular flags) holds the face window of the cut preprocessing only and is addressed by global face id through a shifted
// pointer; rows, brp, rowptr, val and rhs hold the owned slots only; columns are global unknowns of the interface system.

// face id -> the two cells of a free face and the local index of the face in each (as build_block_rows_kernel); false = boundary face
__device__ __forceinline__ bool ifc_face_cells(const Grid& g, long long f, long long ci[2], long long cj[2], int li[2]) {
    const long long per = 2 * g.Nx + 1, j = f / per, rem = f - j * per;
    if (j >= g.Ny) return false;                              // top boundary
    if (rem == 2 * g.Nx) return false;                        // right boundary
    if ((rem & 1) == 0) { const long long i = rem / 2; if (j == 0) return false; ci[0] = i; cj[0] = j - 1; li[0] = 2; ci[1] = i; cj[1] = j; li[1] = 0; return true; }
    const long long i = (rem - 1) / 2; if (i == 0) return false;
    ci[0] = i - 1; cj[0] = j; li[0] = 1; ci[1] = i; cj[1] = j; li[1] = 3; return true;
}
// weight of the faces [f0, f1) in face_table: 0 Dirichlet (the generator's boundary faces), 2 cut, 1 otherwise
__global__ void ifc_face_weight_kernel(Grid g, long long f0, long long f1, const int* __restrict__ face_loc, int* __restrict__ w) {
    for (long long f = f0 + blockIdx.x * (long long)blockDim.x + threadIdx.x; f < f1; f += (long long)gridDim.x * blockDim.x) {
        long long ci[2], cj[2]; int li[2];
        w[f] = !ifc_face_cells(g, f, ci, cj, li) ? 0 : (face_loc[f] == PB_ON_INTERFACE ? 2 : 1);
    }
}
__global__ void ifc_add_base_kernel(long long f0, long long f1, int base, int* __restrict__ ftab) {
    for (long long f = f0 + blockIdx.x * (long long)blockDim.x + threadIdx.x; f <= f1; f += (long long)gridDim.x * blockDim.x) ftab[f] += base;
}
// one key per local entry: row * n + col of the condensed system, or the sentinel n*n for entries that are not assembled
// (Dirichlet rows / columns; the stale S block of a cut cell, whose entries come from the second list)
__global__ void ifc_triplets_uncut_kernel(Grid g, int F, const int* __restrict__ cell_loc, const int* __restrict__ ftab, long long n,
                                          const double* __restrict__ S, unsigned long long* __restrict__ keys, double* __restrict__ vals) {
    const int NF = 4 * F; const long long nc = g.Nx * g.Ny, per = (long long)NF * NF;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < nc * per; t += (long long)gridDim.x * blockDim.x) {
        const long long c = t / per; const int e = (int)(t - c * per), j = e / NF, i = e - j * NF;     // S is column-major
        unsigned long long key = (unsigned long long)n * (unsigned long long)n;
        double v = 0.0;
        if (cell_loc[c] != PB_ON_INTERFACE) {
            const long long ci = c % g.Nx, cj = c / g.Nx;
            const int li = i / F, lj = j / F;
            if (!cell_face_dirichlet(g, ci, cj, li) && !cell_face_dirichlet(g, ci, cj, lj)) {
                long long f[4]; cell_faces(g, ci, cj, f);
                const long long row = (long long)ftab[f[li]] * F + (i - li * F), col = (long long)ftab[f[lj]] * F + (j - lj * F);
                key = (unsigned long long)row * (unsigned long long)n + (unsigned long long)col; v = S[t];
            }
        }
        keys[t] = key; vals[t] = v;
    }
}
__device__ __forceinline__ long long ifc_slot_dof(const Grid& g, int F, const int* face_loc, const int* ftab, const long long* f, int t) {
    const int NF = 4 * F, side = t / NF, r = t - side * NF, li = r / F, a = r - li * F;
    const long long fid = f[li];
    return (long long)ftab[fid] * F + a + ((side == 1 && face_loc[fid] == PB_ON_INTERFACE) ? F : 0);
}
__global__ void ifc_triplets_cut_kernel(Grid g, int F, int ncut, const int* __restrict__ cut_cells, const int* __restrict__ face_loc,
                                        const int* __restrict__ ftab, long long n, const double* __restrict__ Sx,
                                        unsigned long long* __restrict__ keys, double* __restrict__ vals) {
    const int NF2 = 8 * F; const long long per = (long long)NF2 * NF2;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < ncut * per; t += (long long)gridDim.x * blockDim.x) {
        const long long slot = t / per; const int e = (int)(t - slot * per), j = e / NF2, i = e - j * NF2;
        const long long c = cut_cells[slot], ci = c % g.Nx, cj = c / g.Nx;
        long long f[4]; cell_faces(g, ci, cj, f);
        const long long row = ifc_slot_dof(g, F, face_loc, ftab, f, i), col = ifc_slot_dof(g, F, face_loc, ftab, f, j);
        keys[t] = (unsigned long long)row * (unsigned long long)n + (unsigned long long)col; vals[t] = Sx[t];
    }
}
// rhs: every face unknown receives at most two non-zero contributions, added to an exact zero: the sum does not depend on the order
__global__ void ifc_rhs_kernel(Grid g, int F, int ncut, const int* __restrict__ cut_cells, const int* __restrict__ cell_loc, const int* __restrict__ face_loc,
                               const int* __restrict__ ftab, const double* __restrict__ gv, const double* __restrict__ gx, double* __restrict__ rhs) {
    const int NF = 4 * F; const long long nc = g.Nx * g.Ny;
    const long long total = nc * NF + (long long)ncut * 2 * NF;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        if (t < nc * NF) {
            const long long c = t / NF; const int i = (int)(t - c * NF), li = i / F;
            if (cell_loc[c] == PB_ON_INTERFACE) continue;
            const long long ci = c % g.Nx, cj = c / g.Nx;
            if (cell_face_dirichlet(g, ci, cj, li)) continue;
            long long f[4]; cell_faces(g, ci, cj, f);
            atomicAdd(rhs + (long long)ftab[f[li]] * F + (i - li * F), gv[t]);
        } else {
            const long long u = t - nc * NF, slot = u / (2 * NF); const int i = (int)(u - slot * 2 * NF);
            const long long c = cut_cells[slot], ci = c % g.Nx, cj = c / g.Nx;
            long long f[4]; cell_faces(g, ci, cj, f);
            const double v = gx[u];
            if (v != 0.0) atomicAdd(rhs + ifc_slot_dof(g, F, face_loc, ftab, f, i), v);
        }
    }
}
__global__ void ifc_split_keys_kernel(long long nnz, long long n, const unsigned long long* __restrict__ ukeys, int* __restrict__ col) {
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < nnz; t += (long long)gridDim.x * blockDim.x)
        col[t] = (int)(ukeys[t] % (unsigned long long)n);
}
__global__ void ifc_rowptr_kernel(long long nnz, long long n, const unsigned long long* __restrict__ ukeys, long long* __restrict__ rowptr) {
    for (long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x; r <= n; r += (long long)gridDim.x * blockDim.x) {
        const unsigned long long want = (unsigned long long)r * (unsigned long long)n;
        long long lo = 0, hi = nnz;
        while (lo < hi) { const long long mid = (lo + hi) / 2; if (ukeys[mid] < want) lo = mid + 1; else hi = mid; }
        rowptr[r] = lo;
    }
}

template <int CD, int FD>
static cudaError_t launch_ifc_t(pb_ctx* c, const IfcArgs& A) {
    const size_t smem = sizeof(IfcSmem<CD, FD>);
    auto kern = local_cut_interface_kernel<CD, FD>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int grid = A.ncut < 148 * 8 ? A.ncut : 148 * 8;
    kern<<<grid, 32, smem, c->stream>>>(A);
    return cudaGetLastError();
}

// cut cells of the interface problem -> Sx, gx, recAx, recbx (and the local matrices when keep_local)
static pb_status launch_interface_cut(pb_ctx* ctx, const pb_assembly_params* p, int cd, int fd) {
    const int C = ctx->C, NF = ctx->NF, N2 = 2 * C + 2 * NF;
    const size_t nc = (size_t)(ctx->ncut > 0 ? ctx->ncut : 1);
    if (!ctx->Sx || ctx->ifc_ncut_alloc != (long long)nc) {          // kept across assemblies of the same cut mesh
        CK(dalloc(ctx, &ctx->Sx, nc * 4 * NF * NF)); CK(dalloc(ctx, &ctx->gx, nc * 2 * NF));
        CK(dalloc(ctx, &ctx->recAx, nc * 4 * NF * C)); CK(dalloc(ctx, &ctx->recbx, nc * 2 * C));
        dfree(&ctx->lcx); dfree(&ctx->fx);
        ctx->ifc_ncut_alloc = (long long)nc;
    }
    if (p->keep_local) { if (!ctx->lcx) { CK(dalloc(ctx, &ctx->lcx, nc * N2 * N2)); CK(dalloc(ctx, &ctx->fx, nc * 2 * C)); } }
    else { dfree(&ctx->lcx); dfree(&ctx->fx); }
    if (ctx->ncut == 0) return PB_OK;
    IfcArgs A{};
    A.grid = ctx->grid; A.ls = ctx->ls; A.points = ctx->points; A.node_loc = ctx->node_loc; A.face_loc = ctx->face_loc; A.face_ip = ctx->face_ip;
    A.cut_cells = ctx->cut_cells; A.iface = ctx->iface; A.ncut = ctx->ncut; A.NI = ctx->NI;
    A.fun = p->fun; A.eta = p->eta; A.kappa1 = 1.0; A.kappa2 = 1.0;             // params<T>(): kappa_1 = kappa_2 = 1 (cuthho_square.cpp:293-299, 1659)
    A.Sx = ctx->Sx; A.gx = ctx->gx; A.recAx = ctx->recAx; A.recbx = ctx->recbx; A.lcx = ctx->lcx; A.fx = ctx->fx;
    cudaError_t e;
    if (cd == 1 && fd == 0) e = launch_ifc_t<1, 0>(ctx, A);
    else if (cd == 2 && fd == 1) e = launch_ifc_t<2, 1>(ctx, A);
    else if (cd == 3 && fd == 2) e = launch_ifc_t<3, 2>(ctx, A);
    else return fail(ctx, PB_ERR_UNSUPPORTED, "pb_assemble: the interface problem is built for HHO(k+1,k), k <= 2 (SURVEY F4)");
    if (e != cudaSuccess) return fail(ctx, PB_ERR_CUDA, "local_cut_interface_kernel launch failed: %s", cudaGetErrorString(e));
    return PB_OK;
}

// condensed system in interface_assembler's numbering: triplets -> radix sort -> sum of duplicates -> CSR
static pb_status build_interface_system_sorted(pb_ctx* ctx) {
    const Grid& g = ctx->grid; const int F = ctx->F, NF = ctx->NF;
    NEED(ctx->strip.cf_begin == 0 && ctx->strip.cf_end == num_free_faces(ctx->grid), PB_ERR_UNSUPPORTED, "pb_assemble: PB_IFC_SORT_ALL is a whole-mesh path");
    ctx->ifc_struct_ok = false; dfree(&ctx->brows); dfree(&ctx->brp);          // no block-row table on this path: pb_pcg takes the scalar-CSR SpMV
    const long long nc = g.Nx * g.Ny, nf = ctx->nfaces;
    cudaStream_t sm = ctx->stream;
    int* w = nullptr; CK(cudaMalloc((void**)&w, (size_t)(nf + 1) * 4));
    CK(cudaMemsetAsync(w, 0, (size_t)(nf + 1) * 4, sm));
    ifc_face_weight_kernel<<<148 * 8, 256, 0, sm>>>(g, 0, nf, ctx->face_loc, w);
    CK(dalloc(ctx, &ctx->ftab_a, (size_t)nf + 1)); ctx->ftab = ctx->ftab_a;
    void* tmp = nullptr; size_t tb = 0;
    CK(cub::DeviceScan::ExclusiveSum(nullptr, tb, w, ctx->ftab, (int)(nf + 1), sm));
    CK(cudaMalloc(&tmp, tb));
    CK(cub::DeviceScan::ExclusiveSum(tmp, tb, w, ctx->ftab, (int)(nf + 1), sm));
    int nslots = 0;
    CK(cudaMemcpyAsync(&nslots, ctx->ftab + nf, 4, cudaMemcpyDeviceToHost, sm));
    CK(cudaStreamSynchronize(sm));
    cudaFree(tmp); cudaFree(w);
    const long long n = (long long)nslots * F;
    NEED(n < (1ll << 31), PB_ERR_UNSUPPORTED, "pb_assemble: interface system too large for 32-bit column indices");
    const long long T = nc * NF * NF + (long long)ctx->ncut * 4 * NF * NF;
    unsigned long long *k0 = nullptr, *k1 = nullptr; double *v0 = nullptr, *v1 = nullptr;
    cudaError_t e = cudaMalloc((void**)&k0, (size_t)T * 8);
    if (e == cudaSuccess) e = cudaMalloc((void**)&k1, (size_t)T * 8);
    if (e == cudaSuccess) e = cudaMalloc((void**)&v0, (size_t)T * 8);
    if (e == cudaSuccess) e = cudaMalloc((void**)&v1, (size_t)T * 8);
    auto release = [&]() { cudaFree(k0); cudaFree(k1); cudaFree(v0); cudaFree(v1); };
    if (e != cudaSuccess) { release(); return fail(ctx, PB_ERR_CUDA, "pb_assemble: no memory for %lld interface triplets (%s)", T, cudaGetErrorString(e)); }
    ifc_triplets_uncut_kernel<<<148 * 16, 256, 0, sm>>>(g, F, ctx->cell_loc_g, ctx->ftab, n, ctx->S, k0, v0);
    if (ctx->ncut > 0)
        ifc_triplets_cut_kernel<<<148 * 4, 256, 0, sm>>>(g, F, ctx->ncut, ctx->cut_cells, ctx->face_loc, ctx->ftab, n, ctx->Sx, k0 + nc * NF * NF, v0 + nc * NF * NF);
    int bits = 1; while (bits < 64 && ((unsigned long long)n * (unsigned long long)n >> bits) != 0) bits++;
    cub::DoubleBuffer<unsigned long long> dk(k0, k1); cub::DoubleBuffer<double> dv(v0, v1);
    tb = 0; tmp = nullptr;
    e = cub::DeviceRadixSort::SortPairs(nullptr, tb, dk, dv, T, 0, bits, sm);
    if (e == cudaSuccess) e = cudaMalloc(&tmp, tb);
    if (e == cudaSuccess) e = cub::DeviceRadixSort::SortPairs(tmp, tb, dk, dv, T, 0, bits, sm);      // stable: equal keys keep the generation order
    cudaFree(tmp); tmp = nullptr;
    if (e != cudaSuccess) { release(); return fail(ctx, PB_ERR_CUDA, "pb_assemble: sorting the interface triplets failed: %s", cudaGetErrorString(e)); }
    unsigned long long* uk = dk.Alternate(); double* uv = dv.Alternate();
    long long* dnum = nullptr; e = cudaMalloc((void**)&dnum, 8);
    tb = 0;
    if (e == cudaSuccess) e = cub::DeviceReduce::ReduceByKey(nullptr, tb, dk.Current(), uk, dv.Current(), uv, dnum, cuda::std::plus<>(), T, sm);
    if (e == cudaSuccess) e = cudaMalloc(&tmp, tb);
    if (e == cudaSuccess) e = cub::DeviceReduce::ReduceByKey(tmp, tb, dk.Current(), uk, dv.Current(), uv, dnum, cuda::std::plus<>(), T, sm);
    long long nuniq = 0; unsigned long long lastkey = 0;
    if (e == cudaSuccess) e = cudaMemcpyAsync(&nuniq, dnum, 8, cudaMemcpyDeviceToHost, sm);
    if (e == cudaSuccess) e = cudaStreamSynchronize(sm);
    if (e == cudaSuccess && nuniq > 0) e = cudaMemcpy(&lastkey, uk + nuniq - 1, 8, cudaMemcpyDeviceToHost);
    cudaFree(tmp); cudaFree(dnum);
    if (e != cudaSuccess) { release(); return fail(ctx, PB_ERR_CUDA, "pb_assemble: reducing the interface triplets failed: %s", cudaGetErrorString(e)); }
    const long long nnz = (nuniq > 0 && lastkey == (unsigned long long)n * (unsigned long long)n) ? nuniq - 1 : nuniq;      // drop the sentinel
    ctx->nrows = n; ctx->nnz = nnz; ctx->nrows_blk = nslots; ctx->ifc_nglob = n;
    ctx->rb_begin = ctx->rb_lo = 0; ctx->rb_end = ctx->rb_hi = nslots; ctx->halo_prev = ctx->halo_next = 0;
    e = dalloc(ctx, &ctx->rowptr, (size_t)n + 1);
    if (e == cudaSuccess) e = dalloc(ctx, &ctx->col, (size_t)nnz);
    if (e == cudaSuccess) e = dalloc(ctx, &ctx->val, (size_t)nnz);
    if (e == cudaSuccess) e = dalloc(ctx, &ctx->rhs, (size_t)n);
    if (e == cudaSuccess) e = cudaMemcpyAsync(ctx->val, uv, (size_t)nnz * 8, cudaMemcpyDeviceToDevice, sm);
    if (e == cudaSuccess) e = cudaMemsetAsync(ctx->rhs, 0, (size_t)n * 8, sm);
    if (e != cudaSuccess) { release(); return fail(ctx, PB_ERR_CUDA, "pb_assemble: interface system allocation failed: %s", cudaGetErrorString(e)); }
    ifc_split_keys_kernel<<<148 * 8, 256, 0, sm>>>(nnz, n, uk, ctx->col);
    ifc_rowptr_kernel<<<(unsigned)((n + 256) / 256), 256, 0, sm>>>(nnz, n, uk, ctx->rowptr);
    ifc_rhs_kernel<<<148 * 8, 256, 0, sm>>>(g, F, ctx->ncut, ctx->cut_cells, ctx->cell_loc_g, ctx->face_loc, ctx->ftab, ctx->g, ctx->gx, ctx->rhs);
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaStreamSynchronize(sm);
    release();
    if (e != cudaSuccess) return fail(ctx, PB_ERR_CUDA, "pb_assemble: interface system kernels failed: %s", cudaGetErrorString(e));
    return PB_OK;
}


// ------------------------------------------------------------------------------------------------------------------
// Hybrid system build (default).  Rows of a face whose two cells are uncut keep the closed-form block row of the plain
// problems — only their column blocks move, by the number of cut faces before them — so K3 (and the block SpMV) serve
// them unchanged from a BlockRow table in interface order.  Only the rows of the cut cells' faces ("irregular": their
// four faces, both sets of the cut ones; about 4 ncut block rows) are built from sorted triplets.  All rows still come in
// groups of fbs with fbs-wide column blocks, so one brp / rowptr layout covers both kinds.

// *ndirichlet counts the Dirichlet faces met on cut cells: the reference refuses them (:1292-1293)
__global__ void ifc_flag_irregular_kernel(Grid g, int ncut, const int* __restrict__ cut_cells, int* __restrict__ irr, int* __restrict__ ndirichlet) {
    for (int s = blockIdx.x * blockDim.x + threadIdx.x; s < ncut; s += gridDim.x * blockDim.x) {
        const long long c = cut_cells[s], ci = c % g.Nx, cj = c / g.Nx;
        long long f[4]; cell_faces(g, ci, cj, f);
        for (int li = 0; li < 4; li++) { if (!cell_face_dirichlet(g, ci, cj, li)) irr[f[li]] = 1; else atomicAdd(ndirichlet, 1); }
    }
}
constexpr unsigned long long IFC_NOKEY = ~0ull;
// triplets of the irregular rows: per (irregular face, adjacent cell) a fixed window of 4*F*NF entries
__global__ void ifc_irregular_triplets_kernel(Grid g, int F, int nirr, const int* __restrict__ irr_faces, const int* __restrict__ cell_loc, int ncut,
                                              const int* __restrict__ cut_cells, const int* __restrict__ face_loc, const int* __restrict__ ftab, long long n,
                                              long long c0, const double* __restrict__ S, const double* __restrict__ Sx, unsigned long long* __restrict__ keys,
                                              double* __restrict__ vals) {      // c0: first local cell (S holds the local cells)
    const int NF = 4 * F, NF2 = 2 * NF, win = 4 * F * NF;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < (long long)nirr * 2 * win; t += (long long)gridDim.x * blockDim.x) {
        const long long q = t / (2 * win); const int r = (int)(t - q * 2 * win), k = r / win, e = r - k * win;
        unsigned long long key = IFC_NOKEY; double v = 0.0;
        long long ci[2], cj[2]; int li[2];
        if (ifc_face_cells(g, irr_faces[q], ci, cj, li)) {
            const long long c = cj[k] * g.Nx + ci[k];
            long long f[4]; cell_faces(g, ci[k], cj[k], f);
            if (cell_loc[c] != PB_ON_INTERFACE) {
                if (e < F * NF) {             // rows (li, a), all columns
                    const int a = e / NF, jj = e - a * NF, lj = jj / F;
                    if (!cell_face_dirichlet(g, ci[k], cj[k], lj)) {
                        const long long row = (long long)ftab[f[li[k]]] * F + a, col = (long long)ftab[f[lj]] * F + (jj - lj * F);
                        key = (unsigned long long)row * (unsigned long long)n + (unsigned long long)col;
                        v = S[(c - c0) * (long long)(NF * NF) + (li[k] * F + a) + (long long)jj * NF];
                    }
                }
            } else {
                int lo = 0, hi = ncut - 1;
                while (lo < hi) { const int mid = (lo + hi) / 2; if (cut_cells[mid] < c) lo = mid + 1; else hi = mid; }
                const int sd = e / (F * NF2), r2 = e - sd * (F * NF2), a = r2 / NF2, tc = r2 - a * NF2, tr = sd * NF + li[k] * F + a;
                const long long row = ifc_slot_dof(g, F, face_loc, ftab, f, tr), col = ifc_slot_dof(g, F, face_loc, ftab, f, tc);
                key = (unsigned long long)row * (unsigned long long)n + (unsigned long long)col;
                v = Sx[lo * (long long)(NF2 * NF2) + tr + (long long)tc * NF2];
            }
        }
        keys[t] = key; vals[t] = v;
    }
}
__device__ __forceinline__ long long ifc_lower_bound(const unsigned long long* k, long long n, unsigned long long want) {
    long long lo = 0, hi = n;
    while (lo < hi) { const long long mid = (lo + hi) / 2; if (k[mid] < want) lo = mid + 1; else hi = mid; }
    return lo;
}
// one thread per face: block row(s) of the face in interface order.  Regular face: the closed-form block row with the
// column blocks renumbered by face_table; irregular face: an empty BlockRow (K3 and the block SpMV skip it) and the
// number of column blocks counted in the sorted irregular entries.
// (faces [f0, f1) = the owned faces; rows / nb_all are addressed by global slot through pointers shifted back by the first owned slot)
__global__ void ifc_block_rows_kernel(Grid g, int F, long long f0, long long f1, const int* __restrict__ ftab, const int* __restrict__ irr, const int* __restrict__ face_loc,
                                      long long n, const unsigned long long* __restrict__ ukeys, long long nuniq, long long c0, BlockRow* __restrict__ rows, int* __restrict__ nb_all) {
    for (long long f = f0 + blockIdx.x * (long long)blockDim.x + threadIdx.x; f < f1; f += (long long)gridDim.x * blockDim.x) {
        if (ftab[f + 1] == ftab[f]) continue;                // Dirichlet
        const long long s = ftab[f];
        long long ci[2], cj[2]; int li[2];
        ifc_face_cells(g, f, ci, cj, li);
        BlockRow br; memset(&br, 0, sizeof br);
        if (irr[f]) {
            const int sets = face_loc[f] == PB_ON_INTERFACE ? 2 : 1;
            for (int k = 0; k < sets; k++) {
                const unsigned long long r0 = (unsigned long long)((s + k) * F) * (unsigned long long)n;
                const long long len = ifc_lower_bound(ukeys, nuniq, r0 + (unsigned long long)n) - ifc_lower_bound(ukeys, nuniq, r0);
                for (int m = 0; m < 7; m++) { br.bcol[m] = -1; br.src[0][m] = 255; br.src[1][m] = 255; }
                rows[s + k] = br; nb_all[s + k] = (int)(len / F);
            }
            continue;
        }
        long long ids[8]; unsigned char s0[8], s1[8]; int cnt = 0;
        for (int k = 0; k < 2; k++) {
            br.cell[k] = (int)(cj[k] * g.Nx + ci[k] - c0); br.li[k] = (unsigned char)li[k];
            long long fc[4]; cell_faces(g, ci[k], cj[k], fc);
            for (int lf = 0; lf < 4; lf++) {
                if (cell_face_dirichlet(g, ci[k], cj[k], lf)) continue;
                const long long id = ftab[fc[lf]];
                int pos = -1;
                for (int m = 0; m < cnt; m++) if (ids[m] == id) pos = m;
                if (pos < 0) { pos = cnt++; ids[pos] = id; s0[pos] = 255; s1[pos] = 255; }
                if (k == 0) s0[pos] = (unsigned char)lf; else s1[pos] = (unsigned char)lf;
            }
        }
        for (int a = 1; a < cnt; a++) {
            long long id = ids[a]; unsigned char x0 = s0[a], x1 = s1[a]; int b = a - 1;
            while (b >= 0 && ids[b] > id) { ids[b + 1] = ids[b]; s0[b + 1] = s0[b]; s1[b + 1] = s1[b]; b--; }
            ids[b + 1] = id; s0[b + 1] = x0; s1[b + 1] = x1;
        }
        br.nb = (unsigned char)cnt;
        for (int m = 0; m < 7; m++) { br.bcol[m] = m < cnt ? (int)ids[m] : -1; br.src[0][m] = m < cnt ? s0[m] : 255; br.src[1][m] = m < cnt ? s1[m] : 255; }
        rows[s] = br; nb_all[s] = cnt;
    }
}
// rowptr of every scalar row; columns of the regular rows
__global__ void ifc_pattern_kernel(long long nslots, int F, const BlockRow* __restrict__ rows, const int* __restrict__ nb_all, const long long* __restrict__ brp,
                                   long long* __restrict__ rowptr, int* __restrict__ col) {
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t <= nslots * F; t += (long long)gridDim.x * blockDim.x) {
        if (t == nslots * F) { rowptr[t] = brp[nslots] * F * F; continue; }
        const long long s = t / F; const int a = (int)(t - s * F);
        const long long base = brp[s] * F * F + (long long)a * nb_all[s] * F;
        rowptr[t] = base;
        const BlockRow& R = rows[s];
        for (int b = 0; b < R.nb; b++) for (int bc = 0; bc < F; bc++) col[base + b * F + bc] = R.bcol[b] * F + bc;
    }
}
// columns and values of the irregular rows from the sorted, summed entries
__global__ void ifc_irregular_fill_kernel(long long nuniq, long long n, const unsigned long long* __restrict__ ukeys, const double* __restrict__ uvals,
                                          long long row0, const long long* __restrict__ rowptr, int* __restrict__ col, double* __restrict__ val) {
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < nuniq; e += (long long)gridDim.x * blockDim.x) {
        const unsigned long long key = ukeys[e];
        if (key == IFC_NOKEY) continue;
        const unsigned long long r = key / (unsigned long long)n;
        const long long lo = ifc_lower_bound(ukeys, nuniq, r * (unsigned long long)n);
        const long long dst = rowptr[(long long)r - row0] + (e - lo);      // rowptr holds the owned rows, which start at global row row0
        col[dst] = (int)(key - r * (unsigned long long)n); val[dst] = uvals[e];
    }
}
// rows of the irregular block rows, for the scalar SpMV pass
__global__ void ifc_irregular_rows_kernel(long long nslots, int F, const BlockRow* __restrict__ rows, const int* __restrict__ nb_all, int* __restrict__ flags) {
    for (long long s = blockIdx.x * (long long)blockDim.x + threadIdx.x; s < nslots; s += (long long)gridDim.x * blockDim.x)
        flags[s] = (rows[s].nb == 0 && nb_all[s] > 0) ? 1 : 0;
}

// rhs of the irregular rows: zeroed, then the <= 2 cells of every irregular face add their part (K3 has written the others)
__global__ void ifc_zero_rhs_rows_kernel(int nrows_irr, int F, const int* __restrict__ irr_rows, double* __restrict__ rhs) {
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < nrows_irr * F; t += gridDim.x * blockDim.x) rhs[(long long)irr_rows[t / F] * F + (t % F)] = 0.0;
}
__global__ void ifc_irregular_rhs_kernel(Grid g, int F, int nirr, const int* __restrict__ irr_faces, const int* __restrict__ cell_loc, int ncut,
                                         const int* __restrict__ cut_cells, const int* __restrict__ face_loc, const int* __restrict__ ftab,
                                         long long c0, const double* __restrict__ gv, const double* __restrict__ gx, double* __restrict__ rhs) {
    // (gv holds the local cells, the first of which is c0; rhs is addressed by global unknown through a shifted pointer)
    const int NF = 4 * F, win = 2 * F;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < (long long)nirr * 2 * win; t += (long long)gridDim.x * blockDim.x) {
        const long long q = t / (2 * win); const int r = (int)(t - q * 2 * win), k = r / win, e = r - k * win;
        long long ci[2], cj[2]; int li[2];
        if (!ifc_face_cells(g, irr_faces[q], ci, cj, li)) continue;
        const long long c = cj[k] * g.Nx + ci[k];
        long long f[4]; cell_faces(g, ci[k], cj[k], f);
        if (cell_loc[c] != PB_ON_INTERFACE) {
            if (e < F) atomicAdd(rhs + (long long)ftab[f[li[k]]] * F + e, gv[(c - c0) * NF + li[k] * F + e]);
        } else {
            int lo = 0, hi = ncut - 1;
            while (lo < hi) { const int mid = (lo + hi) / 2; if (cut_cells[mid] < c) lo = mid + 1; else hi = mid; }
            const int sd = e / F, a = e - sd * F, tr = sd * NF + li[k] * F + a;
            const double v = gx[(long long)lo * 2 * NF + tr];
            if (v != 0.0) atomicAdd(rhs + ifc_slot_dof(g, F, face_loc, ftab, f, tr), v);
        }
    }
}

// sorted, summed entries of the irregular rows (values of the current S / Sx) in buffers kept with the structure
struct IfcSorted { unsigned long long* uk = nullptr; double* uv = nullptr; long long nuniq = 0; void release() {} };
static pb_status ifc_sorted_irregular(pb_ctx* ctx, IfcSorted& R) {
    const Grid& g = ctx->grid; const int F = ctx->F, NF = ctx->NF; cudaStream_t sm = ctx->stream;
    const int nirr = ctx->n_irr_faces; const long long n = ctx->ifc_nglob;      // keys: global row * n + global column
    const long long T = (long long)(nirr > 0 ? nirr : 1) * 2 * 4 * F * NF;
    if (ctx->ifc_T != T || !ctx->ifc_k0) {
        dfree(&ctx->ifc_k0); dfree(&ctx->ifc_k1); dfree(&ctx->ifc_v0); dfree(&ctx->ifc_v1); dfree(&ctx->ifc_dnum);
        if (ctx->ifc_tmp) { cudaFree(ctx->ifc_tmp); ctx->ifc_tmp = nullptr; }
        CK(dalloc(ctx, &ctx->ifc_k0, (size_t)T)); CK(dalloc(ctx, &ctx->ifc_k1, (size_t)T)); CK(dalloc(ctx, &ctx->ifc_v0, (size_t)T)); CK(dalloc(ctx, &ctx->ifc_v1, (size_t)T));
        CK(dalloc(ctx, &ctx->ifc_dnum, 1));
        cub::DoubleBuffer<unsigned long long> dk(ctx->ifc_k0, ctx->ifc_k1); cub::DoubleBuffer<double> dv(ctx->ifc_v0, ctx->ifc_v1);
        size_t t1 = 0, t2 = 0;
        CK(cub::DeviceRadixSort::SortPairs(nullptr, t1, dk, dv, T, 0, 64, sm));
        CK(cub::DeviceReduce::ReduceByKey(nullptr, t2, ctx->ifc_k0, ctx->ifc_k1, ctx->ifc_v0, ctx->ifc_v1, ctx->ifc_dnum, cuda::std::plus<>(), T, sm));
        ctx->ifc_tmp_bytes = t1 > t2 ? t1 : t2;
        CK(cudaMalloc(&ctx->ifc_tmp, ctx->ifc_tmp_bytes));
        ctx->ifc_T = T;
    }
    CK(cudaMemsetAsync(ctx->ifc_k0, 0xff, (size_t)T * 8, sm)); CK(cudaMemsetAsync(ctx->ifc_v0, 0, (size_t)T * 8, sm));
    if (nirr > 0)
        ifc_irregular_triplets_kernel<<<148 * 4, 256, 0, sm>>>(g, F, nirr, ctx->irr_faces, ctx->cell_loc_g, ctx->ncut, ctx->cut_cells, ctx->face_loc, ctx->ftab, n, ctx->strip.cell_row0 * g.Nx, ctx->S, ctx->Sx,
                                                              ctx->ifc_k0, ctx->ifc_v0);
    cub::DoubleBuffer<unsigned long long> dk(ctx->ifc_k0, ctx->ifc_k1); cub::DoubleBuffer<double> dv(ctx->ifc_v0, ctx->ifc_v1);
    size_t tb = ctx->ifc_tmp_bytes;
    CK(cub::DeviceRadixSort::SortPairs(ctx->ifc_tmp, tb, dk, dv, T, 0, 64, sm));          // stable: equal keys keep the generation order
    R.uk = dk.Alternate(); R.uv = dv.Alternate();
    tb = ctx->ifc_tmp_bytes;
    CK(cub::DeviceReduce::ReduceByKey(ctx->ifc_tmp, tb, dk.Current(), R.uk, dv.Current(), R.uv, ctx->ifc_dnum, cuda::std::plus<>(), T, sm));
    CK(cudaMemcpyAsync(&R.nuniq, ctx->ifc_dnum, 8, cudaMemcpyDeviceToHost, sm));
    CK(cudaStreamSynchronize(sm));
    return PB_OK;
}

// cut faces among the owned faces of this context (what the strips above need to know of it); needs the cut preprocessing only
static pb_status ifc_count_owned_cut_faces(pb_ctx* ctx, long long* count) {
    const Grid& g = ctx->grid; const long long per = 2 * g.Nx + 1;
    const long long fo0 = g.row_begin * per, fo1 = g.row_end < g.Ny ? g.row_end * per : ctx->nfaces, nfo = fo1 - fo0;
    int* w = nullptr; long long* dsum = nullptr; void* tmp = nullptr; size_t tb = 0;
    struct Cleanup { int*& w; long long*& d; void*& t; ~Cleanup() { cudaFree(w); cudaFree(d); cudaFree(t); } } cleanup{w, dsum, tmp};
    CK(cudaMalloc((void**)&w, (size_t)(nfo + 1) * 4)); CK(cudaMalloc((void**)&dsum, 8));
    ifc_face_weight_kernel<<<148 * 4, 256, 0, ctx->stream>>>(g, fo0, fo1, ctx->face_loc, w - fo0);
    CK(cub::DeviceReduce::Sum(nullptr, tb, w, dsum, (int)nfo, ctx->stream));
    CK(cudaMalloc(&tmp, tb));
    CK(cub::DeviceReduce::Sum(tmp, tb, w, dsum, (int)nfo, ctx->stream));
    long long slots = 0;
    CK(cudaMemcpyAsync(&slots, dsum, 8, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    *count = slots - (ctx->strip.cf_end - ctx->strip.cf_begin);      // slots = free faces + cut faces
    return PB_OK;
}
static pb_status dist_gather_counts(pb_ctx* ctx, long long mine, long long* before, long long* total);      // pb_api_dist.inl

static pb_status build_interface_system(pb_ctx* ctx) {
    const Grid& g = ctx->grid; const int F = ctx->F; const Strip& strip = ctx->strip;
    const bool whole = strip.cf_begin == 0 && strip.cf_end == num_free_faces(g);
    if (whole && getenv("PB_IFC_SORT_ALL")) return build_interface_system_sorted(ctx);        // the first version, kept for comparison
    cudaStream_t sm = ctx->stream;
    const long long per = 2 * g.Nx + 1;
    const long long fw0 = ctx->fw0, fw1 = ctx->fw1, nfw = fw1 - fw0;                                  // faces with cut data
    const long long fo0 = g.row_begin * per, fo1 = g.row_end < g.Ny ? g.row_end * per : ctx->nfaces;    // owned faces
    const long long c0 = strip.cell_row0 * g.Nx;                                                        // first local cell
    IfcSorted R;
    if (!ctx->ifc_struct_ok) {
        // ---- structure, once per mesh / level set / degree: face_table, irregular faces, block rows, brp, CSR pattern
        int *w = nullptr, *irr = nullptr, *dcount = nullptr, *nb_all = nullptr, *fl = nullptr; void* tmp = nullptr;
        auto release = [&]() { cudaFree(w); cudaFree(irr); cudaFree(dcount); cudaFree(nb_all); cudaFree(fl); cudaFree(tmp); R.release(); };
#define IFC(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { release(); return fail(ctx, PB_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); } } while (0)
        IFC(cudaMalloc((void**)&w, (size_t)(nfw + 1) * 4)); IFC(cudaMalloc((void**)&irr, (size_t)(nfw + 1) * 4)); IFC(cudaMalloc((void**)&dcount, 8));
        IFC(dalloc(ctx, &ctx->irr_faces, (size_t)(fo1 - fo0) + 1));
        IFC(cudaMemsetAsync(w, 0, (size_t)(nfw + 1) * 4, sm)); IFC(cudaMemsetAsync(irr, 0, (size_t)(nfw + 1) * 4, sm)); IFC(cudaMemsetAsync(dcount, 0, 8, sm));
        int* const wg = w - fw0; int* const irrg = irr - fw0;      // addressed by global face id
        ifc_face_weight_kernel<<<148 * 8, 256, 0, sm>>>(g, fw0, fw1, ctx->face_loc, wg);
        if (ctx->ncut > 0) ifc_flag_irregular_kernel<<<(ctx->ncut + 255) / 256, 256, 0, sm>>>(g, ctx->ncut, ctx->cut_cells, irrg, dcount + 1);
        IFC(dalloc(ctx, &ctx->ftab_a, (size_t)nfw + 1)); ctx->ftab = ctx->ftab_a - fw0;
        size_t tb = 0;
        IFC(cub::DeviceScan::ExclusiveSum(nullptr, tb, w, ctx->ftab_a, (int)(nfw + 1), sm));
        IFC(cudaMalloc(&tmp, tb));
        IFC(cub::DeviceScan::ExclusiveSum(tmp, tb, w, ctx->ftab_a, (int)(nfw + 1), sm));
        cudaFree(tmp); tmp = nullptr; tb = 0;
        IFC(cub::DeviceSelect::Flagged(nullptr, tb, cub::CountingInputIterator<int>((int)fo0), irrg + fo0, ctx->irr_faces, dcount, (int)(fo1 - fo0), sm));
        IFC(cudaMalloc(&tmp, tb));
        IFC(cub::DeviceSelect::Flagged(tmp, tb, cub::CountingInputIterator<int>((int)fo0), irrg + fo0, ctx->irr_faces, dcount, (int)(fo1 - fo0), sm));
        // window-relative slots of the six faces that delimit the ranges (all multiples of the face-row length inside the window)
        const long long flast = std::max(fo0, (g.row_end - 1) * per), ffirst_end = std::min(fo0 + per, fw1);
        const long long probe[6] = {fw0, fo0, ffirst_end, flast, fo1, fw1};
        int rel[6] = {0, 0, 0, 0, 0, 0}, nirr = 0, ndir_cut = 0;
        for (int i = 0; i < 6; i++) IFC(cudaMemcpyAsync(&rel[i], ctx->ftab + probe[i], 4, cudaMemcpyDeviceToHost, sm));
        IFC(cudaMemcpyAsync(&nirr, dcount, 4, cudaMemcpyDeviceToHost, sm));
        IFC(cudaMemcpyAsync(&ndir_cut, dcount + 1, 4, cudaMemcpyDeviceToHost, sm));
        IFC(cudaStreamSynchronize(sm));
        cudaFree(tmp); tmp = nullptr;
        if (ndir_cut > 0) { release(); return fail(ctx, PB_ERR_INVALID, "Dirichlet boundary on cut cell not supported."); }      // interface_assembler::assemble_cut, :1292-1293
        // the one global piece: cut faces of the strips below, and of all strips
        const long long cut_ghost = (rel[1] - rel[0]) - (strip.cf_begin - strip.col_lo);                  // cut faces of the ghost face row
        const long long cut_owned = (rel[4] - rel[1]) - (strip.cf_end - strip.cf_begin);
        long long cut_before = 0, cut_total = cut_owned;
        if (ctx->world > 1 && ctx->comm) { pb_status st = dist_gather_counts(ctx, cut_owned, &cut_before, &cut_total); if (st != PB_OK) { release(); return st; } }
        else if (!whole) {
            if (ctx->ifc_cut_before < 0 || ctx->ifc_cut_total < ctx->ifc_cut_before + cut_owned) {
                release();
                return fail(ctx, PB_ERR_STATE, "pb_assemble: an interface problem on a strip needs the cut faces of the other strips: pb_dist_init, or pb_interface_set_cut_offsets "
                                               "(this strip owns %lld, from pb_interface_cut_faces_owned)", cut_owned);
            }
            cut_before = ctx->ifc_cut_before; cut_total = ctx->ifc_cut_total;
        }
        const long long base = strip.col_lo + cut_before - cut_ghost;                                    // slot of the first face of the window
        const long long nglob = (num_free_faces(g) + cut_total) * F;
        if (nglob >= (1ll << 31)) { release(); return fail(ctx, PB_ERR_UNSUPPORTED, "pb_assemble: interface system too large for 32-bit column indices"); }
        if (base) ifc_add_base_kernel<<<148 * 4, 256, 0, sm>>>(fw0, fw1, (int)base, ctx->ftab);
        const long long sb = base + rel[1], se = base + rel[4];
        const int nslots = (int)(se - sb);
        ctx->rb_lo = base + rel[0]; ctx->rb_begin = sb; ctx->rb_end = se; ctx->rb_hi = base + rel[5];
        ctx->halo_prev = rel[2] - rel[1]; ctx->halo_next = rel[4] - rel[3];
        ctx->ifc_nglob = nglob;
        const long long n = (long long)nslots * F;
        ctx->nrows = n; ctx->nrows_blk = nslots; ctx->n_irr_faces = nirr;
        { pb_status st = ifc_sorted_irregular(ctx, R); if (st != PB_OK) { release(); return st; } }
        IFC(dalloc(ctx, &ctx->brows, (size_t)nslots + 1)); IFC(dalloc(ctx, &ctx->brp, (size_t)nslots + 1));
        IFC(cudaMalloc((void**)&nb_all, (size_t)(nslots + 1) * 4));
        IFC(cudaMemsetAsync(nb_all, 0, (size_t)(nslots + 1) * 4, sm));
        ifc_block_rows_kernel<<<148 * 8, 256, 0, sm>>>(g, F, fo0, fo1, ctx->ftab, irrg, ctx->face_loc, nglob, R.uk, R.nuniq, c0, ctx->brows - sb, nb_all - sb);
        tb = 0;
        IFC(cub::DeviceScan::ExclusiveScan(nullptr, tb, nb_all, ctx->brp, cuda::std::plus<>(), 0LL, nslots + 1, sm));
        IFC(cudaMalloc(&tmp, tb));
        IFC(cub::DeviceScan::ExclusiveScan(tmp, tb, nb_all, ctx->brp, cuda::std::plus<>(), 0LL, nslots + 1, sm));
        long long nblk = 0;
        IFC(cudaMemcpyAsync(&nblk, ctx->brp + nslots, 8, cudaMemcpyDeviceToHost, sm));
        IFC(cudaStreamSynchronize(sm));
        cudaFree(tmp); tmp = nullptr;
        ctx->nnz = nblk * F * F;
        IFC(dalloc(ctx, &ctx->rowptr, (size_t)n + 1)); IFC(dalloc(ctx, &ctx->col, (size_t)(ctx->nnz > 0 ? ctx->nnz : 1)));
        IFC(dalloc(ctx, &ctx->val, (size_t)(ctx->nnz > 0 ? ctx->nnz : 1))); IFC(dalloc(ctx, &ctx->rhs, (size_t)(n > 0 ? n : 1)));
        ifc_pattern_kernel<<<148 * 8, 256, 0, sm>>>(nslots, F, ctx->brows, nb_all, ctx->brp, ctx->rowptr, ctx->col);
        // list of the irregular block rows (rhs fix-up, scalar SpMV pass of pb_pcg), relative to the first owned slot
        IFC(cudaMalloc((void**)&fl, (size_t)(nslots + 1) * 4));
        ifc_irregular_rows_kernel<<<148 * 4, 256, 0, sm>>>(nslots, F, ctx->brows, nb_all, fl);
        IFC(dalloc(ctx, &ctx->irr_rows, (size_t)(2 * (nirr > 0 ? nirr : 1))));
        tb = 0;
        IFC(cub::DeviceSelect::Flagged(nullptr, tb, cub::CountingInputIterator<int>(0), fl, ctx->irr_rows, dcount, nslots, sm));
        IFC(cudaMalloc(&tmp, tb));
        IFC(cub::DeviceSelect::Flagged(tmp, tb, cub::CountingInputIterator<int>(0), fl, ctx->irr_rows, dcount, nslots, sm));
        int nir = 0;
        IFC(cudaMemcpyAsync(&nir, dcount, 4, cudaMemcpyDeviceToHost, sm));
        IFC(cudaStreamSynchronize(sm));
        ctx->n_irr_rows = nir;
        cudaFree(w); cudaFree(irr); cudaFree(dcount); cudaFree(nb_all); cudaFree(fl); cudaFree(tmp);
#undef IFC
        ctx->ifc_struct_ok = true;
    } else {
        pb_status st = ifc_sorted_irregular(ctx, R); if (st != PB_OK) return st;
    }
    // ---- values: irregular rows from the sorted entries (their columns too), regular rows and their rhs by K3, rhs of the irregular rows
    const long long row0 = ctx->rb_begin * F;
    ifc_irregular_fill_kernel<<<148 * 4, 256, 0, sm>>>(R.nuniq, ctx->ifc_nglob, R.uk, R.uv, row0, ctx->rowptr, ctx->col, ctx->val);
    { pb_status st = scatter_system(ctx); if (st != PB_OK) { R.release(); return st; } }
    if (ctx->n_irr_rows > 0) {
        ifc_zero_rhs_rows_kernel<<<(ctx->n_irr_rows * F + 255) / 256, 256, 0, sm>>>(ctx->n_irr_rows, F, ctx->irr_rows, ctx->rhs);
        ifc_irregular_rhs_kernel<<<148 * 2, 256, 0, sm>>>(g, F, ctx->n_irr_faces, ctx->irr_faces, ctx->cell_loc_g, ctx->ncut, ctx->cut_cells, ctx->face_loc, ctx->ftab, c0, ctx->g, ctx->gx,
                                                         ctx->rhs - row0);
    }
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaStreamSynchronize(sm);
    R.release();
    if (e != cudaSuccess) return fail(ctx, PB_ERR_CUDA, "pb_assemble: interface system kernels failed: %s", cudaGetErrorString(e));
    return PB_OK;
}

/* strips: cut faces among the faces this context owns, and what the host tells a strip context without a communicator about the
 * others (cut faces owned by the strips below it, by all strips).  Ranks of a communicator do this themselves in pb_assemble. */
extern "C" pb_status pb_interface_cut_faces_owned(pb_ctx* ctx, long long* count) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->have_mesh && ctx->have_cut && count, PB_ERR_STATE, "pb_interface_cut_faces_owned: needs pb_cut_preprocess / pb_cut_upload first");
    CK(cudaSetDevice(ctx->device));
    return ifc_count_owned_cut_faces(ctx, count);
}
extern "C" pb_status pb_interface_set_cut_offsets(pb_ctx* ctx, long long cut_faces_before, long long cut_faces_total) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(cut_faces_before >= 0 && cut_faces_total >= cut_faces_before, PB_ERR_INVALID, "pb_interface_set_cut_offsets: 0 <= before <= total");
    ctx->ifc_cut_before = cut_faces_before; ctx->ifc_cut_total = cut_faces_total; ctx->ifc_struct_ok = false;
    return PB_OK;
}
/* out = { first owned slot, one past the last, first / one-past-last referenced slot, slots of the whole system, block rows the
 * lower / the upper neighbour needs of this strip } of the assembled interface problem */
extern "C" pb_status pb_interface_strip_layout(pb_ctx* ctx, long long out[7]) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->assembled && ctx->prm.problem == PB_PROBLEM_INTERFACE && out, PB_ERR_STATE, "pb_interface_strip_layout: no interface problem assembled");
    out[0] = ctx->rb_begin; out[1] = ctx->rb_end; out[2] = ctx->rb_lo; out[3] = ctx->rb_hi; out[4] = ctx->ifc_nglob / ctx->F; out[5] = ctx->halo_prev; out[6] = ctx->halo_next;
    return PB_OK;
}

/* test / driver accessors: per-cut-cell arrays in ascending cell order (what: 0 condensed block (8 fbs)^2 column-major,
 * 1 condensed rhs, 2 local matrix (2 cbs + 8 fbs)^2 [keep_local], 3 local load [keep_local]) and face_table (-1 = Dirichlet) */
extern "C" pb_status pb_interface_get_local(pb_ctx* ctx, int what, double* out) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->assembled && ctx->prm.problem == PB_PROBLEM_INTERFACE && out, PB_ERR_STATE, "pb_interface_get_local: no interface problem assembled");
    CK(cudaSetDevice(ctx->device));
    const int C = ctx->C, NF = ctx->NF, N2 = 2 * C + 2 * NF;
    const double* src = nullptr; long long per = 0;
    switch (what) {
    case 0: src = ctx->Sx; per = 4ll * NF * NF; break;
    case 1: src = ctx->gx; per = 2 * NF; break;
    case 2: src = ctx->lcx; per = (long long)N2 * N2; break;
    case 3: src = ctx->fx; per = 2 * C; break;
    default: return fail(ctx, PB_ERR_INVALID, "pb_interface_get_local: unknown selector %d", what);
    }
    NEED(src, PB_ERR_STATE, "pb_interface_get_local: assemble with keep_local = 1 first");
    if (ctx->ncut > 0) CK(cudaMemcpy(out, src, (size_t)ctx->ncut * per * 8, cudaMemcpyDeviceToHost));
    return PB_OK;
}
extern "C" pb_status pb_interface_face_table(pb_ctx* ctx, long long* table) {
    if (!ctx) return PB_ERR_INVALID;
    NEED(ctx->assembled && ctx->prm.problem == PB_PROBLEM_INTERFACE && table && ctx->ftab, PB_ERR_STATE, "pb_interface_face_table: no interface problem assembled");
    CK(cudaSetDevice(ctx->device));
    // (strips: the faces outside the window of this context are reported as -2)
    std::vector<int> t((size_t)(ctx->fw1 - ctx->fw0) + 1);
    CK(cudaMemcpy(t.data(), ctx->ftab + ctx->fw0, t.size() * 4, cudaMemcpyDeviceToHost));
    for (long long f = 0; f < ctx->nfaces; f++) table[f] = (f < ctx->fw0 || f >= ctx->fw1) ? -2 : (t[f - ctx->fw0 + 1] == t[f - ctx->fw0] ? -1 : t[f - ctx->fw0]);
    return PB_OK;
}
