// C ABI — the two-material interface problem (`cuthho_square -i`): cut cells by K2i (pb_interface.cuh), uncut
// cells by K1, and the condensed face system in the numbering of interface_assembler (apps/cuthho/cuthho_square.cpp:1091-
// 1436): every cut face carries two sets of unknowns, face_table counts them in face order (:1161-1175), the second set
// of face f sits fbs after the first (:1301).  First version of this row: the global matrix is built the way the reference
// builds it (one triplet per local entry, sorted and summed — setFromTriplets, :1430-1434) with cub on the device, and the
// solve runs on the scalar-CSR SpMV; the closed-form block structure of the plain problems does not hold around cut faces.
// Whole-mesh contexts only.

// weight of every face in face_table: 0 Dirichlet, 2 cut, 1 otherwise
__global__ void ifc_face_weight_kernel(Grid g, const int* __restrict__ face_loc, int* __restrict__ w) {
    const long long nc = g.Nx * g.Ny;
    for (long long c = blockIdx.x * (long long)blockDim.x + threadIdx.x; c < nc; c += (long long)gridDim.x * blockDim.x) {
        const long long ci = c % g.Nx, cj = c / g.Nx;
        long long f[4]; cell_faces(g, ci, cj, f);
        for (int li = 0; li < 4; li++) w[f[li]] = cell_face_dirichlet(g, ci, cj, li) ? 0 : (face_loc[f[li]] == PB_ON_INTERFACE ? 2 : 1);
    }
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
    int grid = A.ncut < 148 * 4 ? A.ncut : 148 * 4;
    kern<<<grid, 32, smem, c->stream>>>(A);
    return cudaGetLastError();
}

// cut cells of the interface problem -> Sx, gx, recAx, recbx (and the local matrices when keep_local)
static pb_status launch_interface_cut(pb_ctx* ctx, const pb_assembly_params* p, int cd, int fd) {
    const int C = ctx->C, NF = ctx->NF, N2 = 2 * C + 2 * NF;
    const size_t nc = (size_t)(ctx->ncut > 0 ? ctx->ncut : 1);
    CK(dalloc(ctx, &ctx->Sx, nc * 4 * NF * NF)); CK(dalloc(ctx, &ctx->gx, nc * 2 * NF));
    CK(dalloc(ctx, &ctx->recAx, nc * 4 * NF * C)); CK(dalloc(ctx, &ctx->recbx, nc * 2 * C));
    if (p->keep_local) { CK(dalloc(ctx, &ctx->lcx, nc * N2 * N2)); CK(dalloc(ctx, &ctx->fx, nc * 2 * C)); }
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
static pb_status build_interface_system(pb_ctx* ctx) {
    const Grid& g = ctx->grid; const int F = ctx->F, NF = ctx->NF;
    const long long nc = g.Nx * g.Ny, nf = ctx->nfaces;
    cudaStream_t sm = ctx->stream;
    int* w = nullptr; CK(cudaMalloc((void**)&w, (size_t)(nf + 1) * 4));
    CK(cudaMemsetAsync(w, 0, (size_t)(nf + 1) * 4, sm));
    ifc_face_weight_kernel<<<148 * 8, 256, 0, sm>>>(g, ctx->face_loc, w);
    CK(dalloc(ctx, &ctx->ftab, (size_t)nf + 1));
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
    ctx->nrows = n; ctx->nnz = nnz; ctx->nrows_blk = nslots;
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
    std::vector<int> t((size_t)ctx->nfaces + 1);
    CK(cudaMemcpy(t.data(), ctx->ftab, t.size() * 4, cudaMemcpyDeviceToHost));
    for (long long f = 0; f < ctx->nfaces; f++) table[f] = t[f + 1] == t[f] ? -1 : t[f];
    return PB_OK;
}
