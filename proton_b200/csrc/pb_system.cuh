// K3 — structure of the condensed face system and the scatter of the per-cell condensed blocks into it.
// Replaces assembler::assemble / finalize (hho.hpp:344-406, 451-455): instead of pushing n_loc^2 triplets per cell
// and sorting them, the CSR pattern (union of the dense local blocks, explicit zeros kept — SURVEY App. A.9) is
// derived from the closed-form numbering and every CSR entry GATHERS its <= 2 contributions through a precomputed
// insertion map, so the result is deterministic and stores are coalesced.
#pragma once
#include "pb_device.cuh"

namespace pb {

// which rows this context owns: compressed face ids [cf_begin, cf_end); local cells = rows [cell_row0, row_end)
struct Strip {
    long long cf_begin, cf_end;      // owned block rows
    long long cell_row0;             // first local cell row (row_begin - 1 when a ghost row below exists)
    long long col_lo, col_hi;        // compressed face ids referenced by owned rows: [col_lo, col_hi)
};

// compressed face id -> (horizontal?, i, j)   (inverse of compress_h / compress_v)
PB_HD void decompress_face(const Grid& g, long long cf, bool& horiz, long long& i, long long& j) {
    if (cf < g.Nx - 1) { horiz = false; i = cf + 1; j = 0; return; }
    long long t = cf - (g.Nx - 1), per = 2 * g.Nx - 1;
    j = 1 + t / per; long long l = t % per;
    if (l == 0) { horiz = true; i = 0; }
    else if (l & 1) { horiz = true; i = (l + 1) / 2; }
    else { horiz = false; i = l / 2; }
}
// first compressed id of face-row j (faces with ids j*(2Nx+1)...)
PB_HD long long compress_row_start(const Grid& g, long long j) { return j == 0 ? 0 : (g.Nx - 1) + (j - 1) * (2 * g.Nx - 1); }

struct BlockRow {            // one per owned block row (free face)
    int cell[2];             // local cell index of the two cells sharing the face (always two: free faces are interior)
    unsigned char li[2];     // local face index of this face inside each cell
    unsigned char nb;        // number of column blocks
    unsigned char pad;
    int bcol[7];             // compressed face ids of the column blocks, ascending
    unsigned char src[2][7]; // local face index inside cell k that feeds column block b, 255 = none
};

__global__ void build_block_rows_kernel(Grid g, Strip st, BlockRow* rows, int* nb_out) {
    long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    long long nrows = st.cf_end - st.cf_begin;
    if (t >= nrows) return;
    long long cf = st.cf_begin + t;
    bool horiz; long long i, j; decompress_face(g, cf, horiz, i, j);
    long long ci[2], cj[2]; int li[2];
    if (horiz) { ci[0] = i; cj[0] = j - 1; li[0] = 2; ci[1] = i; cj[1] = j; li[1] = 0; }
    else       { ci[0] = i - 1; cj[0] = j; li[0] = 1; ci[1] = i; cj[1] = j; li[1] = 3; }
    BlockRow br;
    long long ids[8]; unsigned char s0[8], s1[8]; int n = 0;
    for (int k = 0; k < 2; k++) {
        br.cell[k] = (int)((cj[k] - st.cell_row0) * g.Nx + ci[k]);
        br.li[k] = (unsigned char)li[k];
        for (int lf = 0; lf < 4; lf++) {
            if (cell_face_dirichlet(g, ci[k], cj[k], lf)) continue;
            long long id = cell_face_compressed(g, ci[k], cj[k], lf);
            int pos = -1;
            for (int m = 0; m < n; m++) if (ids[m] == id) pos = m;
            if (pos < 0) { pos = n++; ids[pos] = id; s0[pos] = 255; s1[pos] = 255; }
            if (k == 0) s0[pos] = (unsigned char)lf; else s1[pos] = (unsigned char)lf;
        }
    }
    // insertion sort by compressed id (n <= 7)
    for (int a = 1; a < n; a++) {
        long long id = ids[a]; unsigned char x0 = s0[a], x1 = s1[a]; int b = a - 1;
        while (b >= 0 && ids[b] > id) { ids[b + 1] = ids[b]; s0[b + 1] = s0[b]; s1[b + 1] = s1[b]; b--; }
        ids[b + 1] = id; s0[b + 1] = x0; s1[b + 1] = x1;
    }
    br.nb = (unsigned char)n; br.pad = 0;
    for (int m = 0; m < 7; m++) { br.bcol[m] = m < n ? (int)ids[m] : -1; br.src[0][m] = m < n ? s0[m] : 255; br.src[1][m] = m < n ? s1[m] : 255; }
    rows[t] = br; nb_out[t] = n;
}

// rowptr / col of the scalar CSR from the block structure.  brp = exclusive scan of nb.
__global__ void fill_csr_pattern_kernel(long long nrows_blk, int F, const BlockRow* rows, const long long* brp,
                                        long long* rowptr, int* col) {
    long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;   // one thread per scalar row
    if (t > nrows_blk * F) return;
    if (t == nrows_blk * F) { rowptr[t] = brp[nrows_blk] * F * F; return; }
    long long br = t / F; int a = (int)(t - br * F);
    const BlockRow& R = rows[br];
    long long base = brp[br] * F * F + (long long)a * R.nb * F;
    rowptr[t] = base;
    for (int b = 0; b < R.nb; b++)
        for (int bc = 0; bc < F; bc++) col[base + b * F + bc] = R.bcol[b] * F + bc;
}

// values: one warp per block row.  The F scalar rows of a block row are consecutive in the CSR arrays, so a block row
// is one contiguous run of nb*F*F values: lanes walk it with unit stride (coalesced stores) and gather their <= 2 cell
// contributions from the two S blocks of the face (1152 B each at k=2: L1/L2 hits after the first touch).
// S is [cell][NF*NF] column-major (symmetric); entry (row li*F+a, col lj*F+bc).
__global__ void scatter_values_kernel(long long nrows_blk, int F, const BlockRow* __restrict__ rows, const long long* __restrict__ brp,
                                      const double* __restrict__ S, double* __restrict__ val) {
    const int lane = threadIdx.x & 31;
    const long long nwarps = (long long)gridDim.x * (blockDim.x >> 5);
    const int NF = 4 * F;
    for (long long br = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5; br < nrows_blk; br += nwarps) {
        const BlockRow R = rows[br];
        const int rowlen = R.nb * F, len = rowlen * F;
        double* out = val + brp[br] * F * F;
        const double* S0 = S + (long long)R.cell[0] * (NF * NF) + R.li[0] * F * NF;
        const double* S1 = S + (long long)R.cell[1] * (NF * NF) + R.li[1] * F * NF;
        for (int e = lane; e < len; e += 32) {
            const int a = e / rowlen, rem = e - a * rowlen, b = rem / F, bc = rem - b * F;
            const unsigned char l0 = R.src[0][b], l1 = R.src[1][b];
            double v = 0.0;
            // S is symmetric up to rounding: read the transposed entry so that bc runs over contiguous doubles
            if (l0 != 255) v += S0[a * NF + l0 * F + bc];
            if (l1 != 255) v += S1[a * NF + l1 * F + bc];
            out[e] = v;
        }
    }
}

// values + rhs with the face degree known at compile time: one warp per block row, lane = position (b, bc) inside a
// scalar row (<= 7*F entries, F <= 4), loop over the F scalar rows.  The lane's source offsets are fixed per block row,
// so the inner loop is two loads, one add, one store.  U block rows per warp pass; measured on B200 at k = 2, 2048^2:
// U = 1 with >= 32 blocks per SM is fastest (2.55 ms, 4.2 TB/s); U > 1 strides the rows of a warp apart and loses the
// L2 reuse of the S blocks between the four faces of a cell.  Lanes 28.. write the rhs rows.
template <int F, int U>
__global__ void __launch_bounds__(256) scatter_kernel_t(long long nrows_blk, const BlockRow* __restrict__ rows, const long long* __restrict__ brp,
                                                        const double* __restrict__ S, const double* __restrict__ g,
                                                        double* __restrict__ val, double* __restrict__ rhs) {
    static_assert(7 * F <= 28, "lanes 28..31 carry the rhs");
    constexpr int NF = 4 * F;
    const int lane = threadIdx.x & 31;
    const int b = lane / F, bc = lane - b * F;
    const long long nwarps = (long long)gridDim.x * (blockDim.x >> 5);
    for (long long br0 = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5; br0 < nrows_blk; br0 += U * nwarps) {
        double v[U][F];
        double* out[U]; int rowlen[U]; bool on[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const long long br = br0 + u * nwarps;
            on[u] = false; rowlen[u] = 0; out[u] = nullptr;
#pragma unroll
            for (int a = 0; a < F; a++) v[u][a] = 0.0;
            if (br < nrows_blk) {
                const BlockRow& R = rows[br];
                const int nb = R.nb;
                rowlen[u] = nb * F; out[u] = val + brp[br] * (F * F) + lane;
                const double* S0 = S + (long long)R.cell[0] * (NF * NF) + R.li[0] * F * NF;
                const double* S1 = S + (long long)R.cell[1] * (NF * NF) + R.li[1] * F * NF;
                if (lane < rowlen[u]) {
                    on[u] = true;
                    const int l0 = R.src[0][b], l1 = R.src[1][b];
                    // S is symmetric up to rounding: read the transposed entry so that bc runs over contiguous doubles
#pragma unroll
                    for (int a = 0; a < F; a++) {
                        double v0 = 0.0, v1 = 0.0;
                        if (l0 != 255) v0 = S0[a * NF + l0 * F + bc];
                        if (l1 != 255) v1 = S1[a * NF + l1 * F + bc];
                        v[u][a] = v0 + v1;
                    }
                } else if (lane >= 28 && lane - 28 < F) {
                    const int a = lane - 28;
                    v[u][0] = g[(long long)R.cell[0] * NF + R.li[0] * F + a] + g[(long long)R.cell[1] * NF + R.li[1] * F + a];
                }
            }
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (on[u]) {
#pragma unroll
                for (int a = 0; a < F; a++) out[u][a * rowlen[u]] = v[u][a];
            } else if (lane >= 28 && lane - 28 < F && br0 + u * nwarps < nrows_blk) {
                rhs[(br0 + u * nwarps) * F + (lane - 28)] = v[u][0];
            }
        }
    }
}

// the same for small face degrees (F = 1, 2): a scalar row has only 7F <= 14 entries, so 28/(7F) block rows sit side by side
// in one warp, lane = (row slot, position (b, bc) inside the scalar row).  (For F >= 3 this variant measured slower
// than the one-row kernel above: 3.5 vs 2.6 ms at k = 2.)  The lane's source offsets are fixed per block row,
// so the inner loop is two loads, one add, one store.  U block rows per warp pass; measured on B200 at k = 2, 2048^2:
// U = 1 with >= 32 blocks per SM is fastest (2.55 ms, 4.2 TB/s); U > 1 strides the rows of a warp apart and loses the
// L2 reuse of the S blocks between the four faces of a cell.  Lanes 28.. write the rhs rows.
template <int F, int U>
__global__ void __launch_bounds__(256) scatter_packed_kernel_t(long long nrows_blk, const BlockRow* __restrict__ rows, const long long* __restrict__ brp,
                                                        const double* __restrict__ S, const double* __restrict__ g,
                                                        double* __restrict__ val, double* __restrict__ rhs) {
    static_assert(7 * F <= 28, "lanes 28..31 carry the rhs");
    constexpr int NF = 4 * F, W = 7 * F, RPW = 28 / W;          // RPW block rows side by side in one warp (4 / 2 / 1 / 1 for F = 1..4)
    const int lane = threadIdx.x & 31;
    const int slot = lane < 28 ? lane / W : 0, pos = lane - slot * W;          // lanes < 28: (block row slot, position in the scalar row)
    const int b = pos / F, bc = pos - b * F;
    const int rs = lane >= 28 ? (lane - 28) / F : 0, ra = lane >= 28 ? (lane - 28) - rs * F : 0;      // lanes >= 28: (slot, scalar row) of a rhs entry
    const long long nwarps = (long long)gridDim.x * (blockDim.x >> 5);
    for (long long br0 = ((blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5) * RPW; br0 < nrows_blk; br0 += U * nwarps * RPW) {
        double v[U][F];
        double* out[U]; int rowlen[U]; bool on[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const long long br = br0 + u * nwarps * RPW + (lane < 28 ? slot : rs);
            on[u] = false; rowlen[u] = 0; out[u] = nullptr;
#pragma unroll
            for (int a = 0; a < F; a++) v[u][a] = 0.0;
            if (br < nrows_blk) {
                const BlockRow& R = rows[br];
                if (lane < 28) {
                    rowlen[u] = R.nb * F;
                    if (slot < RPW && pos < rowlen[u]) {
                        on[u] = true;
                        out[u] = val + brp[br] * (F * F) + pos;
                        const double* S0 = S + (long long)R.cell[0] * (NF * NF) + R.li[0] * F * NF;
                        const double* S1 = S + (long long)R.cell[1] * (NF * NF) + R.li[1] * F * NF;
                        const int l0 = R.src[0][b], l1 = R.src[1][b];
                        // S is symmetric up to rounding: read the transposed entry so that bc runs over contiguous doubles
#pragma unroll
                        for (int a = 0; a < F; a++) {
                            double v0 = 0.0, v1 = 0.0;
                            if (l0 != 255) v0 = S0[a * NF + l0 * F + bc];
                            if (l1 != 255) v1 = S1[a * NF + l1 * F + bc];
                            v[u][a] = v0 + v1;
                        }
                    }
                } else if (lane - 28 < RPW * F) {
                    v[u][0] = g[(long long)R.cell[0] * NF + R.li[0] * F + ra] + g[(long long)R.cell[1] * NF + R.li[1] * F + ra];
                    rowlen[u] = -1;          // marks a live rhs lane
                }
            }
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (on[u]) {
#pragma unroll
                for (int a = 0; a < F; a++) out[u][a * rowlen[u]] = v[u][a];
            } else if (rowlen[u] == -1) {
                rhs[(br0 + u * nwarps * RPW + rs) * F + ra] = v[u][0];
            }
        }
    }
}

// values + rhs, tile version (F = 2..4): a block of 256 threads owns T consecutive block rows.  Every S chunk a block row needs
// (the F columns of the face inside each of its two cells: F*NF contiguous doubles, 16-byte aligned) is read exactly once in
// the whole launch, so the read side is already minimal; what the one-warp-per-row kernel above loses is (a) memory-level
// parallelism — record -> chunk -> store are three dependent round trips per warp — and (b) sector fills: a block row's
// nb*F*F*8 bytes (504 at k = 2) start and end inside 32-byte sectors that another warp completes later, and ncu showed
// 0.77 GB of DRAM reads beyond the 5.77 GB the kernel asks for.  Here the records arrive as one coalesced copy, all 2T chunks
// are in flight together as 16-byte cp.async copies into shared memory, the tile's values are formed in shared memory and leave
// as ONE contiguous, fully coalesced span [brp[br0], brp[br0+T]) * F*F (16 KB at k = 2): only the two ends of a span are partial
// sectors.  Tiles that contain rows the interface problem fills elsewhere (nb = 0, or a span longer than nb blocks) take the
// per-row path, which writes exactly the entries scatter_kernel_t writes.
PB_D void cp_async16(void* smem_dst, const void* gsrc) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(gsrc) : "memory");
}
PB_D void cp_async_wait_all() { asm volatile("cp.async.wait_all;\n" ::: "memory"); }

template <int F, int T>
struct ScatterTileSmem {
    static constexpr int NF = 4 * F, CH = F * NF;          // doubles per chunk
    double chunk[T][2][CH];
    double outv[T * 7 * F * F];
    double rhsv[T * F];
    long long off[T + 1];
    BlockRow rec[T];
    int regular;
};

template <int F, int T>
__global__ void __launch_bounds__(256) scatter_tile_kernel(long long nrows_blk, const BlockRow* __restrict__ rows, const long long* __restrict__ brp,
                                                           const double* __restrict__ S, const double* __restrict__ g,
                                                           double* __restrict__ val, double* __restrict__ rhs) {
    using SM = ScatterTileSmem<F, T>;
    constexpr int NF = 4 * F, CH = SM::CH, V = CH / 2;      // V 16-byte vectors per chunk
    static_assert(sizeof(BlockRow) % 4 == 0 && CH % 2 == 0, "records are copied as ints, chunks as double2");
    extern __shared__ __align__(16) unsigned char scatter_smem_raw[];
    SM& s = *reinterpret_cast<SM*>(scatter_smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long ntiles = (nrows_blk + T - 1) / T;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const long long br0 = tile * T;
        const int cnt = (int)(nrows_blk - br0 < T ? nrows_blk - br0 : T);
        // records and offsets: contiguous in memory
        {
            const int* src = reinterpret_cast<const int*>(rows + br0); int* dst = reinterpret_cast<int*>(s.rec);
            for (int i = tid; i < cnt * (int)(sizeof(BlockRow) / 4); i += 256) dst[i] = src[i];
            for (int i = tid; i <= cnt; i += 256) s.off[i] = brp[br0 + i];
            if (tid == 0) s.regular = 1;
        }
        __syncthreads();
        // all chunks of the tile in flight at once
        for (int i = tid; i < cnt * 2 * V; i += 256) {
            const int t = i / (2 * V), r = i - t * (2 * V), k = r / V, v = r - k * V;
            const BlockRow& R = s.rec[t];
            const double* src = S + (long long)R.cell[k] * (NF * NF) + R.li[k] * CH + 2 * v;
            cp_async16(&s.chunk[t][k][2 * v], src);
        }
        for (int i = tid; i < cnt * F; i += 256) {
            const int t = i / F, a = i - t * F;
            const BlockRow& R = s.rec[t];
            s.rhsv[i] = g[(long long)R.cell[0] * NF + R.li[0] * F + a] + g[(long long)R.cell[1] * NF + R.li[1] * F + a];
        }
        for (int t = tid; t < cnt; t += 256)
            if (s.rec[t].nb == 0 || s.off[t + 1] - s.off[t] != s.rec[t].nb) s.regular = 0;      // benign race: every writer stores 0
        cp_async_wait_all();
        __syncthreads();
        const bool regular = s.regular != 0;
        const long long base = s.off[0] * (F * F);
        // lane = position (b, bc) inside a scalar row; the transposed entry is read so that bc runs over contiguous doubles (S is
        // symmetric up to rounding), exactly the sums of scatter_kernel_t
        const int b = lane / F, bc = lane - b * F;
        for (int t = warp; t < cnt; t += 8) {
            const BlockRow& R = s.rec[t];
            const int rowlen = R.nb * F;
            if (lane < rowlen) {
                const int l0 = R.src[0][b], l1 = R.src[1][b];
                const long long o = s.off[t] * (F * F) - base;
#pragma unroll
                for (int a = 0; a < F; a++) {
                    double v0 = 0.0, v1 = 0.0;
                    if (l0 != 255) v0 = s.chunk[t][0][a * NF + l0 * F + bc];
                    if (l1 != 255) v1 = s.chunk[t][1][a * NF + l1 * F + bc];
                    if (regular) s.outv[o + a * rowlen + lane] = v0 + v1;
                    else val[base + o + a * rowlen + lane] = v0 + v1;
                }
            }
        }
        __syncthreads();
        if (regular) {
            const int nout = (int)(s.off[cnt] * (F * F) - base);
            for (int i = tid; i < nout; i += 256) val[base + i] = s.outv[i];
        }
        for (int i = tid; i < cnt * F; i += 256) rhs[br0 * F + i] = s.rhsv[i];
        __syncthreads();
    }
}

// rhs: one thread per scalar row; g is [cell][NF] (already Dirichlet-lifted)
__global__ void scatter_rhs_kernel(long long nrows_blk, int F, const BlockRow* rows, const double* __restrict__ g, double* __restrict__ rhs) {
    long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (t >= nrows_blk * F) return;
    long long br = t / F; int a = (int)(t - br * F);
    const BlockRow& R = rows[br];
    const int NF = 4 * F;
    rhs[t] = g[(long long)R.cell[0] * NF + R.li[0] * F + a] + g[(long long)R.cell[1] * NF + R.li[1] * F + a];
}

}  // namespace pb
