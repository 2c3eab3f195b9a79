// Line-block Jacobi for the Chebyshev smoother of the multigrid preconditioner (pb_mg.cuh): D = the block-tridiagonal parts of A along
// the mesh lines instead of the fbs x fbs diagonal blocks.  NOT reference behaviour (the whole preconditioner is not).
//
// Why.  The cut / naive stabilisation of the reference is scaled by 1/area instead of 1/diameter, so in the condensed
// system the stabilisation dominates like 1/h.  Its near-kernel couples the two faces of a cell that lie OPPOSITE each other (the
// bottom and top face of a cell carry the same trace of x^2, the left and right face that of y^2): along a column of horizontal faces,
// or a row of vertical faces, the coupling is O(1/h) stronger than across.  Modes that are constant along such a line and oscillate
// from line to line have almost no energy, are invisible to the vertex coarse space (a face mean that alternates between neighbouring
// faces is no vertex interpolant) and are not damped by a point smoother, whose diagonal is dominated by the strong coupling: the
// smallest eigenvalues of M^-1 A are exactly these (tools/scratch/mg_slow_modes.py: 0.666 at 64^2, k = 1, pure dof-0 modes on the
// horizontal OR the vertical faces, alternating sign from column to column), and the CG iteration count grows like sqrt(N).  Solving
// along the lines removes them: 9 / 10 / 9 iterations at 64^2 / 128^2 / 256^2 against 14 / 19 / 22 on the restatement
// (tools/scratch/line_smoother_experiment.py).
//
// Lines.  Horizontal faces h(i, j) of a column i, j ascending (coupled through cell (i, j)); vertical faces v(i, j) of a cell row j,
// i ascending (coupled through cell (i, j)).  Every free face lies on exactly one line.  A strip of cell rows owns whole rows of
// vertical faces; its columns of horizontal faces end at the strip (the segments of a column on different ranks are solved
// independently: block Jacobi between strips, 16 against 14 iterations at 8 strips of 256^2 on the restatement).
// Block Thomas factorisation per line, S_1 = D_1, S_k = D_k - U_{k-1}^T S_{k-1}^{-1} U_{k-1}: stored are Sinv_k = S_k^{-1} and
// G_k = S_k^{-1} U_k (the transposed multiplier is G^T because S is symmetric), as a structure of arrays [step][entry][line] so that every
// load of a warp, one line per thread, is one 128-byte line (sets of few long lines: [line][step][entry], one warp per line).
// Solve: y_k = r_k - G_{k-1}^T y_{k-1}, t_k = Sinv_k y_k, x_k = t_k - G_k x_{k+1}.
#pragma once
#include "pb_mg.cuh"

namespace pb {

struct MgLineSet {          // the lines of one direction on the owned block rows (local block row = compressed face - first owned)
    int nlines, nsteps;     // threads, blocks per line
    int kind;               // 0: columns of horizontal faces (line = i, step = j - j0); 1: rows of vertical faces (line = j - row_begin, step = i - 1)
    long long base, per;    // closed-form numbering (compress_h / compress_v, pb_device.cuh): see mg_line_block_row
    int row0_first;         // kind 1: line 0 is cell row 0, whose vertical faces are numbered densely from 0
    int warp_per_line;      // 1: one warp per line (few long lines), line-major factors; 0: one thread per line, structure-of-arrays factors (mg_line_fac_elem)
};
// element e (< 2 F^2) of block (line, step).  One thread per line: structure of arrays, [step][e][line], so that every load of a warp is one
// 128-byte line (the array-of-blocks layout cost 18 wavefronts per 4-byte load: 3.5 ms per solve at 4096^2).  One warp per line: the blocks of a
// line are contiguous, [line][step][e], and a chunk of 32 steps is copied linearly.
PB_HD long long mg_line_fac_elem(const MgLineSet& S, int FF2, int line, int step, int e) {
    return S.warp_per_line ? ((long long)line * S.nsteps + step) * FF2 + e : ((long long)step * FF2 + e) * S.nlines + line;
}
PB_HD long long mg_line_block_row(const MgLineSet& S, int line, int step) {
    if (S.kind == 0) return S.base + (long long)step * S.per + (line ? 2 * line - 1 : 0);
    if (S.row0_first && line == 0) return step;
    return S.base + (long long)line * S.per + 2 * (step + 1);
}
// the two line sets of a strip of cell rows [row_begin, row_end) whose first owned compressed face is cf_begin
PB_HD void mg_line_sets(const Grid& g, long long row_begin, long long row_end, long long cf_begin, MgLineSet& H, MgLineSet& V) {
    const long long per = 2 * g.Nx - 1, j0 = row_begin > 1 ? row_begin : 1;
    H.kind = 0; H.nlines = (int)g.Nx; H.nsteps = (int)(row_end - j0 > 0 ? row_end - j0 : 0); H.per = per; H.row0_first = 0;
    H.base = compress_row_start(g, j0) - cf_begin;
    V.kind = 1; V.nlines = (int)(row_end - row_begin); V.nsteps = (int)(g.Nx - 1); V.per = per; V.row0_first = row_begin == 0 ? 1 : 0;
    V.base = (row_begin == 0 ? (g.Nx - 1) - per : compress_row_start(g, row_begin)) - cf_begin;
    // few long lines (the rows of vertical faces of a narrow strip: their length does not shrink with the number of ranks): one warp per line
    for (MgLineSet* S : {&H, &V}) S->warp_per_line = (S->nlines < 2048 && S->nsteps >= 256) ? 1 : 0;
}

#ifdef __CUDACC__
// F x F block of the CSR row block `br` (local) at the column block whose first global column is gcol0; false = not stored
template <int F>
__device__ __forceinline__ bool mg_csr_block(const long long* __restrict__ rowptr, const int* __restrict__ col, const double* __restrict__ val, long long br,
                                             long long gcol0, double B[F][F]) {
    const long long k0 = rowptr[br * F], k1 = rowptr[br * F + 1];
    long long off = -1;
    for (long long k = k0; k < k1; k += F) if (col[k] == (int)gcol0) off = k - k0;
    if (off < 0) return false;
#pragma unroll
    for (int a = 0; a < F; a++) {
        const long long ka = rowptr[br * F + a] + off;
#pragma unroll
        for (int c = 0; c < F; c++) B[a][c] = val[ka + c];
    }
    return true;
}
template <int F>
__device__ __forceinline__ void mg_invert_spd(double A[F][F], double B[F][F]) {      // Gauss-Jordan without pivoting
#pragma unroll
    for (int a = 0; a < F; a++)
#pragma unroll
        for (int c = 0; c < F; c++) B[a][c] = a == c ? 1.0 : 0.0;
#pragma unroll
    for (int k = 0; k < F; k++) {
        const double ip = 1.0 / A[k][k];
#pragma unroll
        for (int c = 0; c < F; c++) { A[k][c] *= ip; B[k][c] *= ip; }
#pragma unroll
        for (int a = 0; a < F; a++) {
            if (a == k) continue;
            const double f = A[a][k];
#pragma unroll
            for (int c = 0; c < F; c++) { A[a][c] -= f * A[k][c]; B[a][c] -= f * B[k][c]; }
        }
    }
}

// steps whose operands a thread loads ahead of the recurrence (all loads of a chunk are independent of it and in flight together;
// without the staging every step waited for its own DRAM round trip: 8.5 ms per solve at 4096^2 against 0.9 ms)
// (k = 1: 16 steps with 24 bytes of spills measured 4.29 ms per iteration at 2048^2 against 4.40 with 12 and none; k = 3: 4 steps, no spills)
template <int F> struct MgLineChunk { static constexpr int value = F <= 2 ? 16 : F == 3 ? 8 : 4; };

// factorisation, phase 1 (one thread per step of a line): the raw blocks D_k and U_k into the slots of Sinv_k and G_k
template <int F, typename VT>
__global__ void mg_line_gather_kernel(MgLineSet S, long long row0, const long long* __restrict__ rowptr, const int* __restrict__ col, const double* __restrict__ val,
                                      VT* __restrict__ fac) {
    const long long total = (long long)S.nlines * S.nsteps;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        const int k = (int)(t / S.nlines), line = (int)(t - (long long)k * S.nlines);
        const long long br = mg_line_block_row(S, line, k);
        double D[F][F], U[F][F];
        mg_csr_block<F>(rowptr, col, val, br, row0 + br * F, D);
        bool have_u = false;
        if (k + 1 < S.nsteps) have_u = mg_csr_block<F>(rowptr, col, val, br, row0 + mg_line_block_row(S, line, k + 1) * F, U);
#pragma unroll
        for (int a = 0; a < F; a++)
#pragma unroll
            for (int c = 0; c < F; c++) {
                fac[mg_line_fac_elem(S, 2 * F * F, line, k, a * F + c)] = (VT)D[a][c];
                fac[mg_line_fac_elem(S, 2 * F * F, line, k, F * F + a * F + c)] = (VT)(have_u ? U[a][c] : 0.0);
            }
    }
}
// factorisation, phase 2 (one thread per line): D_k, U_k -> Sinv_k, G_k in place
template <int F, typename VT>
__global__ void mg_line_factor_kernel(MgLineSet S, VT* __restrict__ fac) {
    constexpr int CH = MgLineChunk<F>::value, FF = F * F;
    const int line = blockIdx.x * blockDim.x + threadIdx.x;
    if (line >= S.nlines) return;
    // element e of step k of this line: fac[k * fstride + e * estride]
    const long long fstride = S.warp_per_line ? 2 * FF : (long long)(2 * FF) * S.nlines, estride = S.warp_per_line ? 1 : S.nlines;
    fac += S.warp_per_line ? (long long)line * S.nsteps * (2 * FF) : line;
    double G[F][F], Uprev[F][F];
#pragma unroll
    for (int a = 0; a < F; a++)
#pragma unroll
        for (int c = 0; c < F; c++) { G[a][c] = 0.0; Uprev[a][c] = 0.0; }
    for (int k0 = 0; k0 < S.nsteps; k0 += CH) {
        VT raw[CH][2 * FF];
#pragma unroll
        for (int q = 0; q < CH; q++) {
            const VT* f = fac + (long long)min(k0 + q, S.nsteps - 1) * fstride;
#pragma unroll
            for (int e = 0; e < 2 * FF; e++) raw[q][e] = f[e * estride];
        }
#pragma unroll
        for (int q = 0; q < CH; q++) {
            if (k0 + q >= S.nsteps) break;
            double D[F][F], Si[F][F];
#pragma unroll
            for (int a = 0; a < F; a++)
#pragma unroll
                for (int c = 0; c < F; c++) {
                    double s = (double)raw[q][a * F + c];
#pragma unroll
                    for (int e = 0; e < F; e++) s -= Uprev[e][a] * G[e][c];      // D_k - U_{k-1}^T G_{k-1}
                    D[a][c] = s;
                }
#pragma unroll
            for (int a = 0; a < F; a++)
#pragma unroll
                for (int c = a + 1; c < F; c++) { const double m = 0.5 * (D[a][c] + D[c][a]); D[a][c] = m; D[c][a] = m; }
            mg_invert_spd<F>(D, Si);
            VT* o = fac + (long long)(k0 + q) * fstride;
#pragma unroll
            for (int a = 0; a < F; a++)
#pragma unroll
                for (int c = 0; c < F; c++) {
                    double s = 0.0;
#pragma unroll
                    for (int e = 0; e < F; e++) s += Si[a][e] * (double)raw[q][FF + e * F + c];
                    G[a][c] = s;
                    o[(a * F + c) * estride] = (VT)(0.5 * (Si[a][c] + Si[c][a])); o[(FF + a * F + c) * estride] = (VT)s;
                }
#pragma unroll
            for (int a = 0; a < F; a++)
#pragma unroll
                for (int c = 0; c < F; c++) Uprev[a][c] = (double)raw[q][FF + a * F + c];
        }
    }
}

// t = L^{-1} r on one line, one thread: operands of CH steps are loaded ahead of the recurrence
template <int F, typename VT>
__device__ __forceinline__ void mg_line_solve_thread(const MgLineSet& S, const VT* __restrict__ fac, const double* __restrict__ r, double* __restrict__ t, int line) {
    constexpr int CH = MgLineChunk<F>::value, FF = F * F;
    const long long fstride = (long long)(2 * FF) * S.nlines, estride = S.nlines;      // structure of arrays: [step][e][line]
    fac += line;
    const int m = S.nsteps;
    // forward: y_k = r_k - G_{k-1}^T y_{k-1};  t_k = Sinv_k y_k
    double y[F], gp[FF];      // y_{k-1}, G_{k-1}
#pragma unroll
    for (int a = 0; a < F; a++) y[a] = 0.0;
#pragma unroll
    for (int e = 0; e < FF; e++) gp[e] = 0.0;
    for (int k0 = 0; k0 < m; k0 += CH) {
        VT f[CH][2 * FF]; double rr[CH][F]; long long e0[CH];
#pragma unroll
        for (int q = 0; q < CH; q++) {
            const int k = min(k0 + q, m - 1);
            e0[q] = mg_line_block_row(S, line, k) * F;
            const VT* fk = fac + (long long)k * fstride;
#pragma unroll
            for (int e = 0; e < 2 * FF; e++) f[q][e] = fk[e * estride];
#pragma unroll
            for (int a = 0; a < F; a++) rr[q][a] = r[e0[q] + a];
        }
#pragma unroll
        for (int q = 0; q < CH; q++) {
            if (k0 + q >= m) break;
            double yn[F];
#pragma unroll
            for (int a = 0; a < F; a++) {
                double s = rr[q][a];
#pragma unroll
                for (int e = 0; e < F; e++) s -= gp[e * F + a] * y[e];
                yn[a] = s;
            }
#pragma unroll
            for (int a = 0; a < F; a++) y[a] = yn[a];
#pragma unroll
            for (int a = 0; a < F; a++) {
                double s = 0.0;
#pragma unroll
                for (int e = 0; e < F; e++) s += (double)f[q][a * F + e] * y[e];
                t[e0[q] + a] = s;
            }
#pragma unroll
            for (int e = 0; e < FF; e++) gp[e] = (double)f[q][FF + e];
        }
    }
    // backward: x_k = t_k - G_k x_{k+1}
    double x[F];
#pragma unroll
    for (int a = 0; a < F; a++) x[a] = 0.0;
    for (int k1 = m - 1; k1 >= 0; k1 -= CH) {
        VT g[CH][FF]; double tt[CH][F]; long long e0[CH];
#pragma unroll
        for (int q = 0; q < CH; q++) {
            const int k = max(k1 - q, 0);
            e0[q] = mg_line_block_row(S, line, k) * F;
            const VT* fk = fac + (long long)k * fstride + FF * estride;
#pragma unroll
            for (int e = 0; e < FF; e++) g[q][e] = fk[e * estride];
#pragma unroll
            for (int a = 0; a < F; a++) tt[q][a] = t[e0[q] + a];
        }
#pragma unroll
        for (int q = 0; q < CH; q++) {
            if (k1 - q < 0) break;
            double xn[F];
#pragma unroll
            for (int a = 0; a < F; a++) {
                double s = tt[q][a];
#pragma unroll
                for (int e = 0; e < F; e++) s -= (double)g[q][a * F + e] * x[e];
                xn[a] = s;
            }
#pragma unroll
            for (int a = 0; a < F; a++) { x[a] = xn[a]; t[e0[q] + a] = xn[a]; }
        }
    }
}

// the same on one line by one WARP (few long lines).  Per chunk of 32 steps: the chunk's factor blocks are one contiguous piece, copied with
// coalesced loads (issued a chunk ahead, in flight during the recurrence) and widened to double on the way into shared memory; lane 0 runs the
// bare recurrence (y_k = r_k - G_{k-1}^T y_{k-1}, resp. x_k = t_k - G_k x_{k+1}: 9 multiply-adds per step at k = 2), unrolled so that its
// shared-memory loads run ahead; the products t_k = Sinv_k y_k are then taken by all lanes, one step each.
template <int F, typename VT> struct MgLineWarpSmem { double f[32][2 * F * F]; double r[32][F]; double o[32][F]; };
template <int F, typename VT>
__device__ __forceinline__ void mg_line_solve_warp(const MgLineSet& S, const VT* __restrict__ fac, const double* __restrict__ r, double* __restrict__ t, int line,
                                                   MgLineWarpSmem<F, VT>& sm) {
    constexpr int FF = F * F, W = 2 * FF;      // words of a block; a chunk = 32 W consecutive words = W loads per lane
    const int lane = threadIdx.x & 31, m = S.nsteps, nch = (m + 31) / 32;
    fac += (long long)line * m * W;
    const long long nwords = (long long)m * W;
    VT nf[W]; double nr[F];
    double* const sf = &sm.f[0][0];
    auto fetch = [&](int k_first, int k_mine, const double* vec) {      // words of the 32 blocks from k_first on; vector entries of step k_mine
        const long long w0 = (long long)k_first * W;
#pragma unroll
        for (int j = 0; j < W; j++) { const long long w = w0 + j * 32 + lane; nf[j] = fac[w < nwords ? w : nwords - 1]; }
        const long long e0 = mg_line_block_row(S, line, k_mine) * F;
#pragma unroll
        for (int a = 0; a < F; a++) nr[a] = vec[e0 + a];
    };
    auto stage = [&]() {
        __syncwarp();
#pragma unroll
        for (int j = 0; j < W; j++) sf[j * 32 + lane] = (double)nf[j];
#pragma unroll
        for (int a = 0; a < F; a++) sm.r[lane][a] = nr[a];
        __syncwarp();
    };
    // ---- forward
    double y[F], gp[FF];
#pragma unroll
    for (int a = 0; a < F; a++) y[a] = 0.0;
#pragma unroll
    for (int e = 0; e < FF; e++) gp[e] = 0.0;
    fetch(0, min(lane, m - 1), r);
    for (int c = 0; c < nch; c++) {
        stage();
        if (c + 1 < nch) fetch((c + 1) * 32, min((c + 1) * 32 + lane, m - 1), r);
        const int cnt = min(32, m - c * 32);
        if (lane == 0) {
#pragma unroll 8
            for (int q = 0; q < 32; q++) {
                if (q < cnt) {
                    double yn[F];
#pragma unroll
                    for (int a = 0; a < F; a++) {
                        double s2 = sm.r[q][a];
#pragma unroll
                        for (int e = 0; e < F; e++) s2 -= gp[e * F + a] * y[e];
                        yn[a] = s2;
                    }
#pragma unroll
                    for (int a = 0; a < F; a++) { y[a] = yn[a]; sm.o[q][a] = yn[a]; }
#pragma unroll
                    for (int e = 0; e < FF; e++) gp[e] = sm.f[q][FF + e];
                }
            }
        }
        __syncwarp();
        const int k = c * 32 + lane;
        if (k < m) {
            const long long e0 = mg_line_block_row(S, line, k) * F;
#pragma unroll
            for (int a = 0; a < F; a++) {
                double s2 = 0.0;
#pragma unroll
                for (int e = 0; e < F; e++) s2 += sm.f[lane][a * F + e] * sm.o[lane][e];
                t[e0 + a] = s2;
            }
        }
    }
    __syncwarp();
    // ---- backward: chunk c holds the steps [kb, kb + 32), kb = m - 32 (c + 1) (the last chunk starts below 0: its first -kb slots are idle)
    double x[F];
#pragma unroll
    for (int a = 0; a < F; a++) x[a] = 0.0;
    auto fetch_b = [&](int c) { const int kb = m - 32 * (c + 1); fetch(max(kb, 0), min(max(kb + lane, 0), m - 1), t); };
    fetch_b(0);
    for (int c = 0; c < nch; c++) {
        const int kb = m - 32 * (c + 1), shift = kb < 0 ? -kb : 0;      // slot q of the staged chunk = step max(kb, 0) + q for the factors ...
        stage();                                                        // ... and step kb + lane for the vector (clamped)
        if (c + 1 < nch) fetch_b(c + 1);
        if (lane == 0) {
#pragma unroll 8
            for (int q = 31; q >= 0; q--) {
                const int k = kb + q;
                if (k >= 0) {
                    double xn[F];
#pragma unroll
                    for (int a = 0; a < F; a++) {
                        double s2 = sm.r[q][a];
#pragma unroll
                        for (int e = 0; e < F; e++) s2 -= sm.f[q - shift][FF + a * F + e] * x[e];
                        xn[a] = s2;
                    }
#pragma unroll
                    for (int a = 0; a < F; a++) { x[a] = xn[a]; sm.o[q][a] = xn[a]; }
                }
            }
        }
        __syncwarp();
        const int k = kb + lane;
        if (k >= 0) {
            const long long e0 = mg_line_block_row(S, line, k) * F;
#pragma unroll
            for (int a = 0; a < F; a++) t[e0 + a] = sm.o[lane][a];
        }
    }
}

// t = L^{-1} r on the lines of both directions in one launch of one-warp blocks: blocks [0, nbA) serve set A, the rest set B; each set by one
// thread or one warp per line (MgLineSet::warp_per_line).  One thread per line: only `lanes` lanes of a warp carry a line, so that the few
// thousand lines of a mesh spread over all SMs as independent instruction streams (8192 lines in full 128-thread blocks occupy 64 of the
// 148 SMs).  r, t = owned rows.
PB_HD int mg_line_blocks(const MgLineSet& S, int lanes) { if (S.nlines <= 0 || S.nsteps <= 0) return 0; return S.warp_per_line ? S.nlines : (S.nlines + lanes - 1) / lanes; }
template <int F, typename VT>
__global__ void __launch_bounds__(32) mg_line_solve_kernel(MgLineSet A, const VT* __restrict__ facA, int nbA, MgLineSet B, const VT* __restrict__ facB, int lanes,
                                                           const double* __restrict__ r, double* __restrict__ t, const CgState* st) {
    __shared__ MgLineWarpSmem<F, VT> sm;
    if (st && st->done) return;
    const bool second = (int)blockIdx.x >= nbA;
    const MgLineSet S = second ? B : A;
    const VT* __restrict__ fac = second ? facB : facA;
    const int blk = second ? blockIdx.x - nbA : blockIdx.x;
    if (S.warp_per_line) {
        if (blk < S.nlines) mg_line_solve_warp<F, VT>(S, fac, r, t, blk, sm);
    } else {
        const int line = blk * lanes + threadIdx.x;
        if ((int)threadIdx.x < lanes && line < S.nlines) mg_line_solve_thread<F, VT>(S, fac, r, t, line);
    }
}

// Chebyshev direction update from t = L^{-1} res:  d = (first ? 0 : c1 d) + c2 t;  z = (first && !acc) ? d : z + d
__global__ void mg_cheb_combine_kernel(long long n, int first, int acc, double c1, double c2, const double* __restrict__ t, double* __restrict__ d,
                                       double* __restrict__ z, const CgState* st) {
    if (st && st->done) return;
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < n; e += (long long)gridDim.x * blockDim.x) {
        const double dn = first ? c2 * t[e] : c1 * d[e] + c2 * t[e];
        d[e] = dn;
        z[e] = (first && !acc) ? dn : z[e] + dn;
    }
}
// partial sums of w.w (power iteration with the line solve)
__global__ void mg_norm2_kernel(long long n, const double* __restrict__ w, double* __restrict__ parts) {
    __shared__ double sh[32];
    double s = 0.0;
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < n; e += (long long)gridDim.x * blockDim.x) s += w[e] * w[e];
    const double t = block_reduce_sum(s, sh); if (threadIdx.x == 0) parts[blockIdx.x] = t;
}
#endif  // __CUDACC__

}  // namespace pb
