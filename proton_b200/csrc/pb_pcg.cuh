// K4/K5 — Jacobi-preconditioned conjugate gradient on the condensed face system: the reference's
// conjugated_gradient (solver_cg.hpp:63-144) with the same update order, stopping tests and iteration count,
// as three bandwidth-bound kernels per iteration:
//   dir   : [reduce (r.r, r.z)] -> exit tests -> beta -> d = z + beta d            (iteration > 0)
//   spmv  : y = A d  fused with the partial sums of d.y
//   xrz   : [reduce d.y] -> alpha -> x += alpha d, r -= alpha y, z = M^{-1} r, partial sums of r.r and r.z
// Reductions are two-stage (per-block partials, then every block re-reduces the same partial array in the same
// order), so results are deterministic and no host round trip is needed inside an iteration.
#pragma once
#include "pb_device.cuh"
#include "pb_system.cuh"

namespace pb {

struct CgState {               // lives in device memory
    double rho;                // r . M^{-1} r of the current residual
    double nr0;                // ||r0||
    double relres;             // nr / nr0 at exit
    long long iters;           // completed direction updates at exit (solver_cg.hpp `iter`)
    int done, reason;
};

constexpr int PCG_THREADS = 256;
constexpr int PCG_MAX_BLOCKS = 148 * 8;     // persistent grid: multiple of the SM count

PB_D double block_reduce_sum(double v, double* sh) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) sh[w] = v;
    __syncthreads();
    double t = 0.0;
    if (threadIdx.x < 32) {
        t = (threadIdx.x < (blockDim.x >> 5)) ? sh[threadIdx.x] : 0.0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_down_sync(0xffffffffu, t, o);
    }
    __syncthreads();
    return t;    // valid in thread 0
}
// every block sums the same nparts partials in the same order -> identical result everywhere
PB_D double reduce_partials(const double* parts, int nparts, double* sh) {
    double v = 0.0;
    for (int i = threadIdx.x; i < nparts; i += blockDim.x) v += parts[i];
    double t = block_reduce_sum(v, sh);
    __shared__ double bc;
    if (threadIdx.x == 0) bc = t;
    __syncthreads();
    double r = bc;
    __syncthreads();
    return r;
}

// multi-rank runs: fold the per-block partials into one scalar per quantity so that the rank sums can be all-reduced
__global__ void pcg_sum_partials_kernel(int nparts, const double* a, double* out_a, const double* b, double* out_b) {
    __shared__ double sh[32];
    double va = reduce_partials(a, nparts, sh);
    if (threadIdx.x == 0) *out_a = va;
    if (b) { double vb = reduce_partials(b, nparts, sh); if (threadIdx.x == 0) *out_b = vb; }
}

// iA = 1/diag(A) (solver_cg.hpp:78-81); A is scalar CSR with global column ids, rows [row0, row0+n)
__global__ void pcg_diag_kernel(long long n, long long row0, const long long* __restrict__ rowptr, const int* __restrict__ col,
                                const double* __restrict__ val, double* __restrict__ iA) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        double d = 0.0;
        for (long long k = rowptr[i]; k < rowptr[i + 1]; k++) if (col[k] == row0 + i) d = val[k];
        iA[i] = 1.0 / d;
    }
}

// x = 0, r = b, z = M^{-1} r, d = z; partials of r.r and r.z          (solver_cg.hpp:74-85)
__global__ void pcg_init_kernel(long long n, int precond, const double* __restrict__ b, const double* __restrict__ iA,
                                double* __restrict__ x, double* __restrict__ r, double* __restrict__ z, double* __restrict__ d,
                                double* __restrict__ parts_rr, double* __restrict__ parts_rz) {
    __shared__ double sh[32];
    double rr = 0.0, rz = 0.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        double ri = b[i], zi = precond ? iA[i] * ri : ri;
        x[i] = 0.0; r[i] = ri; z[i] = zi; d[i] = zi;
        rr += ri * ri; rz += ri * zi;
    }
    double t = block_reduce_sum(rr, sh); if (threadIdx.x == 0) parts_rr[blockIdx.x] = t;
    t = block_reduce_sum(rz, sh); if (threadIdx.x == 0) parts_rz[blockIdx.x] = t;
}
__global__ void pcg_init_state_kernel(int nparts, const double* parts_rr, const double* parts_rz, CgState* st, double* hist = nullptr) {
    if (hist && threadIdx.x == 0) hist[0] = 1.0;                 // nr/nr0 printed at the top of iteration 0 (solver_cg.hpp:101-102)
    __shared__ double sh[32];
    double rr = reduce_partials(parts_rr, nparts, sh), rz = reduce_partials(parts_rz, nparts, sh);
    if (threadIdx.x == 0) { st->rho = rz; st->nr0 = sqrt(rr); st->relres = 1.0; st->iters = 0; st->done = 0; st->reason = PB_CG_MAX_ITER_REACHED; }
}

// y = A d (+ partials of d.y).  LPR lanes cooperate on one row (rows have <= 7*F entries), so a warp reads
// 32/LPR consecutive rows = one contiguous run of val/col.  xoff: d is stored for columns [xoff, ...).
template <int LPR>
__global__ void __launch_bounds__(PCG_THREADS, 8) pcg_spmv_kernel(long long n, long long xoff, const long long* __restrict__ rowptr, const int* __restrict__ col,
                                const double* __restrict__ val, const double* __restrict__ dvec, long long down,
                                double* __restrict__ y, double* __restrict__ parts_dy, const CgState* st) {
    __shared__ double sh[32];
    if (st->done) return;
    const int sub = threadIdx.x % LPR;
    const long long rows_per_pass = (long long)gridDim.x * blockDim.x / LPR;
    double dy = 0.0;
    for (long long row = (blockIdx.x * (long long)blockDim.x + threadIdx.x) / LPR; row < ((n + rows_per_pass - 1) / rows_per_pass) * rows_per_pass; row += rows_per_pass) {
        double acc = 0.0;
        if (row < n) {
            long long k0 = rowptr[row], k1 = rowptr[row + 1];
            for (long long k = k0 + sub; k < k1; k += LPR) acc += val[k] * dvec[col[k] - xoff];
        }
#pragma unroll
        for (int o = LPR / 2; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o, LPR);
        if (sub == 0 && row < n) { y[row] = acc; dy += acc * dvec[down + row]; }
    }
    double t = block_reduce_sum(dy, sh); if (threadIdx.x == 0) parts_dy[blockIdx.x] = t;
}

// The same product on the implicit block structure: the F scalar rows of a free face share their column blocks, so the
// column of entry k of a row is bcol[k / F] * F + k % F from the 56-byte BlockRow record of the face (cached, shared by
// F rows) instead of a 4-byte index per entry from HBM: 8 + 64/F^2 bytes per non-zero instead of 12.  The value
// array is the scalar CSR one (rows of a face are contiguous).  CS = 1 streams the values with evict-first loads;
// measured slower (a 504-byte block row is consumed by one lane group over several instructions and needs L1 to merge
// the partially used sectors), so CS = 0 is what runs.
// VT = float: the same product on a single-precision COPY of the values (vectors and accumulation stay double): half the matrix
// bytes.  Only the multigrid preconditioner uses it (pb_api_mg.inl); the A d of the CG recurrence always runs on the double values.
template <int LPR, int F, int CS, typename VT = double>
__global__ void pcg_spmv_blk_kernel(long long n, long long xoff, const BlockRow* __restrict__ rows, const long long* __restrict__ brp,
                                    const VT* __restrict__ val, const double* __restrict__ dvec, long long down,
                                    double* __restrict__ y, double* __restrict__ parts_dy, const CgState* st) {
    // one group of LPR lanes per BLOCK row: a lane gathers the vector entry of its columns once and feeds the F scalar
    // rows of the face with it (F x fewer gathers, F independent value loads in flight per column)
    __shared__ double sh[32];
    if (st->done) return;
    const int sub = threadIdx.x % LPR;
    const long long nbr = n / F;
    const long long groups = (long long)gridDim.x * blockDim.x / LPR;
    const long long npass = (nbr + groups - 1) / groups;
    double dy = 0.0;
    long long br = (blockIdx.x * (long long)blockDim.x + threadIdx.x) / LPR;
    for (long long pass = 0; pass < npass; pass++, br += groups) {
        double acc[F];
#pragma unroll
        for (int a = 0; a < F; a++) acc[a] = 0.0;
        if (br < nbr) {
            const BlockRow* R = rows + br;
            const int len = R->nb * F;
            const VT* v = val + brp[br] * (F * F);
#pragma unroll
            for (int t = 0; t < (7 * F + LPR - 1) / LPR; t++) {
                const int k = sub + t * LPR;
                if (k < len) {
                    const int b = k / F, bc = k - b * F;
                    const double xv = dvec[(long long)R->bcol[b] * F + bc - xoff];
#pragma unroll
                    for (int a = 0; a < F; a++) acc[a] += (double)(CS ? __ldcs(v + a * len + k) : v[a * len + k]) * xv;
                }
            }
        }
#pragma unroll
        for (int a = 0; a < F; a++)
#pragma unroll
            for (int o = LPR / 2; o > 0; o >>= 1) acc[a] += __shfl_down_sync(0xffffffffu, acc[a], o, LPR);
        if (sub == 0 && br < nbr && rows[br].nb) {          // empty block rows (interface problem) belong to pcg_spmv_rows_kernel
#pragma unroll
            for (int a = 0; a < F; a++) { y[br * F + a] = acc[a]; dy += acc[a] * dvec[down + br * F + a]; }
        }
    }
    double t = block_reduce_sum(dy, sh); if (threadIdx.x == 0) parts_dy[blockIdx.x] = t;
}

// scalar rows of a short list of block rows (interface problem: the rows of the cut cells' faces, which the block kernel
// leaves at zero); one thread per scalar row, partial sums of d.y appended after the block kernel's
__global__ void pcg_spmv_rows_kernel(int nbrows, int F, const int* __restrict__ brow_list, const long long* __restrict__ rowptr, const int* __restrict__ col,
                                     const double* __restrict__ val, const double* __restrict__ dvec, long long row0, double* __restrict__ y,
                                     double* __restrict__ parts_dy, const CgState* st) {
    // dvec is indexed by GLOBAL unknown (the caller shifts the halo'd vector back by its first column); the listed rows and y are
    // local to the owned rows, which start at global unknown row0
    __shared__ double sh[32];
    if (st->done) return;
    const int lane = threadIdx.x & 31, nwarps = (gridDim.x * blockDim.x) >> 5;
    double dy = 0.0;
    for (int t = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; t < nbrows * F; t += nwarps) {        // one warp per scalar row
        const long long r = (long long)brow_list[t / F] * F + (t % F);
        double acc = 0.0;
        for (long long k = rowptr[r] + lane; k < rowptr[r + 1]; k += 32) acc += val[k] * dvec[col[k]];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
        if (lane == 0) { y[r] = acc; dy += acc * dvec[row0 + r]; }
    }
    double t = block_reduce_sum(dy, sh); if (threadIdx.x == 0) parts_dy[blockIdx.x] = t;
}

// alpha = rho / (d.y); x += alpha d; r -= alpha y; z = M^{-1} r; partials of r.r, r.z   (solver_cg.hpp:106-110, 127)
__global__ void pcg_xrz_kernel(long long n, int precond, int nparts, const double* __restrict__ parts_dy, const double* __restrict__ iA,
                               const double* __restrict__ d, const double* __restrict__ y, double* __restrict__ x, double* __restrict__ r,
                               double* __restrict__ z, double* __restrict__ parts_rr, double* __restrict__ parts_rz, const CgState* st) {
    __shared__ double sh[32];
    if (st->done) return;
    double dy = reduce_partials(parts_dy, nparts, sh);
    double alpha = st->rho / dy;
    double rr = 0.0, rz = 0.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        double di = d[i];
        x[i] = x[i] + alpha * di;
        double ri = r[i] - alpha * y[i];
        double zi = precond ? iA[i] * ri : ri;
        r[i] = ri; z[i] = zi;
        rr += ri * ri; rz += ri * zi;
    }
    double t = block_reduce_sum(rr, sh); if (threadIdx.x == 0) parts_rr[blockIdx.x] = t;
    t = block_reduce_sum(rz, sh); if (threadIdx.x == 0) parts_rz[blockIdx.x] = t;
}

// ---- optional F x F block-Jacobi preconditioner (apply_preconditioner = 2; NOT reference behaviour: the reference only
// knows the diagonal, solver_cg.hpp:78-81).  iAb[br] = inverse of the diagonal block of face br, row-major.
template <int F>
__global__ void pcg_bdiag_kernel(long long nbr, long long row0, const long long* __restrict__ rowptr, const int* __restrict__ col,
                                 const double* __restrict__ val, double* __restrict__ iAb) {
    for (long long br = blockIdx.x * (long long)blockDim.x + threadIdx.x; br < nbr; br += (long long)gridDim.x * blockDim.x) {
        double A[F][F], B[F][F];
        const int want = (int)(row0 + br * F);
#pragma unroll
        for (int a = 0; a < F; a++) {
            const long long k0 = rowptr[br * F + a], k1 = rowptr[br * F + a + 1];
            long long kd = k0;
            for (long long k = k0; k < k1; k += F) if (col[k] == want) kd = k;      // column blocks are F wide and sorted
#pragma unroll
            for (int c = 0; c < F; c++) { A[a][c] = val[kd + c]; B[a][c] = a == c ? 1.0 : 0.0; }
        }
        // Gauss-Jordan without pivoting (the block is symmetric positive definite)
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
#pragma unroll
        for (int a = 0; a < F; a++)
#pragma unroll
            for (int c = 0; c < F; c++) iAb[br * (F * F) + a * F + c] = 0.5 * (B[a][c] + B[c][a]);    // exactly symmetric
    }
}
// one thread per face: x = 0, r = b, z = iAb r, d = z
template <int F>
__global__ void pcg_init_blk_kernel(long long nbr, const double* __restrict__ b, const double* __restrict__ iAb,
                                    double* __restrict__ x, double* __restrict__ r, double* __restrict__ z, double* __restrict__ d,
                                    double* __restrict__ parts_rr, double* __restrict__ parts_rz) {
    __shared__ double sh[32];
    double rr = 0.0, rz = 0.0;
    for (long long br = blockIdx.x * (long long)blockDim.x + threadIdx.x; br < nbr; br += (long long)gridDim.x * blockDim.x) {
        double ri[F];
#pragma unroll
        for (int a = 0; a < F; a++) ri[a] = b[br * F + a];
#pragma unroll
        for (int a = 0; a < F; a++) {
            double zi = 0.0;
#pragma unroll
            for (int c = 0; c < F; c++) zi += iAb[br * (F * F) + a * F + c] * ri[c];
            x[br * F + a] = 0.0; r[br * F + a] = ri[a]; z[br * F + a] = zi; d[br * F + a] = zi;
            rr += ri[a] * ri[a]; rz += ri[a] * zi;
        }
    }
    double t = block_reduce_sum(rr, sh); if (threadIdx.x == 0) parts_rr[blockIdx.x] = t;
    t = block_reduce_sum(rz, sh); if (threadIdx.x == 0) parts_rz[blockIdx.x] = t;
}
// x += alpha d, r -= alpha y, z = iAb r: a block walks chunks of blockDim * F consecutive unknowns (whole faces) with
// unit-stride loads, stages the new residual in shared memory and applies the F x F blocks from there
template <int F>
__global__ void __launch_bounds__(PCG_THREADS) pcg_xrz_blk_kernel(long long nbr, int nparts, const double* __restrict__ parts_dy, const double* __restrict__ iAb,
                                   const double* __restrict__ d, const double* __restrict__ y, double* __restrict__ x, double* __restrict__ r,
                                   double* __restrict__ z, double* __restrict__ parts_rr, double* __restrict__ parts_rz, const CgState* st) {
    __shared__ double sh[32];
    __shared__ double sr[PCG_THREADS * F];
    if (st->done) return;
    double dy = reduce_partials(parts_dy, nparts, sh);
    double alpha = st->rho / dy;
    double rr = 0.0, rz = 0.0;
    const long long n = nbr * F, chunk = (long long)PCG_THREADS * F;
    for (long long base = blockIdx.x * chunk; base < n; base += (long long)gridDim.x * chunk) {
#pragma unroll
        for (int j = 0; j < F; j++) {
            const int le = j * PCG_THREADS + threadIdx.x; const long long e = base + le;
            if (e < n) { x[e] = x[e] + alpha * d[e]; const double ri = r[e] - alpha * y[e]; r[e] = ri; sr[le] = ri; }
        }
        __syncthreads();
#pragma unroll
        for (int j = 0; j < F; j++) {
            const int le = j * PCG_THREADS + threadIdx.x; const long long e = base + le;
            if (e < n) {
                const int fb = (le / F) * F;
                double zi = 0.0;
#pragma unroll
                for (int c = 0; c < F; c++) zi += iAb[e * F + c] * sr[fb + c];
                z[e] = zi;
                const double ri = sr[le];
                rr += ri * ri; rz += ri * zi;
            }
        }
        __syncthreads();
    }
    double t = block_reduce_sum(rr, sh); if (threadIdx.x == 0) parts_rr[blockIdx.x] = t;
    t = block_reduce_sum(rz, sh); if (threadIdx.x == 0) parts_rz[blockIdx.x] = t;
}

// exit tests in the reference's order (solver_cg.hpp:112-125), then beta = (r.z)/rho, d = z + beta d, iter++.
// st->iters = number of completed direction updates before this call.
__global__ void pcg_dir_kernel(long long n, int nparts, const double* __restrict__ parts_rr, const double* __restrict__ parts_rz,
                               const double* __restrict__ z, double* __restrict__ d, double tol, double divtol,
                               long long max_iter, CgState* st) {
    __shared__ double sh[32];
    if (st->done) return;
    const long long iter = st->iters;       // read before pcg_commit_kernel (the next launch) advances it: the launch
                                            // sequence carries no host-side iteration number and can be replayed as a graph
    double rr = reduce_partials(parts_rr, nparts, sh), rz = reduce_partials(parts_rz, nparts, sh);
    double rel = sqrt(rr) / st->nr0;
    int reason = -1;
    if (rel < tol) reason = PB_CG_CONVERGED;
    else if (iter > max_iter) reason = PB_CG_MAX_ITER_REACHED;
    else if (rel > divtol) reason = PB_CG_DIVERGED;
    double rho = st->rho;
    // all blocks must have read st before block 0 may overwrite it: the update is deferred to pcg_commit_kernel
    if (reason >= 0) return;
    double beta = rz / rho;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        d[i] = z[i] + beta * d[i];
}
// single-block bookkeeping after pcg_dir_kernel: same decisions, written once
__global__ void pcg_commit_kernel(int nparts, const double* parts_rr, const double* parts_rz, double tol, double divtol,
                                  long long max_iter, CgState* st, double* hist = nullptr, long long hist_cap = 0) {
    __shared__ double sh[32];
    if (st->done) return;
    const long long iter = st->iters;
    double rr = reduce_partials(parts_rr, nparts, sh), rz = reduce_partials(parts_rz, nparts, sh);
    if (threadIdx.x != 0) return;
    double rel = sqrt(rr) / st->nr0;
    int reason = -1;
    if (rel < tol) reason = PB_CG_CONVERGED;
    else if (iter > max_iter) reason = PB_CG_MAX_ITER_REACHED;
    else if (rel > divtol) reason = PB_CG_DIVERGED;
    st->relres = rel;
    if (hist && iter + 1 < hist_cap) hist[iter + 1] = rel;       // the histfile line of iteration iter+1 (or the closing line)
    if (reason >= 0) { st->done = 1; st->reason = reason; st->iters = iter; }
    else { st->rho = rz; st->iters = iter + 1; }
}

}  // namespace pb
