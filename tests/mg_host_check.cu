// TEST INFRASTRUCTURE ONLY — runs the __host__ __device__ index logic of proton_b200/csrc/pb_mg.cuh (the multigrid
// preconditioner's transfer operators, Galerkin products and stencil assembly) on host arrays, so that tests/test_mg_cpu.py can
// check it against scipy's P^T A P without a GPU.  The library itself has no host path; nothing here is linked into it.
#include "../proton_b200/csrc/pb_mg.cuh"
#include "../proton_b200/csrc/pb_mg_lines.cuh"
using namespace pb;

static MgGrid whole(int Nx, int Ny) {
    MgGrid L{}; L.Nx = Nx; L.Ny = Ny; L.cx = Nx > 2; L.cy = Ny > 2; L.row0 = 0; L.nrows = Ny + 1; L.own0 = 0; L.own1 = Ny + 1; L.crow0 = 0; L.ncrows = Ny;
    return L;
}
static MgGrid strip(int Nx, int Ny, int rb, int re) {      // level 0 of a rank owning cell rows [rb, re)
    MgGrid L = whole(Nx, Ny);
    L.row0 = rb > 0 ? rb - 1 : 0; L.nrows = (re + 1 <= Ny ? re + 1 : Ny + 1) - L.row0; L.own0 = rb; L.own1 = re;
    L.crow0 = rb > 0 ? rb - 1 : 0; L.ncrows = re - L.crow0;
    return L;
}
static Grid grid_of(int Nx, int Ny) { Grid g{}; g.Nx = Nx; g.Ny = Ny; g.row_begin = 0; g.row_end = Ny; g.hx = 1.0 / Nx; g.hy = 1.0 / Ny; return g; }

extern "C" {
// element matrices of all cells from their condensed blocks S[cell][NF*NF] (column-major)
void mgh_elements(int Nx, int Ny, int F, const double* S, const int* node_loc, const int* face_loc, const double* face_ip, const double* points, double* E) {
    Grid g = grid_of(Nx, Ny);
    MgCut C{}; if (node_loc) { C.node_loc = node_loc; C.face_loc = face_loc; C.face_ip = (const double2*)face_ip; C.points = (const double2*)points; }
    const int NF = 4 * F;
    for (long long c = 0; c < (long long)Nx * Ny; c++) mg_element_from_S(g, C, F, c % Nx, c / Nx, S + c * NF * NF, E + c * 16);
}
// stencils [node][9] of the vertex rows [rb, re) of a strip level (E = element matrices of cell rows [crow0, ...))
void mgh_stencils(int Nx, int Ny, int rb, int re, const double* E, double* st) {
    MgGrid L = strip(Nx, Ny, rb, re);
    for (int j = rb; j < re; j++) for (int i = 0; i <= Nx; i++) {
        double s[9];
        if (mg_boundary(L, i, j)) { for (int k = 0; k < 9; k++) s[k] = 0; s[4] = 1; } else mg_node_stencil(L, E, i, j, s);
        for (int k = 0; k < 9; k++) st[((long long)(j - rb) * (Nx + 1) + i) * 9 + k] = s[k];
    }
}
void mgh_coarse_dims(int Nx, int Ny, int* out) { out[0] = mg_coarse_cells(Nx); out[1] = mg_coarse_cells(Ny); }
// Galerkin element matrices of the next level; fine cell rows [flo, fhi) only (partial sums of a strip)
void mgh_galerkin(int Nx, int Ny, const double* Ef, int flo, int fhi, double* Ec) {
    MgGrid Lf = whole(Nx, Ny), Lc = whole(mg_coarse_cells(Nx), mg_coarse_cells(Ny));
    for (int J = 0; J < Lc.Ny; J++) for (int I = 0; I < Lc.Nx; I++) mg_galerkin_cell(Lf, Lc, Ef, I, J, flo, fhi, Ec + ((long long)J * Lc.Nx + I) * 16);
}
// rc = R rf with the fine rows [own0, own1) only; whole-grid vectors
void mgh_restrict(int Nx, int Ny, int own0, int own1, const double* rf, double* rc) {
    MgGrid Lf = whole(Nx, Ny), Lc = whole(mg_coarse_cells(Nx), mg_coarse_cells(Ny)); Lf.own0 = own0; Lf.own1 = own1;
    for (int J = 0; J <= Lc.Ny; J++) for (int I = 0; I <= Lc.Nx; I++) rc[mg_node(Lc, I, J)] = mg_boundary(Lc, I, J) ? 0.0 : mg_restrict_node(Lf, Lc, rf, I, J);
}
void mgh_prolong(int Nx, int Ny, const double* xc, double* xf) {
    MgGrid Lf = whole(Nx, Ny), Lc = whole(mg_coarse_cells(Nx), mg_coarse_cells(Ny));
    for (int j = 0; j <= Ny; j++) for (int i = 0; i <= Nx; i++) xf[mg_node(Lf, i, j)] = mg_boundary(Lf, i, j) ? 0.0 : mg_prolong_node(Lf, Lc, xc, i, j);
}
// face level: b0 = P^T r (r = all free-face unknowns, compressed numbering), z = P x0
void mgh_face_restrict(int Nx, int Ny, int F, const double* r, const int* node_loc, const int* face_loc, const double* face_ip, const double* points, double* b0) {
    Grid g = grid_of(Nx, Ny); MgGrid L0 = whole(Nx, Ny);
    MgCut C{}; if (node_loc) { C.node_loc = node_loc; C.face_loc = face_loc; C.face_ip = (const double2*)face_ip; C.points = (const double2*)points; }
    for (int j = 0; j <= Ny; j++) for (int i = 0; i <= Nx; i++) b0[mg_node(L0, i, j)] = mg_boundary(L0, i, j) ? 0.0 : mg_face_restrict_node(g, C, F, 0, r, i, j);
}
void mgh_face_prolong(int Nx, int Ny, int F, const double* x0, const int* node_loc, const int* face_loc, const double* face_ip, const double* points, double* z) {
    Grid g = grid_of(Nx, Ny); MgGrid L0 = whole(Nx, Ny);
    MgCut C{}; if (node_loc) { C.node_loc = node_loc; C.face_loc = face_loc; C.face_ip = (const double2*)face_ip; C.points = (const double2*)points; }
    const long long nfree = num_free_faces(g);
    for (long long cf = 0; cf < nfree; cf++) {
        double o[2]; mg_face_prolong(g, C, F, L0, x0, cf, o);
        for (int a = 0; a < F; a++) z[cf * F + a] = a < 2 ? o[a] : 0.0;
    }
}
// line sets of the strip of cell rows [rb, re) (pb_mg_lines.cuh): owner[local block row] = kind * 2^40 + line * 2^20 + step for every (line, step),
// hits[local block row] counts how many (line, step) pairs map to it; dims = { nlines_h, nsteps_h, nlines_v, nsteps_v, warp_per_line_h, warp_per_line_v }
void mgh_line_sets(int Nx, int Ny, int rb, int re, long long* owner, int* hits, int* dims) {
    Grid g = grid_of(Nx, Ny);
    const long long cf_begin = rb == 0 ? 0 : compress_row_start(g, rb);
    MgLineSet H, V; mg_line_sets(g, rb, re, cf_begin, H, V);
    dims[0] = H.nlines; dims[1] = H.nsteps; dims[2] = V.nlines; dims[3] = V.nsteps; dims[4] = H.warp_per_line; dims[5] = V.warp_per_line;
    for (int kind = 0; kind < 2; kind++) {
        const MgLineSet& S = kind ? V : H;
        for (int line = 0; line < S.nlines; line++) for (int step = 0; step < S.nsteps; step++) {
            const long long br = mg_line_block_row(S, line, step);
            owner[br] = ((long long)kind << 40) + ((long long)line << 20) + step; hits[br]++;
        }
    }
}
}
