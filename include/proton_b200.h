/* proton_b200 — C ABI of the B200-native HHO / cutHHO hot path.
 *
 * The reference (OmarDuran/ProtoN) is a header-only C++ template library with no FFI layer: its drivers call
 * free functions per cell from a serial loop.  This ABI is what a binding for that path would bind instead:
 * the same stages, batched over all cells of a Cartesian mesh strip and executed by hand-written sm_100a
 * kernels.  Each entry point cites the reference interface it replaces (paths relative to the reference
 * root).  All pointers are caller-owned HOST arrays unless a name ends in `_dev`.  Every call returns a
 * pb_status; pb_last_error() gives the message.  A context is bound to one CUDA device and is
 * thread-compatible (one thread at a time).  There is no CPU fallback: without a CUDA device
 * pb_ctx_create fails with PB_ERR_CUDA.
 */
#ifndef PROTON_B200_H
#define PROTON_B200_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct pb_ctx pb_ctx;

typedef enum { PB_OK = 0, PB_ERR_INVALID = 1, PB_ERR_CUDA = 2, PB_ERR_STATE = 3, PB_ERR_UNSUPPORTED = 4 } pb_status;

/* which integrate(msh, cl, deg) overload the mesh type selects (src/core/core_bits/quadratures.hpp:311-375
 * for quad_mesh, :377-402 for poly_mesh / cuthho_poly_mesh) */
typedef enum { PB_QUAD_TENSOR_GAUSS = 0, PB_QUAD_POLY_DUNAVANT = 1 } pb_quad_kind;
/* make_hho_naive_stabilization (src/methods/hho_bits/hho.hpp:99-148), make_hho_fancy_stabilization
 * (:155-237), make_hho_cut_stabilization (apps/cuthho/cuthho_square.cpp:566-621) */
typedef enum { PB_STAB_NAIVE = 0, PB_STAB_FANCY = 1, PB_STAB_CUT = 2 } pb_stab_kind;
/* plain HHO Poisson (apps/convergence_test, apps/obstacle) or fictitious-domain cutHHO
 * (run_cuthho_fictdom, apps/cuthho/cuthho_square.cpp:806-1080) */
typedef enum { PB_PROBLEM_HHO = 0, PB_PROBLEM_FICTDOM = 1, PB_PROBLEM_INTERFACE = 2 } pb_problem_kind;
/* the point-wise host lambdas of the drivers, as device functors: rhs_fun / sol_fun / sol_grad of
 * convergence_test.cpp:100-113 and cuthho_square.cpp:847-866 (SINSIN), obstacle.cpp:65-89 (OBSTACLE), the
 * commented-out tanh problem of convergence_test.cpp:115-138 (TANH) */
typedef enum { PB_FUN_SINSIN = 0, PB_FUN_OBSTACLE = 1, PB_FUN_TANH = 2 } pb_fun_kind;
/* element_location, src/methods/cuthho_bits/cuthho_mesh.hpp:31-36 */
typedef enum { PB_IN_NEGATIVE_SIDE = 0, PB_IN_POSITIVE_SIDE = 1, PB_ON_INTERFACE = 2, PB_UNDEF = 3 } pb_location;
/* cg_exit_reason, src/core/core_bits/solver_cg.hpp:37-42 */
typedef enum { PB_CG_CONVERGED = 0, PB_CG_DIVERGED = 1, PB_CG_MAX_ITER_REACHED = 2 } pb_cg_exit_reason;

/* hho_degree_info(cd, fd) + the per-driver choices (src/core/core_bits/utils.hpp:62-111) */
typedef struct {
    int cell_degree, face_degree;      /* invalid pairs revert to equal order like the reference */
    int stab_kind;                     /* pb_stab_kind */
    int problem;                       /* pb_problem_kind */
    int fun;                           /* pb_fun_kind: source term, Dirichlet / Nitsche data, exact solution */
    int rhs_degree_increase;           /* `di` of make_rhs (utils.hpp:153-174); obstacle.cpp:102 uses 1 */
    double eta;                        /* Nitsche penalty, cell_eta() = 5 (cuthho_square.cpp:301-306) */
    int keep_local;                    /* 1: also keep oper / (data+stab) / rhs per cell for pb_get_local; 2: additionally gr.second
                                        * (`data`) apart from the stabilisation (selector 7) — the per-cell views of the C++ shim;
                                        * this routes the uncut cells through the literal, untuned kernel (pb_generic.cuh) */
} pb_assembly_params;

/* cg_params (solver_cg.hpp:44-61).  apply_preconditioner: 0 = none, 1 = the reference's diagonal preconditioner
 * (solver_cg.hpp:78-81); extensions with no reference counterpart (same solution to the tolerance, fewer iterations):
 * 2 = fbs x fbs block-Jacobi (inverse of every face's diagonal block), 3 = multigrid V-cycle (Chebyshev/block-Jacobi
 * smoothing on the face system, Galerkin vertex-based coarse levels; plain and fictitious-domain problems, face degree
 * <= 3, whole meshes and strips; iteration counts independent of the mesh size on the plain problem). */
typedef struct { double convergence_threshold, divergence_threshold; long long max_iter; int apply_preconditioner; } pb_cg_params;

typedef struct { long long iterations; double relative_residual; int exit_reason; double seconds; } pb_cg_result;

/* ---- context ---- */
pb_status pb_ctx_create(int device, pb_ctx** ctx);
pb_status pb_ctx_destroy(pb_ctx* ctx);
const char* pb_last_error(const pb_ctx* ctx);      /* ctx may be NULL: message of the last failed create */
int pb_abi_version(void);

/* ---- mesh: quad_mesh<T>(mesh_init_params) / cuthho_poly_mesh<T>(mip), src/core/core_bits/basic_mesh.hpp:178-197,
 * 230-298, 321-403.  The Cartesian generator's sorted cell / face arrays are closed forms (cell (i,j) = j*Nx+i;
 * faces ordered by (ptid0, ptid1)), so only the point coordinates live in memory.  [row_begin,row_end) is the
 * strip of cell rows this context owns (multi-GPU: one context per rank); 0,Ny = the whole mesh. ---- */
pb_status pb_mesh_cartesian(pb_ctx* ctx, long long Nx, long long Ny, double min_x, double max_x, double min_y, double max_y,
                            int quad_kind, long long row_begin, long long row_end);
/* all (Nx+1)*(Ny+1) points, xy interleaved (msh.points); set = after move_nodes (cuthho_geom.hpp:466-543) */
pb_status pb_mesh_get_points(pb_ctx* ctx, double* xy);
pb_status pb_mesh_set_points(pb_ctx* ctx, const double* xy);
/* offset(msh, fc) of the 4 faces of every cell, in faces(msh, cl) order (basic_geom.hpp:183-212), and the
 * assembler's compress_table (hho.hpp:313-323; -1 for Dirichlet faces) — for binding-side checks */
pb_status pb_mesh_cell_faces(pb_ctx* ctx, long long* face_ids4);
pb_status pb_mesh_compress_table(pb_ctx* ctx, long long* table);

/* ---- cutHHO: level set and mesh preprocessing.
 * pb_cut_levelset: circle_level_set / line_level_set (apps/cuthho/cuthho_square.cpp:56-124): kind 0 = circle of
 * `radius` centred (cx, cy), kind 1 = horizontal line y = cut_y.
 * pb_cut_preprocess: detect_node_position, detect_cut_faces, move_nodes, detect_cut_faces, detect_cut_cells,
 * refine_interface(refine_steps) in the order of cuthho_square.cpp:2030-2052 (src/methods/cuthho_bits/
 * cuthho_geom.hpp:118-161, 275-340, 466-543, 609-673), on the device.  Moves mesh points.  refine_steps <= 6 (the reference's -r is
 * uncapped and its table has r = 7 rows: PB_ERR_UNSUPPORTED here; pb_cut_upload has the same limit of 65 points per cell).  The reference throws
 * for >2 cuts per cell, concave displaced cells or a lost interface: PB_ERR_INVALID with the same message.
 * pb_cut_upload: the same user data computed by the caller (all arrays for the WHOLE mesh; iface_off has
 * ncells+1 entries, every cut cell must carry the same number of interface points).
 * pb_cut_download: read the device-side user data back (any pointer may be NULL); iface_xy receives, cell by
 * cell in ascending order, the interface polylines; *n_iface_pts the number of points per cut cell.
 * A strip context (one rank of a multi-GPU job) classifies and keeps ONLY the rows it needs — its own cell rows, the ghost row below,
 * their faces and vertices — so preprocessing costs each rank its share: pb_mesh_get_points returns displaced points for those rows,
 * pb_cut_download fills the rest of the whole-mesh arrays with PB_UNDEF / zeros, *n_cut_cells and iface_xy cover the cut cells of
 * those rows.  pb_cut_counts: cut cells held (ghost row included) and cut cells of the OWNED rows (sum it over the ranks). ---- */
pb_status pb_cut_levelset(pb_ctx* ctx, int kind, double radius, double cx, double cy, double cut_y);
pb_status pb_cut_preprocess(pb_ctx* ctx, int refine_steps);
pb_status pb_cut_upload(pb_ctx* ctx, const int* node_location, const int* face_location, const double* face_intersection_xy,
                        const int* cell_location, const long long* iface_off, const double* iface_xy);
pb_status pb_cut_download(pb_ctx* ctx, int* node_location, int* face_location, double* face_intersection_xy, int* cell_location,
                          long long* n_cut_cells, int* n_iface_pts, double* iface_xy);
pb_status pb_cut_counts(pb_ctx* ctx, long long* n_cut_cells_held, long long* n_cut_cells_owned);

/* ---- stages 1+2: the drivers' per-cell loop (make_hho_laplacian + make_hho_*_stabilization + make_rhs, then
 * static condensation) and assembler.assemble()/finalize() (hho.hpp:344-406, 451-455) onto the condensed face
 * system: row/column of face dof (f,a) = compress_table[f]*fbs + a. ---- */
pb_status pb_assemble(pb_ctx* ctx, const pb_assembly_params* params);
/* The same step for a caller that streams meshes through the device (the driver loop of convergence_test.cpp:186-199
 * repeated over a mesh sequence; no counterpart in the serial reference).  Returns after enqueueing:
 *   host_points (page-locked, the whole-mesh array of 2*num_points doubles, may be NULL; a strip context reads only
 *   pb_strip_point_range of it) is uploaded on a copy stream as soon as the previous
 *   step's local-operator kernel has released the vertex array;
 *   host_rhs (page-locked, n_rows doubles, may be NULL) receives assembler.RHS on a second copy stream after the scatter;
 *   the next step's scatter waits for that read-back.  The caller must not reuse host_rhs before pb_pipeline_sync or
 *   two further steps, and calls pb_pipeline_sync before reading results (the other entry points do it themselves). */
pb_status pb_assemble_pipelined(pb_ctx* ctx, const pb_assembly_params* params, const double* host_points, double* host_rhs);
pb_status pb_pipeline_sync(pb_ctx* ctx);
/* the vertices a strip context reads (those of its own cell rows and the ghost row): points [first, first+count) of the
 * whole-mesh array; pb_assemble_pipelined uploads exactly this range from host_points */
pb_status pb_strip_point_range(pb_ctx* ctx, long long* first, long long* count);
pb_status pb_system_sizes(pb_ctx* ctx, long long* n_rows, long long* nnz, long long* n_cells, long long* n_local);
/* assembler.LHS / assembler.RHS of the condensed system, CSR with sorted columns */
pb_status pb_get_csr(pb_ctx* ctx, long long* rowptr, int* col, double* val, double* rhs);
/* per-cell dense results, column-major like Eigen, for cells [cell_begin, cell_begin+count):
 * what = 0: lc = gr.second + stab (n x n); 1: oper = gr.first ((rbs-1) x n, or rbs x n rows padded layout, see
 * DESIGN.md); 2: cell rhs (cbs); 3: condensed block S (4fbs x 4fbs); 4: condensed rhs g (4fbs); 5: K_TT^-1 K_TF (cbs x 4fbs);
 * 6: K_TT^-1 f_T (cbs); 7: data = gr.second (n x n).  0-2 need keep_local, 7 needs keep_local = 2. */
pb_status pb_get_local(pb_ctx* ctx, int what, long long cell_begin, long long count, double* out);

/* ---- stage 3: conjugated_gradient(A, b, x, cg_params) (solver_cg.hpp:63-144) on the condensed system ---- */
pb_status pb_pcg(pb_ctx* ctx, const pb_cg_params* params, pb_cg_result* result);
pb_status pb_get_face_solution(pb_ctx* ctx, double* x);
pb_status pb_set_face_solution(pb_ctx* ctx, const double* x);
/* back-substitution: the reference's FULL solution vector (cell dofs c*cbs+i, then face dofs
 * cbs*ncells + compress[f]*fbs + a), hho.hpp:362-379 */
pb_status pb_recover(pb_ctx* ctx, double* full_solution);
/* assembler.take_local_data(msh, cl, sol, bf) (hho.hpp:408-449) for the owned cells [cell_begin, cell_begin + count) (0 = the first
 * owned cell): n_local = cbs + 4 fbs values per cell, the cell dofs, then the dofs of faces(msh, cl) with the Dirichlet faces
 * replaced by the projection of the datum */
pb_status pb_get_local_solution(pb_ctx* ctx, long long cell_begin, long long count, double* local_dofs);
/* error norms of the drivers: out[0..2] = sums of squares of convergence_test.cpp:254-306 (L2, mass-norm, energy)
 * for PB_PROBLEM_HHO; out[0] = H1 error^2 of cuthho_square.cpp:1031-1050 for PB_PROBLEM_FICTDOM.  A strip context
 * returns the sums over its owned cells (add them over the ranks). */
pb_status pb_error_norms(pb_ctx* ctx, double out[3]);

/* ---- obstacle problem (apps/obstacle/obstacle.cpp:47-218): HHO(0, face_degree), face_degree in {0, 1}, full HHO
 * stabilisation, load with quadrature increase `rhs_degree_increase` (the driver uses 1), data of obstacle.cpp:65-89 on
 * the mesh the caller created (the driver uses [-1,1]^2).  The reference rebuilds make_obstacle_assembler(msh, in_A, hdi)
 * + assemble(msh, cl, lc, f, gamma, bcs) (src/methods/hho_bits/hho.hpp:471-695) and solves the saddle system with
 * SparseLU in every outer iteration, then expand_solution (:698-744).  Here the cell rows are eliminated in closed form
 * (cell degree 0) and each outer iteration is condense -> scatter -> PCG -> update on cached local matrices.
 * pb_obstacle_iterate performs ONE outer iteration (in_A from the current alpha, beta; solve; new alpha, beta) and
 * returns ||alpha_prev - alpha|| (the driver's stopping quantity, obstacle.cpp:193), the size of the active set and
 * the inner CG iterations.  pb_obstacle_error: the driver's final error^2 (obstacle.cpp:201-218). ---- */
pb_status pb_obstacle_begin(pb_ctx* ctx, int face_degree, int rhs_degree_increase);
pb_status pb_obstacle_iterate(pb_ctx* ctx, const pb_cg_params* cg, double* delta_alpha_norm, long long* n_active, long long* cg_iterations);
pb_status pb_obstacle_error(pb_ctx* ctx, double* error_squared);
pb_status pb_obstacle_get(pb_ctx* ctx, double* alpha_cells, double* beta, unsigned char* in_A);

/* ---- multi-GPU: one context per rank, 1-D strip partition by cell rows.  The reference is a single
 * serial process; these entry points have no counterpart there.  Stages 1+2 need no communication (every rank also
 * assembles the ghost row of cells below its strip).  The solve exchanges one halo of (2Nx-1)*fbs doubles with each
 * neighbour per SpMV and all-reduces 1 + 2 scalars per iteration over NCCL (loaded with dlopen at pb_dist_init, so
 * the library has no link-time NCCL dependency).
 * pb_strip_layout is pure host arithmetic (works without a GPU): out = { first owned compressed face, one past the
 * last, first / one-past-last compressed face referenced by the owned rows, first local cell row, owned scalar
 * rows, lower halo doubles, upper halo doubles }. ---- */
pb_status pb_strip_layout(long long Nx, long long Ny, long long row_begin, long long row_end, int fbs, long long out[8]);
pb_status pb_dist_unique_id(unsigned char id[128]);
pb_status pb_dist_init(pb_ctx* ctx, int rank, int world, const unsigned char id[128]);
/* back-substitution on a strip: cell dofs of the OWNED cell rows only ([row_begin,row_end) x Nx x cbs) and the owned
 * face dofs.  Interface problem: the cell part is interface_assembler's packed layout of the owned cells, (owned cells + owned cut
 * cells, pb_cut_counts) x cbs with a cut cell's negative set followed by its positive one; the ranks' cell parts, then their face parts,
 * concatenated in rank order are the vector of pb_recover */
pb_status pb_recover_strip(pb_ctx* ctx, double* cell_dofs, double* face_dofs);

/* ---- two-material interface problem, `cuthho_square -i` (PB_PROBLEM_INTERFACE).  pb_assemble replaces the
 * loop of run_cuthho_interface (apps/cuthho/cuthho_square.cpp:1667-1716): make_hho_laplacian + make_hho_naive_stabilization
 * on uncut cells, make_hho_laplacian_interface (:390-502) + make_hho_cut_stabilization on both sides (:1689-1706) + make_rhs
 * on both sides on cut cells, then interface_assembler::assemble / assemble_cut / finalize (:1197-1340, 1430-1434) onto the
 * condensed face system in interface_assembler's numbering: unknown (f, a) = face_table[f]*fbs + a, the second set of a cut
 * face fbs further (:1161-1175, 1301).  HHO(k+1,k), k <= 2, kappa_1 = kappa_2 = 1 (params<T>, :293-299).
 * Strips: face_table is monotone in the face id, so a strip of cell rows owns the contiguous slots of its faces and
 * pb_assemble / pb_get_csr / pb_pcg / pb_get_face_solution / pb_error_norms (partial sum over the owned cells) work on strip
 * contexts as for the other problems; columns are global unknowns of the interface system.  The one global quantity, the number of
 * cut faces owned by the strips below, is all-reduced inside pb_assemble by the ranks of a communicator (pb_dist_init); strip
 * contexts WITHOUT one (several strips driven by one process) are told: pb_interface_cut_faces_owned after the cut preprocessing,
 * the host's prefix sum, pb_interface_set_cut_offsets(before, total) before pb_assemble.  pb_interface_strip_layout: out =
 * { first owned slot, one past the last, first / one-past-last referenced slot, slots of the whole system, block rows the lower /
 * the upper neighbour reads of this strip }.  pb_recover_strip returns the strip's part of the reference's vector, pb_sample_solution the rows
 * of interface_uT.dat of the owned cells (the ranks of a communicator call these together: they fetch the halo of the solution); pb_recover and
 * pb_interface_get_local stay whole-mesh calls;
 * pb_interface_face_table reports the faces outside the strip's window as -2.
 * pb_get_csr / pb_pcg / pb_get_face_solution work as for the other problems.  pb_recover returns interface_assembler's solution
 * layout: (ncells + ncut) cell blocks of cbs in cell_table order (a cut cell carries its negative, then its positive set,
 * :1143-1160), then the face unknowns; pb_error_norms out[0] = the driver's H1 error^2 over both parts (:1775-1833);
 * pb_sample_solution the rows of interface_uT.dat.  pb_interface_get_local: per-cut-cell arrays in ascending cell order (what = 0: condensed block,
 * (8 fbs)^2 column-major in the driver's local order [faces-, faces+]; 1: condensed rhs; with keep_local 2: the local matrix
 * (2 cbs + 8 fbs)^2 in the order [cell-, cell+, faces-, faces+]; 3: the local load, 2 cbs).  pb_interface_face_table:
 * face_table per face, -1 for Dirichlet faces. ---- */
pb_status pb_interface_get_local(pb_ctx* ctx, int what, double* out);
pb_status pb_interface_face_table(pb_ctx* ctx, long long* table);
pb_status pb_interface_cut_faces_owned(pb_ctx* ctx, long long* count);
pb_status pb_interface_set_cut_offsets(pb_ctx* ctx, long long cut_faces_before, long long cut_faces_total);
pb_status pb_interface_strip_layout(pb_ctx* ctx, long long out[7]);

/* ---- driver outputs.
 * pb_pcg_history(capacity > 0): the following pb_pcg calls record nr/nr0 exactly where conjugated_gradient writes a line
 * to cg_params::histfile (solver_cg.hpp:92-93, 101-102, 133-137): entry 0 = 1, entry i = relative residual after i
 * updates, the closing line included, iterations + 2 entries in all; capacity 0 switches it off.
 * pb_sample_solution: the rows "x y value" of the driver's solution_N_<N>_k_<k>.dat (convergence_test.cpp:250-252,
 * 280-286) for the owned cells: cell solution at the degree-2*cell_degree quadrature points, cells ascending, points in
 * integrate()'s order; xyv = 3 doubles per point (NULL: only *n_points is set).  On the interface problem: the rows of
 * interface_uT.dat (cuthho_square.cpp:1775-1833): cut cells contribute their negative, then their positive part. ---- */
pb_status pb_pcg_history(pb_ctx* ctx, long long capacity);
pb_status pb_pcg_get_history(pb_ctx* ctx, double* out, long long capacity, long long* n);
pb_status pb_sample_solution(pb_ctx* ctx, double* xyv, long long capacity_points, long long* n_points);

/* ---- multigrid preconditioner (apply_preconditioner = 3), inspection for tests and bench; no reference counterpart.
 * pb_mg_info builds the hierarchy for the current system if needed: number of vertex levels, the upper end of the smoothing
 * range (1.1 x the power-iteration estimate of lambda_max(D^-1 A)), milliseconds this call spent building (0 when cached).
 * pb_mg_level: dims = { Nx, Ny, first stored vertex row, stored vertex rows, first stored cell row, stored cell rows };
 * stencil9 (may be NULL): 9 x stored vertices, entry k = (dj+1)*3 + di+1 of vertex v at [k * nvertices + v];
 * elements16 (may be NULL): 4 x 4 Galerkin element matrix per stored cell, row-major, corners (i,j) (i+1,j) (i+1,j+1) (i,j+1).
 * pb_mg_apply: z = M^-1 r for host vectors of the owned rows (all ranks of a communicator call it together). ---- */
pb_status pb_mg_info(pb_ctx* ctx, int* n_levels, double* lmax, double* setup_ms);
pb_status pb_mg_level(pb_ctx* ctx, int level, long long dims[6], double* stencil9, double* elements16);
pb_status pb_mg_apply(pb_ctx* ctx, const double* r, double* z);

/* ---- allocation-free variant of pb_error_norms.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same kernels,
 * same launch and summation order, same out[3] as pb_error_norms, but the scratch (padded face vector, recovered cell dofs,
 * per-block partial sums) lives in one file-scope device array of pb_static_scratch_bytes() = 16 MiB that every GPU loading
 * the library holds until exit; the call performs no cudaMalloc / cudaFree and no allocating library call.  Limits: whole-mesh
 * contexts on one GPU, plain and fictitious-domain problems, and 8*(face unknowns + cells*cbs) + 57 KiB <= 16 MiB (e.g. k=2 up
 * to about 260 x 260 cells); beyond that PB_ERR_UNSUPPORTED with the sizes in pb_last_error.  The first call on a device
 * resolves the array's device address (cudaGetSymbolAddress): make it outside a graph capture.  All contexts of a device share
 * the array and it keeps the last call's contents (the part that must start at zero is cleared by every call): calls that
 * overlap on the GPU — other contexts / streams, other host threads — race on it; serialise them. ---- */
pb_status pb_error_norms_static(pb_ctx* ctx, double out[3]);
long long pb_static_scratch_bytes(void);

/* ---- measurement hooks (bench.py): device-side timing of the last pb_assemble / pb_pcg, milliseconds ---- */
pb_status pb_last_timings(pb_ctx* ctx, double* ms_local, double* ms_scatter, double* ms_pcg);
/* register-resident DFMA peak probe (FP64 roofline denominator), TFLOP/s */
pb_status pb_probe_fp64_peak(pb_ctx* ctx, double* tflops);

#ifdef __cplusplus
}
#endif
#endif
