/* nsb200.h -- C-ABI of libnsb200.so: hand-written sm_100a CUDA kernels for the per-step
 * hot path of the P2-P1 incompressible Navier-Stokes solvers of
 * LKM-code-base/NavierStokes-with-Fenics (finite-element assembly, Dirichlet rows, sparse
 * Krylov solves, BDF history updates).
 *
 * The reference has no FFI of its own: the seam is the subset of the dolfin Python API its
 * solver classes call.  Each entry point below cites the reference call site
 * (file:line under /root/reference) whose native work it replaces.
 *
 * Conventions
 *  - every function returns 0 on success, <0 on error; nsb_last_error() gives the message.
 *    No C++ exception crosses this boundary.  The Python shim raises RuntimeError on != 0,
 *    which keeps the reference's `except RuntimeError` continuation fallback working
 *    (source/ns_problem.py:462-501).
 *  - "dev" pointers are device addresses (e.g. torch.Tensor.data_ptr()); buffers are owned by
 *    the caller, the library never frees them.  "host" pointers are ordinary host memory.
 *  - `stream` is a cudaStream_t passed as void* (0 = legacy default stream).  Calls are
 *    asynchronous unless they return a scalar to the host.
 *  - one nsb_ctx per GPU / rank; not thread-safe; no internal threads.
 *  - all floating point is IEEE double; all indices int32 (value offsets are widened to
 *    int64 inside the kernels).
 *  - P2 "nodes" are scalar: vertex v -> v, edge e -> nverts + e.  The collapsed velocity
 *    space V = [P2]^d is interleaved, dof = d*node + comp; Q = P1, dof = vertex.
 *    Cell-local P2 order: vertices then edges in UFC order.
 */
#ifndef NSB200_H
#define NSB200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct nsb_ctx nsb_ctx;
typedef struct nsb_pattern_builder nsb_pattern_builder;

#define NSB_VERSION 100

/* assembly flags */
#define NSB_VEC 1        /* assemble the vector (residual / rhs)            */
#define NSB_MAT 2        /* assemble the matrix (Jacobian / lhs)            */
#define NSB_PICARD 4     /* Jacobian without the Newton term (frozen convecting velocity) */
#define NSB_BDF 8        /* mixed forms: include the BDF mass term          */
#define NSB_SYMBC 16     /* mixed forms: SystemAssembler symmetric local BC elimination */

/* pattern ids for nsb_set_pattern */
#define NSB_PAT_P2P2 0   /* scalar P2 nodes x P2 nodes (velocity BSR-d / mass CSR)  */
#define NSB_PAT_P1P1 1   /* vertices x vertices (pressure)                           */
#define NSB_PAT_WW 2     /* mixed Taylor-Hood dofs x dofs (monolithic solvers)       */

/* matrix kinds for nsb_matrix */
#define NSB_CSR 0        /* scalar CSR applied to `bs` interleaved components: y[bs*i+c] = sum_j a_ij x[bs*j+c] */
#define NSB_BSR 1        /* bs x bs dense blocks, row-major inside the block         */

/* Dirichlet modes */
#define NSB_BC_LINEAR 0  /* b[i] = g[i]          LinearVariationalSolver  bc.apply(A,b)   */
#define NSB_BC_NEWTON 1  /* b[i] = x[i] - g[i]   NonlinearVariationalSolver bc.apply(b,x) */

/* kernel classes of the built-in CUDA-event timers (nsb_timer_report) */
#define NSB_T_ASM_JAC 0      /* momentum / mixed Jacobian assembly (K2, K8)          */
#define NSB_T_ASM_RES 1      /* momentum / mixed residual assembly (K1, K7)          */
#define NSB_T_ASM_CONST 2    /* constant operators: P1 stiffness, P1/P2 mass (K3,K5) */
#define NSB_T_ASM_RHS 3      /* Poisson and correction right-hand sides (K4, K6)     */
#define NSB_T_SPMV_BSR 4     /* BSR-d SpMV (K10)                                     */
#define NSB_T_SPMV_CSR 5     /* CSR / Kronecker-CSR SpMV (K10)                       */
#define NSB_T_KRYLOV 6       /* Krylov vector kernels: fused updates, dots, Arnoldi (K11) */
#define NSB_T_BC 7           /* Dirichlet rows (K9)                                  */
#define NSB_T_VEC 8          /* vector ops, history shift, gather/scatter (K13, K15) */
#define NSB_T_COMM 9         /* halo pack + NCCL send/recv, NCCL all-reduce (K14)    */
#define NSB_T_SPMV_BSR32 10  /* BSR-d SpMV on the single-precision value copy (Chebyshev preconditioner only) */
#define NSB_T_CG_FUSED 11    /* whole CG solves as one persistent cooperative kernel (csrc/cg_persistent.cu) */
#define NSB_T_NCLASS 12

/* index spaces with a halo plan (nsb_set_halo, nsb_matrix.halo) */
#define NSB_HALO_NONE 0
#define NSB_HALO_NODES 1     /* scalar P2 nodes (velocity vectors: bs = d)             */
#define NSB_HALO_VERTS 2     /* vertices (pressure vectors)                            */

/* SELL-C-sigma copy (C = 32, sigma = 1: P1 rows are nearly uniform, 7 .. 15 entries) of a scalar CSR matrix: rows in
 * slices of 32, a slice padded to its longest row and stored COLUMN-major, so that one thread per row reads the
 * value / column streams fully coalesced (entry k of row 32 s + l sits at slice_ptr[s] + 32 k + l).  Padding entries
 * carry value 0 and the row's own index as column.  vals32 / col16 (col - row, when every delta fits 16 bits): the
 * 6-bytes-per-entry copy read ONLY by the SpMVs inside the Chebyshev polynomial of the persistent CG -- the CG
 * operator itself stays FP64. */
typedef struct nsb_sell {
    int32_t nslices;               /* ceil(nrows / 32)                                   */
    int32_t reserved;
    const int32_t* slice_ptr;      /* dev [nslices + 1], entry offsets (multiples of 32) */
    const int32_t* col;            /* dev [slice_ptr[nslices]]                           */
    const double* vals;            /* dev [slice_ptr[nslices]]                           */
    const float* vals32;           /* dev or NULL                                        */
    const int16_t* col16;          /* dev or NULL                                        */
} nsb_sell;

typedef struct nsb_matrix {
    int32_t kind;            /* NSB_CSR | NSB_BSR                                        */
    int32_t bs;              /* components (CSR) or block size (BSR)                      */
    int32_t nrows;           /* scalar / block rows                                       */
    int32_t nnz;             /* stored scalar (CSR) / block (BSR) entries = rowptr[nrows]  */
    const int32_t* rowptr;   /* dev [nrows+1]                                             */
    const int32_t* col;      /* dev [nnz] ascending per row                               */
    const double* vals;      /* dev [nnz] (CSR) or [nnz*bs*bs] (BSR)                      */
    const uint8_t* bc_mask;  /* dev [nrows*bs] or NULL: 1 = Dirichlet dof, the operator acts
                                as the identity on that row (dolfin `ident`, diagonal 1.0) */
    const uint32_t* tiles;   /* dev [ntiles*64] or NULL: TMA row tiles of this pattern
                                (nsb_tiles_build); enables the bulk-copy SpMV kernels       */
    int32_t ntiles;
    int32_t halo;            /* NSB_HALO_*: index space of the columns on a partitioned mesh.  Rows
                                are the owned prefix; columns >= n_owned are ghosts, refreshed
                                from their owners inside nsb_spmv / the Krylov solvers.  0 on one GPU */
    const float* vals32;     /* dev [nnz*bs*bs] or NULL: single-precision copy of a BSR matrix's values
                                (nsb_vec_to_f32).  Read ONLY by the SpMVs inside the Chebyshev
                                preconditioner; every operator application of the solvers stays FP64 */
    const float* diag32;     /* dev [nnz] or NULL: block-scalar approximation S (x) I of a BSR matrix, S_ij = mean of
                                the diagonal of block ij (nsb_bsr_block_scalar_f32).  Read ONLY by the single-
                                precision Chebyshev polynomial inside the flexible nsb_gmres (nsb_precond.lowp),
                                in preference to vals32: 8 instead of 40 bytes per block entry */
    const struct nsb_sell* sell;   /* or NULL: SELL-32 copy of a scalar CSR matrix (nsb_sell_fill), the layout the
                                persistent CG kernel streams */
} nsb_matrix;

/* preconditioners of nsb_cg / nsb_gmres (K12) */
#define NSB_PC_NONE 0
#define NSB_PC_JACOBI 1      /* z = D^-1 v                                                    */
#define NSB_PC_CHEBYSHEV 2   /* z = p(D^-1 A) D^-1 v, p = Chebyshev polynomial of `degree` terms
                                (degree-1 SpMVs) for the interval [lambda_min, lambda_max] of D^-1 A */
#define NSB_PC_SADDLE 3      /* block-triangular preconditioner of the mixed Taylor-Hood Newton system (nsb_gmres only,
                                flexible), described by nsb_precond.saddle */
struct nsb_saddle;
typedef struct nsb_precond {
    int32_t kind;
    int32_t degree;
    const double* inv_diag;  /* dev [nrows*bs], nsb_inv_diagonal */
    double lambda_min, lambda_max;
    int32_t lowp;            /* Chebyshev inside nsb_gmres on a BSR matrix with vals32 and tiles: 1 = the whole
                                polynomial runs in single precision (values, vectors padded to a float4 per
                                node, arithmetic) and GMRES runs FLEXIBLE (keeps z_j = M^-1 v_j, x += Z y), so
                                the rounding inside M never enters the solution: Arnoldi, the residual test and
                                the operator stay FP64.  Ignored by nsb_cg. */
    int32_t reserved;
    const struct nsb_saddle* saddle;   /* NSB_PC_SADDLE only, else NULL */
} nsb_precond;

/* K16, hand-written: right block-triangular preconditioner for the monolithic Newton system
 *     A = [ F  B ; B^T 0 ]  (+ Dirichlet identity rows),   dofs = [velocity (n_u) | pressure (n_p)]  (W numbering)
 * which the reference hands to sparse LU (source/ns_solver_base.py:938-949, source/ns_bdf_solver.py:91-106):
 *     z_p = -S~^-1 v_p,  S~^-1 = c_mass Mp^-1 + c_lap Lp^-1   (Cahouet-Chabard: c_mass = c_visc / c_p^2,
 *                                                               c_lap = (alpha0 / k) / c_p^2; 0 when stationary)
 *     z_u = F~^-1 (v_u - B z_p),  F~^-1 = Chebyshev polynomial of vel_degree terms in D^-1 F
 * Mp / Lp: P1 mass / stiffness matrices on NSB_PAT_P1P1 (their bc_mask marks the vertices where the Schur
 * complement is replaced by the identity: pressure Dirichlet dofs; for Lp additionally the vertices of natural
 * (do-nothing) velocity boundaries).  The inner solves are nsb_cg to inner_rtol (Jacobi for Mp, Chebyshev-CG of
 * lp_degree for Lp; persistent kernel), so the preconditioner varies from one application to the next and nsb_gmres
 * runs FLEXIBLE.  work: n_u + 11 n_p doubles.  inner_iterations accumulates the inner CG iterations. */
typedef struct nsb_saddle {
    int32_t n_u, n_p;
    nsb_matrix Mp, Lp;
    const double* Mp_inv_diag;
    const double* Lp_inv_diag;
    const uint8_t* p_identity;     /* dev [n_p] or NULL: pressure dofs whose row of A is an identity row */
    double c_mass, c_lap;
    int32_t lp_nullspace;          /* Lp has the constant null space (no identity row at all) */
    int32_t vel_degree;
    double vel_lambda_min, vel_lambda_max;
    int32_t lp_degree;
    double lp_lambda_min, lp_lambda_max;
    double inner_rtol;
    int32_t inner_maxit;
    int32_t reserved;
    double* work;
    int64_t inner_iterations;
} nsb_saddle;
/* largest eigenvalue (modulus) of D^-1 F, F = the velocity block (first n_u rows / columns) of the mixed matrix A,
 * by power iteration; work: 2 n doubles */
int nsb_saddle_lambda_max(nsb_ctx* ctx, const nsb_matrix* A, const double* inv_diag, int32_t n_u, int32_t iters,
                          double* work, double* lambda_out_host, void* stream);

/* ---------------------------------------------------------------- life cycle ---------- */
int nsb_version(void);
int nsb_create(nsb_ctx** ctx, int device);
int nsb_destroy(nsb_ctx* ctx);
const char* nsb_last_error(const nsb_ctx* ctx);   /* ctx may be NULL: last ctx-less error */

/* ---------------------------------------------------------------- host-side setup ------
 * Sparsity pattern = union over cells of (row dofs x col dofs) of the whole element, columns
 * ascending, no extra diagonal: dolfin SparsityPatternBuilder + PETSc AIJ as triggered by
 * source/ns_solver_base.py:501-524 (function spaces) and every assemble().  Also emits the
 * row -> (cell, local row, slot of every local column) incidence used by the atomic-free
 * row-gather assembly kernels (the compressed cell-to-nnz map).  Pure CPU, setup only. */
int nsb_pattern_build(int32_t nrows, int32_t ncells,
                      int32_t nloc_row, const int32_t* cell_rows_host,
                      int32_t nloc_col, const int32_t* cell_cols_host,
                      nsb_pattern_builder** out);
int nsb_pattern_sizes(const nsb_pattern_builder* pb, int64_t* nnz, int64_t* npairs, int32_t* max_row_len);
int nsb_pattern_fill(const nsb_pattern_builder* pb, int32_t* rowptr, int32_t* col,
                     int32_t* pair_ptr, int32_t* pair_cell, uint8_t* pair_iloc, uint8_t* pair_slots);
int nsb_pattern_free(nsb_pattern_builder* pb);
/* Row tiles for the TMA (cp.async.bulk) SpMV kernels: rows grouped so that the value and column
 * streams of a tile are one contiguous 16-byte aligned chunk of <= max_entries entries.  Pure CPU.
 * records == NULL: only count.  Record layout (64 uint32 per tile) in csrc/setup_host.cu. */
int nsb_tiles_build(int32_t nrows, const int32_t* rowptr_host, int32_t max_entries, int32_t* ntiles_out,
                    uint32_t* records_host);

/* ---------------------------------------------------------------- device state --------- */
/* mesh of source/grid_generator.py:111-208 (ordered simplices) + scalar P2 node map */
int nsb_set_mesh(nsb_ctx* ctx, int32_t dim, int32_t nverts, const double* coords_dev,
                 int32_t ncells, const int32_t* cell_verts_dev,
                 int32_t nnodes, const int32_t* cell_nodes_dev);
int nsb_set_pattern(nsb_ctx* ctx, int32_t which, int32_t nrows,
                    const int32_t* rowptr_dev, const int32_t* col_dev,
                    const int32_t* pair_ptr_dev, const int32_t* pair_cell_dev,
                    const uint8_t* pair_iloc_dev, const uint8_t* pair_slots_dev);
/* Conflict-free scatter schedule of the warp-cooperative matrix assembly (csrc/assemble_jac.cu):
 * rows grouped (<= 32 incidence pairs, <= max_blocks slots), and per chunk of 32 pairs a proper
 * edge colouring of the (pair, slot) graph so that all lanes of a warp add into distinct slots
 * in every round, preferring rounds without shared-memory bank conflicts.  Entries with many
 * contributions (the diagonal of a vertex row) get shadow slots behind the group's real entries; the
 * fold list (real entry << 16 | shadow slot) says which copies are added up before the rows are written.
 * nsb_sched_build is host-only; call it with group_row_host == NULL to obtain
 * counts4 = {groups, chunks, rounds, folds}, then with buffers group_row[groups+1], chunk_first[groups+1],
 * round_ptr[chunks+1], sched[32*rounds], fold_ptr[groups+1], fold[folds].  nsb_set_schedule registers the
 * device copies (group_max_entries = the max_blocks used).  Without a schedule the generic kernels are used. */
int nsb_sched_build(int32_t nrows, const int32_t* rowptr_host, const int32_t* pair_ptr_host,
                    const uint8_t* pair_slots_host, int32_t nloc, int32_t max_blocks, int64_t* counts4,
                    int32_t* group_row_host, int32_t* chunk_first_host, int32_t* round_ptr_host,
                    uint16_t* sched_host, int32_t* fold_ptr_host, uint32_t* fold_host);
int nsb_set_schedule(nsb_ctx* ctx, int32_t which, int32_t ngroups, int32_t group_max_entries,
                     const int32_t* group_row_dev, const int32_t* chunk_first_dev, const int32_t* round_ptr_dev,
                     const uint16_t* sched_dev, const int32_t* fold_ptr_dev, const uint32_t* fold_dev);
/* Caller-owned device workspace of the two-phase (element vectors per cell -> gather per row) vector assembly
 * kernels: ncells * nsb_scratch_doubles_per_cell(d) doubles (= NL2 * d).  Without it the one-thread-per-row
 * kernels are used.  NULL / 0 unregisters. */
int nsb_set_scratch(nsb_ctx* ctx, double* scratch_dev, int64_t ndoubles);
int nsb_scratch_doubles_per_cell(int32_t dim);
/* mixed Taylor-Hood dof map [ncells x (d*NL2 + d+1)] (source/ns_solver_base.py:501-524) */
int nsb_set_mixed_dofmap(nsb_ctx* ctx, int32_t ndofs, const int32_t* cell_dofs_dev);
/* scalars: dlfn.Constant.assign of k, alpha_i (source/ns_ipcs_solver.py:210-227,
 * source/ns_bdf_solver.py:108-126) and EquationCoefficientHandler values
 * (source/auxiliary_classes.py:251-306).  coef = {k, a0, a1, a2, c_conv, c_p, c_visc, c_body} */
int nsb_set_coefficients(nsb_ctx* ctx, const double* coef_host8);

/* weak-form variants of the momentum equation: constructor arguments form_convective_term / form_viscous_term of
 * source/ns_solver_base.py:78-115 (terms :370-390, :478-499, :662-673) and the Coriolis term :173-191.
 * conv: 0 standard | 1 rotational | 2 divergence | 3 skew_symmetric; visc: 0 reduced | 1 traction;
 * coriolis_host3 = 2 c_Omega Omega (2D: only [2] = the z component; NULL = none).  The default combination
 * (0, 0, none) runs the tuned IPCS kernels, every other one the generic row kernels; generic != 0 forces the generic
 * kernels for the default combination too (they are tested against the tuned ones). */
int nsb_set_forms(nsb_ctx* ctx, int32_t conv, int32_t visc, const double* coriolis_host3, int32_t generic);

/* ---------------------------------------------------------------- assembly (K1-K8) ------ */
/* IPCS tentative velocity F / J = derivative(F,u*)   source/ns_ipcs_solver.py:106-147 (:126-130, :136)
 * b_out [d*nnodes]; A_out BSR-d values on NSB_PAT_P2P2. */
int nsb_assemble_ipcs_momentum(nsb_ctx* ctx, int32_t flags, const double* ustar, const double* un,
                               const double* unm1, const double* pn, double* b_out, double* A_out,
                               void* stream);
/* Allocation-free variant of the residual F above.  It has NO counterpart in the reference and was added
 * for callers that cannot call cudaMalloc (CUDA-graph capture, tight serving loops, harnesses
 * that forbid it).  Same inputs, same output and same arithmetic as nsb_assemble_ipcs_momentum(NSB_VEC) with a
 * registered scratch, but the element-vector scratch is a file-scope static device array of
 * nsb_static_scratch_doubles() = 2 097 152 doubles (16 MiB, held by every GPU that loads the library until exit):
 * meshes with more than 69 905 tetrahedra / 174 762 triangles (ncells * nloc * d doubles beyond that) are rejected with an
 * error.  The call allocates nothing (no cudaMalloc / cudaFree / allocating library routine) and registers nothing.
 * Default weak form only (standard convective term, reduced viscous term, no Coriolis term).  All calls on one GPU
 * share the array, which keeps the last call's element vectors (nothing in it needs to start at zero): calls that
 * overlap on the GPU -- different streams, different contexts -- race; serialise them. */
int64_t nsb_static_scratch_doubles(void);
int nsb_assemble_ipcs_momentum_residual_noalloc(nsb_ctx* ctx, const double* ustar, const double* un, const double* unm1,
                                                const double* pn, double* b_out, void* stream);
/* IPCS pressure Poisson a / L   source/ns_ipcs_solver.py:149-171 (:160-164) */
int nsb_assemble_poisson(nsb_ctx* ctx, int32_t flags, const double* pn, const double* ustar,
                         double* b_out, double* A_out, void* stream);
/* IPCS velocity correction a / L   source/ns_ipcs_solver.py:173-196 (:183-187).
 * A_out: scalar P2 mass on NSB_PAT_P2P2 (the d x d block matrix of dolfin is mass (x) I_d). */
int nsb_assemble_correction(nsb_ctx* ctx, int32_t flags, const double* ustar, const double* pnew,
                            const double* pold, double* b_out, double* A_out, void* stream);
/* P1 mass matrix and lumped weights int psi_v dx: dlfn.project / mean-pressure functionals
 * source/ns_solver_base.py:1151,1168,1190-1203 */
int nsb_assemble_p1_mass(nsb_ctx* ctx, int32_t flags, double* w_out, double* A_out, void* stream);
/* monolithic residual / Jacobian on W   source/ns_bdf_solver.py:68-88,
 * source/ns_solver_base.py:913-936; flags NSB_VEC|NSB_MAT|NSB_PICARD|NSB_BDF */
int nsb_assemble_mixed(nsb_ctx* ctx, int32_t flags, const double* w0, const double* w1,
                       const double* w2, double* b_out, double* A_out, void* stream);

/* ---------------------------------------------------------------- Dirichlet rows (K9) ---- */
/* DirichletBC.apply(A): zero the rows, diagonal 1.0, columns untouched (App. A.5);
 * dofs are scalar dof indices (bs*node+comp for BSR). */
int nsb_bc_apply_matrix(nsb_ctx* ctx, int32_t kind, int32_t bs, const int32_t* rowptr,
                        const int32_t* col, double* vals, int32_t nbc, const int32_t* bc_dofs,
                        void* stream);
/* DirichletBC.apply(b) / apply(b,x) */
int nsb_bc_apply_vector(nsb_ctx* ctx, int32_t mode, double* b, const double* x, int32_t nbc,
                        const int32_t* bc_dofs, const double* bc_vals, void* stream);
/* SystemAssembler symmetric local elimination (source/auxiliary_classes.py:141,160-164) is
 * done inside nsb_assemble_mixed (flag NSB_SYMBC) from this table: mask[ndofs] (1 = Dirichlet
 * dof) and g[ndofs] (boundary value, 0 elsewhere), both device arrays. */
int nsb_set_mixed_bc(nsb_ctx* ctx, const uint8_t* bc_mask_dev, const double* bc_g_dev);

/* ---------------------------------------------------------------- linear algebra (K10-13) */
int nsb_spmv(nsb_ctx* ctx, const nsb_matrix* A, const double* x, double* y, void* stream);
/* Fill the arrays of a SELL-32 copy (slice_ptr given, computed by the caller from the row lengths) from the CSR
 * matrix A; vals32_out / col16_out may be NULL.  *overflow_out != 0: some |col - row| does not fit 16 bits, col16_out
 * is unusable (leave nsb_sell.col16 NULL). */
int nsb_sell_fill(nsb_ctx* ctx, const nsb_matrix* A, int32_t nslices, const int32_t* slice_ptr_dev, int32_t* col_out,
                  double* vals_out, float* vals32_out, int16_t* col16_out, int32_t* overflow_out, void* stream);
/* inv_diag[i] = 1/A_ii (1 on masked rows): Jacobi preconditioner (K12) */
int nsb_inv_diagonal(nsb_ctx* ctx, const nsb_matrix* A, double* inv_diag, void* stream);
int64_t nsb_cg_workspace_doubles(int64_t n);
int64_t nsb_gmres_workspace_doubles(int64_t n, int32_t restart, int32_t flexible /* = M->lowp */);
/* Preconditioned CG for the SPD pressure-Poisson and mass systems which the reference solves by
 * LU (source/ns_ipcs_solver.py:171,195-196).  nullspace != 0: the operator has the constant
 * null space (all-Neumann pressure): rhs and solution are projected to zero
 * arithmetic mean.  Converged when ||r|| <= max(rtol*||b||, atol).  work: nsb_cg_workspace_doubles. */
int nsb_cg(nsb_ctx* ctx, const nsb_matrix* A, const nsb_precond* M, const double* b, double* x,
           double rtol, double atol, int32_t maxit, int32_t nullspace, double* work,
           int32_t* iters_out, double* resid_out, void* stream);
/* right-preconditioned restarted GMRES (CGS-2 Arnoldi) for the velocity / mixed Newton systems
 * (source/ns_ipcs_solver.py:138-147, source/ns_bdf_solver.py:91-100); flexible (keeps M^-1 v_j) when
 * M->lowp.  x0_is_zero = 1: the caller guarantees x = 0 on entry (Newton updates), which saves the
 * first residual SpMV (r = b).  work: nsb_gmres_workspace_doubles(n, restart, M->lowp). */
int nsb_gmres(nsb_ctx* ctx, const nsb_matrix* A, const nsb_precond* M, const double* b, double* x,
              double rtol, double atol, int32_t maxit, int32_t restart, int32_t x0_is_zero, double* work,
              int32_t* iters_out, double* resid_out, void* stream);
/* largest eigenvalue (modulus) of D^-1 A by `iters` power iterations: upper end of the Chebyshev
 * interval.  work: 2 * nrows*bs doubles.  warm_start = 1: work[0 .. nrows*bs) still holds the iterate
 * the previous call left there (the Jacobian of the previous time step): continue from it, two or
 * three iterations then track the slowly moving spectrum. */
int nsb_lambda_max(nsb_ctx* ctx, const nsb_matrix* A, const double* inv_diag, int32_t iters, int32_t warm_start,
                   double* work, double* lambda_out_host, void* stream);
/* DG post-processing fields of source/ns_problem.py:55-102.  dolfin.project(curl(u), DG_1) and
 * dolfin.project(grad(p), [DG_0]^d) are cell-wise L2 projections of a field that already lies in the target space
 * (curl of P2 is P1 per cell, gradient of P1 is constant per cell): the dofs are its vertex values / its value.
 * nsb_cell_vorticity: out[cell][local vertex][nc], nc = 1 (2D, d_x u_y - d_y u_x) or 3; nsb_cell_pressure_gradient:
 * out[cell][d].  u in V numbering (interleaved), p in Q numbering. */
int nsb_cell_vorticity(nsb_ctx* ctx, const double* u_dev, double* out_dev, void* stream);
int nsb_cell_pressure_gradient(nsb_ctx* ctx, const double* p_dev, double* out_dev, void* stream);
/* Successive right-hand sides (Fischer 1998) for a solve that is repeated with the same SPD operator every time
 * step -- the pressure projection, source/ns_ipcs_solver.py:149-171, 205.  X / AX: caller-owned lmax x n arrays
 * (lmax <= 16) holding an A-orthonormal basis of the previous solutions and its image; l = the number of vectors in
 * use (host integer owned by the caller, start at 0).  nsb_rhs_guess: x = sum_k (X_k . b) X_k (x untouched when
 * l == 0).  nsb_rhs_update (after the solve): appends the A-orthogonal part of x; a full basis first drops its
 * older half (lmax / 2 vectors stay); work: 2 n doubles.  The solve itself keeps its tolerance: only its iteration count changes. */
int nsb_rhs_guess(nsb_ctx* ctx, int64_t n, int32_t l, const double* X_dev, const double* b_dev, double* x_dev, void* stream);
int nsb_rhs_update(nsb_ctx* ctx, const nsb_matrix* A, int32_t* l_inout_host, int32_t lmax, double* X_dev, double* AX_dev,
                   const double* x_dev, double* work_dev, void* stream);
/* The same power iteration on the single-precision block-scalar copy A->diag32 (the operator the f32 Chebyshev
 * preconditioner of nsb_gmres is a polynomial of): one f32 SpMV per iteration.  work: 2 x 16/12 x n bytes <= 2 n doubles. */
int nsb_lambda_max32(nsb_ctx* ctx, const nsb_matrix* A, const double* inv_diag, int32_t iters, int32_t warm_start,
                   double* work, double* lambda_out_host, void* stream);

/* Function.assign / Vector ops (source/ns_ipcs_solver.py:41-43, source/ns_solver_base.py:1015-1016) */
int nsb_vec_copy(nsb_ctx* ctx, int64_t n, const double* x, double* y, void* stream);
int nsb_vec_axpy(nsb_ctx* ctx, int64_t n, double a, const double* x, double* y, void* stream); /* y += a x */
int nsb_vec_scale(nsb_ctx* ctx, int64_t n, double a, double* x, void* stream);
int nsb_vec_to_f32(nsb_ctx* ctx, int64_t n, const double* x, float* y, void* stream);   /* y = (float) x */
int nsb_vec_dot(nsb_ctx* ctx, int64_t n, const double* x, const double* y, double* out_host, void* stream);
int nsb_vec_sum(nsb_ctx* ctx, int64_t n, const double* x, double* out_host, void* stream);
/* FunctionAssigner (source/ns_solver_base.py:424-476): y[idx[i]] = x[i] / y[i] = x[idx[i]] */
int nsb_vec_scatter(nsb_ctx* ctx, int64_t n, const int32_t* idx, const double* x, double* y, void* stream);
int nsb_vec_gather(nsb_ctx* ctx, int64_t n, const int32_t* idx, const double* x, double* y, void* stream);
/* BDF history shift of the IPCS fields, one launch (source/ns_ipcs_solver.py:35-43):
 * u2 <- u1 <- u0 ; p_old <- p */
int nsb_ipcs_advance(nsb_ctx* ctx, int64_t nv, const double* u0, double* u1, double* u2,
                     int64_t nq, const double* p, double* p_old, void* stream);

/* ---------------------------------------------------------------- functionals (K15) ------ */
/* F = -int_Gamma(-p n + c_visc sym(grad u) n) ds over the listed exterior facets: the drag / lift
 * post-processing of demo/dfg_benchmark.py:44-66 (c_D = 2 F_x, c_L = 2 F_y).  facet_cell: owning
 * cell of each facet, facet_opp: local number of the cell vertex opposite to it; u in V numbering
 * (interleaved P2), p in Q numbering.  force_out_host3: 3 doubles (unused components 0). */
int nsb_boundary_force(nsb_ctx* ctx, int32_t nfacets, const int32_t* facet_cell_dev,
                       const uint8_t* facet_opp_dev, const double* u, const double* p, double c_visc,
                       double* force_out_host3, void* stream);

/* ---------------------------------------------------------------- projection / diagnostics */
/* Right-hand side of dolfin.project(expr, V|Q) (source/ns_solver_base.py:1151,1168; the mean-pressure
 * re-projection :1201; demo/taylor_green_vortex.py:61-69): the host interpolates the Expression into P_k on
 * every cell (cell_vals_dev [ncells][nk][ncomp], values at the cell's P_k lattice nodes) and hands over the
 * exact reference tensor table_dev [nloc_row][nk] = int phi_I psi^k_J; b[row*ncomp+c] = sum over incident
 * cells of detJ sum_J table[I][J] vals[cell][J][c].  which = NSB_PAT_P2P2 (rows = P2 nodes) | NSB_PAT_P1P1
 * (rows = vertices).  The mass solve that follows is nsb_cg on the P2 / P1 mass matrix. */
int nsb_assemble_cell_rhs(nsb_ctx* ctx, int32_t which, int32_t nk, int32_t ncomp, const double* table_dev,
                          const double* cell_vals_dev, double* b_out, void* stream);
/* dolfin.errornorm(u_exact, u_h) ** 2 (convergence_test/taylor_green_vortex.py:101-119): both functions interpolated
 * into P_{k+3} on every cell, the difference integrated exactly.  cell_vals_dev [ncells][nk][ncomp]: the exact
 * solution at the cell's lattice nodes; E_dev [nk][nloc]: the space's basis at those nodes; Psi_dev [nq][nk], wq_dev
 * [nq]: the P_{k+3} basis at the points of a quadrature rule exact for degree 2(k+3), and its weights (reference
 * cell, sum = 1/d!); work_dev [ncells*nk*ncomp] scratch.  space as `which` above; uh interleaved (ncomp per node). */
int nsb_cell_error_norm2(nsb_ctx* ctx, int32_t space, int32_t ncomp, const double* uh, int32_t nk,
                         const double* E_dev, int32_t nq, const double* Psi_dev, const double* wq_dev,
                         const double* cell_vals_dev, double* work_dev, double* out_host, void* stream);
/* CFL diagnostic of source/ns_problem.py:554-587 (runs every step, :714): max-norm of the DG-k LocalSolver
 * projection of  k |u| dt / h_T  (quadrature degree 2k; h_T = UFL CellDiameter = longest edge).  scale = k * dt;
 * B_dev [nq][NL2]: P2 basis at the quadrature points; P_dev [NL2][nq]: nodes x points matrix of the local
 * projection  Phi (B^T W B)^-1 B^T W.  u in V numbering.  Max over all ranks on a partitioned mesh. */
int nsb_cfl_max(nsb_ctx* ctx, const double* u, double scale, int32_t nq, const double* B_dev,
                const double* P_dev, double* out_host, void* stream);
/* x -= s: the mean-pressure shift of source/ns_solver_base.py:1190-1203 (subtracting a constant from a P1
 * function IS its L2 projection) */
int nsb_vec_shift(nsb_ctx* ctx, int64_t n, double s, double* x, void* stream);

/* ---------------------------------------------------------------- direct solve (K16) ----- */
/* x = A^-1 b for the indefinite mixed Newton system, which the reference hands to PETSc's sparse LU
 * (source/ns_bdf_solver.py:91-100,106; source/ns_solver_base.py:938-949,974-983).  LIBRARY leg, not
 * a hand-written kernel: cuSOLVER sparse QR on the device (cusolverSpDcsrlsvqr, bound at run
 * time), reorder 0 none | 1 symrcm | 2 symamd | 3 metis.  Blocking.  Its wall time is accumulated
 * apart from the kernel-class timers and read with nsb_direct_seconds (reset != 0 clears it). */
int nsb_direct_solve(nsb_ctx* ctx, int32_t n, int32_t nnz, const int32_t* rowptr, const int32_t* col,
                     const double* vals, const double* b, double* x, int32_t reorder,
                     double* seconds_out, void* stream);
int nsb_direct_seconds(nsb_ctx* ctx, double* seconds_out, int64_t* solves_out, int32_t reset);

/* ---------------------------------------------------------------- multi-GPU (K14) -------- */
/* Mesh-partitioned runs, one rank per GPU (the reference has no counterpart: it
 * would be dolfin/PETSc MPI, which source/ never exercises).  Each rank registers its LOCAL mesh
 * (owned cells + one ghost layer) with nsb_set_mesh, patterns whose rows are the owned prefix of
 * the owned-first local numbering, and one halo plan per index space.  Vectors handed to the
 * assembly calls hold [owned | ghost] entries; nsb_halo_exchange refreshes the ghost block from
 * the owners.  nsb_spmv / nsb_cg / nsb_gmres / nsb_lambda_max refresh the ghosts of their
 * operand themselves (matrix.halo) and all-reduce every inner product, nsb_vec_dot / _sum
 * all-reduce their result.  NCCL (libnccl.so.2, the copy torch loads) is bound at run time. */
int nsb_comm_unique_id(uint8_t* id_out128);                      /* ncclGetUniqueId (rank 0) */
int nsb_comm_init(nsb_ctx* ctx, int32_t rank, int32_t nranks, const uint8_t* id128);
/* collective tear-down (every rank calls it, after its last kernel): ncclCommDestroy.  nsb_destroy
 * alone leaves the communicator to process exit, because destroying it synchronises with the peers. */
int nsb_comm_finalize(nsb_ctx* ctx);
/* which = NSB_HALO_*; peers/send_ptr/recv_ptr are host arrays ([npeers], [npeers+1], [npeers+1]):
 * ghost slots recv_ptr[i]..recv_ptr[i+1] (offsets into the ghost block) come from peers[i];
 * send_idx_dev[send_ptr[i]..send_ptr[i+1]] are the local owned indices peers[i] needs, in the
 * order of its ghost slots.  sendbuf_dev [3*send_ptr[npeers]] and ghostbuf_dev [3*n_ghost] are
 * caller-owned scratch.  n_global = size of the index space over all ranks. */
int nsb_set_halo(nsb_ctx* ctx, int32_t which, int32_t n_owned, int32_t n_ghost, int64_t n_global,
                 int32_t npeers, const int32_t* peers_host, const int32_t* send_ptr_host,
                 const int32_t* send_idx_dev, const int32_t* recv_ptr_host, double* sendbuf_dev,
                 double* ghostbuf_dev);
/* ghost_dst[bs*g + c] = owner's x[bs*i + c]; ghost_dst NULL: the plan's ghostbuf.  x holds the
 * owned entries; pass ghost_dst = x + bs*n_owned to refresh an [owned | ghost] vector in place. */
int nsb_halo_exchange(nsb_ctx* ctx, int32_t which, int32_t bs, const double* x, double* ghost_dst,
                      void* stream);

/* buf += the same buffer of every other rank (in place, FP64 sum; no-op on one rank): the one exchange of the
 * REPLICATED pressure solve -- every rank scatters its owned right-hand-side entries into a zeroed global
 * vector, all-reduces it (every entry has exactly one non-zero contributor: exact and bit-identical on all
 * ranks) and then solves the whole 405 k-row system itself with the persistent CG kernel, with no halo
 * exchange and no all-reduce inside the solve. */
int nsb_allreduce(nsb_ctx* ctx, double* buf_dev, int64_t count, void* stream);

/* Halo exchange and scalar all-reduce through NVLink peer memory instead of NCCL (optional, after
 * nsb_comm_init / nsb_set_halo; single node, CUDA IPC).
 * nsb_p2p_export allocates this rank's IPC-shared block for `which` and returns its
 * cudaIpcMemHandle (64 bytes):
 *   which = NSB_HALO_NODES / NSB_HALO_VERTS: acknowledgement words + a ring of receive slots shaped
 *           like the ghost block of that index space;
 *   which = NSB_HALO_NONE: the all-reduce block of the Krylov scalars (2..8 ranks).
 * The host gathers the handles of all ranks and calls nsb_p2p_connect
 *   - for a halo plan with, per neighbour in plan order: its handle, its ghost count, the offset
 *     (entries) of this rank's part inside its ghost block (its recv_ptr for this rank) and this
 *     rank's index in its peer list;
 *   - for NSB_HALO_NONE with npeers = nranks handles in rank order (own entry ignored), other
 *     tables NULL.
 * From then on nsb_halo_exchange (and the exchanges inside nsb_spmv / the Krylov solvers) is ONE
 * kernel: gather the boundary values and store them, each double with the exchange's epoch beside
 * it, into the neighbours' ring slots over NVLink; poll the local slot until every line carries the
 * epoch, write the ghosts, acknowledge.  The all-reduce of the inner products (<= 64 doubles) is one
 * single-block kernel of the same kind, summing in rank order (bit-identical on all ranks).
 * Waits are bounded; nsb_p2p_status returns a non-zero error word if one ever timed out (1: a
 * neighbour stopped consuming, 2: a neighbour stopped sending, 3: all-reduce).  All ranks must use
 * the same transport: nsb_p2p_disconnect (after a host barrier) returns to NCCL;
 * nsb_comm_finalize unmaps and frees the blocks (call it after a host barrier). */
/* out[k] = (float) mean of the diagonal of block k of a BSR value array (nsb_matrix.diag32) */
int nsb_bsr_block_scalar_f32(nsb_ctx* ctx, int64_t nnz, int32_t bs, const double* vals_dev, float* out_dev, void* stream);
/* The same fused into the assembly: while out_dev is registered (NULL unregisters), nsb_assemble_ipcs_momentum(NSB_MAT)
 * writes the block-scalar copy of the Jacobian it assembles alongside A_out (in the row-group kernel, from shared
 * memory: the 6.5 GB re-read at C5 goes away).  nnz_blocks = entries of the P2 x P2 pattern.  After
 * nsb_bc_apply_matrix changed Dirichlet rows, nsb_bsr_block_scalar_rows_f32 refreshes the listed block rows only. */
int nsb_set_block_scalar_out(nsb_ctx* ctx, float* out_dev, int64_t nnz_blocks);
int nsb_bsr_block_scalar_rows_f32(nsb_ctx* ctx, int32_t nlisted, const int32_t* block_rows_dev, const int32_t* rowptr_dev,
                                  int32_t bs, const double* vals_dev, float* out_dev, void* stream);

int nsb_p2p_export(nsb_ctx* ctx, int32_t which, uint8_t* handle_out64);
int nsb_p2p_connect(nsb_ctx* ctx, int32_t which, int32_t npeers, const uint8_t* handles64,
                    const int32_t* peer_n_ghost, const int32_t* peer_off, const int32_t* my_slot);
int nsb_p2p_status(nsb_ctx* ctx, int32_t* error_out);
int nsb_p2p_disconnect(nsb_ctx* ctx);   /* back to NCCL send/recv (all ranks, after a host barrier) */

/* ---------------------------------------------------------------- instrumentation ------- */
/* Per-kernel-class CUDA-event timing on the launching stream (SURVEY.md 8b nsb_timer_*).
 * enable = 1 starts recording an event pair around every launch group of every class; enable =
 * (mask << 1) with bit c of mask set times class c only (the event pairs cost about a microsecond per
 * launch group: bench.py times only the large kernels inside its headline region); 0 stops;
 * nsb_timer_report synchronises the device, writes accumulated milliseconds and the number of
 * timed regions per class (arrays of NSB_T_NCLASS; either may be NULL) and resets both. */
int nsb_timer_enable(nsb_ctx* ctx, int32_t enable);
int nsb_timer_report(nsb_ctx* ctx, double* ms_out, int64_t* regions_out);
/* kernels launched by this context since creation (every class) */
int nsb_launch_count(nsb_ctx* ctx, int64_t* count_out);

/* measured FP64 FMA rate (TFLOP/s) of this device: roofline denominator for element kernels */
int nsb_fp64_peak(nsb_ctx* ctx, double* tflops_out);

#ifdef __cplusplus
}
#endif
#endif
