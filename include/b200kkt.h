/* b200kkt — C ABI of the B200-native sparse symmetric-indefinite KKT solver.
 *
 * This is the drop-in boundary for the path behind Uno's
 * DirectSymmetricIndefiniteLinearSolver<size_t,double>
 * (reference: uno/ingredients/subproblem_solvers/DirectSymmetricIndefiniteLinearSolver.hpp:13-25,
 *  uno/ingredients/subproblem_solvers/SymmetricIndefiniteLinearSolver.hpp:20-33).
 * Every entry point below replaces one dmumps_c() call site or INFOG read of the reference's
 * MUMPSSolver (uno/ingredients/subproblem_solvers/MUMPS/MUMPSSolver.cpp); the line cited beside
 * each function is the reference interface it stands in for.
 *
 * Conventions: plain pointers and sizes only.  Matrix = upper-triangular COO, row <= col, 0-based
 * 64-bit indices, duplicates are summed, explicit zeros are structural, the first reg_size entries
 * are the (k,k) regularization slots (uno/linear_algebra/COOFormat.hpp:101-125).  Every function
 * returns 0 on success, a negative B200KKT_E* code on failure (message via b200kkt_last_error).
 * Singularity is NOT an error: it is reported through the inertia triple, as in the reference.
 * Handles are independent (own stream, plan, workspaces); calls on one handle must not overlap.
 * Host pointers are only borrowed for the duration of a call.
 */
#ifndef B200KKT_H
#define B200KKT_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200kkt_handle b200kkt_handle;

enum {
   B200KKT_OK = 0,
   B200KKT_EINVAL = -1,      /* bad handle / argument / call order */
   B200KKT_ECUDA = -2,       /* CUDA runtime error (no device, launch failure, ...) */
   B200KKT_ENOMEM = -3,      /* host or device allocation failed */
   B200KKT_EINTERNAL = -4    /* symbolic-phase failure */
};

/* Work figures and timings.  All counts are exact (computed from the plan), times in seconds. */
typedef struct b200kkt_stats {
   int64_t n, nnz, reg_size;
   int64_t num_fronts, num_levels, max_front, max_pivots, num_pairs;
   int64_t nnz_factor;         /* sum_s k_s f_s - k_s(k_s-1)/2 */
   int64_t pool_doubles;       /* device factor pool size (sum of ld x k panels) */
   int64_t assembly_distinct;  /* distinct destination slots of the scatter map */
   double flops_factor;        /* sum_s k^3/3 + k^2 r + k r^2 */
   double flops_tensor;        /* k^2 r + k r^2 over fronts with k >= 128 */
   double t_ordering, t_symbolic, t_maps, t_upload;       /* analysis */
   double t_factor_last, t_solve_last;                    /* device time of the last call (CUDA events) */
   int64_t refine_steps_last;
   double backward_error_last;
   int64_t kernel_launches;    /* cumulative count of this library's kernel launches */
   int64_t num_null_pivots_last, num_two_by_two_last;
   /* option "profile_updates" = 1: CUDA-event time, algorithmic flops and launch count of the K4
    * trailing-update launches on fronts of order >= 1024 during the last factorization */
   double t_update_profiled, flops_update_profiled;
   int64_t n_update_profiled;
   double t_assemble_profiled;   /* K1 (memset excluded) under the same option */
   /* subtree partition (num_gpus > 1): this rank's share, the estimated critical path, and the bytes this rank sends and
    * receives over NVLink per factorization / per solve sweep */
   int64_t num_parts, part_rank, num_cut_fronts;
   double flops_own, flops_critical;
   int64_t comm_bytes_factor, comm_bytes_sweep;
   /* option "profile_updates" on a multi-GPU handle: main-stream time inside the shared-front batches of the last factorization,
    * their flops (whole fronts, all ranks), and the part of that time spent waiting for a factored panel to arrive */
   double t_shared_profiled, flops_shared_profiled, t_shared_stall_profiled;
   /* same option: time from the start of the factorization to the first shared batch (this rank's subtrees + transfers), and the
    * part of the factorization the main stream spent waiting for contribution blocks from other ranks */
   double t_before_shared_profiled, t_transfer_wait_profiled;
   /* delayed pivots: columns moved into an ancestor front since the analysis, host re-plans since the analysis and during the
    * last factorization (each re-plan is followed by a new factorization attempt) */
   int64_t num_delayed_total, num_replans_total, num_replans_last;
   /* componentwise backward error of the last solve: max_i |b - K x|_i / (|K| |x| + |b|)_i (Oettli-Prager; MUMPS' omega_1) */
   double backward_error_componentwise_last;
   /* device storage in doubles.  pool_doubles (above) and cb_arena_doubles are what THIS rank allocates; with num_parts > 1 and the
    * option partition_memory (default 1) that is the fronts it owns plus the shared ones, the *_global figures are the whole system */
   int64_t cb_arena_doubles, pool_doubles_global, cb_arena_doubles_global;
} b200kkt_stats;

/* Handle life cycle.  Replaces JOB=-1 / JOB=-2 (MUMPSSolver.cpp:16-49).
 * device >= 0: CUDA device ordinal.  device == -1: host-only handle (symbolic phase and plan
 * getters work, every numeric call fails with B200KKT_ECUDA) — there is no CPU numeric path. */
int b200kkt_create(b200kkt_handle** h, int device, int num_gpus);
int b200kkt_destroy(b200kkt_handle* h);

/* Multi-GPU (num_gpus = 2, 4, 8 in b200kkt_create): one process per GPU, every process creates a handle with the same
 * num_gpus and runs the same call sequence on the same matrix (SPMD, as the reference's MUMPS does over MPI ranks:
 * MUMPSSolver.cpp:20 "comm_fortran = USE_COMM_WORLD").  Independent elimination subtrees are mapped to the ranks, the
 * contribution blocks of the cut fronts travel over NVLink (NCCL send/recv), inertia and solution are summed over ranks, so
 * every rank returns the same inertia and the same x.  Rank 0 calls b200kkt_comm_unique_id, the host program hands the 128
 * bytes to the other ranks by whatever means it has (MPI_Bcast, torch.distributed, a file), every rank calls
 * b200kkt_comm_init before b200kkt_analyse.  libnccl.so.2 is opened at run time by these two calls only. */
int b200kkt_comm_unique_id(char id[128]);
int b200kkt_comm_init(b200kkt_handle* h, int rank, int world, const char id[128]);
/* The subtree map of the last analysis: owner[s] per front (length = number of fronts), shared[s] = 1 for the big fronts above
 * the cut that all ranks factorize together (column chunks of the trailing matrix per rank, factored panels broadcast; owner[s]
 * then names the rank that solves with it), the cut fronts (owner differs from the parent's; at most number of fronts entries,
 * *num_cut receives the count), factor flops per rank (length num_gpus), flops on the critical path.
 * Works on host-only handles; any pointer may be NULL.
 * Options:  "share_top_fronts" (1), "share_min_pivots" (1024), "share_min_flops" (5e10), "share_chunk_panels" (2), "partition_slack" (1.15),
 * "partition_memory" (1): a rank allocates factor columns only for the fronts it owns and the shared ones, and room in the
 * contribution-block arena only for the blocks it produces, consumes or helps to broadcast (b200kkt_stats: pool_doubles,
 * cb_arena_doubles against the *_global figures); 0 = every rank allocates the whole system's pools (round-1 behaviour).
 * "debug_rank" (host-only handles): the rank whose storage the plan of the next analysis describes. */
int b200kkt_plan_partition(const b200kkt_handle* h, int32_t* owner, int8_t* shared, int64_t* num_cut, int32_t* cut_fronts,
                           double* part_flops, double* critical_flops);

/* Options (defaults need no user action).  Names: "ordering" (0 METIS, 1 natural),
 * "kkt_pairing", "relax_small", "relax_mid", "relax_big", "scaling_sweeps", "null_pivot_tol",
 * "refine_max_steps", "refine_tol", "use_graph", "profile_updates", and for the delayed pivots
 * "positive_definite_target" (0; 1 = Cholesky-type LDL^T without pivot search for the Hessian-convexification path of
 * PrimalRegularization.hpp:90-129, SURVEY 8(f) rank 4: a pivot that is not positive shows up in the inertia), and
 * "delay_tol" (1e-12), "growth_limit" (0 = off), "max_replans" (8), "adaptive_growth" (1), "adaptive_growth_limit" (1e8).
 *
 * Delayed pivots (what dmumps_c JOB=2 does with threshold pivoting; MUMPSSolver.cpp:85-89 with ICNTL(24)=1 at :36): the pivot
 * search of a front is confined to its fully-summed block.  A column whose candidates there are all below null_pivot_tol is a
 * zero eigenvalue only if the rest of its front column is below delay_tol as well; otherwise the factorization reports it, the
 * host moves it into an ancestor front (same elimination order for everything else, plan rebuilt, values stay on the device) and
 * factorizes again.  The moves persist, so later factorizations of the same pattern start from the adapted plan.
 * growth_limit > 0 treats a column with an L entry above the limit the same way (what the threshold test CNTL(1) is for);
 * adaptive_growth arms that guard for one factorization after a solve whose first sweep missed a normwise backward error of
 * "adaptive_growth_trigger" (1e-12).
 * b200kkt_get_permutation / plan getters return the adapted plan.
 *
 * Kernel and schedule variants (all default to the fastest measured; every one has a parity case in tests/test_gpu_parity.py):
 * "fused_assembly" (1: K1 fused with the zero-fill of the pool), "fused_leaves" (2: small fronts whole by one CTA, K2),
 * "gather_lists" (1: forward gather of the big fronts through lists built at analysis), "small_solve_ctas" (1: 64-thread CTAs for
 * levels of small fronts in the solve sweeps), "lookahead", "lookahead_min_front", "pair_panels", "pair_min_front", "multi_stream",
 * "update_tile_mode".  Host analysis: "parallel_ordering_depth" (3: levels of the nested dissection split by the library, halves
 * ordered by forked processes; 0 = METIS_NodeND alone), "parallel_ordering_min_nodes" (150000), "metis_niter/nseps/ufactor". */
int b200kkt_set_option(b200kkt_handle* h, const char* name, double value);
/* Optional user ordering for the next analyse (perm[new] = old, length N); NULL clears it. */
int b200kkt_set_user_permutation(b200kkt_handle* h, const int64_t* perm, uint64_t n);

/* Optional hint for the next b200kkt_analyse: the matrix values (COO order, length nnz) as they are when the pattern is analysed -
 * Uno hands do_symbolic_analysis a filled matrix (PrimalDualRegularization.hpp:145-149, Preprocessing.cpp:60), MUMPS ignores its
 * values at JOB=1 (MUMPSSolver.cpp:72-83).  The KKT pairing then matches every multiplier row with the primal column of its
 * largest equilibrated Jacobian entry, which keeps the 2 x 2 pivots well conditioned (far fewer refinement sweeps on problems
 * whose weakly coupled variables - controls, slacks - have the lowest degree).  Copied; used by one analysis; NULL clears it. */
int b200kkt_set_pairing_values(b200kkt_handle* h, const double* values, uint64_t nnz);

/* Symbolic analysis.  Replaces do_symbolic_analysis -> JOB=1 (MUMPSSolver.cpp:72-83,149-157).
 * Re-callable with another pattern/dimension (Preprocessing.cpp:60 then
 * PrimalDualRegularization.hpp:147).  num_primal = number of primal rows n (rows >= n are
 * multiplier rows); pass N (or 0) if unknown. */
int b200kkt_analyse(b200kkt_handle* h, uint64_t n, uint64_t nnz, uint64_t reg_size, uint64_t num_primal,
                    const uint64_t* row, const uint64_t* col);

/* Numeric factorization.  Replaces do_numerical_factorization -> JOB=2 (MUMPSSolver.cpp:85-89).
 * values: nnz doubles in analysis order (matrix.data_pointer()). */
int b200kkt_factorize(b200kkt_handle* h, const double* values);
/* Same, when only the leading reg_size slots changed since the last factorize (the inertia
 * correction loop, PrimalDualRegularization.hpp:178-186,210-211): uploads reg_size doubles. */
int b200kkt_refactorize_shifted(b200kkt_handle* h, const double* reg_values);
/* Same as factorize, values already resident in device memory. */
int b200kkt_factorize_device(b200kkt_handle* h, const double* d_values);

/* ---- the host steps either side of the factorization, on the device --------------------------------------------
 * No counterpart inside the reference's linear solver: these replace HOST loops of its callers, for models whose derivative
 * values live in device memory.  All pointers are device pointers; calls are enqueued on the handle's stream and, unless a
 * scalar comes back, return without waiting.  Sums run in the order of the reference's loops without FMA contraction. */

/* Layout of Uno's augmented-system COO (COOFormat.hpp:119-125, Subproblem.cpp:57-70, PrimalDualInteriorPointProblem.cpp:56-78):
 * [num_primal_slots x (k,k) | num_dual_slots x (k,k) | model Lagrangian Hessian | one barrier diagonal per bounded variable,
 *  ascending variable index | Jacobian entries (var, n + j), constraint by constraint]. */
typedef struct b200kkt_kkt_layout {
   uint64_t num_primal_slots, num_dual_slots, nnz_hessian, num_barrier, nnz_jacobian;
} b200kkt_kkt_layout;

/* Fills the handle's resident value array in one kernel: slots = (+delta_w, -delta_c) (COOFormat.hpp:101-110 as called from
 * PrimalDualRegularization.hpp:178-179), Hessian and Jacobian values copied, barrier diagonals z_L/(x-l) + z_U/(x-u) computed
 * (PrimalDualInteriorPointProblem.cpp:62-76; d_lower/d_upper hold +-inf where a bound is missing; d_barrier_variable[t] is the
 * variable of barrier slot t).  Replaces the host fill + the 8 nnz byte upload of b200kkt_factorize. */
int b200kkt_assemble_kkt_values_device(b200kkt_handle* h, const b200kkt_kkt_layout* layout, const double* d_hessian_values,
                                       const int32_t* d_barrier_variable, const double* d_x, const double* d_lower, const double* d_upper,
                                       const double* d_zl, const double* d_zu, const double* d_jacobian_values, double delta_w, double delta_c);
/* One step of the inertia-correction loop with no PCIe traffic: slots [0, num_primal_slots) <- +delta_w, the other
 * regularization slots <- -delta_c (PrimalDualRegularization.hpp:178-179,210-211); follow with b200kkt_factorize_resident. */
int b200kkt_shift_diagonal_device(b200kkt_handle* h, uint64_t num_primal_slots, double delta_w, double delta_c);
/* Numeric factorization of the resident values (b200kkt_factorize without the upload). */
int b200kkt_factorize_resident(b200kkt_handle* h);
/* The resident value array (nnz doubles, analysis order), e.g. for a model that writes its derivatives in place. */
int b200kkt_get_values_device(b200kkt_handle* h, double** d_values);
/* rhs = [-(gradient - J^T multipliers) ; -constraints] (Subproblem::assemble_augmented_rhs, Subproblem.cpp:80-99), the Jacobian
 * read from the resident values.  gradient: n, constraints and multipliers: m, rhs: n + m (n = num_primal of the analysis). */
int b200kkt_assemble_rhs_device(b200kkt_handle* h, const double* d_gradient, const double* d_constraints, const double* d_multipliers, double* d_rhs);
/* y = K x with the resident values (SymmetricMatrix::product, SymmetricMatrix.hpp:99-109). */
int b200kkt_symmetric_product_device(b200kkt_handle* h, const double* d_x, double* d_y);
/* PrimalDualInteriorPointProblem::assemble_primal_dual_direction (PrimalDualInteriorPointProblem.cpp:173-194): dx = solution[0:n],
 * dy = -solution[n:n+m], bound-dual directions (:262-277), the primal and dual fraction-to-boundary step lengths (:281-325,
 * tau = max(tau_min, 1 - mu) is the caller's), and the scaling of (dx, dy) and (dz_L, dz_U) by them.  Blocking (two scalars return). */
int b200kkt_unpack_direction_device(b200kkt_handle* h, double barrier_parameter, double tau, const double* d_solution, const double* d_x,
                                    const double* d_lower, const double* d_upper, const double* d_zl, const double* d_zu, double* d_dx, double* d_dy,
                                    double* d_dzl, double* d_dzu, double* primal_step_length, double* dual_step_length);

/* Inertia of the last factorization.  Replaces get_inertia / number_negative_eigenvalues /
 * matrix_is_singular / rank, i.e. INFOG(12) and INFOG(28) (MUMPSSolver.cpp:124-147). */
int b200kkt_inertia(const b200kkt_handle* h, uint64_t* n_pos, uint64_t* n_neg, uint64_t* n_zero);

/* Solve with the last factorization plus iterative refinement.  Replaces
 * solve_indefinite_system(matrix, rhs, result) -> JOB=3 (MUMPSSolver.cpp:91-96).
 * backward_error (may be NULL) receives ||b-Kx||_inf / (||K||_inf ||x||_inf + ||b||_inf); the componentwise measure
 * max_i |b-Kx|_i / (|K||x|+|b|)_i (Arioli-Demmel-Duff omega_1, omega_2 for rows with a denominator at rounding level) is in b200kkt_stats.backward_error_componentwise_last.  Refinement runs until both are below
 * their tolerances ("refine_tol" 1e-13, "refine_tol_componentwise" 1e-12) or stagnate.  A non-finite residual or pivot is an
 * error (B200KKT_EINTERNAL), as MUMPS' INFOG(1) < 0 would be. */
int b200kkt_solve(b200kkt_handle* h, const double* rhs, double* x, double* backward_error);
int b200kkt_solve_device(b200kkt_handle* h, const double* d_rhs, double* d_x, double* backward_error);

/* Backward errors of a given x for the last factorized matrix, WITHOUT ANY ALLOCATION: for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  No counterpart in the reference.
 * d_x, d_rhs: device pointers, length N.  normwise receives
 * ||b-Kx||_inf / (||K||_inf ||x||_inf + ||b||_inf), componentwise max_i |b-Kx|_i / (|K||x|+|b|)_i (either may be NULL) - the
 * two numbers b200kkt_solve reports for its own x.  The residual and the norm accumulators live in a static device array of
 * the library (B200KKT_NOALLOC_MAX_N doubles + 256 bytes, held by every GPU that loads the library until exit): N above
 * B200KKT_NOALLOC_MAX_N is rejected with B200KKT_EINVAL.  All handles share that array: calls that overlap on the GPU race.
 * Blocking; calls no cudaMalloc, cudaFree or allocating library routine. */
#define B200KKT_NOALLOC_MAX_N 1048576
int b200kkt_backward_error_noalloc(b200kkt_handle* h, const double* d_x, const double* d_rhs, double* normwise, double* componentwise);

/* Parity hooks: the assembly permutation.  perm[new] = old (length N); scatter_map[slot] = offset
 * of COO slot in the front pool (length nnz).  Either pointer may be NULL. */
int b200kkt_get_permutation(const b200kkt_handle* h, uint64_t* perm, uint64_t* scatter_map);
int b200kkt_get_stats(const b200kkt_handle* h, b200kkt_stats* stats);
const char* b200kkt_last_error(const b200kkt_handle* h);
/* The CUDA stream (cudaStream_t) every kernel of this handle is launched on, for event timing. */
int b200kkt_get_stream(const b200kkt_handle* h, void** stream);

/* Plan getters (integer plan of the symbolic phase).  sizes[0..7] = number of fronts, total row
 * entries, factor-pool doubles, levels, sum of boundary rows, distinct assembly destinations, contribution-block
 * arena doubles, 0. */
int b200kkt_plan_sizes(const b200kkt_handle* h, int64_t sizes[8]);
/* per front (length = number of fronts): pivots k, order f, leading dimension and pool offset of the ld x k factor panel,
 * parent front (-1 root), level, first pivot column (new numbering).  Any pointer may be NULL. */
int b200kkt_plan_fronts(const b200kkt_handle* h, int32_t* k, int32_t* f, int32_t* ld, int64_t* pool_offset,
                        int32_t* parent, int32_t* level, int32_t* first_col);
/* rows_off: number of fronts + 1; rows: total row entries (new numbering, pivots first). */
int b200kkt_plan_rows(const b200kkt_handle* h, int64_t* rows_off, int32_t* rows);

/* Debug/parity: run scaling + assembly only and copy the assembled front pool (pool_doubles) and
 * the row scaling (length N, may be NULL) back to the host. */
int b200kkt_debug_assemble(b200kkt_handle* h, const double* values, double* pool_out, double* scale_out);
/* Debug/parity: move the pivot columns at `positions` (current numbering) into the parent front (further up on repeated calls for
 * the same column), exactly as a factorization does for the columns it flags.  Works on host-only handles. */
int b200kkt_debug_delay_columns(b200kkt_handle* h, uint64_t count, const int64_t* positions);
/* Debug/parity: after factorize, copy back the factored pool, D (diag, offdiag: length N each, new
 * numbering incl. in-front pivoting), pivot types (0 1x1, 1/2 first/second of 2x2, 3 null) and the
 * in-front pivot permutation.  Any pointer may be NULL. */
int b200kkt_debug_factor(b200kkt_handle* h, double* pool_out, double* d_diag, double* d_off, int32_t* piv_type,
                         int32_t* local_perm);

/* Profiling hook: times `reps` launches of the K4 trailing update (FP64 tensor-core Schur update, K = k <= 256)
 * on one synthetic front of order f; returns seconds and algorithmic flops per launch. */
int b200kkt_debug_bench_update(int device, int f, int k, int reps, int big_tiles, double* seconds_per_launch, double* flops_per_launch);

/* Profiling hook: cycle counters of the panel-kernel phases; returns 0 unless the library was built with -DB200KKT_PHASES. */
int b200kkt_debug_phase_cycles(unsigned long long* out16, int reset);

/* Profiling hook: cycles per 32 x 32 Bunch-Kaufman block factorization (one warp), for three kinds of blocks. */
int b200kkt_debug_bench_bk(int variant, int reps, long long* cycles);

#ifdef __cplusplus
}
#endif
#endif /* B200KKT_H */
