// Launch wrappers of the sm_100a numeric phase.  Everything here is asynchronous on the stream
// passed in; no wrapper synchronizes.  Each wrapper returns the number of kernels it launched.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace b200kkt {

constexpr int NBI = 32;    // inner block: pivot search window of the in-front Bunch-Kaufman
constexpr int NBO = 256;   // outer panel: K of the trailing FP64 tensor-core update
constexpr int NBS = 256;   // panel width of the multi-CTA triangular solves
// per inner block of a front factored by the panel kernels: X = L11^{-1} and X2 = D^{-1} L11^{-1}, row-major NBI x NBI each;
// kept after the factorization: the triangular solves of big fronts multiply by X instead of substituting
constexpr int XINV_BLOCK_DOUBLES = 2 * NBI * NBI;

// device-resident integer plan (SoA over fronts)
struct FrontArrays {
   const int64_t* off;        // factor panel offset in lpool (ld x k, column-major, rows = all f front rows)
   const int* k;              // pivots
   const int* f;              // front order
   const int* ld;             // leading dimension of the factor panel
   const int64_t* cboff;      // contribution block offset in cbpool (ldc x r, lower triangle used)
   const int* ldc;
   const int* first;          // first pivot column (new numbering)
   const int64_t* rows_off;
   const int* rows;
   const int64_t* rel_off;
   const int* relmap;
   const int* child_ptr;
   const int* child_idx;
   const int64_t* woff;       // offset of the front's ld x NBO W panel in wpool
   const int64_t* xoff;       // offset of the front's inverted diagonal blocks in xinv, -1: none (front factored by the light kernels)
};

struct AssemblyArrays {
   int64_t ndst;
   const int64_t* dst;
   const int* seg;
   const int* src;
   const int* row;
   const int* col;
   // fused zero-fill + assembly: the pool in tiles of ASM_TILE doubles; destinations [tile_ptr[t], tile_ptr[t + 1]) fall into tile t
   int64_t ntiles;
   int64_t pool_doubles;
   const int* tile_ptr;
};
constexpr int ASM_TILE = 2048;   // doubles per tile (16 KB): what one CTA builds in shared memory and streams out

struct CsrArrays {
   int64_t n;
   const int* ptr;
   const int* col;
   const int* slot;
};

struct FactorArrays {
   double* lpool;    // factor panels
   double* cbpool;   // contribution-block arena
   double* wpool;    // W = L*D panels of the level being factorized
   double* xinv;     // inverted diagonal blocks (XINV_BLOCK_DOUBLES per inner block)
   double* dd;       // D diagonal, per pivot position (new numbering + in-front pivoting)
   double* de;       // D off-diagonal of 2x2 blocks (stored at the first column of the pair)
   int* ptype;       // 0: 1x1, 1/2: first/second column of a 2x2, 3: null pivot (row/column dropped)
   int* lperm;       // pivot position q = first+j holds original local column lperm[q] (0..k)
   int* blkflag;     // per inner block (indexed by its first pivot position): non-identity permutation
   // delayed pivots: delay[c] = 1 when plan column c (current numbering) could not be eliminated inside its front - every
   // fully-summed candidate in its pivot block is below the null-pivot tolerance while the column still has entries above
   // delay_tol in rows outside the block (or an L entry above growth_limit, when that guard is on); *ndelay counts them.
   // The host then moves these columns into an ancestor front and factorizes again (solver.cpp: run_factor).
   int* delay;
   int* ndelay;
   double delay_tol;
   double growth_limit;   // 0: guard off
   // 1: positive-definite target (the byrd preset's Hessian convexification, PrimalRegularization.hpp:90-129):
   // pivots are taken in the given order, 1 x 1 only, no search, no interchanges - a Cholesky-type LDL^T.  A pivot that is not
   // positive simply shows up in the inertia, which is all that caller asks ("is H + delta I positive definite yet?").
   int no_pivot;
};

// K1' equilibration sweep (power-of-two Ruiz step) and K-infinity norm
int launch_scale_sweep(const CsrArrays& A, const double* values, const double* sc_in, double* sc_out, cudaStream_t st);
int launch_fill(double* p, int64_t n, double v, cudaStream_t st);
int launch_matrix_inf_norm(const CsrArrays& A, const double* values, unsigned long long* out_bits, cudaStream_t st);
// K1 assembly: scatter COO values (scaled) into the zeroed factor pool, duplicates summed in slot order
int launch_assemble(const AssemblyArrays& M, const double* values, const double* scale, double* lpool, cudaStream_t st);
// K1 fused with the zero-fill of the factor pool: every tile of the pool is written exactly once, zeros and assembled entries
// together (full 32-byte sectors, no read-modify-write of a pool that a separate memset has just written)
int launch_assemble_fill(const AssemblyArrays& M, const double* values, const double* scale, double* lpool, cudaStream_t st);

// K5: zero the contribution blocks of the listed fronts and extend-add all their children into them
int launch_prepare(const FrontArrays& F, const int* d_list, int nfronts, int max_f, const FactorArrays& W, cudaStream_t st);
// K3 panel step for the outer panel at pivot column J0: (A) one CTA per front factors the NBO x NBO
// diagonal block (Bunch-Kaufman confined to NBI x NBI inner blocks), in place; (B) the rows below it:
// writes L in place and W = L*D into the W panel
int launch_diagpanel(const FrontArrays& F, const int* d_list, int nfronts, int max_k, int J0, const FactorArrays& W, double null_tol,
                     cudaStream_t st);
int launch_trsmpanel(const FrontArrays& F, const int* d_list, int nfronts, int max_f, int max_k, int J0, const FactorArrays& W,
                     cudaStream_t st);
// K2: leaf fronts (no children, k <= NBI, f - k <= leaf_front_max_boundary()) factored completely by one CTA each, contribution
// block included; replaces prepare + diagpanel + trsmpanel + update for such a batch
int leaf_front_max_boundary();
// (accumulate: the batch went through launch_prepare first - fronts with children - and the block is updated, not written)
int launch_leaf_fronts(const FrontArrays& F, const int* d_list, int nfronts, int max_r, const FactorArrays& W, double null_tol, bool accumulate,
                       cudaStream_t st);
// rows of the L columns left of each inner block follow the block's interchanges (all blocks at once)
int launch_permute_rows(const FrontArrays& F, const int* d_list, int nfronts, int max_k, const FactorArrays& W, cudaStream_t st);
// K4 Schur-complement update C -= W * L^T with FP64 tensor-core MMA.
// mode 0: inside the outer panel (after inner block j0); mode 1: trailing matrix behind the K range that starts at outer panel J0.
// The K range is one outer panel or two taken together (the second one's W panel in another buffer: wpool2); the columns of the
// trailing matrix can be restricted to [clo, chi) outer panels behind the K range (chi < 0: to the end) - look-ahead splits.
// policy decides what happens to fronts whose last panel lies in the K range ("last" fronts): 0 = the column range applies
// to every front alike, 1 = last fronts get their whole trailing matrix, 2 = last fronts are skipped, 3 = only last fronts.
struct UpdateRange {
   int kpanels = 1;
   int clo = 0, chi = -1;
   const double* wpool2 = nullptr;
   int policy = 0;
   int clip_to_pivots = 0;   // 1: the column range ends at the front's last pivot column (only factor-panel columns)
   // abs_hi > 0: additionally only columns [abs_lo, abs_hi) of the front (absolute front columns, multiples of NBO) - the column
   // chunks one rank owns when a front is shared by several GPUs
   int abs_lo = 0, abs_hi = 0;
   // cyc_cols > 0 (mode 1): only the column chunks of cyc_cols columns that belong to this rank - chunk c with
   // (c + cyc_shift) mod cyc_g == cyc_me - all of them in ONE launch (grid z runs over cyc_nchunks own chunks from the first one at
   // or behind the start of the column range); cyc_cols is a multiple of the tile width
   int cyc_cols = 0, cyc_g = 1, cyc_me = 0, cyc_shift = 0, cyc_nchunks = 0;
};
int launch_update(const FrontArrays& F, const int* d_list, int nfronts, int max_f, int mode, int j0, int J0, const FactorArrays& W,
                  int tile_mode, cudaStream_t st, const UpdateRange& R = UpdateRange());
// tile_mode 0: 64x64 (4 warps), 1: 128x128 (8 warps), 2: 128x64 (4 warps, two CTAs per SM), 3: warp-specialised persistent,
// 4: 128x64 (8 warps, two CTAs per SM), 5: 128x128 (16 warps), 6: 128x64 (8 warps), K tiles of 32 double-buffered
int init_kernel_attributes();
// phase cycle counters of the panel kernels (only when built with -DB200KKT_PHASES); returns 0 when not compiled in
int debug_phase_cycles(unsigned long long* out16, bool reset);
// micro-benchmark of the 32 x 32 Bunch-Kaufman block kernel: cycles per block (variant 0 dominant diagonal, 1 KKT-like, 2 random)
int debug_bench_bk(int variant, int reps, long long* cycles_out, int* n2x2);
// K6 inertia from the D blocks: counters[0..4] = positive, negative, zero, number of 2x2, unused
// col_owner != nullptr (subtree partition): only pivot positions of fronts this rank owns are counted
int launch_inertia(int64_t n, const FactorArrays& W, int* counters, cudaStream_t st, const signed char* col_owner = nullptr, int rank = 0);

// K7 solves, one level at a time.  Small fronts: one CTA per front.
int launch_forward_level(const FrontArrays& F, const int* d_list, int nfronts, const FactorArrays& W, double* zg, double* z2,
                         double* cpool, cudaStream_t st, int threads = 256);
int launch_backward_level(const FrontArrays& F, const int* d_list, int nfronts, const FactorArrays& W, double* z2, double* xg,
                          double* cpool, cudaStream_t st, int threads = 256);
// Big fronts: many CTAs per front, NBS-wide panel steps.
int launch_fwd_big_gather(const FrontArrays& F, const int* d_list, int nfronts, int max_f, const FactorArrays& W, double* zg, double* z2,
                          double* cpool, cudaStream_t st);
// The same gather from a list precomputed at analysis: for row d of big front s (gather_off[s] >= 0) the child entries that land on
// it are src[ptr[gather_off[s] + d] .. ptr[gather_off[s] + d + 1]) (offsets into cpool, children in ascending order: the summation
// order of k_fwd_big_gather, hence bit-identical results) - one thread per row, no loop over the children, no barriers.
struct GatherArrays {
   const int64_t* off;   // per front; -1: no list
   const int* ptr;
   const int* src;
};
int launch_fwd_big_gather_list(const FrontArrays& F, const GatherArrays& Gt, const int* d_list, int nfronts, int max_f, double* zg, double* cpool,
                               cudaStream_t st);
int launch_fwd_big_permute(const FrontArrays& F, const int* d_list, int nfronts, int max_k, const FactorArrays& W, const double* zg,
                           double* z2, cudaStream_t st);
int launch_fwd_big_diag(const FrontArrays& F, const int* d_list, int nfronts, int jp, const FactorArrays& W, double* z2, cudaStream_t st);
int launch_fwd_big_gemv(const FrontArrays& F, const int* d_list, int nfronts, int max_f, int jp, const FactorArrays& W, double* z2,
                        double* cpool, cudaStream_t st);
int launch_bwd_big_begin(const FrontArrays& F, const int* d_list, int nfronts, int max_f, const FactorArrays& W, double* z2,
                         const double* xg, double* cpool, cudaStream_t st);
int launch_bwd_big_gemv(const FrontArrays& F, const int* d_list, int nfronts, int jp, const FactorArrays& W, double* z2,
                        const double* cpool, double* partial, cudaStream_t st);
int launch_bwd_big_diag(const FrontArrays& F, const int* d_list, int nfronts, int jp, const FactorArrays& W, double* z2,
                        const double* partial, cudaStream_t st);
int bwd_partial_doubles_per_front();   // size of the partial-sum scratch of the backward panel product, per front of a level
int launch_bwd_big_scatter(const FrontArrays& F, const int* d_list, int nfronts, int max_k, const FactorArrays& W, const double* z2,
                           double* xg, cudaStream_t st);
// permute+scale in, scale+permute out (accumulating into x)
int launch_solve_gather(int64_t n, const int* perm, const double* scale, const double* r, double* zg, cudaStream_t st);
int launch_solve_scatter_add(int64_t n, const int* perm, const double* scale, const double* xg, double* x, cudaStream_t st);
// K9 (subtree partition): boundary solution of cut fronts into / out of their slot of the boundary pool; zero the solution
// entries of pivots owned by other ranks (col_owner: rank per pivot position)
int launch_boundary_pack(const FrontArrays& F, const int* d_list, int nfronts, const double* xg, double* cpool, cudaStream_t st);
int launch_boundary_unpack(const FrontArrays& F, const int* d_list, int nfronts, const double* cpool, double* xg, cudaStream_t st);
int launch_keep_owned(int64_t n, const signed char* col_owner, int rank, double* xg, cudaStream_t st);
// rows [row0, row0 + nrows) of ncols panel columns (leading dimension ld) <-> dense nrows x ncols staging block
int launch_rows_pack(const double* panel, int64_t ld, int row0, int nrows, int ncols, double* staging, cudaStream_t st);
int launch_rows_unpack(const double* staging, int64_t ld, int row0, int nrows, int ncols, double* panel, cudaStream_t st);
// K8 residual r = b - K x on the COO-order values, and infinity norms as ordered bit patterns
// (with the componentwise backward error max_i |r_i| / (|K||x| + |b|)_i: atomicMax on *omega_bits)
int launch_residual(const CsrArrays& A, const double* values, const double* x, const double* b, double* r,
                    const unsigned long long* xnorm_bits, unsigned long long* omega_bits, cudaStream_t st);
int launch_absmax(const double* v, int64_t n, unsigned long long* out_bits, cudaStream_t st);
// Allocation-free residual + backward errors on a file-scope device scratch array (kernels.cu: g_noalloc_scratch); blocking.
// out_bits = |K|_inf, |r|_inf, |x|_inf, |b|_inf, componentwise omega as ordered bit patterns.  Returns a cudaError_t value
// (cudaErrorInvalidValue when A.n exceeds noalloc_max_n()).
int64_t noalloc_max_n();
int launch_backward_error_noalloc(int device, const CsrArrays& A, const double* values, const double* x, const double* b,
                                  unsigned long long out_bits[5], int* launches, cudaStream_t st);

}  // namespace b200kkt
