// Integer plan produced by the host symbolic phase and consumed by the CUDA numeric phase.
//
// Replaces what MUMPS builds internally at JOB=1 (reference call site:
// uno/ingredients/subproblem_solvers/MUMPS/MUMPSSolver.cpp:72-83).  The plan is immutable once
// built; the numeric phase never changes front sizes (pivoting is confined to the fully-summed
// block of each front).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace b200kkt {

struct AnalyseOptions {
   int ordering = 0;          // 0 = METIS nested dissection on the KKT-compressed graph, 1 = natural, 2 = user permutation
   int kkt_pairing = 1;       // match each constraint row with a primal column (2x2 pivot candidates)
   int relax_small = 16;      // always merge a child when the merged pivot count is <= this (fronts up to relax_small_front rows)
   int relax_small_front = 256;
   double relax_chain_zeros = 0.02;   // wide fronts: merge when this step alone adds at most this fraction of explicit zeros ...
   int relax_chain_max_k = 256;       // ... and the merged front keeps at most this many pivots
   int relax_mid = 64;        // merged pivot count <= this: accept up to relax_mid_zeros explicit zeros
   double relax_mid_zeros = 0.5;
   int relax_big = 256;
   double relax_big_zeros = 0.15;
   double relax_any_zeros = 0.03;
   const int64_t* user_perm = nullptr;   // ordering == 2: perm[new] = old, length N
   // optional hint for the KKT pairing: matrix values in COO order (length nnz) as they are when the pattern is analysed (Uno
   // analyses a filled matrix: PrimalDualRegularization.hpp:145-149).  A multiplier row is then matched with the primal column
   // that carries its largest equilibrated Jacobian entry - the 2 x 2 pivot [[h, a], [a, -dc]] with the largest |a| - instead of
   // the primal of lowest degree.  Only the pairing uses it; nullptr = structural rule.
   const double* pair_values = nullptr;
   // METIS_NodeND tuning (0 = METIS default): separators tried per bisection, refinement iterations, allowed imbalance (1/1000)
   int metis_nseps = 0, metis_niter = 0, metis_ufactor = 0;
   // graphs of at least parallel_ordering_min_nodes (compressed) nodes: the top parallel_ordering_depth levels of the nested
   // dissection are split here and their halves ordered by parallel processes (2^depth of them at the bottom); 0 = METIS_NodeND alone
   int parallel_ordering_depth = 3;
   int64_t parallel_ordering_min_nodes = 150000;
};

struct Plan {
   int64_t N = 0, nnz = 0, reg_size = 0, num_primal = 0;
   int32_t nsn = 0;        // number of supernodes == number of fronts
   int32_t nlevels = 0;

   // perm[new] = old, iperm[old] = new.  "new" is the final elimination numbering in which every
   // supernode is a contiguous column range.
   std::vector<int32_t> perm, iperm;
   std::vector<int32_t> sn_first;   // nsn+1
   std::vector<int32_t> sn_of;      // N, by new index

   // per front
   std::vector<int32_t> k, f, ld;   // pivots, front order, leading dimension of the factor panel (f rounded up to 4)
   std::vector<int64_t> front_off;  // offset of the ld x k column-major factor panel [L11; L21] in the factor pool (doubles)
   std::vector<int32_t> ldc;        // leading dimension of the r x r contribution block (r = f - k rounded up to 4)
   std::vector<int64_t> cb_off;     // offset of the contribution block in the transient arena (lifetime: own level .. parent's level)
   std::vector<int64_t> rows_off;   // nsn+1, into rows
   std::vector<int32_t> rows;       // front row structure in new numbering; first k entries are the pivots
   std::vector<int32_t> parent;     // supernodal parent (-1 for roots)
   std::vector<int64_t> rel_off;    // nsn+1, into relmap (r = f - k entries per front)
   std::vector<int32_t> relmap;     // boundary row i of front s -> local index in parent(s)
   std::vector<int32_t> child_ptr, child_idx;     // children CSR, ascending child index
   std::vector<int32_t> level;                    // per front; leaves are 0
   std::vector<int32_t> level_ptr, level_fronts;  // fronts grouped by level, big fronts first

   int64_t pool_size = 0;           // factor pool, doubles: sum ld*k
   int64_t cb_pool_size = 0;        // contribution-block arena, doubles (peak of the level-order schedule)
   int64_t pool_size_global = 0, cb_pool_size_global = 0;   // the same for the whole system, before localize_storage

   // Assembly ("assembly permutation"): COO slot -> pool offset, and the same map grouped by
   // destination so that duplicates are summed in ascending slot order without atomics.
   std::vector<int64_t> scatter;    // nnz
   std::vector<int64_t> asm_dst;    // ndst
   std::vector<int32_t> asm_seg;    // ndst+1
   std::vector<int32_t> asm_src;    // nnz, slots grouped by destination
   std::vector<int32_t> asm_row, asm_col;   // nnz, ORIGINAL indices of asm_src entries (for scaling)

   // Full symmetric CSR of the COO in ORIGINAL numbering, duplicates kept, entries in ascending
   // slot order inside a row.  Used for the residual SpMV and for equilibration.
   std::vector<int32_t> csr_ptr;    // N+1
   std::vector<int32_t> csr_col;    // 2*nnz - ndiag
   std::vector<int32_t> csr_slot;

   // kept for re-planning with delayed pivots: symmetric off-diagonal adjacency (original numbering, sorted, unique) and the
   // COO pattern in 32-bit indices
   std::vector<int64_t> adj_ptr;
   std::vector<int32_t> adj_idx;
   std::vector<int32_t> coo_row, coo_col;

   // exact work figures
   int64_t nnz_L = 0;               // sum k*f - k(k-1)/2
   double flops = 0.0;              // sum k^3/3 + k^2 r + k r^2
   double flops_tc = 0.0;           // k^2 r + k r^2 over fronts with k >= 128
   int64_t sum_r = 0;
   int32_t max_f = 0, max_k = 0;
   int64_t num_pairs = 0;

   // Subtree-to-GPU mapping; filled by map_subtrees.  owner[s] = rank that assembles, factorizes and solves
   // front s.  A "cut edge" is a front whose parent lives on another rank: its contribution block (factorization), its boundary
   // partial sums (forward solve) and its boundary solution (backward solve) cross NVLink there and nowhere else.
   int32_t num_parts = 1;
   std::vector<int32_t> owner;       // nsn
   // shared[s] = 1: a big front above the cut, factorized by ALL ranks together: every rank holds the whole front, owns a
   // cyclic set of column chunks of its trailing matrix and factors the outer panels inside its chunks; each factored panel
   // is broadcast.  owner[s] of a shared front is the rank that counts its inertia and runs its triangular solves.
   std::vector<char> shared;         // nsn
   int64_t share_min_pivots = 1024;  // a front above the cut is shared when it has at least this many pivots ...
   double share_min_flops = 5e10;    // ... and this much work
   double partition_slack = 1.15;    // the tree is opened until the heaviest open subtree is at most this multiple of an even share
   std::vector<int32_t> cut_edges;   // fronts s with owner[s] != owner[parent[s]], ascending (level[s], s)
   std::vector<double> part_flops;   // factor flops per rank
   double critical_flops = 0.0;      // factor flops on the longest chain of dependent work (subtree + top fronts above it)
   double t_order = 0, t_symbolic = 0, t_maps = 0;
   std::string ordering_used;
};

// Throws std::runtime_error on malformed input (index out of range, lower-triangular entry).
void analyse(int64_t N, int64_t nnz, int64_t reg_size, int64_t num_primal,
             const uint64_t* row, const uint64_t* col, const AnalyseOptions& opt, Plan& plan);

// Rebuilds everything that follows from (P.perm, P.sn_first, P.nsn); see symbolic.cpp.
void build_structures(Plan& plan);
// Delayed pivots: the columns at positions cols[i] (current numbering) move into the ancestor front up[i] levels above their own
// and become its last pivots; the plan is rebuilt.  map_subtrees must be called again afterwards.
void delay_columns(Plan& plan, const std::vector<int32_t>& cols, const std::vector<int32_t>& up);

// Proportional mapping of the assembly tree onto `parts` ranks: the tree is opened from the roots (heaviest subtree first)
// until the open subtrees can be packed evenly, the subtrees are packed by decreasing flop count (LPT), and every front above
// the cut goes to the owner of its heaviest child.  Deterministic: every rank computes the same map from the same plan.
void map_subtrees(Plan& plan, int parts, bool share_top_fronts = true);
void localize_storage(Plan& plan, int rank);

}  // namespace b200kkt
