// Solver handle: owns the plan, the device copy of it, the front pool and all workspaces.
// One handle == one instance of Uno's DirectSymmetricIndefiniteLinearSolver (several coexist:
// uno/ingredients/constraint_relaxation_strategies/FeasibilityRestoration.cpp:33-34), so there is
// no global device state: every handle has its own stream and buffers.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "../../include/b200kkt.h"
#include "comm.hpp"
#include "kernels.hpp"
#include "plan.hpp"

namespace b200kkt {

// fronts of one elimination-tree level with similar order f, sorted by pivot count descending
struct Batch {
   int level = 0;
   int begin = 0, count = 0;      // range in the device batch list
   int max_f = 0, max_k = 0;
   bool has_children = false;
   bool has_cb = false;
   int max_r = 0;         // largest boundary (f - k) of the batch
   bool shared = false;   // fronts factorized by all ranks together (subtree partition, fronts above the cut)
   // shared batch: its fronts dealt round-robin into num_gpus sub-lists (each k descending, contiguous in the batch list);
   // sub-list g rotates the chunk ownership by g, so the panels of one step are factored by different ranks side by side
   std::vector<Batch> subs;
   int shift = 0;         // of a sub-list: chunk c belongs to rank (c + shift) mod num_gpus
   int big_tiles = 0;   // K4 tile mode (see launch_update)
   std::vector<int> k;            // per front of the batch (descending)
   std::vector<int> pmaxf;        // prefix maximum of f in batch order
   int count_k_greater(int j) const;   // number of fronts with k > j (a prefix of the batch)
};

}  // namespace b200kkt

struct b200kkt_handle {
   int device = -1;
   int num_gpus = 1;
   bool has_device = false;
   mutable std::string err;

   // options
   b200kkt::AnalyseOptions aopt;
   std::vector<int64_t> user_perm;
   std::vector<double> pair_values;          // pairing hint for the next analyse (b200kkt_set_pairing_values); cleared by it
   int scaling_sweeps = 3;
   double null_pivot_tol = 1e-14;
   // delayed pivots (kernels.hpp: FactorArrays::delay): a column that cannot be eliminated inside its front is moved into an
   // ancestor front by re-planning on the host and the factorization is repeated; the moves persist for later factorizations
   double delay_tol = 1e-12;         // scaled Schur entry outside the pivot block above which a null pivot is not null but delayed
   double growth_limit = 0.0;        // > 0: also delay a column with an L entry above this (guard against unbounded growth)
   // the guard switches itself on for ONE factorization (limit adaptive_growth_limit) after a solve whose first sweep missed
   // adaptive_growth_trigger in the normwise backward error: restricted pivoting let L grow; the columns responsible move up
   // and stay there, so the following factorizations of the pattern deliver the accuracy without the extra sweeps
   int positive_definite_target = 0;   // kernels.hpp: FactorArrays::no_pivot
   int adaptive_growth = 1;
   double adaptive_growth_limit = 1e8;
   double adaptive_growth_trigger = 1e-12;
   bool growth_armed = false;
   double growth_baked = 0.0;        // the limit the captured factor graph was built with
   int max_replans = 8;              // per factorization; beyond it the remaining columns are reported as null pivots
   std::vector<unsigned char> delay_count;   // per ORIGINAL column: times delayed so far (1st: parent, 2nd: 4 levels, then the root)
   int64_t delayed_total = 0, replans_total = 0, replans_last = 0;
   double* keep_values = nullptr;    // the resident value array survives a re-plan
   int refine_max_steps = 6;
   double refine_tol = 1e-15;
   double refine_tol_componentwise = 1e-12;   // on max_i |r_i| / (|K||x| + |b|)_i
   int use_graph = 1;
   int profile_updates = 0;
   struct ProfiledLaunch { cudaEvent_t a, b; double flops; int kind; };
   std::vector<ProfiledLaunch> profiled;

   b200kkt::Plan plan;
   bool analysed = false, values_resident = false, factorized = false;
   std::vector<b200kkt::Batch> batches;
   std::vector<int> level_batch_ptr;       // batches grouped by level
   std::vector<int> batch_fronts_host;     // host copy of d_batch_list
   // solve schedule: per level, a list of small fronts (one CTA each) and a list of big fronts
   // (k descending; NBS-wide panel steps over many CTAs)
   struct SolveLevel {
      int small_begin = 0, small_count = 0;
      int small_max_f = 0;                 // largest small front of the level: picks the CTA size of the one-CTA-per-front sweeps
      int big_begin = 0, big_count = 0;
      int big_max_f = 0, big_max_k = 0;
      std::vector<int> big_k;              // descending
      std::vector<int> big_pmaxf;          // prefix maximum of f
   };
   std::vector<SolveLevel> solve_levels;

   cudaStream_t stream = nullptr;
   static constexpr int NUM_AUX = 3;       // side streams: independent batches of one tree level run concurrently
   cudaStream_t aux[NUM_AUX] = {nullptr, nullptr, nullptr};
   cudaEvent_t ev_fork = nullptr, ev_join[NUM_AUX] = {nullptr, nullptr, nullptr};
   // look-ahead: the next outer panel of a front is factored on a panel stream while the rest of the trailing update runs
   cudaStream_t panel_stream[NUM_AUX + 1] = {nullptr, nullptr, nullptr, nullptr};
   cudaEvent_t ev_next_cols[NUM_AUX + 1] = {nullptr, nullptr, nullptr, nullptr}, ev_panel[NUM_AUX + 1] = {nullptr, nullptr, nullptr, nullptr};
   int fused_leaves = 2;     // small-front batches go through the one-CTA-per-front kernel (kernels.cu: k_leaf_front): 1 = leaf batches only, 2 = all
   int fused_assembly = 1;   // K1 and the zero-fill of the factor pool in one pass (kernels.cu: k_assemble_fill)
   int lookahead = 1;
   int lookahead_min_front = 2048;   // fronts of at least this order (with more than one outer panel) factor the next panel ahead
   int pair_panels = 1;      // trailing updates of very big fronts take two outer panels at once (K <= 512): +3 % on 128^3, -3 % on 64^3
   int pair_min_front = 16384;
   int debug_skip = 0;
   cudaGraphExec_t solve_graph = nullptr;   // one solve sweep, captured on first use
   double* solve_graph_x = nullptr;
   const double* solve_graph_scale = nullptr;
   int64_t solve_graph_launches = 0;
   int64_t xinv_size = 0;    // doubles of inverted diagonal blocks
   int64_t wpool_size = 0;   // doubles per W-panel buffer (four buffers: panel p uses buffer p & 3)
   int multi_stream = 1;
   int update_tile_mode = 4;   // K4 variant for fronts wider than 256: 1 = 128x128 tiles (8 warps), 2 = 128x64 (4 warps, two CTAs per SM), 3 = warp-specialised persistent, 4 = 128x64 (8 warps, two CTAs per SM), 5 = 128x128 (16 warps)
   cudaEvent_t ev0 = nullptr, ev1 = nullptr;
   cudaGraphExec_t factor_graph = nullptr;
   std::vector<void*> plan_allocs;         // freed on re-analysis / destroy

   b200kkt::FrontArrays F{};
   b200kkt::AssemblyArrays M{};
   b200kkt::CsrArrays C{};
   b200kkt::FactorArrays W{};
   b200kkt::GatherArrays Gt{};               // gather lists of the big fronts' forward solve (kernels.hpp)
   int gather_lists = 1;
   int small_solve_ctas = 1;  // levels whose small fronts have at most 96 rows run the one-CTA-per-front sweeps with 64-thread CTAs (same arithmetic, more fronts in flight)
   int* d_perm = nullptr;
   int* d_batch_list = nullptr;
   int* d_solve_list = nullptr;
   double* d_values = nullptr;
   double* d_scale = nullptr;
   double* d_scale_tmp = nullptr;
   double* d_zg = nullptr;
   double* d_z2 = nullptr;
   double* d_xg = nullptr;
   double* d_cpool = nullptr;
   double* d_partial = nullptr;            // partial sums of the backward panel products
   double* d_b = nullptr;
   double* d_x = nullptr;
   double* d_r = nullptr;
   // Jacobian entries of the analysed pattern grouped by variable (ipm.cu: right-hand side on the device); built on first use
   int* d_jt_ptr = nullptr;
   int* d_jt_slot = nullptr;
   int* d_jt_con = nullptr;
   int* d_counters = nullptr;              // 8 ints
   unsigned long long* d_norms = nullptr;  // 4 x u64: |K|_inf, |r|_inf, |x|_inf, |b|_inf
   int* h_counters = nullptr;              // pinned mirrors
   unsigned long long* h_norms = nullptr;
   double* d_scale_final = nullptr;        // whichever of d_scale/d_scale_tmp holds the result

   // ---- subtree partition over num_gpus ranks (one process per GPU) ------------------------------------
   // Every rank holds the whole integer plan and owns the fronts plan.owner assigns to it.  Contribution blocks of cut fronts
   // travel on comm_stream (NCCL send/recv straight between the two arenas: same plan, same offsets), so a rank keeps working
   // on its own levels while a block is in flight; the solve ships boundary vectors on the main stream.
   b200kkt::Comm comm;
   int rank = 0;
   bool partitioned() const { return num_gpus > 1; }
   struct CutLevel {                      // cut fronts of one tree level that touch this rank, ascending front index
      std::vector<int> fronts;            // solve: point to point between owner(front) and owner(parent)
      std::vector<int> factor_p2p;        // factorization: the subset whose parent is not a shared front
      std::vector<int> factor_bcast;      // factorization: fronts below a shared parent, broadcast by their owner (every rank takes part)
      int out_begin = 0, out_count = 0;   // ranges in d_cut_list: fronts this rank owns (sends CB / boundary sums, receives x)
      int in_begin = 0, in_count = 0;     // fronts whose parent this rank owns
   };
   std::vector<CutLevel> cut_levels;      // per level
   std::vector<int> wait_level;           // per level L: last child level with a block received for a front of level L (-1: none)
   std::vector<cudaEvent_t> ev_comm;      // per level: transfers of that level's cut fronts done (null where there are none)
   cudaStream_t comm_stream = nullptr;
   cudaEvent_t ev_level_done = nullptr;
   int partition_graph = 0;               // experimental, off: capture the partitioned factorization / solve sweep (NCCL calls included) in CUDA graphs
   int share_one_launch = 1;              // a rank updates all its column chunks of one step in one K4 launch (8 GPUs, 128^3: 1.407 -> 1.359 s with both)
   int partition_memory = 1;              // multi-GPU: a rank allocates the factor pool and the contribution-block arena for its own and the shared fronts only
   int share_trim_rows = 1;               // broadcast only the rows of a factored panel at and below its diagonal block
   int share_top_fronts = 1;              // big fronts above the cut are factorized by all ranks together
   int share_chunk_panels = 2;            // column chunk of a shared front, in outer panels; chunk c belongs to rank c mod num_gpus
   int share_min_pivots = 1024;
   double share_min_flops = 5e10;
   double partition_slack = 1.15;
   cudaEvent_t ev_sh_main = nullptr, ev_sh_panel = nullptr, ev_sh_comm = nullptr, ev_sh_bcast[2] = {nullptr, nullptr};
   std::vector<int64_t> woff_host, xoff_host;   // this rank's W-panel and inverted-block offsets per front
   int* d_cut_list = nullptr;
   signed char* d_col_owner = nullptr;
   int64_t comm_bytes_factor = 0, comm_bytes_solve = 0;   // per factorization / per sweep, sent + received by this rank

   uint64_t inertia[3] = {0, 0, 0};
   b200kkt_stats stats{};
   int64_t launches = 0;
   int64_t graph_launches = 0;             // kernels inside one replay of factor_graph
};
