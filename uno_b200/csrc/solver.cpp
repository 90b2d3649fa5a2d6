// Host orchestration of the B200 KKT solver and its C ABI (include/b200kkt.h).
//
// Mirrors the call sequence of the reference's MUMPSSolver
// (uno/ingredients/subproblem_solvers/MUMPS/MUMPSSolver.cpp): create <-> JOB=-1 (:16-44),
// analyse <-> JOB=1 (:72-83), factorize <-> JOB=2 (:85-89), solve <-> JOB=3 (:91-96),
// inertia <-> INFOG(12)/INFOG(28) (:124-147), destroy <-> JOB=-2 (:46-49).
// There is no CPU numeric path: a handle without a CUDA device can only run the symbolic phase.
#include "solver.hpp"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <new>
#include <stdexcept>

using namespace b200kkt;

namespace b200kkt {
int Batch::count_k_greater(int j) const {
   // k is sorted descending: first index with k <= j
   int lo = 0, hi = static_cast<int>(k.size());
   while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (k[mid] > j) lo = mid + 1; else hi = mid;
   }
   return lo;
}
}  // namespace b200kkt

namespace {

struct CudaError {
   cudaError_t code;
   const char* what;
};

#define CK(expr)                                                     \
   do {                                                              \
      const cudaError_t e__ = (expr);                                \
      if (e__ != cudaSuccess) throw CudaError{e__, #expr};           \
   } while (0)

// NCCL wrappers return non-zero with a message; surface them as an internal error of the call in progress
#define CKN(expr)                                                    \
   do {                                                              \
      std::string e__;                                               \
      if ((expr) != 0) throw std::runtime_error(e__);                \
   } while (0)

// an NCCL group that is closed on every path out of the scope (an exception between start and end must not leave it open)
struct CommGroup {
   bool open = false;
   CommGroup() { std::string e; if (comm_group_start(e) != 0) throw std::runtime_error(e); open = true; }
   void end() { if (open) { open = false; std::string e; if (comm_group_end(e) != 0) throw std::runtime_error(e); } }
   ~CommGroup() { if (open) { std::string e; comm_group_end(e); } }
   CommGroup(const CommGroup&) = delete;
   CommGroup& operator=(const CommGroup&) = delete;
};

int fail(const b200kkt_handle* h, int code, const std::string& msg) {
   if (h != nullptr) h->err = msg;
   return code;
}

template <typename T>
T* dev_upload(b200kkt_handle* h, const std::vector<T>& v, size_t min_count = 1) {
   T* p = nullptr;
   const size_t count = std::max(v.size(), min_count);
   CK(cudaMalloc(reinterpret_cast<void**>(&p), count * sizeof(T)));
   h->plan_allocs.push_back(p);
   if (!v.empty()) CK(cudaMemcpyAsync(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, h->stream));
   return p;
}

template <typename T>
T* dev_alloc(b200kkt_handle* h, size_t count) {
   T* p = nullptr;
   CK(cudaMalloc(reinterpret_cast<void**>(&p), std::max<size_t>(count, 1) * sizeof(T)));
   h->plan_allocs.push_back(p);
   return p;
}

void free_plan_device(b200kkt_handle* h) {
   if (h->factor_graph != nullptr) { cudaGraphExecDestroy(h->factor_graph); h->factor_graph = nullptr; }
   if (h->solve_graph != nullptr) { cudaGraphExecDestroy(h->solve_graph); h->solve_graph = nullptr; }
   for (void* p : h->plan_allocs) cudaFree(p);
   h->plan_allocs.clear();
   h->d_jt_ptr = h->d_jt_slot = h->d_jt_con = nullptr;
   for (cudaEvent_t e : h->ev_comm) if (e != nullptr) cudaEventDestroy(e);
   h->ev_comm.clear();
   h->analysed = false;
   h->values_resident = false;
   h->factorized = false;
}

int size_class(int f) {
   int c = 0;
   int bound = 64;
   while (f > bound) { bound *= 2; ++c; }
   return c;
}

// group the fronts of every level into batches of similar order, pivot count descending
void build_batches(b200kkt_handle* h, std::vector<int>& batch_list, std::vector<int64_t>& woff, int64_t& wpool_size, std::vector<int64_t>& xoff,
                   int64_t& xinv_size) {
   const Plan& P = h->plan;
   h->batches.clear();
   h->level_batch_ptr.assign(1, 0);
   batch_list.clear();
   woff.assign(static_cast<size_t>(P.nsn), 0);
   xoff.assign(static_cast<size_t>(P.nsn), -1);
   xinv_size = 0;
   wpool_size = 0;
   std::vector<int> ids;
   for (int l = 0; l < P.nlevels; ++l) {
      ids.clear();
      for (int t = P.level_ptr[l]; t < P.level_ptr[l + 1]; ++t)
         if (P.owner[P.level_fronts[t]] == h->rank || P.shared[P.level_fronts[t]]) ids.push_back(P.level_fronts[t]);   // subtree partition: this rank's fronts and the shared ones
      std::stable_sort(ids.begin(), ids.end(), [&](int a, int b) {
         if (P.shared[a] != P.shared[b]) return P.shared[a] > P.shared[b];   // shared batches first, the same on every rank
         const int ca = size_class(P.f[a]), cb = size_class(P.f[b]);
         if (ca != cb) return ca > cb;
         if (P.k[a] != P.k[b]) return P.k[a] > P.k[b];
         return P.f[a] > P.f[b];
      });
      int64_t level_w = 0;
      size_t i = 0;
      while (i < ids.size()) {
         Batch b;
         b.level = l;
         b.begin = static_cast<int>(batch_list.size());
         const int cls = size_class(P.f[ids[i]]);
         b.shared = P.shared[ids[i]] != 0;
         while (i < ids.size() && size_class(P.f[ids[i]]) == cls && (P.shared[ids[i]] != 0) == b.shared) {
            const int s = ids[i];
            batch_list.push_back(s);
            b.k.push_back(P.k[s]);
            b.max_f = std::max(b.max_f, P.f[s]);
            b.pmaxf.push_back(b.max_f);
            b.max_k = std::max(b.max_k, P.k[s]);
            if (P.child_ptr[s + 1] > P.child_ptr[s]) b.has_children = true;
            if (P.f[s] > P.k[s]) b.has_cb = true;
            b.max_r = std::max(b.max_r, P.f[s] - P.k[s]);
            ++i;
         }
         b.count = static_cast<int>(b.k.size());
         if (b.shared) {
            // deal the fronts (k descending) round-robin into num_gpus sub-lists and lay the sub-lists out one after the other
            const int G = std::max(1, h->num_gpus);
            std::vector<int> dealt;
            for (int g = 0; g < G && g < b.count; ++g) {
               Batch sb;
               sb.level = l; sb.shared = true; sb.shift = g;
               sb.begin = b.begin + static_cast<int>(dealt.size());
               for (int t = g; t < b.count; t += G) {
                  const int s = batch_list[static_cast<size_t>(b.begin + t)];
                  dealt.push_back(s);
                  sb.k.push_back(P.k[s]);
                  sb.max_f = std::max(sb.max_f, P.f[s]);
                  sb.pmaxf.push_back(sb.max_f);
                  sb.max_k = std::max(sb.max_k, P.k[s]);
                  if (P.f[s] > P.k[s]) sb.has_cb = true;
               }
               sb.count = static_cast<int>(sb.k.size());
               b.subs.push_back(std::move(sb));
            }
            std::copy(dealt.begin(), dealt.end(), batch_list.begin() + b.begin);
         }
         // W panels; a batch that runs the panel kernels (some front with more than NBI pivots) also gets room for the
         // inverted diagonal blocks of the current outer panel behind every W panel
         for (int t = 0; t < b.count; ++t) {
            const int s = batch_list[static_cast<size_t>(b.begin + t)];
            woff[s] = level_w;
            level_w += static_cast<int64_t>(P.ld[s]) * std::min(NBO, (P.k[s] + 3) & ~3);
            if (b.max_k > NBI) {
               xoff[s] = xinv_size;
               xinv_size += static_cast<int64_t>((P.k[s] + NBI - 1) / NBI) * XINV_BLOCK_DOUBLES;
            }
         }
         b.big_tiles = (b.max_f > 256) ? h->update_tile_mode : 0;
         for (Batch& sb : b.subs) sb.big_tiles = b.big_tiles;
         h->batches.push_back(std::move(b));
      }
      wpool_size = std::max(wpool_size, level_w);
      h->level_batch_ptr.push_back(static_cast<int>(h->batches.size()));
   }
}

int64_t cb_doubles(const Plan& P, int s) {
   const int64_t r = P.f[s] - P.k[s];
   return r * (r + 1) / 2;
}

// Subtree partition: which cut fronts this rank sends or receives at which tree level, the device lists for the boundary
// pack/unpack kernels, and the rank of every pivot position (for the sum of the solution over ranks).
void build_cut_schedule(b200kkt_handle* h) {
   const Plan& P = h->plan;
   for (cudaEvent_t e : h->ev_comm) if (e != nullptr) cudaEventDestroy(e);
   h->ev_comm.assign(static_cast<size_t>(P.nlevels), nullptr);
   h->cut_levels.assign(static_cast<size_t>(P.nlevels), b200kkt_handle::CutLevel());
   h->wait_level.assign(static_cast<size_t>(P.nlevels), -1);
   h->comm_bytes_factor = h->comm_bytes_solve = 0;
   h->d_cut_list = nullptr;
   h->d_col_owner = nullptr;
   if (!h->partitioned()) return;
   for (const int s : P.cut_edges) {   // ascending (level, front)
      const int p = P.parent[s];
      if (P.owner[s] != h->rank && P.owner[p] != h->rank) continue;
      h->cut_levels[P.level[s]].fronts.push_back(s);
      h->comm_bytes_solve += 2 * 8 * static_cast<int64_t>(P.f[s] - P.k[s]);
   }
   // factorization: a front below a shared parent is needed by every rank (broadcast by its owner); below a parent owned by
   // another rank it goes there alone; a shared front spreads its own contribution block when it is done (enqueue_shared_batch)
   for (int32_t l = 0; l < P.nlevels; ++l) {
      for (int32_t t = P.level_ptr[l]; t < P.level_ptr[l + 1]; ++t) {
         const int s = P.level_fronts[t];
         const int p = P.parent[s];
         if (p < 0 || P.shared[s] || P.f[s] == P.k[s]) continue;
         if (P.shared[p]) {
            h->cut_levels[l].factor_bcast.push_back(s);
            h->comm_bytes_factor += 8 * cb_doubles(P, s);
            if (P.owner[s] != h->rank) h->wait_level[P.level[p]] = std::max(h->wait_level[P.level[p]], l);
         } else if (P.owner[s] != P.owner[p] && (P.owner[s] == h->rank || P.owner[p] == h->rank)) {
            h->cut_levels[l].factor_p2p.push_back(s);
            h->comm_bytes_factor += 8 * cb_doubles(P, s);
            // the receiver assembles the block at the parent's level; the sender must not pass that level either before the
            // block has left: the arena region is free for its later fronts from there on
            h->wait_level[P.level[p]] = std::max(h->wait_level[P.level[p]], l);
         }
      }
      std::sort(h->cut_levels[l].factor_p2p.begin(), h->cut_levels[l].factor_p2p.end());
      std::sort(h->cut_levels[l].factor_bcast.begin(), h->cut_levels[l].factor_bcast.end());
   }
   std::vector<int> cut_list;
   for (int l = 0; l < P.nlevels; ++l) {
      b200kkt_handle::CutLevel& CL = h->cut_levels[l];
      if (!CL.factor_p2p.empty() || !CL.factor_bcast.empty()) CK(cudaEventCreateWithFlags(&h->ev_comm[l], cudaEventDisableTiming));
      if (CL.fronts.empty()) continue;
      CL.out_begin = static_cast<int>(cut_list.size());
      for (const int s : CL.fronts) if (P.owner[s] == h->rank) cut_list.push_back(s);
      CL.out_count = static_cast<int>(cut_list.size()) - CL.out_begin;
      CL.in_begin = static_cast<int>(cut_list.size());
      for (const int s : CL.fronts) if (P.owner[s] != h->rank) cut_list.push_back(s);
      CL.in_count = static_cast<int>(cut_list.size()) - CL.in_begin;
   }
   h->d_cut_list = dev_upload(h, cut_list);
   std::vector<signed char> col_owner(static_cast<size_t>(P.N));
   for (int32_t s = 0; s < P.nsn; ++s)
      for (int32_t q = P.sn_first[s]; q < P.sn_first[s + 1]; ++q) col_owner[q] = static_cast<signed char>(P.owner[s]);
   h->d_col_owner = dev_upload(h, col_owner);
}

// NCCL group with the transfers of the cut fronts of tree level l: `what` 0 = contribution blocks child -> parent (factorization;
// to every rank when the parent is a shared front), 1 = boundary partial sums child -> parent (forward solve), 2 = boundary
// solution parent -> child (backward solve)
void enqueue_cut_transfers(b200kkt_handle* h, int l, int what, double* pool, cudaStream_t st) {
   const Plan& P = h->plan;
   const b200kkt_handle::CutLevel& CL = h->cut_levels[l];
   const std::vector<int>& list = what == 0 ? CL.factor_p2p : CL.fronts;
   if (list.empty() && (what != 0 || CL.factor_bcast.empty())) return;
   CommGroup group;
   for (const int s : list) {
      const bool mine = P.owner[s] == h->rank;
      const int peer = mine ? P.owner[P.parent[s]] : P.owner[s];
      const int64_t count = what == 0 ? cb_doubles(P, s) : static_cast<int64_t>(P.f[s] - P.k[s]);
      if (count <= 0) continue;
      double* ptr = pool + (what == 0 ? P.cb_off[s] : P.rel_off[s]);
      const bool send = (what == 2) ? !mine : mine;
      if (send) CKN(comm_send_f64(h->comm, ptr, static_cast<size_t>(count), peer, st, e__));
      else CKN(comm_recv_f64(h->comm, ptr, static_cast<size_t>(count), peer, st, e__));
   }
   if (what == 0)
      for (const int s : CL.factor_bcast)
         CKN(comm_bcast_bytes(h->comm, pool + P.cb_off[s], static_cast<size_t>(cb_doubles(P, s)) * sizeof(double), P.owner[s], st, e__));
   group.end();
}

// K2-K4 of a batch of SHARED fronts (subtree partition; every rank runs this with the same batch).  The trailing matrix of a
// front is cut into column chunks of share_chunk_panels outer panels, chunk c belongs to rank c mod num_gpus.  A rank updates only
// its chunks; the rank that owns the chunk of outer panel p factors that panel and broadcasts it (L columns, W panel, inverted
// diagonal blocks, D, pivot types and orders) over NVLink, one panel ahead of the updates that need it, so the transfer and the
// latency-bound panel chain hide behind the FP64 tensor-core work of the other ranks.  At the end the contribution block is put
// together on every rank (each chunk broadcast by its owner): the parent front, shared too, assembles it everywhere.
void enqueue_shared_batch(b200kkt_handle* h, const Batch& b, const int* /*list*/, cudaStream_t st) {
   const Plan& P = h->plan;
   int64_t& L = h->launches;
   const int G = h->num_gpus, me = h->rank, CH = std::max(1, h->share_chunk_panels);
   cudaStream_t pst = h->panel_stream[0], cs = h->comm_stream;
   FactorArrays Wbuf[2] = {h->W, h->W};
   Wbuf[1].wpool += h->wpool_size;
   const int npan = (b.max_k + NBO - 1) / NBO;
   auto sub_list = [&](const Batch& sb) { return h->d_batch_list + sb.begin; };
   auto has_panel = [&](const Batch& sb, int p) { return p * NBO < sb.max_k; };
   auto panel_owner = [&](const Batch& sb, int p) { return (p / CH + sb.shift) % G; };
   auto chain = [&](const Batch& sb, int p, cudaStream_t s_) {
      const int J0 = p * NBO, nf = sb.count_k_greater(J0);
      L += launch_diagpanel(h->F, sub_list(sb), nf, sb.max_k, J0, Wbuf[p & 1], h->null_pivot_tol, s_);
      L += launch_trsmpanel(h->F, sub_list(sb), nf, sb.pmaxf[nf - 1], sb.max_k, J0, Wbuf[p & 1], s_);
   };
   // one NCCL group per step: panel p of every sub-list, each from the rank that factored it
   // option share_trim_rows (experimental, off): only rows >= J0 of the L columns and of the W panel travel, through two staging
   // blocks in the W buffers this path does not use (2 and 3); pack on the root before the group, unpack on the others behind it
   const bool trim = h->share_trim_rows != 0;
   double* stage_l = h->W.wpool + 2 * h->wpool_size;
   double* stage_w = h->W.wpool + 3 * h->wpool_size;
   auto for_panel_fronts = [&](int p, auto&& fn) {
      const int J0 = p * NBO;
      for (const Batch& sb : b.subs) {
         if (!has_panel(sb, p)) continue;
         const int nf = sb.count_k_greater(J0), root = panel_owner(sb, p);
         for (int t = 0; t < nf; ++t) fn(h->batch_fronts_host[sb.begin + t], root);
      }
   };
   auto bcast_panels = [&](int p) {
      const int J0 = p * NBO;
      if (trim)
         for_panel_fronts(p, [&](int s, int root) {
            if (root != me) return;
            const int kc = std::min(NBO, P.k[s] - J0), nrows = P.f[s] - J0;
            L += launch_rows_pack(h->W.lpool + P.front_off[s] + static_cast<int64_t>(P.ld[s]) * J0, P.ld[s], J0, nrows, kc, stage_l + h->woff_host[s], cs);
            L += launch_rows_pack(Wbuf[p & 1].wpool + h->woff_host[s], P.ld[s], J0, nrows, kc, stage_w + h->woff_host[s], cs);
         });
      CommGroup group;
      for (const Batch& sb : b.subs) {
         if (!has_panel(sb, p)) continue;
         const int nf = sb.count_k_greater(J0), root = panel_owner(sb, p);
         for (int t = 0; t < nf; ++t) {
            const int s = h->batch_fronts_host[sb.begin + t];
            const size_t kc = static_cast<size_t>(std::min(NBO, P.k[s] - J0)), ld = static_cast<size_t>(P.ld[s]);
            const int64_t q = P.sn_first[s] + J0;
            if (trim) {
               const size_t count = static_cast<size_t>(P.f[s] - J0) * kc;
               CKN(comm_bcast_bytes(h->comm, stage_l + h->woff_host[s], count * sizeof(double), root, cs, e__));
               CKN(comm_bcast_bytes(h->comm, stage_w + h->woff_host[s], count * sizeof(double), root, cs, e__));
            }
            else {
               CKN(comm_bcast_bytes(h->comm, h->W.lpool + P.front_off[s] + static_cast<int64_t>(ld) * J0, ld * kc * sizeof(double), root, cs, e__));
               CKN(comm_bcast_bytes(h->comm, Wbuf[p & 1].wpool + h->woff_host[s], ld * kc * sizeof(double), root, cs, e__));
            }
            if (h->xoff_host[s] >= 0)
               CKN(comm_bcast_bytes(h->comm, h->W.xinv + h->xoff_host[s] + static_cast<int64_t>(J0 / NBI) * XINV_BLOCK_DOUBLES,
                                    ((kc + NBI - 1) / NBI) * XINV_BLOCK_DOUBLES * sizeof(double), root, cs, e__));
            CKN(comm_bcast_bytes(h->comm, h->W.dd + q, kc * sizeof(double), root, cs, e__));
            CKN(comm_bcast_bytes(h->comm, h->W.de + q, kc * sizeof(double), root, cs, e__));
            CKN(comm_bcast_bytes(h->comm, h->W.ptype + q, kc * sizeof(int), root, cs, e__));
            CKN(comm_bcast_bytes(h->comm, h->W.lperm + q, kc * sizeof(int), root, cs, e__));
            CKN(comm_bcast_bytes(h->comm, h->W.blkflag + q, kc * sizeof(int), root, cs, e__));
         }
      }
      group.end();
      if (trim)
         for_panel_fronts(p, [&](int s, int root) {
            if (root == me) return;
            const int kc = std::min(NBO, P.k[s] - J0), nrows = P.f[s] - J0;
            L += launch_rows_unpack(stage_l + h->woff_host[s], P.ld[s], J0, nrows, kc, h->W.lpool + P.front_off[s] + static_cast<int64_t>(P.ld[s]) * J0, cs);
            L += launch_rows_unpack(stage_w + h->woff_host[s], P.ld[s], J0, nrows, kc, Wbuf[p & 1].wpool + h->woff_host[s], cs);
         });
   };
   auto update = [&](const Batch& sb, int p, int lo, int hi) {
      if (lo >= hi || lo >= sb.max_f) return;
      const int J0 = p * NBO, nf = sb.count_k_greater(J0);
      UpdateRange R;
      R.abs_lo = lo;
      R.abs_hi = hi;
      L += launch_update(h->F, sub_list(sb), nf, sb.pmaxf[nf - 1], 1, 0, J0, Wbuf[p & 1], sb.big_tiles, st, R);
   };
   auto link = [&](cudaEvent_t ev, cudaStream_t from, cudaStream_t to) {
      CK(cudaEventRecord(ev, from));
      CK(cudaStreamWaitEvent(to, ev, 0));
   };
   // option profile_updates: time of the whole batch (kind 2) and of the waits for a broadcast panel (kind 3) on the main stream
   const bool prof = h->profile_updates != 0;
   auto prof_begin = [&](double flops, int kind) {
      if (!prof) return;
      b200kkt_handle::ProfiledLaunch pl{nullptr, nullptr, flops, kind};
      CK(cudaEventCreate(&pl.a));
      CK(cudaEventCreate(&pl.b));
      CK(cudaEventRecord(pl.a, st));
      h->profiled.push_back(pl);
   };
   auto prof_end = [&](size_t idx) { if (prof) CK(cudaEventRecord(h->profiled[idx].b, st)); };
   double batch_flops = 0.0;
   for (int t = 0; t < b.count; ++t) {
      const int s = h->batch_fronts_host[b.begin + t];
      const double kd = P.k[s], rd = P.f[s] - P.k[s];
      batch_flops += kd * kd * kd / 3.0 + kd * kd * rd + kd * rd * rd;
   }
   prof_begin(batch_flops, 2);
   const size_t prof_batch = h->profiled.size() - 1;
   // panel 0 of every sub-list
   for (const Batch& sb : b.subs) if (panel_owner(sb, 0) == me) chain(sb, 0, st);
   link(h->ev_sh_main, st, cs);
   bcast_panels(0);
   link(h->ev_sh_comm, cs, st);
   for (int p = 0; p < npan; ++p) {
      const bool next = p + 1 < npan;
      bool chained = false;
      if (next) {
         // the sub-lists whose panel p + 1 is mine: that panel's own columns first, then the panel on the panel stream while
         // the rest is updated here
         for (const Batch& sb : b.subs)
            if (has_panel(sb, p + 1) && panel_owner(sb, p + 1) == me) { update(sb, p, (p + 1) * NBO, (p + 2) * NBO); chained = true; }
         if (chained) {
            link(h->ev_sh_main, st, pst);
            CK(cudaStreamWaitEvent(pst, h->ev_sh_bcast[(p + 1) & 1], 0));   // the broadcast of panel p - 1 has left this W buffer
            for (const Batch& sb : b.subs)
               if (has_panel(sb, p + 1) && panel_owner(sb, p + 1) == me) chain(sb, p + 1, pst);
            link(h->ev_sh_panel, pst, cs);
         }
         link(h->ev_sh_main, st, cs);                  // the W buffer of panel p + 1 is free: the updates of panel p - 1 are done
         bcast_panels(p + 1);
         CK(cudaEventRecord(h->ev_sh_bcast[(p + 1) & 1], cs));
      }
      for (const Batch& sb : b.subs) {
         if (!has_panel(sb, p)) continue;
         const bool next_mine = next && has_panel(sb, p + 1) && panel_owner(sb, p + 1) == me;
         // a front whose last panel is p and whose pivot count is not a multiple of NBO has contribution-block columns inside
         // panel p's own column range
         bool stragglers = false;
         for (const int kf : sb.k) stragglers = stragglers || (kf > p * NBO && kf < (p + 1) * NBO);
         const int nchunks = (sb.max_f + CH * NBO - 1) / (CH * NBO);
         // option share_one_launch (experimental, off): every own chunk behind the one that holds panel p + 1 in ONE launch (the W
         // panel is then read once per step, not once per chunk); the chunks up to there go through the range launches below
         int range_end = nchunks;
         if (h->share_one_launch != 0) {
            const int cfirst = (next_mine ? (p + 1) / CH : p / CH) + (next_mine ? 1 : 0);
            int own = 0;
            for (int c = cfirst; c < nchunks; ++c) own += ((c + sb.shift) % G == me) ? 1 : 0;
            if (own > 0) {
               const int J0 = p * NBO, nf = sb.count_k_greater(J0);
               UpdateRange R;
               R.abs_lo = cfirst * CH * NBO;
               R.abs_hi = nchunks * CH * NBO;
               R.cyc_cols = CH * NBO; R.cyc_g = G; R.cyc_me = me; R.cyc_shift = sb.shift; R.cyc_nchunks = own;
               L += launch_update(h->F, sub_list(sb), nf, sb.pmaxf[nf - 1], 1, 0, J0, Wbuf[p & 1], sb.big_tiles, st, R);
            }
            range_end = std::min(nchunks, cfirst);
         }
         for (int c = p / CH; c < range_end; ++c) {
            if ((c + sb.shift) % G != me) continue;
            const int lo = std::max(c * CH, p) * NBO, hi = (c + 1) * CH * NBO;   // the kernel starts at the end of the K range
            if (next_mine && lo <= (p + 1) * NBO && (p + 2) * NBO <= hi) {
               if (stragglers) update(sb, p, lo, (p + 1) * NBO);
               update(sb, p, (p + 2) * NBO, hi);
            } else if (lo >= (p + 1) * NBO || stragglers || hi > (p + 1) * NBO) {
               update(sb, p, lo, hi);
            }
         }
      }
      // the next step needs panel p + 1 of every sub-list: a rank that factored all it needs itself (one sub-list) goes on as
      // soon as its panel chain is through (the broadcast goes on behind), the others when the broadcast has arrived
      if (next) {
         prof_begin(0.0, 3);
         const bool all_mine = chained && b.subs.size() == 1;
         if (chained) CK(cudaStreamWaitEvent(st, h->ev_sh_panel, 0));
         if (!all_mine) CK(cudaStreamWaitEvent(st, h->ev_sh_bcast[(p + 1) & 1], 0));
         prof_end(h->profiled.size() - 1);
      }
   }
   for (const Batch& sb : b.subs) L += launch_permute_rows(h->F, sub_list(sb), sb.count_k_greater(NBO), sb.max_k, h->W, st);
   // contribution blocks: every chunk to every rank
   if (b.has_cb) {
      link(h->ev_sh_main, st, cs);
      CommGroup group;
      for (const Batch& sb : b.subs) {
         const int nchunks = (sb.max_f + CH * NBO - 1) / (CH * NBO);
         for (int t = 0; t < sb.count; ++t) {
            const int s = h->batch_fronts_host[sb.begin + t];
            const int64_t k = P.k[s], r = P.f[s] - k;
            if (r <= 0) continue;
            auto col_start = [&](int64_t j) { return j * r - (j * (j - 1)) / 2; };   // packed lower triangle: first entry of column j
            for (int c = static_cast<int>(k / (CH * NBO)); c < nchunks; ++c) {
               const int64_t a = std::max<int64_t>(static_cast<int64_t>(c) * CH * NBO - k, 0), e = std::min<int64_t>(static_cast<int64_t>(c + 1) * CH * NBO - k, r);
               if (a >= e) continue;
               CKN(comm_bcast_bytes(h->comm, h->W.cbpool + P.cb_off[s] + col_start(a), static_cast<size_t>(col_start(e) - col_start(a)) * sizeof(double),
                                    (c + sb.shift) % G, cs, e__));
            }
         }
      }
      group.end();
      link(h->ev_sh_comm, cs, st);
   }
   prof_end(prof_batch);
}

void upload_plan(b200kkt_handle* h) {
   const Plan& P = h->plan;
   const auto t0 = std::chrono::steady_clock::now();
   std::vector<int> batch_list;
   std::vector<int64_t> woff;
   std::vector<int64_t> xoff;
   int64_t wpool_size = 0, xinv_size = 0;
   build_batches(h, batch_list, woff, wpool_size, xoff, xinv_size);

   h->F.off = dev_upload(h, P.front_off);
   h->F.k = dev_upload(h, P.k);
   h->F.f = dev_upload(h, P.f);
   h->F.ld = dev_upload(h, P.ld);
   h->F.cboff = dev_upload(h, P.cb_off);
   h->F.ldc = dev_upload(h, P.ldc);
   {
      std::vector<int32_t> first(P.sn_first.begin(), P.sn_first.end() - 1);
      h->F.first = dev_upload(h, first);
   }
   h->F.rows_off = dev_upload(h, P.rows_off);
   h->F.rows = dev_upload(h, P.rows);
   h->F.rel_off = dev_upload(h, P.rel_off);
   h->F.relmap = dev_upload(h, P.relmap);
   h->F.child_ptr = dev_upload(h, P.child_ptr);
   h->F.child_idx = dev_upload(h, P.child_idx);
   h->F.woff = dev_upload(h, woff);
   h->F.xoff = dev_upload(h, xoff);
   h->woff_host = woff;
   h->xoff_host = xoff;
   h->d_perm = dev_upload(h, P.perm);
   h->d_batch_list = dev_upload(h, batch_list);
   h->batch_fronts_host = batch_list;
   {
      // solve schedule: per level small fronts first, then big fronts by descending pivot count
      std::vector<int> solve_list;
      solve_list.reserve(static_cast<size_t>(P.nsn));
      h->solve_levels.assign(static_cast<size_t>(P.nlevels), b200kkt_handle::SolveLevel());
      std::vector<int> big;
      for (int l = 0; l < P.nlevels; ++l) {
         b200kkt_handle::SolveLevel& SL = h->solve_levels[l];
         SL.small_begin = static_cast<int>(solve_list.size());
         big.clear();
         for (int t = P.level_ptr[l]; t < P.level_ptr[l + 1]; ++t) {
            const int s = P.level_fronts[t];
            if (P.owner[s] != h->rank) continue;
            const bool is_big = P.k[s] > NBS || static_cast<int64_t>(P.f[s]) * P.k[s] > 98304;
            if (is_big) big.push_back(s); else { solve_list.push_back(s); SL.small_max_f = std::max(SL.small_max_f, P.f[s]); }
         }
         SL.small_count = static_cast<int>(solve_list.size()) - SL.small_begin;
         std::stable_sort(big.begin(), big.end(), [&](int a, int b) { return P.k[a] > P.k[b]; });
         SL.big_begin = static_cast<int>(solve_list.size());
         SL.big_count = static_cast<int>(big.size());
         for (int s : big) {
            solve_list.push_back(s);
            SL.big_k.push_back(P.k[s]);
            SL.big_max_f = std::max(SL.big_max_f, P.f[s]);
            SL.big_max_k = std::max(SL.big_max_k, P.k[s]);
            SL.big_pmaxf.push_back(SL.big_max_f);
         }
      }
      h->d_solve_list = dev_upload(h, solve_list);
      {
         // gather lists of the big fronts (forward solve): per row of the front, where its children's boundary sums sit in cpool
         std::vector<int64_t> goff(static_cast<size_t>(P.nsn), -1);
         std::vector<int> gptr, gsrc, count;
         int64_t total_rows = 0;
         for (const auto& SL : h->solve_levels)
            for (int t = 0; t < SL.big_count; ++t) {
               const int s = solve_list[static_cast<size_t>(SL.big_begin + t)];
               goff[s] = total_rows;
               total_rows += P.f[s] + 1;
            }
         gptr.assign(static_cast<size_t>(total_rows) + 1, 0);
         // pass 1: counts; pass 2: fill, children in ascending order (child_idx is ascending)
         for (int pass = 0; pass < 2; ++pass) {
            if (pass == 1) {
               int64_t run = 0;
               for (int32_t s = 0; s < P.nsn; ++s) {
                  if (goff[s] < 0) continue;
                  for (int d = 0; d <= P.f[s]; ++d) { const int c = gptr[static_cast<size_t>(goff[s] + d)]; gptr[static_cast<size_t>(goff[s] + d)] = static_cast<int>(run); run += (d < P.f[s]) ? c : 0; }
               }
               gsrc.resize(static_cast<size_t>(run));
               count.assign(static_cast<size_t>(total_rows), 0);
            }
            for (int32_t s = 0; s < P.nsn; ++s) {
               if (goff[s] < 0) continue;
               for (int32_t ci = P.child_ptr[s]; ci < P.child_ptr[s + 1]; ++ci) {
                  const int32_t c = P.child_idx[ci];
                  const int64_t r0 = P.rel_off[c];
                  for (int32_t i = 0; i < P.f[c] - P.k[c]; ++i) {
                     const int64_t d = goff[s] + P.relmap[static_cast<size_t>(r0 + i)];
                     if (pass == 0) ++gptr[static_cast<size_t>(d)];
                     else gsrc[static_cast<size_t>(gptr[static_cast<size_t>(d)] + count[static_cast<size_t>(d)]++)] = static_cast<int>(r0 + i);
                  }
               }
            }
         }
         if (P.sum_r >= (int64_t(1) << 31)) h->gather_lists = 0;   // 32-bit offsets into cpool
         h->Gt.off = dev_upload(h, goff);
         h->Gt.ptr = dev_upload(h, gptr);
         h->Gt.src = dev_upload(h, gsrc);
      }
      int max_big = 1;
      for (const auto& SL : h->solve_levels) max_big = std::max(max_big, SL.big_count);
      h->d_partial = dev_alloc<double>(h, static_cast<size_t>(max_big) * bwd_partial_doubles_per_front());
   }

   h->M.ndst = static_cast<int64_t>(P.asm_dst.size());
   h->M.dst = dev_upload(h, P.asm_dst);
   h->M.seg = dev_upload(h, P.asm_seg);
   h->M.src = dev_upload(h, P.asm_src);
   h->M.row = dev_upload(h, P.asm_row);
   h->M.col = dev_upload(h, P.asm_col);
   {
      // tiles of the pool for the fused zero-fill + assembly; asm_dst ascends
      h->M.pool_doubles = P.pool_size;
      h->M.ntiles = (P.pool_size + ASM_TILE - 1) / ASM_TILE;
      std::vector<int> tile_ptr(static_cast<size_t>(h->M.ntiles) + 1, 0);
      size_t d = 0;
      for (int64_t t = 0; t < h->M.ntiles; ++t) {
         tile_ptr[static_cast<size_t>(t)] = static_cast<int>(d);
         const int64_t end = (t + 1) * ASM_TILE;
         while (d < P.asm_dst.size() && P.asm_dst[d] < end) ++d;
      }
      tile_ptr[static_cast<size_t>(h->M.ntiles)] = static_cast<int>(P.asm_dst.size());
      h->M.tile_ptr = dev_upload(h, tile_ptr);
   }
   h->C.n = P.N;
   h->C.ptr = dev_upload(h, P.csr_ptr);
   h->C.col = dev_upload(h, P.csr_col);
   h->C.slot = dev_upload(h, P.csr_slot);

   const size_t n = static_cast<size_t>(P.N);
   h->W.lpool = dev_alloc<double>(h, static_cast<size_t>(P.pool_size));
   h->W.cbpool = dev_alloc<double>(h, static_cast<size_t>(P.cb_pool_size));
   h->wpool_size = (wpool_size + 1) & ~int64_t(1);
   h->W.wpool = dev_alloc<double>(h, static_cast<size_t>(4 * h->wpool_size));
   h->W.xinv = dev_alloc<double>(h, static_cast<size_t>(std::max<int64_t>(xinv_size, 1)));
   h->xinv_size = xinv_size;
   h->W.dd = dev_alloc<double>(h, n + 1);
   h->W.de = dev_alloc<double>(h, n + 1);
   h->W.ptype = dev_alloc<int>(h, n + 1);
   h->W.lperm = dev_alloc<int>(h, n + 1);
   h->W.blkflag = dev_alloc<int>(h, n + 1);
   h->W.delay = dev_alloc<int>(h, n + 1);
   if (h->keep_values != nullptr) { h->d_values = h->keep_values; h->plan_allocs.push_back(h->keep_values); h->keep_values = nullptr; }
   else h->d_values = dev_alloc<double>(h, static_cast<size_t>(P.nnz));
   h->d_scale = dev_alloc<double>(h, n);
   h->d_scale_tmp = dev_alloc<double>(h, n);
   h->d_zg = dev_alloc<double>(h, n);
   h->d_z2 = dev_alloc<double>(h, n);
   h->d_xg = dev_alloc<double>(h, n);
   h->d_cpool = dev_alloc<double>(h, static_cast<size_t>(P.sum_r));
   h->d_b = dev_alloc<double>(h, n);
   h->d_x = dev_alloc<double>(h, n);
   h->d_r = dev_alloc<double>(h, n);
   h->d_counters = dev_alloc<int>(h, 8);
   h->W.ndelay = h->d_counters + 5;
   h->d_norms = dev_alloc<unsigned long long>(h, 8);
   h->d_scale_final = h->d_scale;
   build_cut_schedule(h);
   CK(cudaMemsetAsync(h->W.wpool, 0, static_cast<size_t>(std::max<int64_t>(4 * h->wpool_size, 1)) * sizeof(double), h->stream));
   CK(cudaStreamSynchronize(h->stream));
   h->stats.t_upload = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

void fill_plan_stats(b200kkt_handle* h) {
   const Plan& P = h->plan;
   b200kkt_stats& s = h->stats;
   const int64_t keep_launches = s.kernel_launches;
   s = b200kkt_stats{};
   s.kernel_launches = keep_launches;
   s.n = P.N; s.nnz = P.nnz; s.reg_size = P.reg_size;
   s.num_fronts = P.nsn; s.num_levels = P.nlevels; s.max_front = P.max_f; s.max_pivots = P.max_k;
   s.num_pairs = P.num_pairs;
   s.nnz_factor = P.nnz_L; s.pool_doubles = P.pool_size;
   s.cb_arena_doubles = P.cb_pool_size; s.pool_doubles_global = P.pool_size_global; s.cb_arena_doubles_global = P.cb_pool_size_global;
   s.assembly_distinct = static_cast<int64_t>(P.asm_dst.size());
   s.flops_factor = P.flops; s.flops_tensor = P.flops_tc;
   s.t_ordering = P.t_order; s.t_symbolic = P.t_symbolic; s.t_maps = P.t_maps;
   s.num_parts = P.num_parts; s.part_rank = h->rank; s.num_cut_fronts = static_cast<int64_t>(P.cut_edges.size());
   s.flops_own = P.part_flops.empty() ? P.flops : P.part_flops[static_cast<size_t>(h->rank)];
   s.flops_critical = P.critical_flops;
}

// ---- numeric factorization: enqueue everything on the handle's stream -------------------------
void enqueue_factor(b200kkt_handle* h) {
   const Plan& P = h->plan;
   cudaStream_t st = h->stream;
   int64_t& L = h->launches;
   h->W.delay_tol = std::max(h->delay_tol, h->null_pivot_tol);   // what is null inside the block is null outside it too
   h->W.no_pivot = h->positive_definite_target != 0 ? 1 : 0;
   if (h->W.no_pivot) h->W.delay_tol = 1e300;                    // nothing is postponed: a failed pivot is the verdict "not positive definite"
   CK(cudaMemsetAsync(h->d_counters, 0, 8 * sizeof(int), st));
   CK(cudaMemsetAsync(h->W.delay, 0, static_cast<size_t>(P.N) * sizeof(int), st));
   // K1' equilibration
   L += launch_fill(h->d_scale, P.N, 1.0, st);
   double* cur = h->d_scale;
   double* nxt = h->d_scale_tmp;
   for (int sw = 0; sw < h->scaling_sweeps; ++sw) {
      L += launch_scale_sweep(h->C, h->d_values, cur, nxt, st);
      std::swap(cur, nxt);
   }
   h->d_scale_final = cur;
   CK(cudaMemsetAsync(h->d_norms, 0, 4 * sizeof(unsigned long long), st));
   L += launch_matrix_inf_norm(h->C, h->d_values, h->d_norms + 0, st);
   // K1 assembly: one pass writes the whole factor pool, zeros and assembled entries together (option fused_assembly = 0: memset
   // + scatter, the round-1 path)
   if (!h->fused_assembly) CK(cudaMemsetAsync(h->W.lpool, 0, static_cast<size_t>(P.pool_size) * sizeof(double), st));

   const bool prof = h->profile_updates != 0 && !h->use_graph;
   auto prof_begin = [&](double flops, int kind) {
      b200kkt_handle::ProfiledLaunch p{nullptr, nullptr, flops, kind};
      CK(cudaEventCreate(&p.a));
      CK(cudaEventCreate(&p.b));
      CK(cudaEventRecord(p.a, st));
      h->profiled.push_back(p);
   };
   auto prof_end = [&] { CK(cudaEventRecord(h->profiled.back().b, st)); };
   if (prof) prof_begin(0.0, 1);
   L += h->fused_assembly ? launch_assemble_fill(h->M, h->d_values, cur, h->W.lpool, st) : launch_assemble(h->M, h->d_values, cur, h->W.lpool, st);
   if (prof) prof_end();
   // K2-K5 level by level; the batches of one level are independent and go round-robin over the
   // handle's streams (fork/join with events, which stream capture turns into graph edges)
   const bool multi = h->multi_stream != 0 && !prof;
   for (int l = 0; l < P.nlevels; ++l) {
      // contribution blocks of cut fronts that a front of this level assembles: wait until they have arrived
      if (h->partitioned() && h->wait_level[l] >= 0) {
         if (prof) prof_begin(0.0, 4);   // kind 4: main stream waiting for contribution blocks from other ranks
         CK(cudaStreamWaitEvent(h->stream, h->ev_comm[h->wait_level[l]], 0));
         if (prof) prof_end();
      }
      const int nb = h->level_batch_ptr[l + 1] - h->level_batch_ptr[l];
      const bool level_shared = nb > 0 && h->batches[h->level_batch_ptr[l]].shared;   // shared batches come first; they own the streams
      const int nside = (multi && !level_shared) ? std::min(nb - 1, static_cast<int>(b200kkt_handle::NUM_AUX)) : 0;
      if (nside > 0) {
         CK(cudaEventRecord(h->ev_fork, h->stream));
         for (int a = 0; a < nside; ++a) CK(cudaStreamWaitEvent(h->aux[a], h->ev_fork, 0));
      }
      for (int bi = h->level_batch_ptr[l]; bi < h->level_batch_ptr[l + 1]; ++bi) {
         const Batch& b = h->batches[bi];
         const int* list = h->d_batch_list + b.begin;
         const int lane = (bi - h->level_batch_ptr[l]) % (nside + 1);
         cudaStream_t st = (lane == 0) ? h->stream : h->aux[lane - 1];
         if (h->fused_leaves != 0 && !b.shared && b.max_k <= NBI && b.max_r <= leaf_front_max_boundary() && h->debug_skip == 0 &&
             (!b.has_children || h->fused_leaves > 1)) {
            // K2: the whole front by one CTA, contribution block included; fronts with children are extend-added first
            if (b.has_children) L += launch_prepare(h->F, list, b.count, b.max_f, h->W, st);
            L += launch_leaf_fronts(h->F, list, b.count, b.max_r, h->W, h->null_pivot_tol, b.has_children, st);
            continue;
         }
         if ((b.has_cb || b.has_children) && !(h->debug_skip & 4)) L += launch_prepare(h->F, list, b.count, b.max_f, h->W, st);
         // W panels: panel p of a front lives in buffer p & 3 (two panels of a pair + the pair factored ahead)
         FactorArrays Wbuf[4] = {h->W, h->W, h->W, h->W};
         for (int t = 1; t < 4; ++t) Wbuf[t].wpool += t * h->wpool_size;
         if (b.shared) {
            enqueue_shared_batch(h, b, list, st);
            continue;
         }
         const bool ahead = h->lookahead != 0 && !prof && b.max_k > NBO && b.max_f >= h->lookahead_min_front;
         cudaStream_t pst = h->panel_stream[lane];
         const bool prof_this = prof && b.max_f >= 1024;
         auto trailing_flops = [&](int Jbeg, int kpanels, int nf) {
            double fl = 0.0;
            for (int t = 0; t < nf; ++t) {
               const int s_id = h->batch_fronts_host[b.begin + t];
               if (kpanels == 2 && P.k[s_id] <= Jbeg + NBO) continue;
               const double J1 = std::min(Jbeg + kpanels * NBO, P.k[s_id]), rem = P.f[s_id] - J1;
               if (rem > 0) fl += 2.0 * (J1 - Jbeg) * rem * (rem + 1.0) * 0.5;
            }
            return fl;
         };
         auto panel = [&](int J0, int nf, const FactorArrays& Wp, cudaStream_t s_) {
            if (h->debug_skip & 2) return;
            if (!(h->debug_skip & 16)) L += launch_diagpanel(h->F, list, nf, b.max_k, J0, Wp, h->null_pivot_tol, s_);
            if (!(h->debug_skip & 8)) L += launch_trsmpanel(h->F, list, nf, b.pmaxf[nf - 1], b.max_k, J0, Wp, s_);
         };
         auto update = [&](int J0, int nf, const FactorArrays& Wp, cudaStream_t s_, const UpdateRange& R) {
            if (h->debug_skip & 1) return;
            L += launch_update(h->F, list, nf, b.pmaxf[nf - 1], 1, 0, J0, Wp, b.big_tiles, s_, R);
         };
         if (h->pair_panels != 0 && b.max_k > NBO && b.max_f >= h->pair_min_front) {
            // Outer panels are taken in pairs: panel p0, the update of panel p1's columns (K = 256, narrow), panel p1, then ONE
            // trailing update with K <= 512 - half the read-modify-write traffic and half the per-tile fill/epilogue overhead of
            // two K = 256 updates.  With look-ahead the next pair's chain runs on the panel stream under the bulk of that update.
            auto chain = [&](int q, cudaStream_t s_) {
               const int p0 = 2 * q, J0 = p0 * NBO, J0b = J0 + NBO;
               const int nf0 = b.count_k_greater(J0);
               panel(J0, nf0, Wbuf[p0 & 3], s_);
               if (J0b < b.max_k) {
                  update(J0, nf0, Wbuf[p0 & 3], s_, UpdateRange{1, 0, 1, nullptr, 0, 1});   // panel p1's pivot columns only
                  panel(J0b, b.count_k_greater(J0b), Wbuf[(p0 + 1) & 3], s_);
               }
            };
            chain(0, st);
            for (int q = 0; 2 * q * NBO < b.max_k; ++q) {
               const int p0 = 2 * q, J0 = p0 * NBO, J0b = J0 + NBO;
               const int nf0 = b.count_k_greater(J0);
               if (J0b >= b.max_k) {
                  // single last panel
                  if (prof_this) prof_begin(trailing_flops(J0, 1, nf0), 0);
                  update(J0, nf0, Wbuf[p0 & 3], st, UpdateRange{1, 0, -1, nullptr, 0});
                  if (prof_this) prof_end();
                  break;
               }
               const int nf1 = b.count_k_greater(J0b);
               // fronts whose last panel was p0: their whole trailing matrix (the chain only touched pivot columns)
               if (nf0 > nf1) update(J0, nf0, Wbuf[p0 & 3], st, UpdateRange{1, 0, -1, nullptr, 3});
               const double* w2 = Wbuf[(p0 + 1) & 3].wpool;
               if (ahead && J0 + 2 * NBO < b.max_k) {
                  update(J0, nf1, Wbuf[p0 & 3], st, UpdateRange{2, 0, 2, w2, 1});
                  CK(cudaEventRecord(h->ev_next_cols[lane], st));
                  CK(cudaStreamWaitEvent(pst, h->ev_next_cols[lane], 0));
                  chain(q + 1, pst);
                  CK(cudaEventRecord(h->ev_panel[lane], pst));
                  update(J0, nf1, Wbuf[p0 & 3], st, UpdateRange{2, 2, -1, w2, 2});
                  CK(cudaStreamWaitEvent(st, h->ev_panel[lane], 0));
               } else {
                  if (prof_this) prof_begin(trailing_flops(J0, 2, nf1), 0);
                  update(J0, nf1, Wbuf[p0 & 3], st, UpdateRange{2, 0, -1, w2, 0});
                  if (prof_this) prof_end();
                  if (J0 + 2 * NBO < b.max_k) chain(q + 1, st);
               }
            }
         } else {
            int p = 0;
            for (int J0 = 0; J0 < b.max_k; J0 += NBO, ++p) {
               const int nfP = b.count_k_greater(J0);
               const FactorArrays& Wp = Wbuf[p & 1];
               if (!ahead || J0 == 0) panel(J0, nfP, Wp, st);
               if (prof_this) prof_begin(trailing_flops(J0, 1, nfP), 0);
               if (ahead && J0 + NBO < b.max_k) {
                  // look-ahead: update the next panel's columns, factor that panel on the panel stream while the rest of
                  // the trailing matrix is updated here
                  update(J0, nfP, Wp, st, UpdateRange{1, 0, 1, nullptr, 0});
                  CK(cudaEventRecord(h->ev_next_cols[lane], st));
                  CK(cudaStreamWaitEvent(pst, h->ev_next_cols[lane], 0));
                  panel(J0 + NBO, b.count_k_greater(J0 + NBO), Wbuf[(p + 1) & 1], pst);
                  CK(cudaEventRecord(h->ev_panel[lane], pst));
                  update(J0, nfP, Wp, st, UpdateRange{1, 1, -1, nullptr, 0});
                  CK(cudaStreamWaitEvent(st, h->ev_panel[lane], 0));
               } else {
                  update(J0, nfP, Wp, st, UpdateRange{1, 0, -1, nullptr, 0});
               }
               if (prof_this) prof_end();
            }
         }
         L += launch_permute_rows(h->F, list, b.count_k_greater(NBO), b.max_k, h->W, st);
      }
      for (int a = 0; a < nside; ++a) {
         CK(cudaEventRecord(h->ev_join[a], h->aux[a]));
         CK(cudaStreamWaitEvent(h->stream, h->ev_join[a], 0));
      }
      // K9: the cut fronts of this level hand their contribution blocks to the parent's rank over NVLink, on the
      // communication stream, while this rank goes on with its next levels
      if (h->partitioned() && h->ev_comm[l] != nullptr) {
         CK(cudaEventRecord(h->ev_level_done, h->stream));
         CK(cudaStreamWaitEvent(h->comm_stream, h->ev_level_done, 0));
         enqueue_cut_transfers(h, l, 0, h->W.cbpool, h->comm_stream);
         CK(cudaEventRecord(h->ev_comm[l], h->comm_stream));
      }
   }
   if (h->partitioned()) {
      // nothing of this factorization may still be in flight when the caller starts the next one
      CK(cudaEventRecord(h->ev_level_done, h->comm_stream));
      CK(cudaStreamWaitEvent(h->stream, h->ev_level_done, 0));
   }
   // K6 inertia (partitioned: this rank's pivots, then summed over ranks); counter 5 = columns flagged for a delay
   L += launch_inertia(P.N, h->W, h->d_counters, st, h->partitioned() ? h->d_col_owner : nullptr, h->rank);
   if (h->partitioned()) CKN(comm_allreduce_sum_i32(h->comm, h->d_counters, 8, st, e__));
   CK(cudaMemcpyAsync(h->h_counters, h->d_counters, 8 * sizeof(int), cudaMemcpyDeviceToHost, st));
   CK(cudaMemcpyAsync(h->h_norms + 7, h->d_norms + 0, sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));   // |K|_inf: inf <=> NaN/Inf in the values
}

}  // namespace

namespace {

double bits_to_double(unsigned long long b) {
   double d;
   std::memcpy(&d, &b, sizeof(d));
   return d;
}

// run the numeric factorization on values already resident in h->d_values
void run_factor_once(b200kkt_handle* h) {
   cudaStream_t st = h->stream;
   CK(cudaEventRecord(h->ev0, st));
   // partitioned: NCCL calls between the levels are launched directly unless partition_graph asks for their capture too
   if (h->use_graph && (!h->partitioned() || h->partition_graph != 0)) {
      if (h->factor_graph == nullptr) {
         const int64_t before = h->launches;
         cudaGraph_t graph = nullptr;
         CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
         try {
            enqueue_factor(h);
         } catch (...) {
            cudaStreamEndCapture(st, &graph);
            if (graph != nullptr) cudaGraphDestroy(graph);
            throw;
         }
         CK(cudaStreamEndCapture(st, &graph));
         const cudaError_t e = cudaGraphInstantiate(&h->factor_graph, graph, 0);
         cudaGraphDestroy(graph);
         if (e != cudaSuccess) throw CudaError{e, "cudaGraphInstantiate"};
         h->graph_launches = h->launches - before;
         h->launches = before;
      }
      CK(cudaGraphLaunch(h->factor_graph, st));
      h->launches += h->graph_launches;
   } else {
      enqueue_factor(h);
   }
   CK(cudaEventRecord(h->ev1, st));
   CK(cudaStreamSynchronize(st));
   CK(cudaGetLastError());
   float ms = 0.f;
   CK(cudaEventElapsedTime(&ms, h->ev0, h->ev1));
   h->stats.t_factor_last = ms * 1e-3;
   if (!h->profiled.empty()) {
      double tu = 0.0, fu = 0.0, ta = 0.0, tsh = 0.0, fsh = 0.0, tstall = 0.0, twait = 0.0;
      h->stats.t_before_shared_profiled = 0.0;
      int64_t nu = 0;
      for (auto& p : h->profiled) {
         float pm = 0.f;
         cudaEventElapsedTime(&pm, p.a, p.b);
         if (p.kind == 0) { tu += pm * 1e-3; fu += p.flops; ++nu; }
         else if (p.kind == 1) ta += pm * 1e-3;
         else if (p.kind == 2) {
            if (tsh == 0.0) { float t0 = 0.f; cudaEventElapsedTime(&t0, h->ev0, p.a); h->stats.t_before_shared_profiled = t0 * 1e-3; }
            tsh += pm * 1e-3; fsh += p.flops;
         }
         else if (p.kind == 3) tstall += pm * 1e-3;
         else twait += pm * 1e-3;
         cudaEventDestroy(p.a);
         cudaEventDestroy(p.b);
      }
      h->profiled.clear();
      h->stats.t_update_profiled = tu; h->stats.flops_update_profiled = fu; h->stats.n_update_profiled = nu;
      h->stats.t_assemble_profiled = ta;
      h->stats.t_transfer_wait_profiled = twait;
      h->stats.t_shared_profiled = tsh; h->stats.flops_shared_profiled = fsh; h->stats.t_shared_stall_profiled = tstall;
   }
   h->inertia[0] = static_cast<uint64_t>(h->h_counters[0]);
   h->inertia[1] = static_cast<uint64_t>(h->h_counters[1]);
   h->inertia[2] = static_cast<uint64_t>(h->h_counters[2]);
   h->factorized = true;
   // NaN/Inf in the values ends up in the D blocks: an error (MUMPS: INFOG(1) < 0), never a sign of an eigenvalue
   if (!std::isfinite(bits_to_double(h->h_norms[7]))) h->h_counters[6] = 1;
   if (h->h_counters[6] > 0) { h->factorized = false; throw std::runtime_error("factorize: non-finite pivots (the matrix values hold NaN/Inf)"); }
}

// Move the given columns (positions in the current numbering) into ancestor fronts and rebuild plan, schedules and device
// arrays; the resident values stay.  How far a column travels grows with the number of times it has been delayed before:
// parent, then 4 levels, then the root of its tree (where every row is fully summed and the verdict is final).
void replan_with_delays(b200kkt_handle* h, const std::vector<int32_t>& cols) {
   Plan& P = h->plan;
   if (h->delay_count.size() != static_cast<size_t>(P.N)) h->delay_count.assign(static_cast<size_t>(P.N), 0);
   std::vector<int32_t> up(cols.size());
   for (size_t i = 0; i < cols.size(); ++i) {
      unsigned char& c = h->delay_count[static_cast<size_t>(P.perm[cols[i]])];
      up[i] = (c == 0) ? 1 : ((c == 1) ? 4 : (1 << 30));
      if (c < 255) ++c;
   }
   const double t_order = P.t_order;
   const std::string ordering_used = P.ordering_used;
   const int64_t num_pairs = P.num_pairs;
   delay_columns(P, cols, up);
   P.t_order = t_order; P.ordering_used = ordering_used; P.num_pairs = num_pairs;
   P.share_min_pivots = h->share_min_pivots;
   P.share_min_flops = h->share_min_flops;
   P.partition_slack = h->partition_slack;
   map_subtrees(P, h->num_gpus, h->share_top_fronts != 0);
   if (h->partitioned() && h->partition_memory) localize_storage(P, h->rank);
   const bool resident = h->values_resident;
   if (h->has_device) {
      CK(cudaStreamSynchronize(h->stream));
      // keep the value array: take it out of the list free_plan_device walks
      for (auto it = h->plan_allocs.begin(); it != h->plan_allocs.end(); ++it)
         if (*it == h->d_values) { h->plan_allocs.erase(it); break; }
      h->keep_values = h->d_values;
      free_plan_device(h);
   }
   const double t_upload = h->stats.t_upload;
   fill_plan_stats(h);
   if (h->has_device) {
      upload_plan(h);
      h->stats.t_upload += t_upload;
   }
   h->analysed = true;
   h->values_resident = resident;
   h->delayed_total += static_cast<int64_t>(cols.size());
   ++h->replans_total;
}

void run_factor(b200kkt_handle* h) {
   h->replans_last = 0;
   const double limit = h->positive_definite_target != 0 ? 0.0 : h->growth_limit > 0.0 ? h->growth_limit : ((h->growth_armed && h->adaptive_growth != 0) ? h->adaptive_growth_limit : 0.0);
   h->growth_armed = false;
   if (limit != h->growth_baked && h->factor_graph != nullptr) { cudaGraphExecDestroy(h->factor_graph); h->factor_graph = nullptr; }
   h->growth_baked = limit;
   h->W.growth_limit = limit;
   for (int attempt = 0;; ++attempt) {
      run_factor_once(h);
      const int flagged = h->h_counters[5];
      if (flagged <= 0 || attempt >= h->max_replans) break;
      // columns that could not be eliminated in their fronts: collect, re-plan, factorize again
      const Plan& P = h->plan;
      cudaStream_t st = h->stream;
      if (h->partitioned()) CKN(comm_allreduce_sum_i32(h->comm, h->W.delay, static_cast<size_t>(P.N), st, e__));
      std::vector<int> flags(static_cast<size_t>(P.N));
      CK(cudaMemcpyAsync(flags.data(), h->W.delay, flags.size() * sizeof(int), cudaMemcpyDeviceToHost, st));
      CK(cudaStreamSynchronize(st));
      std::vector<int32_t> cols;
      for (int64_t q = 0; q < P.N; ++q)
         if (flags[static_cast<size_t>(q)] != 0 && P.parent[P.sn_of[q]] >= 0) cols.push_back(static_cast<int32_t>(q));
      if (cols.empty()) break;   // only root fronts: the null pivots are final
      replan_with_delays(h, cols);
      ++h->replans_last;
   }
}

}  // namespace

namespace {

// number of entries > j in a descending list
int count_greater(const std::vector<int>& k, int j) {
   int lo = 0, hi = static_cast<int>(k.size());
   while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (k[mid] > j) lo = mid + 1; else hi = mid;
   }
   return lo;
}

// x = K^{-1} b with iterative refinement; b and x are device pointers (x may alias nothing else)
void run_solve(b200kkt_handle* h, const double* d_b, double* d_x, double* backward_error) {
   const Plan& P = h->plan;
   cudaStream_t st = h->stream;
   int64_t& L = h->launches;
   const size_t nbytes = static_cast<size_t>(P.N) * sizeof(double);
   CK(cudaEventRecord(h->ev0, st));
   CK(cudaMemsetAsync(d_x, 0, nbytes, st));
   CK(cudaMemcpyAsync(h->d_r, d_b, nbytes, cudaMemcpyDeviceToDevice, st));
   CK(cudaMemsetAsync(h->d_norms + 1, 0, 3 * sizeof(unsigned long long), st));
   L += launch_absmax(d_b, P.N, h->d_norms + 3, st);
   // one sweep: permute+scale in, forward, block-diagonal, backward, scale+permute out (accumulating into x)
   auto enqueue_sweep = [&] {
      const bool parts = h->partitioned();
      L += launch_solve_gather(P.N, h->d_perm, h->d_scale_final, h->d_r, h->d_zg, st);
      for (int l = 0; l < P.nlevels; ++l) {
         const b200kkt_handle::SolveLevel& SL = h->solve_levels[l];
         L += launch_forward_level(h->F, h->d_solve_list + SL.small_begin, SL.small_count, h->W, h->d_zg, h->d_z2, h->d_cpool, st,
                                   (h->small_solve_ctas != 0 && SL.small_max_f <= 96) ? 64 : 256);
         if (SL.big_count > 0) {
            const int* list = h->d_solve_list + SL.big_begin;
            if (h->gather_lists != 0) L += launch_fwd_big_gather_list(h->F, h->Gt, list, SL.big_count, SL.big_max_f, h->d_zg, h->d_cpool, st);
            else L += launch_fwd_big_gather(h->F, list, SL.big_count, SL.big_max_f, h->W, h->d_zg, h->d_z2, h->d_cpool, st);
            L += launch_fwd_big_permute(h->F, list, SL.big_count, SL.big_max_k, h->W, h->d_zg, h->d_z2, st);
            for (int jp = 0; jp < SL.big_max_k; jp += NBS) {
               const int nf = count_greater(SL.big_k, jp);
               L += launch_fwd_big_diag(h->F, list, nf, jp, h->W, h->d_z2, st);
               L += launch_fwd_big_gemv(h->F, list, nf, SL.big_pmaxf[nf - 1], jp, h->W, h->d_z2, h->d_cpool, st);
            }
         }
         // K9: boundary partial sums of this level's cut fronts go to the parent's rank
         if (parts) enqueue_cut_transfers(h, l, 1, h->d_cpool, st);
      }
      for (int l = P.nlevels - 1; l >= 0; --l) {
         const b200kkt_handle::SolveLevel& SL = h->solve_levels[l];
         if (parts && !h->cut_levels[l].fronts.empty()) {
            // K9: the parent's rank ships the solution on the cut front's boundary rows
            const b200kkt_handle::CutLevel& CL = h->cut_levels[l];
            L += launch_boundary_pack(h->F, h->d_cut_list + CL.in_begin, CL.in_count, h->d_xg, h->d_cpool, st);
            enqueue_cut_transfers(h, l, 2, h->d_cpool, st);
            L += launch_boundary_unpack(h->F, h->d_cut_list + CL.out_begin, CL.out_count, h->d_cpool, h->d_xg, st);
         }
         L += launch_backward_level(h->F, h->d_solve_list + SL.small_begin, SL.small_count, h->W, h->d_z2, h->d_xg, h->d_cpool, st,
                                    (h->small_solve_ctas != 0 && SL.small_max_f <= 96) ? 64 : 256);
         if (SL.big_count > 0) {
            const int* list = h->d_solve_list + SL.big_begin;
            L += launch_bwd_big_begin(h->F, list, SL.big_count, SL.big_max_f, h->W, h->d_z2, h->d_xg, h->d_cpool, st);
            const int last = ((SL.big_max_k - 1) / NBS) * NBS;
            for (int jp = last; jp >= 0; jp -= NBS) {
               const int nf = count_greater(SL.big_k, jp);
               L += launch_bwd_big_gemv(h->F, list, nf, jp, h->W, h->d_z2, h->d_cpool, h->d_partial, st);
               L += launch_bwd_big_diag(h->F, list, nf, jp, h->W, h->d_z2, h->d_partial, st);
            }
            L += launch_bwd_big_scatter(h->F, list, SL.big_count, SL.big_max_k, h->W, h->d_z2, h->d_xg, st);
         }
      }
      if (parts) {
         // every rank holds the solution of its own pivots: zero the rest, sum over ranks, and all ranks carry the same x
         L += launch_keep_owned(P.N, h->d_col_owner, h->rank, h->d_xg, st);
         CKN(comm_allreduce_sum_f64(h->comm, h->d_xg, static_cast<size_t>(P.N), st, e__));
      }
      L += launch_solve_scatter_add(P.N, h->d_perm, h->d_scale_final, h->d_xg, d_x, st);
   };
   double berr = 0.0, omega = 0.0, prev = 0.0, prev_omega = 0.0, first_berr = 0.0;
   static const bool trace = std::getenv("B200KKT_TRACE_REFINE") != nullptr;
   int steps = 0;
   bool finite = true;
   for (int it = 0; it <= h->refine_max_steps; ++it) {
      if (h->use_graph && (!h->partitioned() || h->partition_graph != 0)) {
         // the sweep is a long chain of short kernels: replayed as a CUDA graph (re-captured when the output vector moves)
         if (h->solve_graph == nullptr || h->solve_graph_x != d_x || h->solve_graph_scale != h->d_scale_final) {
            if (h->solve_graph != nullptr) { cudaGraphExecDestroy(h->solve_graph); h->solve_graph = nullptr; }
            const int64_t before = L;
            cudaGraph_t graph = nullptr;
            CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
            enqueue_sweep();
            CK(cudaStreamEndCapture(st, &graph));
            CK(cudaGraphInstantiate(&h->solve_graph, graph, 0));
            cudaGraphDestroy(graph);
            h->solve_graph_launches = L - before;
            L = before;
            h->solve_graph_x = d_x;
            h->solve_graph_scale = h->d_scale_final;
         }
         CK(cudaGraphLaunch(h->solve_graph, st));
         L += h->solve_graph_launches;
      } else {
         enqueue_sweep();
      }
      // K8 residual and backward error
      CK(cudaMemsetAsync(h->d_norms + 1, 0, 2 * sizeof(unsigned long long), st));
      CK(cudaMemsetAsync(h->d_norms + 4, 0, sizeof(unsigned long long), st));
      L += launch_absmax(d_x, P.N, h->d_norms + 2, st);
      L += launch_residual(h->C, h->d_values, d_x, d_b, h->d_r, h->d_norms + 2, h->d_norms + 4, st);
      L += launch_absmax(h->d_r, P.N, h->d_norms + 1, st);
      CK(cudaMemcpyAsync(h->h_norms, h->d_norms, 5 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
      CK(cudaStreamSynchronize(st));
      const double knorm = bits_to_double(h->h_norms[0]), rnorm = bits_to_double(h->h_norms[1]);
      const double xnorm = bits_to_double(h->h_norms[2]), bnorm = bits_to_double(h->h_norms[3]);
      const double denom = knorm * xnorm + bnorm;
      berr = (denom > 0.0) ? rnorm / denom : 0.0;
      omega = bits_to_double(h->h_norms[4]);
      steps = it;
      // both measures drive the refinement: the normwise one (north star) and the componentwise one (the measure MUMPS reports)
      if (!std::isfinite(berr) || !std::isfinite(omega)) { finite = false; break; }
      if (it == 0) first_berr = berr;
      if (trace) std::fprintf(stderr, "[b200kkt] refine step %d: normwise %.3e componentwise %.3e\n", it, berr, omega);
      const bool need_n = berr > h->refine_tol, need_c = omega > h->refine_tol_componentwise;
      if (!need_n && !need_c) break;                       // converged (also leaves on a zero residual)
      // stagnation: no measure that still has to come down has halved
      if (it > 0 && (!need_n || berr > 0.5 * prev) && (!need_c || omega > 0.5 * prev_omega)) break;
      prev = berr;
      prev_omega = omega;
   }
   CK(cudaEventRecord(h->ev1, st));
   CK(cudaStreamSynchronize(st));
   CK(cudaGetLastError());
   float ms = 0.f;
   CK(cudaEventElapsedTime(&ms, h->ev0, h->ev1));
   h->stats.t_solve_last = ms * 1e-3;
   h->stats.refine_steps_last = steps;
   h->stats.backward_error_last = berr;
   h->stats.backward_error_componentwise_last = omega;
   if (first_berr > h->adaptive_growth_trigger) h->growth_armed = true;   // see adaptive_growth
   if (backward_error != nullptr) *backward_error = berr;
   // a factorization or right-hand side with NaN/Inf in it is an error, not a converged solve (MUMPS: INFOG(1) < 0)
   if (!finite) throw std::runtime_error("solve: non-finite residual (the factorization or the right-hand side holds NaN/Inf)");
}

template <typename Fn>
int guarded(b200kkt_handle* h, Fn&& fn) {
   try {
      fn();
      return B200KKT_OK;
   } catch (const CudaError& e) {
      // a failed capture leaves the stream capturing: make sure it is terminated
      cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
      if (h->stream != nullptr && cudaStreamIsCapturing(h->stream, &cs) == cudaSuccess && cs != cudaStreamCaptureStatusNone) {
         cudaGraph_t g = nullptr;
         cudaStreamEndCapture(h->stream, &g);
         if (g != nullptr) cudaGraphDestroy(g);
      }
      cudaGetLastError();
      const bool oom = e.code == cudaErrorMemoryAllocation;
      return fail(h, oom ? B200KKT_ENOMEM : B200KKT_ECUDA, std::string("CUDA error: ") + cudaGetErrorString(e.code) + " in " + e.what);
   } catch (const std::bad_alloc&) {
      return fail(h, B200KKT_ENOMEM, "host allocation failed");
   } catch (const std::exception& e) {
      return fail(h, B200KKT_EINTERNAL, e.what());
   }
}

int need_device(const b200kkt_handle* h) {
   if (h == nullptr) return B200KKT_EINVAL;
   if (!h->has_device) return fail(h, B200KKT_ECUDA, "handle has no CUDA device: the numeric phase has no CPU path");
   return B200KKT_OK;
}

}  // namespace

// ================================================================================================
// C ABI
// ================================================================================================
extern "C" {

int b200kkt_create(b200kkt_handle** out, int device, int num_gpus) {
   if (out == nullptr) return B200KKT_EINVAL;
   *out = nullptr;
   b200kkt_handle* h = new (std::nothrow) b200kkt_handle();
   if (h == nullptr) return B200KKT_ENOMEM;
   h->device = device;
   h->num_gpus = num_gpus < 1 ? 1 : num_gpus;
   h->refine_tol = 1e-13;
   if (device >= 0) {
      const int rc = guarded(h, [&] {
         CK(cudaSetDevice(device));
         CK(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
         for (int a = 0; a < b200kkt_handle::NUM_AUX; ++a) {
            CK(cudaStreamCreateWithFlags(&h->aux[a], cudaStreamNonBlocking));
            CK(cudaEventCreateWithFlags(&h->ev_join[a], cudaEventDisableTiming));
         }
         CK(cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming));
         int prio_least = 0, prio_greatest = 0;
         CK(cudaDeviceGetStreamPriorityRange(&prio_least, &prio_greatest));
         for (int a = 0; a <= b200kkt_handle::NUM_AUX; ++a) {
            // the panel chain is the critical path: its CTAs go first whenever an SM frees up under a running update
            CK(cudaStreamCreateWithPriority(&h->panel_stream[a], cudaStreamNonBlocking, prio_greatest));
            CK(cudaEventCreateWithFlags(&h->ev_next_cols[a], cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&h->ev_panel[a], cudaEventDisableTiming));
         }
         CK(cudaEventCreate(&h->ev0));
         CK(cudaEventCreate(&h->ev1));
         if (h->num_gpus > 1) {
            CK(cudaStreamCreateWithPriority(&h->comm_stream, cudaStreamNonBlocking, prio_greatest));
            CK(cudaEventCreateWithFlags(&h->ev_level_done, cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&h->ev_sh_main, cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&h->ev_sh_panel, cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&h->ev_sh_comm, cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&h->ev_sh_bcast[0], cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&h->ev_sh_bcast[1], cudaEventDisableTiming));
         }
         CK(cudaMallocHost(reinterpret_cast<void**>(&h->h_counters), 8 * sizeof(int)));
         CK(cudaMallocHost(reinterpret_cast<void**>(&h->h_norms), 8 * sizeof(unsigned long long)));
         const int e = init_kernel_attributes();
         if (e != 0) throw CudaError{static_cast<cudaError_t>(e), "cudaFuncSetAttribute(k_update)"};
         h->has_device = true;
      });
      if (rc != B200KKT_OK) {
         // keep the handle so that the caller can read last_error, but it stays host-only
         h->has_device = false;
         *out = h;
         return rc;
      }
   }
   *out = h;
   return B200KKT_OK;
}

int b200kkt_destroy(b200kkt_handle* h) {
   if (h == nullptr) return B200KKT_EINVAL;
   if (h->has_device) {
      cudaSetDevice(h->device);
      if (h->stream != nullptr) cudaStreamSynchronize(h->stream);
      free_plan_device(h);
      b200kkt::comm_destroy(h->comm);
      if (h->comm_stream != nullptr) cudaStreamDestroy(h->comm_stream);
      if (h->ev_level_done != nullptr) cudaEventDestroy(h->ev_level_done);
      for (cudaEvent_t e : {h->ev_sh_main, h->ev_sh_panel, h->ev_sh_comm, h->ev_sh_bcast[0], h->ev_sh_bcast[1]}) if (e != nullptr) cudaEventDestroy(e);
      if (h->h_counters != nullptr) cudaFreeHost(h->h_counters);
      if (h->h_norms != nullptr) cudaFreeHost(h->h_norms);
      if (h->ev0 != nullptr) cudaEventDestroy(h->ev0);
      if (h->ev1 != nullptr) cudaEventDestroy(h->ev1);
      for (int a = 0; a < b200kkt_handle::NUM_AUX; ++a) {
         if (h->ev_join[a] != nullptr) cudaEventDestroy(h->ev_join[a]);
         if (h->aux[a] != nullptr) cudaStreamDestroy(h->aux[a]);
      }
      if (h->ev_fork != nullptr) cudaEventDestroy(h->ev_fork);
      for (int a = 0; a <= b200kkt_handle::NUM_AUX; ++a) {
         if (h->ev_next_cols[a] != nullptr) cudaEventDestroy(h->ev_next_cols[a]);
         if (h->ev_panel[a] != nullptr) cudaEventDestroy(h->ev_panel[a]);
         if (h->panel_stream[a] != nullptr) cudaStreamDestroy(h->panel_stream[a]);
      }
      if (h->stream != nullptr) cudaStreamDestroy(h->stream);
   }
   delete h;
   return B200KKT_OK;
}

int b200kkt_comm_unique_id(char id[128]) {
   if (id == nullptr) return B200KKT_EINVAL;
   std::string err;
   return b200kkt::comm_unique_id(id, err) == 0 ? B200KKT_OK : B200KKT_ECUDA;
}

int b200kkt_comm_init(b200kkt_handle* h, int rank, int world, const char id[128]) {
   if (int rc = need_device(h)) return rc;
   if (id == nullptr || world != h->num_gpus || rank < 0 || rank >= world)
      return fail(h, B200KKT_EINVAL, "comm_init: world must equal the handle's num_gpus and 0 <= rank < world");
   if (h->analysed) return fail(h, B200KKT_EINVAL, "comm_init after analyse");
   return guarded(h, [&] {
      CK(cudaSetDevice(h->device));
      std::string err;
      if (b200kkt::comm_init(h->comm, rank, world, id, err) != 0) throw std::runtime_error(err);
      h->rank = rank;
   });
}

int b200kkt_plan_partition(const b200kkt_handle* h, int32_t* owner, int8_t* shared, int64_t* num_cut, int32_t* cut_fronts, double* part_flops,
                           double* critical_flops) {
   if (h == nullptr) return B200KKT_EINVAL;
   if (!h->analysed) return fail(h, B200KKT_EINVAL, "plan getter before analyse");
   const Plan& P = h->plan;
   if (owner != nullptr) std::copy(P.owner.begin(), P.owner.end(), owner);
   if (shared != nullptr) std::copy(P.shared.begin(), P.shared.end(), shared);
   if (num_cut != nullptr) *num_cut = static_cast<int64_t>(P.cut_edges.size());
   if (cut_fronts != nullptr) std::copy(P.cut_edges.begin(), P.cut_edges.end(), cut_fronts);
   if (part_flops != nullptr) std::copy(P.part_flops.begin(), P.part_flops.end(), part_flops);
   if (critical_flops != nullptr) *critical_flops = P.critical_flops;
   return B200KKT_OK;
}

int b200kkt_set_option(b200kkt_handle* h, const char* name, double value) {
   if (h == nullptr || name == nullptr) return B200KKT_EINVAL;
   const std::string n(name);
   if (n == "ordering") h->aopt.ordering = static_cast<int>(value);
   else if (n == "kkt_pairing") h->aopt.kkt_pairing = static_cast<int>(value);
   else if (n == "relax_small") h->aopt.relax_small = static_cast<int>(value);
   else if (n == "relax_mid") h->aopt.relax_mid = static_cast<int>(value);
   else if (n == "relax_big") h->aopt.relax_big = static_cast<int>(value);
   else if (n == "relax_mid_zeros") h->aopt.relax_mid_zeros = value;
   else if (n == "relax_big_zeros") h->aopt.relax_big_zeros = value;
   else if (n == "relax_any_zeros") h->aopt.relax_any_zeros = value;
   else if (n == "relax_small_front") h->aopt.relax_small_front = static_cast<int>(value);
   else if (n == "relax_chain_zeros") h->aopt.relax_chain_zeros = value;
   else if (n == "relax_chain_max_k") h->aopt.relax_chain_max_k = static_cast<int>(value);
   else if (n == "metis_nseps") h->aopt.metis_nseps = static_cast<int>(value);
   else if (n == "metis_niter") h->aopt.metis_niter = static_cast<int>(value);
   else if (n == "metis_ufactor") h->aopt.metis_ufactor = static_cast<int>(value);
   else if (n == "parallel_ordering_depth") h->aopt.parallel_ordering_depth = std::max(0, std::min(6, static_cast<int>(value)));
   else if (n == "parallel_ordering_min_nodes") h->aopt.parallel_ordering_min_nodes = static_cast<int64_t>(value);
   else if (n == "scaling_sweeps") h->scaling_sweeps = std::max(0, static_cast<int>(value));
   else if (n == "null_pivot_tol") h->null_pivot_tol = value;
   else if (n == "delay_tol") h->delay_tol = value;
   else if (n == "growth_limit") h->growth_limit = value;
   else if (n == "max_replans") h->max_replans = std::max(0, static_cast<int>(value));
   else if (n == "positive_definite_target") h->positive_definite_target = static_cast<int>(value);
   else if (n == "adaptive_growth") h->adaptive_growth = static_cast<int>(value);
   else if (n == "adaptive_growth_limit") h->adaptive_growth_limit = value;
   else if (n == "adaptive_growth_trigger") h->adaptive_growth_trigger = value;
   else if (n == "refine_tol_componentwise") h->refine_tol_componentwise = value;
   else if (n == "refine_max_steps") h->refine_max_steps = std::max(0, static_cast<int>(value));
   else if (n == "refine_tol") h->refine_tol = value;
   else if (n == "use_graph") h->use_graph = static_cast<int>(value);
   else if (n == "profile_updates") h->profile_updates = static_cast<int>(value);
   else if (n == "multi_stream") h->multi_stream = static_cast<int>(value);
   else if (n == "update_tile_mode") h->update_tile_mode = static_cast<int>(value);
   else if (n == "lookahead") h->lookahead = static_cast<int>(value);
   else if (n == "fused_assembly") h->fused_assembly = static_cast<int>(value);
   else if (n == "fused_leaves") h->fused_leaves = static_cast<int>(value);
   else if (n == "gather_lists") h->gather_lists = static_cast<int>(value);
   else if (n == "small_solve_ctas") h->small_solve_ctas = static_cast<int>(value);
   else if (n == "lookahead_min_front") h->lookahead_min_front = static_cast<int>(value);
   else if (n == "pair_panels") h->pair_panels = static_cast<int>(value);
   else if (n == "pair_min_front") h->pair_min_front = static_cast<int>(value);
   else if (n == "partition_graph") h->partition_graph = static_cast<int>(value);
   else if (n == "share_trim_rows") h->share_trim_rows = static_cast<int>(value);
   else if (n == "partition_memory") h->partition_memory = static_cast<int>(value);
   else if (n == "debug_rank") {   // host-only handles (plan inspection in the CPU tests): the rank whose storage the plan describes
      if (h->has_device) return fail(h, B200KKT_EINVAL, "debug_rank: only for host-only handles (a device handle gets its rank from comm_init)");
      h->rank = std::max(0, std::min(h->num_gpus - 1, static_cast<int>(value)));
   }
   else if (n == "share_one_launch") h->share_one_launch = static_cast<int>(value);
   else if (n == "share_top_fronts") h->share_top_fronts = static_cast<int>(value);
   else if (n == "share_chunk_panels") h->share_chunk_panels = std::max(1, static_cast<int>(value));
   else if (n == "share_min_pivots") h->share_min_pivots = std::max(2 * NBO, static_cast<int>(value));
   else if (n == "share_min_flops") h->share_min_flops = value;
   else if (n == "partition_slack") h->partition_slack = std::max(1.0, value);
   else if (n == "debug_skip") h->debug_skip = static_cast<int>(value);   // timing experiments only: 1 no K4, 2 no panel kernels, 4 no K5
   else return fail(h, B200KKT_EINVAL, "unknown option: " + n);
   // numeric options are baked into the captured graphs
   if (h->factor_graph != nullptr) { cudaGraphExecDestroy(h->factor_graph); h->factor_graph = nullptr; }
   if (h->solve_graph != nullptr) { cudaGraphExecDestroy(h->solve_graph); h->solve_graph = nullptr; }
   return B200KKT_OK;
}

int b200kkt_set_user_permutation(b200kkt_handle* h, const int64_t* perm, uint64_t n) {
   if (h == nullptr) return B200KKT_EINVAL;
   if (perm == nullptr) { h->user_perm.clear(); if (h->aopt.ordering == 2) h->aopt.ordering = 0; return B200KKT_OK; }
   h->user_perm.assign(perm, perm + n);
   h->aopt.ordering = 2;
   return B200KKT_OK;
}

int b200kkt_set_pairing_values(b200kkt_handle* h, const double* values, uint64_t nnz) {
   if (h == nullptr) return B200KKT_EINVAL;
   if (values == nullptr) { h->pair_values.clear(); return B200KKT_OK; }
   try { h->pair_values.assign(values, values + nnz); } catch (const std::bad_alloc&) { return fail(h, B200KKT_ENOMEM, "host allocation failed"); }
   return B200KKT_OK;
}

int b200kkt_analyse(b200kkt_handle* h, uint64_t n, uint64_t nnz, uint64_t reg_size, uint64_t num_primal, const uint64_t* row,
                    const uint64_t* col) {
   if (h == nullptr || (nnz > 0 && (row == nullptr || col == nullptr))) return B200KKT_EINVAL;
   if (h->has_device) { cudaSetDevice(h->device); free_plan_device(h); }
   h->analysed = false;
   h->delay_count.clear();
   h->delayed_total = h->replans_total = h->replans_last = 0;
   if (h->has_device && h->partitioned() && !h->comm.active())
      return fail(h, B200KKT_EINVAL, "multi-GPU handle: b200kkt_comm_init must be called before b200kkt_analyse");
   const int rc = guarded(h, [&] {
      AnalyseOptions opt = h->aopt;
      if (opt.ordering == 2) {
         if (h->user_perm.size() != n) throw std::runtime_error("analyse: user permutation length differs from N");
         opt.user_perm = h->user_perm.data();
      }
      if (h->pair_values.size() == nnz && nnz > 0) opt.pair_values = h->pair_values.data();
      analyse(static_cast<int64_t>(n), static_cast<int64_t>(nnz), static_cast<int64_t>(reg_size), static_cast<int64_t>(num_primal), row, col,
              opt, h->plan);
      h->plan.share_min_pivots = h->share_min_pivots;
      h->plan.share_min_flops = h->share_min_flops;
      h->plan.partition_slack = h->partition_slack;
      map_subtrees(h->plan, h->num_gpus, h->share_top_fronts != 0);
      if (h->partitioned() && h->partition_memory) localize_storage(h->plan, h->rank);
      fill_plan_stats(h);
      if (h->has_device) upload_plan(h);
      h->analysed = true;
   });
   if (rc != B200KKT_OK && h->has_device) free_plan_device(h);
   std::vector<double>().swap(h->pair_values);   // a hint belongs to one analysis
   return rc;
}

int b200kkt_factorize(b200kkt_handle* h, const double* values) {
   if (int rc = need_device(h)) return rc;
   if (!h->analysed) return fail(h, B200KKT_EINVAL, "factorize before analyse");
   if (values == nullptr) return B200KKT_EINVAL;
   return guarded(h, [&] {
      CK(cudaSetDevice(h->device));
      CK(cudaMemcpyAsync(h->d_values, values, static_cast<size_t>(h->plan.nnz) * sizeof(double), cudaMemcpyHostToDevice, h->stream));
      h->values_resident = true;
      run_factor(h);
   });
}

int b200kkt_refactorize_shifted(b200kkt_handle* h, const double* reg_values) {
   if (int rc = need_device(h)) return rc;
   if (!h->analysed || !h->values_resident) return fail(h, B200KKT_EINVAL, "refactorize_shifted before factorize");
   if (reg_values == nullptr && h->plan.reg_size > 0) return B200KKT_EINVAL;
   return guarded(h, [&] {
      CK(cudaSetDevice(h->device));
      if (h->plan.reg_size > 0)
         CK(cudaMemcpyAsync(h->d_values, reg_values, static_cast<size_t>(h->plan.reg_size) * sizeof(double), cudaMemcpyHostToDevice, h->stream));
      run_factor(h);
   });
}

int b200kkt_factorize_device(b200kkt_handle* h, const double* d_values) {
   if (int rc = need_device(h)) return rc;
   if (!h->analysed) return fail(h, B200KKT_EINVAL, "factorize before analyse");
   if (d_values == nullptr) return B200KKT_EINVAL;
   return guarded(h, [&] {
      CK(cudaSetDevice(h->device));
      if (d_values != h->d_values)
         CK(cudaMemcpyAsync(h->d_values, d_values, static_cast<size_t>(h->plan.nnz) * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
      h->values_resident = true;
      run_factor(h);
   });
}

int b200kkt_factorize_resident(b200kkt_handle* h) {
   if (int rc = need_device(h)) return rc;
   if (!h->analysed || !h->values_resident) return fail(h, B200KKT_EINVAL, "factorize_resident: no resident values (b200kkt_assemble_kkt_values_device or a factorize first)");
   return guarded(h, [&] {
      CK(cudaSetDevice(h->device));
      run_factor(h);
   });
}

int b200kkt_inertia(const b200kkt_handle* h, uint64_t* n_pos, uint64_t* n_neg, uint64_t* n_zero) {
   if (h == nullptr) return B200KKT_EINVAL;
   if (!h->factorized) return fail(h, B200KKT_EINVAL, "inertia before factorize");
   if (n_pos != nullptr) *n_pos = h->inertia[0];
   if (n_neg != nullptr) *n_neg = h->inertia[1];
   if (n_zero != nullptr) *n_zero = h->inertia[2];
   return B200KKT_OK;
}

int b200kkt_solve(b200kkt_handle* h, const double* rhs, double* x, double* backward_error) {
   if (int rc = need_device(h)) return rc;
   if (!h->factorized) return fail(h, B200KKT_EINVAL, "solve before factorize");
   if (rhs == nullptr || x == nullptr) return B200KKT_EINVAL;
   return guarded(h, [&] {
      CK(cudaSetDevice(h->device));
      const size_t nbytes = static_cast<size_t>(h->plan.N) * sizeof(double);
      CK(cudaMemcpyAsync(h->d_b, rhs, nbytes, cudaMemcpyHostToDevice, h->stream));
      run_solve(h, h->d_b, h->d_x, backward_error);
      CK(cudaMemcpyAsync(x, h->d_x, nbytes, cudaMemcpyDeviceToHost, h->stream));
      CK(cudaStreamSynchronize(h->stream));
   });
}

int b200kkt_solve_device(b200kkt_handle* h, const double* d_rhs, double* d_x, double* backward_error) {
   if (int rc = need_device(h)) return rc;
   if (!h->factorized) return fail(h, B200KKT_EINVAL, "solve before factorize");
   if (d_rhs == nullptr || d_x == nullptr) return B200KKT_EINVAL;
   return guarded(h, [&] {
      CK(cudaSetDevice(h->device));
      run_solve(h, d_rhs, d_x, backward_error);
   });
}

int b200kkt_backward_error_noalloc(b200kkt_handle* h, const double* d_x, const double* d_rhs, double* normwise, double* componentwise) {
   if (int rc = need_device(h)) return rc;
   if (!h->analysed || !h->values_resident) return fail(h, B200KKT_EINVAL, "backward_error_noalloc before factorize (the matrix values must be resident)");
   if (d_x == nullptr || d_rhs == nullptr) return B200KKT_EINVAL;
   if (h->plan.N > noalloc_max_n())
      return fail(h, B200KKT_EINVAL, "backward_error_noalloc: N = " + std::to_string(h->plan.N) + " exceeds B200KKT_NOALLOC_MAX_N = " +
                                     std::to_string(noalloc_max_n()) + " (size of the static device scratch array); use b200kkt_solve_device");
   // no cudaMalloc / cudaFree / allocating routine on this path: static device scratch, results through a stack buffer
   if (cudaSetDevice(h->device) != cudaSuccess) return fail(h, B200KKT_ECUDA, "cudaSetDevice failed");
   unsigned long long bits[5] = {0, 0, 0, 0, 0};
   int launches = 0;
   const int e = launch_backward_error_noalloc(h->device, h->C, h->d_values, d_x, d_rhs, bits, &launches, h->stream);
   h->launches += launches;
   if (e != 0) {
      cudaGetLastError();
      return fail(h, B200KKT_ECUDA, std::string("backward_error_noalloc: CUDA error: ") + cudaGetErrorString(static_cast<cudaError_t>(e)));
   }
   const double knorm = bits_to_double(bits[0]), rnorm = bits_to_double(bits[1]), xnorm = bits_to_double(bits[2]), bnorm = bits_to_double(bits[3]);
   const double denom = knorm * xnorm + bnorm;
   if (normwise != nullptr) *normwise = (denom > 0.0) ? rnorm / denom : 0.0;
   if (componentwise != nullptr) *componentwise = bits_to_double(bits[4]);
   return B200KKT_OK;
}

int b200kkt_get_permutation(const b200kkt_handle* h, uint64_t* perm, uint64_t* scatter_map) {
   if (h == nullptr) return B200KKT_EINVAL;
   if (!h->analysed) return fail(h, B200KKT_EINVAL, "get_permutation before analyse");
   if (perm != nullptr) for (int64_t i = 0; i < h->plan.N; ++i) perm[i] = static_cast<uint64_t>(h->plan.perm[i]);
   if (scatter_map != nullptr) for (int64_t e = 0; e < h->plan.nnz; ++e) scatter_map[e] = static_cast<uint64_t>(h->plan.scatter[e]);
   return B200KKT_OK;
}

int b200kkt_get_stats(const b200kkt_handle* h, b200kkt_stats* stats) {
   if (h == nullptr || stats == nullptr) return B200KKT_EINVAL;
   *stats = h->stats;
   stats->kernel_launches = h->launches;
   stats->comm_bytes_factor = h->comm_bytes_factor;
   stats->comm_bytes_sweep = h->comm_bytes_solve;
   if (h->factorized) { stats->num_null_pivots_last = h->h_counters[2]; stats->num_two_by_two_last = h->h_counters[3]; }
   stats->num_delayed_total = h->delayed_total;
   stats->num_replans_total = h->replans_total;
   stats->num_replans_last = h->replans_last;
   return B200KKT_OK;
}

const char* b200kkt_last_error(const b200kkt_handle* h) { return h == nullptr ? "null handle" : h->err.c_str(); }

int b200kkt_get_stream(const b200kkt_handle* h, void** stream) {
   if (h == nullptr || stream == nullptr) return B200KKT_EINVAL;
   *stream = static_cast<void*>(h->stream);
   return B200KKT_OK;
}

int b200kkt_plan_sizes(const b200kkt_handle* h, int64_t sizes[8]) {
   if (h == nullptr || sizes == nullptr) return B200KKT_EINVAL;
   if (!h->analysed) return fail(h, B200KKT_EINVAL, "plan getter before analyse");
   const Plan& P = h->plan;
   sizes[0] = P.nsn; sizes[1] = static_cast<int64_t>(P.rows.size()); sizes[2] = P.pool_size; sizes[3] = P.nlevels;
   sizes[4] = P.sum_r; sizes[5] = static_cast<int64_t>(P.asm_dst.size()); sizes[6] = P.cb_pool_size; sizes[7] = 0;
   return B200KKT_OK;
}

int b200kkt_plan_fronts(const b200kkt_handle* h, int32_t* k, int32_t* f, int32_t* ld, int64_t* pool_offset, int32_t* parent, int32_t* level,
                        int32_t* first_col) {
   if (h == nullptr) return B200KKT_EINVAL;
   if (!h->analysed) return fail(h, B200KKT_EINVAL, "plan getter before analyse");
   const Plan& P = h->plan;
   for (int32_t s = 0; s < P.nsn; ++s) {
      if (k != nullptr) k[s] = P.k[s];
      if (f != nullptr) f[s] = P.f[s];
      if (ld != nullptr) ld[s] = P.ld[s];
      if (pool_offset != nullptr) pool_offset[s] = P.front_off[s];
      if (parent != nullptr) parent[s] = P.parent[s];
      if (level != nullptr) level[s] = P.level[s];
      if (first_col != nullptr) first_col[s] = P.sn_first[s];
   }
   return B200KKT_OK;
}

int b200kkt_plan_rows(const b200kkt_handle* h, int64_t* rows_off, int32_t* rows) {
   if (h == nullptr) return B200KKT_EINVAL;
   if (!h->analysed) return fail(h, B200KKT_EINVAL, "plan getter before analyse");
   const Plan& P = h->plan;
   if (rows_off != nullptr) std::copy(P.rows_off.begin(), P.rows_off.end(), rows_off);
   if (rows != nullptr) std::copy(P.rows.begin(), P.rows.end(), rows);
   return B200KKT_OK;
}

int b200kkt_debug_assemble(b200kkt_handle* h, const double* values, double* pool_out, double* scale_out) {
   if (int rc = need_device(h)) return rc;
   if (!h->analysed || values == nullptr) return B200KKT_EINVAL;
   return guarded(h, [&] {
      const Plan& P = h->plan;
      cudaStream_t st = h->stream;
      CK(cudaSetDevice(h->device));
      CK(cudaMemcpyAsync(h->d_values, values, static_cast<size_t>(P.nnz) * sizeof(double), cudaMemcpyHostToDevice, st));
      h->values_resident = true;
      h->factorized = false;
      h->launches += launch_fill(h->d_scale, P.N, 1.0, st);
      double* cur = h->d_scale;
      double* nxt = h->d_scale_tmp;
      for (int sw = 0; sw < h->scaling_sweeps; ++sw) {
         h->launches += launch_scale_sweep(h->C, h->d_values, cur, nxt, st);
         std::swap(cur, nxt);
      }
      h->d_scale_final = cur;
      if (!h->fused_assembly) CK(cudaMemsetAsync(h->W.lpool, 0, static_cast<size_t>(P.pool_size) * sizeof(double), st));
      h->launches += h->fused_assembly ? launch_assemble_fill(h->M, h->d_values, cur, h->W.lpool, st) : launch_assemble(h->M, h->d_values, cur, h->W.lpool, st);
      if (pool_out != nullptr) CK(cudaMemcpyAsync(pool_out, h->W.lpool, static_cast<size_t>(P.pool_size) * sizeof(double), cudaMemcpyDeviceToHost, st));
      if (scale_out != nullptr) CK(cudaMemcpyAsync(scale_out, cur, static_cast<size_t>(P.N) * sizeof(double), cudaMemcpyDeviceToHost, st));
      CK(cudaStreamSynchronize(st));
      CK(cudaGetLastError());
   });
}

int b200kkt_debug_delay_columns(b200kkt_handle* h, uint64_t count, const int64_t* positions) {
   if (h == nullptr || (count > 0 && positions == nullptr)) return B200KKT_EINVAL;
   if (!h->analysed) return fail(h, B200KKT_EINVAL, "debug_delay_columns before analyse");
   return guarded(h, [&] {
      if (h->has_device) CK(cudaSetDevice(h->device));
      std::vector<int32_t> cols;
      for (uint64_t i = 0; i < count; ++i) {
         if (positions[i] < 0 || positions[i] >= h->plan.N) throw std::runtime_error("debug_delay_columns: position out of range");
         if (h->plan.parent[h->plan.sn_of[positions[i]]] >= 0) cols.push_back(static_cast<int32_t>(positions[i]));
      }
      if (!cols.empty()) replan_with_delays(h, cols);
      h->factorized = false;
   });
}

int b200kkt_debug_factor(b200kkt_handle* h, double* pool_out, double* d_diag, double* d_off, int32_t* piv_type, int32_t* local_perm) {
   if (int rc = need_device(h)) return rc;
   if (!h->factorized) return fail(h, B200KKT_EINVAL, "debug_factor before factorize");
   return guarded(h, [&] {
      const Plan& P = h->plan;
      cudaStream_t st = h->stream;
      CK(cudaSetDevice(h->device));
      const size_t n = static_cast<size_t>(P.N);
      if (pool_out != nullptr) CK(cudaMemcpyAsync(pool_out, h->W.lpool, static_cast<size_t>(P.pool_size) * sizeof(double), cudaMemcpyDeviceToHost, st));
      if (d_diag != nullptr) CK(cudaMemcpyAsync(d_diag, h->W.dd, n * sizeof(double), cudaMemcpyDeviceToHost, st));
      if (d_off != nullptr) CK(cudaMemcpyAsync(d_off, h->W.de, n * sizeof(double), cudaMemcpyDeviceToHost, st));
      if (piv_type != nullptr) CK(cudaMemcpyAsync(piv_type, h->W.ptype, n * sizeof(int), cudaMemcpyDeviceToHost, st));
      if (local_perm != nullptr) CK(cudaMemcpyAsync(local_perm, h->W.lperm, n * sizeof(int), cudaMemcpyDeviceToHost, st));
      CK(cudaStreamSynchronize(st));
   });
}

}  // extern "C"

// ================================================================================================
// Kernel microbenchmark hook (profiling only): one synthetic front of order f with k pivots, random
// panel data, times `reps` launches of the K4 trailing update (mode 1, J0 = 0) with CUDA events.
// ================================================================================================
extern "C" int b200kkt_debug_bench_update(int device, int f, int k, int reps, int big_tiles, double* seconds_per_launch, double* flops_per_launch) {
   if (f <= 0 || k <= 0 || k > NBO || k >= f || reps <= 0) return B200KKT_EINVAL;
   try {
      CK(cudaSetDevice(device));
      if (init_kernel_attributes() != 0) return B200KKT_ECUDA;
      const int ld = (f + 3) & ~3, r = f - k, ldc = r;
      std::vector<double> hl(static_cast<size_t>(ld) * k), hw(static_cast<size_t>(ld) * NBO);
      uint64_t sd = 88172645463325252ull;
      auto rnd = [&] { sd ^= sd << 13; sd ^= sd >> 7; sd ^= sd << 17; return (static_cast<double>(sd >> 11) / 9007199254740992.0) - 0.5; };
      for (auto& v : hl) v = rnd();
      for (auto& v : hw) v = rnd();
      double *lpool, *cbpool, *wpool;
      CK(cudaMalloc(reinterpret_cast<void**>(&lpool), hl.size() * sizeof(double)));
      CK(cudaMalloc(reinterpret_cast<void**>(&wpool), hw.size() * sizeof(double)));
      CK(cudaMalloc(reinterpret_cast<void**>(&cbpool), static_cast<size_t>(ldc) * r * sizeof(double)));
      CK(cudaMemcpy(lpool, hl.data(), hl.size() * sizeof(double), cudaMemcpyHostToDevice));
      CK(cudaMemcpy(wpool, hw.data(), hw.size() * sizeof(double), cudaMemcpyHostToDevice));
      CK(cudaMemset(cbpool, 0, static_cast<size_t>(ldc) * r * sizeof(double)));
      const int64_t zero64 = 0;
      const int hk = k, hf = f, hld = ld, hldc = ldc, hlist = 0;
      int64_t* d64; int* d32;
      CK(cudaMalloc(reinterpret_cast<void**>(&d64), 8 * sizeof(int64_t)));
      CK(cudaMalloc(reinterpret_cast<void**>(&d32), 8 * sizeof(int)));
      CK(cudaMemcpy(d64, &zero64, sizeof(int64_t), cudaMemcpyHostToDevice));
      const int h32[5] = {hk, hf, hld, hldc, hlist};
      CK(cudaMemcpy(d32, h32, sizeof(h32), cudaMemcpyHostToDevice));
      FrontArrays F{};
      F.off = d64; F.cboff = d64; F.woff = d64; F.xoff = d64;
      F.k = d32; F.f = d32 + 1; F.ld = d32 + 2; F.ldc = d32 + 3;
      FactorArrays W{};
      W.lpool = lpool; W.cbpool = cbpool; W.wpool = wpool;
      cudaStream_t st;
      CK(cudaStreamCreate(&st));
      cudaEvent_t e0, e1;
      CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
      for (int i = 0; i < 2; ++i) launch_update(F, d32 + 4, 1, f, 1, 0, 0, W, big_tiles, st);
      CK(cudaEventRecord(e0, st));
      for (int i = 0; i < reps; ++i) launch_update(F, d32 + 4, 1, f, 1, 0, 0, W, big_tiles, st);
      CK(cudaEventRecord(e1, st));
      CK(cudaStreamSynchronize(st));
      CK(cudaGetLastError());
      float ms = 0.f;
      CK(cudaEventElapsedTime(&ms, e0, e1));
      if (seconds_per_launch != nullptr) *seconds_per_launch = ms * 1e-3 / reps;
      if (flops_per_launch != nullptr) *flops_per_launch = 2.0 * k * (static_cast<double>(r) * (r + 1.0) * 0.5);
      cudaEventDestroy(e0); cudaEventDestroy(e1); cudaStreamDestroy(st);
      cudaFree(lpool); cudaFree(wpool); cudaFree(cbpool); cudaFree(d64); cudaFree(d32);
      return B200KKT_OK;
   } catch (const CudaError&) {
      cudaGetLastError();
      return B200KKT_ECUDA;
   }
}

extern "C" int b200kkt_debug_phase_cycles(unsigned long long* out16, int reset) { return debug_phase_cycles(out16, reset != 0); }

extern "C" int b200kkt_debug_bench_bk(int variant, int reps, long long* cycles) { return debug_bench_bk(variant, reps, cycles, nullptr); }
