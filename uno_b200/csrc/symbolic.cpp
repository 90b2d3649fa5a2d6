// Host symbolic phase: upper-triangular COO pattern -> integer plan.
//
// Takes the place of MUMPS JOB=1 behind MUMPSSolver::do_symbolic_analysis
// (reference: uno/ingredients/subproblem_solvers/MUMPS/MUMPSSolver.cpp:72-83 and the index copy at
// :149-157).  The input convention is the reference's COO: row <= col, 0-based, duplicates legal,
// the first reg_size entries are (k,k) regularization placeholders
// (uno/linear_algebra/COOFormat.hpp:101-125).  Explicit zeros are structural (never pruned).
//
// Pipeline: adjacency -> KKT pairing (constraint row matched to a primal column so that a 2x2
// pivot is always available inside one front) -> METIS_NodeND on the pair-compressed graph ->
// elimination tree -> postorder -> column counts -> fundamental supernodes -> relaxed
// amalgamation -> final numbering (pairs first inside each supernode) -> front row structures ->
// extend-add maps, level schedule, assembly scatter map, CSR for the residual.
#include "plan.hpp"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cerrno>
#include <cstring>
#include <limits>
#include <map>
#include <set>
#include <numeric>
#include <stdexcept>
#include <sys/types.h>
#include <sys/wait.h>
#include <unistd.h>

extern "C" {
// CUDA ships libmetis_static.a (METIS 5.x, 64-bit idx_t) without a header.
int METIS_NodeND(int64_t* nvtxs, int64_t* xadj, int64_t* adjncy, int64_t* vwgt, int64_t* options,
                 int64_t* perm, int64_t* iperm);
int METIS_SetDefaultOptions(int64_t* options);
int METIS_ComputeVertexSeparator(int64_t* nvtxs, int64_t* xadj, int64_t* adjncy, int64_t* vwgt, int64_t* options, int64_t* sepsize,
                                 int64_t* part);
}

namespace b200kkt {
namespace {

using clk = std::chrono::steady_clock;
double secs(clk::time_point a, clk::time_point b) { return std::chrono::duration<double>(b - a).count(); }

struct Graph {
   std::vector<int64_t> ptr;
   std::vector<int32_t> idx;
};

// symmetric off-diagonal adjacency, sorted, duplicates removed
Graph build_adjacency(int64_t N, int64_t nnz, const uint64_t* row, const uint64_t* col) {
   Graph g;
   g.ptr.assign(static_cast<size_t>(N) + 1, 0);
   for (int64_t e = 0; e < nnz; ++e) {
      const uint64_t r = row[e], c = col[e];
      if (r >= static_cast<uint64_t>(N) || c >= static_cast<uint64_t>(N)) throw std::runtime_error("analyse: index out of range");
      if (r > c) throw std::runtime_error("analyse: entry below the diagonal (COO must be upper triangular)");
      if (r != c) { ++g.ptr[r + 1]; ++g.ptr[c + 1]; }
   }
   for (int64_t i = 0; i < N; ++i) g.ptr[i + 1] += g.ptr[i];
   g.idx.resize(static_cast<size_t>(g.ptr[N]));
   std::vector<int64_t> fill(g.ptr.begin(), g.ptr.end() - 1);
   for (int64_t e = 0; e < nnz; ++e) {
      const uint64_t r = row[e], c = col[e];
      if (r != c) {
         g.idx[fill[r]++] = static_cast<int32_t>(c);
         g.idx[fill[c]++] = static_cast<int32_t>(r);
      }
   }
   // sort + unique per row, compact in place
   int64_t out = 0;
   int64_t begin = 0;
   for (int64_t i = 0; i < N; ++i) {
      const int64_t end = g.ptr[i + 1];
      std::sort(g.idx.begin() + begin, g.idx.begin() + end);
      const int64_t row_start = out;
      for (int64_t p = begin; p < end; ++p) {
         if (p == begin || g.idx[p] != g.idx[p - 1]) g.idx[out++] = g.idx[p];
      }
      begin = end;
      g.ptr[i] = row_start;
      if (i + 1 == N) g.ptr[N] = out;
   }
   // ptr[i] currently holds row starts for i<N and ptr[N]=out; nothing else to fix
   g.idx.resize(static_cast<size_t>(out));
   g.idx.shrink_to_fit();
   return g;
}

// Match every constraint-like node with a distinct adjacent primal node (bipartite matching:
// greedy by ascending primal degree, then augmenting paths).  partner[i] = matched node or -1.
void kkt_pairing(int64_t N, const Graph& g, const std::vector<char>& is_con, std::vector<int32_t>& partner, const std::vector<float>& wgt) {
   partner.assign(static_cast<size_t>(N), -1);
   const bool weighted = !wgt.empty();
   auto degree = [&](int32_t v) { return g.ptr[v + 1] - g.ptr[v]; };
   std::vector<int32_t> unmatched;
   int64_t first_con = 0;
   while (first_con < N && !is_con[first_con]) ++first_con;
   for (int64_t j = 0; j < N; ++j) {
      if (!is_con[j]) continue;
      // lowest primal degree first; ties go to the primal whose index is closest to the constraint's ordinal
      // (states and their equations are usually numbered alike, which keeps the compressed graph grid-like)
      int32_t best = -1;
      int64_t best_deg = 0, best_dist = 0;
      float best_w = 0.f;
      const int64_t ordinal = j - first_con;
      for (int64_t p = g.ptr[j]; p < g.ptr[j + 1]; ++p) {
         const int32_t v = g.idx[p];
         if (is_con[v] || partner[v] != -1) continue;
         const int64_t d = degree(v);
         const int64_t dist = (v > ordinal) ? v - ordinal : ordinal - v;
         if (weighted) {
            // largest equilibrated entry first (within a factor of two: the weights are estimates), then the structural rule
            const float wv = wgt[static_cast<size_t>(p)];
            const bool better = best == -1 || wv > 2.f * best_w || (wv >= 0.5f * best_w && (d < best_deg || (d == best_deg && dist < best_dist)));
            if (better) { best = v; best_deg = d; best_dist = dist; best_w = std::max(best_w, wv); }
            continue;
         }
         if (best == -1 || d < best_deg || (d == best_deg && dist < best_dist)) { best = v; best_deg = d; best_dist = dist; }
      }
      if (best != -1) { partner[j] = best; partner[best] = static_cast<int32_t>(j); }
      else unmatched.push_back(static_cast<int32_t>(j));
   }
   if (unmatched.empty()) return;
   // Augmenting paths (iterative Kuhn) for the constraints the greedy pass left unmatched.  The DFS
   // stack holds constraint nodes; partner[stack[d]] (for d > 0) is the primal through which
   // stack[d] was reached.  Work per root is bounded so analysis stays near-linear.
   std::vector<int32_t> stamp(static_cast<size_t>(N), -1);
   std::vector<int64_t> cursor(static_cast<size_t>(N), 0);
   std::vector<int32_t> stack;
   int32_t round = 0;
   for (int32_t root : unmatched) {
      ++round;
      int64_t budget = 200000;
      stack.assign(1, root);
      cursor[root] = g.ptr[root];
      bool found = false;
      while (!stack.empty() && budget > 0 && !found) {
         const int32_t j = stack.back();
         bool descended = false;
         while (cursor[j] < g.ptr[j + 1]) {
            --budget;
            const int32_t v = g.idx[cursor[j]++];
            if (is_con[v] || stamp[v] == round) continue;
            stamp[v] = round;
            if (partner[v] == -1) {
               // flip the alternating path: every constraint on the stack hands its old primal down
               int32_t free_v = v;
               for (int64_t d = static_cast<int64_t>(stack.size()) - 1; d >= 0; --d) {
                  const int32_t cj = stack[d];
                  const int32_t old_v = partner[cj];   // -1 for the root
                  partner[cj] = free_v;
                  partner[free_v] = cj;
                  free_v = old_v;
               }
               found = true;
            } else {
               const int32_t next_con = partner[v];
               stack.push_back(next_con);
               cursor[next_con] = g.ptr[next_con];
               descended = true;
            }
            break;
         }
         if (!found && !descended) stack.pop_back();
      }
   }
}

}  // namespace

// fronts grouped by level (P.level given), big fronts first inside a level
// ------------------------------------------------------------------------------------------------
// Nested dissection with the top levels of the recursion done here, the halves of every split side by side.
// METIS keeps process-wide state (its random number generator), so concurrent calls from threads would make the ordering
// depend on the interleaving - and every rank of a partitioned run must compute the SAME ordering.  The halves therefore run
// in forked PROCESSES: each has a private copy of that state, the result is a function of the graph alone.  A child only runs
// METIS on host memory and writes its answer to a pipe; it never touches the CUDA context it inherited and leaves with _exit.
// ------------------------------------------------------------------------------------------------
struct NdGraph {
   std::vector<int64_t> xadj, adjncy, vwgt;
   std::vector<int64_t> ids;   // node of the top-level graph behind each local node
};

bool nd_serial(NdGraph& G, const std::vector<int64_t>& options, std::vector<int64_t>& order) {
   int64_t n = static_cast<int64_t>(G.ids.size());
   order.resize(static_cast<size_t>(n));
   if (n == 0) return true;
   if (n < 3 || G.adjncy.empty()) { order = G.ids; return true; }
   std::vector<int64_t> perm(static_cast<size_t>(n)), iperm(static_cast<size_t>(n)), opts(options);
   if (METIS_NodeND(&n, G.xadj.data(), G.adjncy.data(), G.vwgt.data(), opts.data(), perm.data(), iperm.data()) != 1) return false;
   for (int64_t i = 0; i < n; ++i) order[static_cast<size_t>(i)] = G.ids[static_cast<size_t>(perm[static_cast<size_t>(i)])];
   return true;
}

void nd_induce(const NdGraph& G, const std::vector<int64_t>& part, int64_t which, NdGraph& out) {
   const int64_t n = static_cast<int64_t>(G.ids.size());
   std::vector<int64_t> local(static_cast<size_t>(n), -1);
   int64_t m = 0;
   for (int64_t v = 0; v < n; ++v) if (part[static_cast<size_t>(v)] == which) local[static_cast<size_t>(v)] = m++;
   out.xadj.assign(static_cast<size_t>(m) + 1, 0);
   out.vwgt.resize(static_cast<size_t>(m));
   out.ids.resize(static_cast<size_t>(m));
   out.adjncy.clear();
   for (int64_t v = 0; v < n; ++v) {
      const int64_t lv = local[static_cast<size_t>(v)];
      if (lv < 0) continue;
      out.vwgt[static_cast<size_t>(lv)] = G.vwgt[static_cast<size_t>(v)];
      out.ids[static_cast<size_t>(lv)] = G.ids[static_cast<size_t>(v)];
      for (int64_t p = G.xadj[static_cast<size_t>(v)]; p < G.xadj[static_cast<size_t>(v) + 1]; ++p) {
         const int64_t lw = local[static_cast<size_t>(G.adjncy[static_cast<size_t>(p)])];
         if (lw >= 0) out.adjncy.push_back(lw);
      }
      out.xadj[static_cast<size_t>(lv) + 1] = static_cast<int64_t>(out.adjncy.size());
   }
}

bool write_all(int fd, const void* buf, size_t bytes) {
   const char* p = static_cast<const char*>(buf);
   while (bytes > 0) {
      const ssize_t w = ::write(fd, p, bytes);
      if (w < 0) { if (errno == EINTR) continue; return false; }
      p += w; bytes -= static_cast<size_t>(w);
   }
   return true;
}

bool read_all(int fd, void* buf, size_t bytes) {
   char* p = static_cast<char*>(buf);
   while (bytes > 0) {
      const ssize_t r = ::read(fd, p, bytes);
      if (r < 0) { if (errno == EINTR) continue; return false; }
      if (r == 0) return false;
      p += r; bytes -= static_cast<size_t>(r);
   }
   return true;
}

// elimination order of G (top-level node ids): separator last, the two sides before it, `depth` levels of splits done here
bool nd_process_parallel(NdGraph& G, int depth, const std::vector<int64_t>& options, std::vector<int64_t>& order) {
   int64_t n = static_cast<int64_t>(G.ids.size());
   if (depth <= 0 || n < 20000) return nd_serial(G, options, order);
   std::vector<int64_t> part(static_cast<size_t>(n)), opts(options);
   int64_t sepsize = 0;
   if (METIS_ComputeVertexSeparator(&n, G.xadj.data(), G.adjncy.data(), G.vwgt.data(), opts.data(), &sepsize, part.data()) != 1) return false;
   NdGraph A, B;
   nd_induce(G, part, 0, A);
   nd_induce(G, part, 1, B);
   std::vector<int64_t> sep;
   for (int64_t v = 0; v < n; ++v) if (part[static_cast<size_t>(v)] == 2) sep.push_back(G.ids[static_cast<size_t>(v)]);
   if (A.ids.empty() || B.ids.empty()) return nd_serial(G, options, order);   // no real split: leave it to METIS
   { NdGraph empty; std::swap(G.xadj, empty.xadj); std::swap(G.adjncy, empty.adjncy); std::swap(G.vwgt, empty.vwgt); }   // the halves carry on
   std::vector<int64_t> order_a, order_b;
   int fds[2];
   pid_t child = -1;
   if (::pipe(fds) == 0) {
      child = ::fork();
      if (child == 0) {
         ::close(fds[0]);
         std::vector<int64_t> mine;
         const bool ok = nd_process_parallel(A, depth - 1, options, mine) && mine.size() == A.ids.size();
         const int64_t head = ok ? static_cast<int64_t>(mine.size()) : -1;
         const bool sent = write_all(fds[1], &head, sizeof head) && (!ok || write_all(fds[1], mine.data(), mine.size() * sizeof(int64_t)));
         ::close(fds[1]);
         ::_exit(sent && ok ? 0 : 1);
      }
      if (child < 0) { ::close(fds[0]); ::close(fds[1]); }
   }
   bool ok_b = nd_process_parallel(B, depth - 1, options, order_b);
   bool ok_a = false;
   if (child > 0) {
      ::close(fds[1]);
      int64_t head = -1;
      if (read_all(fds[0], &head, sizeof head) && head == static_cast<int64_t>(A.ids.size())) {
         order_a.resize(static_cast<size_t>(head));
         ok_a = read_all(fds[0], order_a.data(), order_a.size() * sizeof(int64_t));
      }
      ::close(fds[0]);
      int status = 0;
      while (::waitpid(child, &status, 0) < 0 && errno == EINTR) {}
      ok_a = ok_a && WIFEXITED(status) && WEXITSTATUS(status) == 0;
   }
   if (!ok_a) ok_a = nd_process_parallel(A, 0, options, order_a);   // no child (or it failed): this process does that half too
   if (!ok_a || !ok_b) return false;
   order.clear();
   order.reserve(static_cast<size_t>(n));
   order.insert(order.end(), order_a.begin(), order_a.end());
   order.insert(order.end(), order_b.begin(), order_b.end());
   order.insert(order.end(), sep.begin(), sep.end());
   return true;
}

// B200KKT_TRACE_ANALYSE=1: wall clock of the sections of the symbolic phase on stderr
struct Trace {
   bool on = std::getenv("B200KKT_TRACE_ANALYSE") != nullptr;
   clk::time_point t = clk::now();
   void mark(const char* what) {
      if (!on) return;
      const auto n = clk::now();
      std::fprintf(stderr, "[b200kkt analyse] %-28s %8.3f s\n", what, secs(t, n));
      t = n;
   }
};

void build_level_lists(Plan& P) {
   const int32_t nsn = P.nsn;
   P.nlevels = 0;
   for (int32_t s = 0; s < nsn; ++s) P.nlevels = std::max(P.nlevels, P.level[s] + 1);
   P.level_ptr.assign(static_cast<size_t>(P.nlevels) + 1, 0);
   for (int32_t s = 0; s < nsn; ++s) ++P.level_ptr[P.level[s] + 1];
   for (int32_t l = 0; l < P.nlevels; ++l) P.level_ptr[l + 1] += P.level_ptr[l];
   P.level_fronts.resize(nsn);
   {
      std::vector<int32_t> fillp(P.level_ptr.begin(), P.level_ptr.end() - 1);
      for (int32_t s = 0; s < nsn; ++s) P.level_fronts[fillp[P.level[s]]++] = s;
      for (int32_t l = 0; l < P.nlevels; ++l) {
         std::stable_sort(P.level_fronts.begin() + P.level_ptr[l], P.level_fronts.begin() + P.level_ptr[l + 1],
                          [&](int32_t a, int32_t b) { return P.f[a] != P.f[b] ? P.f[a] > P.f[b] : P.k[a] > P.k[b]; });
      }
   }
}

// offsets of the contribution blocks in the transient arena (P.level_ptr / P.level_fronts given); `keep` (optional) marks the
// fronts whose block this rank ever holds - the others get no room
void build_cb_arena(Plan& P, const std::vector<char>* keep = nullptr) {
   const int32_t nsn = P.nsn;
   // ---- contribution-block arena: best fit over the level-order schedule -------------------------
   // A block lives from the level of its front until the level of its parent (where it is consumed by
   // the extend-add); the numeric phase runs level by level, so the offsets can be fixed here.
   {
      P.cb_off.assign(nsn, 0);
      // free blocks, coalesced: by offset (to merge neighbours) and by size (best fit in O(log n): the level-order schedule
      // frees and takes hundreds of thousands of small blocks, a first-fit walk over the list was most of the analysis time)
      std::map<int64_t, int64_t> free_list;                    // offset -> size
      std::set<std::pair<int64_t, int64_t>> by_size;           // (size, offset): the smallest block that fits, lowest address first
      int64_t top = 0;
      auto cb_size = [&](int32_t s) {
         const int64_t r = P.f[s] - P.k[s];
         return (r * (r + 1) / 2 + 1) & ~int64_t(1);   // packed lower triangle, even number of doubles
      };
      auto drop = [&](std::map<int64_t, int64_t>::iterator it) {
         by_size.erase({it->second, it->first});
         free_list.erase(it);
      };
      auto add = [&](int64_t o, int64_t size) {
         free_list.emplace(o, size);
         by_size.emplace(size, o);   // (size, offset)
      };
      auto alloc = [&](int64_t size) {
         auto fit = by_size.lower_bound({size, std::numeric_limits<int64_t>::min()});
         if (fit != by_size.end()) {
            const int64_t o = fit->second, rest = fit->first - size;
            by_size.erase(fit);
            free_list.erase(o);
            if (rest > 0) add(o + size, rest);
            return o;
         }
         // grow: extend a free block that touches the top, if any
         if (!free_list.empty()) {
            auto last = std::prev(free_list.end());
            if (last->first + last->second == top) {
               const int64_t o = last->first;
               drop(last);
               top = o + size;
               return o;
            }
         }
         const int64_t o = top;
         top += size;
         return o;
      };
      auto release = [&](int64_t o, int64_t size) {
         auto nx = free_list.lower_bound(o);
         if (nx != free_list.end() && o + size == nx->first) { size += nx->second; drop(nx); }
         auto pv = free_list.lower_bound(o);
         if (pv != free_list.begin()) {
            --pv;
            if (pv->first + pv->second == o) { o = pv->first; size += pv->second; drop(pv); }
         }
         add(o, size);
      };
      for (int32_t l = 0; l < P.nlevels; ++l) {
         for (int32_t t = P.level_ptr[l]; t < P.level_ptr[l + 1]; ++t) {
            const int32_t s = P.level_fronts[t];
            const int64_t sz = (keep == nullptr || (*keep)[s]) ? cb_size(s) : 0;
            if (sz > 0) P.cb_off[s] = alloc(sz);
         }
         for (int32_t t = P.level_ptr[l]; t < P.level_ptr[l + 1]; ++t) {
            const int32_t s = P.level_fronts[t];
            for (int32_t ci = P.child_ptr[s]; ci < P.child_ptr[s + 1]; ++ci) {
               const int32_t c = P.child_idx[ci];
               const int64_t sz = (keep == nullptr || (*keep)[c]) ? cb_size(c) : 0;
               if (sz > 0) release(P.cb_off[c], sz);
            }
         }
      }
      P.cb_pool_size = top;
   }
}

void analyse(int64_t N, int64_t nnz, int64_t reg_size, int64_t num_primal, const uint64_t* row, const uint64_t* col,
             const AnalyseOptions& opt, Plan& P) {
   if (N <= 0) throw std::runtime_error("analyse: empty matrix");
   if (N >= (int64_t(1) << 31) - 8 || nnz >= (int64_t(1) << 31) - 8) throw std::runtime_error("analyse: dimension exceeds int32 plan indices");
   if (reg_size > nnz) throw std::runtime_error("analyse: reg_size > nnz");
   P = Plan();
   P.N = N; P.nnz = nnz; P.reg_size = reg_size;
   if (num_primal <= 0 || num_primal > N) num_primal = N;
   P.num_primal = num_primal;
   const auto t0 = clk::now();

   Trace tro;
   Graph g = build_adjacency(N, nnz, row, col);
   tro.mark("adjacency");

   // ---- constraint-like nodes and KKT pairing --------------------------------------------------
   std::vector<char> is_con(static_cast<size_t>(N), 0);
   bool any_con = false;
   if (num_primal < N) {
      for (int64_t i = num_primal; i < N; ++i) is_con[i] = 1;
      any_con = true;
   } else {
      // no hint: a row without any structural diagonal entry behaves like a multiplier row
      std::vector<char> has_diag(static_cast<size_t>(N), 0);
      for (int64_t e = 0; e < nnz; ++e) if (row[e] == col[e]) has_diag[row[e]] = 1;
      for (int64_t i = 0; i < N; ++i) if (!has_diag[i] && g.ptr[i + 1] > g.ptr[i]) { is_con[i] = 1; any_con = true; }
   }
   std::vector<int32_t> partner(static_cast<size_t>(N), -1);
   std::vector<float> wgt;
   if (opt.kkt_pairing && any_con && opt.pair_values != nullptr) {
      // |a_ij| s_i s_j per adjacency entry, s = 1 / sqrt(row maximum) (one equilibration sweep on the hint values)
      std::vector<double> rmax(static_cast<size_t>(N), 0.0);
      for (int64_t e = 0; e < nnz; ++e) {
         const double a = std::fabs(opt.pair_values[e]);
         if (!(a <= 1e300)) continue;
         rmax[row[e]] = std::max(rmax[row[e]], a);
         rmax[col[e]] = std::max(rmax[col[e]], a);
      }
      std::vector<double> acc(g.idx.size(), 0.0);
      auto slot = [&](uint64_t r, uint64_t c) {
         const auto b = g.idx.begin() + g.ptr[r], e2 = g.idx.begin() + g.ptr[r + 1];
         return static_cast<size_t>(std::lower_bound(b, e2, static_cast<int32_t>(c)) - g.idx.begin());
      };
      for (int64_t e = 0; e < nnz; ++e) {
         if (row[e] == col[e]) continue;
         const double a = std::fabs(opt.pair_values[e]);
         if (!(a <= 1e300)) continue;
         acc[slot(row[e], col[e])] += a;
         acc[slot(col[e], row[e])] += a;
      }
      wgt.resize(g.idx.size());
      for (int64_t i = 0; i < N; ++i) {
         const double si = rmax[i] > 0.0 ? 1.0 / std::sqrt(rmax[i]) : 1.0;
         for (int64_t p = g.ptr[i]; p < g.ptr[i + 1]; ++p) {
            const int32_t v = g.idx[p];
            const double sv = rmax[v] > 0.0 ? 1.0 / std::sqrt(rmax[v]) : 1.0;
            wgt[static_cast<size_t>(p)] = static_cast<float>(acc[static_cast<size_t>(p)] * si * sv);
         }
      }
   }
   if (opt.kkt_pairing && any_con) kkt_pairing(N, g, is_con, partner, wgt);
   tro.mark("KKT pairing");
   for (int64_t i = 0; i < N; ++i) if (is_con[i] && partner[i] != -1) ++P.num_pairs;

   // ---- fill-reducing ordering on the pair-compressed graph -----------------------------------
   // compressed node = {primal, matched constraint} or a singleton
   std::vector<int32_t> cid(static_cast<size_t>(N), -1);
   std::vector<int32_t> cfirst;   // representative (primal first) of each compressed node
   for (int64_t i = 0; i < N; ++i) {
      if (cid[i] != -1) continue;
      const int32_t p = partner[i];
      const int32_t id = static_cast<int32_t>(cfirst.size());
      if (p == -1) { cid[i] = id; cfirst.push_back(static_cast<int32_t>(i)); }
      else {
         const int32_t var = is_con[i] ? p : static_cast<int32_t>(i);
         cid[var] = id; cid[partner[var]] = id; cfirst.push_back(var);
      }
   }
   // A vertex left alone whose only neighbour sits in a pair (a control or a slack that enters one constraint only) is a leaf of
   // the elimination tree wherever it is ordered before that neighbour: it is taken out of the graph METIS sees (which is then
   // the plain grid of pairs) and ordered first.
   std::vector<int32_t> leaves_first;
   if (opt.kkt_pairing && any_con && opt.ordering == 0) {
      for (int64_t i = 0; i < N; ++i) {
         if (partner[i] != -1 || g.ptr[i + 1] - g.ptr[i] != 1) continue;
         if (partner[g.idx[g.ptr[i]]] == -1) continue;
         leaves_first.push_back(static_cast<int32_t>(i));
      }
      if (!leaves_first.empty()) {
         std::vector<char> gone(static_cast<size_t>(N), 0);
         for (const int32_t v : leaves_first) gone[v] = 1;
         std::vector<int32_t> remap(cfirst.size(), -1), cfirst2;
         for (size_t c = 0; c < cfirst.size(); ++c) {
            if (gone[cfirst[c]]) continue;
            remap[c] = static_cast<int32_t>(cfirst2.size());
            cfirst2.push_back(cfirst[c]);
         }
         for (int64_t i = 0; i < N; ++i) cid[i] = gone[i] ? -1 : remap[static_cast<size_t>(cid[i])];
         cfirst.swap(cfirst2);
      }
   }
   const int64_t cN = static_cast<int64_t>(cfirst.size());
   std::vector<int64_t> corder(static_cast<size_t>(cN));   // corder[new] = compressed node
   std::iota(corder.begin(), corder.end(), 0);
   P.ordering_used = "natural";
   if (opt.ordering == 2 && opt.user_perm != nullptr) {
      // user permutation on ORIGINAL nodes; compressed order = first appearance of each compressed node
      std::vector<char> seen(static_cast<size_t>(cN), 0);
      int64_t out = 0;
      std::vector<char> check(static_cast<size_t>(N), 0);
      for (int64_t i = 0; i < N; ++i) {
         const int64_t o = opt.user_perm[i];
         if (o < 0 || o >= N || check[o]) throw std::runtime_error("analyse: user permutation is not a permutation");
         check[o] = 1;
         if (!seen[cid[o]]) { seen[cid[o]] = 1; corder[out++] = cid[o]; }
      }
      P.ordering_used = "user";
   } else if (opt.ordering == 0 && cN > 2) {
      std::vector<int64_t> xadj(static_cast<size_t>(cN) + 1, 0), adjncy, vwgt(static_cast<size_t>(cN), 1);
      adjncy.reserve(g.idx.size());
      std::vector<int32_t> mark(static_cast<size_t>(cN), -1);
      for (int64_t c = 0; c < cN; ++c) {
         const int32_t a = cfirst[c];
         const int32_t members[2] = {a, partner[a]};
         mark[c] = static_cast<int32_t>(c);
         for (int32_t mnode : members) {
            if (mnode < 0) continue;
            if (mnode != a) vwgt[c] = 2;
            for (int64_t p = g.ptr[mnode]; p < g.ptr[mnode + 1]; ++p) {
               const int32_t t = cid[g.idx[p]];
               if (t < 0) continue;   // a leaf ordered first
               if (mark[t] != c) { mark[t] = static_cast<int32_t>(c); adjncy.push_back(t); }
            }
         }
         xadj[c + 1] = static_cast<int64_t>(adjncy.size());
      }
      if (!adjncy.empty()) {
         std::vector<int64_t> options(40), mperm(static_cast<size_t>(cN)), miperm(static_cast<size_t>(cN));
         METIS_SetDefaultOptions(options.data());
         // option slots of METIS 5.1 (metis.h: NITER = 6, NSEPS = 15, UFACTOR = 16)
         if (opt.metis_niter > 0) options[6] = opt.metis_niter;
         if (opt.metis_nseps > 0) options[15] = opt.metis_nseps;
         if (opt.metis_ufactor > 0) options[16] = opt.metis_ufactor;
         int64_t n = cN;
         tro.mark("compressed graph");
         bool done = false;
         if (opt.parallel_ordering_depth > 0 && cN >= opt.parallel_ordering_min_nodes) {
            // big graphs: the top levels of the dissection here, their halves in parallel processes (same ordering on every run)
            NdGraph G;
            G.xadj = xadj; G.adjncy = adjncy; G.vwgt = vwgt;
            G.ids.resize(static_cast<size_t>(cN));
            std::iota(G.ids.begin(), G.ids.end(), 0);
            std::vector<int64_t> ord;
            if (nd_process_parallel(G, opt.parallel_ordering_depth, options, ord) && static_cast<int64_t>(ord.size()) == cN) {
               std::vector<char> seen(static_cast<size_t>(cN), 0);
               bool valid = true;
               for (const int64_t v : ord) { if (v < 0 || v >= cN || seen[static_cast<size_t>(v)]) { valid = false; break; } seen[static_cast<size_t>(v)] = 1; }
               if (valid) { corder = ord; P.ordering_used = "metis_nested_dissection_process_parallel"; done = true; }
            }
            tro.mark("nested dissection (parallel)");
         }
         if (!done) {
            const int rc = METIS_NodeND(&n, xadj.data(), adjncy.data(), vwgt.data(), options.data(), mperm.data(), miperm.data());
            tro.mark("METIS_NodeND");
            if (rc == 1) { corder = mperm; P.ordering_used = "metis_nodend"; }
         }
      }
   }
   // expand: primal first, then its matched constraint
   std::vector<int32_t> order(static_cast<size_t>(N)), ipos(static_cast<size_t>(N));
   {
      int64_t out = 0;
      for (const int32_t v : leaves_first) order[out++] = v;
      for (int64_t c = 0; c < cN; ++c) {
         const int32_t a = cfirst[corder[c]];
         order[out++] = a;
         if (partner[a] != -1) order[out++] = partner[a];
      }
      for (int64_t i = 0; i < N; ++i) ipos[order[i]] = static_cast<int32_t>(i);
   }
   const auto t1 = clk::now();
   P.t_order = secs(t0, t1);
   Trace tra;
   tra.t = t0;
   tra.mark("pairing + ordering");

   // ---- elimination tree (Liu) ----------------------------------------------------------------
   std::vector<int32_t> parent(static_cast<size_t>(N), -1);
   {
      std::vector<int32_t> ancestor(static_cast<size_t>(N), -1);
      for (int64_t j = 0; j < N; ++j) {
         const int32_t oj = order[j];
         for (int64_t p = g.ptr[oj]; p < g.ptr[oj + 1]; ++p) {
            int32_t i = ipos[g.idx[p]];
            while (i != -1 && i < j) {
               const int32_t inext = ancestor[i];
               ancestor[i] = static_cast<int32_t>(j);
               if (inext == -1) parent[i] = static_cast<int32_t>(j);
               i = inext;
            }
         }
      }
   }
   tra.mark("elimination tree");
   // ---- postorder and relabel -----------------------------------------------------------------
   {
      std::vector<int32_t> head(static_cast<size_t>(N), -1), next(static_cast<size_t>(N), -1), post(static_cast<size_t>(N));
      for (int64_t j = N - 1; j >= 0; --j) {
         if (parent[j] == -1) continue;
         next[j] = head[parent[j]];
         head[parent[j]] = static_cast<int32_t>(j);
      }
      int64_t kpos = 0;
      std::vector<int32_t> stack;
      for (int64_t root = 0; root < N; ++root) {
         if (parent[root] != -1) continue;
         stack.push_back(static_cast<int32_t>(root));
         while (!stack.empty()) {
            const int32_t p = stack.back();
            const int32_t i = head[p];
            if (i == -1) { stack.pop_back(); post[kpos++] = p; }
            else { head[p] = next[i]; stack.push_back(i); }
         }
      }
      std::vector<int32_t> ipost(static_cast<size_t>(N));
      for (int64_t kk = 0; kk < N; ++kk) ipost[post[kk]] = static_cast<int32_t>(kk);
      std::vector<int32_t> order2(static_cast<size_t>(N)), parent2(static_cast<size_t>(N));
      for (int64_t kk = 0; kk < N; ++kk) {
         order2[kk] = order[post[kk]];
         const int32_t pp = parent[post[kk]];
         parent2[kk] = (pp == -1) ? -1 : ipost[pp];
      }
      order.swap(order2);
      parent.swap(parent2);
      for (int64_t i = 0; i < N; ++i) ipos[order[i]] = static_cast<int32_t>(i);
   }
   tra.mark("postorder");
   // ---- column counts (Gilbert-Ng-Peyton, postordered so post = identity) ----------------------
   std::vector<int32_t> colcount(static_cast<size_t>(N), 0);
   {
      std::vector<int32_t> ancestor(static_cast<size_t>(N)), maxfirst(static_cast<size_t>(N), -1), prevleaf(static_cast<size_t>(N), -1),
         first(static_cast<size_t>(N), -1);
      std::vector<int32_t>& delta = colcount;
      for (int64_t kk = 0; kk < N; ++kk) {
         int32_t j = static_cast<int32_t>(kk);
         delta[j] = (first[j] == -1) ? 1 : 0;
         for (; j != -1 && first[j] == -1; j = parent[j]) first[j] = static_cast<int32_t>(kk);
      }
      std::iota(ancestor.begin(), ancestor.end(), 0);
      for (int64_t j = 0; j < N; ++j) {
         if (parent[j] != -1) --delta[parent[j]];
         const int32_t oj = order[j];
         for (int64_t p = g.ptr[oj]; p < g.ptr[oj + 1]; ++p) {
            const int32_t i = ipos[g.idx[p]];
            if (i <= j || first[j] <= maxfirst[i]) continue;   // j is not a leaf of the i-th row subtree
            maxfirst[i] = first[j];
            const int32_t jprev = prevleaf[i];
            prevleaf[i] = static_cast<int32_t>(j);
            if (jprev == -1) { ++delta[j]; continue; }          // first leaf
            int32_t q = jprev;
            while (q != ancestor[q]) q = ancestor[q];
            for (int32_t s = jprev; s != q;) { const int32_t sp = ancestor[s]; ancestor[s] = q; s = sp; }
            ++delta[j];
            --delta[q];
         }
         if (parent[j] != -1) ancestor[j] = parent[j];
      }
      for (int64_t j = 0; j < N; ++j) if (parent[j] != -1) colcount[parent[j]] += colcount[j];
   }
   tra.mark("column counts");
   // ---- fundamental supernodes ----------------------------------------------------------------
   std::vector<int32_t> sfirst0;   // first column of each fundamental supernode
   {
      std::vector<int32_t> nchild(static_cast<size_t>(N), 0);
      for (int64_t j = 0; j < N; ++j) if (parent[j] != -1) ++nchild[parent[j]];
      for (int64_t j = 0; j < N; ++j) {
         const bool chain = j > 0 && parent[j - 1] == j && colcount[j - 1] == colcount[j] + 1 && nchild[j] == 1;
         if (!chain) sfirst0.push_back(static_cast<int32_t>(j));
      }
      sfirst0.push_back(static_cast<int32_t>(N));
   }
   const int32_t ns0 = static_cast<int32_t>(sfirst0.size()) - 1;
   std::vector<int32_t> sn0_of(static_cast<size_t>(N));
   for (int32_t s = 0; s < ns0; ++s) for (int32_t j = sfirst0[s]; j < sfirst0[s + 1]; ++j) sn0_of[j] = s;

   // ---- relaxed amalgamation on the supernodal tree --------------------------------------------
   std::vector<int64_t> kk(static_cast<size_t>(ns0)), ff(static_cast<size_t>(ns0)), zeros(static_cast<size_t>(ns0), 0);
   std::vector<int32_t> sparent(static_cast<size_t>(ns0)), merged_into(static_cast<size_t>(ns0), -1);
   std::vector<char> force(static_cast<size_t>(ns0), 0);
   for (int32_t s = 0; s < ns0; ++s) {
      kk[s] = sfirst0[s + 1] - sfirst0[s];
      ff[s] = colcount[sfirst0[s]];
      const int32_t last = sfirst0[s + 1] - 1;
      sparent[s] = (parent[last] == -1) ? -1 : sn0_of[parent[last]];
      // a paired primal column must end up in the same front as its constraint row (its etree parent)
      const int32_t olast = order[last];
      if (partner[olast] != -1 && !is_con[olast] && sparent[s] != -1 && ipos[partner[olast]] > last) force[s] = 1;
   }
   {
      std::vector<std::vector<int32_t>> kids(static_cast<size_t>(ns0));
      for (int32_t s = 0; s < ns0; ++s) if (sparent[s] != -1) kids[sparent[s]].push_back(s);
      std::vector<int32_t> cand;
      for (int32_t s = 0; s < ns0; ++s) {
         cand = kids[s];
         // best structural match first: fewest explicit zeros per merged column
         std::sort(cand.begin(), cand.end(), [&](int32_t a, int32_t b) {
            const int64_t za = kk[a] + ff[s] - ff[a], zb = kk[b] + ff[s] - ff[b];
            if (force[a] != force[b]) return force[a] > force[b];
            if (za != zb) return za < zb;
            return a < b;
         });
         for (int32_t c : cand) {
            const int64_t kn = kk[c] + kk[s], fn = kk[c] + ff[s];
            const int64_t znew = zeros[c] + zeros[s] + kk[c] * (kk[c] + ff[s] - ff[c]);
            const double nnzl = static_cast<double>(kn) * static_cast<double>(fn) - 0.5 * static_cast<double>(kn) * static_cast<double>(kn - 1);
            const double ratio = static_cast<double>(znew) / nnzl;
            bool merge = force[c] != 0;
            if (!merge) {
               // the zeros this merge adds by itself, against the merged front: a nearly nested chain of wide fronts (many small
               // fronts hanging off a big separator) must not be kept apart by the zeros its leaves brought in earlier
               const double step = static_cast<double>(kk[c]) * static_cast<double>(kk[c] + ff[s] - ff[c]) / nnzl;
               if (kn <= opt.relax_small && fn <= opt.relax_small_front) merge = true;
               else if (kn <= opt.relax_mid) merge = ratio <= opt.relax_mid_zeros;
               else if (kn <= opt.relax_big) merge = ratio <= opt.relax_big_zeros;
               else merge = ratio <= opt.relax_any_zeros;
               if (!merge && fn > opt.relax_small_front && kn <= opt.relax_chain_max_k) merge = step <= opt.relax_chain_zeros;
            }
            if (!merge) continue;
            merged_into[c] = s;
            kk[s] = kn; ff[s] = fn; zeros[s] = znew;
            for (int32_t gc : kids[c]) { sparent[gc] = s; kids[s].push_back(gc); }
            kids[c].clear();
         }
      }
   }
   tra.mark("supernodes, amalgamation");
   // ---- final numbering: postorder of the merged tree, pairs first inside each supernode -------
   std::vector<int32_t> rep(static_cast<size_t>(ns0));
   for (int32_t s = ns0 - 1; s >= 0; --s) rep[s] = (merged_into[s] == -1) ? s : rep[merged_into[s]];
   std::vector<int32_t> final_order(static_cast<size_t>(N));
   {
      // members of each representative, ascending
      std::vector<int32_t> mcount(static_cast<size_t>(ns0) + 1, 0);
      for (int32_t s = 0; s < ns0; ++s) ++mcount[rep[s] + 1];
      for (int32_t s = 0; s < ns0; ++s) mcount[s + 1] += mcount[s];
      std::vector<int32_t> members(static_cast<size_t>(ns0)), mfill(mcount.begin(), mcount.end() - 1);
      for (int32_t s = 0; s < ns0; ++s) members[mfill[rep[s]]++] = s;
      // representatives are visited in ascending index = a postorder of the merged tree
      // (a representative is the topmost = largest-index member of its group, and children < parents)
      int32_t nsn = 0;
      int64_t out = 0;
      P.sn_first.clear();
      std::vector<int32_t> singles;
      for (int32_t s = 0; s < ns0; ++s) {
         if (rep[s] != s) continue;
         P.sn_first.push_back(static_cast<int32_t>(out));
         singles.clear();
         for (int32_t mi = mcount[s]; mi < mcount[s + 1]; ++mi) {
            const int32_t ms = members[mi];
            for (int32_t j = sfirst0[ms]; j < sfirst0[ms + 1]; ++j) {
               const int32_t oj = order[j];
               const int32_t pj = partner[oj];
               if (pj == -1 || rep[sn0_of[ipos[pj]]] != s) { singles.push_back(oj); continue; }
               if (!is_con[oj]) { final_order[out++] = oj; final_order[out++] = pj; }
            }
         }
         for (int32_t oj : singles) final_order[out++] = oj;
         ++nsn;
      }
      P.sn_first.push_back(static_cast<int32_t>(out));
      if (out != N) throw std::runtime_error("analyse: internal error (final numbering incomplete)");
      P.nsn = nsn;
   }
   P.perm = final_order;
   // kept for re-planning (delayed pivots): the adjacency and the COO pattern in 32-bit indices
   P.adj_ptr = std::move(g.ptr);
   P.adj_idx = std::move(g.idx);
   P.coo_row.resize(static_cast<size_t>(nnz));
   P.coo_col.resize(static_cast<size_t>(nnz));
   for (int64_t e = 0; e < nnz; ++e) { P.coo_row[e] = static_cast<int32_t>(row[e]); P.coo_col[e] = static_cast<int32_t>(col[e]); }
   tra.mark("final numbering");
   build_structures(P);
   P.t_symbolic = secs(t1, clk::now()) - P.t_maps;   // etree, supernodes and front structures; the maps are timed apart
}

// Everything that follows from the elimination order and the supernode partition (P.perm, P.sn_first, P.nsn): front row
// structures, extend-add maps, level schedule, pool layout, work figures, the assembly scatter map and the CSR of the residual.
// Valid for ANY partition of the order into contiguous column ranges (a front is treated as dense), which is what lets the
// numeric phase move a pivot column into an ancestor front (delay_columns) without a new ordering.
void build_structures(Plan& P) {
   const int64_t N = P.N, nnz = P.nnz;
   const auto t1 = clk::now();
   struct { const std::vector<int64_t>& ptr; const std::vector<int32_t>& idx; } g{P.adj_ptr, P.adj_idx};
   const std::vector<int32_t>& row = P.coo_row;
   const std::vector<int32_t>& col = P.coo_col;
   P.nnz_L = 0; P.flops = 0.0; P.flops_tc = 0.0; P.sum_r = 0; P.max_f = 0; P.max_k = 0;
   P.iperm.resize(static_cast<size_t>(N));
   for (int64_t i = 0; i < N; ++i) P.iperm[P.perm[i]] = static_cast<int32_t>(i);
   P.sn_of.resize(static_cast<size_t>(N));
   for (int32_t s = 0; s < P.nsn; ++s) for (int32_t j = P.sn_first[s]; j < P.sn_first[s + 1]; ++j) P.sn_of[j] = s;

   // ---- front row structures (children before parents) ----------------------------------------
   const int32_t nsn = P.nsn;
   P.k.resize(nsn); P.f.resize(nsn); P.ld.resize(nsn); P.parent.assign(nsn, -1);
   P.rows_off.assign(static_cast<size_t>(nsn) + 1, 0);
   P.rows.clear();
   {
      std::vector<int32_t> mark(static_cast<size_t>(N), -1);
      std::vector<std::vector<int32_t>> kids(static_cast<size_t>(nsn));
      std::vector<int32_t> bnd;
      for (int32_t s = 0; s < nsn; ++s) {
         const int32_t c0 = P.sn_first[s], c1 = P.sn_first[s + 1];
         bnd.clear();
         for (int32_t j = c0; j < c1; ++j) {
            const int32_t oj = P.perm[j];
            for (int64_t p = g.ptr[oj]; p < g.ptr[oj + 1]; ++p) {
               const int32_t i = P.iperm[g.idx[p]];
               if (i >= c1 && mark[i] != s) { mark[i] = s; bnd.push_back(i); }
            }
         }
         for (int32_t c : kids[s]) {
            const int64_t b0 = P.rows_off[c] + P.k[c], b1 = P.rows_off[c + 1];
            for (int64_t p = b0; p < b1; ++p) {
               const int32_t i = P.rows[p];
               if (i >= c1 && mark[i] != s) { mark[i] = s; bnd.push_back(i); }
            }
         }
         std::sort(bnd.begin(), bnd.end());
         P.k[s] = c1 - c0;
         P.f[s] = P.k[s] + static_cast<int32_t>(bnd.size());
         P.ld[s] = (P.f[s] + 3) & ~3;
         for (int32_t j = c0; j < c1; ++j) P.rows.push_back(j);
         P.rows.insert(P.rows.end(), bnd.begin(), bnd.end());
         P.rows_off[s + 1] = static_cast<int64_t>(P.rows.size());
         if (!bnd.empty()) {
            P.parent[s] = P.sn_of[bnd[0]];
            kids[P.parent[s]].push_back(s);
         }
      }
      P.child_ptr.assign(static_cast<size_t>(nsn) + 1, 0);
      for (int32_t s = 0; s < nsn; ++s) P.child_ptr[s + 1] = P.child_ptr[s] + static_cast<int32_t>(kids[s].size());
      P.child_idx.resize(static_cast<size_t>(P.child_ptr[nsn]));
      for (int32_t s = 0; s < nsn; ++s) std::copy(kids[s].begin(), kids[s].end(), P.child_idx.begin() + P.child_ptr[s]);
   }
   Trace tr;
   tr.t = t1;
   tr.mark("front row structures");
   // ---- extend-add maps -------------------------------------------------------------------------
   P.rel_off.assign(static_cast<size_t>(nsn) + 1, 0);
   for (int32_t s = 0; s < nsn; ++s) P.rel_off[s + 1] = P.rel_off[s] + (P.f[s] - P.k[s]);
   P.relmap.resize(static_cast<size_t>(P.rel_off[nsn]));
   for (int32_t s = 0; s < nsn; ++s) {
      const int32_t p = P.parent[s];
      if (p == -1) continue;
      const int32_t pc0 = P.sn_first[p], pc1 = P.sn_first[p + 1];
      const int32_t* pb = P.rows.data() + P.rows_off[p] + P.k[p];
      const int32_t* pe = P.rows.data() + P.rows_off[p + 1];
      const int32_t* cur = pb;
      for (int32_t i = 0; i < P.f[s] - P.k[s]; ++i) {
         const int32_t grow = P.rows[P.rows_off[s] + P.k[s] + i];
         int32_t loc;
         if (grow < pc1) loc = grow - pc0;
         else {
            cur = std::lower_bound(cur, pe, grow);
            if (cur == pe || *cur != grow) throw std::runtime_error("analyse: internal error (child row missing from parent front)");
            loc = P.k[p] + static_cast<int32_t>(cur - pb);
         }
         P.relmap[P.rel_off[s] + i] = loc;
      }
   }
   tr.mark("extend-add maps");
   // ---- level schedule ------------------------------------------------------------------------
   P.level.assign(nsn, 0);
   for (int32_t s = 0; s < nsn; ++s) {
      const int32_t p = P.parent[s];
      if (p != -1) P.level[p] = std::max(P.level[p], P.level[s] + 1);
   }
   build_level_lists(P);
   // ---- factor pool layout and work figures ----------------------------------------------------
   P.front_off.resize(nsn);
   P.ldc.resize(nsn);
   int64_t off = 0;
   for (int32_t s = 0; s < nsn; ++s) {
      P.front_off[s] = off;
      off += static_cast<int64_t>(P.ld[s]) * P.k[s];
      P.ldc[s] = P.f[s] - P.k[s];   // contribution blocks are packed lower triangles: no padding
      const double kd = P.k[s], rd = P.f[s] - P.k[s];
      P.nnz_L += static_cast<int64_t>(P.k[s]) * P.f[s] - static_cast<int64_t>(P.k[s]) * (P.k[s] - 1) / 2;
      const double fl = kd * kd * kd / 3.0 + kd * kd * rd + kd * rd * rd;
      P.flops += fl;
      if (P.k[s] >= 128) P.flops_tc += kd * kd * rd + kd * rd * rd;
      P.sum_r += P.f[s] - P.k[s];
      P.max_f = std::max(P.max_f, P.f[s]);
      P.max_k = std::max(P.max_k, P.k[s]);
   }
   P.pool_size = off;
   P.pool_size_global = off;
   tr.mark("levels, pool layout");
   build_cb_arena(P);
   tr.mark("contribution-block arena");
   P.cb_pool_size_global = P.cb_pool_size;
   const auto t2 = clk::now();
   P.t_symbolic = secs(t1, t2);

   // ---- assembly scatter map --------------------------------------------------------------------
   P.scatter.resize(static_cast<size_t>(nnz));
   std::vector<int32_t> front_count(static_cast<size_t>(nsn) + 1, 0);
   for (int64_t e = 0; e < nnz; ++e) {
      const int32_t pr = P.iperm[row[e]], pc = P.iperm[col[e]];
      const int32_t lo = std::min(pr, pc), hi = std::max(pr, pc);
      const int32_t s = P.sn_of[lo];
      const int32_t c1 = P.sn_first[s + 1];
      int32_t lrow;
      if (hi < c1) lrow = hi - P.sn_first[s];
      else {
         const int32_t* pb = P.rows.data() + P.rows_off[s] + P.k[s];
         const int32_t* pe = P.rows.data() + P.rows_off[s + 1];
         const int32_t* it = std::lower_bound(pb, pe, hi);
         if (it == pe || *it != hi) throw std::runtime_error("analyse: internal error (entry outside its front)");
         lrow = P.k[s] + static_cast<int32_t>(it - pb);
      }
      P.scatter[e] = P.front_off[s] + static_cast<int64_t>(lo - P.sn_first[s]) * P.ld[s] + lrow;
      ++front_count[static_cast<size_t>(s) + 1];
   }
   {
      // entries by destination, ties by slot: the destinations of a front are one range of the pool, so the entries are first
      // dealt to their fronts (in slot order), then ordered inside each front - short sorts that stay in cache
      std::vector<int32_t> idx(static_cast<size_t>(nnz));
      for (int32_t s = 0; s < nsn; ++s) front_count[static_cast<size_t>(s) + 1] += front_count[s];
      {
         std::vector<int32_t> fillp(front_count.begin(), front_count.end() - 1);
         for (int64_t e = 0; e < nnz; ++e) idx[static_cast<size_t>(fillp[P.sn_of[std::min(P.iperm[row[e]], P.iperm[col[e]])]]++)] = static_cast<int32_t>(e);
      }
      for (int32_t s = 0; s < nsn; ++s)
         std::stable_sort(idx.begin() + front_count[s], idx.begin() + front_count[static_cast<size_t>(s) + 1],
                          [&](int32_t a, int32_t b) { return P.scatter[a] < P.scatter[b]; });
      P.asm_src = idx;
      P.asm_row.resize(static_cast<size_t>(nnz));
      P.asm_col.resize(static_cast<size_t>(nnz));
      P.asm_dst.clear();
      P.asm_seg.clear();
      for (int64_t t = 0; t < nnz; ++t) {
         const int32_t e = idx[t];
         P.asm_row[t] = static_cast<int32_t>(row[e]);
         P.asm_col[t] = static_cast<int32_t>(col[e]);
         if (t == 0 || P.scatter[e] != P.scatter[idx[t - 1]]) {
            P.asm_dst.push_back(P.scatter[e]);
            P.asm_seg.push_back(static_cast<int32_t>(t));
         }
      }
      P.asm_seg.push_back(static_cast<int32_t>(nnz));
   }
   // ---- symmetric CSR (original numbering, duplicates kept, slot order) ------------------------
   P.csr_ptr.assign(static_cast<size_t>(N) + 1, 0);
   for (int64_t e = 0; e < nnz; ++e) {
      ++P.csr_ptr[row[e] + 1];
      if (row[e] != col[e]) ++P.csr_ptr[col[e] + 1];
   }
   for (int64_t i = 0; i < N; ++i) P.csr_ptr[i + 1] += P.csr_ptr[i];
   P.csr_col.resize(static_cast<size_t>(P.csr_ptr[N]));
   P.csr_slot.resize(static_cast<size_t>(P.csr_ptr[N]));
   {
      std::vector<int32_t> fillp(P.csr_ptr.begin(), P.csr_ptr.end() - 1);
      for (int64_t e = 0; e < nnz; ++e) {
         const int32_t r = static_cast<int32_t>(row[e]), c = static_cast<int32_t>(col[e]);
         P.csr_col[fillp[r]] = c; P.csr_slot[fillp[r]++] = static_cast<int32_t>(e);
         if (r != c) { P.csr_col[fillp[c]] = r; P.csr_slot[fillp[c]++] = static_cast<int32_t>(e); }
      }
   }
   P.t_maps = secs(t2, clk::now());
   tr.mark("scatter map, CSR");
}

// Delayed pivots.  Column cols[i] (a position in the current numbering) leaves its front and becomes the LAST pivot of the
// ancestor front up[i] levels above (capped at the root of its tree; a column of a root front stays).  Fronts that lose all
// their pivots disappear.  Everything derived from the partition is rebuilt; the elimination order of all other columns is kept.
void delay_columns(Plan& P, const std::vector<int32_t>& cols, const std::vector<int32_t>& up) {
   const int32_t nsn = P.nsn;
   std::vector<std::vector<int32_t>> arrivals(static_cast<size_t>(nsn));
   std::vector<char> leaves(static_cast<size_t>(P.N), 0);
   for (size_t i = 0; i < cols.size(); ++i) {
      const int32_t q = cols[i];
      if (q < 0 || q >= P.N) throw std::runtime_error("delay_columns: position out of range");
      if (leaves[q]) continue;
      int32_t t = P.sn_of[q];
      for (int32_t u = 0; u < std::max(1, up[i]) && P.parent[t] >= 0; ++u) t = P.parent[t];
      if (t == P.sn_of[q]) continue;   // root front: nowhere to go
      leaves[q] = 1;
      arrivals[t].push_back(P.perm[q]);
   }
   std::vector<int32_t> order;
   order.reserve(static_cast<size_t>(P.N));
   std::vector<int32_t> first;
   for (int32_t s = 0; s < nsn; ++s) {
      const size_t before = order.size();
      for (int32_t q = P.sn_first[s]; q < P.sn_first[s + 1]; ++q) if (!leaves[q]) order.push_back(P.perm[q]);
      order.insert(order.end(), arrivals[s].begin(), arrivals[s].end());
      if (order.size() > before) first.push_back(static_cast<int32_t>(before));
   }
   if (static_cast<int64_t>(order.size()) != P.N) throw std::runtime_error("delay_columns: internal error (order incomplete)");
   first.push_back(static_cast<int32_t>(P.N));
   P.perm = std::move(order);
   P.sn_first = std::move(first);
   P.nsn = static_cast<int32_t>(P.sn_first.size()) - 1;
   build_structures(P);
}

// ------------------------------------------------------------------------------------------------
// Subtree-to-GPU mapping: independent elimination subtrees go to different ranks,
// the separators above the cut to the owner of their heaviest child.
// ------------------------------------------------------------------------------------------------
void map_subtrees(Plan& P, int parts, bool share_top_fronts) {
   const int32_t nsn = P.nsn;
   parts = std::max(1, parts);
   P.num_parts = parts;
   P.owner.assign(static_cast<size_t>(nsn), 0);
   P.shared.assign(static_cast<size_t>(nsn), 0);
   P.cut_edges.clear();
   P.part_flops.assign(static_cast<size_t>(parts), 0.0);
   std::vector<double> own(static_cast<size_t>(nsn)), sub(static_cast<size_t>(nsn));
   for (int32_t s = 0; s < nsn; ++s) {
      const double kd = P.k[s], rd = P.f[s] - P.k[s];
      own[s] = kd * kd * kd / 3.0 + kd * kd * rd + kd * rd * rd + 1.0;   // + 1: empty fronts still cost a launch slot
      sub[s] = own[s];
   }
   // children precede parents in the final numbering (postorder)
   for (int32_t s = 0; s < nsn; ++s) if (P.parent[s] >= 0) sub[P.parent[s]] += sub[s];
   if (parts > 1 && nsn > 0) {
      // open the tree from the roots: `open` holds the roots of the subtrees below the cut
      auto heavier = [&](int32_t a, int32_t b) { return sub[a] != sub[b] ? sub[a] < sub[b] : a > b; };   // max-heap on weight, ties by index
      std::vector<int32_t> open;
      std::vector<char> above(static_cast<size_t>(nsn), 0);
      double total = 0.0;
      for (int32_t s = 0; s < nsn; ++s) if (P.parent[s] < 0) { open.push_back(s); total += sub[s]; }
      std::make_heap(open.begin(), open.end(), heavier);
      double open_weight = total;
      // stop when the heaviest open subtree fits its share (with 5 % slack) or cannot be opened any further
      while (!open.empty()) {
         const int32_t top = open.front();
         const bool enough = static_cast<int>(open.size()) >= parts && sub[top] <= P.partition_slack * open_weight / parts;
         if (enough || P.child_ptr[top + 1] == P.child_ptr[top]) break;
         std::pop_heap(open.begin(), open.end(), heavier);
         open.pop_back();
         above[top] = 1;
         open_weight -= own[top];
         for (int32_t ci = P.child_ptr[top]; ci < P.child_ptr[top + 1]; ++ci) {
            open.push_back(P.child_idx[ci]);
            std::push_heap(open.begin(), open.end(), heavier);
         }
      }
      // LPT packing of the open subtrees that are worth a transfer; a light one (under 1 % of an even share) hanging off a
      // separator stays with that separator
      std::sort(open.begin(), open.end(), [&](int32_t a, int32_t b) { return sub[a] != sub[b] ? sub[a] > sub[b] : a < b; });
      std::vector<double> load(static_cast<size_t>(parts), 0.0);
      std::fill(P.owner.begin(), P.owner.end(), -1);
      const double light = 0.01 * total / parts;
      for (const int32_t s : open) {
         if (sub[s] < light && P.parent[s] >= 0) continue;
         int best = 0;
         for (int p = 1; p < parts; ++p) if (load[p] < load[best]) best = p;
         load[best] += sub[s];
         P.owner[s] = best;
      }
      // bottom-up: a front above the cut goes where its heaviest placed child is
      for (int32_t s = 0; s < nsn; ++s) {
         if (!above[s]) continue;
         int32_t best = -1;
         for (int32_t ci = P.child_ptr[s]; ci < P.child_ptr[s + 1]; ++ci) {
            const int32_t c = P.child_idx[ci];
            if (P.owner[c] >= 0 && (best < 0 || sub[c] > sub[best])) best = c;
         }
         if (best >= 0) P.owner[s] = P.owner[best];
      }
      if (share_top_fronts)
         for (int32_t s = 0; s < nsn; ++s)
            if (above[s] && P.k[s] >= P.share_min_pivots && own[s] >= P.share_min_flops) P.shared[s] = 1;
      // top-down (parents have higher numbers): everything not placed yet follows its parent
      for (int32_t s = nsn - 1; s >= 0; --s) {
         if (P.owner[s] >= 0) continue;
         if (P.parent[s] >= 0) { P.owner[s] = P.owner[P.parent[s]]; continue; }
         int best = 0;
         for (int p = 1; p < parts; ++p) if (load[p] < load[best]) best = p;
         load[best] += sub[s];
         P.owner[s] = best;
      }
   }
   // Level schedule of a partitioned run: every rank walks the levels in ascending order, so a shared front must not sit at a
   // lower level than some rank's unfinished private subtree (that rank would stop for the shared front, the others would wait
   // for its subtree afterwards).  Fronts with a shared front at or below them are lifted above all purely private fronts.
   if (parts > 1) {
      std::vector<char> top(static_cast<size_t>(nsn), 0);
      bool any = false;
      for (int32_t s = 0; s < nsn; ++s) {
         if (P.shared[s]) top[s] = 1;
         if (top[s]) { any = true; if (P.parent[s] >= 0) top[P.parent[s]] = 1; }
      }
      if (any) {
         int32_t lsub = -1;
         for (int32_t s = 0; s < nsn; ++s) if (!top[s]) lsub = std::max(lsub, P.level[s]);
         for (int32_t s = 0; s < nsn; ++s) {   // children first
            if (!top[s]) continue;
            int32_t lv = lsub + 1;
            for (int32_t ci = P.child_ptr[s]; ci < P.child_ptr[s + 1]; ++ci) lv = std::max(lv, P.level[P.child_idx[ci]] + 1);
            P.level[s] = lv;
         }
         build_level_lists(P);
         build_cb_arena(P);
      }
   }
   // cut edges in the order every rank walks them, work per rank, critical path
   for (int32_t s = 0; s < nsn; ++s) {
      if (P.shared[s]) for (int p = 0; p < parts; ++p) P.part_flops[p] += (own[s] - 1.0) / parts;
      else P.part_flops[P.owner[s]] += own[s] - 1.0;
      if (P.parent[s] >= 0 && P.owner[P.parent[s]] != P.owner[s]) P.cut_edges.push_back(s);
   }
   // critical path estimate: a front can start when all its children are done; children on the same rank run one after the
   // other, children on different ranks side by side:  finish(s) = flops(s) + max over ranks q of sum of finish(children on q)
   {
      P.critical_flops = 0.0;
      std::vector<double> finish(static_cast<size_t>(nsn), 0.0);
      std::vector<double> by_rank(static_cast<size_t>(parts), 0.0);
      for (int32_t s = 0; s < nsn; ++s) {
         double longest = 0.0;
         if (P.child_ptr[s + 1] > P.child_ptr[s]) {
            for (int32_t ci = P.child_ptr[s]; ci < P.child_ptr[s + 1]; ++ci) by_rank[P.owner[P.child_idx[ci]]] = 0.0;
            for (int32_t ci = P.child_ptr[s]; ci < P.child_ptr[s + 1]; ++ci) {
               const int32_t c = P.child_idx[ci];
               by_rank[P.owner[c]] += finish[c];
               longest = std::max(longest, by_rank[P.owner[c]]);
            }
         }
         finish[s] = (own[s] - 1.0) / (P.shared[s] ? parts : 1) + longest;
         if (P.parent[s] < 0) P.critical_flops = std::max(P.critical_flops, finish[s]);
      }
      // never shorter than the busiest rank's own load
      for (const double w : P.part_flops) P.critical_flops = std::max(P.critical_flops, w);
   }
   std::stable_sort(P.cut_edges.begin(), P.cut_edges.end(), [&](int32_t a, int32_t b) {
      return P.level[a] != P.level[b] ? P.level[a] < P.level[b] : a < b;
   });
}

// Per-rank storage of a partitioned run (after map_subtrees).  A rank stores the factor columns of the fronts it owns and of
// the shared fronts (every rank keeps those: it applies them in the solves), and holds a contribution block only if it
// produces it, consumes it (parent owned here) or takes part in its broadcast (front or parent shared).  The factor pool is
// compacted over the stored fronts - their order is kept, so the assembly destinations stay ascending - and the scatter map
// loses the entries of the other fronts.  A front without storage gets an offset far outside the pool: touching it faults.
void localize_storage(Plan& P, int rank) {
   const int32_t nsn = P.nsn;
   if (P.num_parts <= 1 || nsn == 0) return;
   std::vector<char> hold(static_cast<size_t>(nsn)), cbkeep(static_cast<size_t>(nsn));
   for (int32_t s = 0; s < nsn; ++s) {
      hold[s] = (P.owner[s] == rank || P.shared[s]) ? 1 : 0;
      const int32_t p = P.parent[s];
      cbkeep[s] = (hold[s] || (p >= 0 && (P.shared[p] || P.owner[p] == rank))) ? 1 : 0;
   }
   const std::vector<int64_t> old_off = P.front_off;
   constexpr int64_t NO_STORAGE = -(int64_t(1) << 45);
   int64_t off = 0;
   for (int32_t s = 0; s < nsn; ++s) {
      if (!hold[s]) { P.front_off[s] = NO_STORAGE; continue; }
      P.front_off[s] = off;
      off += static_cast<int64_t>(P.ld[s]) * P.k[s];
   }
   P.pool_size = off;
   // assembly map: segments of the fronts stored here, re-based
   {
      std::vector<int64_t> dst;
      std::vector<int32_t> seg, src, arow, acol;
      const size_t nd = P.asm_dst.size();
      int32_t s = 0;
      for (size_t d = 0; d < nd; ++d) {
         const int64_t o = P.asm_dst[d];
         while (s + 1 < nsn && old_off[s + 1] <= o) ++s;          // destinations ascend; empty fronts share an offset with their successor
         if (!hold[s]) continue;
         dst.push_back(o - old_off[s] + P.front_off[s]);
         seg.push_back(static_cast<int32_t>(src.size()));
         for (int32_t t = P.asm_seg[d]; t < P.asm_seg[d + 1]; ++t) {
            src.push_back(P.asm_src[t]);
            arow.push_back(P.asm_row[t]);
            acol.push_back(P.asm_col[t]);
         }
      }
      seg.push_back(static_cast<int32_t>(src.size()));
      P.asm_dst.swap(dst); P.asm_seg.swap(seg); P.asm_src.swap(src); P.asm_row.swap(arow); P.asm_col.swap(acol);
      for (int64_t e = 0; e < P.nnz; ++e) {     // the per-entry view of the same map (b200kkt_get_permutation)
         const int32_t fs = P.sn_of[std::min(P.iperm[P.coo_row[e]], P.iperm[P.coo_col[e]])];
         P.scatter[e] = hold[fs] ? P.scatter[e] - old_off[fs] + P.front_off[fs] : -1;
      }
   }
   P.cb_pool_size_global = P.cb_pool_size;   // map_subtrees may have re-levelled the shared fronts
   build_cb_arena(P, &cbkeep);
}

}  // namespace b200kkt
