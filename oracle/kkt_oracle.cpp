// ORACLE — TEST INFRASTRUCTURE ONLY.  Never linked into, imported by or called from the product
// (uno_b200/).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs may build or call this file.
//
// CPU restatement of the reference's numeric path: what dmumps_c does for
// MUMPSSolver::do_symbolic_analysis / do_numerical_factorization / solve_indefinite_system /
// get_inertia (reference: uno/ingredients/subproblem_solvers/MUMPS/MUMPSSolver.cpp:72-96,124-147).
// The arithmetic itself lives in MUMPS 5.8.0 (pinned by .github/workflows/unit-tests-ubuntu.yml:55),
// a Fortran dependency that is NOT vendored under /root/reference and cannot be built here (no
// gfortran).  This file therefore restates MUMPS' *published* algorithm (Amestoy, Duff, L'Excellent,
// "Multifrontal parallel distributed symmetric and unsymmetric solvers", CMAME 2000; Duff & Reid,
// "The multifrontal solution of indefinite sparse symmetric linear systems", TOMS 1983):
//   * symmetric equilibration of the input (ICNTL(8): MUMPSSolver.cpp:82),
//   * elimination tree + supernodes of the permuted pattern (ordering handed in, as with
//     ICNTL(7)=1/PERM_IN, or METIS_NodeND when none is given),
//   * multifrontal LDL^T (SYM=2: MUMPSSolver.cpp:17) with 1x1 and 2x2 pivots, threshold partial
//     pivoting (CNTL(1), default 0.01) tested against the WHOLE front column, and delayed pivots
//     passed to the parent front,
//   * null-pivot detection at the root (ICNTL(24)=1: MUMPSSolver.cpp:36) -> INFOG(28),
//     negative pivot count -> INFOG(12) (MUMPSSolver.cpp:131-139),
//   * forward / diagonal / backward solve without iterative refinement (ICNTL(10)=0).
// Parity pins: the three known-answer tests the reference keeps (commented out) in
// unotest/functional_tests/MUMPSSolverTests.cpp:11-83 are checked in tests/test_oracle.py.
// Parity against a real MUMPS binary is UNPINNED (none exists in this environment).
//
// It deliberately shares no code with the product: own symbolic phase, own pivoting (stronger:
// full-column threshold test + delays, where the GPU path confines the search to 32x32 blocks).
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <dlfcn.h>
#include <map>
#include <memory>
#include <numeric>
#include <stdexcept>
#include <string>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

extern "C" {
int METIS_NodeND(int64_t*, int64_t*, int64_t*, int64_t*, int64_t*, int64_t*, int64_t*);
int METIS_SetDefaultOptions(int64_t*);
}

namespace {

// ---- optional BLAS (OpenBLAS shipped inside scipy), loaded at run time ---------------------------
typedef void (*cblas_dgemm_t)(int, int, int, int, int, int, double, const double*, int, const double*, int, double, double*, int);
typedef void (*cblas_dtrsm_t)(int, int, int, int, int, int, int, double, const double*, int, double*, int);
typedef void (*set_threads_t)(int);
cblas_dgemm_t g_dgemm = nullptr;
cblas_dtrsm_t g_dtrsm = nullptr;
set_threads_t g_set_threads = nullptr;

// C[m x n] -= A[m x k] * B[n x k]^T, all column-major
void gemm_nt_sub(int m, int n, int k, const double* A, int lda, const double* B, int ldb, double* C, int ldc) {
   if (m <= 0 || n <= 0 || k <= 0) return;
   // a BLAS call costs a thread wake-up: only products that are worth it go there
   if (g_dgemm != nullptr && static_cast<double>(m) * n * k >= 2.0e6) {
      g_dgemm(102 /*ColMajor*/, 111 /*NoTrans*/, 112 /*Trans*/, m, n, k, -1.0, A, lda, B, ldb, 1.0, C, ldc);
      return;
   }
   for (int j = 0; j < n; ++j) {
      double* c = C + static_cast<size_t>(ldc) * j;
      for (int p = 0; p < k; ++p) {
         const double b = B[j + static_cast<size_t>(ldb) * p];
         if (b == 0.0) continue;
         const double* a = A + static_cast<size_t>(lda) * p;
         for (int i = 0; i < m; ++i) c[i] -= a[i] * b;
      }
   }
}

// B[m x n] <- B * T^{-T}, T unit lower triangular n x n (its diagonal is not read), all column-major
void trsm_right_lower_trans_unit(int m, int n, const double* T, int ldt, double* B, int ldb) {
   if (m <= 0 || n <= 0) return;
   if (g_dtrsm != nullptr && static_cast<double>(m) * n * n >= 2.0e6) {
      g_dtrsm(102 /*ColMajor*/, 142 /*Right*/, 122 /*Lower*/, 112 /*Trans*/, 132 /*Unit*/, m, n, 1.0, T, ldt, B, ldb);
      return;
   }
   // X T^T = B: column p of X is final once columns q < p are; then it leaves all later columns
   for (int p = 0; p < n; ++p) {
      const double* xp = B + static_cast<size_t>(ldb) * p;
      for (int q = p + 1; q < n; ++q) {
         const double t = T[q + static_cast<size_t>(ldt) * p];
         if (t == 0.0) continue;
         double* xq = B + static_cast<size_t>(ldb) * q;
         for (int i = 0; i < m; ++i) xq[i] -= xp[i] * t;
      }
   }
}

#ifdef _OPENMP
inline bool in_team() { return omp_in_parallel() != 0; }
#else
inline bool in_team() { return true; }
#endif

// The same product cut into tiles that the OpenMP team works through, single-threaded BLAS inside each tile: ONE thread pool.
// (A threaded BLAS keeps its own workers spinning after every call, and they fight the OpenMP loops around it for the cores.)
// lower: C is a diagonal block of the front (rows and columns from the same origin) and only its lower part is needed.
void gemm_nt_sub_team(int m, int n, int k, const double* A, int lda, const double* B, int ldb, double* C, int ldc, bool lower) {
   if (m <= 0 || n <= 0 || k <= 0) return;
   const int TM = 768, TN = 256;
   const int nj = (n + TN - 1) / TN, ni = (m + TM - 1) / TM;
   const bool team = !in_team() && static_cast<double>(m) * n * k >= 4.0e7;
#pragma omp parallel for collapse(2) schedule(dynamic, 1) if (team)
   for (int tj = 0; tj < nj; ++tj) {
      for (int ti = 0; ti < ni; ++ti) {
         const int j0 = tj * TN, j1 = std::min(n, j0 + TN);
         int i0 = ti * TM;
         const int i1 = std::min(m, i0 + TM);
         if (lower) { if (i1 <= j0) continue; i0 = std::max(i0, j0); }
         gemm_nt_sub(i1 - i0, j1 - j0, k, A + i0, lda, B + j0, ldb, C + i0 + static_cast<size_t>(ldc) * j0, ldc);
      }
   }
}

// rows of B are independent in B T^{-T}: the team takes them in chunks
void trsm_right_lower_trans_unit_team(int m, int n, const double* T, int ldt, double* B, int ldb) {
   if (m <= 0 || n <= 0) return;
   const int TM = 512;
   const int ni = (m + TM - 1) / TM;
   const bool team = !in_team() && static_cast<double>(m) * n * n >= 4.0e7;
#pragma omp parallel for schedule(dynamic, 1) if (team)
   for (int ti = 0; ti < ni; ++ti) trsm_right_lower_trans_unit(std::min(TM, m - ti * TM), n, T, ldt, B + ti * TM, ldb);
}

struct Front {
   // static structure
   int first = 0, ncol = 0;            // own columns [first, first+ncol) in elimination numbering
   std::vector<int> bnd;               // static boundary rows (sorted)
   int parent = -1;
   std::vector<int> children;
   // numeric result
   int nf = 0, nelim = 0;
   std::vector<int> idx;               // global ids of the front rows in final local order
   std::unique_ptr<double[]> L;        // columns of L packed one after the other, column j holding rows j .. nf-1 (unit diagonal implied)
   size_t L_size = 0;
   std::vector<double> d, e;           // D blocks: d diagonal, e[j] != 0 -> 2x2 block (j, j+1)
   std::vector<char> ptype;            // 0 1x1, 1/2 first/second of 2x2, 3 null
   // contribution block handed to the parent (freed once consumed)
   int ncb = 0, ndelay = 0;
   std::vector<int> cb_idx;
   std::unique_ptr<double[]> cb;       // packed lower triangle of the ncb x ncb block, by columns; from / back to the buffer pool
   size_t cb_cap = 0;
};

// Contribution blocks come and go with every front: fresh allocations of that size are mmap'ed and page-faulted in each time,
// which costs more than the arithmetic.  Buffers are recycled instead (best fit by capacity).
// One pool per thread (a block is taken by the thread that factors the front and given to the pool of the thread that consumes
// it); big blocks are not kept - their allocation is nothing next to the work on such a front, and the memory goes back at once.
struct BufferPool {
   static constexpr size_t KEEP_BELOW = size_t(1) << 22;   // doubles (32 MB)
   std::multimap<size_t, std::unique_ptr<double[]>> free_list;
   std::unique_ptr<double[]> take(size_t need, size_t& cap) {
      std::unique_ptr<double[]> out;
      auto it = free_list.lower_bound(need);
      if (it != free_list.end() && it->first <= 2 * need + 4096) { cap = it->first; out = std::move(it->second); free_list.erase(it); }
      if (!out) { cap = need; out.reset(new double[need > 0 ? need : 1]); }
      return out;
   }
   void give(std::unique_ptr<double[]>& buf, size_t cap) {
      if (!buf) return;
      if (cap >= KEEP_BELOW) { buf.reset(); return; }
      free_list.emplace(cap, std::move(buf));
   }
};

// offset of column j of a packed factor block of nf rows: entry (i, j), i >= j, sits at lpos(nf, j) + i
inline size_t lpos(int nf, int j) { return static_cast<size_t>(j) * nf - static_cast<size_t>(j) * (j + 1) / 2; }

// per-thread scratch, kept from one factorization to the next.  The dense buffers are raw memory: they are never cleared as a
// whole (only the lower triangle of a front is), and they grow geometrically so that a rising sequence of front sizes does not
// re-allocate (and page-fault) at every step.
struct Work {
   std::vector<int> loc;
   BufferPool pool;
   std::unique_ptr<double[]> A, W;
   size_t capA = 0, capW = 0;
   static double* ensure(std::unique_ptr<double[]>& buf, size_t& cap, size_t need) {
      if (need > cap) { cap = need + (need < (size_t(1) << 27) ? need / 2 : need / 16); buf.reset(); buf.reset(new double[cap]); }
      return buf.get();
   }
};

struct Oracle {
   int64_t N = 0, nnz = 0;
   std::vector<int> perm, iperm;       // perm[new] = old
   std::vector<int64_t> cptr;          // permuted lower CSC of the pattern with slot references
   std::vector<int> crow, cslot;
   std::vector<int> urow, ucol;        // original COO
   std::vector<Front> fronts;
   std::vector<int> order;             // fronts, children before parents
   std::vector<std::vector<int>> levels;
   std::vector<double> scale;
   std::vector<Work> works;
   std::vector<int> first_desc;        // fronts are numbered in postorder: the subtree of s is [first_desc[s], s]
   std::vector<double> sub_flops;      // static estimate of the work in the subtree of s
   std::vector<int> sub_maxnf;         // largest front in it
   double min_free_gb = 12.0;          // ORACLE_MIN_FREE_GB: a factorization stops (with an error) before the host runs out of memory
   uint64_t pos = 0, neg = 0, zero = 0;
   int64_t n2x2 = 0, ndelayed = 0;
   double nnzL = 0, flops = 0, t_analyse = 0, t_factor = 0, t_solve = 0;
   double u = 0.01, null_tol = 1e-20;
   // 0: threshold test against the whole front column (MUMPS' rule, many delayed pivots on KKT matrices);
   // 1: against the fully-summed rows only ("restricted" pivoting: a column is delayed only when the fully-summed block has no
   //    acceptable pivot for it) - used where mode 0 is too slow (BASELINE sizes) and as the same-flops CPU baseline
   int pivot_mode = 0;
   std::string err;
};

}  // namespace

namespace {

using clk = std::chrono::steady_clock;
double since(clk::time_point t0) { return std::chrono::duration<double>(clk::now() - t0).count(); }

// ORACLE_PROFILE=1: where the wall clock of a factorization goes (big fronts run one at a time, so plain sums are exact for them)
enum { T_SIDE, T_ASM, T_EXTADD, T_PIVOT, T_INNER_GEMM, T_ROWS, T_OUTER_GEMM, T_COPY, T_COUNT };
double g_prof[T_COUNT];
double g_prof_cls[5]; long g_prof_cnt[5];
double g_prof_nf2 = 0, g_prof_nfne = 0, g_prof_cb2 = 0; int g_prof_big = 0;
bool g_prof_on = false;
struct Tick {
   int cat; clk::time_point t0; bool on;
   Tick(int c, bool enable) : cat(c), on(enable && g_prof_on) { if (on) t0 = clk::now(); }
   ~Tick() { if (on) g_prof[cat] += since(t0); }
};

// ---- symbolic phase (independent of the product's) -----------------------------------------------
void analyse(Oracle& o, int64_t N, int64_t nnz, const uint64_t* row, const uint64_t* col, const int64_t* user_perm) {
   const auto t0 = clk::now();
   o.N = N; o.nnz = nnz;
   o.urow.resize(nnz); o.ucol.resize(nnz);
   std::vector<std::vector<int>> adj(N);
   for (int64_t e = 0; e < nnz; ++e) {
      if (row[e] >= static_cast<uint64_t>(N) || col[e] >= static_cast<uint64_t>(N) || row[e] > col[e]) throw std::runtime_error("oracle: bad COO entry");
      o.urow[e] = static_cast<int>(row[e]); o.ucol[e] = static_cast<int>(col[e]);
      if (row[e] != col[e]) { adj[row[e]].push_back(o.ucol[e]); adj[col[e]].push_back(o.urow[e]); }
   }
   for (auto& a : adj) { std::sort(a.begin(), a.end()); a.erase(std::unique(a.begin(), a.end()), a.end()); }
   std::vector<int> perm(N);
   if (user_perm != nullptr) {
      for (int64_t i = 0; i < N; ++i) perm[i] = static_cast<int>(user_perm[i]);
   } else {
      std::iota(perm.begin(), perm.end(), 0);
      std::vector<int64_t> xadj(N + 1, 0), adjncy;
      for (int64_t i = 0; i < N; ++i) { for (int v : adj[i]) adjncy.push_back(v); xadj[i + 1] = static_cast<int64_t>(adjncy.size()); }
      if (N > 2 && !adjncy.empty()) {
         std::vector<int64_t> opt(40), mp(N), mip(N);
         METIS_SetDefaultOptions(opt.data());
         int64_t n = N;
         if (METIS_NodeND(&n, xadj.data(), adjncy.data(), nullptr, opt.data(), mp.data(), mip.data()) == 1)
            for (int64_t i = 0; i < N; ++i) perm[i] = static_cast<int>(mp[i]);
      }
   }
   std::vector<int> ip(N);
   for (int64_t i = 0; i < N; ++i) ip[perm[i]] = static_cast<int>(i);
   // elimination tree (Liu 1990)
   std::vector<int> parent(N, -1), anc(N, -1);
   for (int j = 0; j < N; ++j)
      for (int v : adj[perm[j]]) {
         int i = ip[v];
         while (i != -1 && i < j) { const int nx = anc[i]; anc[i] = j; if (nx == -1) parent[i] = j; i = nx; }
      }
   // postorder (children in ascending order), then relabel
   {
      std::vector<int> head(N, -1), next(N, -1), post; post.reserve(N);
      for (int j = static_cast<int>(N) - 1; j >= 0; --j) if (parent[j] != -1) { next[j] = head[parent[j]]; head[parent[j]] = j; }
      std::vector<int> st;
      for (int r = 0; r < N; ++r) {
         if (parent[r] != -1) continue;
         st.push_back(r);
         while (!st.empty()) {
            const int p = st.back(), c = head[p];
            if (c == -1) { st.pop_back(); post.push_back(p); } else { head[p] = next[c]; st.push_back(c); }
         }
      }
      std::vector<int> ipost(N), perm2(N), parent2(N);
      for (int k = 0; k < N; ++k) ipost[post[k]] = k;
      for (int k = 0; k < N; ++k) { perm2[k] = perm[post[k]]; parent2[k] = parent[post[k]] == -1 ? -1 : ipost[parent[post[k]]]; }
      perm.swap(perm2); parent.swap(parent2);
      for (int64_t i = 0; i < N; ++i) ip[perm[i]] = static_cast<int>(i);
   }
   o.perm = perm; o.iperm = ip;
   // column counts (Gilbert, Ng, Peyton 1994) on the postordered tree
   std::vector<int> cc(N, 0);
   {
      std::vector<int> first(N, -1), maxfirst(N, -1), prevleaf(N, -1), uf(N);
      for (int k = 0; k < N; ++k) { int j = k; cc[j] = first[j] == -1 ? 1 : 0; for (; j != -1 && first[j] == -1; j = parent[j]) first[j] = k; }
      std::iota(uf.begin(), uf.end(), 0);
      for (int j = 0; j < N; ++j) {
         if (parent[j] != -1) --cc[parent[j]];
         for (int v : adj[perm[j]]) {
            const int i = ip[v];
            if (i <= j || first[j] <= maxfirst[i]) continue;
            maxfirst[i] = first[j];
            const int jp = prevleaf[i];
            prevleaf[i] = j;
            ++cc[j];
            if (jp != -1) {
               int q = jp;
               while (q != uf[q]) q = uf[q];
               for (int s = jp; s != q;) { const int sp = uf[s]; uf[s] = q; s = sp; }
               --cc[q];
            }
         }
         if (parent[j] != -1) uf[j] = parent[j];
      }
      for (int j = 0; j < N; ++j) if (parent[j] != -1) cc[parent[j]] += cc[j];
   }
   // fundamental supernodes, then merge short chains (last child only: numbering stays contiguous)
   std::vector<int> nchild(N, 0);
   for (int j = 0; j < N; ++j) if (parent[j] != -1) ++nchild[parent[j]];
   std::vector<int> sfirst;
   double sn_zeros = 0.0;   // explicit zeros the current supernode already stores
   for (int j = 0; j < N; ++j) {
      bool glue = false;
      if (j > 0 && parent[j - 1] == j && !sfirst.empty()) {
         // column j joins the supernode that ends at its child j-1.  struct(j) holds struct(j-1) \ {j-1}; the `extra` rows it
         // brings on top (from its other children or its own entries) become explicit zeros in every column already there.
         const double k_prev = j - sfirst.back();
         const double extra = cc[j] - (cc[j - 1] - 1);
         const double znew = sn_zeros + k_prev * extra;
         const double kn = k_prev + 1.0, fn = k_prev + cc[j];
         const double nnz_merged = kn * fn - 0.5 * kn * (kn - 1.0);
         if (extra == 0.0) glue = true;                                   // nested: a fundamental supernode (or as good as one)
         else if (kn <= 24 && extra <= 8) glue = true;                    // small blocks: anything for fewer, wider fronts
         else if (znew <= 0.12 * nnz_merged) glue = true;                 // relaxed amalgamation: separators broken up by small side trees
         if (glue) sn_zeros = znew;
      }
      if (!glue) { sfirst.push_back(j); sn_zeros = 0.0; }
   }
   sfirst.push_back(static_cast<int>(N));
   const int ns = static_cast<int>(sfirst.size()) - 1;
   std::vector<int> snof(N);
   for (int s = 0; s < ns; ++s) for (int j = sfirst[s]; j < sfirst[s + 1]; ++j) snof[j] = s;
   o.fronts.clear();
   o.fronts.resize(ns);
   std::vector<int> mark(N, -1);
   for (int s = 0; s < ns; ++s) {
      Front& F = o.fronts[s];
      F.first = sfirst[s]; F.ncol = sfirst[s + 1] - sfirst[s];
      const int c1 = sfirst[s + 1];
      for (int j = F.first; j < c1; ++j)
         for (int v : adj[perm[j]]) { const int i = ip[v]; if (i >= c1 && mark[i] != s) { mark[i] = s; F.bnd.push_back(i); } }
      for (int c : F.children)
         for (int i : o.fronts[c].bnd) if (i >= c1 && mark[i] != s) { mark[i] = s; F.bnd.push_back(i); }
      std::sort(F.bnd.begin(), F.bnd.end());
      if (!F.bnd.empty()) { F.parent = snof[F.bnd[0]]; o.fronts[F.parent].children.push_back(s); }
   }
   // levels (children before parents) for the threaded numeric phase
   std::vector<int> lev(ns, 0);
   int nlev = 0;
   for (int s = 0; s < ns; ++s) { if (o.fronts[s].parent != -1) lev[o.fronts[s].parent] = std::max(lev[o.fronts[s].parent], lev[s] + 1); }
   for (int s = 0; s < ns; ++s) nlev = std::max(nlev, lev[s] + 1);
   o.levels.assign(nlev, {});
   for (int s = 0; s < ns; ++s) o.levels[lev[s]].push_back(s);
   // permuted lower-triangular CSC with references to the COO slots (duplicates kept)
   o.cptr.assign(N + 1, 0);
   for (int64_t e = 0; e < nnz; ++e) ++o.cptr[std::min(ip[o.urow[e]], ip[o.ucol[e]]) + 1];
   for (int64_t j = 0; j < N; ++j) o.cptr[j + 1] += o.cptr[j];
   o.crow.resize(nnz); o.cslot.resize(nnz);
   std::vector<int64_t> fill(o.cptr.begin(), o.cptr.end() - 1);
   for (int64_t e = 0; e < nnz; ++e) {
      const int a = ip[o.urow[e]], b = ip[o.ucol[e]];
      const int64_t p = fill[std::min(a, b)]++;
      o.crow[p] = std::max(a, b); o.cslot[p] = static_cast<int>(e);
   }
   o.nnzL = 0; o.flops = 0;
   o.first_desc.resize(ns); o.sub_flops.assign(ns, 0.0); o.sub_maxnf.assign(ns, 0);
   for (int s = 0; s < ns; ++s) {
      const Front& F = o.fronts[s];
      const double k = F.ncol, r = static_cast<double>(F.bnd.size());
      const double fl = k * k * k / 3 + k * k * r + k * r * r;
      o.nnzL += k * (k + r) - k * (k - 1) / 2; o.flops += fl;
      // supernodes follow the postordered columns, so children come before parents and a subtree is a range of indices
      o.first_desc[s] = s; o.sub_flops[s] += fl + 1.0; o.sub_maxnf[s] = std::max(o.sub_maxnf[s], F.ncol + static_cast<int>(F.bnd.size()));
      for (const int c : F.children) { o.first_desc[s] = std::min(o.first_desc[s], o.first_desc[c]); }
      if (F.parent != -1) {
         if (F.parent <= s) throw std::runtime_error("oracle: fronts are not in postorder");
         o.sub_flops[F.parent] += o.sub_flops[s];
         o.sub_maxnf[F.parent] = std::max(o.sub_maxnf[F.parent], o.sub_maxnf[s]);
      }
   }
   o.t_analyse = since(t0);
}

}  // namespace

namespace {

// ---- numeric phase: multifrontal LDL^T, threshold pivoting, delayed pivots -----------------------
inline double& lower(double* A, int nf, int i, int j) { return (i >= j) ? A[i + static_cast<size_t>(nf) * j] : A[j + static_cast<size_t>(nf) * i]; }

// symmetric interchange of indices a < b in a lower-stored nf x nf matrix
void swap_sym(double* A, int nf, int a, int b, std::vector<int>& idx) {
   if (a == b) return;
   if (a > b) std::swap(a, b);
   const size_t ld = static_cast<size_t>(nf);
   for (int c = 0; c < a; ++c) std::swap(A[a + ld * c], A[b + ld * c]);
   for (int i = a + 1; i < b; ++i) std::swap(A[i + ld * a], A[b + ld * i]);
   std::swap(A[a + ld * a], A[b + ld * b]);
   for (int i = b + 1; i < nf; ++i) std::swap(A[i + ld * a], A[i + ld * b]);
   std::swap(idx[a], idx[b]);
}

// what the host can still give: MemAvailable, and the room left under a cgroup limit if there is one
double host_free_gb() {
   double avail = 1e30;
   if (FILE* f = std::fopen("/proc/meminfo", "r")) {
      char line[256];
      while (std::fgets(line, sizeof line, f) != nullptr) {
         long long kb = 0;
         if (std::sscanf(line, "MemAvailable: %lld kB", &kb) == 1) { avail = 1e-6 * static_cast<double>(kb); break; }
      }
      std::fclose(f);
   }
   long long limit = -1, current = -1;
   if (FILE* f = std::fopen("/sys/fs/cgroup/memory.max", "r")) { if (std::fscanf(f, "%lld", &limit) != 1) limit = -1; std::fclose(f); }
   if (FILE* f = std::fopen("/sys/fs/cgroup/memory.current", "r")) { if (std::fscanf(f, "%lld", &current) != 1) current = -1; std::fclose(f); }
   if (limit > 0 && current >= 0 && limit < (1LL << 60)) avail = std::min(avail, 1e-9 * static_cast<double>(limit - current));
   return avail;
}

void factor_front(Oracle& o, int s, const double* val, Work& w) {
   Front& F = o.fronts[s];
   const bool is_root = F.parent == -1;
   F.idx.clear();
   for (int c : F.children) { const Front& C = o.fronts[c]; F.idx.insert(F.idx.end(), C.cb_idx.begin(), C.cb_idx.begin() + C.ndelay); }
   for (int j = F.first; j < F.first + F.ncol; ++j) F.idx.push_back(j);
   const int nfs = static_cast<int>(F.idx.size());
   F.idx.insert(F.idx.end(), F.bnd.begin(), F.bnd.end());
   const int nf = static_cast<int>(F.idx.size());
   F.nf = nf;
   const size_t ld = static_cast<size_t>(nf);
   const bool big = nf >= 1024;
#ifdef _OPENMP
   const bool solo = !omp_in_parallel();
#else
   const bool solo = true;
#endif
   // a front of this size asks for gigabytes: stop with an error while the host can still take it (a box driven out of memory
   // loses everything; the caller falls back to a smaller configuration)
   if (solo && nf >= 4096) {
      const double need_gb = 8e-9 * (static_cast<double>(nf) * nf * (ld * nf > w.capA ? 1.7 : 0.6));
      const double free_gb = host_free_gb();
      if (free_gb - need_gb < o.min_free_gb)
         throw std::runtime_error("oracle: host memory nearly exhausted (" + std::to_string(free_gb) + " GB available, front of " + std::to_string(nf) +
                                  " rows needs about " + std::to_string(need_gb) + " GB more): factorization stopped");
   }
   // only the lower triangle is ever read: it alone is cleared (the BLAS strips may leave anything above the diagonal)
   double* A = Work::ensure(w.A, w.capA, ld * nf);
   std::unique_ptr<Tick> tick(new Tick(T_ASM, solo));
#pragma omp parallel for schedule(static, 16) if (big)
   for (int j = 0; j < nf; ++j) std::memset(A + ld * j + j, 0, sizeof(double) * static_cast<size_t>(nf - j));
   for (int t = 0; t < nf; ++t) w.loc[F.idx[t]] = t;
   // original entries of the own columns (equilibrated)
   for (int j = F.first; j < F.first + F.ncol; ++j) {
      const int lj = w.loc[j];
      for (int64_t p = o.cptr[j]; p < o.cptr[j + 1]; ++p) {
         const int e = o.cslot[p];
         lower(A, nf, w.loc[o.crow[p]], lj) += val[e] * o.scale[o.urow[e]] * o.scale[o.ucol[e]];
      }
   }
   tick.reset(); tick.reset(new Tick(T_EXTADD, solo));
   // extend-add of the children's contribution blocks (delayed pivots included)
   for (int c : F.children) {
      Front& C = o.fronts[c];
      const int ncb = C.ncb;
      // positions of the child's rows in this front; when they ascend (the rule, unless delayed pivots came along) every
      // entry (ci >= cj) stays below the diagonal and a child column is one indexed add into a parent column
      std::vector<int> lc(static_cast<size_t>(ncb));
      bool ascending = true;
      for (int ci = 0; ci < ncb; ++ci) { lc[ci] = w.loc[C.cb_idx[ci]]; if (ci > 0 && lc[ci] <= lc[ci - 1]) ascending = false; }
      const int* lcp = lc.data();
      const double* cbase = C.cb.get();
      // every (ci, cj) lands on its own entry of the parent: the child's columns can be added side by side
#pragma omp parallel for schedule(dynamic, 16) if (big && ncb >= 256)
      for (int cj = 0; cj < ncb; ++cj) {
         const int gj = lcp[cj];
         // column cj of the packed triangle starts at cj ncb - cj (cj - 1) / 2 and holds rows cj .. ncb - 1
         const double* col = cbase + (static_cast<size_t>(cj) * ncb - static_cast<size_t>(cj) * (cj - 1) / 2) - cj;
         if (ascending) {
            double* dst = A + ld * gj;
            for (int ci = cj; ci < ncb; ++ci) dst[lcp[ci]] += col[ci];
         } else {
            for (int ci = cj; ci < ncb; ++ci) lower(A, nf, lcp[ci], gj) += col[ci];
         }
      }
      w.pool.give(C.cb, C.cb_cap);
   }
   tick.reset();
   F.d.assign(nfs, 1.0); F.e.assign(nfs, 0.0); F.ptype.assign(nfs, 0);
   // big fronts: wider blocks (fewer passes over the trailing matrix) and the in-block column work spread over the threads;
   // small fronts are factored one per thread by the level loop (nested regions stay serial)
   const int NB = nf >= 2048 ? 256 : (nf >= 256 ? 128 : 64);   // outer block: K of the BLAS-3 trailing update
   const int NBI = 32;                                         // inner block
   const bool par = nf >= 1024;
   double* W = Work::ensure(w.W, w.capW, ld * NB);
   // Restricted pivoting never looks at the rows that are not fully summed (R = [rlim, nf)): they stay out of the column-by-column
   // work and are brought in once per outer block with BLAS-3 calls (W_R = A_R L_PP^{-T}, L_R = W_R D^{-1}).
   const int rlim = (o.pivot_mode == 1) ? nfs : nf;
   int ne = 0, active_end = nfs, since_retry = 0;
   double u = o.u;
   bool forcing = false;
   const double tol = o.null_tol;
   while (true) {
      // Two-level blocking.  Outer block [ostart, obend): its pivots update everything right of it in ONE BLAS-3 call per strip.
      // Inner block [ne, ibend): factored right-looking, column by column, on NBI columns that stay in cache; when it is done its
      // pivots update the rest of the outer block (a small dgemm), so every candidate column is up to date when it is tested.
      // Interchanges only ever involve positions below obend, hence the rows of W used by the outer update never move.
      while (ne < active_end) {
         const int ostart = ne, obend = std::min(ne + NB, active_end);
         int oend = obend;            // candidates that failed with nothing eliminated around them are parked in [oend, obend)
         while (ne < oend) {
            const int istart = ne, ibend = std::min(ne + NBI, oend);
            int cand = ne;
            tick.reset(new Tick(T_PIVOT, solo));
            while (cand < ibend) {
               // column maxima of candidate c over the not yet eliminated part of the front
               const int c = cand;
               double lam_all = 0.0, lam_fs = 0.0;
               int r = -1;
               for (int i = ne; i < ibend; ++i) {
                  if (i == c) continue;
                  const double v = std::fabs(lower(A, nf, i, c));
                  if (v > lam_fs) { lam_fs = v; r = i; }
               }
               lam_all = lam_fs;
               const int scan_end = o.pivot_mode == 1 ? active_end : nf;
               {
                  const double* colc = A + ld * c;
                  double m = 0.0;
   #pragma omp parallel for reduction(max : m) if (par)
                  for (int i = ibend; i < scan_end; ++i) m = std::max(m, std::fabs(colc[i]));
                  lam_all = std::max(lam_all, m);
               }
               const double acc = A[c + ld * c];
               int step = 0;
               if (std::fabs(acc) > tol && std::fabs(acc) >= u * lam_all) {
                  step = 1;
               } else if (r >= 0 && lam_fs > 0.0) {
                  const double b = lower(A, nf, r, c), dd = A[r + ld * r];
                  const double det = acc * dd - b * b;
                  double lc = 0.0, lr = 0.0;
                  for (int i = ne; i < scan_end; ++i) {
                     if (i == c || i == r) continue;
                     lc = std::max(lc, std::fabs(lower(A, nf, i, c)));
                     lr = std::max(lr, std::fabs(lower(A, nf, i, r)));
                  }
                  const double g1 = std::fabs(dd) * lc + std::fabs(b) * lr, g2 = std::fabs(b) * lc + std::fabs(acc) * lr;
                  if (std::fabs(det) > tol * tol && std::fabs(b) > tol && u * std::max(g1, g2) <= std::fabs(det)) step = 2;
               }
               if (step == 0) { ++cand; continue; }
               if (step == 1) {
                  swap_sym(A, nf, ne, c, F.idx);
                  const double piv = A[ne + ld * ne];
                  double* wc = W + ld * (ne - ostart);
                  for (int i = ne + 1; i < rlim; ++i) { wc[i] = A[i + ld * ne]; A[i + ld * ne] = wc[i] / piv; }
   #pragma omp parallel for schedule(static) if (par)
                  for (int j = ne + 1; j < ibend; ++j) {
                     const double wj = wc[j];
                     if (wj == 0.0) continue;
                     double* cj = A + ld * j;
                     const double* lcol = A + ld * ne;
                     for (int i = j; i < rlim; ++i) cj[i] -= lcol[i] * wj;
                  }
                  F.d[ne] = piv; F.e[ne] = 0.0; F.ptype[ne] = 0;
                  ne += 1;
               } else {
                  int rr = r;
                  swap_sym(A, nf, ne, c, F.idx);
                  if (rr == ne) rr = c;            // the partner sat at position ne and moved to c
                  swap_sym(A, nf, ne + 1, rr, F.idx);
                  const double a = A[ne + ld * ne], b = A[ne + 1 + ld * ne], dd = A[ne + 1 + ld * (ne + 1)];
                  const double det = a * dd - b * b;
                  double* w1 = W + ld * (ne - ostart);
                  double* w2 = w1 + ld;
                  for (int i = ne + 2; i < rlim; ++i) {
                     w1[i] = A[i + ld * ne]; w2[i] = A[i + ld * (ne + 1)];
                     A[i + ld * ne] = (dd * w1[i] - b * w2[i]) / det;
                     A[i + ld * (ne + 1)] = (a * w2[i] - b * w1[i]) / det;
                  }
                  w1[ne + 1] = 0.0; w2[ne + 1] = 0.0;
                  A[ne + 1 + ld * ne] = 0.0;
   #pragma omp parallel for schedule(static) if (par)
                  for (int j = ne + 2; j < ibend; ++j) {
                     const double x1 = w1[j], x2 = w2[j];
                     double* cj = A + ld * j;
                     const double* l1 = A + ld * ne;
                     const double* l2 = A + ld * (ne + 1);
                     for (int i = j; i < rlim; ++i) cj[i] -= l1[i] * x1 + l2[i] * x2;
                  }
                  F.d[ne] = a; F.d[ne + 1] = dd; F.e[ne] = b; F.e[ne + 1] = 0.0; F.ptype[ne] = 1; F.ptype[ne + 1] = 2;
                  ne += 2;
               }
               cand = std::max(cand + 1, ne);   // go on with untested candidates; failed ones are retried in the next block
            }
            const int inbw = ne - istart;
            tick.reset(); tick.reset(new Tick(T_INNER_GEMM, solo));
            if (inbw > 0) {
               // the inner block's pivots on the rest of the outer block; failed candidates [ne, ibend) stay in front (up to date)
               gemm_nt_sub_team(rlim - ibend, obend - ibend, inbw, W + ibend + ld * (istart - ostart), nf, A + ibend + ld * istart, nf, A + ibend + ld * ibend, nf,
                                false);
            } else {
               const int cnt = ibend - ne;
               for (int t = 0; t < cnt; ++t) { const int a = ne + t, b = oend - 1 - t; if (a < b) swap_sym(A, nf, a, b, F.idx); }
               oend -= cnt;
            }
         }
         const int nbw = ne - ostart;
         tick.reset(); tick.reset(new Tick(T_ROWS, solo));
         if (nbw > 0 && rlim < nf) {
            // rows R of the block's pivot columns: W_R = A_R L_PP^{-T} (A_R still as the outer block found it), L_R = W_R D^{-1}
            const int mr = nf - rlim;
            trsm_right_lower_trans_unit_team(mr, nbw, A + ostart + ld * ostart, nf, A + rlim + ld * ostart, nf);
#pragma omp parallel for schedule(static) if (big)
            for (int p = ostart; p < ne; ++p) {
               if (F.ptype[p] == 2) continue;
               double* wp = W + ld * (p - ostart);
               double* ap = A + ld * p;
               if (F.ptype[p] == 0) {
                  const double piv = F.d[p];
                  for (int i = rlim; i < nf; ++i) { wp[i] = ap[i]; ap[i] = wp[i] / piv; }
               } else {
                  double* wq = wp + ld;
                  double* aq = ap + ld;
                  const double a = F.d[p], b = F.e[p], dd = F.d[p + 1], det = a * dd - b * b;
                  for (int i = rlim; i < nf; ++i) {
                     wp[i] = ap[i]; wq[i] = aq[i];
                     ap[i] = (dd * wp[i] - b * wq[i]) / det;
                     aq[i] = (a * wq[i] - b * wp[i]) / det;
                  }
               }
            }
            // and rows R of the columns of this block that were not eliminated
            gemm_nt_sub_team(mr, obend - ne, nbw, W + rlim, nf, A + ne + ld * ostart, nf, A + rlim + ld * ne, nf, false);
         }
         tick.reset(); tick.reset(new Tick(T_OUTER_GEMM, solo));
         if (nbw > 0) {
            since_retry += nbw;
            // rank-nbw update of everything right of the outer block: C -= W * L^T, tile by tile (lower part only)
            gemm_nt_sub_team(nf - obend, nf - obend, nbw, W + obend, nf, A + obend + ld * ostart, nf, A + obend + ld * obend, nf, true);
            // parked candidates [ne, obend) stay in front of the next block (already up to date)
         } else {
            // nothing could be eliminated in this block: set its columns aside and go on
            const int cnt = obend - ne;
            for (int t = 0; t < cnt; ++t) { const int a = ne + t, b = active_end - 1 - t; if (a < b) swap_sym(A, nf, a, b, F.idx); }
            active_end -= cnt;
         }
      }
      tick.reset();
      if (active_end < nfs && since_retry > 0) { active_end = nfs; since_retry = 0; continue; }   // retry postponed columns
      if (is_root && ne < nfs && !forcing) { forcing = true; u = 0.0; active_end = nfs; since_retry = 0; continue; }
      break;
   }
   if (is_root) {
      // what is left at the root has no acceptable pivot even without threshold: null pivots
      for (int j = ne; j < nfs; ++j) {
         for (int i = j + 1; i < nf; ++i) A[i + ld * j] = 0.0;
         F.d[j] = 1.0; F.e[j] = 0.0; F.ptype[j] = 3;
      }
      ne = nfs;
   }
   F.nelim = ne;
   tick.reset(new Tick(T_COPY, solo));
   if (solo && g_prof_on) { g_prof_nf2 += double(nf) * nf; g_prof_nfne += double(nf) * ne; g_prof_big += 1; for (int c : F.children) g_prof_cb2 += double(o.fronts[c].ncb) * o.fronts[c].ncb; }
   // (only the part of a column at and below its diagonal is ever read: nothing is cleared)
   const size_t lsize = lpos(nf, ne) + static_cast<size_t>(ne);   // ne nf - ne (ne - 1) / 2
   if (F.L_size != lsize) { F.L.reset(); F.L.reset(new double[lsize > 0 ? lsize : 1]); F.L_size = lsize; }
   double* Lp = F.L.get();
#pragma omp parallel for schedule(static, 8) if (big)
   for (int j = 0; j < ne; ++j) std::memcpy(Lp + lpos(nf, j) + j, A + ld * j + j, sizeof(double) * (nf - j));
   F.d.resize(ne); F.e.resize(ne); F.ptype.resize(ne);
   F.ncb = nf - ne; F.ndelay = nfs - ne;
   F.cb_idx.assign(F.idx.begin() + ne, F.idx.end());
   w.pool.give(F.cb, F.cb_cap);   // a root's block of the previous factorization
   F.cb = w.pool.take(static_cast<size_t>(F.ncb) * (F.ncb + 1) / 2, F.cb_cap);
   double* cbp = F.cb.get();
   const size_t ncb_ = static_cast<size_t>(F.ncb);
#pragma omp parallel for schedule(static, 8) if (big)
   for (int j = ne; j < nf; ++j) {
      const size_t cj = static_cast<size_t>(j - ne);
      std::memcpy(cbp + (cj * ncb_ - cj * (cj - 1) / 2), A + ld * j + j, sizeof(double) * (nf - j));
   }
}

void equilibrate(Oracle& o, const double* val, int sweeps) {
   const int64_t N = o.N;
   o.scale.assign(N, 1.0);
   std::vector<double> m(N);
   for (int it = 0; it < sweeps; ++it) {
      std::fill(m.begin(), m.end(), 0.0);
      for (int64_t e = 0; e < o.nnz; ++e) {
         const int r = o.urow[e], c = o.ucol[e];
         const double v = std::fabs(val[e]) * o.scale[r] * o.scale[c];
         if (v > m[r]) m[r] = v;
         if (v > m[c]) m[c] = v;
      }
      for (int64_t i = 0; i < N; ++i) if (m[i] > 0.0 && std::isfinite(m[i])) o.scale[i] /= std::sqrt(m[i]);
   }
}

void factorize(Oracle& o, const double* val, int nthreads) {
   const auto t0 = clk::now();
   g_prof_on = std::getenv("ORACLE_PROFILE") != nullptr;
   std::fill(g_prof, g_prof + T_COUNT, 0.0); g_prof_nf2 = g_prof_nfne = g_prof_cb2 = 0; g_prof_big = 0;
   for (int c = 0; c < 5; ++c) { g_prof_cls[c] = 0; g_prof_cnt[c] = 0; }
   equilibrate(o, val, 3);
   if (nthreads < 1) nthreads = 1;
#ifdef _OPENMP
   omp_set_num_threads(nthreads);
#endif
   if (g_set_threads != nullptr) g_set_threads(1);   // every BLAS call runs in the calling OpenMP thread (see gemm_nt_sub_team)
   std::vector<Work>& works = o.works;
   if (static_cast<int>(works.size()) != nthreads) { works.clear(); works.resize(nthreads); }
   for (auto& w : works) if (static_cast<int64_t>(w.loc.size()) != o.N) w.loc.assign(o.N, -1);
   if (const char* e = std::getenv("ORACLE_MIN_FREE_GB")) o.min_free_gb = std::atof(e);
   // Two phases.  Below a cut, whole subtrees go one per thread, each swept in postorder with single-threaded BLAS: no
   // synchronisation, the contribution blocks stay in that thread's cache and pool.  Above it, the fronts follow one another in
   // postorder with every thread inside the front (tiled BLAS, threaded assembly and copies): at most one stack of contribution
   // blocks is alive, which is what lets the 128^3 system fit the host.  A subtree qualifies if it is a small share of the work
   // and its largest front fits a per-thread workspace.
   const int ns = static_cast<int>(o.fronts.size());
   const double share = o.flops / (8.0 * nthreads);
   auto below = [&](int f) { return nthreads > 1 && o.sub_flops[f] <= share && o.sub_maxnf[f] <= 8192; };
   std::vector<int> roots;
   for (int f = 0; f < ns; ++f) if (below(f) && (o.fronts[f].parent == -1 || !below(o.fronts[f].parent))) roots.push_back(f);
   std::sort(roots.begin(), roots.end(), [&](int a, int b) { return o.sub_flops[a] != o.sub_flops[b] ? o.sub_flops[a] > o.sub_flops[b] : a < b; });
   const int nroots = static_cast<int>(roots.size());
   if (nroots > 0) {
      Tick side_tick(T_SIDE, true);
      bool starved = false;
#pragma omp parallel for schedule(dynamic, 1)
      for (int t = 0; t < nroots; ++t) {
#ifdef _OPENMP
         Work& w = works[omp_get_thread_num()];
#else
         Work& w = works[0];
#endif
         if (starved) continue;
         if (o.sub_maxnf[roots[t]] >= 2048 && host_free_gb() < o.min_free_gb) { starved = true; continue; }
         for (int f = o.first_desc[roots[t]]; f <= roots[t]; ++f) {
            if (g_prof_on) {
               const auto tf = clk::now();
               factor_front(o, f, val, w);
               const double dt = since(tf);
               const int nf = o.fronts[f].nf;
               const int cls = nf < 32 ? 0 : nf < 128 ? 1 : nf < 512 ? 2 : nf < 1500 ? 3 : 4;
#pragma omp critical(oracle_prof)
               { g_prof_cls[cls] += dt; g_prof_cnt[cls] += 1; }
            } else {
               factor_front(o, f, val, w);
            }
         }
      }
      if (starved) throw std::runtime_error("oracle: host memory nearly exhausted during the subtree phase: factorization stopped");
      // the per-thread workspaces of the subtree phase are not needed above the cut
      if (o.flops > 2.0e13) for (size_t t = 1; t < works.size(); ++t) { works[t].A.reset(); works[t].capA = 0; works[t].W.reset(); works[t].capW = 0; }
   }
   for (int f = 0; f < ns; ++f) if (!below(f)) factor_front(o, f, val, works[0]);
   o.pos = o.neg = o.zero = 0; o.n2x2 = 0; o.ndelayed = 0;
   for (const Front& F : o.fronts) {
      o.ndelayed += F.ndelay;
      for (int j = 0; j < F.nelim; ++j) {
         if (F.ptype[j] == 0) { if (F.d[j] > 0) ++o.pos; else ++o.neg; }
         else if (F.ptype[j] == 1) {
            const double det = F.d[j] * F.d[j + 1] - F.e[j] * F.e[j];
            if (det < 0) { ++o.pos; ++o.neg; } else if (F.d[j] > 0) o.pos += 2; else o.neg += 2;
            ++o.n2x2;
         } else if (F.ptype[j] == 3) ++o.zero;
      }
   }
   o.t_factor = since(t0);
   if (g_prof_on)
      std::fprintf(stderr, "oracle profile: total %.3f s | levels of small fronts %.3f | big fronts: assemble %.3f extend-add %.3f pivot columns %.3f "
                   "inner gemm %.3f boundary rows %.3f outer gemm %.3f copy-out %.3f\n", o.t_factor, g_prof[T_SIDE], g_prof[T_ASM], g_prof[T_EXTADD],
                   g_prof[T_PIVOT], g_prof[T_INNER_GEMM], g_prof[T_ROWS], g_prof[T_OUTER_GEMM], g_prof[T_COPY]);
   if (g_prof_on) std::fprintf(stderr, "oracle profile: small fronts by size (<32, <128, <512, <1500, rest): thread-seconds %.3f %.3f %.3f %.3f %.3f, counts %ld %ld %ld %ld %ld\n",
      g_prof_cls[0], g_prof_cls[1], g_prof_cls[2], g_prof_cls[3], g_prof_cls[4], g_prof_cnt[0], g_prof_cnt[1], g_prof_cnt[2], g_prof_cnt[3], g_prof_cnt[4]);
   if (g_prof_on) std::fprintf(stderr, "oracle profile: %d big fronts, sum nf^2 %.3g, sum nf*ne %.3g, sum child ncb^2 %.3g doubles\n", g_prof_big, g_prof_nf2, g_prof_nfne, g_prof_cb2);
}

// One right-hand side: every factor entry is read once per sweep (memory-bound).  Small fronts run as they are; a big front is
// swept in blocks of SB pivot columns - the triangle of the block by one thread, the rows below it shared out over the team.
void solve(Oracle& o, const double* rhs, double* x) {
   const auto t0 = clk::now();
   const int64_t N = o.N;
   constexpr int SB = 64;
   std::vector<double> y(N), xs, part;
   for (int64_t g = 0; g < N; ++g) y[g] = rhs[o.perm[g]] * o.scale[o.perm[g]];
   const int ns = static_cast<int>(o.fronts.size());
#ifdef _OPENMP
   const int nth = omp_get_max_threads();
#else
   const int nth = 1;
#endif
   for (int s = 0; s < ns; ++s) {
      const Front& F = o.fronts[s];
      const int nf = F.nf, ne = F.nelim;
      xs.resize(nf);
      for (int t = 0; t < nf; ++t) xs[t] = y[F.idx[t]];
      const double* L = F.L.get();
      double* xp = xs.data();
      if (static_cast<double>(nf) * ne < 2.0e5 || nth == 1) {
         for (int j = 0; j < ne; ++j) {
            const double yj = xp[j];
            if (yj == 0.0) continue;
            const double* l = L + lpos(nf, j);
            for (int i = j + 1; i < nf; ++i) xp[i] -= l[i] * yj;
         }
      } else {
         for (int jb = 0; jb < ne; jb += SB) {
            const int je = std::min(jb + SB, ne);
            for (int j = jb; j < je; ++j) {
               const double yj = xp[j];
               const double* l = L + lpos(nf, j);
               for (int i = j + 1; i < je; ++i) xp[i] -= l[i] * yj;
            }
#pragma omp parallel for schedule(static)
            for (int c = 0; c < nth; ++c) {
               const int64_t rows = nf - je;
               const int i0 = je + static_cast<int>(rows * c / nth), i1 = je + static_cast<int>(rows * (c + 1) / nth);
               for (int j = jb; j < je; ++j) {
                  const double yj = xp[j];
                  const double* l = L + lpos(nf, j);
                  for (int i = i0; i < i1; ++i) xp[i] -= l[i] * yj;
               }
            }
         }
      }
      for (int t = 0; t < nf; ++t) y[F.idx[t]] = xs[t];
   }
   for (int s = 0; s < ns; ++s) {
      const Front& F = o.fronts[s];
      for (int j = 0; j < F.nelim; ++j) {
         double& a = y[F.idx[j]];
         if (F.ptype[j] == 0) a /= F.d[j];
         else if (F.ptype[j] == 1) {
            double& b = y[F.idx[j + 1]];
            const double det = F.d[j] * F.d[j + 1] - F.e[j] * F.e[j];
            const double y1 = a, y2 = b;
            a = (F.d[j + 1] * y1 - F.e[j] * y2) / det;
            b = (F.d[j] * y2 - F.e[j] * y1) / det;
         } else if (F.ptype[j] == 3) a = 0.0;
      }
   }
   part.resize(static_cast<size_t>(nth) * SB);
   for (int s = ns - 1; s >= 0; --s) {
      const Front& F = o.fronts[s];
      const int nf = F.nf, ne = F.nelim;
      xs.resize(nf);
      for (int t = 0; t < nf; ++t) xs[t] = y[F.idx[t]];
      const double* L = F.L.get();
      double* xp = xs.data();
      if (static_cast<double>(nf) * ne < 2.0e5 || nth == 1) {
         for (int j = ne - 1; j >= 0; --j) {
            const double* l = L + lpos(nf, j);
            double acc = xp[j];
            for (int i = j + 1; i < nf; ++i) acc -= l[i] * xp[i];
            xp[j] = acc;
         }
      } else {
         const int nblk = (ne + SB - 1) / SB;
         for (int b = nblk - 1; b >= 0; --b) {
            const int jb = b * SB, je = std::min(jb + SB, ne);
            double* pp = part.data();
            // the rows below the block: every thread's share of the je - jb dot products
#pragma omp parallel for schedule(static)
            for (int c = 0; c < nth; ++c) {
               const int64_t rows = nf - je;
               const int i0 = je + static_cast<int>(rows * c / nth), i1 = je + static_cast<int>(rows * (c + 1) / nth);
               for (int j = jb; j < je; ++j) {
                  const double* l = L + lpos(nf, j);
                  double acc = 0.0;
                  for (int i = i0; i < i1; ++i) acc += l[i] * xp[i];
                  pp[static_cast<size_t>(c) * SB + (j - jb)] = acc;
               }
            }
            for (int j = je - 1; j >= jb; --j) {
               const double* l = L + lpos(nf, j);
               double acc = xp[j];
               for (int c = 0; c < nth; ++c) acc -= pp[static_cast<size_t>(c) * SB + (j - jb)];
               for (int i = j + 1; i < je; ++i) acc -= l[i] * xp[i];
               xp[j] = acc;
            }
         }
      }
      for (int j = 0; j < ne; ++j) y[F.idx[j]] = xs[j];
   }
   for (int64_t g = 0; g < N; ++g) x[o.perm[g]] = y[g] * o.scale[o.perm[g]];
   o.t_solve = since(t0);
}

}  // namespace

extern "C" {

void* oracle_create() { return new Oracle(); }
void oracle_destroy(void* p) { delete static_cast<Oracle*>(p); }
const char* oracle_last_error(void* p) { return static_cast<Oracle*>(p)->err.c_str(); }

// path of a shared library exporting cblas_dgemm (optionally with a prefix, e.g. scipy_); returns 1 if bound
int oracle_set_blas(const char* path, const char* prefix) {
   void* lib = dlopen(path, RTLD_NOW | RTLD_LOCAL);
   if (lib == nullptr) return 0;
   const std::string pre = prefix == nullptr ? "" : prefix;
   g_dgemm = reinterpret_cast<cblas_dgemm_t>(dlsym(lib, (pre + "cblas_dgemm").c_str()));
   g_dtrsm = reinterpret_cast<cblas_dtrsm_t>(dlsym(lib, (pre + "cblas_dtrsm").c_str()));
   g_set_threads = reinterpret_cast<set_threads_t>(dlsym(lib, (pre + "openblas_set_num_threads").c_str()));
   return g_dgemm != nullptr ? 1 : 0;
}

int oracle_analyse(void* p, int64_t n, int64_t nnz, const uint64_t* row, const uint64_t* col, const int64_t* perm) {
   Oracle& o = *static_cast<Oracle*>(p);
   try { analyse(o, n, nnz, row, col, perm); return 0; } catch (const std::exception& e) { o.err = e.what(); return -1; }
}

int oracle_set_pivot_mode(void* p, int mode) {
   static_cast<Oracle*>(p)->pivot_mode = mode;
   return 0;
}

int oracle_set_parameters(void* p, double threshold_u, double null_tol) {
   Oracle& o = *static_cast<Oracle*>(p);
   if (threshold_u >= 0.0) o.u = threshold_u;
   if (null_tol >= 0.0) o.null_tol = null_tol;
   return 0;
}

int oracle_factorize(void* p, const double* values, int nthreads) {
   Oracle& o = *static_cast<Oracle*>(p);
   try { factorize(o, values, nthreads); return 0; } catch (const std::exception& e) { o.err = e.what(); return -1; }
}

int oracle_inertia(void* p, uint64_t* pos, uint64_t* neg, uint64_t* zero) {
   const Oracle& o = *static_cast<Oracle*>(p);
   *pos = o.pos; *neg = o.neg; *zero = o.zero;
   return 0;
}

int oracle_solve(void* p, const double* rhs, double* x) {
   Oracle& o = *static_cast<Oracle*>(p);
   try { solve(o, rhs, x); return 0; } catch (const std::exception& e) { o.err = e.what(); return -1; }
}

int oracle_get_permutation(void* p, int64_t* perm) {
   const Oracle& o = *static_cast<Oracle*>(p);
   for (int64_t i = 0; i < o.N; ++i) perm[i] = o.perm[i];
   return 0;
}

// out[0..9] = nnz(L) estimate, flops estimate, delayed pivots, 2x2 pivots, t_analyse, t_factor, t_solve, fronts, blas bound, threads
int oracle_stats(void* p, double* out) {
   const Oracle& o = *static_cast<Oracle*>(p);
   out[0] = o.nnzL; out[1] = o.flops; out[2] = static_cast<double>(o.ndelayed); out[3] = static_cast<double>(o.n2x2);
   out[4] = o.t_analyse; out[5] = o.t_factor; out[6] = o.t_solve; out[7] = static_cast<double>(o.fronts.size());
   out[8] = g_dgemm != nullptr ? 1.0 : 0.0;
#ifdef _OPENMP
   out[9] = omp_get_max_threads();
#else
   out[9] = 1.0;
#endif
   return 0;
}

}  // extern "C"
