// sm_100a numeric phase of the B200 KKT solver: assembly (K1/K1'), supernodal multifrontal
// LDL^T with in-front Bunch-Kaufman pivoting (K3), FP64 tensor-core Schur updates (K4),
// extend-add (K5), inertia (K6), level-scheduled solves (K7) and the residual for refinement (K8).
//
// Stands in for what dmumps_c does at JOB=2 / JOB=3 behind
// uno/ingredients/subproblem_solvers/MUMPS/MUMPSSolver.cpp:85-96.  Nothing here is derived from
// MUMPS; the algorithm is the textbook multifrontal method laid out for the GPU:
// all fronts of an elimination-tree level are processed by the same batched launches.
#include "kernels.hpp"

#include <cstdio>

namespace b200kkt {

namespace {

constexpr double BK_ALPHA = 0.6403882032022076;   // (1 + sqrt(17)) / 8
constexpr int PANEL_THREADS = 256;
constexpr int DIAG_THREADS = 128;
constexpr int EA_CBW = 16;                        // parent columns owned by one extend-add CTA

__device__ __forceinline__ unsigned long long ordered_bits(double v) {
   // v >= 0: IEEE bit pattern is order preserving
   return static_cast<unsigned long long>(__double_as_longlong(v));
}

// ------------------------------------------------------------------------------------------------
// K1' equilibration, norms, K1 assembly, K8 residual
// ------------------------------------------------------------------------------------------------
__global__ void k_fill(double* p, int64_t n, double v) {
   const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
   if (i < n) p[i] = v;
}

// One Ruiz step with power-of-two factors: sc_out = sc_in * 2^(-ceil(e/2)), e = floor(log2(scaled row max)).
// Powers of two keep the scaled entries exactly representable, so assembly stays bit-reproducible.
__global__ void k_scale_sweep(CsrArrays A, const double* __restrict__ values, const double* __restrict__ sc_in,
                              double* __restrict__ sc_out) {
   const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
   if (i >= A.n) return;
   double m = 0.0;
   for (int t = A.ptr[i]; t < A.ptr[i + 1]; ++t) m = fmax(m, fabs(values[A.slot[t]]) * sc_in[A.col[t]]);
   const double si = sc_in[i];
   m *= si;
   double out = si;
   if (m > 0.0 && m < 1.0e300) {
      int ex;
      frexp(m, &ex);              // m = fr * 2^ex, fr in [0.5, 1)  ->  floor(log2 m) = ex - 1
      const int e = ex - 1;
      const int sh = -((e + 1) >> 1);
      out = ldexp(si, sh);
   }
   sc_out[i] = out;
}

__global__ void k_matrix_inf_norm(CsrArrays A, const double* __restrict__ values, unsigned long long* out_bits) {
   const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
   double s = 0.0;
   if (i < A.n) for (int t = A.ptr[i]; t < A.ptr[i + 1]; ++t) s += fabs(values[A.slot[t]]);
   if (s != s) s = __longlong_as_double(0x7ff0000000000000LL);   // NaN in the values: reported as an infinite norm
   for (int o = 16; o; o >>= 1) s = fmax(s, __shfl_xor_sync(0xffffffffu, s, o));
   if ((threadIdx.x & 31) == 0 && s > 0.0) atomicMax(out_bits, ordered_bits(s));
}

__global__ void k_absmax(const double* __restrict__ v, int64_t n, unsigned long long* out_bits) {
   double s = 0.0;
   for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < n; i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
      const double a = fabs(v[i]);
      s = (a > s || a != a) ? a : s;   // NaN propagates
   }
   if (s != s) s = __longlong_as_double(0x7ff0000000000000LL);   // report NaN as +inf
   for (int o = 16; o; o >>= 1) s = fmax(s, __shfl_xor_sync(0xffffffffu, s, o));
   if ((threadIdx.x & 31) == 0 && s > 0.0) atomicMax(out_bits, ordered_bits(s));
}

// One thread per distinct destination; its sources are summed in ascending COO slot order.
__global__ void k_assemble(AssemblyArrays M, const double* __restrict__ values, const double* __restrict__ scale,
                           double* __restrict__ pool) {
   const int64_t d = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
   if (d >= M.ndst) return;
   double s = 0.0;
   for (int t = M.seg[d]; t < M.seg[d + 1]; ++t) s += values[M.src[t]] * scale[M.row[t]] * scale[M.col[t]];
   pool[M.dst[d]] = s;
}

// K1 + zero-fill in one pass over the pool.  Persistent CTAs take tiles of ASM_TILE doubles.  A tile without destinations (the
// rule: one pool entry in a few hundred is an original entry) is streamed out as zeros; a tile with destinations is built in
// shared memory - zeros, then each destination by one thread that sums its sources in ascending slot order - and streamed out.
// Stores are 16 bytes per lane, 512 contiguous bytes per warp.
__global__ void __launch_bounds__(256) k_assemble_fill(AssemblyArrays M, const double* __restrict__ values, const double* __restrict__ scale,
                                                       double* __restrict__ pool) {
   __shared__ __align__(16) double tile[ASM_TILE];
   const int tid = threadIdx.x;
   const double2 zero2 = make_double2(0.0, 0.0);
   for (int64_t t = blockIdx.x; t < M.ntiles; t += gridDim.x) {
      const int64_t base = t * ASM_TILE;
      const int64_t left = M.pool_doubles - base;
      const int count = left < ASM_TILE ? static_cast<int>(left) : ASM_TILE;
      const int d0 = M.tile_ptr[t], d1 = M.tile_ptr[t + 1];
      double2* out = reinterpret_cast<double2*>(pool + base);   // tiles start on 16 KB boundaries of an allocation aligned to 256 bytes
      if (d0 == d1) {
         if (count == ASM_TILE) {
#pragma unroll
            for (int i = 0; i < ASM_TILE / 2 / 256; ++i) out[tid + 256 * i] = zero2;
         } else {
            for (int i = tid; i < count; i += 256) pool[base + i] = 0.0;
         }
         continue;
      }
#pragma unroll
      for (int i = 0; i < ASM_TILE / 2 / 256; ++i) reinterpret_cast<double2*>(tile)[tid + 256 * i] = zero2;
      __syncthreads();
      for (int d = d0 + tid; d < d1; d += 256) {
         double s = 0.0;
         for (int q = M.seg[d]; q < M.seg[d + 1]; ++q) s += values[M.src[q]] * scale[M.row[q]] * scale[M.col[q]];
         tile[M.dst[d] - base] = s;
      }
      __syncthreads();
      if (count == ASM_TILE) {
#pragma unroll
         for (int i = 0; i < ASM_TILE / 2 / 256; ++i) out[tid + 256 * i] = reinterpret_cast<const double2*>(tile)[tid + 256 * i];
      } else {
         for (int i = tid; i < count; i += 256) pool[base + i] = tile[i];
      }
      __syncthreads();   // the tile buffer is reused
   }
}

// r = b - K x, and the componentwise backward error of Arioli, Demmel and Duff (SIAM J. Matrix Anal. 1989) - what MUMPS reports
// as RINFOG(7)/(8): rows whose denominator (|K||x| + |b|)_i is safely above rounding level give omega_1 = |r_i| / (|K||x| + |b|)_i,
// the others (denominator below 1000 n eps (|K_i|_inf |x|_inf + |b_i|)) give omega_2 = |r_i| / ((|K||x|)_i + |K_i|_1 |x|_inf).
// The maximum of both over all rows goes to *omega_bits as an ordered bit pattern; |x|_inf is read from *xnorm_bits.
__global__ void k_residual(CsrArrays A, const double* __restrict__ values, const double* __restrict__ x,
                           const double* __restrict__ b, double* __restrict__ r, const unsigned long long* __restrict__ xnorm_bits,
                           unsigned long long* __restrict__ omega_bits) {
   const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
   double w = 0.0;
   if (i < A.n) {
      const double xnorm = __longlong_as_double(static_cast<long long>(*xnorm_bits));
      double s = 0.0, a = 0.0, rowmax = 0.0, rowsum = 0.0;
      for (int t = A.ptr[i]; t < A.ptr[i + 1]; ++t) {
         const double v = values[A.slot[t]], xv = x[A.col[t]], av = fabs(v);
         s += v * xv;
         a = fma(av, fabs(xv), a);
         rowmax = fmax(rowmax, av);
         rowsum += av;
      }
      const double bi = b[i], ri = bi - s;
      r[i] = ri;
      const double d1 = a + fabs(bi);
      const double tau = 1000.0 * static_cast<double>(A.n) * 2.220446049250313e-16 * (rowmax * xnorm + fabs(bi));
      const double den = (d1 > tau) ? d1 : (a + rowsum * xnorm);
      if (den > 0.0) w = fabs(ri) / den;
      if (w != w || ri != ri) w = __longlong_as_double(0x7ff0000000000000LL);
   }
   for (int o = 16; o; o >>= 1) w = fmax(w, __shfl_xor_sync(0xffffffffu, w, o));
   if ((threadIdx.x & 31) == 0 && w > 0.0) atomicMax(omega_bits, ordered_bits(w));
}

__global__ void k_solve_gather(int64_t n, const int* __restrict__ perm, const double* __restrict__ scale,
                               const double* __restrict__ r, double* __restrict__ zg) {
   const int64_t g = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
   if (g >= n) return;
   const int o = perm[g];
   zg[g] = scale[o] * r[o];
}

__global__ void k_solve_scatter_add(int64_t n, const int* __restrict__ perm, const double* __restrict__ scale,
                                    const double* __restrict__ xg, double* __restrict__ x) {
   const int64_t g = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
   if (g >= n) return;
   const int o = perm[g];
   x[o] += scale[o] * xg[g];
}

// ------------------------------------------------------------------------------------------------
// K9 helpers of the subtree-partitioned solve (one rank per GPU).  At a cut edge (front s whose parent lives on another
// rank) the backward sweep ships the solution on s's boundary rows: the parent's rank packs it into s's slot of the boundary
// pool, NCCL moves the slot, s's rank unpacks it into its own solution vector.  One CTA per cut front.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_boundary_pack(FrontArrays F, const int* __restrict__ list, const double* __restrict__ xg,
                                                       double* __restrict__ cpool) {
   const int s = list[blockIdx.x];
   const int k = F.k[s], r = F.f[s] - k;
   const int* __restrict__ rows = F.rows + F.rows_off[s] + k;
   double* slot = cpool + F.rel_off[s];
   for (int i = threadIdx.x; i < r; i += 256) slot[i] = xg[rows[i]];
}

__global__ void __launch_bounds__(256) k_boundary_unpack(FrontArrays F, const int* __restrict__ list, const double* __restrict__ cpool,
                                                         double* __restrict__ xg) {
   const int s = list[blockIdx.x];
   const int k = F.k[s], r = F.f[s] - k;
   const int* __restrict__ rows = F.rows + F.rows_off[s] + k;
   const double* slot = cpool + F.rel_off[s];
   for (int i = threadIdx.x; i < r; i += 256) xg[rows[i]] = slot[i];
}

// rows [row0, row0 + nrows) of ncols columns of a column-major panel (leading dimension ld) to / from a dense nrows x ncols
// staging block: a factored panel of a shared front is broadcast without the rows above its diagonal block
__global__ void __launch_bounds__(256) k_rows_pack(const double* __restrict__ src, const int64_t ld, const int row0, const int nrows,
                                                   double* __restrict__ dst) {
   const int i = blockIdx.x * 256 + threadIdx.x, c = blockIdx.y;
   if (i < nrows) dst[static_cast<int64_t>(c) * nrows + i] = src[static_cast<int64_t>(c) * ld + row0 + i];
}

__global__ void __launch_bounds__(256) k_rows_unpack(const double* __restrict__ src, const int64_t ld, const int row0, const int nrows,
                                                     double* __restrict__ dst) {
   const int i = blockIdx.x * 256 + threadIdx.x, c = blockIdx.y;
   if (i < nrows) dst[static_cast<int64_t>(c) * ld + row0 + i] = src[static_cast<int64_t>(c) * nrows + i];
}

// solution entries of pivots owned by other ranks are zeroed before the sum over ranks puts the vector together
__global__ void k_keep_owned(int64_t n, const signed char* __restrict__ col_owner, const int rank, double* __restrict__ xg) {
   const int64_t g = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
   if (g < n && col_owner[g] != rank) xg[g] = 0.0;
}

// ------------------------------------------------------------------------------------------------
// K5 prepare: zero the front's contribution block and extend-add the children into the front.
// CTA (parent, column block) owns EA_CBW parent columns and pulls the matching columns of every
// child's contribution block, children in ascending order -> no atomics, fixed summation order.
// Parent column < k lands in the factor panel, column >= k in the parent's contribution block.
// ------------------------------------------------------------------------------------------------
// Contribution blocks are stored packed: the lower triangle of the r x r block by columns, column j holding rows j .. r-1.
// cb_col(r, j) + i is the position of entry (i, j), i >= j.
__device__ __forceinline__ int64_t cb_col(const int r, const int j) {
   return static_cast<int64_t>(j) * r - (static_cast<int64_t>(j) * (j - 1)) / 2 - j;
}

__global__ void __launch_bounds__(256) k_prepare(FrontArrays F, const int* __restrict__ list, FactorArrays W) {
   const int p = list[blockIdx.x];
   const int fp = F.f[p], kp = F.k[p];
   const int cb0 = blockIdx.y * EA_CBW;
   if (cb0 >= fp) return;
   const int cb1 = min(cb0 + EA_CBW, fp);
   const int ldp = F.ld[p], rp = fp - kp;
   double* Lp = W.lpool + F.off[p];
   double* Cp = W.cbpool + F.cboff[p];
   const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
   for (int col = max(cb0, kp) + warp; col < cb1; col += 8) {
      double* dst = Cp + cb_col(rp, col - kp) - kp;
      for (int i = col + lane; i < fp; i += 32) dst[i] = 0.0;
   }
   if (F.child_ptr[p] == F.child_ptr[p + 1]) return;
   __syncthreads();
   for (int ci = F.child_ptr[p]; ci < F.child_ptr[p + 1]; ++ci) {
      const int c = F.child_idx[ci];
      const int rc = F.f[c] - F.k[c];
      const int* __restrict__ rel = F.relmap + F.rel_off[c];
      // rel is strictly increasing: columns of the child that land in [cb0, cb1) are a contiguous range
      int lo = 0, hi = rc;
      while (lo < hi) { const int mid = (lo + hi) >> 1; if (rel[mid] < cb0) lo = mid + 1; else hi = mid; }
      const int jlo = lo;
      hi = rc;
      while (lo < hi) { const int mid = (lo + hi) >> 1; if (rel[mid] < cb1) lo = mid + 1; else hi = mid; }
      const int jhi = lo;
      const double* Ac = W.cbpool + F.cboff[c];
      for (int j = jlo + warp; j < jhi; j += 8) {
         const int pc = rel[j];
         double* dst = (pc < kp) ? (Lp + static_cast<int64_t>(ldp) * pc) : (Cp + cb_col(rp, pc - kp) - kp);
         const double* src = Ac + cb_col(rc, j);
         int i = j + lane;
         for (; i + 96 < rc; i += 128) {
            const int r0 = rel[i], r1 = rel[i + 32], r2 = rel[i + 64], r3 = rel[i + 96];
            const double s0 = src[i], s1 = src[i + 32], s2 = src[i + 64], s3 = src[i + 96];
            const double d0 = dst[r0], d1 = dst[r1], d2 = dst[r2], d3 = dst[r3];
            dst[r0] = d0 + s0; dst[r1] = d1 + s1; dst[r2] = d2 + s2; dst[r3] = d3 + s3;
         }
         for (; i < rc; i += 32) dst[rel[i]] += src[i];
      }
      __syncthreads();   // next child may touch the same parent entries from another warp
   }
}

// ------------------------------------------------------------------------------------------------
// K3 panel kernels
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void dmma8x8x4(double& c0, double& c1, const double a, const double b) {
   asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// a pivot column that has to leave its front (see FactorArrays::delay); counted once
__device__ __forceinline__ void flag_delay(const FactorArrays& W, const int plan_col) {
   if (atomicExch(W.delay + plan_col, 1) == 0) atomicAdd(W.ndelay, 1);
}

struct alignas(16) BkShared {
   double B[NBI][NBI + 2];   // full symmetric diagonal block (rows 16-byte aligned); after factorization: strict lower = L11
   double d[NBI], e[NBI];
   int p[NBI], type[NBI];
   int nontrivial;
};

#ifdef B200KKT_PHASES
__device__ unsigned long long g_phase[16];
#define PHASE_BEGIN() long long ph_t = clock64()
#define PHASE_MARK(i) do { if (threadIdx.x == 0) { const long long ph_n = clock64(); atomicAdd(g_phase + (i), static_cast<unsigned long long>(ph_n - ph_t)); ph_t = ph_n; } } while (0)
#else
#define PHASE_BEGIN() do {} while (0)
#define PHASE_MARK(i) do {} while (0)
#endif

// max over the warp of non-negative doubles (negative inputs count as "absent" and give -1):
// the IEEE order of non-negative doubles is the order of their bit patterns, so two 32-bit REDUX
// operations (high words, then low words among the lanes holding the maximal high word) replace
// a five-step 64-bit shuffle butterfly.
__device__ __forceinline__ double warp_max(const double v) {
   const bool present = v >= 0.0;
   const unsigned hi = present ? static_cast<unsigned>(__double2hiint(v)) + 1u : 0u;   // +1: distinguishes +0.0 from absent
   const unsigned mhi = __reduce_max_sync(0xffffffffu, hi);
   if (mhi == 0u) return -1.0;
   const unsigned lo = (hi == mhi) ? static_cast<unsigned>(__double2loint(v)) : 0u;
   const unsigned mlo = __reduce_max_sync(0xffffffffu, lo);
   return __hiloint2double(static_cast<int>(mhi - 1u), static_cast<int>(mlo));
}

// Bunch-Kaufman (partial pivoting, pivot search confined to this block) by one warp.
// Null pivots (whole remaining column <= tol) are dropped MUMPS-style: unit pivot, zero column,
// reported as a zero eigenvalue.
// Lane i keeps row i of the (full, symmetric) Schur complement in registers; shared memory holds the mirror
// M[c][i] = a_i[c] (conflict-free stores), which by symmetry is also the matrix itself.  Everything indexed by a
// run-time column (pivot search, pivot row broadcast, interchanges) goes through the mirror, everything indexed
// statically through the registers: one elimination step costs 16 broadcast 128-bit loads, 32 fused
// multiply-adds and 32 stores per lane instead of 64 loads + 32 stores, and the division overlaps the loads.
// During the factorization column j of L lives in M[j][i] (i > j); the block is transposed at the end so that
// callers find L11 in the strict lower triangle S.B[i][j].
__device__ void bk_factor_block(BkShared& S, const int n, const int lane, const double tol, const bool nopiv = false) {
   S.p[lane] = lane;
   S.d[lane] = 1.0;
   S.e[lane] = 0.0;
   S.type[lane] = 0;
   if (lane == 0) S.nontrivial = 0;
   double a[NBI];
#pragma unroll
   for (int c = 0; c < NBI; ++c) a[c] = S.B[c][lane];
   __syncwarp();
   int j = 0;
   PHASE_BEGIN();
   while (j < n) {
      PHASE_MARK(8);
      const bool below = lane > j && lane < n;
      const double vj = S.B[j][lane];
      const double absajj = fabs(S.B[j][j]);
      // lambda = largest off-diagonal magnitude of column j, r = its first row
      const double v = below ? fabs(vj) : -1.0;
      double lam = warp_max(v);
      int r = j;
      if (lam >= 0.0) r = __ffs(__ballot_sync(0xffffffffu, v == lam)) - 1; else lam = 0.0;
      int kp = j, step = 1;
      bool isnull = false;
      if (nopiv) {
         isnull = !(absajj > tol);   // pivots in the given order; a non-positive one is the caller's answer
      } else if (!(fmax(absajj, lam) > tol)) {
         isnull = true;
      } else if (absajj >= BK_ALPHA * lam) {
         // 1x1 pivot, no interchange
      } else {
         const double sigma = warp_max((lane >= j && lane < n && lane != r) ? fabs(S.B[r][lane]) : -1.0);
         if (absajj * sigma >= BK_ALPHA * lam * lam) {
            // 1x1 pivot, no interchange
         } else if (fabs(S.B[r][r]) >= BK_ALPHA * sigma) {
            kp = r;
         } else {
            kp = r;
            step = 2;
         }
      }
      PHASE_MARK(9);
      const int kk = j + step - 1;
      if (kp != kk) {
         // symmetric interchange kk <-> kp on the mirror (its columns carry the already computed L rows), then reload the rows
         double t1 = S.B[kk][lane], t2 = S.B[kp][lane];
         S.B[kk][lane] = t2;
         S.B[kp][lane] = t1;
         __syncwarp();
         t1 = S.B[lane][kk];
         t2 = S.B[lane][kp];
         S.B[lane][kk] = t2;
         S.B[lane][kp] = t1;
         if (lane == 0) {
            const int tp = S.p[kk];
            S.p[kk] = S.p[kp];
            S.p[kp] = tp;
            S.nontrivial = 1;
         }
         __syncwarp();
#pragma unroll
         for (int c = 0; c < NBI; ++c) a[c] = S.B[c][lane];
      }
      PHASE_MARK(10);
      if (step == 1) {
         const double d = S.B[j][j];
         if (isnull) {
            __syncwarp();
            if (lane > j) S.B[j][lane] = 0.0;
            if (lane == 0) { S.d[j] = 1.0; S.e[j] = 0.0; S.type[j] = 3; }
         } else {
            const double li = below ? S.B[j][lane] / d : 0.0;
            const double2* prow = reinterpret_cast<const double2*>(S.B[j]);
#pragma unroll
            for (int c = 0; c < NBI; c += 2) {
               const double2 pc = prow[c >> 1];
               a[c] = a[c] - li * pc.x;
               a[c + 1] = a[c + 1] - li * pc.y;
            }
            __syncwarp();   // every lane has read the pivot row before the L column replaces it
#pragma unroll
            for (int c = 0; c < NBI; ++c) if (below && c > j && c < n) S.B[c][lane] = a[c];
            if (below) S.B[j][lane] = li;
            if (lane == 0) { S.d[j] = d; S.e[j] = 0.0; S.type[j] = 0; }
         }
      } else {
         const double pa = S.B[j][j], pb = S.B[j][j + 1], pc = S.B[j + 1][j + 1];
         const double det = pa * pc - pb * pb;
         const bool act = lane > j + 1 && lane < n;
         const double w1 = S.B[j][lane], w2 = S.B[j + 1][lane];
         const double l1 = act ? (pc * w1 - pb * w2) / det : 0.0;
         const double l2 = act ? (pa * w2 - pb * w1) / det : 0.0;
         const double2* prow1 = reinterpret_cast<const double2*>(S.B[j]);
         const double2* prow2 = reinterpret_cast<const double2*>(S.B[j + 1]);
#pragma unroll
         for (int cc = 0; cc < NBI; cc += 2) {
            const double2 c1 = prow1[cc >> 1], c2 = prow2[cc >> 1];
            a[cc] = a[cc] - (l1 * c1.x + l2 * c2.x);
            a[cc + 1] = a[cc + 1] - (l1 * c1.y + l2 * c2.y);
         }
         __syncwarp();
#pragma unroll
         for (int cc = 0; cc < NBI; ++cc) if (act && cc > j + 1 && cc < n) S.B[cc][lane] = a[cc];
         if (act) { S.B[j][lane] = l1; S.B[j + 1][lane] = l2; }
         if (lane == 0) {
            S.B[j][j + 1] = 0.0;
            S.d[j] = pa; S.d[j + 1] = pc; S.e[j] = pb; S.e[j + 1] = 0.0;
            S.type[j] = 1; S.type[j + 1] = 2;
         }
      }
      __syncwarp();
      PHASE_MARK(step == 1 ? 11 : 12);
      j += step;
   }
   // transpose: S.B[i][c] = L(i, c) for c < i
#pragma unroll
   for (int c = 0; c < NBI; ++c) a[c] = S.B[c][lane];
   __syncwarp();
#pragma unroll
   for (int c = 0; c < NBI; ++c) S.B[lane][c] = a[c];
   __syncwarp();
}

// Row-per-thread triangular solve against one factored NBI x NBI diagonal block held in shared memory:
//   w <- w L11^{-T},  l = w D^{-1}   (null pivots: both zero)
// Returns a bit mask of the block's pivot positions that must be delayed because of this row: a null pivot whose Schur entry in
// this row is above delay_tol, or (growth guard) an L entry above growth_limit.
__device__ __forceinline__ unsigned row_trsm(double (&w)[NBI], double (&l)[NBI], const double (*L11)[NBI + 1], const double* sd, const double* se,
                                             const int* stype, const double delay_tol, const double growth_limit) {
   unsigned mask = 0u;
   // right-looking order: once w[t] is final it is eliminated from all later columns at once, so the
   // dependent chain is NBI fused multiply-adds long instead of NBI^2 / 2
#pragma unroll
   for (int t = 0; t < NBI - 1; ++t) {
      const double wt = w[t];
#pragma unroll
      for (int c = t + 1; c < NBI; ++c) w[c] = fma(-wt, L11[c][t], w[c]);
   }
#pragma unroll
   for (int c = 0; c < NBI; ++c) {
      const int ty = stype[c];
      if (ty == 0) {
         l[c] = w[c] / sd[c];
      } else if (ty == 1) {
         if (c + 1 < NBI) {
            const double a = sd[c], b = se[c], cc = sd[c + 1];
            const double det = a * cc - b * b;
            l[c] = (cc * w[c] - b * w[c + 1]) / det;
            l[c + 1] = (a * w[c + 1] - b * w[c]) / det;
         }
      } else if (ty == 3) {
         if (fabs(w[c]) > delay_tol) mask |= 1u << c;
         l[c] = 0.0;
         w[c] = 0.0;
      }
   }
   if (growth_limit > 0.0) {
#pragma unroll
      for (int c = 0; c < NBI; ++c) if (fabs(l[c]) > growth_limit) mask |= 1u << c;
   }
   return mask;
}

// ------------------------------------------------------------------------------------------------
// Kernel A of a panel step, warp-specialised.
//   warps 0-3 ("pivot group"): Bunch-Kaufman of one NBI x NBI inner block after the other, the inverse
//     X = L11^{-1} carried along by the same row operations, X2 = D^{-1} X;
//   warps 4-7 ("update group"): for the block just factored, the rows of the NBO x NBO diagonal block
//     below it (w = a X^T, l = a X2^T) and the rank-NBI update of the rest of the diagonal block, both on
//     the FP64 tensor pipe; the next inner block is updated first and handed to the pivot group through
//     shared memory, so the Bunch-Kaufman chain - the latency-bound critical path - never waits for the bulk.
// Named barriers: 1 pivot group, 2 "X of block b ready", 3 "diagonal block b+1 ready", 4 update group.
// ------------------------------------------------------------------------------------------------
constexpr int BK4_THREADS = 128;
constexpr int XB_LD = NBI + 4;   // row stride of the X / X2 blocks: conflict-free B fragments

struct alignas(16) Bk4Shared {
   double B[NBI][NBI + 2];    // mirror M[c][i] = row i, column c of the (symmetric) Schur complement
   double Lm[NBI][NBI + 2];   // Lm[j][i] = L(i, j)
   double d[NBI], e[NBI];
   int p[NBI], type[NBI];
   int nontrivial;
};

template <int ID, int COUNT>
__device__ __forceinline__ void named_sync() { asm volatile("bar.sync %0, %1;" ::"n"(ID), "n"(COUNT) : "memory"); }
template <int ID, int COUNT>
__device__ __forceinline__ void named_arrive() { asm volatile("bar.arrive %0, %1;" ::"n"(ID), "n"(COUNT) : "memory"); }

// branch-free reciprocal / quotient for normal arguments well inside the exponent range (accepted pivots
// exceed the null-pivot tolerance in the equilibrated matrix); the quotient carries one correction step
__device__ __forceinline__ double fast_rcp(const double d) {
   double r;
   asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
   r = fma(fma(-d, r, 1.0), r, r);
   r = fma(fma(-d, r, 1.0), r, r);
   return r;
}
__device__ __forceinline__ double fast_div(const double x, const double d, const double r) {
   const double q = x * r;
   return fma(fma(-q, d, x), r, q);
}

// Bunch-Kaufman by four warps: lane i owns row i, warp w the columns [8w, 8w + 8) of the Schur complement
// (registers a[]) and of the inverse (registers x[]).  The mirror in shared memory serves everything that is
// indexed by a run-time column: pivot search, pivot row broadcast, interchanges.  One barrier per step.
__device__ void bk_factor_block4(Bk4Shared& S, const int n, const int tid, const double tol, double (&x)[8], const bool nopiv) {
   const int lane = tid & 31, w = tid >> 5, c0 = 8 * w;
   if (w == 0) {
      S.p[lane] = lane; S.d[lane] = 1.0; S.e[lane] = 0.0; S.type[lane] = 0;
      if (lane == 0) S.nontrivial = 0;
   }
   double a[8];
#pragma unroll
   for (int c = 0; c < 8; ++c) {
      a[c] = S.B[c0 + c][lane];
      x[c] = (c0 + c == lane) ? 1.0 : 0.0;
   }
   named_sync<1, BK4_THREADS>();
   int j = 0;
   while (j < n) {
      const bool below = lane > j && lane < n;
      const double vj = S.B[j][lane];
      const double djj = S.B[j][j];
      double rd = fast_rcp(djj);
      const double absajj = fabs(djj);
      // lambda = largest off-diagonal magnitude of column j, r = its first row
      const double v = below ? fabs(vj) : -1.0;
      double lam = warp_max(v);
      int r = j;
      if (lam >= 0.0) r = __ffs(__ballot_sync(0xffffffffu, v == lam)) - 1; else lam = 0.0;
      int kp = j, step = 1;
      bool isnull = false, looked = false;
      if (nopiv) {
         isnull = !(absajj > tol);
      } else if (!(fmax(absajj, lam) > tol)) {
         isnull = true;
      } else if (absajj >= BK_ALPHA * lam) {
         // 1x1 pivot, no interchange
      } else {
         looked = true;
         const double sigma = warp_max((lane >= j && lane < n && lane != r) ? fabs(S.B[r][lane]) : -1.0);
         if (absajj * sigma >= BK_ALPHA * lam * lam) {
            // 1x1 pivot, no interchange
         } else if (fabs(S.B[r][r]) >= BK_ALPHA * sigma) {
            kp = r;
         } else {
            kp = r;
            step = 2;
         }
      }
      const int kk = j + step - 1;
      if (kp != kk) {
         named_sync<1, BK4_THREADS>();   // every warp is done reading the block for the search
         if (w == 0) {
            // symmetric interchange kk <-> kp on the mirror; the L rows follow
            double t1 = S.B[kk][lane], t2 = S.B[kp][lane];
            S.B[kk][lane] = t2;
            S.B[kp][lane] = t1;
            __syncwarp();
            t1 = S.B[lane][kk];
            t2 = S.B[lane][kp];
            S.B[lane][kk] = t2;
            S.B[lane][kp] = t1;
            t1 = S.Lm[lane][kk];
            t2 = S.Lm[lane][kp];
            S.Lm[lane][kk] = t2;
            S.Lm[lane][kp] = t1;
            if (lane == 0) {
               const int tp = S.p[kk];
               S.p[kk] = S.p[kp];
               S.p[kp] = tp;
               S.nontrivial = 1;
            }
         }
         // rows kk and kp of the inverse: only the columns < j differ from the identity
         if (c0 < j) {
            const int partner = (lane == kk) ? kp : ((lane == kp) ? kk : lane);
#pragma unroll
            for (int c = 0; c < 8; ++c) {
               const double other = __shfl_sync(0xffffffffu, x[c], partner);
               if (c0 + c < j) x[c] = other;
            }
         }
         named_sync<1, BK4_THREADS>();
#pragma unroll
         for (int c = 0; c < 8; ++c) a[c] = S.B[c0 + c][lane];
         if (step == 1) rd = fast_rcp(S.B[j][j]);
      } else if (looked) {
         named_sync<1, BK4_THREADS>();   // row r has been read by everybody before its owner overwrites it
      }
      if (step == 1) {
         if (isnull) {
            if (w == 0) {
               S.Lm[j][lane] = 0.0;
               if (lane == 0) { S.d[j] = 1.0; S.e[j] = 0.0; S.type[j] = 3; }
            }
         } else {
            const double d = S.B[j][j];
            const double li = below ? fast_div(S.B[j][lane], d, rd) : 0.0;
            const double2* prow = reinterpret_cast<const double2*>(&S.B[j][c0]);
#pragma unroll
            for (int c = 0; c < 8; c += 2) {
               const double2 pc = prow[c >> 1];
               a[c] = a[c] - li * pc.x;
               a[c + 1] = a[c + 1] - li * pc.y;
            }
#pragma unroll
            for (int c = 0; c < 8; ++c) if (below && c0 + c > j && c0 + c < n) S.B[c0 + c][lane] = a[c];
            if (w == 0) {
               S.Lm[j][lane] = li;
               if (lane == 0) { S.d[j] = d; S.e[j] = 0.0; S.type[j] = 0; }
            }
            if (c0 <= j) {   // row j of the inverse is zero right of column j
#pragma unroll
               for (int c = 0; c < 8; ++c) x[c] = fma(-li, __shfl_sync(0xffffffffu, x[c], j), x[c]);
            }
         }
      } else {
         const double pa = S.B[j][j], pb = S.B[j][j + 1], pc = S.B[j + 1][j + 1];
         const double det = pa * pc - pb * pb;
         const double rdet = fast_rcp(det);
         const bool act = lane > j + 1 && lane < n;
         const double w1 = S.B[j][lane], w2 = S.B[j + 1][lane];
         const double l1 = act ? fast_div(pc * w1 - pb * w2, det, rdet) : 0.0;
         const double l2 = act ? fast_div(pa * w2 - pb * w1, det, rdet) : 0.0;
         const double2* prow1 = reinterpret_cast<const double2*>(&S.B[j][c0]);
         const double2* prow2 = reinterpret_cast<const double2*>(&S.B[j + 1][c0]);
#pragma unroll
         for (int c = 0; c < 8; c += 2) {
            const double2 q1 = prow1[c >> 1], q2 = prow2[c >> 1];
            a[c] = a[c] - (l1 * q1.x + l2 * q2.x);
            a[c + 1] = a[c + 1] - (l1 * q1.y + l2 * q2.y);
         }
#pragma unroll
         for (int c = 0; c < 8; ++c) if (act && c0 + c > j + 1 && c0 + c < n) S.B[c0 + c][lane] = a[c];
         if (w == 0) {
            S.Lm[j][lane] = l1;
            S.Lm[j + 1][lane] = l2;
            if (lane == 0) {
               S.d[j] = pa; S.d[j + 1] = pc; S.e[j] = pb; S.e[j + 1] = 0.0;
               S.type[j] = 1; S.type[j + 1] = 2;
            }
         }
         if (c0 <= j + 1) {
#pragma unroll
            for (int c = 0; c < 8; ++c) {
               const double xa = __shfl_sync(0xffffffffu, x[c], j), xb = __shfl_sync(0xffffffffu, x[c], j + 1);
               x[c] = x[c] - (l1 * xa + l2 * xb);
            }
         }
      }
      named_sync<1, BK4_THREADS>();
      j += step;
   }
}

// X (rows of the inverse in registers) -> X2 = D^{-1} X; both to shared memory ([c][t], row stride XB_LD) for the
// update group and, row-major NBI x NBI each, to the front's scratch for kernel B.  Null pivots: zero row in X2 only.
__device__ void publish_inverse(const Bk4Shared& S, const int n, const int tid, double (&x)[8], double* Xx, double* Xl, double* Xg) {
   const int lane = tid & 31, c0 = 8 * (tid >> 5);
   const int ty = S.type[lane];
   const int first = (ty == 2) ? lane - 1 : lane;   // first row of the pivot block this row belongs to
   const double pa = S.d[first], pb = S.e[first], pc = S.d[(first + 1) & 31];
   const double det = pa * pc - pb * pb;
#pragma unroll
   for (int c = 0; c < 8; ++c) {
      const double dn = __shfl_down_sync(0xffffffffu, x[c], 1), up = __shfl_up_sync(0xffffffffu, x[c], 1);
      double v = x[c], v2;
      if (ty == 0) v2 = v / pa;
      else if (ty == 1) v2 = (pc * v - pb * dn) / det;
      else if (ty == 2) v2 = (pa * v - pb * up) / det;
      else v2 = 0.0;   // null pivot: X keeps its row (w = a X^T is then the column's Schur entry, tested for a delay), X2 = 0
      if (lane >= n || c0 + c >= n) { v = 0.0; v2 = 0.0; }
      Xx[lane * XB_LD + c0 + c] = v;
      Xl[lane * XB_LD + c0 + c] = v2;
      Xg[lane * NBI + c0 + c] = v;
      Xg[NBI * NBI + lane * NBI + c0 + c] = v2;
   }
}

constexpr int DP_THREADS = 256;
constexpr int DP_LD = NBO - NBI + 4;   // rows of the in-panel W/L staging ([c][row]); DP_LD mod 16 == 4: conflict-free fragments
constexpr int DP_PB = 2 * NBI + 1;     // per buffer: pivot order of the block, its "nontrivial" flag, its pivot types
constexpr int DP_SMEM = static_cast<int>((sizeof(Bk4Shared) + 15) & ~size_t(15)) +
                        (4 * NBI * XB_LD + 2 * NBI * DP_LD) * static_cast<int>(sizeof(double)) + 2 * DP_PB * static_cast<int>(sizeof(int)) + 16;

// rows [row0, row0 + 8 MT) of the diagonal block against the inner block at j0: w = a X^T, l = a X2^T with
// a = A[row, j0 + p[.]] (the block's pivot order); results to the staging (row index rel0 + ...) and l to the factor panel
template <int MT>
__device__ __forceinline__ void panel_rows_trsm(double* __restrict__ A, const int ld, const int row0, const int rowend, const int j0, const int nbw,
                                                const int* __restrict__ pb, const double* __restrict__ Xx, const double* __restrict__ Xl,
                                                double* __restrict__ Ws, double* __restrict__ Ls, const int rel0, const int lane,
                                                const FactorArrays& W, const int colbase) {
   const int* __restrict__ tyb = pb + NBI + 1;   // pivot types of the block
   const int g = lane >> 2, q = lane & 3;
   double af[MT][8];
#pragma unroll
   for (int ks = 0; ks < 8; ++ks) {
      const int t = 4 * ks + q;
      const double* colp = A + static_cast<int64_t>(ld) * (j0 + pb[t]);
#pragma unroll
      for (int mt = 0; mt < MT; ++mt) {
         const int row = row0 + 8 * mt + g;
         af[mt][ks] = (row < rowend && t < nbw) ? colp[row] : 0.0;
      }
   }
#pragma unroll
   for (int nt = 0; nt < 4; ++nt) {
      double cw[MT][2], cl[MT][2];
#pragma unroll
      for (int mt = 0; mt < MT; ++mt) { cw[mt][0] = 0.0; cw[mt][1] = 0.0; cl[mt][0] = 0.0; cl[mt][1] = 0.0; }
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) {
         // X is lower triangular, X2 reaches one column further (2x2 pivots)
         if (ks <= 2 * nt + 2) {
            const double bl = Xl[(8 * nt + g) * XB_LD + 4 * ks + q];
#pragma unroll
            for (int mt = 0; mt < MT; ++mt) dmma8x8x4(cl[mt][0], cl[mt][1], af[mt][ks], bl);
            if (ks <= 2 * nt + 1) {
               const double bx = Xx[(8 * nt + g) * XB_LD + 4 * ks + q];
#pragma unroll
               for (int mt = 0; mt < MT; ++mt) dmma8x8x4(cw[mt][0], cw[mt][1], af[mt][ks], bx);
            }
         }
      }
#pragma unroll
      for (int mt = 0; mt < MT; ++mt)
#pragma unroll
         for (int e = 0; e < 2; ++e) {
            const int cc = 8 * nt + 2 * q + e, rel = rel0 + 8 * mt + g;
            double wv = cw[mt][e];
            const bool rin = row0 + 8 * mt + g < rowend && cc < nbw;
            if (tyb[cc] == 3) {
               if (rin && fabs(wv) > W.delay_tol) flag_delay(W, colbase + pb[cc]);
               wv = 0.0;
            } else if (rin && W.growth_limit > 0.0 && fabs(cl[mt][e]) > W.growth_limit) {
               flag_delay(W, colbase + pb[cc]);
            }
            Ws[cc * DP_LD + rel] = wv;
            Ls[cc * DP_LD + rel] = cl[mt][e];
         }
   }
   __syncwarp();
   // l back to the factor panel, rows contiguous (every a of these rows has been consumed by now)
   constexpr int R = 8 * MT;
   for (int idx = lane; idx < R * NBI; idx += 32) {
      const int rr = idx % R, cc = idx / R;
      if (row0 + rr < rowend && cc < nbw) A[(row0 + rr) + static_cast<int64_t>(ld) * (j0 + cc)] = Ls[cc * DP_LD + rel0 + rr];
   }
}

__global__ void __launch_bounds__(DP_THREADS) k_diagpanel(FrontArrays F, const int* __restrict__ list, const int J0, FactorArrays W,
                                                          const double null_tol) {
   extern __shared__ __align__(16) unsigned char dp_smem[];
   Bk4Shared& S = *reinterpret_cast<Bk4Shared*>(dp_smem);
   double* XB = reinterpret_cast<double*>(dp_smem + ((sizeof(Bk4Shared) + 15) & ~size_t(15)));   // [buffer][X | X2][c][t]
   double* Ws = XB + 4 * NBI * XB_LD;                                                           // [c][row]
   double* Ls = Ws + NBI * DP_LD;
   int* pbuf = reinterpret_cast<int*>(Ls + NBI * DP_LD);   // [buffer][DP_PB]: pivot order of the block, its "nontrivial" flag, its pivot types
   const int s = list[blockIdx.x];
   const int k = F.k[s], ld = F.ld[s], first = F.first[s];
   const int J1 = min(J0 + NBO, k);
   double* A = W.lpool + F.off[s];
   double* Xg = W.xinv + F.xoff[s] + static_cast<int64_t>(J0 / NBI) * XINV_BLOCK_DOUBLES;
   const int tid = threadIdx.x;
   if (tid < BK4_THREADS) {
      // ---------------- pivot group ----------------
      int b = 0;
      for (int j0 = J0; j0 < J1; j0 += NBI, ++b) {
         const int nbw = min(NBI, k - j0);
         if (b == 0) {
            for (int idx = tid; idx < NBI * NBI; idx += BK4_THREADS) {
               const int r = idx % NBI, c = idx / NBI;
               double v;
               if (r < nbw && c < nbw) {
                  v = (r >= c) ? A[(j0 + r) + static_cast<int64_t>(ld) * (j0 + c)] : A[(j0 + c) + static_cast<int64_t>(ld) * (j0 + r)];
               } else {
                  v = (r == c) ? 1.0 : 0.0;
               }
               S.B[c][r] = v;
            }
            named_sync<1, BK4_THREADS>();
         } else {
            named_sync<3, DP_THREADS>();   // the update group has written this block into S.B
         }
         double x[8];
         bk_factor_block4(S, nbw, tid, null_tol, x, W.no_pivot != 0);
         double* Xx = XB + (b & 1) * 2 * NBI * XB_LD;
         publish_inverse(S, nbw, tid, x, Xx, Xx + NBI * XB_LD, Xg + b * 2 * NBI * NBI);
         // the factored block back to the panel: D on the diagonal, L11 below it
         for (int idx = tid; idx < NBI * NBI; idx += BK4_THREADS) {
            const int r = idx % NBI, c = idx / NBI;
            if (r < nbw && c < nbw && r >= c) A[(j0 + r) + static_cast<int64_t>(ld) * (j0 + c)] = (r == c) ? S.d[c] : S.Lm[c][r];
         }
         if (tid < nbw) {
            const int q = first + j0 + tid;
            W.dd[q] = S.d[tid];
            W.de[q] = S.e[tid];
            W.ptype[q] = S.type[tid];
            W.lperm[q] = j0 + S.p[tid];
         }
         int* pb = pbuf + (b & 1) * DP_PB;
         if (tid < NBI) { pb[tid] = S.p[tid]; pb[NBI + 1 + tid] = (tid < nbw) ? S.type[tid] : 0; }
         if (tid == 0) { W.blkflag[first + j0] = S.nontrivial; pb[NBI] = S.nontrivial; }
         named_arrive<2, DP_THREADS>();
      }
   } else {
      // ---------------- update group ----------------
      const int wt = tid - BK4_THREADS, ww = wt >> 5, lane = wt & 31, g = lane >> 2, q = lane & 3;
      int b = 0;
      for (int j0 = J0; j0 < J1; j0 += NBI, ++b) {
         const int nbw = min(NBI, k - j0), j1 = j0 + nbw;
         const int rem = J1 - j1;   // rows (and columns) of the diagonal block still to be factored
         named_sync<2, DP_THREADS>();
         const double* Xx = XB + (b & 1) * 2 * NBI * XB_LD;
         const double* Xl = Xx + NBI * XB_LD;
         const int* pb = pbuf + (b & 1) * DP_PB;
         if (rem > 0) {
            // the next NBI rows first: 8 per warp
            panel_rows_trsm<1>(A, ld, j1 + 8 * ww, J1, j0, nbw, pb, Xx, Xl, Ws, Ls, 8 * ww, lane, W, first + j0);
            named_sync<4, BK4_THREADS>();
            {
               // next inner block: S.B = A[j1.., j1..] - Ws Ls^T, exactly symmetric (the lower triangle is mirrored)
               const int nbn = min(NBI, k - j1);
               double acc[4][2];
#pragma unroll
               for (int nt = 0; nt < 4; ++nt) { acc[nt][0] = 0.0; acc[nt][1] = 0.0; }
               double cur[4][2];
#pragma unroll
               for (int nt = 0; nt < 4; ++nt)
#pragma unroll
                  for (int e = 0; e < 2; ++e) {
                     const int r = 8 * ww + g, c = 8 * nt + 2 * q + e;
                     cur[nt][e] = (r >= c && r < nbn) ? A[(j1 + r) + static_cast<int64_t>(ld) * (j1 + c)] : 0.0;
                  }
#pragma unroll
               for (int ks = 0; ks < 8; ++ks) {
                  const double af = Ws[(4 * ks + q) * DP_LD + 8 * ww + g];
#pragma unroll
                  for (int nt = 0; nt < 4; ++nt) dmma8x8x4(acc[nt][0], acc[nt][1], af, Ls[(4 * ks + q) * DP_LD + 8 * nt + g]);
               }
#pragma unroll
               for (int nt = 0; nt < 4; ++nt)
#pragma unroll
                  for (int e = 0; e < 2; ++e) {
                     const int r = 8 * ww + g, c = 8 * nt + 2 * q + e;
                     if (r >= c) {
                        const double v = (r < nbn) ? cur[nt][e] - acc[nt][e] : ((r == c) ? 1.0 : 0.0);
                        S.B[r][c] = v;
                        S.B[c][r] = v;
                     }
                  }
            }
            named_arrive<3, DP_THREADS>();
            // the remaining rows, 32 per warp pass
            for (int slab = 1 + ww; 32 * slab < rem; slab += 4)
               panel_rows_trsm<4>(A, ld, j1 + 32 * slab, J1, j0, nbw, pb, Xx, Xl, Ws, Ls, 32 * slab, lane, W, first + j0);
            named_sync<4, BK4_THREADS>();
            // rank-NBI update of the rest of the diagonal block: 32 x 32 tiles of the lower triangle, tile (0,0) is done
            const int T32 = (rem + 31) >> 5;
            int cnt = 0;
            for (int tm = 1; tm < T32; ++tm)
               for (int tn = 0; tn <= tm; ++tn) {
                  if ((cnt++ & 3) != ww) continue;
                  double* tile = A + (j1 + 32 * tm) + static_cast<int64_t>(ld) * (j1 + 32 * tn);
                  double cur[4][4][2];
#pragma unroll
                  for (int mt = 0; mt < 4; ++mt)
#pragma unroll
                     for (int nt = 0; nt < 4; ++nt)
#pragma unroll
                        for (int e = 0; e < 2; ++e) {
                           const int row = 32 * tm + 8 * mt + g, col = 32 * tn + 8 * nt + 2 * q + e;
                           cur[mt][nt][e] = (row < rem && row >= col) ? tile[(8 * mt + g) + static_cast<int64_t>(ld) * (8 * nt + 2 * q + e)] : 0.0;
                        }
                  double acc[4][4][2];
#pragma unroll
                  for (int mt = 0; mt < 4; ++mt)
#pragma unroll
                     for (int nt = 0; nt < 4; ++nt) { acc[mt][nt][0] = 0.0; acc[mt][nt][1] = 0.0; }
#pragma unroll
                  for (int ks = 0; ks < 8; ++ks) {
                     double af[4], bf[4];
#pragma unroll
                     for (int mt = 0; mt < 4; ++mt) af[mt] = Ws[(4 * ks + q) * DP_LD + 32 * tm + 8 * mt + g];
#pragma unroll
                     for (int nt = 0; nt < 4; ++nt) bf[nt] = Ls[(4 * ks + q) * DP_LD + 32 * tn + 8 * nt + g];
#pragma unroll
                     for (int mt = 0; mt < 4; ++mt)
#pragma unroll
                        for (int nt = 0; nt < 4; ++nt) dmma8x8x4(acc[mt][nt][0], acc[mt][nt][1], af[mt], bf[nt]);
                  }
#pragma unroll
                  for (int mt = 0; mt < 4; ++mt)
#pragma unroll
                     for (int nt = 0; nt < 4; ++nt)
#pragma unroll
                        for (int e = 0; e < 2; ++e) {
                           const int row = 32 * tm + 8 * mt + g, col = 32 * tn + 8 * nt + 2 * q + e;
                           if (row < rem && row >= col) tile[(8 * mt + g) + static_cast<int64_t>(ld) * (8 * nt + 2 * q + e)] = cur[mt][nt][e] - acc[mt][nt][e];
                        }
               }
         }
         // the block's interchanges on the panel's earlier L columns (columns left of J0 are deferred)
         if (pb[NBI]) {
            for (int c = J0 + wt; c < j0; c += BK4_THREADS) {
               double* colp = A + static_cast<int64_t>(ld) * c + j0;
               double tmp[NBI];
#pragma unroll
               for (int t = 0; t < NBI; ++t) tmp[t] = (t < nbw) ? colp[pb[t]] : 0.0;
#pragma unroll
               for (int t = 0; t < NBI; ++t) if (t < nbw) colp[t] = tmp[t];
            }
         }
         named_sync<4, BK4_THREADS>();   // panel writes of this block are visible to the group before the next block's reads
      }
   }
}

constexpr int TP_THREADS = 128;
constexpr int TP_LDP = NBO - NBI + 2;   // [c][kk] staging of the block rows' earlier L entries
constexpr int TP_SMEM_BASE = static_cast<int>(sizeof(double)) * (NBI * (NBI + 1) + 2 * NBI) + 2 * NBI * static_cast<int>(sizeof(int));
constexpr int TP_SMEM = TP_SMEM_BASE + static_cast<int>(sizeof(double)) * NBI * TP_LDP;

// Kernel B of a panel step.  grid = (front, row chunk): rows below the diagonal block of the outer
// panel at J0, one row per thread, all inner blocks in sequence (left-looking inside the panel):
//   a(b) -= sum over earlier blocks of w(b') L[block b rows, block b']^T ; w(b) = a(b) L11(b)^{-T} ; l(b) = w(b) D(b)^{-1}
// Writes L in place and W = L*D into the W panel for the trailing update.
// Launched with TP_SMEM_BASE bytes when no front of the batch has more than NBI pivots (no earlier blocks).
__global__ void __launch_bounds__(TP_THREADS) k_trsmpanel(FrontArrays F, const int* __restrict__ list, const int J0, FactorArrays W) {
   extern __shared__ __align__(16) unsigned char tp_smem[];
   double (*L11)[NBI + 1] = reinterpret_cast<double (*)[NBI + 1]>(tp_smem);
   double* sd = reinterpret_cast<double*>(tp_smem) + NBI * (NBI + 1);
   double* se = sd + NBI;
   int* sp = reinterpret_cast<int*>(se + NBI);
   int* stype = sp + NBI;
   double* Lp = reinterpret_cast<double*>(stype + NBI);          // [c][kk], kk < j0 - J0
   const int s = list[blockIdx.x];
   const int k = F.k[s], f = F.f[s], ld = F.ld[s], first = F.first[s];
   const int J1 = min(J0 + NBO, k);
   const int ibeg = J1 + blockIdx.y * TP_THREADS;
   if (ibeg >= f) return;
   double* A = W.lpool + F.off[s];
   double* Wp = W.wpool + F.woff[s];
   const int tid = threadIdx.x;
   const int i = ibeg + tid;
   const bool act = i < f;
   for (int j0 = J0; j0 < J1; j0 += NBI) {
      const int nbw = min(NBI, k - j0);
      const int kprev = j0 - J0;
      __syncthreads();
      for (int idx = tid; idx < NBI * NBI; idx += TP_THREADS) {
         const int r = idx % NBI, c = idx / NBI;
         L11[r][c] = (r < nbw && c < nbw && r > c) ? A[(j0 + r) + static_cast<int64_t>(ld) * (j0 + c)] : 0.0;
      }
      {
         // TP_THREADS / NBI columns per pass, 8 passes in flight
         const int c = tid % NBI, kq = tid / NBI;
         constexpr int KSTEP = TP_THREADS / NBI;
         const double* src = A + (j0 + c) + static_cast<int64_t>(ld) * J0;
         for (int kk = kq; kk < kprev; kk += 8 * KSTEP) {
            double t[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) t[u] = (c < nbw && kk + u * KSTEP < kprev) ? src[static_cast<int64_t>(ld) * (kk + u * KSTEP)] : 0.0;
#pragma unroll
            for (int u = 0; u < 8; ++u) if (kk + u * KSTEP < kprev) Lp[c * TP_LDP + kk + u * KSTEP] = t[u];
         }
      }
      if (tid < NBI) {
         const int q = first + j0 + tid;
         const bool in = tid < nbw;
         sd[tid] = in ? W.dd[q] : 1.0;
         se[tid] = in ? W.de[q] : 0.0;
         stype[tid] = in ? W.ptype[q] : 0;
         sp[tid] = in ? (W.lperm[q] - j0) : tid;
      }
      __syncthreads();
      if (!act) continue;
      double w[NBI], l[NBI];
#pragma unroll
      for (int c = 0; c < NBI; ++c) w[c] = (c < nbw) ? A[i + static_cast<int64_t>(ld) * (j0 + sp[c])] : 0.0;
      if (kprev > 0) {
         // the row's W entries of the earlier blocks, 8 at a time, next batch in flight while this one is used
         const double* wrow = Wp + i;
         double wn[8];
#pragma unroll
         for (int u = 0; u < 8; ++u) wn[u] = wrow[static_cast<int64_t>(ld) * u];
         for (int kk = 0; kk < kprev; kk += 8) {
            double wv[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) wv[u] = wn[u];
            if (kk + 8 < kprev) {
#pragma unroll
               for (int u = 0; u < 8; ++u) wn[u] = wrow[static_cast<int64_t>(ld) * (kk + 8 + u)];
            }
#pragma unroll
            for (int u = 0; u < 8; u += 2) {
#pragma unroll
               for (int c = 0; c < NBI; ++c) {
                  const double2 lv = *reinterpret_cast<const double2*>(Lp + c * TP_LDP + kk + u);
                  w[c] = fma(-wv[u + 1], lv.y, fma(-wv[u], lv.x, w[c]));
               }
            }
         }
      }
      unsigned dmask = row_trsm(w, l, L11, sd, se, stype, W.delay_tol, W.growth_limit);
      while (dmask) {
         const int c = __ffs(dmask) - 1;
         dmask &= dmask - 1u;
         if (c < nbw) flag_delay(W, first + j0 + sp[c]);
      }
#pragma unroll
      for (int c = 0; c < NBI; ++c) {
         if (c < nbw) {
            A[i + static_cast<int64_t>(ld) * (j0 + c)] = l[c];
            Wp[i + static_cast<int64_t>(ld) * (kprev + c)] = w[c];
         }
      }
   }
}


// Kernel B, tensor-core variant for fronts with more than NBI pivots.  CTA = 64 rows below the diagonal
// block, 4 warps.  The CTA's W rows of the whole outer panel stay in shared memory; per inner block:
//   T = A[rows, block] - Wsm[rows, 0:kprev] * L[block rows, 0:kprev]^T     (DMMA.8x8x4, 16 rows per warp)
//   w = T L11^{-T}, l = w D^{-1}                                            (row per lane, 16 lanes per warp)
constexpr int TD_ROWS = 64;
constexpr int TD_THREADS = 128;
constexpr int TD_LDW = NBO + 4;          // row stride of Wsm: 4 mod 16 -> conflict-free fragment loads
constexpr int TD_LDP = NBO - NBI + 4;    // row stride of the staged L rows of the block
constexpr int TD_LDT = NBI + 4;          // row stride of T, X, X2
constexpr int TD_SMEM = static_cast<int>(sizeof(double)) * (TD_ROWS * TD_LDW + NBI * TD_LDP + TD_ROWS * TD_LDT + 2 * NBI * TD_LDT) +
                        2 * NBI * static_cast<int>(sizeof(int));

__global__ void __launch_bounds__(TD_THREADS) k_trsmpanel_mma(FrontArrays F, const int* __restrict__ list, const int J0, FactorArrays W) {
   extern __shared__ __align__(16) unsigned char td_smem[];
   double* Wsm = reinterpret_cast<double*>(td_smem);                 // [row][kk]
   double* Lp = Wsm + TD_ROWS * TD_LDW;                              // [c][kk]
   double* T = Lp + NBI * TD_LDP;                                    // [row][c]
   double* Xb = T + TD_ROWS * TD_LDT;                                // [c][t]: X = L11^{-1}
   double* X2b = Xb + NBI * TD_LDT;                                  // D^{-1} X
   int* sp = reinterpret_cast<int*>(X2b + NBI * TD_LDT);
   int* sty = sp + NBI;
   const int s = list[blockIdx.x];
   const int k = F.k[s], f = F.f[s], ld = F.ld[s], first = F.first[s];
   const int J1 = min(J0 + NBO, k);
   const int i0 = J1 + blockIdx.y * TD_ROWS;
   if (i0 >= f) return;
   double* A = W.lpool + F.off[s];
   double* Wp = W.wpool + F.woff[s];
   const double* Xg = W.xinv + F.xoff[s] + static_cast<int64_t>(J0 / NBI) * XINV_BLOCK_DOUBLES;
   const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
   for (int j0 = J0; j0 < J1; j0 += NBI) {
      const int nbw = min(NBI, k - j0);
      const int kprev = j0 - J0;
      __syncthreads();
      if (tid < NBI) {
         sp[tid] = (tid < nbw) ? (W.lperm[first + j0 + tid] - j0) : tid;
         sty[tid] = (tid < nbw) ? W.ptype[first + j0 + tid] : 0;
      }
      __syncthreads();
      // stage: inverted diagonal block, the block rows' earlier L entries, the CTA's rows of the block columns (pivot order)
      {
         const double* xg = Xg + (kprev / NBI) * 2 * NBI * NBI;
         for (int idx = tid; idx < 2 * NBI * NBI; idx += TD_THREADS) {
            const int which = idx / (NBI * NBI), c = (idx / NBI) % NBI, t = idx % NBI;
            (which ? X2b : Xb)[c * TD_LDT + t] = xg[idx];
         }
      }
      {
         const int c = tid % NBI, kq = tid / NBI;
         constexpr int KSTEP = TD_THREADS / NBI;
         const double* src = A + (j0 + c) + static_cast<int64_t>(ld) * J0;
         for (int kk = kq; kk < kprev; kk += 8 * KSTEP) {
            double t[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) t[u] = (c < nbw && kk + u * KSTEP < kprev) ? src[static_cast<int64_t>(ld) * (kk + u * KSTEP)] : 0.0;
#pragma unroll
            for (int u = 0; u < 8; ++u) if (kk + u * KSTEP < kprev) Lp[c * TD_LDP + kk + u * KSTEP] = t[u];
         }
      }
      {
         const int r = tid % TD_ROWS, cq = tid / TD_ROWS;   // 2 column phases
         const bool rin = i0 + r < f;
         double t[NBI / 2];
#pragma unroll
         for (int u = 0; u < NBI / 2; ++u) {
            const int c = cq + 2 * u;
            t[u] = (rin && c < nbw) ? A[(i0 + r) + static_cast<int64_t>(ld) * (j0 + sp[c])] : 0.0;
         }
#pragma unroll
         for (int u = 0; u < NBI / 2; ++u) T[r * TD_LDT + cq + 2 * u] = t[u];
      }
      __syncthreads();
      const int r0 = warp * 16 + (lane >> 2);   // this lane's fragment rows: r0, r0 + 8
      // T[16 rows of this warp] -= Wsm * Lp^T
      if (kprev > 0) {
         double acc[2][4][2];
#pragma unroll
         for (int a = 0; a < 2; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) { acc[a][b][0] = 0.0; acc[a][b][1] = 0.0; }
         const double* wa = Wsm + r0 * TD_LDW + (lane & 3);
         const double* lb = Lp + (lane >> 2) * TD_LDP + (lane & 3);
#pragma unroll 4
         for (int kk = 0; kk < kprev; kk += 4) {
            double af[2], bf[4];
#pragma unroll
            for (int a = 0; a < 2; ++a) af[a] = wa[a * 8 * TD_LDW + kk];
#pragma unroll
            for (int b = 0; b < 4; ++b) bf[b] = lb[b * 8 * TD_LDP + kk];
#pragma unroll
            for (int a = 0; a < 2; ++a)
#pragma unroll
               for (int b = 0; b < 4; ++b) dmma8x8x4(acc[a][b][0], acc[a][b][1], af[a], bf[b]);
         }
#pragma unroll
         for (int a = 0; a < 2; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b)
#pragma unroll
               for (int h = 0; h < 2; ++h) T[(r0 + a * 8) * TD_LDT + b * 8 + 2 * (lane & 3) + h] -= acc[a][b][h];
         __syncwarp();
      }
      // w = T X^T, l = T X2^T (both on the tensor pipe)
      double aw[2][4][2], al[2][4][2];
#pragma unroll
      for (int a = 0; a < 2; ++a)
#pragma unroll
         for (int b = 0; b < 4; ++b) { aw[a][b][0] = 0.0; aw[a][b][1] = 0.0; al[a][b][0] = 0.0; al[a][b][1] = 0.0; }
      {
         const double* ta = T + r0 * TD_LDT + (lane & 3);
         const double* xb = Xb + (lane >> 2) * TD_LDT + (lane & 3);
         const double* x2b = X2b + (lane >> 2) * TD_LDT + (lane & 3);
#pragma unroll
         for (int kk = 0; kk < NBI; kk += 4) {
            double af[2], bf[4], b2f[4];
#pragma unroll
            for (int a = 0; a < 2; ++a) af[a] = ta[a * 8 * TD_LDT + kk];
#pragma unroll
            for (int b = 0; b < 4; ++b) { bf[b] = xb[b * 8 * TD_LDT + kk]; b2f[b] = x2b[b * 8 * TD_LDT + kk]; }
#pragma unroll
            for (int a = 0; a < 2; ++a)
#pragma unroll
               for (int b = 0; b < 4; ++b) {
                  dmma8x8x4(aw[a][b][0], aw[a][b][1], af[a], bf[b]);
                  dmma8x8x4(al[a][b][0], al[a][b][1], af[a], b2f[b]);
               }
         }
      }
#pragma unroll
      for (int a = 0; a < 2; ++a) {
         const int r = r0 + a * 8;
         const int i = i0 + r;
#pragma unroll
         for (int b = 0; b < 4; ++b)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
               const int c = b * 8 + 2 * (lane & 3) + h;
               double wv = aw[a][b][h];
               const bool in = i < f && c < nbw;
               if (sty[c] == 3) {
                  // null pivot: X kept its row, so w is the column's Schur entry in this row
                  if (in && fabs(wv) > W.delay_tol) flag_delay(W, first + j0 + sp[c]);
                  wv = 0.0;
               } else if (in && W.growth_limit > 0.0 && fabs(al[a][b][h]) > W.growth_limit) {
                  flag_delay(W, first + j0 + sp[c]);
               }
               Wsm[r * TD_LDW + kprev + c] = wv;
               if (in) {
                  A[i + static_cast<int64_t>(ld) * (j0 + c)] = al[a][b][h];
                  Wp[i + static_cast<int64_t>(ld) * (kprev + c)] = wv;
               }
            }
      }
   }
}

// Light variant of kernel A for batches whose fronts have at most NBI pivots (the bulk of the tree):
// one warp per front, several fronts per CTA would need more shared memory than it saves; 64 threads.
__global__ void __launch_bounds__(64) k_diag32(FrontArrays F, const int* __restrict__ list, FactorArrays W, const double null_tol) {
   __shared__ BkShared S;
   const int s = list[blockIdx.x];
   const int k = F.k[s], ld = F.ld[s];
   const int nbw = min(NBI, k);
   double* A = W.lpool + F.off[s];
   const int tid = threadIdx.x;
   for (int idx = tid; idx < NBI * NBI; idx += 64) {
      const int r = idx % NBI, c = idx / NBI;
      double v;
      if (r < nbw && c < nbw) {
         v = (r >= c) ? A[r + static_cast<int64_t>(ld) * c] : A[c + static_cast<int64_t>(ld) * r];
      } else {
         v = (r == c) ? 1.0 : 0.0;
      }
      S.B[r][c] = v;
   }
   __syncthreads();
   if (tid < 32) bk_factor_block(S, nbw, tid, null_tol, W.no_pivot != 0);
   __syncthreads();
   for (int idx = tid; idx < NBI * NBI; idx += 64) {
      const int r = idx % NBI, c = idx / NBI;
      if (r < nbw && c < nbw && r >= c) A[r + static_cast<int64_t>(ld) * c] = (r == c) ? S.d[c] : S.B[r][c];
   }
   if (tid < nbw) {
      const int q = F.first[s] + tid;
      W.dd[q] = S.d[tid];
      W.de[q] = S.e[tid];
      W.ptype[q] = S.type[tid];
      W.lperm[q] = S.p[tid];
   }
   if (tid == 0) W.blkflag[F.first[s]] = S.nontrivial;
}

// K2: a whole LEAF front (no children, at most NBI pivots, at most LEAF_RMAX boundary rows) by one CTA of 64 threads, everything
// in shared memory: Bunch-Kaufman of the pivot block (warp 0), the rows below it (one row per thread: w = a L11^{-T}, l = w D^{-1}),
// and the Schur complement -W L^T written straight into the packed contribution block (a leaf's block starts from zero, so it is
// written, not updated: no k_prepare, no W panel in global memory, no separate update launch).  The bulk of every tree is such
// fronts; the level-0 batches of the BASELINE configs cost four launches and ~1 ms (config 2) before.
constexpr int LEAF_RMAX = 56;   // keeps the shared memory under 48 KB
// The W / L staging ([c][i], row stride rmax + 1) is dynamic shared memory sized from the batch's largest boundary: the leaf levels
// (boundaries of 10 - 20 rows, thousands of fronts per launch) then fit 8 CTAs per SM instead of 4.
// accumulate = 1: the batch has fronts with children; k_prepare has zeroed the contribution blocks and extend-added the children
// (into the factor panel and into the block), so the Schur complement is subtracted from the block instead of written.
__global__ void __launch_bounds__(64) k_leaf_front(FrontArrays F, const int* __restrict__ list, FactorArrays W, const double null_tol, const int accumulate,
                                                   const int rmax) {
   __shared__ BkShared S;
   __shared__ double L11[NBI][NBI + 1];
   extern __shared__ __align__(16) double leaf_dyn[];
   const int lds = rmax + 1;
   double* Ws = leaf_dyn;                 // [c][i]
   double* Ls = leaf_dyn + NBI * lds;
   const int s = list[blockIdx.x];
   const int k = F.k[s], f = F.f[s], ld = F.ld[s], first = F.first[s];
   const int r = f - k;
   double* A = W.lpool + F.off[s];
   const int tid = threadIdx.x;
   for (int idx = tid; idx < NBI * NBI; idx += 64) {
      const int rr = idx % NBI, c = idx / NBI;
      double v;
      if (rr < k && c < k) v = (rr >= c) ? A[rr + static_cast<int64_t>(ld) * c] : A[c + static_cast<int64_t>(ld) * rr];
      else v = (rr == c) ? 1.0 : 0.0;
      S.B[rr][c] = v;
   }
   __syncthreads();
   if (tid < 32) bk_factor_block(S, k, tid, null_tol, W.no_pivot != 0);
   __syncthreads();
   for (int idx = tid; idx < NBI * NBI; idx += 64) {
      const int rr = idx % NBI, c = idx / NBI;
      const bool in = rr < k && c < k;
      if (in && rr >= c) A[rr + static_cast<int64_t>(ld) * c] = (rr == c) ? S.d[c] : S.B[rr][c];
      L11[rr][c] = (in && rr > c) ? S.B[rr][c] : 0.0;
   }
   if (tid < k) {
      const int q = first + tid;
      W.dd[q] = S.d[tid];
      W.de[q] = S.e[tid];
      W.ptype[q] = S.type[tid];
      W.lperm[q] = S.p[tid];
   }
   if (tid == 0) W.blkflag[first] = S.nontrivial;
   __syncthreads();
   if (tid < r) {
      const int i = k + tid;
      double w[NBI], l[NBI];
#pragma unroll
      for (int c = 0; c < NBI; ++c) w[c] = (c < k) ? A[i + static_cast<int64_t>(ld) * S.p[c]] : 0.0;
      unsigned dmask = row_trsm(w, l, L11, S.d, S.e, S.type, W.delay_tol, W.growth_limit);
      while (dmask) {
         const int c = __ffs(dmask) - 1;
         dmask &= dmask - 1u;
         if (c < k) flag_delay(W, first + S.p[c]);
      }
#pragma unroll
      for (int c = 0; c < NBI; ++c) {
         if (c < k) A[i + static_cast<int64_t>(ld) * c] = l[c];
         Ws[c * lds + tid] = (c < k) ? w[c] : 0.0;
         Ls[c * lds + tid] = (c < k) ? l[c] : 0.0;
      }
   }
   __syncthreads();
   if (r > 0) {
      double* Cb = W.cbpool + F.cboff[s];
      // lanes along the rows of a column of the packed lower triangle: contiguous stores
      for (int j = 0; j < r; ++j) {
         const int64_t col0 = cb_col(r, j);
         for (int i = j + tid; i < r; i += 64) {
            double acc = 0.0;
            for (int c = 0; c < k; ++c) acc = fma(Ws[c * lds + i], Ls[c * lds + j], acc);
            Cb[col0 + i] = accumulate ? Cb[col0 + i] - acc : -acc;
         }
      }
   }
}

// Deferred row interchanges: the rows of an inner block were permuted among themselves by its
// Bunch-Kaufman; the L columns left of the block must follow.  Blocks touch disjoint rows, so all
// of them are applied in one launch: grid = (front, inner block, 32-column chunk); one warp moves
// the 32 rows of the block in one column (lanes along rows: coalesced).
__global__ void __launch_bounds__(256) k_permute_rows(FrontArrays F, const int* __restrict__ list, FactorArrays W) {
   const int s = list[blockIdx.x];
   const int k = F.k[s];
   const int j0 = (blockIdx.y + 1) * NBI;
   if (j0 >= k) return;
   const int first = F.first[s];
   if (!W.blkflag[first + j0]) return;
   const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
   const int nbw = min(NBI, k - j0);
   const int src = (lane < nbw) ? (W.lperm[first + j0 + lane] - j0) : lane;
   const int ld = F.ld[s];
   double* A = W.lpool + F.off[s];
   const int cend = (j0 / NBO) * NBO;   // columns of the block's own outer panel were permuted by k_diagpanel
   for (int c = blockIdx.z * 64 + warp; c < min(cend, blockIdx.z * 64 + 64); c += 8) {
      double* colp = A + static_cast<int64_t>(ld) * c + j0;
      const double v = (lane < nbw) ? colp[src] : 0.0;
      __syncwarp();
      if (lane < nbw) colp[lane] = v;
   }
}

// ------------------------------------------------------------------------------------------------
// K4 Schur-complement update on the FP64 tensor pipe:  C[m,n] -= sum_k W[m,k] * L[n,k]
// (W = L*D panel, both operands column-major with m/n contiguous -> "NT" GEMM).
// Tiles BM x BN x 16, 3-stage cp.async pipeline, warp-level mma.sync m8n8k4 f64 (SASS DMMA.8x8x4;
// tcgen05 has no f64 kind).  Only tiles touching the lower triangle are computed.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src, bool pred) {
   const unsigned saddr = static_cast<unsigned>(__cvta_generic_to_shared(smem_dst));
   const int bytes = pred ? 16 : 0;
   asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(saddr), "l"(gmem_src), "r"(bytes));
}
__device__ __forceinline__ void cp_async16_n(void* smem_dst, const void* gmem_src, const int bytes) {
   const unsigned saddr = static_cast<unsigned>(__cvta_generic_to_shared(smem_dst));
   asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(saddr), "l"(gmem_src), "r"(bytes));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }



template <int BM, int BN, int WARPS_M, int WARPS_N, int UPD_BK, int MINB = 1, int UPD_STAGES = 3>
__global__ void __launch_bounds__(WARPS_M* WARPS_N * 32, MINB) k_update(FrontArrays F, const int* __restrict__ list, const int mode, const int j0,
                                                                  const int J0, FactorArrays W, const UpdateRange R) {
   constexpr int NT = WARPS_M * WARPS_N * 32;
   constexpr int LDA = BM + 4, LDB = BN + 4;   // +4 doubles: conflict-free fragment loads (LD mod 16 == 4)
   constexpr int WTM = BM / WARPS_M, WTN = BN / WARPS_N, MF = WTM / 8, NF = WTN / 8;
   constexpr int STAGE_DOUBLES = UPD_BK * (LDA + LDB);
   extern __shared__ __align__(16) double smem[];

   const int s = list[blockIdx.x];
   const int k = F.k[s], f = F.f[s], ld = F.ld[s];
   const int J1 = min(J0 + R.kpanels * NBO, k);   // end of the K range: one outer panel, or two taken together
   int r0, c0, c1, kbeg, K;
   if (mode == 0) {
      const int j1 = min(j0 + NBI, k);
      r0 = j1; c0 = j1; c1 = min(J0 + NBO, k); kbeg = j0 - J0; K = j1 - j0;
   } else {
      // trailing matrix behind the K range, columns [J1 + clo NBO, J1 + chi NBO); R.policy: see UpdateRange.
      // A front without the second panel of a pair is not part of a two-panel update.
      if (R.kpanels == 2 && k <= J0 + NBO) return;
      r0 = J1; c0 = J1; c1 = f; kbeg = 0; K = J1 - J0;
      const bool last = k <= J0 + R.kpanels * NBO;
      if ((R.policy == 2 && last) || (R.policy == 3 && !last)) return;
      if (!(R.policy == 1 && last)) {
         c0 = J1 + R.clo * NBO;
         if (R.chi >= 0) c1 = min(f, J1 + R.chi * NBO);
         if (R.clip_to_pivots) c1 = min(c1, k);
      }
      if (R.abs_hi > 0) { c0 = max(c0, R.abs_lo); c1 = min(c1, R.abs_hi); }
   }
   if (c0 >= c1 || r0 >= f || K <= 0) return;
   int ntile = (c0 & ~1) + blockIdx.z * BN;
   if (mode != 0 && R.cyc_cols > 0) {
      // this rank's column chunks only: grid z = (own chunk, tile inside the chunk)
      const int tiles_per_chunk = R.cyc_cols / BN;
      int chunk = c0 / R.cyc_cols;
      chunk += ((R.cyc_me - R.cyc_shift - chunk) % R.cyc_g + R.cyc_g) % R.cyc_g;   // first own chunk at or behind the range start
      chunk += static_cast<int>(blockIdx.z / tiles_per_chunk) * R.cyc_g;
      const int lo = chunk * R.cyc_cols;
      if (lo >= c1) return;
      ntile = lo + static_cast<int>(blockIdx.z % tiles_per_chunk) * BN;
      c0 = max(c0, lo);
      c1 = min(c1, lo + R.cyc_cols);
      if (ntile + BN <= c0) return;
   }
   const int m0 = (r0 & ~1) + blockIdx.y * BM;
   const int n0 = ntile;
   if (m0 >= f || n0 >= c1) return;
   if (m0 + BM <= n0) return;   // tile entirely above the diagonal

   double* A = W.lpool + F.off[s];
   // columns >= k of the trailing matrix live in the contribution block (rows and columns shifted by k)
   double* Cb = W.cbpool + F.cboff[s] - k;
   const int rcb = f - k;
   // column n of the trailing matrix: in the factor panel (n < k) or in the packed contribution block
   auto col_ptr = [&](const int n) -> double* { return (n < k) ? (A + static_cast<int64_t>(ld) * n) : (Cb + cb_col(rcb, n - k)); };
   const double* Wp = W.wpool + F.woff[s] + static_cast<int64_t>(ld) * kbeg;   // operand A: rows m, K columns
   // second panel of a pair: its W panel lives in another buffer; k-tiles from NBO on read it
   const int64_t w2delta = (R.kpanels == 2) ? ((R.wpool2 + F.woff[s] - static_cast<int64_t>(ld) * NBO) - Wp) : 0;
   const double* Lp = A + static_cast<int64_t>(ld) * (J0 + kbeg);               // operand B: rows n, K columns

   const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
   const int wm0 = (warp % WARPS_M) * WTM, wn0 = (warp / WARPS_M) * WTN;

   // every thread moves the same chunks of every stage: source pointers and row predicates are fixed per thread,
   // a stage refill is a handful of pointer increments
   constexpr int A_CHUNKS = UPD_BK * (BM / 2) / NT, B_CHUNKS = UPD_BK * (BN / 2) / NT;
   static_assert(UPD_BK * (BM / 2) % NT == 0 && UPD_BK * (BN / 2) % NT == 0, "stage chunks must divide evenly over the threads");
   const double* a_src[A_CHUNKS];
   const double* b_src[B_CHUNKS];
   int a_dst[A_CHUNKS], b_dst[B_CHUNKS], a_kk[A_CHUNKS], b_kk[B_CHUNKS];
#pragma unroll
   for (int c = 0; c < A_CHUNKS; ++c) {
      const int q = tid + c * NT, kk = q / (BM / 2), mm = (q % (BM / 2)) * 2;
      const bool rowok = m0 + mm < ld;
      a_kk[c] = rowok ? kk : (1 << 20);   // out-of-range rows never pass the k test
      a_src[c] = Wp + (rowok ? (m0 + mm) : 0) + static_cast<int64_t>(ld) * kk;
      a_dst[c] = kk * LDA + mm;
   }
#pragma unroll
   for (int c = 0; c < B_CHUNKS; ++c) {
      const int q = tid + c * NT, kk = q / (BN / 2), nn = (q % (BN / 2)) * 2;
      const bool rowok = n0 + nn < ld;
      b_kk[c] = rowok ? kk : (1 << 20);
      b_src[c] = Lp + (rowok ? (n0 + nn) : 0) + static_cast<int64_t>(ld) * kk;
      b_dst[c] = UPD_BK * LDA + kk * LDB + nn;
   }
   const int64_t kstride = static_cast<int64_t>(ld) * UPD_BK;
   auto load_stage = [&](const int stage, const int kt) {
      double* base = smem + stage * STAGE_DOUBLES;
      const int kbase = kt * UPD_BK;
#pragma unroll
      for (int c = 0; c < A_CHUNKS; ++c) {
         const bool pred = kbase + a_kk[c] < K;
         cp_async16(base + a_dst[c], pred ? (a_src[c] + kstride * kt + ((kbase >= NBO) ? w2delta : 0)) : Wp, pred);
      }
#pragma unroll
      for (int c = 0; c < B_CHUNKS; ++c) {
         const bool pred = kbase + b_kk[c] < K;
         cp_async16(base + b_dst[c], pred ? (b_src[c] + kstride * kt) : Lp, pred);
      }
   };

   // pull the C tile towards L2 while the K loop runs: the epilogue's read-modify-write then hits L2
   for (int q = tid; q < BN * (BM / 16); q += NT) {
      const int n = n0 + q / (BM / 16), m = m0 + (q % (BM / 16)) * 16;
      if (n < c1 && m < f && m + 15 >= n) {
         const double* pc = col_ptr(n) + m;
         asm volatile("prefetch.global.L2 [%0];" ::"l"(pc));
      }
   }

   double acc[MF][NF][2];
#pragma unroll
   for (int a = 0; a < MF; ++a)
#pragma unroll
      for (int b = 0; b < NF; ++b) { acc[a][b][0] = 0.0; acc[a][b][1] = 0.0; }

   const int num_kt = (K + UPD_BK - 1) / UPD_BK;
#pragma unroll
   for (int st = 0; st < UPD_STAGES - 1; ++st) {
      if (st < num_kt) load_stage(st, st);
      cp_async_commit();
   }
   for (int kt = 0; kt < num_kt; ++kt) {
      cp_async_wait<UPD_STAGES - 2>();
      __syncthreads();
      const double* As = smem + (kt % UPD_STAGES) * STAGE_DOUBLES;
      const double* Bs = As + UPD_BK * LDA;
#pragma unroll
      for (int k4 = 0; k4 < UPD_BK / 4; ++k4) {
         const int krow = k4 * 4 + (lane & 3);
         double af[MF], bf[NF];
#pragma unroll
         for (int a = 0; a < MF; ++a) af[a] = As[krow * LDA + wm0 + a * 8 + (lane >> 2)];
#pragma unroll
         for (int b = 0; b < NF; ++b) bf[b] = Bs[krow * LDB + wn0 + b * 8 + (lane >> 2)];
#pragma unroll
         for (int a = 0; a < MF; ++a)
#pragma unroll
            for (int b = 0; b < NF; ++b) dmma8x8x4(acc[a][b][0], acc[a][b][1], af[a], bf[b]);
         if (k4 == 0) {
            // refill the stage consumed in the previous iteration behind the first DMMA group: the address arithmetic
            // issues in the pipe's 16-cycle slots instead of holding the tensor pipe up after the barrier
            const int nk = kt + UPD_STAGES - 1;
            if (nk < num_kt) load_stage(nk % UPD_STAGES, nk);
            cp_async_commit();
         }
      }
   }
   cp_async_wait<0>();

   // Interior tiles (fully inside the region, strictly below the diagonal, on one side of column k):
   // no per-element predicates, loads batched 2*MF deep so that the read-modify-write is not a chain
   // of dependent round trips.
   const bool interior = m0 >= r0 && m0 + BM <= f && n0 >= c0 && n0 + BN <= c1 && m0 >= n0 + BN;
   if (interior) {
      const int mrow = m0 + wm0 + (lane >> 2);
#pragma unroll
      for (int b = 0; b < NF; ++b) {
         double* cp[2];
#pragma unroll
         for (int h = 0; h < 2; ++h) cp[h] = col_ptr(n0 + wn0 + b * 8 + 2 * (lane & 3) + h) + mrow;
         double v[MF][2];
#pragma unroll
         for (int a = 0; a < MF; ++a)
#pragma unroll
            for (int h = 0; h < 2; ++h) v[a][h] = cp[h][a * 8];
#pragma unroll
         for (int a = 0; a < MF; ++a)
#pragma unroll
            for (int h = 0; h < 2; ++h) cp[h][a * 8] = v[a][h] - acc[a][b][h];
      }
      return;
   }
#pragma unroll
   for (int b = 0; b < NF; ++b) {
      double v[MF][2];
      bool ok[MF][2];
      double* dst[2];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
         const int n = n0 + wn0 + b * 8 + 2 * (lane & 3) + h;
         dst[h] = col_ptr(n);
#pragma unroll
         for (int a = 0; a < MF; ++a) {
            const int m = m0 + wm0 + a * 8 + (lane >> 2);
            ok[a][h] = m >= r0 && m < f && n >= c0 && n < c1 && m >= n;
            v[a][h] = ok[a][h] ? dst[h][m] : 0.0;
         }
      }
#pragma unroll
      for (int h = 0; h < 2; ++h)
#pragma unroll
         for (int a = 0; a < MF; ++a) {
            const int m = m0 + wm0 + a * 8 + (lane >> 2);
            if (ok[a][h]) dst[h][m] = v[a][h] - acc[a][b][h];
         }
   }
}

// ------------------------------------------------------------------------------------------------
// K4, warp-specialised persistent variant for the trailing update of big fronts (mode 1).
//   CTA = 4 consumer warps (2 x 2, warp tile 64 x 32, DMMA) + 1 producer warp; tile 128 x 64 x 16, 4-stage ring.
//   The producer moves one operand row (contiguous in the column-major panels) per lane and stage with
//   cp.async.bulk, completion counted on an mbarrier; consumers never touch addresses of the operands, never
//   hit a CTA-wide barrier, and the producer runs ahead into the next tile while they read-modify-write C.
//   Two CTAs are co-resident per SM, so one tile's epilogue overlaps the other's K loop.
//   Tiles of the lower triangle are enumerated directly (no empty CTAs), statically strided over the grid.
// ------------------------------------------------------------------------------------------------
constexpr int WS_BM = 128, WS_BN = 64, WS_BK = 16, WS_STAGES = 4;
constexpr int WS_LDA = WS_BM + 4, WS_LDB = WS_BN + 4;
constexpr int WS_STAGE_DOUBLES = WS_BK * (WS_LDA + WS_LDB);
constexpr int WS_CONSUMER_WARPS = 4;
constexpr int WS_THREADS = (WS_CONSUMER_WARPS + 1) * 32;
constexpr int WS_SMEM = WS_STAGES * WS_STAGE_DOUBLES * static_cast<int>(sizeof(double)) + 2 * WS_STAGES * 8;

__device__ __forceinline__ void mbar_init(const unsigned bar, const int count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count)); }
__device__ __forceinline__ void mbar_arrive(const unsigned bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory"); }
__device__ __forceinline__ void mbar_arrive_expect_tx(const unsigned bar, const unsigned bytes) {
   asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(const unsigned bar, const unsigned parity) {
   asm volatile(
       "{\n"
       ".reg .pred p;\n"
       "WAIT_LOOP:\n"
       "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
       "@p bra WAIT_DONE;\n"
       "bra WAIT_LOOP;\n"
       "WAIT_DONE:\n"
       "}\n" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(const unsigned dst, const void* src, const unsigned bytes, const unsigned bar) {
   asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

// tile t of a lower triangle whose row tm holds the column tiles 0 .. 2 tm + 1
__device__ __forceinline__ void decode_tile(const int t, int& tm, int& tn) {
   int r = static_cast<int>((sqrtf(4.0f * static_cast<float>(t) + 1.0f) - 1.0f) * 0.5f);
   while (r * (r + 1) > t) --r;
   while ((r + 1) * (r + 2) <= t) ++r;
   tm = r;
   tn = t - r * (r + 1);
}

__global__ void __maxnreg__(200) k_update_ws(FrontArrays F, const int* __restrict__ list, const int nfronts, const int J0, FactorArrays W) {
   extern __shared__ __align__(128) double smem[];
   const unsigned smem_base = static_cast<unsigned>(__cvta_generic_to_shared(smem));
   const unsigned bar_full = smem_base + WS_STAGES * WS_STAGE_DOUBLES * static_cast<unsigned>(sizeof(double));
   const unsigned bar_empty = bar_full + WS_STAGES * 8;
   const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
   if (tid == 0) {
      for (int i = 0; i < WS_STAGES; ++i) {
         mbar_init(bar_full + 8 * i, 32);                  // every producer lane arrives (with its byte count)
         mbar_init(bar_empty + 8 * i, WS_CONSUMER_WARPS);  // one arrival per consumer warp
      }
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
   }
   __syncthreads();
   int stage = 0;
   unsigned phase = 0;
   int base = 0;   // tiles of the fronts before the current one
   for (int fi = 0; fi < nfronts; ++fi) {
      const int s = list[fi];
      const int k = F.k[s], f = F.f[s], ld = F.ld[s];
      const int J1 = min(J0 + NBO, k);
      const int K = J1 - J0;
      if (K <= 0 || J1 >= f) continue;
      const int mbase = J1 & ~1;
      const int Tm = (f - mbase + WS_BM - 1) / WS_BM, Tn = (f - mbase + WS_BN - 1) / WS_BN;
      const int ntiles = Tm * (Tm + 1) - (2 * Tm - Tn);
      double* A = W.lpool + F.off[s];
      double* Cb = W.cbpool + F.cboff[s] - k;   // columns >= k live in the packed contribution block
      const int rcb = f - k;
      auto col_ptr = [&](const int n) -> double* { return (n < k) ? (A + static_cast<int64_t>(ld) * n) : (Cb + cb_col(rcb, n - k)); };
      const double* Wp = W.wpool + F.woff[s];                  // operand A: rows m, K columns
      const double* Lp = A + static_cast<int64_t>(ld) * J0;    // operand B: rows n, K columns
      const int num_kt = (K + WS_BK - 1) / WS_BK;
      int g = base + ((static_cast<int>(blockIdx.x) - base) % static_cast<int>(gridDim.x) + static_cast<int>(gridDim.x)) % static_cast<int>(gridDim.x);
      base += ntiles;
      for (; g < base; g += gridDim.x) {
         int tm, tn;
         decode_tile(g - (base - ntiles), tm, tn);
         const int m0 = mbase + tm * WS_BM, n0 = mbase + tn * WS_BN;
         if (warp == WS_CONSUMER_WARPS) {
            // ---------------- producer ----------------
            const bool isA = lane < WS_BK;
            const int kk = isA ? lane : lane - WS_BK;
            const unsigned bytes = static_cast<unsigned>(isA ? min(WS_BM, ld - m0) : min(WS_BN, ld - n0)) * 8u;
            const double* src = isA ? (Wp + m0 + static_cast<int64_t>(ld) * kk) : (Lp + n0 + static_cast<int64_t>(ld) * kk);
            const unsigned dst_off = static_cast<unsigned>(isA ? kk * WS_LDA : WS_BK * WS_LDA + kk * WS_LDB) * 8u;
            // pull the C tile towards L2: the epilogue's read-modify-write then hits L2
            for (int q = lane; q < WS_BN * (WS_BM / 16); q += 32) {
               const int n = n0 + q / (WS_BM / 16), m = m0 + (q % (WS_BM / 16)) * 16;
               if (n < f && m < f && m + 15 >= n) {
                  const double* pc = col_ptr(n) + m;
                  asm volatile("prefetch.global.L2 [%0];" ::"l"(pc));
               }
            }
            for (int kt = 0; kt < num_kt; ++kt) {
               mbar_wait(bar_empty + 8 * stage, phase ^ 1u);
               const unsigned dst = smem_base + static_cast<unsigned>(stage * WS_STAGE_DOUBLES) * 8u + dst_off;
               if (kt * WS_BK + kk < K) {
                  mbar_arrive_expect_tx(bar_full + 8 * stage, bytes);
                  bulk_g2s(dst, src + static_cast<int64_t>(ld) * (kt * WS_BK), bytes, bar_full + 8 * stage);
               } else {
                  // K tail: this operand row does not exist, the consumers must see zeros
                  double* z = smem + stage * WS_STAGE_DOUBLES + (dst_off >> 3);
                  for (int i = 0; i < (isA ? WS_BM : WS_BN); i += 2) *reinterpret_cast<double2*>(z + i) = make_double2(0.0, 0.0);
                  mbar_arrive(bar_full + 8 * stage);
               }
               if (++stage == WS_STAGES) { stage = 0; phase ^= 1u; }
            }
         } else {
            // ---------------- consumers ----------------
            constexpr int MF = 8, NF = 4;
            const int wm0 = (warp & 1) * 64, wn0 = (warp >> 1) * 32;
            double acc[MF][NF][2];
#pragma unroll
            for (int a = 0; a < MF; ++a)
#pragma unroll
               for (int b = 0; b < NF; ++b) { acc[a][b][0] = 0.0; acc[a][b][1] = 0.0; }
            for (int kt = 0; kt < num_kt; ++kt) {
               mbar_wait(bar_full + 8 * stage, phase);
               const double* As = smem + stage * WS_STAGE_DOUBLES;
               const double* Bs = As + WS_BK * WS_LDA;
#pragma unroll
               for (int k4 = 0; k4 < WS_BK / 4; ++k4) {
                  const int krow = k4 * 4 + (lane & 3);
                  double af[MF], bf[NF];
#pragma unroll
                  for (int a = 0; a < MF; ++a) af[a] = As[krow * WS_LDA + wm0 + a * 8 + (lane >> 2)];
#pragma unroll
                  for (int b = 0; b < NF; ++b) bf[b] = Bs[krow * WS_LDB + wn0 + b * 8 + (lane >> 2)];
#pragma unroll
                  for (int a = 0; a < MF; ++a)
#pragma unroll
                     for (int b = 0; b < NF; ++b) dmma8x8x4(acc[a][b][0], acc[a][b][1], af[a], bf[b]);
               }
               __syncwarp();
               if (lane == 0) mbar_arrive(bar_empty + 8 * stage);
               if (++stage == WS_STAGES) { stage = 0; phase ^= 1u; }
            }
            // epilogue: C -= acc.  Interior tiles (strictly below the diagonal, fully inside, on one side of column k)
            // take the path without per-element predicates; loads are batched so that the read-modify-write is not a
            // chain of dependent round trips.
            const bool interior = m0 >= J1 && m0 + WS_BM <= f && n0 >= J1 && n0 + WS_BN <= f && m0 >= n0 + WS_BN;
            if (interior) {
               const int mrow = m0 + wm0 + (lane >> 2);
#pragma unroll
               for (int b = 0; b < NF; ++b) {
                  double* cp[2];
#pragma unroll
                  for (int h = 0; h < 2; ++h) cp[h] = col_ptr(n0 + wn0 + b * 8 + 2 * (lane & 3) + h) + mrow;
                  double v[MF][2];
#pragma unroll
                  for (int a = 0; a < MF; ++a)
#pragma unroll
                     for (int h = 0; h < 2; ++h) v[a][h] = cp[h][a * 8];
#pragma unroll
                  for (int a = 0; a < MF; ++a)
#pragma unroll
                     for (int h = 0; h < 2; ++h) cp[h][a * 8] = v[a][h] - acc[a][b][h];
               }
            } else {
#pragma unroll
               for (int b = 0; b < NF; ++b) {
                  double v[MF][2];
                  bool ok[MF][2];
                  double* dst[2];
#pragma unroll
                  for (int h = 0; h < 2; ++h) {
                     const int n = n0 + wn0 + b * 8 + 2 * (lane & 3) + h;
                     dst[h] = col_ptr(n);
#pragma unroll
                     for (int a = 0; a < MF; ++a) {
                        const int m = m0 + wm0 + a * 8 + (lane >> 2);
                        ok[a][h] = m >= J1 && m < f && n >= J1 && n < f && m >= n;
                        v[a][h] = ok[a][h] ? dst[h][m] : 0.0;
                     }
                  }
#pragma unroll
                  for (int h = 0; h < 2; ++h)
#pragma unroll
                     for (int a = 0; a < MF; ++a) {
                        const int m = m0 + wm0 + a * 8 + (lane >> 2);
                        if (ok[a][h]) dst[h][m] = v[a][h] - acc[a][b][h];
                     }
               }
            }
         }
      }
   }
}

// ------------------------------------------------------------------------------------------------
// K6 inertia from the D blocks
// ------------------------------------------------------------------------------------------------
__global__ void k_inertia(int64_t n, FactorArrays W, int* counters, const signed char* __restrict__ col_owner, const int rank) {
   int pos = 0, neg = 0, zero = 0, two = 0, bad = 0;
   for (int64_t q = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; q < n; q += static_cast<int64_t>(gridDim.x) * blockDim.x) {
      if (col_owner != nullptr && col_owner[q] != rank) continue;
      const int ty = W.ptype[q];
      if (ty == 0) {
         const double d = W.dd[q];
         if (d > 0.0) ++pos; else if (d < 0.0) ++neg; else ++bad;   // NaN (or an exact zero that slipped through): an error, not a sign
      } else if (ty == 1) {
         const double a = W.dd[q], b = W.de[q], c = W.dd[q + 1];
         const double det = a * c - b * b;
         if (det < 0.0) { ++pos; ++neg; }
         else if (!(det >= 0.0)) bad += 2;
         else if (a > 0.0) pos += 2;
         else neg += 2;
         ++two;
      } else if (ty == 3) {
         ++zero;
      }
   }
   for (int o = 16; o; o >>= 1) {
      pos += __shfl_xor_sync(0xffffffffu, pos, o);
      neg += __shfl_xor_sync(0xffffffffu, neg, o);
      zero += __shfl_xor_sync(0xffffffffu, zero, o);
      two += __shfl_xor_sync(0xffffffffu, two, o);
      bad += __shfl_xor_sync(0xffffffffu, bad, o);
   }
   if ((threadIdx.x & 31) == 0) {
      if (pos) atomicAdd(counters + 0, pos);
      if (neg) atomicAdd(counters + 1, neg);
      if (zero) atomicAdd(counters + 2, zero);
      if (two) atomicAdd(counters + 3, two);
      if (bad) atomicAdd(counters + 6, bad);
   }
}

// ------------------------------------------------------------------------------------------------
// K7 solves.  One CTA per front; fronts of one level are independent.
// Forward: v = [rhs pivots ; 0] + sum children boundary vectors, y1 = L11^{-1} P v1, boundary -= L21 y1.
// Backward: y1 = D^{-1} y1 - L21^T x_boundary, x1 = L11^{-T} y1, scattered back through P^T.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_forward(FrontArrays F, const int* __restrict__ list, FactorArrays W, double* __restrict__ zg,
                                                 double* __restrict__ z2, double* __restrict__ cpool) {
   __shared__ double Ls[NBI][NBI + 1];
   __shared__ double yp[NBI];
   const int s = list[blockIdx.x];
   const int k = F.k[s], f = F.f[s], ld = F.ld[s], first = F.first[s];
   const int r = f - k;
   const double* A = W.lpool + F.off[s];
   double* cb = cpool + F.rel_off[s];
   const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nt = blockDim.x;

   for (int i = tid; i < r; i += nt) cb[i] = 0.0;
   __syncthreads();
   for (int ci = F.child_ptr[s]; ci < F.child_ptr[s + 1]; ++ci) {
      const int c = F.child_idx[ci];
      const int rc = F.f[c] - F.k[c];
      const int* __restrict__ rel = F.relmap + F.rel_off[c];
      const double* cc = cpool + F.rel_off[c];
      for (int i = tid; i < rc; i += nt) {
         const int loc = rel[i];
         const double v = cc[i];
         if (loc < k) zg[first + loc] += v; else cb[loc - k] += v;
      }
      __syncthreads();
   }
   double* y = z2 + first;
   for (int j = tid; j < k; j += nt) y[j] = zg[first + W.lperm[first + j]];
   __syncthreads();

   for (int jb = 0; jb < k; jb += NBI) {
      const int nb = min(NBI, k - jb);
      for (int idx = tid; idx < NBI * NBI; idx += nt) {
         const int rr = idx % NBI, c = idx / NBI;
         Ls[rr][c] = (rr < nb && c < nb && rr > c) ? A[(jb + rr) + static_cast<int64_t>(ld) * (jb + c)] : 0.0;
      }
      __syncthreads();
      if (warp == 0) {
         double yv = (lane < nb) ? y[jb + lane] : 0.0;
         for (int c = 0; c < nb; ++c) {
            const double yc = __shfl_sync(0xffffffffu, yv, c);
            if (lane > c) yv -= Ls[lane][c] * yc;
         }
         if (lane < nb) y[jb + lane] = yv;
         yp[lane] = (lane < nb) ? yv : 0.0;
      }
      __syncthreads();
      for (int i = jb + nb + tid; i < f; i += nt) {
         double acc = 0.0;
         for (int c = 0; c < nb; ++c) acc += A[i + static_cast<int64_t>(ld) * (jb + c)] * yp[c];
         if (i < k) y[i] -= acc; else cb[i - k] -= acc;
      }
      __syncthreads();
   }
}

__global__ void __launch_bounds__(256) k_backward(FrontArrays F, const int* __restrict__ list, FactorArrays W, double* __restrict__ z2,
                                                  double* __restrict__ xg, double* __restrict__ cpool) {
   __shared__ double Ls[NBI][NBI + 1];
   __shared__ double red[NBI];
   const int s = list[blockIdx.x];
   const int k = F.k[s], f = F.f[s], ld = F.ld[s], first = F.first[s];
   const int r = f - k;
   const double* A = W.lpool + F.off[s];
   double* xb = cpool + F.rel_off[s];
   const int* __restrict__ rows = F.rows + F.rows_off[s];
   double* y = z2 + first;
   const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nt = blockDim.x, nwarps = blockDim.x >> 5;

   for (int i = tid; i < r; i += nt) xb[i] = xg[rows[k + i]];
   for (int j = tid; j < k; j += nt) {
      const int q = first + j;
      const int ty = W.ptype[q];
      if (ty == 0) {
         y[j] = y[j] / W.dd[q];
      } else if (ty == 1) {
         const double a = W.dd[q], b = W.de[q], c = W.dd[q + 1];
         const double det = a * c - b * b;
         const double y1 = y[j], y2 = y[j + 1];
         y[j] = (c * y1 - b * y2) / det;
         y[j + 1] = (a * y2 - b * y1) / det;
      } else if (ty == 3) {
         y[j] = 0.0;
      }
   }
   __syncthreads();

   const int nblocks = (k + NBI - 1) / NBI;
   for (int bi = nblocks - 1; bi >= 0; --bi) {
      const int jb = bi * NBI;
      const int nb = min(NBI, k - jb);
      // red[c] = sum_{i >= jb+nb} L[i][jb+c] * x(i)
      for (int c = warp; c < nb; c += nwarps) {
         const double* col = A + static_cast<int64_t>(ld) * (jb + c);
         double acc = 0.0;
         for (int i = jb + nb + lane; i < f; i += 32) acc += col[i] * ((i < k) ? y[i] : xb[i - k]);
         for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
         if (lane == 0) red[c] = acc;
      }
      for (int idx = tid; idx < NBI * NBI; idx += nt) {
         const int rr = idx % NBI, c = idx / NBI;
         Ls[rr][c] = (rr < nb && c < nb && rr > c) ? A[(jb + rr) + static_cast<int64_t>(ld) * (jb + c)] : 0.0;
      }
      __syncthreads();
      if (warp == 0) {
         double yv = (lane < nb) ? (y[jb + lane] - red[lane]) : 0.0;
         for (int c = nb - 1; c >= 0; --c) {
            const double yc = __shfl_sync(0xffffffffu, yv, c);
            if (lane < c) yv -= Ls[c][lane] * yc;
         }
         if (lane < nb) y[jb + lane] = yv;
      }
      __syncthreads();
   }
   for (int j = tid; j < k; j += nt) xg[first + W.lperm[first + j]] = y[j];
}

// ------------------------------------------------------------------------------------------------
// K7 for big fronts: many CTAs per front, NBS-wide panel steps (HBM-bound: every L entry is read
// once per sweep, coalesced along the rows).
// ------------------------------------------------------------------------------------------------
constexpr int GATHER_SPAN = 1024;   // parent-local destinations owned by one gather CTA

// v = [rhs pivots ; 0] + sum over children of their boundary vectors.  CTA (front, destination range):
// children in ascending order, each child hits a destination at most once -> no atomics.
__global__ void __launch_bounds__(256) k_fwd_big_gather(FrontArrays F, const int* __restrict__ list, double* __restrict__ zg,
                                                         double* __restrict__ cpool) {
   const int s = list[blockIdx.x];
   const int k = F.k[s], f = F.f[s], first = F.first[s];
   const int d0 = blockIdx.y * GATHER_SPAN;
   if (d0 >= f) return;
   const int d1 = min(d0 + GATHER_SPAN, f);
   double* cb = cpool + F.rel_off[s];
   const int tid = threadIdx.x;
   for (int d = max(d0, k) + tid; d < d1; d += 256) cb[d - k] = 0.0;
   __syncthreads();
   for (int ci = F.child_ptr[s]; ci < F.child_ptr[s + 1]; ++ci) {
      const int c = F.child_idx[ci];
      const int rc = F.f[c] - F.k[c];
      const int* __restrict__ rel = F.relmap + F.rel_off[c];
      const double* cc = cpool + F.rel_off[c];
      int lo = 0, hi = rc;
      while (lo < hi) { const int mid = (lo + hi) >> 1; if (rel[mid] < d0) lo = mid + 1; else hi = mid; }
      const int ilo = lo;
      hi = rc;
      while (lo < hi) { const int mid = (lo + hi) >> 1; if (rel[mid] < d1) lo = mid + 1; else hi = mid; }
      const int ihi = lo;
      for (int i = ilo + tid; i < ihi; i += 256) {
         const int loc = rel[i];
         const double v = cc[i];
         if (loc < k) zg[first + loc] += v; else cb[loc - k] += v;
      }
      __syncthreads();
   }
}

__global__ void __launch_bounds__(256) k_fwd_big_gather_list(FrontArrays F, GatherArrays Gt, const int* __restrict__ list, double* __restrict__ zg,
                                                              double* __restrict__ cpool) {
   const int s = list[blockIdx.x];
   const int k = F.k[s], f = F.f[s];
   const int d = blockIdx.y * 256 + threadIdx.x;
   if (d >= f) return;
   const int* __restrict__ ptr = Gt.ptr + Gt.off[s] + d;
   // pivot rows start from their right-hand side entry, boundary rows from zero: the running sums of k_fwd_big_gather
   double v = (d < k) ? zg[F.first[s] + d] : 0.0;
   for (int t = ptr[0]; t < ptr[1]; ++t) v += cpool[Gt.src[t]];
   if (d < k) zg[F.first[s] + d] = v;
   else cpool[F.rel_off[s] + (d - k)] = v;
}

__global__ void __launch_bounds__(256) k_fwd_big_permute(FrontArrays F, const int* __restrict__ list, FactorArrays W,
                                                          const double* __restrict__ zg, double* __restrict__ z2) {
   const int s = list[blockIdx.x];
   const int j = blockIdx.y * 256 + threadIdx.x;
   if (j >= F.k[s]) return;
   const int first = F.first[s];
   z2[first + j] = zg[first + W.lperm[first + j]];
}

// the strict block-lower triangle of an NBS x NBS diagonal block in NBI-wide strips (28 blocks of NBI x NBI) plus the panel's vector
constexpr int SD_STRIP_DOUBLES = NBI * NBI * ((NBS / NBI) * (NBS / NBI - 1) / 2);
constexpr int SD_SMEM = static_cast<int>(sizeof(double)) * (SD_STRIP_DOUBLES + (NBS / NBI) * 2 + NBS);
static_assert(SD_SMEM <= 232448, "solve panel kernels: shared memory");
static_assert(SD_STRIP_DOUBLES >= (NBS / NBI) * NBI * (NBI + 1), "fallback path of the solve panel kernels uses the same buffer");

// prefetch the lower triangle of the NBS x NBS diagonal block of the panel at jp towards L2
__device__ __forceinline__ void prefetch_diag_block(const double* A, const int ld, const int jp, const int nb) {
   for (int q = threadIdx.x; q < nb * (NBS / 16); q += blockDim.x) {
      const int c = q / (NBS / 16), r = (q % (NBS / 16)) * 16;
      if (r + 15 >= c && r < nb) asm volatile("prefetch.global.L2 [%0];" ::"l"(A + (jp + r) + static_cast<int64_t>(ld) * (jp + c)));
   }
}

// all NBI x NBI diagonal blocks (strict lower parts) of the panel into shared memory with one batch of loads
__device__ __forceinline__ void load_diag_blocks(double* Ls, const double* A, const int ld, const int jp, const int nb) {
   for (int idx = threadIdx.x; idx < (NBS / NBI) * NBI * NBI; idx += blockDim.x) {
      const int sb = idx / (NBI * NBI), rr = idx % NBI, c = (idx / NBI) % NBI;
      const int gr = sb * NBI + rr, gc = sb * NBI + c;
      Ls[(sb * NBI + rr) * (NBI + 1) + c] = (gr < nb && gc < nb && rr > c) ? A[(jp + gr) + static_cast<int64_t>(ld) * (jp + gc)] : 0.0;
   }
}

// inverted diagonal blocks X of the panel at jp into shared memory: [block][r][c], row stride NBI + 1
__device__ __forceinline__ void load_inverse_blocks(double* Xs, const double* Xg, const int nsb) {
   for (int idx = threadIdx.x; idx < nsb * NBI * NBI; idx += blockDim.x) {
      const int sb = idx / (NBI * NBI), r = (idx / NBI) % NBI, c = idx % NBI;
      Xs[(sb * NBI + r) * (NBI + 1) + c] = Xg[static_cast<int64_t>(sb) * XINV_BLOCK_DOUBLES + r * NBI + c];
   }
}

// One CTA per front: y_P <- L11[P,P]^{-1} y_P for the NBS-wide pivot panel starting at jp.
// Fronts factored by the panel kernels carry X = inverse of every NBI x NBI diagonal block: a block step is a
// matrix-vector product instead of a 32-step substitution, and the strip of L below the next block is in flight
// (registers) while the current block is applied.
__global__ void __launch_bounds__(256) k_fwd_big_diag(FrontArrays F, const int* __restrict__ list, const int jp, FactorArrays W,
                                                       double* __restrict__ z2) {
   extern __shared__ __align__(16) double sd_smem[];
   double* Ls = sd_smem;                                   // strips of L (or, fallback, [sb][rr][c] with row stride NBI + 1)
   double* ys = sd_smem + SD_STRIP_DOUBLES + (NBS / NBI) * 2;
   const int s = list[blockIdx.x];
   const int k = F.k[s], ld = F.ld[s];
   const int nb = min(NBS, k - jp);
   const double* A = W.lpool + F.off[s];
   double* y = z2 + F.first[s] + jp;
   const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
   const int64_t xoff = F.xoff[s];
   if (xoff >= 0) {
      // everything the panel needs is requested at once: the strips of L below every inner block (cp.async, rows contiguous)
      // and, by warp 0, the rows of the inverted diagonal blocks one block ahead; then 8 short steps out of shared memory
      const int nsb = (nb + NBI - 1) / NBI;
      const double* Xg = W.xinv + xoff + static_cast<int64_t>(jp / NBI) * XINV_BLOCK_DOUBLES;
      {
         int off = 0;
         for (int b = 0; b + 1 < nsb; ++b) {
            const int r0 = (b + 1) * NBI, R = nb - r0, Rp = (R + 1) & ~1, chunks = Rp >> 1;
            for (int idx = tid; idx < NBI * chunks; idx += 256) {
               const int c = idx / chunks, ch = idx % chunks;
               cp_async16_n(Ls + off + c * Rp + 2 * ch, A + (jp + r0 + 2 * ch) + static_cast<int64_t>(ld) * (jp + b * NBI + c), (2 * ch + 1 < R) ? 16 : 8);
            }
            off += NBI * Rp;
         }
         cp_async_commit();
      }
      ys[tid] = (tid < nb) ? y[tid] : 0.0;
      // warp w holds the inverse of inner block w (lane = row): all blocks are requested at once, block sb is applied by warp sb
      double2 xr[NBI / 2];
      if (warp < nsb) {
         const double2* xp = reinterpret_cast<const double2*>(Xg + static_cast<int64_t>(warp) * XINV_BLOCK_DOUBLES + lane * NBI);
#pragma unroll
         for (int c = 0; c < NBI / 2; ++c) xr[c] = xp[c];
      }
      cp_async_wait<0>();
      __syncthreads();
      int off = 0;
      for (int sb = 0; sb < nsb; ++sb) {
         const int base = sb * NBI;
         if (warp == sb) {
            double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
#pragma unroll
            for (int c = 0; c < NBI / 2; c += 2) {
               a0 = fma(xr[c].x, ys[base + 2 * c], a0);
               a1 = fma(xr[c].y, ys[base + 2 * c + 1], a1);
               a2 = fma(xr[c + 1].x, ys[base + 2 * c + 2], a2);
               a3 = fma(xr[c + 1].y, ys[base + 2 * c + 3], a3);
            }
            __syncwarp();
            ys[base + lane] = (a0 + a1) + (a2 + a3);
         }
         __syncthreads();
         if (sb + 1 < nsb) {
            const int r0 = base + NBI, R = nb - r0, Rp = (R + 1) & ~1;
            if (tid >= r0 && tid < nb) {
               const double* S = Ls + off + (tid - r0);
               double a0 = 0.0, a1 = 0.0;
#pragma unroll
               for (int c = 0; c < NBI; c += 2) {
                  a0 = fma(S[c * Rp], ys[base + c], a0);
                  a1 = fma(S[(c + 1) * Rp], ys[base + c + 1], a1);
               }
               ys[tid] -= a0 + a1;
            }
            off += NBI * Rp;
            __syncthreads();
         }
      }
      if (tid < nb) y[tid] = ys[tid];
      return;
   }
   prefetch_diag_block(A, ld, jp, nb);
   load_diag_blocks(Ls, A, ld, jp, nb);
   ys[tid] = (tid < nb) ? y[tid] : 0.0;
   __syncthreads();
   for (int sb = 0; sb * NBI < nb; ++sb) {
      const int base = sb * NBI;
      const int nbb = min(NBI, nb - base);
      if (warp == 0) {
         const double* Lb = Ls + base * (NBI + 1);
         double yv = (lane < nbb) ? ys[base + lane] : 0.0;
         for (int c = 0; c < nbb; ++c) {
            const double yc = __shfl_sync(0xffffffffu, yv, c);
            if (lane > c) yv -= Lb[lane * (NBI + 1) + c] * yc;
         }
         if (lane < nbb) ys[base + lane] = yv;
      }
      __syncthreads();
      const int i = base + nbb + tid;
      if (i < nb) {
         const double* row = A + (jp + i) + static_cast<int64_t>(ld) * (jp + base);
         double v[NBI];
#pragma unroll
         for (int c = 0; c < NBI; ++c) v[c] = (c < nbb) ? row[static_cast<int64_t>(ld) * c] : 0.0;
         double acc = 0.0;
#pragma unroll
         for (int c = 0; c < NBI; ++c) acc += v[c] * ys[base + c];
         ys[i] -= acc;
      }
      __syncthreads();
   }
   if (tid < nb) y[tid] = ys[tid];
}

// rows below the panel: v(i) -= L[i, P] . y_P.  CTA = 32 rows x 8 column groups of 32: every thread has its 32 loads in
// flight at once (one DRAM round trip per CTA), lanes along the rows (coalesced), partial sums reduced through shared memory.
__global__ void __launch_bounds__(256) k_fwd_big_gemv(FrontArrays F, const int* __restrict__ list, const int jp, FactorArrays W,
                                                       double* __restrict__ z2, double* __restrict__ cpool) {
   __shared__ double ys[NBS];
   __shared__ double red[8][32];
   const int s = list[blockIdx.x];
   const int k = F.k[s], f = F.f[s], ld = F.ld[s], first = F.first[s];
   const int nb = min(NBS, k - jp);
   const int r0 = jp + nb + blockIdx.y * 32;
   if (r0 >= f) return;
   const double* A = W.lpool + F.off[s];
   const int tid = threadIdx.x;
   const int i = r0 + (tid & 31), q = tid >> 5;
   const int c0 = q * 32;
   double v[32];
   const double* row = A + i + static_cast<int64_t>(ld) * (jp + c0);
#pragma unroll
   for (int u = 0; u < 32; ++u) v[u] = (i < f && c0 + u < nb) ? row[static_cast<int64_t>(ld) * u] : 0.0;
   ys[tid] = (tid < nb) ? z2[first + jp + tid] : 0.0;
   __syncthreads();
   double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
#pragma unroll
   for (int u = 0; u < 32; u += 4) {
      a0 = fma(v[u], ys[c0 + u], a0);
      a1 = fma(v[u + 1], ys[c0 + u + 1], a1);
      a2 = fma(v[u + 2], ys[c0 + u + 2], a2);
      a3 = fma(v[u + 3], ys[c0 + u + 3], a3);
   }
   red[q][tid & 31] = (a0 + a1) + (a2 + a3);
   __syncthreads();
   if (q == 0 && i < f) {
      const double sum = ((red[0][tid] + red[1][tid]) + (red[2][tid] + red[3][tid])) + ((red[4][tid] + red[5][tid]) + (red[6][tid] + red[7][tid]));
      if (i < k) z2[first + i] -= sum; else cpool[F.rel_off[s] + (i - k)] -= sum;
   }
}

// backward sweep, big fronts: boundary gather + block-diagonal solve
__global__ void __launch_bounds__(256) k_bwd_big_begin(FrontArrays F, const int* __restrict__ list, FactorArrays W, double* __restrict__ z2,
                                                        const double* __restrict__ xg, double* __restrict__ cpool) {
   const int s = list[blockIdx.x];
   const int k = F.k[s], f = F.f[s], first = F.first[s];
   const int idx = blockIdx.y * 256 + threadIdx.x;
   if (idx >= f) return;
   if (idx >= k) {
      cpool[F.rel_off[s] + (idx - k)] = xg[F.rows[F.rows_off[s] + idx]];
      return;
   }
   const int q = first + idx;
   const int ty = W.ptype[q];
   double* y = z2 + q;
   if (ty == 0) {
      y[0] = y[0] / W.dd[q];
   } else if (ty == 1) {
      const double a = W.dd[q], b = W.de[q], c = W.dd[q + 1];
      const double det = a * c - b * b;
      const double y1 = y[0], y2 = y[1];
      y[0] = (c * y1 - b * y2) / det;
      y[1] = (a * y2 - b * y1) / det;
   } else if (ty == 3) {
      y[0] = 0.0;
   }
}

constexpr int BW_SPLITS = 16;   // row splits of the backward panel product (partial sums, reduced in fixed order)

// partial(split, c) = sum over the split's rows i below the panel of L[i, c] x(i), for the panel columns c.
// grid = (front, NBS / 8 column groups, BW_SPLITS): one warp per column, lanes along the rows (coalesced).
__global__ void __launch_bounds__(256) k_bwd_big_gemv(FrontArrays F, const int* __restrict__ list, const int jp, FactorArrays W,
                                                       const double* __restrict__ z2, const double* __restrict__ cpool,
                                                       double* __restrict__ partial) {
   const int s = list[blockIdx.x];
   const int k = F.k[s], f = F.f[s], ld = F.ld[s], first = F.first[s];
   const int nb = min(NBS, k - jp);
   const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
   const int c = blockIdx.y * 8 + warp;
   if (c >= nb) return;
   const int rbeg = jp + nb, rows = f - rbeg;
   const int chunk = ((rows + BW_SPLITS - 1) / BW_SPLITS + 31) & ~31;
   const int i0 = rbeg + blockIdx.z * chunk, i1 = min(f, i0 + chunk);
   const double* col = W.lpool + F.off[s] + static_cast<int64_t>(ld) * (jp + c);
   const double* y = z2 + first;
   const double* xb = cpool + F.rel_off[s] - k;
   double acc = 0.0;
   int i = i0 + lane;
   for (; i + 224 < i1; i += 256) {
      double l[8], x[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) l[u] = col[i + 32 * u];
#pragma unroll
      for (int u = 0; u < 8; ++u) x[u] = (i + 32 * u < k) ? y[i + 32 * u] : xb[i + 32 * u];
#pragma unroll
      for (int u = 0; u < 8; ++u) acc = fma(l[u], x[u], acc);
   }
   for (; i + 96 < i1; i += 128) {
      const double l0 = col[i], l1 = col[i + 32], l2 = col[i + 64], l3 = col[i + 96];
      const double x0 = (i < k) ? y[i] : xb[i], x1 = (i + 32 < k) ? y[i + 32] : xb[i + 32];
      const double x2 = (i + 64 < k) ? y[i + 64] : xb[i + 64], x3 = (i + 96 < k) ? y[i + 96] : xb[i + 96];
      acc += l0 * x0 + l1 * x1 + l2 * x2 + l3 * x3;
   }
   for (; i < i1; i += 32) acc += col[i] * ((i < k) ? y[i] : xb[i]);
   for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
   if (lane == 0) partial[(static_cast<int64_t>(blockIdx.x) * BW_SPLITS + blockIdx.z) * NBS + c] = acc;
}

// One CTA per front: y_P -= sum of the partial products, then x_P <- L11[P,P]^{-T} y_P, sub-blocks from last to first.
__global__ void __launch_bounds__(256) k_bwd_big_diag(FrontArrays F, const int* __restrict__ list, const int jp, FactorArrays W,
                                                       double* __restrict__ z2, const double* __restrict__ partial) {
   extern __shared__ __align__(16) double sd_smem[];
   double* Ls = sd_smem;
   double* ys = sd_smem + SD_STRIP_DOUBLES + (NBS / NBI) * 2;
   const int s = list[blockIdx.x];
   const int k = F.k[s], f = F.f[s], ld = F.ld[s];
   const int nb = min(NBS, k - jp);
   const double* A = W.lpool + F.off[s];
   double* y = z2 + F.first[s] + jp;
   const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
   const int64_t xoff = F.xoff[s];
   const int nsb = (nb + NBI - 1) / NBI;
   const double* Xg = W.xinv + xoff + static_cast<int64_t>(jp / NBI) * XINV_BLOCK_DOUBLES;
   if (xoff >= 0) {
      // block rows of L left of every inner block: [column][pair of rows], 16-byte pairs XOR-swizzled by the column so that
      // the 128-bit reads of a quarter warp (one column per lane) fall into distinct banks
      for (int b = 1; b < nsb; ++b) {
         const int base = b * NBI, nbb = min(NBI, nb - base);
         double* T = Ls + NBI * NBI * (b * (b - 1) / 2);
         for (int idx = tid; idx < base * (NBI / 2); idx += 256) {
            const int c = idx / (NBI / 2), pr = idx % (NBI / 2);
            const int bytes = (2 * pr + 1 < nbb) ? 16 : ((2 * pr < nbb) ? 8 : 0);
            cp_async16_n(T + c * NBI + 2 * (pr ^ (c & 15)), bytes ? (A + (jp + base + 2 * pr) + static_cast<int64_t>(ld) * (jp + c)) : A, bytes);
         }
      }
      cp_async_commit();
   } else {
      prefetch_diag_block(A, ld, jp, nb);
      load_diag_blocks(Ls, A, ld, jp, nb);
   }
   {
      double v = 0.0;
      if (tid < nb) {
         v = y[tid];
         if (jp + nb < f) {
            const double* pp = partial + static_cast<int64_t>(blockIdx.x) * BW_SPLITS * NBS + tid;
            double sum = 0.0;
#pragma unroll
            for (int q = 0; q < BW_SPLITS; ++q) sum += pp[q * NBS];
            v -= sum;
         }
      }
      ys[tid] = v;
   }
   if (xoff >= 0) {
      // x_b = X_b^T t_b, blocks from last to first; lane c of warp 0 holds column c of the block's inverse, one block ahead
      // warp w holds the inverse of inner block w (lane = column): all blocks requested at once, block sb applied by warp sb
      double xc[NBI];
      if (warp < nsb) {
         const double* xp = Xg + static_cast<int64_t>(warp) * XINV_BLOCK_DOUBLES + lane;
#pragma unroll
         for (int r = 0; r < NBI; ++r) xc[r] = xp[r * NBI];
      }
      cp_async_wait<0>();
      __syncthreads();
      for (int sb = nsb - 1; sb >= 0; --sb) {
         const int base = sb * NBI;
         if (warp == sb) {
            double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
#pragma unroll
            for (int r = 0; r < NBI; r += 4) {
               a0 = fma(xc[r], ys[base + r], a0);
               a1 = fma(xc[r + 1], ys[base + r + 1], a1);
               a2 = fma(xc[r + 2], ys[base + r + 2], a2);
               a3 = fma(xc[r + 3], ys[base + r + 3], a3);
            }
            __syncwarp();
            ys[base + lane] = (a0 + a1) + (a2 + a3);
         }
         __syncthreads();
         if (sb > 0) {
            if (tid < base) {
               const double* T = Ls + NBI * NBI * (sb * (sb - 1) / 2) + tid * NBI;
               double a0 = 0.0, a1 = 0.0;
#pragma unroll
               for (int pr = 0; pr < NBI / 2; ++pr) {
                  const double2 l2 = *reinterpret_cast<const double2*>(T + 2 * (pr ^ (tid & 15)));
                  a0 = fma(l2.x, ys[base + 2 * pr], a0);
                  a1 = fma(l2.y, ys[base + 2 * pr + 1], a1);
               }
               ys[tid] -= a0 + a1;
            }
            __syncthreads();
         }
      }
      if (tid < nb) y[tid] = ys[tid];
      return;
   }
   __syncthreads();
   for (int sb = nsb - 1; sb >= 0; --sb) {
      const int base = sb * NBI;
      const int nbb = min(NBI, nb - base);
      if (warp == 0) {
         const double* Lb = Ls + base * (NBI + 1);
         double yv = (lane < nbb) ? ys[base + lane] : 0.0;
         for (int c = nbb - 1; c >= 0; --c) {
            const double yc = __shfl_sync(0xffffffffu, yv, c);
            if (lane < c) yv -= Lb[c * (NBI + 1) + lane] * yc;
         }
         if (lane < nbb) ys[base + lane] = yv;
      }
      __syncthreads();
      // earlier columns of the panel: y(c) -= sum_t L[base+t, c] x(base+t)
      if (tid < base) {
         const double* col = A + (jp + base) + static_cast<int64_t>(ld) * (jp + tid);
         double v[NBI];
#pragma unroll
         for (int t = 0; t < NBI; ++t) v[t] = (t < nbb) ? col[t] : 0.0;
         double acc = 0.0;
#pragma unroll
         for (int t = 0; t < NBI; ++t) acc += v[t] * ys[base + t];
         ys[tid] -= acc;
      }
      __syncthreads();
   }
   if (tid < nb) y[tid] = ys[tid];
}

__global__ void __launch_bounds__(256) k_bwd_big_scatter(FrontArrays F, const int* __restrict__ list, FactorArrays W,
                                                          const double* __restrict__ z2, double* __restrict__ xg) {
   const int s = list[blockIdx.x];
   const int j = blockIdx.y * 256 + threadIdx.x;
   if (j >= F.k[s]) return;
   const int first = F.first[s];
   xg[first + W.lperm[first + j]] = z2[first + j];
}

inline int grid1d(int64_t n, int threads) { return static_cast<int>((n + threads - 1) / threads); }

}  // namespace

// ------------------------------------------------------------------------------------------------
// launch wrappers
// ------------------------------------------------------------------------------------------------
int launch_fill(double* p, int64_t n, double v, cudaStream_t st) {
   if (n <= 0) return 0;
   k_fill<<<grid1d(n, 256), 256, 0, st>>>(p, n, v);
   return 1;
}

int launch_scale_sweep(const CsrArrays& A, const double* values, const double* sc_in, double* sc_out, cudaStream_t st) {
   k_scale_sweep<<<grid1d(A.n, 256), 256, 0, st>>>(A, values, sc_in, sc_out);
   return 1;
}

int launch_matrix_inf_norm(const CsrArrays& A, const double* values, unsigned long long* out_bits, cudaStream_t st) {
   k_matrix_inf_norm<<<grid1d(A.n, 256), 256, 0, st>>>(A, values, out_bits);
   return 1;
}

int launch_assemble(const AssemblyArrays& M, const double* values, const double* scale, double* pool, cudaStream_t st) {
   if (M.ndst <= 0) return 0;
   k_assemble<<<grid1d(M.ndst, 256), 256, 0, st>>>(M, values, scale, pool);
   return 1;
}

// number of SMs of the current device (148 on a B200), asked once per device
static int sm_count() {
   static int cached[64] = {0};
   int device = 0;
   if (cudaGetDevice(&device) != cudaSuccess || device < 0 || device >= 64) return 148;
   if (cached[device] == 0) {
      int n = 0;
      cached[device] = (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device) == cudaSuccess && n > 0) ? n : 148;
   }
   return cached[device];
}

int launch_assemble_fill(const AssemblyArrays& M, const double* values, const double* scale, double* pool, cudaStream_t st) {
   if (M.ntiles <= 0) return 0;
   const int64_t resident = static_cast<int64_t>(sm_count()) * 8;   // persistent: 8 CTAs of 256 threads per SM
   const int64_t want = M.ntiles < resident ? M.ntiles : resident;
   k_assemble_fill<<<static_cast<unsigned>(want), 256, 0, st>>>(M, values, scale, pool);
   return 1;
}

int launch_prepare(const FrontArrays& F, const int* d_list, int nfronts, int max_f, const FactorArrays& W, cudaStream_t st) {
   if (nfronts <= 0) return 0;
   dim3 grid(nfronts, (max_f + EA_CBW - 1) / EA_CBW);
   k_prepare<<<grid, 256, 0, st>>>(F, d_list, W);
   return 1;
}

int launch_diagpanel(const FrontArrays& F, const int* d_list, int nfronts, int max_k, int J0, const FactorArrays& W, double null_tol,
                     cudaStream_t st) {
   if (nfronts <= 0) return 0;
   if (max_k <= NBI) k_diag32<<<nfronts, 64, 0, st>>>(F, d_list, W, null_tol);
   else k_diagpanel<<<nfronts, DP_THREADS, DP_SMEM, st>>>(F, d_list, J0, W, null_tol);
   return 1;
}

int launch_trsmpanel(const FrontArrays& F, const int* d_list, int nfronts, int max_f, int max_k, int J0, const FactorArrays& W,
                     cudaStream_t st) {
   if (nfronts <= 0) return 0;
   // a front of the list has k > J0: its rows below the diagonal block start at J0 + 1 at the earliest
   const int rows = max_f - (J0 + 1);
   if (rows < 1) return 0;
   if (max_k <= NBI) {
      dim3 grid(nfronts, (rows + TP_THREADS - 1) / TP_THREADS);
      k_trsmpanel<<<grid, TP_THREADS, TP_SMEM_BASE, st>>>(F, d_list, J0, W);
   } else {
      dim3 grid(nfronts, (rows + TD_ROWS - 1) / TD_ROWS);
      k_trsmpanel_mma<<<grid, TD_THREADS, TD_SMEM, st>>>(F, d_list, J0, W);
   }
   return 1;
}

int leaf_front_max_boundary() { return LEAF_RMAX; }

int launch_leaf_fronts(const FrontArrays& F, const int* d_list, int nfronts, int max_r, const FactorArrays& W, double null_tol, bool accumulate,
                       cudaStream_t st) {
   if (nfronts <= 0) return 0;
   const int rmax = std::max(1, std::min(max_r, LEAF_RMAX));
   const size_t smem = static_cast<size_t>(2 * NBI * (rmax + 1)) * sizeof(double);
   k_leaf_front<<<nfronts, 64, smem, st>>>(F, d_list, W, null_tol, accumulate ? 1 : 0, rmax);
   return 1;
}

int launch_permute_rows(const FrontArrays& F, const int* d_list, int nfronts, int max_k, const FactorArrays& W, cudaStream_t st) {
   if (nfronts <= 0 || max_k <= NBO) return 0;
   const int nblocks = (max_k + NBI - 1) / NBI - 1;   // inner blocks 1 .. last
   dim3 grid(nfronts, nblocks, (max_k + 63) / 64);
   k_permute_rows<<<grid, 256, 0, st>>>(F, d_list, W);
   return 1;
}

namespace {
constexpr int smem_bytes(int bm, int bn, int bk, int stages = 3) { return stages * bk * (bm + 4 + bn + 4) * static_cast<int>(sizeof(double)); }
}

// micro-benchmark: one warp factors the same random symmetric 32 x 32 block `reps` times; cycles per factorization
__global__ void k_bench_bk(const double* __restrict__ M, int reps, long long* cycles, double* sink) {
   __shared__ BkShared S;
   const int lane = threadIdx.x;
   long long total = 0;
   double acc = 0.0;
   for (int r = 0; r < reps; ++r) {
      for (int c = 0; c < NBI; ++c) S.B[lane][c] = M[lane * NBI + c];
      __syncwarp();
      const long long t0 = clock64();
      bk_factor_block(S, NBI, lane, 1e-14);
      total += clock64() - t0;
      acc += S.d[lane] + S.B[lane][(lane + 5) & 31];
      __syncwarp();
   }
   sink[lane] = acc;
   if (lane == 0) *cycles = total / reps;
}

int debug_bench_bk(int variant, int reps, long long* cycles_out, int* n2x2) {
   double h[NBI * NBI];
   unsigned long long sd = 1234567ull + 977ull * static_cast<unsigned long long>(variant);
   auto rnd = [&] { sd ^= sd << 13; sd ^= sd >> 7; sd ^= sd << 17; return (static_cast<double>(sd >> 11) / 9007199254740992.0) - 0.5; };
   for (int i = 0; i < NBI; ++i)
      for (int j = 0; j <= i; ++j) {
         double v = rnd();
         if (variant == 0 && i == j) v += 20.0;                    // diagonally dominant: 1x1 pivots, no interchanges
         if (variant == 1 && i == j) v = (i & 1) ? 1e-9 * v : v;   // KKT-like: every second diagonal tiny
         h[i * NBI + j] = v; h[j * NBI + i] = v;
      }
   double *dM, *dsink; long long* dc;
   cudaMalloc(reinterpret_cast<void**>(&dM), sizeof(h)); cudaMalloc(reinterpret_cast<void**>(&dsink), 32 * sizeof(double));
   cudaMalloc(reinterpret_cast<void**>(&dc), sizeof(long long));
   cudaMemcpy(dM, h, sizeof(h), cudaMemcpyHostToDevice);
   k_bench_bk<<<1, 32>>>(dM, reps, dc, dsink);
   k_bench_bk<<<1, 32>>>(dM, reps, dc, dsink);
   cudaMemcpy(cycles_out, dc, sizeof(long long), cudaMemcpyDeviceToHost);
   if (n2x2 != nullptr) *n2x2 = 0;
   cudaFree(dM); cudaFree(dsink); cudaFree(dc);
   return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

int debug_phase_cycles(unsigned long long* out16, bool reset) {
#ifdef B200KKT_PHASES
   if (out16 != nullptr) cudaMemcpyFromSymbol(out16, g_phase, sizeof(g_phase));
   if (reset) { unsigned long long z[16] = {0}; cudaMemcpyToSymbol(g_phase, z, sizeof(z)); }
   return 1;
#else
   (void)out16; (void)reset;
   return 0;
#endif
}

int init_kernel_attributes() {
   cudaError_t e = cudaFuncSetAttribute(k_diagpanel, cudaFuncAttributeMaxDynamicSharedMemorySize, DP_SMEM);
   if (e != cudaSuccess) return static_cast<int>(e);
   e = cudaFuncSetAttribute(k_fwd_big_diag, cudaFuncAttributeMaxDynamicSharedMemorySize, SD_SMEM);
   if (e != cudaSuccess) return static_cast<int>(e);
   e = cudaFuncSetAttribute(k_bwd_big_diag, cudaFuncAttributeMaxDynamicSharedMemorySize, SD_SMEM);
   if (e != cudaSuccess) return static_cast<int>(e);
   e = cudaFuncSetAttribute(k_trsmpanel, cudaFuncAttributeMaxDynamicSharedMemorySize, TP_SMEM);
   if (e != cudaSuccess) return static_cast<int>(e);
   e = cudaFuncSetAttribute(k_trsmpanel_mma, cudaFuncAttributeMaxDynamicSharedMemorySize, TD_SMEM);
   if (e != cudaSuccess) return static_cast<int>(e);
   e = cudaFuncSetAttribute(k_update<128, 128, 2, 4, 32>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes(128, 128, 32));
   if (e != cudaSuccess) return static_cast<int>(e);
   e = cudaFuncSetAttribute(k_update<64, 64, 2, 2, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes(64, 64, 16));
   if (e != cudaSuccess) return static_cast<int>(e);
   e = cudaFuncSetAttribute(k_update<128, 64, 2, 2, 16, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes(128, 64, 16));
   if (e != cudaSuccess) return static_cast<int>(e);
   e = cudaFuncSetAttribute(k_update<128, 64, 4, 2, 16, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes(128, 64, 16));
   if (e != cudaSuccess) return static_cast<int>(e);
   e = cudaFuncSetAttribute(k_update<128, 128, 4, 4, 32, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes(128, 128, 32));
   if (e != cudaSuccess) return static_cast<int>(e);
   e = cudaFuncSetAttribute(k_update<128, 64, 4, 2, 32, 2, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes(128, 64, 32, 2));
   if (e != cudaSuccess) return static_cast<int>(e);
   e = cudaFuncSetAttribute(k_update_ws, cudaFuncAttributeMaxDynamicSharedMemorySize, WS_SMEM);
   return static_cast<int>(e);
}

int launch_update(const FrontArrays& F, const int* d_list, int nfronts, int max_f, int mode, int j0, int J0, const FactorArrays& W,
                  int tile_mode, cudaStream_t st, const UpdateRange& R) {
   if (nfronts <= 0) return 0;
   // region extent upper bounds (per-front extents are resolved inside the kernel)
   const int r0 = (mode == 0) ? j0 + 1 : J0 + 1;   // smallest possible region start (a front with k = r0)
   const int rows = max_f - (r0 & ~1);
   if (rows <= 0) return 0;
   // columns: a later part starts clo panels in; the first part may have to cover a whole trailing matrix (last-panel fronts)
   int cols = (mode == 0) ? NBO : ((R.policy == 1 || R.chi < 0) ? rows - R.clo * NBO : std::min(rows - R.clo * NBO, (R.chi - R.clo) * NBO + 2));
   if (mode == 1 && R.abs_hi > 0) cols = std::min(cols, R.abs_hi - std::max(R.abs_lo, r0 & ~1) + 2);
   if (mode == 1 && R.cyc_cols > 0) cols = R.cyc_nchunks * R.cyc_cols;
   if (cols <= 0) return 0;
   if (tile_mode == 6) {
      dim3 grid(nfronts, (rows + 127) / 128, (cols + 63) / 64);
      k_update<128, 64, 4, 2, 32, 2, 2><<<grid, 256, smem_bytes(128, 64, 32, 2), st>>>(F, d_list, mode, j0, J0, W, R);
   } else if (tile_mode == 4) {
      dim3 grid(nfronts, (rows + 127) / 128, (cols + 63) / 64);
      k_update<128, 64, 4, 2, 16, 2><<<grid, 256, smem_bytes(128, 64, 16), st>>>(F, d_list, mode, j0, J0, W, R);
   } else if (tile_mode == 5) {
      dim3 grid(nfronts, (rows + 127) / 128, (cols + 127) / 128);
      k_update<128, 128, 4, 4, 32, 1><<<grid, 512, smem_bytes(128, 128, 32), st>>>(F, d_list, mode, j0, J0, W, R);
   } else if (tile_mode == 3 && mode == 1 && R.kpanels == 1 && R.clo == 0 && R.chi < 0) {
      k_update_ws<<<2 * sm_count(), WS_THREADS, WS_SMEM, st>>>(F, d_list, nfronts, J0, W);   // persistent, two CTAs per SM
   } else if (tile_mode == 2 || tile_mode == 3) {
      // two co-resident CTAs per SM: the read-modify-write epilogue of one tile overlaps the K loop of the other
      dim3 grid(nfronts, (rows + 127) / 128, (cols + 63) / 64);
      k_update<128, 64, 2, 2, 16, 2><<<grid, 128, smem_bytes(128, 64, 16), st>>>(F, d_list, mode, j0, J0, W, R);
   } else if (tile_mode == 1) {
      dim3 grid(nfronts, (rows + 127) / 128, (cols + 127) / 128);
      k_update<128, 128, 2, 4, 32><<<grid, 256, smem_bytes(128, 128, 32), st>>>(F, d_list, mode, j0, J0, W, R);
   } else {
      dim3 grid(nfronts, (rows + 63) / 64, (cols + 63) / 64);
      k_update<64, 64, 2, 2, 16><<<grid, 128, smem_bytes(64, 64, 16), st>>>(F, d_list, mode, j0, J0, W, R);
   }
   return 1;
}

int launch_inertia(int64_t n, const FactorArrays& W, int* counters, cudaStream_t st, const signed char* col_owner, int rank) {
   int blocks = grid1d(n, 256);
   if (blocks > 8 * sm_count()) blocks = 8 * sm_count();
   k_inertia<<<blocks, 256, 0, st>>>(n, W, counters, col_owner, rank);
   return 1;
}

// threads: 256, or 64 for a level of small fronts (one CTA per front: four times as many fronts in flight per SM)
int launch_forward_level(const FrontArrays& F, const int* d_list, int nfronts, const FactorArrays& W, double* zg, double* z2, double* cpool,
                         cudaStream_t st, int threads) {
   if (nfronts <= 0) return 0;
   k_forward<<<nfronts, threads, 0, st>>>(F, d_list, W, zg, z2, cpool);
   return 1;
}

int launch_backward_level(const FrontArrays& F, const int* d_list, int nfronts, const FactorArrays& W, double* z2, double* xg, double* cpool,
                          cudaStream_t st, int threads) {
   if (nfronts <= 0) return 0;
   k_backward<<<nfronts, threads, 0, st>>>(F, d_list, W, z2, xg, cpool);
   return 1;
}

int launch_fwd_big_gather(const FrontArrays& F, const int* d_list, int nfronts, int max_f, const FactorArrays&, double* zg, double*,
                          double* cpool, cudaStream_t st) {
   if (nfronts <= 0) return 0;
   dim3 grid(nfronts, (max_f + GATHER_SPAN - 1) / GATHER_SPAN);
   k_fwd_big_gather<<<grid, 256, 0, st>>>(F, d_list, zg, cpool);
   return 1;
}

int launch_fwd_big_gather_list(const FrontArrays& F, const GatherArrays& Gt, const int* d_list, int nfronts, int max_f, double* zg, double* cpool,
                               cudaStream_t st) {
   if (nfronts <= 0) return 0;
   dim3 grid(nfronts, (max_f + 255) / 256);
   k_fwd_big_gather_list<<<grid, 256, 0, st>>>(F, Gt, d_list, zg, cpool);
   return 1;
}

int launch_fwd_big_permute(const FrontArrays& F, const int* d_list, int nfronts, int max_k, const FactorArrays& W, const double* zg,
                           double* z2, cudaStream_t st) {
   if (nfronts <= 0) return 0;
   dim3 grid(nfronts, (max_k + 255) / 256);
   k_fwd_big_permute<<<grid, 256, 0, st>>>(F, d_list, W, zg, z2);
   return 1;
}

int launch_fwd_big_diag(const FrontArrays& F, const int* d_list, int nfronts, int jp, const FactorArrays& W, double* z2, cudaStream_t st) {
   if (nfronts <= 0) return 0;
   k_fwd_big_diag<<<nfronts, 256, SD_SMEM, st>>>(F, d_list, jp, W, z2);
   return 1;
}

int launch_fwd_big_gemv(const FrontArrays& F, const int* d_list, int nfronts, int max_f, int jp, const FactorArrays& W, double* z2,
                        double* cpool, cudaStream_t st) {
   if (nfronts <= 0) return 0;
   const int chunks = (max_f - (jp + 1) + 31) / 32;
   if (chunks < 1) return 0;
   dim3 grid(nfronts, chunks);
   k_fwd_big_gemv<<<grid, 256, 0, st>>>(F, d_list, jp, W, z2, cpool);
   return 1;
}

int launch_bwd_big_begin(const FrontArrays& F, const int* d_list, int nfronts, int max_f, const FactorArrays& W, double* z2,
                         const double* xg, double* cpool, cudaStream_t st) {
   if (nfronts <= 0) return 0;
   dim3 grid(nfronts, (max_f + 255) / 256);
   k_bwd_big_begin<<<grid, 256, 0, st>>>(F, d_list, W, z2, xg, cpool);
   return 1;
}

int launch_bwd_big_gemv(const FrontArrays& F, const int* d_list, int nfronts, int jp, const FactorArrays& W, double* z2,
                        const double* cpool, double* partial, cudaStream_t st) {
   if (nfronts <= 0) return 0;
   dim3 grid(nfronts, NBS / 8, BW_SPLITS);
   k_bwd_big_gemv<<<grid, 256, 0, st>>>(F, d_list, jp, W, z2, cpool, partial);
   return 1;
}

int launch_bwd_big_diag(const FrontArrays& F, const int* d_list, int nfronts, int jp, const FactorArrays& W, double* z2,
                        const double* partial, cudaStream_t st) {
   if (nfronts <= 0) return 0;
   k_bwd_big_diag<<<nfronts, 256, SD_SMEM, st>>>(F, d_list, jp, W, z2, partial);
   return 1;
}

int bwd_partial_doubles_per_front() { return BW_SPLITS * NBS; }

int launch_bwd_big_scatter(const FrontArrays& F, const int* d_list, int nfronts, int max_k, const FactorArrays& W, const double* z2,
                           double* xg, cudaStream_t st) {
   if (nfronts <= 0) return 0;
   dim3 grid(nfronts, (max_k + 255) / 256);
   k_bwd_big_scatter<<<grid, 256, 0, st>>>(F, d_list, W, z2, xg);
   return 1;
}

int launch_solve_gather(int64_t n, const int* perm, const double* scale, const double* r, double* zg, cudaStream_t st) {
   k_solve_gather<<<grid1d(n, 256), 256, 0, st>>>(n, perm, scale, r, zg);
   return 1;
}

int launch_solve_scatter_add(int64_t n, const int* perm, const double* scale, const double* xg, double* x, cudaStream_t st) {
   k_solve_scatter_add<<<grid1d(n, 256), 256, 0, st>>>(n, perm, scale, xg, x);
   return 1;
}

int launch_boundary_pack(const FrontArrays& F, const int* d_list, int nfronts, const double* xg, double* cpool, cudaStream_t st) {
   if (nfronts <= 0) return 0;
   k_boundary_pack<<<nfronts, 256, 0, st>>>(F, d_list, xg, cpool);
   return 1;
}

int launch_boundary_unpack(const FrontArrays& F, const int* d_list, int nfronts, const double* cpool, double* xg, cudaStream_t st) {
   if (nfronts <= 0) return 0;
   k_boundary_unpack<<<nfronts, 256, 0, st>>>(F, d_list, cpool, xg);
   return 1;
}

int launch_rows_pack(const double* panel, int64_t ld, int row0, int nrows, int ncols, double* staging, cudaStream_t st) {
   if (nrows <= 0 || ncols <= 0) return 0;
   k_rows_pack<<<dim3((nrows + 255) / 256, ncols), 256, 0, st>>>(panel, ld, row0, nrows, staging);
   return 1;
}

int launch_rows_unpack(const double* staging, int64_t ld, int row0, int nrows, int ncols, double* panel, cudaStream_t st) {
   if (nrows <= 0 || ncols <= 0) return 0;
   k_rows_unpack<<<dim3((nrows + 255) / 256, ncols), 256, 0, st>>>(staging, ld, row0, nrows, panel);
   return 1;
}

int launch_keep_owned(int64_t n, const signed char* col_owner, int rank, double* xg, cudaStream_t st) {
   k_keep_owned<<<grid1d(n, 256), 256, 0, st>>>(n, col_owner, rank, xg);
   return 1;
}

int launch_residual(const CsrArrays& A, const double* values, const double* x, const double* b, double* r,
                    const unsigned long long* xnorm_bits, unsigned long long* omega_bits, cudaStream_t st) {
   k_residual<<<grid1d(A.n, 256), 256, 0, st>>>(A, values, x, b, r, xnorm_bits, omega_bits);
   return 1;
}

// ------------------------------------------------------------------------------------------------
// Allocation-free backward error (b200kkt_backward_error_noalloc).  The residual vector and the norm accumulators that
// b200kkt_solve keeps in cudaMalloc'ed handle buffers live here in a file-scope device array instead: NOALLOC_HEAD doubles of
// accumulators (five 64-bit words used), then up to NOALLOC_MAX_N doubles of residual.  Every GPU that loads the library holds
// these 8 MiB + 256 B until exit.  All calls share the array and it keeps the last call's contents: the accumulators are
// cleared on the stream before use; calls that overlap on the GPU (two handles, two streams) race.
// In host code the array's name is only a stand-in: its device address comes from cudaGetSymbolAddress, once per device.
// ------------------------------------------------------------------------------------------------
constexpr int64_t NOALLOC_MAX_N = int64_t(1) << 20;
constexpr int NOALLOC_HEAD = 32;
static __device__ __align__(256) double g_noalloc_scratch[NOALLOC_MAX_N + NOALLOC_HEAD];

int64_t noalloc_max_n() { return NOALLOC_MAX_N; }

int launch_backward_error_noalloc(int device, const CsrArrays& A, const double* values, const double* x, const double* b,
                                  unsigned long long out_bits[5], int* launches, cudaStream_t st) {
   constexpr int MAX_DEVICES = 64;
   static double* scratch_of[MAX_DEVICES] = {};
   *launches = 0;
   if (device < 0 || device >= MAX_DEVICES) return static_cast<int>(cudaErrorInvalidDevice);
   if (A.n > NOALLOC_MAX_N) return static_cast<int>(cudaErrorInvalidValue);
   if (scratch_of[device] == nullptr) {
      // set-up, once per device, with that device current (the caller has made it so)
      double* p = nullptr;
      const cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_scratch);
      if (e != cudaSuccess) return static_cast<int>(e);
      scratch_of[device] = p;
   }
   double* scratch = scratch_of[device];
   unsigned long long* acc = reinterpret_cast<unsigned long long*>(scratch);   // |K|_inf, |r|_inf, |x|_inf, |b|_inf, omega
   double* r = scratch + NOALLOC_HEAD;
   cudaError_t e = cudaMemsetAsync(acc, 0, NOALLOC_HEAD * sizeof(double), st);
   if (e != cudaSuccess) return static_cast<int>(e);
   const int blocks = grid1d(A.n, 256), rblocks = blocks > 8 * sm_count() ? 8 * sm_count() : (blocks < 1 ? 1 : blocks);
   k_matrix_inf_norm<<<blocks, 256, 0, st>>>(A, values, acc + 0);
   k_absmax<<<rblocks, 256, 0, st>>>(x, A.n, acc + 2);
   k_residual<<<blocks, 256, 0, st>>>(A, values, x, b, r, acc + 2, acc + 4);
   k_absmax<<<rblocks, 256, 0, st>>>(r, A.n, acc + 1);
   k_absmax<<<rblocks, 256, 0, st>>>(b, A.n, acc + 3);
   *launches = 5;
   e = cudaMemcpyAsync(out_bits, acc, 5 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st);
   if (e != cudaSuccess) return static_cast<int>(e);
   e = cudaStreamSynchronize(st);
   if (e != cudaSuccess) return static_cast<int>(e);
   return static_cast<int>(cudaGetLastError());
}

int launch_absmax(const double* v, int64_t n, unsigned long long* out_bits, cudaStream_t st) {
   int blocks = grid1d(n, 256);
   if (blocks > 8 * sm_count()) blocks = 8 * sm_count();
   if (blocks < 1) blocks = 1;
   k_absmax<<<blocks, 256, 0, st>>>(v, n, out_bits);
   return 1;
}

}  // namespace b200kkt
