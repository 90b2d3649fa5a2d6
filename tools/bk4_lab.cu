// Four-warp Bunch-Kaufman on a 32x32 block: prototype + verification + timing.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bk4_lab tools/bk4_lab.cu
#include <cmath>
#include <cstdio>
#include <cstring>
#include <cuda_runtime.h>
constexpr int NBI = 32;
constexpr double BK_ALPHA = 0.6403882032022076;
constexpr int BK4_THREADS = 128;
constexpr int BK4_BARRIER = 1;

struct alignas(16) BkShared {
   double B[NBI][NBI + 2];
   double Lm[NBI][NBI + 2];
   double d[NBI], e[NBI];
   int p[NBI], type[NBI];
   int nontrivial;
};

__device__ __forceinline__ double warp_max(const double v) {
   const bool present = v >= 0.0;
   const unsigned hi = present ? static_cast<unsigned>(__double2hiint(v)) + 1u : 0u;
   const unsigned mhi = __reduce_max_sync(0xffffffffu, hi);
   if (mhi == 0u) return -1.0;
   const unsigned lo = (hi == mhi) ? static_cast<unsigned>(__double2loint(v)) : 0u;
   const unsigned mlo = __reduce_max_sync(0xffffffffu, lo);
   return __hiloint2double(static_cast<int>(mhi - 1u), static_cast<int>(mlo));
}

__device__ __forceinline__ void group_sync() { asm volatile("bar.sync %0, %1;" ::"n"(BK4_BARRIER), "n"(BK4_THREADS) : "memory"); }

// branch-free reciprocal for normal arguments well inside the exponent range (pivots are > 1e-14 in the scaled matrix)
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

// tid in [0, 128): warp w owns columns [8w, 8w + 8) of every row, lane i owns row i.
template <int PROBE>
__device__ void bk_factor_block4(BkShared& S, const int n, const int tid, const double tol) {
   const int lane = tid & 31, w = tid >> 5, c0 = 8 * w;
   if (w == 0) {
      S.p[lane] = lane; S.d[lane] = 1.0; S.e[lane] = 0.0; S.type[lane] = 0;
      if (lane == 0) S.nontrivial = 0;
   }
   double a[8];
#pragma unroll
   for (int c = 0; c < 8; ++c) a[c] = S.B[c0 + c][lane];
   group_sync();
   int j = 0;
   while (j < n) {
      if (PROBE == 3) { group_sync(); ++j; continue; }
      const bool below = lane > j && lane < n;
      const double vj = S.B[j][lane];
      if (PROBE == 4) { if (vj == 1234.5) S.d[0] = 0.0; group_sync(); ++j; continue; }
      const double djj = S.B[j][j];
      double rd = (PROBE == 2) ? 0.01 : fast_rcp(djj);
      const double absajj = fabs(djj);
      const double v = below ? fabs(vj) : -1.0;
      double lam = (PROBE == 1 || PROBE == 2) ? 0.0 : warp_max(v);
      int r = j;
      if (PROBE == 0) { if (lam >= 0.0) r = __ffs(__ballot_sync(0xffffffffu, v == lam)) - 1; else lam = 0.0; }
      int kp = j, step = 1;
      bool isnull = false, looked = false;
      if (!(fmax(absajj, lam) > tol)) {
         isnull = true;
      } else if (absajj >= BK_ALPHA * lam) {
      } else {
         looked = true;
         const double sigma = warp_max((lane >= j && lane < n && lane != r) ? fabs(S.B[r][lane]) : -1.0);
         if (absajj * sigma >= BK_ALPHA * lam * lam) {
         } else if (fabs(S.B[r][r]) >= BK_ALPHA * sigma) {
            kp = r;
         } else {
            kp = r;
            step = 2;
         }
      }
      const int kk = j + step - 1;
      if (kp != kk) {
         group_sync();   // every warp is done reading the block for the search
         if (w == 0) {
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
         group_sync();
#pragma unroll
         for (int c = 0; c < 8; ++c) a[c] = S.B[c0 + c][lane];
         if (step == 1) rd = fast_rcp(S.B[j][j]);
      } else if (looked) {
         group_sync();   // row r has been read by everybody before its owner overwrites it
      }
      if (step == 1) {
         if (isnull) {
            if (w == 0) {
               S.Lm[j][lane] = 0.0;
               if (lane == 0) { S.d[j] = 1.0; S.e[j] = 0.0; S.type[j] = 3; }
            }
         } else {
            const double d = S.B[j][j];
            const double li = below ? ((PROBE == 2) ? vj * rd : fast_div(S.B[j][lane], d, rd)) : 0.0;
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
            S.Lm[j][lane] = (lane == j + 1) ? 0.0 : l1;
            S.Lm[j + 1][lane] = l2;
            if (lane == 0) {
               S.d[j] = pa; S.d[j + 1] = pc; S.e[j] = pb; S.e[j + 1] = 0.0;
               S.type[j] = 1; S.type[j + 1] = 2;
            }
         }
      }
      group_sync();
      j += step;
   }
   // L11 into the strict lower triangle: S.B[i][c] = L(i, c)
#pragma unroll
   for (int c = 0; c < 8; ++c) if (c0 + c < lane) S.B[lane][c0 + c] = S.Lm[c0 + c][lane];
   group_sync();
}

template <int PROBE>
__global__ void bench(const double* __restrict__ M, int n, int reps, long long* cycles, double* out, int* iout) {
   __shared__ BkShared S;
   const int tid = threadIdx.x, lane = tid & 31;
   long long total = 0;
   for (int r = 0; r < reps; ++r) {
      for (int idx = tid; idx < NBI * NBI; idx += BK4_THREADS) S.B[idx / NBI][idx % NBI] = M[idx];
      __syncthreads();
      const long long t0 = clock64();
      bk_factor_block4<PROBE>(S, n, tid, 1e-14);
      total += clock64() - t0;
      __syncthreads();
   }
   for (int idx = tid; idx < NBI * NBI; idx += BK4_THREADS) out[idx] = S.B[idx / NBI][idx % NBI];
   if (tid < 32) { out[NBI * NBI + lane] = S.d[lane]; out[NBI * NBI + 32 + lane] = S.e[lane]; iout[lane] = S.p[lane]; iout[32 + lane] = S.type[lane]; }
   if (tid == 0) *cycles = total / reps;
}

static double verify(const double* A, const double* out, const int* iout, int n, int* n2, int* nswap, int* nnull) {
   // P A P^T = L D L^T on the leading n x n block
   double L[NBI][NBI] = {}, D[NBI][NBI] = {};
   const double* d = out + NBI * NBI; const double* e = d + 32; const int* p = iout; const int* ty = iout + 32;
   *n2 = 0; *nswap = 0; *nnull = 0;
   for (int i = 0; i < n; ++i) { L[i][i] = 1.0; for (int j = 0; j < i; ++j) L[i][j] = out[i * NBI + j]; if (p[i] != i) ++*nswap; }
   for (int i = 0; i < n; ++i) {
      if (ty[i] == 3) { D[i][i] = 0.0; ++*nnull; continue; }
      D[i][i] = d[i];
      if (ty[i] == 1) { D[i][i + 1] = e[i]; D[i + 1][i] = e[i]; ++*n2; }
   }
   double err = 0.0;
   for (int i = 0; i < n; ++i)
      for (int j = 0; j < n; ++j) {
         double s = 0.0;
         for (int a = 0; a < n; ++a) for (int b = 0; b < n; ++b) s += L[i][a] * D[a][b] * L[j][b];
         err = fmax(err, fabs(s - A[p[i] * NBI + p[j]]));
      }
   return err;
}

int main() {
   double *dM, *dout; long long* dc; int* diout;
   cudaMalloc(&dM, NBI * NBI * 8); cudaMalloc(&dout, (NBI * NBI + 64) * 8); cudaMalloc(&dc, 8); cudaMalloc(&diout, 64 * 4);
   unsigned long long sd = 1234567ull;
   auto rnd = [&] { sd ^= sd << 13; sd ^= sd >> 7; sd ^= sd << 17; return (static_cast<double>(sd >> 11) / 9007199254740992.0) - 0.5; };
   const char* names[] = {"dominant diagonal", "KKT pairs [[h,a],[a,-1e-8]]", "random symmetric", "random, n = 27", "rank deficient (two equal rows)"};
   for (int variant = 0; variant < 5; ++variant) {
      double h[NBI * NBI];
      int n = (variant == 3) ? 27 : NBI;
      for (int i = 0; i < NBI; ++i)
         for (int j = 0; j <= i; ++j) {
            double v = rnd();
            if (variant == 0 && i == j) v += 20.0;
            if (variant == 1 && i == j) v = (i & 1) ? -1e-8 : 1.0 + fabs(v);
            if (variant == 1 && i != j && !(j == i - 1 && (i & 1))) v *= 0.05;
            h[i * NBI + j] = v; h[j * NBI + i] = v;
         }
      if (variant == 4) for (int j = 0; j < NBI; ++j) { h[5 * NBI + j] = h[9 * NBI + j]; h[j * NBI + 5] = h[j * NBI + 9]; h[5 * NBI + 5] = h[9 * NBI + 9]; h[5 * NBI + 9] = h[9 * NBI + 9]; h[9 * NBI + 5] = h[9 * NBI + 9]; }
      if (n < NBI) for (int i = 0; i < NBI; ++i) for (int j = 0; j < NBI; ++j) if (i >= n || j >= n) h[i * NBI + j] = (i == j) ? 1.0 : 0.0;
      cudaMemcpy(dM, h, sizeof(h), cudaMemcpyHostToDevice);
      bench<0><<<1, BK4_THREADS>>>(dM, n, 1, dc, dout, diout);
      double out[NBI * NBI + 64]; int iout[64];
      cudaMemcpy(out, dout, sizeof(out), cudaMemcpyDeviceToHost);
      cudaMemcpy(iout, diout, sizeof(iout), cudaMemcpyDeviceToHost);
      int n2, nswap, nnull;
      const double err = verify(h, out, iout, n, &n2, &nswap, &nnull);
      bench<0><<<1, BK4_THREADS>>>(dM, n, 100, dc, dout, diout);
      long long c = 0;
      cudaMemcpy(&c, dc, sizeof(c), cudaMemcpyDeviceToHost);
      if (variant == 0) {
         long long pc[5] = {c, 0, 0, 0, 0};
         bench<1><<<1, BK4_THREADS>>>(dM, n, 100, dc, dout, diout); cudaMemcpy(&pc[1], dc, 8, cudaMemcpyDeviceToHost);
         bench<2><<<1, BK4_THREADS>>>(dM, n, 100, dc, dout, diout); cudaMemcpy(&pc[2], dc, 8, cudaMemcpyDeviceToHost);
         bench<3><<<1, BK4_THREADS>>>(dM, n, 100, dc, dout, diout); cudaMemcpy(&pc[3], dc, 8, cudaMemcpyDeviceToHost);
         bench<4><<<1, BK4_THREADS>>>(dM, n, 100, dc, dout, diout); cudaMemcpy(&pc[4], dc, 8, cudaMemcpyDeviceToHost);
         printf("probes per column: full %lld, no search %lld, no search no rcp %lld, barrier only %lld, LDS+barrier %lld\n", pc[0] / 32, pc[1] / 32, pc[2] / 32, pc[3] / 32, pc[4] / 32);
      }
      printf("%-34s |PAP'-LDL'|max = %.2e  2x2: %2d  moved: %2d  null: %d   %6lld cycles/block  %4.0f /column\n", names[variant], err, n2, nswap, nnull, c, double(c) / n);
   }
   printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
   return 0;
}
