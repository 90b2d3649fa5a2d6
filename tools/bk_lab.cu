// Where do the cycles of a one-warp 32x32 LDL^T elimination go?  Variants of the inner step, timed with clock64.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bk_lab tools/bk_lab.cu
#include <cstdio>
#include <cuda_runtime.h>
constexpr int NBI = 32;
struct alignas(16) Sh { double B[NBI][NBI + 2]; double d[NBI]; };

__device__ __forceinline__ double warp_max(const double v) {
   const bool present = v >= 0.0;
   const unsigned hi = present ? static_cast<unsigned>(__double2hiint(v)) + 1u : 0u;
   const unsigned mhi = __reduce_max_sync(0xffffffffu, hi);
   if (mhi == 0u) return -1.0;
   const unsigned lo = (hi == mhi) ? static_cast<unsigned>(__double2loint(v)) : 0u;
   const unsigned mlo = __reduce_max_sync(0xffffffffu, lo);
   return __hiloint2double(static_cast<int>(mhi - 1u), static_cast<int>(mlo));
}

// MODE bits: 1 = pivot search (lambda, first row), 2 = exact division (else multiply by a precomputed reciprocal),
//            4 = mirror stores, 8 = pivot row through 128-bit loads (else 64-bit)
template <int MODE>
__device__ void factor(Sh& S, const int lane) {
   double a[NBI];
#pragma unroll
   for (int c = 0; c < NBI; ++c) a[c] = S.B[c][lane];
   __syncwarp();
   for (int j = 0; j < NBI; ++j) {
      const bool below = lane > j;
      const double vj = S.B[j][lane];
      const double d = S.B[j][j];
      double keep = 1.0;
      if (MODE & 1) {
         const double v = below ? fabs(vj) : -1.0;
         double lam = warp_max(v);
         int r = j;
         if (lam >= 0.0) r = __ffs(__ballot_sync(0xffffffffu, v == lam)) - 1; else lam = 0.0;
         if (!(fmax(fabs(d), lam) > 1e-14)) keep = 0.0;
         else if (fabs(d) < 0.64 * lam) keep = (r == 77) ? 2.0 : 1.0;
      }
      double li;
      if (MODE & 2) li = below ? vj / d : 0.0; else li = below ? vj * __drcp_rn(d) : 0.0;
      li *= keep;
      if (MODE & 8) {
         const double2* prow = reinterpret_cast<const double2*>(S.B[j]);
#pragma unroll
         for (int c = 0; c < NBI; c += 2) { const double2 pc = prow[c >> 1]; a[c] = a[c] - li * pc.x; a[c + 1] = a[c + 1] - li * pc.y; }
      } else {
#pragma unroll
         for (int c = 0; c < NBI; ++c) a[c] = a[c] - li * S.B[j][c];
      }
      __syncwarp();
      if (MODE & 4) {
#pragma unroll
         for (int c = 0; c < NBI; ++c) if (below && c > j) S.B[c][lane] = a[c];
      } else {
         // only the next pivot row and column are published: lane j+1 stores its row, everybody its column entry
         if (j + 1 < NBI) {
            double nxt = a[0];
#pragma unroll
            for (int c = 1; c < NBI; ++c) nxt = (c == j + 1) ? a[c] : nxt;
            if (below) S.B[j + 1][lane] = nxt;
         }
      }
      if (below) S.B[j][lane] = li;
      if (lane == 0) S.d[j] = d;
      __syncwarp();
   }
}

template <int MODE>
__global__ void bench(const double* __restrict__ M, int reps, long long* cycles, double* sink) {
   __shared__ Sh S;
   const int lane = threadIdx.x;
   long long total = 0;
   double acc = 0.0;
   for (int r = 0; r < reps; ++r) {
      for (int c = 0; c < NBI; ++c) S.B[lane][c] = M[lane * NBI + c];
      __syncwarp();
      const long long t0 = clock64();
      factor<MODE>(S, lane);
      total += clock64() - t0;
      acc += S.d[lane] + S.B[lane][(lane + 5) & 31];
      __syncwarp();
   }
   sink[lane] = acc;
   if (lane == 0) *cycles = total / reps;
}

template <int MODE>
void run(const double* dM, long long* dc, double* dsink, const char* what) {
   bench<MODE><<<1, 32>>>(dM, 100, dc, dsink);
   bench<MODE><<<1, 32>>>(dM, 100, dc, dsink);
   long long c = 0; double s[32];
   cudaMemcpy(&c, dc, sizeof(c), cudaMemcpyDeviceToHost);
   cudaMemcpy(s, dsink, sizeof(s), cudaMemcpyDeviceToHost);
   printf("mode %2d  %-70s %7lld cycles/block  %5.0f /column   (check %.6e)\n", MODE, what, c, c / 32.0, s[7]);
}

int main() {
   double h[NBI * NBI];
   unsigned long long sd = 1234567ull;
   auto rnd = [&] { sd ^= sd << 13; sd ^= sd >> 7; sd ^= sd << 17; return (static_cast<double>(sd >> 11) / 9007199254740992.0) - 0.5; };
   for (int i = 0; i < NBI; ++i) for (int j = 0; j <= i; ++j) { double v = rnd(); if (i == j) v += 20.0; h[i * NBI + j] = v; h[j * NBI + i] = v; }
   double *dM, *dsink; long long* dc;
   cudaMalloc(&dM, sizeof(h)); cudaMalloc(&dsink, 32 * sizeof(double)); cudaMalloc(&dc, sizeof(long long));
   cudaMemcpy(dM, h, sizeof(h), cudaMemcpyHostToDevice);
   run<0>(dM, dc, dsink, "elimination only: reciprocal, 64-bit row loads, publish next row only");
   run<8>(dM, dc, dsink, "  + 128-bit pivot row loads");
   run<4>(dM, dc, dsink, "full mirror stores, 64-bit loads");
   run<12>(dM, dc, dsink, "full mirror stores, 128-bit loads");
   run<14>(dM, dc, dsink, "  + exact division");
   run<13>(dM, dc, dsink, "  + pivot search (reciprocal)");
   run<15>(dM, dc, dsink, "  + pivot search + exact division");
   run<9>(dM, dc, dsink, "publish next row only + search");
   printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
   return 0;
}
