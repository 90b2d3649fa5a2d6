// Latency microbenchmark (one warp): dependent DFMA chain, DP division, REDUX, SHFL, LDS round trip.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(double* out, long long* cyc, double a, double b, int iters) {
   __shared__ double sm[1024];
   const int lane = threadIdx.x;
   sm[lane] = a; sm[lane + 32] = b;
   __syncwarp();
   double x = a + lane;
   long long t0 = clock64();
   for (int i = 0; i < iters; ++i) x = fma(x, a, b);
   long long t1 = clock64();
   double y = a + lane;
   for (int i = 0; i < iters; ++i) y = b / y + a;
   long long t2 = clock64();
   unsigned u = lane;
   for (int i = 0; i < iters; ++i) u = __reduce_max_sync(0xffffffffu, u + i);
   long long t3 = clock64();
   double z = x;
   for (int i = 0; i < iters; ++i) z = __shfl_xor_sync(0xffffffffu, z, 1) + 1.0;
   long long t4 = clock64();
   double w = y; int idx = lane;
   for (int i = 0; i < iters; ++i) { sm[idx] = w; __syncwarp(); w = sm[(idx + 1) & 31] + 1.0; __syncwarp(); }
   long long t5 = clock64();
   double r = 1.0 + lane;
   for (int i = 0; i < iters; ++i) r = __drcp_rn(r) + a;
   long long t6 = clock64();
   // independent DFMAs, 8 chains
   double c0 = a, c1 = b, c2 = a + 1, c3 = b + 1, c4 = a + 2, c5 = b + 2, c6 = a + 3, c7 = b + 3;
   for (int i = 0; i < iters; ++i) { c0 = fma(c0, a, b); c1 = fma(c1, a, b); c2 = fma(c2, a, b); c3 = fma(c3, a, b); c4 = fma(c4, a, b); c5 = fma(c5, a, b); c6 = fma(c6, a, b); c7 = fma(c7, a, b); }
   long long t7 = clock64();
   out[lane] = x + y + u + z + w + r + c0 + c1 + c2 + c3 + c4 + c5 + c6 + c7;
   if (lane == 0) { cyc[0] = t1 - t0; cyc[1] = t2 - t1; cyc[2] = t3 - t2; cyc[3] = t4 - t3; cyc[4] = t5 - t4; cyc[5] = t6 - t5; cyc[6] = t7 - t6; }
}
int main() {
   double* out; long long* cyc;
   cudaMalloc(&out, 32 * 8); cudaMallocManaged(&cyc, 8 * 8);
   const int iters = 2000;
   k<<<1, 32>>>(out, cyc, 1.0000001, 0.9999999, iters);
   k<<<1, 32>>>(out, cyc, 1.0000001, 0.9999999, iters);
   cudaDeviceSynchronize();
   const char* names[] = {"dependent DFMA", "DP division (+DADD)", "REDUX.MAX u32 (+IADD)", "SHFL.64 (+DADD)", "STS+LDS round trip (+DADD, 2 syncwarp)", "__drcp_rn (+DADD)", "8 independent DFMA"};
   for (int i = 0; i < 7; ++i) printf("%-42s %8.1f cycles / iteration\n", names[i], double(cyc[i]) / iters);
   return 0;
}
