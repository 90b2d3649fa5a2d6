// How many warps per SM sub-partition does DMMA.8x8x4 need, and what do interleaved shared-memory fragment loads cost?
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/dmma_issue tools/dmma_issue.cu
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void dmma(double& c0, double& c1, const double a, const double b) {
   asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
// MF x NF accumulator tiles per warp; LDS = 1: fragments come from shared memory every k-step (MF + NF loads per MF*NF DMMA)
template <int MF, int NF, int LDS>
__global__ void k(double* out, int iters, double a0, double b0) {
   __shared__ double sm[2048];
   for (int i = threadIdx.x; i < 2048; i += blockDim.x) sm[i] = a0 + 1e-9 * i;
   __syncthreads();
   const int lane = threadIdx.x & 31;
   double c[MF][NF][2];
#pragma unroll
   for (int i = 0; i < MF; ++i)
#pragma unroll
      for (int j = 0; j < NF; ++j) { c[i][j][0] = lane; c[i][j][1] = i + j; }
   double af[MF], bf[NF];
#pragma unroll
   for (int i = 0; i < MF; ++i) af[i] = a0 + i;
#pragma unroll
   for (int j = 0; j < NF; ++j) bf[j] = b0 + j;
   for (int it = 0; it < iters; ++it) {
      if (LDS) {
         const int base = (it & 7) * 132 * (lane & 3) + (lane >> 2);
#pragma unroll
         for (int i = 0; i < MF; ++i) af[i] = sm[(base + 8 * i) & 2047];
#pragma unroll
         for (int j = 0; j < NF; ++j) bf[j] = sm[(base + 8 * j + 64) & 2047];
      }
#pragma unroll
      for (int i = 0; i < MF; ++i)
#pragma unroll
         for (int j = 0; j < NF; ++j) dmma(c[i][j][0], c[i][j][1], af[i], bf[j]);
   }
   double s = 0.0;
#pragma unroll
   for (int i = 0; i < MF; ++i)
#pragma unroll
      for (int j = 0; j < NF; ++j) s += c[i][j][0] + c[i][j][1];
   out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int MF, int NF, int LDS>
void run(double* out, int sms, int threads, const char* what) {
   const int iters = 4000;
   cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
   k<MF, NF, LDS><<<sms, threads>>>(out, iters, 1.0000001, 0.9999999);
   cudaEventRecord(e0);
   k<MF, NF, LDS><<<sms, threads>>>(out, iters, 1.0000001, 0.9999999);
   cudaEventRecord(e1); cudaEventSynchronize(e1);
   float ms; cudaEventElapsedTime(&ms, e0, e1);
   const double flops = 512.0 * MF * NF * iters * (threads / 32) * sms;
   printf("%-44s %3d threads/SM (%d warps/SMSP): %6.2f TFLOP/s\n", what, threads, threads / 128, flops / ms / 1e9);
}
int main() {
   int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
   double* out; cudaMalloc(&out, sizeof(double) * sms * 1024);
   for (int threads : {128, 256, 384, 512, 1024}) {
      run<8, 4, 0>(out, sms, threads, "8x4 tiles, register operands");
      run<8, 4, 1>(out, sms, threads, "8x4 tiles, fragments from shared memory");
      run<4, 4, 1>(out, sms, threads, "4x4 tiles, fragments from shared memory");
      run<4, 2, 1>(out, sms, threads, "4x2 tiles, fragments from shared memory");
   }
   printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
   return 0;
}
