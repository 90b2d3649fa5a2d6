// Microbenchmark (run on the GPU box): register-resident DMMA.8x8x4 and DFMA issue-rate peaks and cuBLAS DGEMM,
// to fix the FP64 roofline denominator that MEASURED_PEAKS.json does not carry.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_peak fp64_peak.cu -lcublas
#include <cstdio>
#include <cublas_v2.h>
#include <cuda_runtime.h>

template <int ILP>
__global__ void __launch_bounds__(256) k_dmma(double* out, int iters, double a, double b) {
   double c[ILP][2];
#pragma unroll
   for (int i = 0; i < ILP; ++i) { c[i][0] = threadIdx.x; c[i][1] = i; }
   for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int i = 0; i < ILP; ++i)
         asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
   }
   double s = 0;
#pragma unroll
   for (int i = 0; i < ILP; ++i) s += c[i][0] + c[i][1];
   out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int ILP>
__global__ void __launch_bounds__(256) k_dfma(double* out, int iters, double a, double b) {
   double c[ILP];
#pragma unroll
   for (int i = 0; i < ILP; ++i) c[i] = threadIdx.x + i;
   for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int i = 0; i < ILP; ++i) c[i] = fma(c[i], a, b);
   }
   double s = 0;
#pragma unroll
   for (int i = 0; i < ILP; ++i) s += c[i];
   out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename F>
float time_ms(F&& f, int reps) {
   cudaEvent_t e0, e1;
   cudaEventCreate(&e0); cudaEventCreate(&e1);
   f();
   cudaDeviceSynchronize();
   float best = 1e30f;
   for (int r = 0; r < reps; ++r) {
      cudaEventRecord(e0);
      f();
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      if (ms < best) best = ms;
   }
   return best;
}

int main() {
   cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
   const int sms = p.multiProcessorCount;
   double* out; cudaMalloc(&out, sizeof(double) * sms * 8 * 256);
   const int iters = 20000;
   for (int cps = 1; cps <= 8; cps *= 2) {
      float ms = time_ms([&] { k_dmma<8><<<sms * cps, 256>>>(out, iters, 1.0000001, 0.9999999); }, 5);
      double flops = double(sms) * cps * 8 /*warps*/ * 8 /*ilp*/ * iters * 512.0;
      printf("DMMA.8x8x4  ILP8  %d CTA/SM x 256 thr: %.2f TFLOP/s\n", cps, flops / ms / 1e9);
   }
   for (int cps = 1; cps <= 8; cps *= 2) {
      float ms = time_ms([&] { k_dfma<8><<<sms * cps, 256>>>(out, iters, 1.0000001, 0.9999999); }, 5);
      double flops = double(sms) * cps * 256 * 8 * double(iters) * 2.0;
      printf("DFMA        ILP8  %d CTA/SM x 256 thr: %.2f TFLOP/s\n", cps, flops / ms / 1e9);
   }
   cublasHandle_t h; cublasCreate(&h);
   for (int n : {4096, 8192, 16384}) {
      double *A, *B, *C;
      cudaMalloc(&A, sizeof(double) * n * n); cudaMalloc(&B, sizeof(double) * n * n); cudaMalloc(&C, sizeof(double) * n * n);
      cudaMemset(A, 0, sizeof(double) * n * n); cudaMemset(B, 0, sizeof(double) * n * n);
      const double one = 1.0, zero = 0.0;
      float ms = time_ms([&] { cublasDgemm(h, CUBLAS_OP_N, CUBLAS_OP_T, n, n, n, &one, A, n, B, n, &zero, C, n); }, 6);
      printf("cublasDgemm NT %d^3 best of 6: %.2f TFLOP/s (%.2f ms)\n", n, 2.0 * n * n * n / ms / 1e9, ms);
      // sustained: back to back for ~3 s
      cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
      int reps = (int)(3000.0f / ms) + 1;
      cudaEventRecord(e0);
      for (int r = 0; r < reps; ++r) cublasDgemm(h, CUBLAS_OP_N, CUBLAS_OP_T, n, n, n, &one, A, n, B, n, &zero, C, n);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      float tot; cudaEventElapsedTime(&tot, e0, e1);
      printf("cublasDgemm NT %d^3 sustained %d reps: %.2f TFLOP/s\n", n, reps, 2.0 * n * n * n * reps / tot / 1e9);
      // rank-256 update shape (the K4 trailing update): m = n, k = 256
      ms = time_ms([&] { cublasDgemm(h, CUBLAS_OP_N, CUBLAS_OP_T, n, n, 256, &one, A, n, B, n, &one, C, n); }, 6);
      printf("cublasDgemm NT %dx%dx256 (beta=1) best of 6: %.2f TFLOP/s (%.3f ms)\n", n, n, 2.0 * n * n * 256 / ms / 1e9, ms);
      cudaFree(A); cudaFree(B); cudaFree(C);
   }
   return 0;
}
