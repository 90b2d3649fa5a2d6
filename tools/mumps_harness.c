/* MUMPS side-by-side harness: the reference's own arithmetic for the path, timed on the host cores.
 *
 * Calls dmumps_c exactly as the reference's MUMPSSolver does - control block of
 * uno/ingredients/subproblem_solvers/MUMPS/MUMPSSolver.cpp:16-37 (SYM = 2, PAR = 1, silent output, ICNTL(6) = 0, ICNTL(13) = 1,
 * ICNTL(24) = 1), ICNTL(8) = 8 after the analysis (:82) - plus ICNTL(7) = 1 with PERM_IN so that MUMPS eliminates in the SAME
 * order as the GPU run (north star: "same inputs and the same METIS ordering").  JOB = 1 once, then JOB = 2 + JOB = 3 timed
 * (best of 3 after one warm-up), inertia from INFOG(12) / INFOG(28) (:124-139).
 *
 * MUMPS is not vendored by the reference and cannot be built in this repository's container (no gfortran, no network):
 * oracle/cpu_arm.py:mumps_probe() compiles this file only where a libdmumps and dmumps_c.h are present.
 *
 * Input file (written by oracle/cpu_arm.py:run_mumps): int64 n, nnz; int32 irn[nnz], jcn[nnz] (1-based, upper triangle);
 * double a[nnz]; double rhs[n]; int32 perm_in[n].  Output: one JSON line on stdout.
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <time.h>

#include "dmumps_c.h"

#define ICNTL(I) icntl[(I) - 1]
#define INFOG(I) infog[(I) - 1]
#define USE_COMM_WORLD (-987654)

static double now(void) {
   struct timespec ts;
   clock_gettime(CLOCK_MONOTONIC, &ts);
   return (double) ts.tv_sec + 1e-9 * (double) ts.tv_nsec;
}

static void* must(void* p) {
   if (p == NULL) { fprintf(stderr, "mumps_harness: out of memory\n"); exit(2); }
   return p;
}

int main(int argc, char** argv) {
   if (argc < 2) { fprintf(stderr, "usage: mumps_harness system.bin\n"); return 2; }
   FILE* f = fopen(argv[1], "rb");
   if (f == NULL) { perror("mumps_harness"); return 2; }
   int64_t head[2];
   if (fread(head, sizeof(int64_t), 2, f) != 2) return 2;
   const int64_t n = head[0], nnz = head[1];
   int* irn = must(malloc(sizeof(int) * (size_t) nnz));
   int* jcn = must(malloc(sizeof(int) * (size_t) nnz));
   double* a = must(malloc(sizeof(double) * (size_t) nnz));
   double* b = must(malloc(sizeof(double) * (size_t) n));
   double* x = must(malloc(sizeof(double) * (size_t) n));
   int* perm_in = must(malloc(sizeof(int) * (size_t) n));
   if (fread(irn, sizeof(int), (size_t) nnz, f) != (size_t) nnz || fread(jcn, sizeof(int), (size_t) nnz, f) != (size_t) nnz ||
       fread(a, sizeof(double), (size_t) nnz, f) != (size_t) nnz || fread(b, sizeof(double), (size_t) n, f) != (size_t) n ||
       fread(perm_in, sizeof(int), (size_t) n, f) != (size_t) n) {
      fprintf(stderr, "mumps_harness: short read\n");
      return 2;
   }
   fclose(f);

   DMUMPS_STRUC_C id;
   id.sym = 2;                       /* MUMPSSolver.cpp:17 general symmetric */
   id.par = 1;                       /* :24 host takes part */
   id.comm_fortran = USE_COMM_WORLD; /* :20 */
   id.job = -1;                      /* :25-26 */
   dmumps_c(&id);
   id.ICNTL(1) = -1; id.ICNTL(2) = -1; id.ICNTL(3) = -1; id.ICNTL(4) = 0;   /* :28-31 silent */
   id.ICNTL(6) = 0;                  /* :32 no maximum transversal */
   id.ICNTL(8) = 0;                  /* :33-34 scaling computed at factorization, see below */
   id.ICNTL(13) = 1;                 /* :35 sequential root: exact negative pivot count */
   id.ICNTL(24) = 1;                 /* :36 null pivot detection */
   id.ICNTL(7) = 1;                  /* the one addition: user ordering, the GPU run's permutation */
   id.perm_in = perm_in;
   id.n = (MUMPS_INT) n;
   id.nnz = (MUMPS_INT8) nnz;
   id.irn = irn;
   id.jcn = jcn;
   id.a = a;
   const double t_a0 = now();
   id.job = 1;                       /* :80-81 */
   dmumps_c(&id);
   const double t_analyse = now() - t_a0;
   id.ICNTL(8) = 8;                  /* :82 */
   if (id.INFOG(1) < 0) { printf("{\"error\": \"analysis INFOG(1) = %d\"}\n", id.INFOG(1)); return 1; }

   double best_f = 1e300, best_s = 1e300;
   int neg = 0, zero = 0;
   for (int rep = 0; rep < 4; ++rep) {
      const double t0 = now();
      id.job = 2;                    /* :87-88 */
      dmumps_c(&id);
      const double t1 = now();
      if (id.INFOG(1) < 0) { printf("{\"error\": \"factorization INFOG(1) = %d, INFOG(2) = %d\"}\n", id.INFOG(1), id.INFOG(2)); return 1; }
      neg = id.INFOG(12);            /* :133 */
      zero = id.INFOG(28);           /* :138 */
      for (int64_t i = 0; i < n; ++i) x[i] = b[i];   /* :92 */
      id.rhs = x;
      id.nrhs = 1;
      id.lrhs = (MUMPS_INT) n;
      id.job = 3;                    /* :94-95 */
      dmumps_c(&id);
      const double t2 = now();
      if (rep > 0) {                 /* rep 0 is the warm-up */
         if (t1 - t0 < best_f) best_f = t1 - t0;
         if (t2 - t1 < best_s) best_s = t2 - t1;
      }
   }
   /* residual b - K x on the upper-triangular COO (duplicates summed) */
   double* r = must(malloc(sizeof(double) * (size_t) n));
   double rn = 0.0, xn = 0.0, bn = 0.0;
   double* rowsum = must(calloc((size_t) n, sizeof(double)));
   for (int64_t i = 0; i < n; ++i) r[i] = b[i];
   for (int64_t e = 0; e < nnz; ++e) {
      const int i = irn[e] - 1, j = jcn[e] - 1;
      r[i] -= a[e] * x[j];
      rowsum[i] += fabs(a[e]);
      if (i != j) { r[j] -= a[e] * x[i]; rowsum[j] += fabs(a[e]); }
   }
   double kn = 0.0;
   for (int64_t i = 0; i < n; ++i) {
      if (fabs(r[i]) > rn) rn = fabs(r[i]);
      if (fabs(x[i]) > xn) xn = fabs(x[i]);
      if (fabs(b[i]) > bn) bn = fabs(b[i]);
      if (rowsum[i] > kn) kn = rowsum[i];
   }
   printf("{\"solver\": \"MUMPS %s\", \"n\": %lld, \"analyse_seconds\": %.6f, \"factor_seconds\": %.6f, \"solve_seconds\": %.6f, "
          "\"inertia\": [%lld, %d, %d], \"backward_error\": %.3e, \"flops_estimate\": %.6e}\n",
          MUMPS_VERSION, (long long) n, t_analyse, best_f, best_s, (long long) n - neg - zero, neg, zero, rn / (kn * xn + bn), (double) id.rinfog[2]);
   id.job = -2;                      /* :47-48 */
   dmumps_c(&id);
   return 0;
}
