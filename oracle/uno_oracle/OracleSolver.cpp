// ORACLE — TEST INFRASTRUCTURE ONLY (see OracleSolver.hpp).
#include "OracleSolver.hpp"

extern "C" {
void* oracle_create();
void oracle_destroy(void* p);
const char* oracle_last_error(void* p);
int oracle_analyse(void* p, int64_t n, int64_t nnz, const uint64_t* row, const uint64_t* col, const int64_t* perm);
int oracle_set_parameters(void* p, double threshold_u, double null_tol);
int oracle_factorize(void* p, const double* values, int nthreads);
int oracle_inertia(void* p, uint64_t* pos, uint64_t* neg, uint64_t* zero);
int oracle_solve(void* p, const double* rhs, double* x);
}

namespace uno {
   namespace {
      // the oracle reports failure as a non-zero status; the adapter treats negative values as fatal
      int negate(int status) { return status == 0 ? 0 : -1; }
      int oracle_destroy_(void* handle) { oracle_destroy(handle); return 0; }
      int oracle_analyse_(void* handle, uint64_t dimension, uint64_t number_nonzeros, uint64_t, uint64_t, const uint64_t* rows,
            const uint64_t* columns) {
         return negate(oracle_analyse(handle, static_cast<int64_t>(dimension), static_cast<int64_t>(number_nonzeros), rows, columns, nullptr));
      }
      int oracle_factorize_(void* handle, const double* values) { return negate(oracle_factorize(handle, values, 1)); }
      int oracle_inertia_(const void* handle, uint64_t* positive, uint64_t* negative, uint64_t* zero) {
         return negate(oracle_inertia(const_cast<void*>(handle), positive, negative, zero));
      }
      int oracle_solve_(void* handle, const double* rhs, double* x) { return negate(oracle_solve(handle, rhs, x)); }
      const char* oracle_last_error_(const void* handle) { return oracle_last_error(const_cast<void*>(handle)); }

      KKTBackend make_oracle_backend() {
         KKTBackend backend;
         backend.name = "ORACLE";
         backend.handle = oracle_create();
         oracle_set_parameters(backend.handle, 0.01, 2.2e-21);
         backend.destroy = oracle_destroy_;
         backend.analyse = oracle_analyse_;
         backend.factorize = oracle_factorize_;
         backend.refactorize_shifted = nullptr;
         backend.inertia = oracle_inertia_;
         backend.solve = oracle_solve_;
         backend.last_error = oracle_last_error_;
         return backend;
      }
   } // namespace

   OracleSolver::OracleSolver(): KKTSolverAdapter(make_oracle_backend()) {
   }
} // namespace
