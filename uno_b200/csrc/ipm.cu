// the O(nnz) / O(N) host steps either side of the factorization, on the device.
//
//   rank 1  KKT values generated where they are consumed: the reference fills the COO entry by entry on the host
//           (Subproblem.cpp:57-70, PrimalDualInteriorPointProblem.cpp:56-78, COOFormat.hpp:101-110) and MUMPS ingests the array;
//           here one kernel writes regularization slots, model Hessian, barrier diagonals and Jacobian straight into the handle's
//           resident value array, and a refactorization with other (delta_w, delta_c) rewrites only the slots - no PCIe traffic.
//   rank 3  right-hand side [-(grad f_mu - J^T y); -c] (Subproblem.cpp:80-99), K x (SymmetricMatrix.hpp:99-109), and the direction
//           unpack with bound-dual steps and both fraction-to-boundary rules (PrimalDualInteriorPointProblem.cpp:173-194,262-325).
//
// Every sum runs in the order of the reference's loops (ascending COO slot) with separate multiply and add (no FMA contraction), so
// the results equal the host code's bit for bit when that is compiled without FMA (x86-64 default).
#include "solver.hpp"

#include <cmath>
#include <cstring>
#include <numeric>
#include <stdexcept>
#include <string>

using namespace b200kkt;

namespace {

__device__ __forceinline__ bool finite_bound(const double b) { return fabs(b) < 1.0e300 && b == b; }   // the reference marks missing bounds with +-inf

// one thread per COO entry of the value array
__global__ void k_kkt_values(const b200kkt_kkt_layout L, const double* __restrict__ hess, const int* __restrict__ barrier_var, const double* __restrict__ x,
                             const double* __restrict__ lower, const double* __restrict__ upper, const double* __restrict__ zl, const double* __restrict__ zu,
                             const double* __restrict__ jac, const double delta_w, const double delta_c, double* __restrict__ values) {
   const uint64_t e = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
   const uint64_t r0 = L.num_primal_slots, r1 = r0 + L.num_dual_slots, h1 = r1 + L.nnz_hessian, b1 = h1 + L.num_barrier, j1 = b1 + L.nnz_jacobian;
   if (e >= j1) return;
   double v;
   if (e < r0) v = delta_w;                       // set_regularization(primal, 0, +delta_w)      PrimalDualRegularization.hpp:178
   else if (e < r1) v = -delta_c;                 // set_regularization(dual, n, -delta_c)        :179
   else if (e < h1) v = hess[e - r1];
   else if (e < b1) {
      // PrimalDualInteriorPointProblem.cpp:62-76
      const int i = barrier_var[e - h1];
      const double xi = x[i], lo = lower[i], up = upper[i];
      double t = 0.0;
      if (finite_bound(lo)) t = __dadd_rn(t, __ddiv_rn(zl[i], __dsub_rn(xi, lo)));
      if (finite_bound(up)) t = __dadd_rn(t, __ddiv_rn(zu[i], __dsub_rn(xi, up)));
      v = t;
   }
   else v = jac[e - b1];
   values[e] = v;
}

__global__ void k_shift_slots(const uint64_t num_primal_slots, const uint64_t num_slots, const double delta_w, const double delta_c, double* __restrict__ values) {
   const uint64_t e = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
   if (e < num_slots) values[e] = (e < num_primal_slots) ? delta_w : -delta_c;
}

// rhs[i] = -g[i] + sum over the Jacobian entries (i, n + j) in ascending slot order of y[j] * v   (i < n);  rhs[n + j] = -c[j]
__global__ void k_rhs(const int64_t n, const int64_t m, const int* __restrict__ jt_ptr, const int* __restrict__ jt_slot, const int* __restrict__ jt_con,
                      const double* __restrict__ values, const double* __restrict__ g, const double* __restrict__ c, const double* __restrict__ y,
                      double* __restrict__ rhs) {
   const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
   if (i < n) {
      double s = -g[i];
      for (int t = jt_ptr[i]; t < jt_ptr[i + 1]; ++t) {
         const double yj = y[jt_con[t]];
         if (yj != 0.0) s = __dadd_rn(s, __dmul_rn(yj, values[jt_slot[t]]));
      }
      rhs[i] = s;
   } else if (i < n + m) {
      rhs[i] = -c[i - n];
   }
}

// y = K x: row i takes its entries in ascending slot order, as the reference's loop over the COO does for result[i]
__global__ void k_symmetric_product(CsrArrays A, const double* __restrict__ values, const double* __restrict__ x, double* __restrict__ y) {
   const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
   if (i >= A.n) return;
   double s = 0.0;
   for (int t = A.ptr[i]; t < A.ptr[i + 1]; ++t) s = __dadd_rn(s, __dmul_rn(values[A.slot[t]], x[A.col[t]]));
   y[i] = s;
}

__device__ __forceinline__ void atomic_min_positive(unsigned long long* slot, const double v) {
   // positive doubles: the order of the values is the order of the bit patterns
   atomicMin(slot, static_cast<unsigned long long>(__double_as_longlong(v)));
}

// dx = sol[0:n], dy = -sol[n:n+m], bound-dual directions, and the two step lengths as minima (bit patterns, initialised to 1.0)
__global__ void k_direction(const int64_t n, const int64_t m, const double mu, const double tau, const double* __restrict__ sol, const double* __restrict__ x,
                            const double* __restrict__ lower, const double* __restrict__ upper, const double* __restrict__ zl, const double* __restrict__ zu,
                            double* __restrict__ dx, double* __restrict__ dy, double* __restrict__ dzl, double* __restrict__ dzu, unsigned long long* steps) {
   const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
   double pmin = 1.0, dmin = 1.0;
   if (i < n) {
      const double d = sol[i], xi = x[i], lo = lower[i], up = upper[i];
      dx[i] = d;
      double l = 0.0, u = 0.0;
      if (finite_bound(lo)) {
         const double dist = __dsub_rn(xi, lo), z = zl[i];
         l = __dsub_rn(__ddiv_rn(__dsub_rn(mu, __dmul_rn(d, z)), dist), z);                 // PrimalDualInteriorPointProblem.cpp:266-270
         if (d < 0.0) { const double t = __ddiv_rn(__dmul_rn(-tau, dist), d); if (0.0 < t) pmin = fmin(pmin, t); }          // :284-289
         if (l < 0.0) { const double t = __ddiv_rn(__dmul_rn(-tau, z), l); if (0.0 < t) dmin = fmin(dmin, t); }             // :308-313
      }
      if (finite_bound(up)) {
         const double dist = __dsub_rn(xi, up), z = zu[i];
         u = __dsub_rn(__ddiv_rn(__dsub_rn(mu, __dmul_rn(d, z)), dist), z);                 // :272-276
         if (0.0 < d) { const double t = __ddiv_rn(__dmul_rn(-tau, dist), d); if (0.0 < t) pmin = fmin(pmin, t); }          // :292-297
         if (0.0 < u) { const double t = __ddiv_rn(__dmul_rn(-tau, z), u); if (0.0 < t) dmin = fmin(dmin, t); }             // :316-321
      }
      dzl[i] = l;
      dzu[i] = u;
   } else if (i < n + m) {
      dy[i - n] = -sol[i];
   }
   for (int o = 16; o; o >>= 1) {
      pmin = fmin(pmin, __shfl_xor_sync(0xffffffffu, pmin, o));
      dmin = fmin(dmin, __shfl_xor_sync(0xffffffffu, dmin, o));
   }
   if ((threadIdx.x & 31) == 0) {
      if (pmin < 1.0) atomic_min_positive(steps + 0, pmin);
      if (dmin < 1.0) atomic_min_positive(steps + 1, dmin);
   }
}

__global__ void k_scale_direction(const int64_t n, const int64_t m, const unsigned long long* __restrict__ steps, double* __restrict__ dx, double* __restrict__ dy,
                                  double* __restrict__ dzl, double* __restrict__ dzu) {
   const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
   const double ps = __longlong_as_double(static_cast<long long>(steps[0])), ds = __longlong_as_double(static_cast<long long>(steps[1]));
   if (i < n) {
      dx[i] = __dmul_rn(dx[i], ps);
      dzl[i] = __dmul_rn(dzl[i], ds);
      dzu[i] = __dmul_rn(dzu[i], ds);
   } else if (i < n + m) {
      dy[i - n] = __dmul_rn(dy[i - n], ps);
   }
}

int fail_ipm(b200kkt_handle* h, int code, const std::string& msg) {
   h->err = msg;
   return code;
}

int cuda_fail(b200kkt_handle* h, cudaError_t e, const char* what) {
   cudaGetLastError();
   return fail_ipm(h, e == cudaErrorMemoryAllocation ? B200KKT_ENOMEM : B200KKT_ECUDA, std::string(what) + ": CUDA error: " + cudaGetErrorString(e));
}

#define IPM_CK(expr)                                                \
   do {                                                             \
      const cudaError_t e__ = (expr);                               \
      if (e__ != cudaSuccess) return cuda_fail(h, e__, #expr);      \
   } while (0)

int ready(b200kkt_handle* h, const char* who) {
   if (h == nullptr) return B200KKT_EINVAL;
   if (!h->has_device) return fail_ipm(h, B200KKT_ECUDA, std::string(who) + ": handle has no CUDA device: there is no CPU path");
   if (!h->analysed) return fail_ipm(h, B200KKT_EINVAL, std::string(who) + " before analyse");
   return B200KKT_OK;
}

inline unsigned grid_for(uint64_t n) { return static_cast<unsigned>((n + 255) / 256); }

// Jacobian entries (i, n + j) of the analysed pattern grouped by variable i, ascending slot inside a group; built on first use
int build_jacobian_index(b200kkt_handle* h) {
   if (h->d_jt_ptr != nullptr) return B200KKT_OK;
   const Plan& P = h->plan;
   const int64_t n = P.num_primal;
   std::vector<int> ptr(static_cast<size_t>(n) + 1, 0), slot, con;
   for (int64_t e = 0; e < P.nnz; ++e)
      if (P.coo_row[e] < n && P.coo_col[e] >= n) ++ptr[static_cast<size_t>(P.coo_row[e]) + 1];
   for (int64_t i = 0; i < n; ++i) ptr[i + 1] += ptr[i];
   slot.resize(static_cast<size_t>(ptr[n]));
   con.resize(static_cast<size_t>(ptr[n]));
   std::vector<int> fill(ptr.begin(), ptr.end() - 1);
   for (int64_t e = 0; e < P.nnz; ++e)
      if (P.coo_row[e] < n && P.coo_col[e] >= n) {
         const int t = fill[P.coo_row[e]]++;
         slot[t] = static_cast<int>(e);
         con[t] = static_cast<int>(P.coo_col[e] - n);
      }
   auto upload = [&](const std::vector<int>& v, int*& dst) -> cudaError_t {
      cudaError_t e = cudaMalloc(reinterpret_cast<void**>(&dst), std::max<size_t>(v.size(), 1) * sizeof(int));
      if (e != cudaSuccess) return e;
      h->plan_allocs.push_back(dst);
      return cudaMemcpyAsync(dst, v.data(), v.size() * sizeof(int), cudaMemcpyHostToDevice, h->stream);
   };
   IPM_CK(upload(ptr, h->d_jt_ptr));
   IPM_CK(upload(slot, h->d_jt_slot));
   IPM_CK(upload(con, h->d_jt_con));
   IPM_CK(cudaStreamSynchronize(h->stream));   // the host vectors go out of scope
   return B200KKT_OK;
}

}  // namespace

extern "C" {

int b200kkt_assemble_kkt_values_device(b200kkt_handle* h, const b200kkt_kkt_layout* layout, const double* d_hessian_values, const int32_t* d_barrier_variable,
                                       const double* d_x, const double* d_lower, const double* d_upper, const double* d_zl, const double* d_zu,
                                       const double* d_jacobian_values, double delta_w, double delta_c) {
   if (int rc = ready(h, "assemble_kkt_values_device")) return rc;
   if (layout == nullptr) return B200KKT_EINVAL;
   const b200kkt_kkt_layout& L = *layout;
   const uint64_t total = L.num_primal_slots + L.num_dual_slots + L.nnz_hessian + L.num_barrier + L.nnz_jacobian;
   if (total != static_cast<uint64_t>(h->plan.nnz) || L.num_primal_slots + L.num_dual_slots != static_cast<uint64_t>(h->plan.reg_size))
      return fail_ipm(h, B200KKT_EINVAL, "assemble_kkt_values_device: the layout does not add up to the analysed pattern (nnz " + std::to_string(h->plan.nnz) +
                                            ", regularization slots " + std::to_string(h->plan.reg_size) + ")");
   if ((L.nnz_hessian > 0 && d_hessian_values == nullptr) || (L.nnz_jacobian > 0 && d_jacobian_values == nullptr) ||
       (L.num_barrier > 0 && (d_barrier_variable == nullptr || d_x == nullptr || d_lower == nullptr || d_upper == nullptr || d_zl == nullptr || d_zu == nullptr)))
      return B200KKT_EINVAL;
   IPM_CK(cudaSetDevice(h->device));
   if (total > 0) {
      k_kkt_values<<<grid_for(total), 256, 0, h->stream>>>(L, d_hessian_values, d_barrier_variable, d_x, d_lower, d_upper, d_zl, d_zu, d_jacobian_values, delta_w,
                                                           delta_c, h->d_values);
      ++h->launches;
   }
   IPM_CK(cudaGetLastError());
   h->values_resident = true;
   return B200KKT_OK;
}

int b200kkt_shift_diagonal_device(b200kkt_handle* h, uint64_t num_primal_slots, double delta_w, double delta_c) {
   if (int rc = ready(h, "shift_diagonal_device")) return rc;
   if (!h->values_resident) return fail_ipm(h, B200KKT_EINVAL, "shift_diagonal_device: no resident values (factorize or assemble_kkt_values_device first)");
   if (num_primal_slots > static_cast<uint64_t>(h->plan.reg_size)) return fail_ipm(h, B200KKT_EINVAL, "shift_diagonal_device: more primal slots than regularization slots");
   IPM_CK(cudaSetDevice(h->device));
   if (h->plan.reg_size > 0) {
      k_shift_slots<<<grid_for(static_cast<uint64_t>(h->plan.reg_size)), 256, 0, h->stream>>>(num_primal_slots, static_cast<uint64_t>(h->plan.reg_size), delta_w, delta_c,
                                                                                         h->d_values);
      ++h->launches;
   }
   IPM_CK(cudaGetLastError());
   return B200KKT_OK;
}

int b200kkt_get_values_device(b200kkt_handle* h, double** d_values) {
   if (int rc = ready(h, "get_values_device")) return rc;
   if (d_values == nullptr) return B200KKT_EINVAL;
   *d_values = h->d_values;
   return B200KKT_OK;
}

int b200kkt_assemble_rhs_device(b200kkt_handle* h, const double* d_gradient, const double* d_constraints, const double* d_multipliers, double* d_rhs) {
   if (int rc = ready(h, "assemble_rhs_device")) return rc;
   if (!h->values_resident) return fail_ipm(h, B200KKT_EINVAL, "assemble_rhs_device: no resident values (the Jacobian is read from them)");
   if (d_gradient == nullptr || d_rhs == nullptr) return B200KKT_EINVAL;
   const int64_t n = h->plan.num_primal, m = h->plan.N - n;
   if (m > 0 && (d_constraints == nullptr || d_multipliers == nullptr)) return B200KKT_EINVAL;
   IPM_CK(cudaSetDevice(h->device));
   if (int rc = build_jacobian_index(h)) return rc;
   k_rhs<<<grid_for(static_cast<uint64_t>(h->plan.N)), 256, 0, h->stream>>>(n, m, h->d_jt_ptr, h->d_jt_slot, h->d_jt_con, h->d_values, d_gradient, d_constraints,
                                                                              d_multipliers, d_rhs);
   ++h->launches;
   IPM_CK(cudaGetLastError());
   return B200KKT_OK;
}

int b200kkt_symmetric_product_device(b200kkt_handle* h, const double* d_x, double* d_y) {
   if (int rc = ready(h, "symmetric_product_device")) return rc;
   if (!h->values_resident) return fail_ipm(h, B200KKT_EINVAL, "symmetric_product_device: no resident values");
   if (d_x == nullptr || d_y == nullptr || d_x == d_y) return B200KKT_EINVAL;
   IPM_CK(cudaSetDevice(h->device));
   k_symmetric_product<<<grid_for(static_cast<uint64_t>(h->plan.N)), 256, 0, h->stream>>>(h->C, h->d_values, d_x, d_y);
   ++h->launches;
   IPM_CK(cudaGetLastError());
   return B200KKT_OK;
}

int b200kkt_unpack_direction_device(b200kkt_handle* h, double barrier_parameter, double tau, const double* d_solution, const double* d_x, const double* d_lower,
                                    const double* d_upper, const double* d_zl, const double* d_zu, double* d_dx, double* d_dy, double* d_dzl, double* d_dzu,
                                    double* primal_step_length, double* dual_step_length) {
   if (int rc = ready(h, "unpack_direction_device")) return rc;
   if (d_solution == nullptr || d_x == nullptr || d_lower == nullptr || d_upper == nullptr || d_zl == nullptr || d_zu == nullptr || d_dx == nullptr || d_dzl == nullptr ||
       d_dzu == nullptr)
      return B200KKT_EINVAL;
   const int64_t n = h->plan.num_primal, m = h->plan.N - n;
   if (m > 0 && d_dy == nullptr) return B200KKT_EINVAL;
   IPM_CK(cudaSetDevice(h->device));
   // the two minima start at 1.0 (the reference's initial step lengths); the handle's norm words 5 and 6 are free between solves
   unsigned long long* steps = h->d_norms + 5;
   const double one = 1.0;
   unsigned long long one_bits;
   static_assert(sizeof(one_bits) == sizeof(one), "bit copy");
   memcpy(&one_bits, &one, sizeof(one));
   const unsigned long long init[2] = {one_bits, one_bits};
   IPM_CK(cudaMemcpyAsync(steps, init, sizeof(init), cudaMemcpyHostToDevice, h->stream));
   const unsigned grid = grid_for(static_cast<uint64_t>(h->plan.N));
   k_direction<<<grid, 256, 0, h->stream>>>(n, m, barrier_parameter, tau, d_solution, d_x, d_lower, d_upper, d_zl, d_zu, d_dx, d_dy, d_dzl, d_dzu, steps);
   k_scale_direction<<<grid, 256, 0, h->stream>>>(n, m, steps, d_dx, d_dy, d_dzl, d_dzu);
   h->launches += 2;
   unsigned long long out[2] = {0, 0};
   IPM_CK(cudaMemcpyAsync(out, steps, sizeof(out), cudaMemcpyDeviceToHost, h->stream));
   IPM_CK(cudaStreamSynchronize(h->stream));
   IPM_CK(cudaGetLastError());
   double ps, ds;
   memcpy(&ps, &out[0], sizeof(ps));
   memcpy(&ds, &out[1], sizeof(ds));
   if (primal_step_length != nullptr) *primal_step_length = ps;
   if (dual_step_length != nullptr) *dual_step_length = ds;
   return B200KKT_OK;
}

}  // extern "C"
