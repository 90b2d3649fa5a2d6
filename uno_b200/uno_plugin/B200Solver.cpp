// Adapter between Uno's linear-solver interface and the C ABI of the B200 KKT solver.
// Call-by-call counterpart of uno/ingredients/subproblem_solvers/MUMPS/MUMPSSolver.cpp:
//   constructor/destructor  <-> JOB=-1 / JOB=-2            (:16-49)
//   initialize_memory       <-> :51-70
//   do_symbolic_analysis    <-> JOB=1 + index copy          (:72-83, :149-157)
//   do_numerical_factorization <-> JOB=2                    (:85-89)
//   solve_indefinite_system(matrix, rhs, result) <-> JOB=3  (:91-96)
//   solve_indefinite_system(statistics, subproblem, ...)    (:98-122)
//   inertia queries         <-> INFOG(12), INFOG(28)        (:124-147)
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include "B200Solver.hpp"
#include <sys/stat.h>
#include <ctime>
#include "b200kkt.h"

namespace uno {
   KKTSolverAdapter::KKTSolverAdapter(KKTBackend backend): DirectSymmetricIndefiniteLinearSolver<size_t, double>(), backend(std::move(backend)) {
   }

   KKTSolverAdapter::~KKTSolverAdapter() {
      if (this->backend.handle != nullptr && this->backend.destroy != nullptr) {
         this->backend.destroy(this->backend.handle);
      }
   }

   void KKTSolverAdapter::check(int status, const char* what) const {
      // fatal errors surface as exceptions (caught in Uno::solve -> ALGORITHMIC_ERROR); a singular
      // matrix is never an error, it is reported through the inertia
      if (status < 0) {
         std::string message = this->backend.name + ": " + what + " failed";
         if (this->backend.last_error != nullptr) {
            message.append(": ").append(this->backend.last_error(this->backend.handle));
         }
         throw std::runtime_error(message);
      }
   }

   void KKTSolverAdapter::initialize_memory(size_t number_variables, size_t number_constraints, size_t number_hessian_nonzeros,
         size_t regularization_size) {
      // the third argument is a capacity that excludes the regularization slots; from the interior-point
      // method it already contains the Jacobian nonzeros (PrimalDualInteriorPointMethod.cpp:54-57)
      const size_t dimension = number_variables + number_constraints;
      this->number_primal = number_variables;
      this->analysed_dimension = 0;
      this->row_indices.reserve(number_hessian_nonzeros + regularization_size);
      this->column_indices.reserve(number_hessian_nonzeros + regularization_size);
      this->last_values.reserve(number_hessian_nonzeros + regularization_size);

      this->objective_gradient.resize(number_variables);
      this->constraints.resize(number_constraints);
      this->constraint_jacobian.resize(number_constraints, number_variables);

      this->augmented_matrix = SparseSymmetricMatrix<COOFormat<size_t, double>>(dimension, number_hessian_nonzeros, regularization_size);
      this->analysed_regularization_size = regularization_size;
      this->rhs.resize(dimension);
      this->solution.resize(dimension);
      // An instance sized without constraints is the one PrimalRegularization creates for the Hessian convexification
      // (PrimalRegularization.hpp:82-85: initialize_memory(dimension, 0, ...)): every matrix it sees is meant to become positive
      // definite and only "inertia == (n, 0, 0)?" is asked, so the factorization can skip the pivot search.
      if (this->backend.set_option != nullptr) {
         this->check(this->backend.set_option(this->backend.handle, "positive_definite_target", number_constraints == 0 ? 1. : 0.), "option");
      }
   }

   void KKTSolverAdapter::do_symbolic_analysis(const SymmetricMatrix<size_t, double>& matrix) {
      // the index arrays are only reachable through the iterator (SparseSymmetricMatrix keeps its storage protected)
      this->row_indices.clear();
      this->column_indices.clear();
      for (const auto [row_index, column_index, value]: matrix) {
         (void) value;
         this->row_indices.push_back(static_cast<uint64_t>(row_index));
         this->column_indices.push_back(static_cast<uint64_t>(column_index));
      }
      const size_t dimension = matrix.dimension();
      // Is this the augmented system this instance was sized for, or a foreign matrix (least-square multipliers:
      // Preprocessing.cpp:42-62; Hessian-only convexification: PrimalRegularization.hpp:90-129)?
      const bool own_system = (&matrix == static_cast<const SymmetricMatrix<size_t, double>*>(&this->augmented_matrix));
      const uint64_t number_primal_rows = (own_system && this->number_primal < dimension) ? this->number_primal : 0;
      // leading (k, k) placeholders: only trusted when they really are diagonal entries in positions 0, 1, ...
      size_t regularization_size = 0;
      if (own_system) {
         regularization_size = this->analysed_regularization_size;
      }
      else {
         while (regularization_size < this->row_indices.size() && this->row_indices[regularization_size] == this->column_indices[regularization_size] &&
               this->row_indices[regularization_size] == regularization_size) {
            ++regularization_size;
         }
      }
      if (this->backend.pairing_values != nullptr && matrix.number_nonzeros() == this->row_indices.size()) {
         this->check(this->backend.pairing_values(this->backend.handle, matrix.data_pointer(), this->row_indices.size()), "pairing hint");
      }
      this->check(this->backend.analyse(this->backend.handle, dimension, this->row_indices.size(), regularization_size, number_primal_rows,
            this->row_indices.data(), this->column_indices.data()), "symbolic analysis");
      this->analysed_dimension = dimension;
      this->shifted_prefix = regularization_size;
      this->has_factorization = false;
      ++this->number_analyses;
   }

   void KKTSolverAdapter::do_numerical_factorization(const SymmetricMatrix<size_t, double>& matrix) {
      const double* values = matrix.data_pointer();
      const size_t number_nonzeros = matrix.number_nonzeros();
      if (number_nonzeros != this->row_indices.size()) {
         throw std::runtime_error(this->backend.name + ": the sparsity pattern changed since the symbolic analysis");
      }
      // inside the inertia-correction loop only the leading regularization slots change between two
      // factorizations (PrimalDualRegularization.hpp:178-186, :210-211): ship just those
      const size_t prefix = this->shifted_prefix;
      const bool only_shift = this->has_factorization && this->backend.refactorize_shifted != nullptr && prefix > 0 &&
            this->last_values.size() == number_nonzeros &&
            std::memcmp(values + prefix, this->last_values.data() + prefix, (number_nonzeros - prefix) * sizeof(double)) == 0;
      if (only_shift) {
         this->check(this->backend.refactorize_shifted(this->backend.handle, values), "numerical factorization (shifted)");
         std::memcpy(this->last_values.data(), values, prefix * sizeof(double));
         ++this->number_shifted_refactorizations;
      }
      else {
         this->check(this->backend.factorize(this->backend.handle, values), "numerical factorization");
         this->last_values.assign(values, values + number_nonzeros);
      }
      this->has_factorization = true;
      ++this->number_factorizations;
      this->check(this->backend.inertia(this->backend.handle, &this->inertia_triple[0], &this->inertia_triple[1], &this->inertia_triple[2]),
            "inertia");
   }

   void KKTSolverAdapter::solve_indefinite_system(const SymmetricMatrix<size_t, double>& /*matrix*/, const Vector<double>& rhs,
         Vector<double>& result) {
      // the factors of the last factorization are used, whatever matrix is passed (as with MUMPS JOB=3)
      this->check(this->backend.solve(this->backend.handle, rhs.data(), result.data()), "solve");
      ++this->number_solves;
   }

   void KKTSolverAdapter::solve_indefinite_system(Statistics& statistics, const Subproblem& subproblem, Direction& direction,
         const WarmstartInformation& warmstart_information) {
      const bool new_objective = warmstart_information.objective_changed;
      const bool new_constraints = warmstart_information.constraints_changed;
      if (new_objective) {
         subproblem.evaluate_objective_gradient(this->objective_gradient);
      }
      if (new_constraints) {
         subproblem.evaluate_constraints(this->constraints);
         subproblem.evaluate_jacobian(this->constraint_jacobian);
      }
      if (new_objective || new_constraints) {
         // rebuild K = [[H + Sigma, J^T], [J, 0]] in COO; the regularization strategy then calls back into
         // do_symbolic_analysis (once) / do_numerical_factorization / get_inertia / matrix_is_singular of this object
         this->augmented_matrix.reset();
         subproblem.assemble_augmented_matrix(statistics, this->augmented_matrix, this->constraint_jacobian);
         subproblem.regularize_augmented_matrix(statistics, this->augmented_matrix, subproblem.dual_regularization_factor(), *this);
         subproblem.assemble_augmented_rhs(this->objective_gradient, this->constraints, this->constraint_jacobian, this->rhs);
      }
      // otherwise the previous factorization is reused (only the right-hand side path runs)
      this->solve_indefinite_system(this->augmented_matrix, this->rhs, this->solution);
      subproblem.assemble_primal_dual_direction(this->solution, direction);
   }

   Inertia KKTSolverAdapter::get_inertia() const {
      return {static_cast<size_t>(this->inertia_triple[0]), static_cast<size_t>(this->inertia_triple[1]),
            static_cast<size_t>(this->inertia_triple[2])};
   }

   size_t KKTSolverAdapter::number_negative_eigenvalues() const {
      return static_cast<size_t>(this->inertia_triple[1]);
   }

   size_t KKTSolverAdapter::number_zero_eigenvalues() const {
      return static_cast<size_t>(this->inertia_triple[2]);
   }

   bool KKTSolverAdapter::matrix_is_singular() const {
      return this->inertia_triple[2] > 0;
   }

   size_t KKTSolverAdapter::rank() const {
      // dimension of the matrix that was actually factorized (the reference uses the dimension passed to
      // initialize_memory, which is off after the least-square-multiplier pre-solve: MUMPSSolver.cpp:145-147)
      return this->analysed_dimension - static_cast<size_t>(this->inertia_triple[2]);
   }

   // ---- product binding ---------------------------------------------------------------------------
   namespace {
      int b200_destroy(void* handle) { return b200kkt_destroy(static_cast<b200kkt_handle*>(handle)); }
      int b200_analyse(void* handle, uint64_t dimension, uint64_t number_nonzeros, uint64_t regularization_size, uint64_t number_primal,
            const uint64_t* rows, const uint64_t* columns) {
         return b200kkt_analyse(static_cast<b200kkt_handle*>(handle), dimension, number_nonzeros, regularization_size, number_primal, rows, columns);
      }
      int b200_pairing_values(void* handle, const double* values, uint64_t number_nonzeros) {
         return b200kkt_set_pairing_values(static_cast<b200kkt_handle*>(handle), values, number_nonzeros);
      }
      int b200_set_option(void* handle, const char* name, double value) { return b200kkt_set_option(static_cast<b200kkt_handle*>(handle), name, value); }
      int b200_factorize(void* handle, const double* values) { return b200kkt_factorize(static_cast<b200kkt_handle*>(handle), values); }
      int b200_refactorize(void* handle, const double* values) { return b200kkt_refactorize_shifted(static_cast<b200kkt_handle*>(handle), values); }
      int b200_inertia(const void* handle, uint64_t* positive, uint64_t* negative, uint64_t* zero) {
         return b200kkt_inertia(static_cast<const b200kkt_handle*>(handle), positive, negative, zero);
      }
      int b200_solve(void* handle, const double* rhs, double* x) { return b200kkt_solve(static_cast<b200kkt_handle*>(handle), rhs, x, nullptr); }
      const char* b200_last_error(const void* handle) { return b200kkt_last_error(static_cast<const b200kkt_handle*>(handle)); }

      KKTBackend make_b200_backend() {
         KKTBackend backend;
         backend.name = "B200";
         // Several GPUs: N copies of the Uno process run the same iteration (SPMD, like MUMPS over MPI ranks), one GPU each:
         //   B200KKT_WORLD = N, B200KKT_RANK = 0..N-1, B200KKT_COMM_FILE = a path all ranks can reach (NCCL id hand-over).
         // Every solver instance gets its own communicator; the instances are created in the same order on every rank.
         const char* world_variable = std::getenv("B200KKT_WORLD");
         const char* rank_variable = std::getenv("B200KKT_RANK");
         const int world = (world_variable != nullptr) ? std::max(1, std::atoi(world_variable)) : 1;
         const int rank = (rank_variable != nullptr) ? std::atoi(rank_variable) : 0;
         const char* device_variable = std::getenv("B200KKT_DEVICE");
         const int device = (device_variable != nullptr) ? std::atoi(device_variable) : ((world > 1) ? rank : 0);
         b200kkt_handle* handle = nullptr;
         int status = b200kkt_create(&handle, device, world);
         if (status == B200KKT_OK && world > 1) {
            static int instance_counter = 0;
            const char* file_variable = std::getenv("B200KKT_COMM_FILE");
            if (file_variable == nullptr) {
               b200kkt_destroy(handle);
               throw std::runtime_error("B200: B200KKT_WORLD > 1 needs B200KKT_COMM_FILE (path for the NCCL id hand-over)");
            }
            // One file per solver instance; B200KKT_RUN_ID (a nonce the launcher gives every rank of one run) keeps runs apart.
            // Without it a file left behind by a run that died between publishing and joining is told by its age; a run that
            // ends normally leaves nothing: rank 0 removes the file once every rank has joined (comm_init is collective).
            const char* run_variable = std::getenv("B200KKT_RUN_ID");
            const std::string path = std::string(file_variable) + (run_variable != nullptr ? "." + std::string(run_variable) : std::string()) + "." +
               std::to_string(instance_counter++);
            char id[128];
            if (rank == 0) {
               std::remove(path.c_str());   // a stale id of an earlier run must not be read by a rank that starts before the new one is out
               status = b200kkt_comm_unique_id(id);
               if (status == B200KKT_OK) {
                  const std::string temporary = path + ".tmp";
                  std::FILE* file = std::fopen(temporary.c_str(), "wb");
                  if (file == nullptr || std::fwrite(id, 1, sizeof(id), file) != sizeof(id)) status = B200KKT_EINVAL;
                  if (file != nullptr) std::fclose(file);
                  if (status == B200KKT_OK && std::rename(temporary.c_str(), path.c_str()) != 0) status = B200KKT_EINVAL;   // atomic publish
                  if (status != B200KKT_OK) std::remove(temporary.c_str());
               }
            }
            else {
               status = B200KKT_EINVAL;
               const std::time_t oldest_accepted = std::time(nullptr) - 600;
               for (int attempt = 0; attempt < 6000 && status != B200KKT_OK; ++attempt) {   // up to 60 s
                  struct stat info;
                  if (::stat(path.c_str(), &info) == 0 && info.st_mtime >= oldest_accepted) {
                     std::FILE* file = std::fopen(path.c_str(), "rb");
                     if (file != nullptr) {
                        if (std::fread(id, 1, sizeof(id), file) == sizeof(id)) status = B200KKT_OK;
                        std::fclose(file);
                     }
                  }
                  if (status != B200KKT_OK) std::this_thread::sleep_for(std::chrono::milliseconds(10));
               }
            }
            if (status == B200KKT_OK) status = b200kkt_comm_init(handle, rank, world, id);
            if (rank == 0) std::remove(path.c_str());   // joined (or failed): nobody needs the file any more
         }
         if (status != B200KKT_OK) {
            std::string message = "B200: cannot create a solver handle on CUDA device " + std::to_string(device);
            if (handle != nullptr) {
               message.append(": ").append(b200kkt_last_error(handle));
               b200kkt_destroy(handle);
            }
            throw std::runtime_error(message);   // no CPU fallback
         }
         backend.handle = handle;
         backend.destroy = b200_destroy;
         backend.analyse = b200_analyse;
         backend.pairing_values = b200_pairing_values;
         backend.set_option = b200_set_option;
         backend.factorize = b200_factorize;
         backend.refactorize_shifted = b200_refactorize;
         backend.inertia = b200_inertia;
         backend.solve = b200_solve;
         backend.last_error = b200_last_error;
         return backend;
      }
   } // namespace

   B200Solver::B200Solver(): KKTSolverAdapter(make_b200_backend()) {
   }
} // namespace
