// B200Solver: Uno-side adapter of the B200 KKT solver.  Implements Uno's
// DirectSymmetricIndefiniteLinearSolver<size_t, double>
// (reference: uno/ingredients/subproblem_solvers/DirectSymmetricIndefiniteLinearSolver.hpp:13-25,
//  uno/ingredients/subproblem_solvers/SymmetricIndefiniteLinearSolver.hpp:20-33) on top of the C ABI
// in include/b200kkt.h, so that it registers next to "MUMPS"/"MA57"/"MA27" in
// SymmetricIndefiniteLinearSolverFactory and the ipopt/byrd presets, the inertia-correction loop
// and the globalization code run unchanged.  It mirrors the reference's MUMPSSolver
// (uno/ingredients/subproblem_solvers/MUMPS/MUMPSSolver.hpp:17-70): same ownership of the augmented
// matrix / rhs / solution buffers, same call order, same meaning of every query.
//
// The class is a template over a small "backend" of plain C functions so that the very same
// orchestration code can drive the CUDA library (product) and, in the test driver only, the CPU
// oracle that stands in for MUMPS when reference runs are generated.
#ifndef UNO_B200SOLVER_H
#define UNO_B200SOLVER_H

#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>
#include "ingredients/subproblem_solvers/DirectSymmetricIndefiniteLinearSolver.hpp"
#include "ingredients/subproblem/Subproblem.hpp"
#include "linear_algebra/COOFormat.hpp"
#include "linear_algebra/RectangularMatrix.hpp"
#include "linear_algebra/SparseSymmetricMatrix.hpp"
#include "linear_algebra/SymmetricMatrix.hpp"
#include "linear_algebra/Vector.hpp"
#include "optimization/Direction.hpp"
#include "optimization/WarmstartInformation.hpp"

namespace uno {
   // What a backend has to provide (all return 0 on success, like include/b200kkt.h).
   struct KKTBackend {
      void* handle{nullptr};
      int (*destroy)(void* handle){nullptr};
      int (*analyse)(void* handle, uint64_t dimension, uint64_t number_nonzeros, uint64_t regularization_size, uint64_t number_primal,
         const uint64_t* rows, const uint64_t* columns){nullptr};
      // optional: the values the matrix holds when its pattern is analysed (a hint for the KKT pairing; the reference hands
      // do_symbolic_analysis a filled matrix: PrimalDualRegularization.hpp:145-149)
      int (*pairing_values)(void* handle, const double* values, uint64_t number_nonzeros){nullptr};
      // optional: named numeric options of the backend (include/b200kkt.h: b200kkt_set_option)
      int (*set_option)(void* handle, const char* name, double value){nullptr};
      int (*factorize)(void* handle, const double* values){nullptr};
      // optional fast path: only the leading regularization slots changed since the last factorization
      int (*refactorize_shifted)(void* handle, const double* regularization_values){nullptr};
      int (*inertia)(const void* handle, uint64_t* positive, uint64_t* negative, uint64_t* zero){nullptr};
      int (*solve)(void* handle, const double* rhs, double* x){nullptr};
      const char* (*last_error)(const void* handle){nullptr};
      std::string name{};
   };

   class KKTSolverAdapter : public DirectSymmetricIndefiniteLinearSolver<size_t, double> {
   public:
      explicit KKTSolverAdapter(KKTBackend backend);
      ~KKTSolverAdapter() override;
      KKTSolverAdapter(const KKTSolverAdapter&) = delete;
      KKTSolverAdapter& operator=(const KKTSolverAdapter&) = delete;

      void initialize_memory(size_t number_variables, size_t number_constraints, size_t number_hessian_nonzeros,
         size_t regularization_size) override;

      void do_symbolic_analysis(const SymmetricMatrix<size_t, double>& matrix) override;
      void do_numerical_factorization(const SymmetricMatrix<size_t, double>& matrix) override;
      void solve_indefinite_system(const SymmetricMatrix<size_t, double>& matrix, const Vector<double>& rhs, Vector<double>& result) override;
      void solve_indefinite_system(Statistics& statistics, const Subproblem& subproblem, Direction& direction,
         const WarmstartInformation& warmstart_information) override;

      [[nodiscard]] Inertia get_inertia() const override;
      [[nodiscard]] size_t number_negative_eigenvalues() const override;
      [[nodiscard]] size_t number_zero_eigenvalues() const;
      [[nodiscard]] bool matrix_is_singular() const override;
      [[nodiscard]] size_t rank() const override;

      // counters read by the test driver
      size_t number_analyses{0}, number_factorizations{0}, number_shifted_refactorizations{0}, number_solves{0};

   protected:
      KKTBackend backend;
      size_t number_primal{0};           // variables of the problem this instance was sized for
      size_t analysed_dimension{0};      // dimension of the pattern of the last analysis
      size_t analysed_regularization_size{0};   // regularization slots of this instance's own augmented matrix
      size_t shifted_prefix{0};                 // leading slots that the shifted refactorization uploads
      uint64_t inertia_triple[3]{0, 0, 0};

      // pattern of the last analysis and values of the last factorization (to recognise the inertia loop)
      std::vector<uint64_t> row_indices{}, column_indices{};
      std::vector<double> last_values{};
      bool has_factorization{false};

      // evaluations and augmented system, owned here as in MUMPSSolver.hpp:46-54
      Vector<double> objective_gradient;
      std::vector<double> constraints;
      RectangularMatrix<double> constraint_jacobian;
      SparseSymmetricMatrix<COOFormat<size_t, double>> augmented_matrix{};
      Vector<double> rhs{};
      Vector<double> solution{};

      void check(int status, const char* what) const;
   };

   // the product binding: backend = libb200kkt.so (include/b200kkt.h), device and GPU count from the environment
   // (B200KKT_DEVICE, default 0) since Uno's factory passes nothing but the solver name
   class B200Solver : public KKTSolverAdapter {
   public:
      B200Solver();
   };
} // namespace

#endif // UNO_B200SOLVER_H
