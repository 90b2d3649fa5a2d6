// Replacement for uno/ingredients/subproblem_solvers/SymmetricIndefiniteLinearSolverFactory.cpp of the
// reference (same class, same two static functions: create :31-61, available_solvers :64-84), compiled
// INSTEAD of that file: it registers the B200 solver under the name "B200" exactly the way the
// reference registers "MUMPS"/"MA57"/"MA27" behind HAS_<LIB> macros (CMakeLists.txt:107-117).
// With no other HAS_* macro defined, "B200" is the only and therefore the default linear solver
// (uno/options/DefaultOptions.cpp:203-206).
//
// HAS_KKT_ORACLE is defined only by the test driver build (oracle/uno_oracle): it registers the CPU
// oracle under "ORACLE" so that reference runs of the bundled examples can be generated where MUMPS
// is not installable.  The product build never defines it.
#include <stdexcept>
#include <string>
#include "ingredients/subproblem_solvers/SymmetricIndefiniteLinearSolverFactory.hpp"
#include "ingredients/subproblem_solvers/DirectSymmetricIndefiniteLinearSolver.hpp"
#include "linear_algebra/Vector.hpp"

#ifdef HAS_B200
#include "B200Solver.hpp"
#endif
#ifdef HAS_KKT_ORACLE
#include "OracleSolver.hpp"
#endif

namespace uno {
   std::unique_ptr<DirectSymmetricIndefiniteLinearSolver<size_t, double>> SymmetricIndefiniteLinearSolverFactory::create(
         const std::string& linear_solver) {
#ifdef HAS_B200
      if (linear_solver == "B200") {
         return std::make_unique<B200Solver>();
      }
#endif
#ifdef HAS_KKT_ORACLE
      if (linear_solver == "ORACLE") {
         return std::make_unique<OracleSolver>();
      }
#endif
      std::string message = "The linear solver " + linear_solver + " is unknown\nThe following values are available: ";
      message.append(join(SymmetricIndefiniteLinearSolverFactory::available_solvers(), ", "));
      throw std::invalid_argument(message);
   }

   std::vector<std::string> SymmetricIndefiniteLinearSolverFactory::available_solvers() {
      std::vector<std::string> solvers{};
#ifdef HAS_B200
      solvers.emplace_back("B200");
#endif
#ifdef HAS_KKT_ORACLE
      solvers.emplace_back("ORACLE");
#endif
      return solvers;
   }
} // namespace
