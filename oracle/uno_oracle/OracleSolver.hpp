// ORACLE — TEST INFRASTRUCTURE ONLY.  Uno-side binding of the CPU oracle (oracle/kkt_oracle.cpp), used by the
// test driver to produce the reference run of the bundled examples (unmodified Uno host code + a CPU
// multifrontal LDL^T standing in for MUMPS, which cannot be installed here).  Never part of the product.
#ifndef UNO_ORACLESOLVER_H
#define UNO_ORACLESOLVER_H

#include "B200Solver.hpp"

namespace uno {
   class OracleSolver : public KKTSolverAdapter {
   public:
      OracleSolver();
   };
} // namespace

#endif // UNO_ORACLESOLVER_H
