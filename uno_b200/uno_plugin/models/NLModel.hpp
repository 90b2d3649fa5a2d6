// NLModel: a Model read from an AMPL .nl file (text format, "g" header) without the AMPL Solver Library.
//
// Stands in for bindings/AMPL/AMPLModel (reference: bindings/AMPL/AMPLModel.cpp:19-181) where ASL is not installable:
// the same conventions - Lagrangian Hessian sigma * Hess f - sum_j y_j Hess c_j (AMPLModel.cpp:37-39), upper triangle emitted
// column by column with ascending rows (AMPLModel.cpp:170-178), nonlinear constraints first (AMPLModel.cpp:53-54), maximization
// turned into minimization by the objective sign (AMPLModel.cpp:47,86,99).  Expressions are kept as a DAG in evaluation order;
// gradients come from one reverse sweep, Hessian columns from forward-over-reverse sweeps (exact, no finite differences); the
// sparsity pattern of the Hessian is the structural one (which pairs of variables meet in a nonlinear operation).
// Supported: segments C O V x d r b k J G (S skipped), operators + - * / ^ neg abs sqrt exp log log10 sin cos tan sinh cosh tanh
// asin acos atan, n-ary sum, min, max.  Imported functions, logical constraints, complementarity and discrete variables are rejected.
#ifndef UNO_NLMODEL_H
#define UNO_NLMODEL_H

#include <string>
#include <utility>
#include <vector>
#include "AnalyticModel.hpp"

namespace uno {
   struct NLData;

   class NLModel : public AnalyticModel {
   public:
      explicit NLModel(const std::string& file_name);
      ~NLModel() override;

      [[nodiscard]] double evaluate_objective(const Vector<double>& x) const override;
      void evaluate_objective_gradient(const Vector<double>& x, Vector<double>& gradient) const override;
      void evaluate_constraints(const Vector<double>& x, std::vector<double>& constraints) const override;
      void evaluate_constraint_gradient(const Vector<double>& x, size_t constraint_index, SparseVector<double>& gradient) const override;
      void evaluate_lagrangian_hessian(const Vector<double>& x, double objective_multiplier, const Vector<double>& multipliers,
         SymmetricMatrix<size_t, double>& hessian) const override;
      void compute_hessian_vector_product(const double* vector, double objective_multiplier, const Vector<double>& multipliers,
         double* result) const override;
      void initial_dual_point(Vector<double>& multipliers) const override;
      [[nodiscard]] size_t number_jacobian_nonzeros() const override;
      [[nodiscard]] size_t number_hessian_nonzeros() const override;

   private:
      NLModel(const std::string& file_name, NLData* data);
      NLData* data;   // owned

      void forward(const double* x) const;
      void seed_lagrangian(double objective_multiplier, const Vector<double>& multipliers) const;
      void hessian_vector(const double* vector, double* result) const;
   };
} // namespace

#endif // UNO_NLMODEL_H
