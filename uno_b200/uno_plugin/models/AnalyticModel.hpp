// Models with hand-written derivatives, standing in for bindings/AMPL/AMPLModel (the AMPL Solver Library is not
// installable here).  AnalyticModel implements the bookkeeping half of Uno's Model interface
// (uno/model/Model.hpp:36-88) from plain bound vectors, the way AMPLModel does it from the ASL structures
// (bindings/AMPL/AMPLModel.cpp:314-352); concrete models only supply the functions and their derivatives.
// Conventions follow AMPLModel: Lagrangian Hessian = sigma * Hess f - sum_j y_j Hess c_j (AMPLModel.cpp:37-39),
// emitted upper-triangular, column by column (AMPLModel.cpp:170-178).
#ifndef UNO_ANALYTICMODEL_H
#define UNO_ANALYTICMODEL_H

#include <limits>
#include <string>
#include <vector>
#include "linear_algebra/RectangularMatrix.hpp"
#include "linear_algebra/SparseVector.hpp"
#include "linear_algebra/SymmetricMatrix.hpp"
#include "linear_algebra/Vector.hpp"
#include "model/Model.hpp"
#include "optimization/Iterate.hpp"
#include "symbolic/CollectionAdapter.hpp"
#include "symbolic/Range.hpp"

namespace uno {
   class AnalyticModel : public Model {
   public:
      AnalyticModel(const std::string& name, std::vector<double> variable_lower, std::vector<double> variable_upper,
            std::vector<double> constraint_lower, std::vector<double> constraint_upper, std::vector<double> initial_point,
            size_t first_linear_constraint, double objective_sign = 1.);

      void evaluate_constraint_jacobian(const Vector<double>& x, RectangularMatrix<double>& constraint_jacobian) const override;

      [[nodiscard]] double variable_lower_bound(size_t variable_index) const override { return this->variable_lower[variable_index]; }
      [[nodiscard]] double variable_upper_bound(size_t variable_index) const override { return this->variable_upper[variable_index]; }
      [[nodiscard]] const Collection<size_t>& get_lower_bounded_variables() const override { return this->lower_bounded_collection; }
      [[nodiscard]] const Collection<size_t>& get_upper_bounded_variables() const override { return this->upper_bounded_collection; }
      [[nodiscard]] const SparseVector<size_t>& get_slacks() const override { return this->slacks; }
      [[nodiscard]] const Collection<size_t>& get_single_lower_bounded_variables() const override { return this->single_lower_collection; }
      [[nodiscard]] const Collection<size_t>& get_single_upper_bounded_variables() const override { return this->single_upper_collection; }
      [[nodiscard]] const Vector<size_t>& get_fixed_variables() const override { return this->fixed_variables; }

      [[nodiscard]] double constraint_lower_bound(size_t constraint_index) const override { return this->constraint_lower[constraint_index]; }
      [[nodiscard]] double constraint_upper_bound(size_t constraint_index) const override { return this->constraint_upper[constraint_index]; }
      [[nodiscard]] const Collection<size_t>& get_equality_constraints() const override { return this->equality_collection; }
      [[nodiscard]] const Collection<size_t>& get_inequality_constraints() const override { return this->inequality_collection; }
      [[nodiscard]] const Collection<size_t>& get_linear_constraints() const override { return this->linear_constraints; }

      void initial_primal_point(Vector<double>& x) const override;
      void initial_dual_point(Vector<double>& multipliers) const override;
      void postprocess_solution(Iterate& iterate, IterateStatus termination_status) const override;

   protected:
      static constexpr double infinity = std::numeric_limits<double>::infinity();
      const std::vector<double> variable_lower, variable_upper, constraint_lower, constraint_upper, initial_point;
      ForwardRange linear_constraints;
      std::vector<size_t> equality_constraints{}, inequality_constraints{};
      std::vector<size_t> lower_bounded{}, upper_bounded{}, single_lower{}, single_upper{};
      CollectionAdapter<std::vector<size_t>&> equality_collection, inequality_collection;
      CollectionAdapter<std::vector<size_t>&> lower_bounded_collection, upper_bounded_collection, single_lower_collection, single_upper_collection;
      SparseVector<size_t> slacks{};
      Vector<size_t> fixed_variables{};
   };

   // examples/hs015.mod:1-12 of the reference (Hock-Schittkowski 15)
   class HS015Model : public AnalyticModel {
   public:
      HS015Model();
      [[nodiscard]] double evaluate_objective(const Vector<double>& x) const override;
      void evaluate_objective_gradient(const Vector<double>& x, Vector<double>& gradient) const override;
      void evaluate_constraints(const Vector<double>& x, std::vector<double>& constraints) const override;
      void evaluate_constraint_gradient(const Vector<double>& x, size_t constraint_index, SparseVector<double>& gradient) const override;
      void evaluate_lagrangian_hessian(const Vector<double>& x, double objective_multiplier, const Vector<double>& multipliers,
         SymmetricMatrix<size_t, double>& hessian) const override;
      void compute_hessian_vector_product(const double* vector, double objective_multiplier, const Vector<double>& multipliers,
         double* result) const override;
      [[nodiscard]] size_t number_jacobian_nonzeros() const override { return 4; }
      [[nodiscard]] size_t number_hessian_nonzeros() const override { return 3; }
   };

   // examples/polak5.mod:21-29 of the reference
   class Polak5Model : public AnalyticModel {
   public:
      Polak5Model();
      [[nodiscard]] double evaluate_objective(const Vector<double>& x) const override;
      void evaluate_objective_gradient(const Vector<double>& x, Vector<double>& gradient) const override;
      void evaluate_constraints(const Vector<double>& x, std::vector<double>& constraints) const override;
      void evaluate_constraint_gradient(const Vector<double>& x, size_t constraint_index, SparseVector<double>& gradient) const override;
      void evaluate_lagrangian_hessian(const Vector<double>& x, double objective_multiplier, const Vector<double>& multipliers,
         SymmetricMatrix<size_t, double>& hessian) const override;
      void compute_hessian_vector_product(const double* vector, double objective_multiplier, const Vector<double>& multipliers,
         double* result) const override;
      [[nodiscard]] size_t number_jacobian_nonzeros() const override { return 6; }
      [[nodiscard]] size_t number_hessian_nonzeros() const override { return 3; }
   };

   // Distributed-control problem on a k^dim grid (BASELINE.json configs 2 and 4 as a full optimization problem):
   //   min 1/2 |y - y_d|^2 + alpha/2 |u|^2   s.t.  -Laplace_h y - u = f,   -1 <= u <= 1
   // x = [y ; u]; one linear equality per grid node.
   class ControlModel : public AnalyticModel {
   public:
      ControlModel(size_t grid_size, size_t grid_dimension, double alpha = 1e-3);
      [[nodiscard]] double evaluate_objective(const Vector<double>& x) const override;
      void evaluate_objective_gradient(const Vector<double>& x, Vector<double>& gradient) const override;
      void evaluate_constraints(const Vector<double>& x, std::vector<double>& constraints) const override;
      void evaluate_constraint_gradient(const Vector<double>& x, size_t constraint_index, SparseVector<double>& gradient) const override;
      void evaluate_lagrangian_hessian(const Vector<double>& x, double objective_multiplier, const Vector<double>& multipliers,
         SymmetricMatrix<size_t, double>& hessian) const override;
      void compute_hessian_vector_product(const double* vector, double objective_multiplier, const Vector<double>& multipliers,
         double* result) const override;
      [[nodiscard]] size_t number_jacobian_nonzeros() const override { return this->jacobian_nonzeros; }
      [[nodiscard]] size_t number_hessian_nonzeros() const override { return this->number_variables; }

   private:
      const size_t grid_size, grid_dimension, nodes;
      const double alpha, inverse_h_squared;
      std::vector<double> target, source;
      size_t jacobian_nonzeros{0};
      template <typename Visitor>
      void visit_neighbours(size_t node, Visitor&& visitor) const;
   };
} // namespace

#endif // UNO_ANALYTICMODEL_H
