// See AnalyticModel.hpp.  Derivatives were derived by hand from examples/hs015.mod and examples/polak5.mod.
#include <cmath>
#include "AnalyticModel.hpp"
#include "symbolic/Range.hpp"
#include "tools/Infinity.hpp"

namespace uno {
   AnalyticModel::AnalyticModel(const std::string& name, std::vector<double> variable_lower, std::vector<double> variable_upper,
         std::vector<double> constraint_lower, std::vector<double> constraint_upper, std::vector<double> initial_point,
         size_t first_linear_constraint, double objective_sign) :
         Model(name, variable_lower.size(), constraint_lower.size(), objective_sign),
         variable_lower(std::move(variable_lower)), variable_upper(std::move(variable_upper)),
         constraint_lower(std::move(constraint_lower)), constraint_upper(std::move(constraint_upper)),
         initial_point(std::move(initial_point)),
         linear_constraints(first_linear_constraint, this->number_constraints),
         equality_collection(this->equality_constraints), inequality_collection(this->inequality_constraints),
         lower_bounded_collection(this->lower_bounded), upper_bounded_collection(this->upper_bounded),
         single_lower_collection(this->single_lower), single_upper_collection(this->single_upper) {
      for (size_t variable_index: Range(this->number_variables)) {
         const double lower = this->variable_lower[variable_index], upper = this->variable_upper[variable_index];
         if (lower == upper) {
            this->fixed_variables.emplace_back(variable_index);
         }
         else if (is_finite(lower) && is_finite(upper)) {
            this->lower_bounded.emplace_back(variable_index);
            this->upper_bounded.emplace_back(variable_index);
         }
         else if (is_finite(lower)) {
            this->lower_bounded.emplace_back(variable_index);
            this->single_lower.emplace_back(variable_index);
         }
         else if (is_finite(upper)) {
            this->upper_bounded.emplace_back(variable_index);
            this->single_upper.emplace_back(variable_index);
         }
      }
      for (size_t constraint_index: Range(this->number_constraints)) {
         if (this->constraint_lower[constraint_index] == this->constraint_upper[constraint_index]) {
            this->equality_constraints.emplace_back(constraint_index);
         }
         else {
            this->inequality_constraints.emplace_back(constraint_index);
         }
      }
   }

   void AnalyticModel::evaluate_constraint_jacobian(const Vector<double>& x, RectangularMatrix<double>& constraint_jacobian) const {
      for (size_t constraint_index: Range(this->number_constraints)) {
         constraint_jacobian[constraint_index].clear();
         this->evaluate_constraint_gradient(x, constraint_index, constraint_jacobian[constraint_index]);
      }
   }

   void AnalyticModel::initial_primal_point(Vector<double>& x) const {
      for (size_t variable_index: Range(this->number_variables)) {
         x[variable_index] = this->initial_point[variable_index];
      }
   }

   void AnalyticModel::initial_dual_point(Vector<double>& multipliers) const {
      for (size_t constraint_index: Range(this->number_constraints)) {
         multipliers[constraint_index] = 0.;
      }
   }

   void AnalyticModel::postprocess_solution(Iterate& /*iterate*/, IterateStatus /*termination_status*/) const {
      // minimization problems without solution file: nothing to undo
   }

   // ---- hs015:  min 100 (x2 - x1^2)^2 + (1 - x1)^2   s.t.  x1 x2 >= 1,  x1 + x2^2 >= 0,  x1 <= 1/2;  start (-2, 1) ----------
   HS015Model::HS015Model() : AnalyticModel("hs015", {-infinity, -infinity}, {0.5, infinity}, {1., 0.}, {infinity, infinity}, {-2., 1.}, 2) {
   }

   double HS015Model::evaluate_objective(const Vector<double>& x) const {
      const double t = x[1] - x[0] * x[0];
      return 100. * t * t + (1. - x[0]) * (1. - x[0]);
   }

   void HS015Model::evaluate_objective_gradient(const Vector<double>& x, Vector<double>& gradient) const {
      const double t = x[1] - x[0] * x[0];
      gradient[0] = -400. * x[0] * t - 2. * (1. - x[0]);
      gradient[1] = 200. * t;
   }

   void HS015Model::evaluate_constraints(const Vector<double>& x, std::vector<double>& constraints) const {
      constraints[0] = x[0] * x[1];
      constraints[1] = x[0] + x[1] * x[1];
   }

   void HS015Model::evaluate_constraint_gradient(const Vector<double>& x, size_t constraint_index, SparseVector<double>& gradient) const {
      gradient.clear();
      if (constraint_index == 0) {
         gradient.insert(0, x[1]);
         gradient.insert(1, x[0]);
      }
      else {
         gradient.insert(0, 1.);
         gradient.insert(1, 2. * x[1]);
      }
   }

   void HS015Model::evaluate_lagrangian_hessian(const Vector<double>& x, double objective_multiplier, const Vector<double>& multipliers,
         SymmetricMatrix<size_t, double>& hessian) const {
      const double h00 = objective_multiplier * (1200. * x[0] * x[0] - 400. * x[1] + 2.);
      const double h01 = objective_multiplier * (-400. * x[0]) - multipliers[0];
      const double h11 = objective_multiplier * 200. - 2. * multipliers[1];
      hessian.insert(0, 0, h00);
      hessian.finalize_column(0);
      hessian.insert(0, 1, h01);
      hessian.insert(1, 1, h11);
      hessian.finalize_column(1);
   }

   void HS015Model::compute_hessian_vector_product(const double* vector, double objective_multiplier, const Vector<double>& multipliers,
         double* result) const {
      // not reachable with an explicit Hessian model; kept exact for completeness at the point last evaluated is unknown -> unsupported
      (void) vector; (void) objective_multiplier; (void) multipliers; (void) result;
      throw std::runtime_error("HS015Model: Hessian-vector products are not provided");
   }

   // ---- polak5:  min u   s.t.  -u + 3 x1^2 + 50 (x1 - x2^4 -+ 1)^2 <= 0;  start x = 0.1, u = 0 ---------------------------------
   Polak5Model::Polak5Model() : AnalyticModel("polak5", {-infinity, -infinity, -infinity}, {infinity, infinity, infinity},
         {-infinity, -infinity}, {0., 0.}, {0.1, 0.1, 0.}, 2) {
   }

   double Polak5Model::evaluate_objective(const Vector<double>& x) const {
      return x[2];
   }

   void Polak5Model::evaluate_objective_gradient(const Vector<double>& /*x*/, Vector<double>& gradient) const {
      gradient[0] = 0.;
      gradient[1] = 0.;
      gradient[2] = 1.;
   }

   void Polak5Model::evaluate_constraints(const Vector<double>& x, std::vector<double>& constraints) const {
      const double x2_4 = x[1] * x[1] * x[1] * x[1];
      const double g_minus = x[0] - x2_4 - 1., g_plus = x[0] - x2_4 + 1.;
      constraints[0] = -x[2] + 3. * x[0] * x[0] + 50. * g_minus * g_minus;
      constraints[1] = -x[2] + 3. * x[0] * x[0] + 50. * g_plus * g_plus;
   }

   void Polak5Model::evaluate_constraint_gradient(const Vector<double>& x, size_t constraint_index, SparseVector<double>& gradient) const {
      gradient.clear();
      const double x2_3 = x[1] * x[1] * x[1];
      const double g = x[0] - x2_3 * x[1] + ((constraint_index == 0) ? -1. : 1.);
      gradient.insert(0, 6. * x[0] + 100. * g);
      gradient.insert(1, -400. * x2_3 * g);
      gradient.insert(2, -1.);
   }

   void Polak5Model::evaluate_lagrangian_hessian(const Vector<double>& x, double /*objective_multiplier*/, const Vector<double>& multipliers,
         SymmetricMatrix<size_t, double>& hessian) const {
      // the objective is linear; Hess c_j = [[106, -400 x2^3], [-400 x2^3, 1600 x2^6 - 1200 x2^2 g_j]]
      const double x2_2 = x[1] * x[1], x2_3 = x2_2 * x[1], x2_4 = x2_2 * x2_2;
      double h00 = 0., h01 = 0., h11 = 0.;
      for (size_t constraint_index: Range(2)) {
         const double g = x[0] - x2_4 + ((constraint_index == 0) ? -1. : 1.);
         const double y = multipliers[constraint_index];
         h00 -= y * 106.;
         h01 -= y * (-400. * x2_3);
         h11 -= y * (1600. * x2_3 * x2_3 - 1200. * x2_2 * g);
      }
      hessian.insert(0, 0, h00);
      hessian.finalize_column(0);
      hessian.insert(0, 1, h01);
      hessian.insert(1, 1, h11);
      hessian.finalize_column(1);
      hessian.finalize_column(2);
   }

   void Polak5Model::compute_hessian_vector_product(const double* vector, double objective_multiplier, const Vector<double>& multipliers,
         double* result) const {
      (void) vector; (void) objective_multiplier; (void) multipliers; (void) result;
      throw std::runtime_error("Polak5Model: Hessian-vector products are not provided");
   }

   // ---- distributed control on a grid ----------------------------------------------------------------------------------------
   namespace {
      size_t power(size_t base, size_t exponent) {
         size_t result = 1;
         for (size_t i = 0; i < exponent; ++i) result *= base;
         return result;
      }
      std::vector<double> control_bounds(size_t nodes, double state_bound, double control_bound) {
         std::vector<double> bounds(2 * nodes, state_bound);
         for (size_t i = nodes; i < 2 * nodes; ++i) bounds[i] = control_bound;
         return bounds;
      }
   } // namespace

   ControlModel::ControlModel(size_t grid_size, size_t grid_dimension, double alpha) :
         AnalyticModel("control" + std::to_string(grid_dimension) + "d_k" + std::to_string(grid_size),
            control_bounds(power(grid_size, grid_dimension), -infinity, -1.), control_bounds(power(grid_size, grid_dimension), infinity, 1.),
            std::vector<double>(power(grid_size, grid_dimension), 0.), std::vector<double>(power(grid_size, grid_dimension), 0.),
            std::vector<double>(2 * power(grid_size, grid_dimension), 0.), 0),
         grid_size(grid_size), grid_dimension(grid_dimension), nodes(power(grid_size, grid_dimension)), alpha(alpha),
         inverse_h_squared(static_cast<double>((grid_size + 1) * (grid_size + 1))), target(this->nodes), source(this->nodes) {
      // smooth target state and source term (deterministic)
      for (size_t node = 0; node < this->nodes; ++node) {
         double product = 1.;
         size_t rest = node;
         for (size_t d = 0; d < grid_dimension; ++d) {
            const double coordinate = static_cast<double>(rest % grid_size + 1) / static_cast<double>(grid_size + 1);
            product *= std::sin(3.141592653589793 * coordinate);
            rest /= grid_size;
         }
         this->target[node] = 2. * product;
         this->source[node] = 1. + product;
      }
      for (size_t node = 0; node < this->nodes; ++node) {
         this->jacobian_nonzeros += 2;   // diagonal of the Laplacian and the control
         this->visit_neighbours(node, [&](size_t) { ++this->jacobian_nonzeros; });
      }
   }

   template <typename Visitor>
   void ControlModel::visit_neighbours(size_t node, Visitor&& visitor) const {
      size_t stride = 1, rest = node;
      for (size_t d = 0; d < this->grid_dimension; ++d) {
         const size_t coordinate = rest % this->grid_size;
         if (coordinate > 0) visitor(node - stride);
         if (coordinate + 1 < this->grid_size) visitor(node + stride);
         rest /= this->grid_size;
         stride *= this->grid_size;
      }
   }

   double ControlModel::evaluate_objective(const Vector<double>& x) const {
      double objective = 0.;
      for (size_t node = 0; node < this->nodes; ++node) {
         const double misfit = x[node] - this->target[node];
         objective += 0.5 * misfit * misfit + 0.5 * this->alpha * x[this->nodes + node] * x[this->nodes + node];
      }
      return objective;
   }

   void ControlModel::evaluate_objective_gradient(const Vector<double>& x, Vector<double>& gradient) const {
      for (size_t node = 0; node < this->nodes; ++node) {
         gradient[node] = x[node] - this->target[node];
         gradient[this->nodes + node] = this->alpha * x[this->nodes + node];
      }
   }

   void ControlModel::evaluate_constraints(const Vector<double>& x, std::vector<double>& constraints) const {
      const double diagonal = 2. * static_cast<double>(this->grid_dimension) * this->inverse_h_squared;
      for (size_t node = 0; node < this->nodes; ++node) {
         double value = diagonal * x[node] - x[this->nodes + node] - this->source[node];
         this->visit_neighbours(node, [&](size_t neighbour) { value -= this->inverse_h_squared * x[neighbour]; });
         constraints[node] = value;
      }
   }

   void ControlModel::evaluate_constraint_gradient(const Vector<double>& /*x*/, size_t constraint_index, SparseVector<double>& gradient) const {
      gradient.clear();
      const double diagonal = 2. * static_cast<double>(this->grid_dimension) * this->inverse_h_squared;
      gradient.insert(constraint_index, diagonal);
      this->visit_neighbours(constraint_index, [&](size_t neighbour) { gradient.insert(neighbour, -this->inverse_h_squared); });
      gradient.insert(this->nodes + constraint_index, -1.);
   }

   void ControlModel::evaluate_lagrangian_hessian(const Vector<double>& /*x*/, double objective_multiplier, const Vector<double>& /*multipliers*/,
         SymmetricMatrix<size_t, double>& hessian) const {
      for (size_t variable_index = 0; variable_index < this->number_variables; ++variable_index) {
         hessian.insert(variable_index, variable_index, objective_multiplier * ((variable_index < this->nodes) ? 1. : this->alpha));
         hessian.finalize_column(variable_index);
      }
   }

   void ControlModel::compute_hessian_vector_product(const double* vector, double objective_multiplier, const Vector<double>& /*multipliers*/,
         double* result) const {
      for (size_t variable_index = 0; variable_index < this->number_variables; ++variable_index) {
         result[variable_index] = objective_multiplier * ((variable_index < this->nodes) ? 1. : this->alpha) * vector[variable_index];
      }
   }
} // namespace
