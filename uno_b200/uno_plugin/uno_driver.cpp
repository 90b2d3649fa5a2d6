// uno_driver: runs the UNCHANGED Uno host code (compiled from the reference tree) on a built-in model with the
// linear solver chosen by name, the way bindings/AMPL/uno_ampl.cpp:78-139 does for .nl files:
//   uno_driver <file.nl|hs015|polak5|control2d:<k>|control3d:<k>> [option=value ...]      (e.g. preset=ipopt linear_solver=B200)
// Option precedence is the reference's: defaults -> detected solvers -> option file -> preset -> command line.
// After Uno's own summary it prints one machine-readable line starting with "UNO_DRIVER_RESULT" (JSON).
#include <cstdio>
#include <iostream>
#include <memory>
#include <stdexcept>
#include <string>
#include "models/AnalyticModel.hpp"
#include "models/NLModel.hpp"
#include "linear_algebra/COOFormat.hpp"
#include "linear_algebra/SparseSymmetricMatrix.hpp"
#include "model/ModelFactory.hpp"
#include "optimization/Iterate.hpp"
#include "options/DefaultOptions.hpp"
#include "options/Options.hpp"
#include "options/Presets.hpp"
#include "tools/Logger.hpp"
#include "tools/UserCallbacks.hpp"
#include "Uno.hpp"
#include "ingredients/regularization_strategies/PrimalRegularization.hpp"
#include "ingredients/regularization_strategies/UnstableRegularization.hpp"
#include "ingredients/subproblem_solvers/DirectSymmetricIndefiniteLinearSolver.hpp"
#include "symbolic/Range.hpp"
#include "tools/Statistics.hpp"

namespace uno {
   class SilentCallbacks : public UserCallbacks {
   public:
      void notify_acceptable_iterate(const Vector<double>&, const Multipliers&, double) override { }
      void notify_new_primals(const Vector<double>&) override { }
      void notify_new_multipliers(const Multipliers&) override { }
   };

   std::unique_ptr<Model> make_model(const std::string& name) {
      // an AMPL .nl file, as bindings/AMPL/uno_ampl.cpp:95 takes it (read without the ASL: models/NLModel)
      if (name.size() > 3 && name.compare(name.size() - 3, 3, ".nl") == 0) return std::make_unique<NLModel>(name);
      if (name == "hs015") return std::make_unique<HS015Model>();
      if (name == "polak5") return std::make_unique<Polak5Model>();
      for (size_t dimension: {size_t(2), size_t(3)}) {
         const std::string prefix = "control" + std::to_string(dimension) + "d:";
         if (name.rfind(prefix, 0) == 0) {
            return std::make_unique<ControlModel>(static_cast<size_t>(std::stoul(name.substr(prefix.size()))), dimension);
         }
      }
      throw std::invalid_argument("unknown model " + name + " (<file>.nl, hs015, polak5, control2d:<k>, control3d:<k>)");
   }

   // `uno_driver <model> derivatives y0 y1 ...`: every model function at the model's start point with objective multiplier 1 and
   // the given constraint multipliers, one JSON line (tests/test_zz_nl_model.py checks NLModel against closed forms and differences)
   int print_derivatives(const std::string& model_name, int argc, char* argv[]) {
      std::unique_ptr<Model> model = make_model(model_name);
      const size_t n = model->number_variables, m = model->number_constraints;
      Vector<double> x(n), gradient(n), multipliers(m);
      model->initial_primal_point(x);
      for (size_t j = 0; j < m; ++j) multipliers[j] = (3 + static_cast<int>(j) < argc) ? std::stod(argv[3 + j]) : 0.;
      std::vector<double> constraints(m);
      std::printf("UNO_DRIVER_DERIVATIVES {\"n\": %zu, \"m\": %zu, \"objective_sign\": %.1f, \"x\": [", n, m, model->objective_sign);
      for (size_t i = 0; i < n; ++i) std::printf("%s%.17g", i ? ", " : "", x[i]);
      std::printf("], \"objective\": %.17g, \"gradient\": [", model->evaluate_objective(x));
      model->evaluate_objective_gradient(x, gradient);
      for (size_t i = 0; i < n; ++i) std::printf("%s%.17g", i ? ", " : "", gradient[i]);
      model->evaluate_constraints(x, constraints);
      std::printf("], \"constraints\": [");
      for (size_t j = 0; j < m; ++j) std::printf("%s%.17g", j ? ", " : "", constraints[j]);
      std::printf("], \"jacobian\": [");
      bool first = true;
      for (size_t j = 0; j < m; ++j) {
         SparseVector<double> row(n);
         model->evaluate_constraint_gradient(x, j, row);
         for (const auto [variable, value]: row) { std::printf("%s[%zu, %zu, %.17g]", first ? "" : ", ", j, variable, value); first = false; }
      }
      SparseSymmetricMatrix<COOFormat<size_t, double>> hessian(n, model->number_hessian_nonzeros(), 0);
      model->evaluate_lagrangian_hessian(x, 1., multipliers, hessian);
      std::printf("], \"hessian_nonzeros\": %zu, \"hessian\": [", model->number_hessian_nonzeros());
      first = true;
      for (const auto [row, column, value]: hessian) { std::printf("%s[%zu, %zu, %.17g]", first ? "" : ", ", row, column, value); first = false; }
      std::printf("], \"jacobian_nonzeros\": %zu, \"lower\": [", model->number_jacobian_nonzeros());
      for (size_t i = 0; i < n; ++i) std::printf("%s%.17g", i ? ", " : "", model->variable_lower_bound(i));
      std::printf("], \"upper\": [");
      for (size_t i = 0; i < n; ++i) std::printf("%s%.17g", i ? ", " : "", model->variable_upper_bound(i));
      std::printf("], \"constraint_lower\": [");
      for (size_t j = 0; j < m; ++j) std::printf("%s%.17g", j ? ", " : "", model->constraint_lower_bound(j));
      std::printf("], \"constraint_upper\": [");
      for (size_t j = 0; j < m; ++j) std::printf("%s%.17g", j ? ", " : "", model->constraint_upper_bound(j));
      std::printf("]}\n");
      return 0;
   }

   int run(const std::string& model_name, const Options& options) {
      std::unique_ptr<Model> original_model = make_model(model_name);
      DISCRETE << "Original model " << original_model->name << '\n' << original_model->number_variables << " variables, " <<
         original_model->number_constraints << " constraints\n";
      std::unique_ptr<Model> model = ModelFactory::reformulate(std::move(original_model), options);
      DISCRETE << "Reformulated model " << model->name << '\n' << model->number_variables << " variables, " <<
         model->number_constraints << " constraints\n";

      Iterate initial_iterate(model->number_variables, model->number_constraints);
      model->initial_primal_point(initial_iterate.primals);
      model->project_onto_variable_bounds(initial_iterate.primals);
      model->initial_dual_point(initial_iterate.multipliers.constraints);
      initial_iterate.feasibility_multipliers.reset();

      Uno uno{model->number_constraints, options};
      SilentCallbacks callbacks{};
      const Result result = uno.solve(*model, initial_iterate, options, callbacks);
      std::printf("UNO_DRIVER_RESULT {\"model\": \"%s\", \"linear_solver\": \"%s\", \"preset\": \"%s\", \"iterations\": %zu, "
            "\"optimization_status\": \"%s\", \"iterate_status\": \"%s\", \"objective\": %.12e, \"primal_feasibility\": %.6e, "
            "\"stationarity\": %.6e, \"complementarity\": %.6e, \"subproblems\": %zu, \"hessian_evaluations\": %zu}\n",
            model_name.c_str(), options.get_string("linear_solver").c_str(),
            options.get_string_optional("preset").value_or("default").c_str(), result.iteration,
            optimization_status_to_message(result.optimization_status).c_str(), iterate_status_to_message(result.solution.status).c_str(),
            result.solution.evaluations.objective, result.solution.primal_feasibility, result.solution.residuals.stationarity,
            result.solution.residuals.complementarity, result.number_subproblems_solved, result.hessian_evaluations);
      return 0;
   }
} // namespace

namespace uno {
   // The reference's PrimalRegularization (byrd preset: Hessian-only convexification, PrimalRegularization.hpp:79-129), unchanged,
   // with the two sizes that OptimizationProblem would give it and a window on the solver it creates lazily by name (:82-85).
   class PrimalRegularizationHarness : public PrimalRegularization<double> {
   public:
      explicit PrimalRegularizationHarness(const Options& options): PrimalRegularization<double>(options) { }
      void set_sizes(size_t dimension_, size_t number_nonzeros_) { this->dimension = dimension_; this->number_nonzeros = number_nonzeros_; }
      DirectSymmetricIndefiniteLinearSolver<size_t, double>& solver() { return *this->optional_linear_solver; }
   };

   // `uno_driver hessian:<k> [shift=<s>] linear_solver=<name>`: H = 5-point Laplacian on a k x k grid minus s I (indefinite for s > 0),
   // upper-triangular COO with n leading regularization slots; regularize_hessian is called twice (the second time on 2 H: the
   // symbolic analysis is reused), then the regularized system is solved.  One JSON line.
   int run_primal_regularization(const std::string& model_name, const Options& options) {
      const size_t k = static_cast<size_t>(std::stoul(model_name.substr(std::string("hessian:").size())));
      const double shift = options.get_string_optional("shift").has_value() ? std::stod(*options.get_string_optional("shift")) : 1.5;
      const size_t n = k * k;
      const size_t number_nonzeros = n + 2 * k * (k - 1);
      SparseSymmetricMatrix<COOFormat<size_t, double>> hessian(n, number_nonzeros, n);
      auto fill = [&](double factor) {
         hessian.reset();
         for (size_t j = 0; j < k; ++j) {
            for (size_t i = 0; i < k; ++i) {
               const size_t node = i + k * j;
               // column by column, upper triangle (the order of AMPLModel.cpp:170-178)
               if (j > 0) hessian.insert(node - k, node, -factor);
               if (i > 0) hessian.insert(node - 1, node, -factor);
               hessian.insert(node, node, factor * (4. - shift));
            }
         }
      };
      Statistics statistics;
      statistics.add_column("regulariz", Statistics::double_width - 4, 1);
      PrimalRegularizationHarness strategy(options);
      strategy.set_sizes(n, number_nonzeros);
      const Inertia expected_inertia{n, 0, 0};
      double factors[2] = {0., 0.};
      size_t positive = 0, negative = 0, zero = 0;
      double residual = 0., solution_norm = 0.;
      try {
         for (int pass = 0; pass < 2; ++pass) {
            fill(pass == 0 ? 1. : 2.);
            strategy.regularize_hessian(statistics, hessian, Range(n), expected_inertia);
            factors[pass] = strategy.get_primal_regularization_factor();
         }
         const Inertia inertia = strategy.solver().get_inertia();
         positive = inertia.positive; negative = inertia.negative; zero = inertia.zero;
         Vector<double> rhs(n), x(n), product(n);
         for (size_t i = 0; i < n; ++i) rhs[i] = 1. + static_cast<double>(i % 7);
         strategy.solver().solve_indefinite_system(hessian, rhs, x);
         product.fill(0.);
         for (const auto [row_index, column_index, value]: hessian) {
            product[row_index] += value * x[column_index];
            if (row_index != column_index) product[column_index] += value * x[row_index];
         }
         for (size_t i = 0; i < n; ++i) {
            residual = std::max(residual, std::abs(rhs[i] - product[i]));
            solution_norm = std::max(solution_norm, std::abs(x[i]));
         }
      }
      catch (const UnstableRegularization&) {
         std::printf("UNO_DRIVER_RESULT {\"model\": \"%s\", \"error\": \"UnstableRegularization\"}\n", model_name.c_str());
         return 1;
      }
      std::printf("UNO_DRIVER_RESULT {\"model\": \"%s\", \"linear_solver\": \"%s\", \"n\": %zu, \"shift\": %.17g, \"regularization_factors\": [%.17g, %.17g], "
            "\"inertia\": [%zu, %zu, %zu], \"residual\": %.6e, \"solution_norm\": %.6e}\n", model_name.c_str(), options.get_string("linear_solver").c_str(), n,
            shift, factors[0], factors[1], positive, negative, zero, residual, solution_norm);
      return 0;
   }
} // namespace

int main(int argc, char* argv[]) {
   using namespace uno;
   try {
      if (argc < 2) {
         std::cout << "usage: uno_driver <file.nl|hs015|polak5|control2d:<k>|control3d:<k>> [option=value ...]\n";
         Uno::print_available_strategies();
         return EXIT_SUCCESS;
      }
      const std::string model_name(argv[1]);
      if (argc >= 3 && std::string(argv[2]) == "derivatives") return print_derivatives(model_name, argc, argv);
      Options options = DefaultOptions::load();
      Options solvers_options = DefaultOptions::determine_solvers();
      options.overwrite_with(solvers_options);
      Options command_line_options = Options::get_command_line_options(argc, argv, 2);
      const auto optional_option_file = command_line_options.get_string_optional("option_file");
      if (optional_option_file.has_value()) {
         Options file_options = Options::load_option_file(*optional_option_file);
         options.overwrite_with(file_options);
      }
      const auto optional_preset = command_line_options.get_string_optional("preset");
      Options preset_options = Presets::get_preset_options(optional_preset);
      options.overwrite_with(preset_options);
      options.overwrite_with(command_line_options);
      Logger::set_logger(options.get_string("logger"));
      if (model_name.rfind("hessian:", 0) == 0) return run_primal_regularization(model_name, options);
      return run(model_name, options);
   }
   catch (std::exception& exception) {
      std::cout << "uno_driver: " << exception.what() << '\n';
      return EXIT_FAILURE;
   }
}
