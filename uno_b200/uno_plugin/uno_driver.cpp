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
      return run(model_name, options);
   }
   catch (std::exception& exception) {
      std::cout << "uno_driver: " << exception.what() << '\n';
      return EXIT_FAILURE;
   }
}
