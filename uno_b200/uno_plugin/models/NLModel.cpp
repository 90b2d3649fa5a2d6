// See NLModel.hpp.  Reader of the text .nl format ("Writing .nl Files", D. M. Gay) and exact derivatives on the expression DAG.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <fstream>
#include <limits>
#include <memory>
#include <set>
#include <sstream>
#include <stdexcept>
#include "NLModel.hpp"
#include "symbolic/Range.hpp"

namespace uno {
   namespace {
      // node kinds below zero, .nl operator codes from zero on
      constexpr int OP_CONST = -1, OP_VAR = -2;
      constexpr int O_PLUS = 0, O_MINUS = 1, O_MULT = 2, O_DIV = 3, O_POW = 5, O_MIN = 11, O_MAX = 12, O_ABS = 15, O_NEG = 16, O_TANH = 37,
            O_TAN = 38, O_SQRT = 39, O_SINH = 40, O_SIN = 41, O_LOG10 = 42, O_LOG = 43, O_EXP = 44, O_COSH = 45, O_COS = 46, O_ATAN = 49,
            O_ASIN = 51, O_ACOS = 53, O_SUM = 54;

      bool is_binary(int op) { return op == O_PLUS || op == O_MINUS || op == O_MULT || op == O_DIV || op == O_POW; }
      bool is_nary(int op) { return op == O_MIN || op == O_MAX || op == O_SUM; }
      bool is_unary(int op) {
         switch (op) {
            case O_ABS: case O_NEG: case O_TANH: case O_TAN: case O_SQRT: case O_SINH: case O_SIN: case O_LOG10: case O_LOG: case O_EXP:
            case O_COSH: case O_COS: case O_ATAN: case O_ASIN: case O_ACOS: return true;
            default: return false;
         }
      }

      struct Node {
         int op{OP_CONST};
         int a{-1}, b{-1};         // operands (node indices, smaller than the node's own index); OP_VAR: a = variable index
         double constant{0.};
         std::vector<int> args{};  // n-ary operators
      };

      // first and second partial derivatives of w = f(u, v) (unary: only the u entries)
      struct Partials { double fu{0.}, fv{0.}, fuu{0.}, fuv{0.}, fvv{0.}; };
   }

   struct NLData {
      size_t n_var{0}, n_con{0}, n_obj{0}, nlc{0};
      double objective_sign{1.};
      std::vector<double> variable_lower{}, variable_upper{}, constraint_lower{}, constraint_upper{}, x0{}, pi0{};
      std::vector<Node> nodes{};
      std::vector<int> variable_node{};                      // node of variable i (the n_var variables first, then defined variables)
      std::vector<int> constraint_root{}, constraint_first{}; // per constraint: root (-1: purely linear) and first node of its expression
      int objective_root{-1}, objective_first{0};
      std::vector<std::vector<std::pair<size_t, double>>> jacobian{};   // per constraint: (variable, linear coefficient), file order
      std::vector<std::pair<size_t, double>> objective_linear{};
      bool has_defined_variables{false};
      std::vector<std::vector<size_t>> hessian_rows{};       // per column j: rows i <= j of the structural pattern, ascending
      size_t hessian_nonzeros{0}, jacobian_nonzeros{0};
      // evaluation workspace
      std::vector<double> value{}, dot{}, adjoint{}, adjoint_dot{}, x{}, column{}, unit{};
      bool values_current{false};   // value[] belongs to x[]

      [[nodiscard]] double evaluate(const Node& node) const;
      [[nodiscard]] Partials partials(const Node& node) const;
      [[nodiscard]] int selected_argument(const Node& node) const;
   };

   namespace {
      class Reader {
      public:
         explicit Reader(const std::string& file_name) {
            std::ifstream file(file_name);
            if (!file) throw std::runtime_error("NLModel: cannot open " + file_name);
            std::string line;
            while (std::getline(file, line)) {
               const size_t hash = line.find('#');
               if (hash != std::string::npos) line.erase(hash);
               while (!line.empty() && (line.back() == ' ' || line.back() == '\t' || line.back() == '\r')) line.pop_back();
               this->lines.push_back(line);
            }
         }
         [[nodiscard]] bool done() const { return this->position >= this->lines.size(); }
         const std::string& next() {
            if (this->done()) throw std::runtime_error("NLModel: unexpected end of file");
            return this->lines[this->position++];
         }
         std::vector<double> numbers() {
            std::istringstream stream(this->next());
            std::vector<double> result;
            double v;
            while (stream >> v) result.push_back(v);
            return result;
         }
      private:
         std::vector<std::string> lines{};
         size_t position{0};
      };

      int add_node(NLData& data, Node node) {
         data.nodes.push_back(std::move(node));
         return static_cast<int>(data.nodes.size()) - 1;
      }

      int add_constant(NLData& data, double constant) {
         Node node;
         node.op = OP_CONST;
         node.constant = constant;
         return add_node(data, std::move(node));
      }

      int parse_expression(Reader& reader, NLData& data) {
         const std::string& line = reader.next();
         if (line.empty()) throw std::runtime_error("NLModel: empty expression line");
         const char kind = line[0];
         if (kind == 'n') return add_constant(data, std::strtod(line.c_str() + 1, nullptr));
         if (kind == 'v') {
            const size_t index = static_cast<size_t>(std::strtol(line.c_str() + 1, nullptr, 10));
            if (index >= data.variable_node.size() || data.variable_node[index] < 0) throw std::runtime_error("NLModel: variable index out of range: " + line);
            return data.variable_node[index];
         }
         if (kind != 'o') throw std::runtime_error("NLModel: unsupported expression element '" + line + "' (imported functions and strings are not handled)");
         Node node;
         node.op = static_cast<int>(std::strtol(line.c_str() + 1, nullptr, 10));
         if (is_binary(node.op)) {
            node.a = parse_expression(reader, data);
            node.b = parse_expression(reader, data);
         }
         else if (is_unary(node.op)) {
            node.a = parse_expression(reader, data);
         }
         else if (is_nary(node.op)) {
            const int count = static_cast<int>(reader.numbers().at(0));
            for (int i = 0; i < count; ++i) node.args.push_back(parse_expression(reader, data));
         }
         else {
            throw std::runtime_error("NLModel: operator o" + std::to_string(node.op) + " is not supported");
         }
         return add_node(data, std::move(node));
      }

      // expr + sum_k coefficient_k * variable_k as new nodes (defined variables)
      int add_linear_terms(NLData& data, int expression, const std::vector<std::pair<size_t, double>>& terms) {
         int result = expression;
         for (const auto& [variable, coefficient]: terms) {
            Node product;
            product.op = O_MULT;
            product.a = add_constant(data, coefficient);
            product.b = data.variable_node.at(variable);
            const int term = add_node(data, std::move(product));
            Node sum;
            sum.op = O_PLUS;
            sum.a = result;
            sum.b = term;
            result = add_node(data, std::move(sum));
         }
         return result;
      }

      void read_bounds(Reader& reader, size_t count, std::vector<double>& lower, std::vector<double>& upper, const char* what) {
         const double infinity = std::numeric_limits<double>::infinity();
         lower.assign(count, -infinity);
         upper.assign(count, infinity);
         for (size_t i = 0; i < count; ++i) {
            const std::vector<double> entry = reader.numbers();
            const int code = static_cast<int>(entry.at(0));
            switch (code) {
               case 0: lower[i] = entry.at(1); upper[i] = entry.at(2); break;
               case 1: upper[i] = entry.at(1); break;
               case 2: lower[i] = entry.at(1); break;
               case 3: break;
               case 4: lower[i] = upper[i] = entry.at(1); break;
               default: throw std::runtime_error(std::string("NLModel: unsupported ") + what + " bound type (complementarity constraints are not handled)");
            }
         }
      }

      using IndexSet = std::vector<int>;
      IndexSet set_union(const IndexSet& a, const IndexSet& b) {
         IndexSet result;
         std::set_union(a.begin(), a.end(), b.begin(), b.end(), std::back_inserter(result));
         return result;
      }
      void add_pairs(std::set<std::pair<int, int>>& pairs, const IndexSet& a, const IndexSet& b) {
         for (const int i: a) for (const int j: b) pairs.emplace(std::min(i, j), std::max(i, j));
      }

      // structural sparsity of the Lagrangian Hessian: pairs of variables that meet in a nonlinear operation
      void compute_hessian_pattern(NLData& data) {
         std::vector<IndexSet> depends(data.nodes.size());
         std::set<std::pair<int, int>> pairs;
         for (size_t i = 0; i < data.nodes.size(); ++i) {
            const Node& node = data.nodes[i];
            if (node.op == OP_CONST) continue;
            if (node.op == OP_VAR) { depends[i] = {node.a}; continue; }
            if (is_nary(node.op)) {
               for (const int argument: node.args) depends[i] = set_union(depends[i], depends[argument]);
               continue;
            }
            const IndexSet& u = depends[node.a];
            if (is_unary(node.op)) {
               depends[i] = u;
               if (node.op != O_NEG && node.op != O_ABS) add_pairs(pairs, u, u);
               continue;
            }
            const IndexSet& v = depends[node.b];
            depends[i] = set_union(u, v);
            const bool u_constant = data.nodes[node.a].op == OP_CONST, v_constant = data.nodes[node.b].op == OP_CONST;
            switch (node.op) {
               case O_MULT: if (!u_constant && !v_constant) add_pairs(pairs, u, v); break;
               case O_DIV: if (!v_constant) { add_pairs(pairs, u, v); add_pairs(pairs, v, v); } break;
               case O_POW:
                  if (v_constant) { if (data.nodes[node.b].constant != 1.) add_pairs(pairs, u, u); }
                  else if (u_constant) add_pairs(pairs, v, v);
                  else add_pairs(pairs, depends[i], depends[i]);
                  break;
               default: break;   // plus, minus
            }
         }
         data.hessian_rows.assign(data.n_var, {});
         for (const auto& [row, column]: pairs) data.hessian_rows[static_cast<size_t>(column)].push_back(static_cast<size_t>(row));   // set order: rows ascending
         for (auto& rows: data.hessian_rows) std::sort(rows.begin(), rows.end());
         data.hessian_nonzeros = pairs.size();
      }

      NLData* parse_nl(const std::string& file_name) {
         auto data = std::make_unique<NLData>();
         Reader reader(file_name);
         // ---- header (10 lines) ----
         const std::string& first = reader.next();
         if (first.empty() || first[0] != 'g') throw std::runtime_error("NLModel: only text .nl files (header 'g') are supported");
         const std::vector<double> sizes = reader.numbers();
         data->n_var = static_cast<size_t>(sizes.at(0));
         data->n_con = static_cast<size_t>(sizes.at(1));
         data->n_obj = static_cast<size_t>(sizes.at(2));
         const std::vector<double> nonlinear = reader.numbers();
         data->nlc = static_cast<size_t>(nonlinear.at(0));
         reader.next();   // network constraints
         reader.next();   // nonlinear variables in constraints, objectives, both
         const std::vector<double> misc = reader.numbers();
         if (misc.size() > 1 && misc[1] != 0.) throw std::runtime_error("NLModel: imported functions are not supported");
         for (const double discrete: reader.numbers()) {
            if (discrete != 0.) throw std::runtime_error("NLModel: discrete variables are not supported");   // AMPLModel.cpp:25-28
         }
         reader.next();   // nonzeros in Jacobian, gradients
         reader.next();   // name lengths
         size_t defined_variables = 0;
         for (const double count: reader.numbers()) defined_variables += static_cast<size_t>(count);
         data->has_defined_variables = defined_variables > 0;

         data->variable_node.assign(data->n_var + defined_variables, -1);
         for (size_t i = 0; i < data->n_var; ++i) {
            Node node;
            node.op = OP_VAR;
            node.a = static_cast<int>(i);
            data->variable_node[i] = add_node(*data, std::move(node));
         }
         const double infinity = std::numeric_limits<double>::infinity();
         data->variable_lower.assign(data->n_var, -infinity);
         data->variable_upper.assign(data->n_var, infinity);
         data->constraint_lower.assign(data->n_con, -infinity);
         data->constraint_upper.assign(data->n_con, infinity);
         data->x0.assign(data->n_var, 0.);
         data->pi0.assign(data->n_con, 0.);
         data->constraint_root.assign(data->n_con, -1);
         data->constraint_first.assign(data->n_con, 0);
         data->jacobian.assign(data->n_con, {});

         // ---- segments ----
         while (!reader.done()) {
            const std::string line = reader.next();
            if (line.empty()) continue;
            std::istringstream stream(line.substr(1));
            switch (line[0]) {
               case 'C': {
                  size_t index; stream >> index;
                  data->constraint_first.at(index) = static_cast<int>(data->nodes.size());
                  data->constraint_root.at(index) = parse_expression(reader, *data);
                  break;
               }
               case 'O': {
                  size_t index; int maximize; stream >> index >> maximize;
                  const int first_node = static_cast<int>(data->nodes.size());
                  const int root = parse_expression(reader, *data);
                  if (index == 0) {   // Uno works on the first objective (AMPLModel.cpp:84)
                     data->objective_first = first_node;
                     data->objective_root = root;
                     data->objective_sign = (maximize == 1) ? -1. : 1.;
                  }
                  break;
               }
               case 'V': {
                  size_t index, linear_terms; int kind; stream >> index >> linear_terms >> kind;
                  std::vector<std::pair<size_t, double>> terms;
                  for (size_t t = 0; t < linear_terms; ++t) {
                     const std::vector<double> entry = reader.numbers();
                     terms.emplace_back(static_cast<size_t>(entry.at(0)), entry.at(1));
                  }
                  const int expression = parse_expression(reader, *data);
                  data->variable_node.at(index) = add_linear_terms(*data, expression, terms);
                  break;
               }
               case 'x': case 'd': {
                  size_t count; stream >> count;
                  std::vector<double>& target = (line[0] == 'x') ? data->x0 : data->pi0;
                  for (size_t t = 0; t < count; ++t) {
                     const std::vector<double> entry = reader.numbers();
                     target.at(static_cast<size_t>(entry.at(0))) = entry.at(1);
                  }
                  break;
               }
               case 'r': read_bounds(reader, data->n_con, data->constraint_lower, data->constraint_upper, "constraint"); break;
               case 'b': read_bounds(reader, data->n_var, data->variable_lower, data->variable_upper, "variable"); break;
               case 'k': {
                  size_t count; stream >> count;
                  for (size_t t = 0; t < count; ++t) reader.next();
                  break;
               }
               case 'J': case 'G': {
                  size_t index, count; stream >> index >> count;
                  std::vector<std::pair<size_t, double>> terms;
                  for (size_t t = 0; t < count; ++t) {
                     const std::vector<double> entry = reader.numbers();
                     terms.emplace_back(static_cast<size_t>(entry.at(0)), entry.at(1));
                  }
                  if (line[0] == 'J') data->jacobian.at(index) = std::move(terms);
                  else if (index == 0) data->objective_linear = std::move(terms);
                  break;
               }
               case 'S': {
                  int kind; size_t count; stream >> kind >> count;
                  for (size_t t = 0; t < count; ++t) reader.next();
                  break;
               }
               default:
                  throw std::runtime_error("NLModel: unsupported segment '" + line + "'");
            }
         }
         for (const auto& row: data->jacobian) data->jacobian_nonzeros += row.size();
         compute_hessian_pattern(*data);
         const size_t count = data->nodes.size();
         data->value.assign(count, 0.);
         data->dot.assign(count, 0.);
         data->adjoint.assign(count, 0.);
         data->adjoint_dot.assign(count, 0.);
         data->x.assign(data->n_var, 0.);
         data->column.assign(data->n_var, 0.);
         data->unit.assign(data->n_var, 0.);
         return data.release();
      }
   } // namespace

   int NLData::selected_argument(const Node& node) const {
      int best = node.args.front();
      for (const int argument: node.args) {
         if ((node.op == O_MIN && this->value[argument] < this->value[best]) || (node.op == O_MAX && this->value[argument] > this->value[best])) best = argument;
      }
      return best;
   }

   double NLData::evaluate(const Node& node) const {
      const double u = (node.a >= 0 && node.op != OP_VAR) ? this->value[node.a] : 0.;
      const double v = (node.b >= 0) ? this->value[node.b] : 0.;
      switch (node.op) {
         case OP_CONST: return node.constant;
         case OP_VAR: return this->x[static_cast<size_t>(node.a)];
         case O_PLUS: return u + v;
         case O_MINUS: return u - v;
         case O_MULT: return u * v;
         case O_DIV: return u / v;
         case O_POW: return std::pow(u, v);
         case O_ABS: return std::fabs(u);
         case O_NEG: return -u;
         case O_TANH: return std::tanh(u);
         case O_TAN: return std::tan(u);
         case O_SQRT: return std::sqrt(u);
         case O_SINH: return std::sinh(u);
         case O_SIN: return std::sin(u);
         case O_LOG10: return std::log10(u);
         case O_LOG: return std::log(u);
         case O_EXP: return std::exp(u);
         case O_COSH: return std::cosh(u);
         case O_COS: return std::cos(u);
         case O_ATAN: return std::atan(u);
         case O_ASIN: return std::asin(u);
         case O_ACOS: return std::acos(u);
         case O_SUM: { double sum = 0.; for (const int argument: node.args) sum += this->value[argument]; return sum; }
         case O_MIN: case O_MAX: return this->value[this->selected_argument(node)];
         default: throw std::runtime_error("NLModel: operator not supported");
      }
   }

   Partials NLData::partials(const Node& node) const {
      Partials p;
      const double u = this->value[node.a];
      const double v = (node.b >= 0) ? this->value[node.b] : 0.;
      switch (node.op) {
         case O_PLUS: p.fu = 1.; p.fv = 1.; break;
         case O_MINUS: p.fu = 1.; p.fv = -1.; break;
         case O_MULT: p.fu = v; p.fv = u; p.fuv = 1.; break;
         case O_DIV: p.fu = 1. / v; p.fv = -u / (v * v); p.fuv = -1. / (v * v); p.fvv = 2. * u / (v * v * v); break;
         case O_POW: {
            const bool exponent_constant = this->nodes[node.b].op == OP_CONST, base_constant = this->nodes[node.a].op == OP_CONST;
            if (exponent_constant) {
               // integer exponents stay exact at u = 0 (pow(0, -1) never appears in a derivative that is structurally zero)
               if (v == 2.) { p.fu = 2. * u; p.fuu = 2.; }
               else if (v == 1.) { p.fu = 1.; }
               else { p.fu = v * std::pow(u, v - 1.); p.fuu = v * (v - 1.) * std::pow(u, v - 2.); }
            }
            else {
               const double w = std::pow(u, v), log_u = std::log(u);
               p.fv = w * log_u; p.fvv = w * log_u * log_u;
               if (!base_constant) {
                  p.fu = v * std::pow(u, v - 1.); p.fuu = v * (v - 1.) * std::pow(u, v - 2.);
                  p.fuv = std::pow(u, v - 1.) * (1. + v * log_u);
               }
            }
            break;
         }
         case O_ABS: p.fu = (u >= 0.) ? 1. : -1.; break;
         case O_NEG: p.fu = -1.; break;
         case O_TANH: { const double t = std::tanh(u); p.fu = 1. - t * t; p.fuu = -2. * t * p.fu; break; }
         case O_TAN: { const double t = std::tan(u); p.fu = 1. + t * t; p.fuu = 2. * t * p.fu; break; }
         case O_SQRT: { const double s = std::sqrt(u); p.fu = 0.5 / s; p.fuu = -0.25 / (s * u); break; }
         case O_SINH: p.fu = std::cosh(u); p.fuu = std::sinh(u); break;
         case O_SIN: p.fu = std::cos(u); p.fuu = -std::sin(u); break;
         case O_LOG10: { const double l10 = std::log(10.); p.fu = 1. / (u * l10); p.fuu = -1. / (u * u * l10); break; }
         case O_LOG: p.fu = 1. / u; p.fuu = -1. / (u * u); break;
         case O_EXP: p.fu = std::exp(u); p.fuu = p.fu; break;
         case O_COSH: p.fu = std::sinh(u); p.fuu = std::cosh(u); break;
         case O_COS: p.fu = -std::sin(u); p.fuu = -std::cos(u); break;
         case O_ATAN: { const double d = 1. + u * u; p.fu = 1. / d; p.fuu = -2. * u / (d * d); break; }
         case O_ASIN: { const double d = 1. - u * u; p.fu = 1. / std::sqrt(d); p.fuu = u / (d * std::sqrt(d)); break; }
         case O_ACOS: { const double d = 1. - u * u; p.fu = -1. / std::sqrt(d); p.fuu = -u / (d * std::sqrt(d)); break; }
         default: throw std::runtime_error("NLModel: operator not supported");
      }
      return p;
   }

   // ---- the model -------------------------------------------------------------------------------------------------------------
   NLModel::NLModel(const std::string& file_name) : NLModel(file_name, parse_nl(file_name)) {
   }

   NLModel::NLModel(const std::string& file_name, NLData* data) :
         AnalyticModel(file_name, data->variable_lower, data->variable_upper, data->constraint_lower, data->constraint_upper, data->x0,
            data->nlc, data->objective_sign), data(data) {
   }

   NLModel::~NLModel() {
      delete this->data;
   }

   void NLModel::forward(const double* x) const {
      NLData& d = *this->data;
      if (d.values_current && std::equal(x, x + d.n_var, d.x.begin())) return;   // Uno asks for one constraint gradient at a time
      std::copy_n(x, d.n_var, d.x.begin());
      for (size_t i = 0; i < d.nodes.size(); ++i) d.value[i] = d.evaluate(d.nodes[i]);
      d.values_current = true;
   }

   namespace {
      // reverse sweep of first-order adjoints over nodes [first, root]; adjoints of the touched nodes must be zero on entry
      void reverse_gradient(NLData& d, int first, int root) {
         for (int i = root; i >= first; --i) {
            const double w = d.adjoint[static_cast<size_t>(i)];
            if (w == 0.) continue;
            const Node& node = d.nodes[static_cast<size_t>(i)];
            if (node.op == OP_CONST || node.op == OP_VAR) continue;
            if (node.op == O_SUM) { for (const int argument: node.args) d.adjoint[argument] += w; }
            else if (node.op == O_MIN || node.op == O_MAX) { d.adjoint[d.selected_argument(node)] += w; }
            else {
               const Partials p = d.partials(node);
               d.adjoint[node.a] += w * p.fu;
               if (node.b >= 0) d.adjoint[node.b] += w * p.fv;
            }
         }
      }
   }

   double NLModel::evaluate_objective(const Vector<double>& x) const {
      NLData& d = *this->data;
      this->forward(x.data());
      double result = (d.objective_root >= 0) ? d.value[d.objective_root] : 0.;
      for (const auto& [variable, coefficient]: d.objective_linear) result += coefficient * x[variable];
      return d.objective_sign * result;
   }

   void NLModel::evaluate_objective_gradient(const Vector<double>& x, Vector<double>& gradient) const {
      NLData& d = *this->data;
      this->forward(x.data());
      for (size_t i = 0; i < d.n_var; ++i) gradient[i] = 0.;
      std::fill(d.adjoint.begin(), d.adjoint.end(), 0.);
      if (d.objective_root >= 0) {
         d.adjoint[d.objective_root] = 1.;
         reverse_gradient(d, d.has_defined_variables ? 0 : d.objective_first, d.objective_root);
         for (size_t i = 0; i < d.n_var; ++i) gradient[i] = d.adjoint[i];
      }
      for (const auto& [variable, coefficient]: d.objective_linear) gradient[variable] += coefficient;
      for (size_t i = 0; i < d.n_var; ++i) gradient[i] *= d.objective_sign;
   }

   void NLModel::evaluate_constraints(const Vector<double>& x, std::vector<double>& constraints) const {
      NLData& d = *this->data;
      this->forward(x.data());
      for (size_t j = 0; j < d.n_con; ++j) {
         double result = (d.constraint_root[j] >= 0) ? d.value[d.constraint_root[j]] : 0.;
         for (const auto& [variable, coefficient]: d.jacobian[j]) result += coefficient * x[variable];
         constraints[j] = result;
      }
   }

   void NLModel::evaluate_constraint_gradient(const Vector<double>& x, size_t constraint_index, SparseVector<double>& gradient) const {
      NLData& d = *this->data;
      gradient.clear();
      const int root = d.constraint_root[constraint_index];
      if (root >= 0) {
         this->forward(x.data());
         const int first = d.has_defined_variables ? 0 : d.constraint_first[constraint_index];
         std::fill(d.adjoint.begin() + first, d.adjoint.begin() + root + 1, 0.);
         for (const auto& term: d.jacobian[constraint_index]) d.adjoint[term.first] = 0.;   // every variable of the row is listed in J
         d.adjoint[root] = 1.;
         reverse_gradient(d, first, root);
      }
      for (const auto& [variable, coefficient]: d.jacobian[constraint_index]) {
         gradient.insert(variable, coefficient + ((root >= 0) ? d.adjoint[variable] : 0.));
      }
   }

   // adjoint seeds of the Lagrangian sigma * f - sum_j y_j c_j (AMPLModel.cpp:37-39)
   void NLModel::seed_lagrangian(double objective_multiplier, const Vector<double>& multipliers) const {
      NLData& d = *this->data;
      std::fill(d.adjoint.begin(), d.adjoint.end(), 0.);
      if (d.objective_root >= 0) d.adjoint[d.objective_root] += objective_multiplier * d.objective_sign;
      for (size_t j = 0; j < d.n_con; ++j) {
         if (d.constraint_root[j] >= 0) d.adjoint[d.constraint_root[j]] -= multipliers[j];
      }
   }

   // result = Hessian of the seeded Lagrangian times vector, forward-over-reverse; values must be current, seeds in adjoint_seed
   void NLModel::hessian_vector(const double* vector, double* result) const {
      NLData& d = *this->data;
      const size_t count = d.nodes.size();
      // tangents
      for (size_t i = 0; i < count; ++i) {
         const Node& node = d.nodes[i];
         if (node.op == OP_CONST) d.dot[i] = 0.;
         else if (node.op == OP_VAR) d.dot[i] = vector[node.a];
         else if (node.op == O_SUM) { double sum = 0.; for (const int argument: node.args) sum += d.dot[argument]; d.dot[i] = sum; }
         else if (node.op == O_MIN || node.op == O_MAX) d.dot[i] = d.dot[d.selected_argument(node)];
         else {
            const Partials p = d.partials(node);
            d.dot[i] = p.fu * d.dot[node.a] + ((node.b >= 0) ? p.fv * d.dot[node.b] : 0.);
         }
      }
      // second-order adjoints; first-order adjoints are recomputed from the seeds kept in `column` workspace by the caller
      std::fill(d.adjoint_dot.begin(), d.adjoint_dot.end(), 0.);
      std::vector<double>& adjoint = d.adjoint;
      for (size_t i = count; i-- > d.n_var;) {
         const double w = adjoint[i], wd = d.adjoint_dot[i];
         if (w == 0. && wd == 0.) continue;
         const Node& node = d.nodes[i];
         if (node.op == OP_CONST || node.op == OP_VAR) continue;
         if (node.op == O_SUM) { for (const int argument: node.args) { adjoint[argument] += w; d.adjoint_dot[argument] += wd; } }
         else if (node.op == O_MIN || node.op == O_MAX) { const int argument = d.selected_argument(node); adjoint[argument] += w; d.adjoint_dot[argument] += wd; }
         else {
            const Partials p = d.partials(node);
            const double da = d.dot[node.a], db = (node.b >= 0) ? d.dot[node.b] : 0.;
            adjoint[node.a] += w * p.fu;
            d.adjoint_dot[node.a] += wd * p.fu + w * (p.fuu * da + p.fuv * db);
            if (node.b >= 0) {
               adjoint[node.b] += w * p.fv;
               d.adjoint_dot[node.b] += wd * p.fv + w * (p.fuv * da + p.fvv * db);
            }
         }
      }
      for (size_t i = 0; i < d.n_var; ++i) result[i] = d.adjoint_dot[i];
   }

   void NLModel::evaluate_lagrangian_hessian(const Vector<double>& x, double objective_multiplier, const Vector<double>& multipliers,
         SymmetricMatrix<size_t, double>& hessian) const {
      NLData& d = *this->data;
      this->forward(x.data());
      for (size_t column_index: Range(this->number_variables)) {
         if (!d.hessian_rows[column_index].empty()) {
            this->seed_lagrangian(objective_multiplier, multipliers);
            d.unit[column_index] = 1.;
            this->hessian_vector(d.unit.data(), d.column.data());
            d.unit[column_index] = 0.;
            for (const size_t row_index: d.hessian_rows[column_index]) hessian.insert(row_index, column_index, d.column[row_index]);
         }
         hessian.finalize_column(column_index);
      }
   }

   void NLModel::compute_hessian_vector_product(const double* vector, double objective_multiplier, const Vector<double>& multipliers,
         double* result) const {
      // at the point of the last evaluation, as the ASL does (AMPLModel.cpp:181-189)
      this->seed_lagrangian(objective_multiplier, multipliers);
      this->hessian_vector(vector, result);
   }

   void NLModel::initial_dual_point(Vector<double>& multipliers) const {
      for (size_t constraint_index: Range(this->number_constraints)) multipliers[constraint_index] = this->data->pi0[constraint_index];
   }

   size_t NLModel::number_jacobian_nonzeros() const { return this->data->jacobian_nonzeros; }
   size_t NLModel::number_hessian_nonzeros() const { return this->data->hessian_nonzeros; }
} // namespace
