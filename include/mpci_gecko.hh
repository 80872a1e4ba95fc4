// mpci_gecko.hh -- C++ face of the prediction-matrix kernel (header only).
//
// The reference's one working native routine is
//     gecko::tools::extend_matrices(Eigen::Tensor<double,3>& S, Eigen::Tensor<double,4>& U,
//                                   unsigned N, Eigen::MatrixXd& A, Eigen::MatrixXd& B)
// (cpp/include/mpc-interface/tools.hh:24-25, cpp/src/tools.cc:83-144): it resizes
// S to (N, n, n) and U to (m, N, N, n) and fills
//     S(k, c, i)    = (A^(k+1))(i, c)
//     U(j, k, p, i) = (A^(k-p) B)(i, j)   for p <= k, 0 above the diagonal.
// This header gives the same function, same name and argument order, on top of the
// C ABI of mpci.h (the work is done by the CUDA kernel K1 behind
// mpci_extend_matrices_host).  It is a template over the container types so that it
// does not need Eigen to compile; it uses only the part of Eigen's interface the
// reference itself uses:
//     Matrix:  rows(), cols(), operator()(i, j)
//     Tensor3: resize(d0, d1, d2), operator()(i0, i1, i2)
//     Tensor4: resize(d0, d1, d2, d3), operator()(i0, i1, i2, i3)
// so Eigen::MatrixXd / Eigen::Tensor<double, 3 | 4> (any storage order) drop in.
// Errors: the reference has no error path (Eigen asserts); here a failed call throws
// std::runtime_error carrying mpci_last_error().  There is no CPU fallback.
//
// The reference also declares two bookkeeping classes, gecko::tools::ControlSystem and
// gecko::tools::ExtendedSystem (cpp/include/mpc-interface/dynamics.hh:34-146,
// cpp/src/dynamics.cc); its ExtendedSystem::from_control_system is a commented-out TODO
// (dynamics.cc:63-75) and update_sizes is empty.  Below they are given with the same
// constructor arguments, completed the way the Python classes work
// (python/mpc_interface/dynamics.py:216-267): from_control_system builds S and U with
// extend_matrices above, identify_domain names input j "<input><axis>" -> j and the
// initial state "<state vector>0<axis>" -> number of inputs, states "<state><axis>" -> i,
// set_sizes gives every domain variable its length (N predicted actions per input, n for
// the initial state) and every state N; update(A, B) re-extends a time-variant system.
#ifndef MPCI_GECKO_HH
#define MPCI_GECKO_HH

#include <cstddef>
#include <functional>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "mpci.h"

namespace gecko {
namespace tools {

template <class Tensor3, class Tensor4, class Matrix>
void extend_matrices(Tensor3& S, Tensor4& U, unsigned int N, const Matrix& A, const Matrix& B) {
  const long n = static_cast<long>(B.rows());
  const long m = static_cast<long>(B.cols());
  if (static_cast<long>(A.rows()) != n || static_cast<long>(A.cols()) != n)
    throw std::invalid_argument("extend_matrices: A must be n x n with n = B.rows()");
  if (N == 0 || n == 0 || m == 0)
    throw std::invalid_argument("extend_matrices: empty horizon or system");

  // the C ABI takes row-major matrices (the layout of the reference's NumPy arrays)
  std::vector<double> a(static_cast<std::size_t>(n * n)), b(static_cast<std::size_t>(n * m));
  for (long i = 0; i < n; ++i) {
    for (long c = 0; c < n; ++c) a[static_cast<std::size_t>(i * n + c)] = A(i, c);
    for (long j = 0; j < m; ++j) b[static_cast<std::size_t>(i * m + j)] = B(i, j);
  }
  const long NN = static_cast<long>(N);
  std::vector<double> s(static_cast<std::size_t>(NN * n * n));
  std::vector<double> u(static_cast<std::size_t>(m * NN * NN * n));
  const int rc = mpci_extend_matrices_host(a.data(), b.data(), s.data(), u.data(), 1,
                                           static_cast<int>(n), static_cast<int>(m),
                                           static_cast<int>(N));
  if (rc != 0)
    throw std::runtime_error(std::string("extend_matrices: ") + mpci_last_error());

  // hand the result over in the caller's containers (index order of tools.cc:103-143)
  S.resize(NN, n, n);
  for (long k = 0; k < NN; ++k)
    for (long c = 0; c < n; ++c)
      for (long i = 0; i < n; ++i) S(k, c, i) = s[static_cast<std::size_t>((k * n + c) * n + i)];
  U.resize(m, NN, NN, n);
  for (long j = 0; j < m; ++j)
    for (long k = 0; k < NN; ++k)
      for (long p = 0; p < NN; ++p)
        for (long i = 0; i < n; ++i)
          U(j, k, p, i) = u[static_cast<std::size_t>(((j * NN + k) * NN + p) * n + i)];
}

template <class Tensor3, class Tensor4, class Matrix>
class ExtendedSystem;

// dynamics.hh:34-70.  The update callback receives the extended system to refresh and the
// integer arguments of the tick (step counts and the like).
template <class Tensor3, class Tensor4, class Matrix>
class ControlSystem {
 public:
  using Extended = ExtendedSystem<Tensor3, Tensor4, Matrix>;
  using Updater = std::function<void(std::shared_ptr<Extended>, std::map<std::string, int>&)>;

  ControlSystem(const std::vector<std::string>& input_names,
                const std::vector<std::string>& state_names, const Matrix& A, const Matrix& B,
                const std::vector<std::string>& axes, Updater how_to_update_matrices = Updater(),
                bool time_variant = false)
      : input_names_(input_names), state_names_(state_names), A_(A), B_(B), axes_(axes),
        time_variant_(time_variant), how_to_update_matrices_(std::move(how_to_update_matrices)) {
    if (static_cast<std::size_t>(B.rows()) != state_names.size() ||
        static_cast<std::size_t>(B.cols()) != input_names.size())
      throw std::invalid_argument("ControlSystem: B must be (states x inputs)");
    if (A.rows() != B.rows() || A.cols() != B.rows())
      throw std::invalid_argument("ControlSystem: A must be (states x states)");
  }

  void update_matrices(std::shared_ptr<Extended> extended, std::map<std::string, int>& kargs) {
    if (how_to_update_matrices_) how_to_update_matrices_(std::move(extended), kargs);
  }

  const std::vector<std::string>& input_names() const { return input_names_; }
  const std::vector<std::string>& state_names() const { return state_names_; }
  const std::vector<std::string>& axes() const { return axes_; }
  const Matrix& A() const { return A_; }
  const Matrix& B() const { return B_; }
  Matrix& A() { return A_; }
  Matrix& B() { return B_; }
  bool time_variant() const { return time_variant_; }

 protected:
  std::vector<std::string> input_names_, state_names_;
  Matrix A_, B_;
  std::vector<std::string> axes_;
  bool time_variant_;
  Updater how_to_update_matrices_;
};

// dynamics.hh:83-146: x = S x0 + U u over the horizon, S [N, n, n], U [m, N, N, n].
template <class Tensor3, class Tensor4, class Matrix>
class ExtendedSystem {
 public:
  using Extender = std::function<void(Tensor3&, Tensor4&, unsigned int, const Matrix&, const Matrix&)>;

  ExtendedSystem(const std::vector<std::string>& input_names,
                 const std::vector<std::string>& state_names, const std::string& state_vector_name,
                 Tensor3 S, Tensor4 U, const std::vector<std::string>& axes,
                 Extender how_to_update_ext_matrices = Extender(), bool time_variant = true)
      : input_names_(input_names), state_names_(state_names), state_vector_name_(state_vector_name),
        S_(std::move(S)), U_(std::move(U)), axes_(axes), time_variant_(time_variant),
        how_to_update_ext_matrices_(std::move(how_to_update_ext_matrices)) {
    if (static_cast<std::size_t>(U_.dimension(0)) != input_names.size() ||
        static_cast<std::size_t>(S_.dimension(1)) != state_names.size() ||
        U_.dimension(1) != S_.dimension(0) || U_.dimension(3) != S_.dimension(1))
      throw std::invalid_argument("ExtendedSystem: S must be [N, n, n] and U [m, N, N, n]");
    identify_domain(input_names_, state_names_);
    set_sizes();
  }

  // The constructor the reference left as a TODO (dynamics.cc:63-75; Python:
  // dynamics.py:216-246): S and U come from the prediction-matrix kernel.
  static ExtendedSystem from_control_system(const ControlSystem<Tensor3, Tensor4, Matrix>& system,
                                            const std::string& state_vector_name,
                                            unsigned int horizon_length) {
    Tensor3 S;
    Tensor4 U;
    extend_matrices(S, U, horizon_length, system.A(), system.B());
    Extender again = [](Tensor3& s, Tensor4& u, unsigned int N, const Matrix& A, const Matrix& B) {
      extend_matrices(s, u, N, A, B);
    };
    return ExtendedSystem(system.input_names(), system.state_names(), state_vector_name, std::move(S),
                          std::move(U), system.axes(), system.time_variant() ? again : Extender(),
                          system.time_variant());
  }

  // dynamics.cc:77-105 (an empty axis list stands for the single unnamed axis)
  void identify_domain(const std::vector<std::string>& input_names,
                       const std::vector<std::string>& state_names) {
    domain_ID_.clear();
    state_ID_.clear();
    const std::vector<std::string> unnamed(1, std::string());
    for (const std::string& axis : axes_.empty() ? unnamed : axes_) {
      for (std::size_t j = 0; j < input_names.size(); ++j)
        domain_ID_[input_names[j] + axis] = static_cast<int>(j);
      domain_ID_[state_vector_name_ + "0" + axis] = static_cast<int>(input_names.size());
      for (std::size_t i = 0; i < state_names.size(); ++i)
        state_ID_[state_names[i] + axis] = static_cast<int>(i);
    }
  }

  // Refreshes S and U of a time-variant system from new (A, B) (dynamics.py:340-344).
  void update(const Matrix& A, const Matrix& B) {
    if (!time_variant_ || !how_to_update_ext_matrices_) return;
    how_to_update_ext_matrices_(S_, U_, static_cast<unsigned int>(S_.dimension(0)), A, B);
    update_sizes();
  }

  const Tensor3& S() const { return S_; }
  const Tensor4& U() const { return U_; }
  const std::map<std::string, int>& domain_ID() const { return domain_ID_; }
  const std::map<std::string, int>& state_ID() const { return state_ID_; }
  // length of every variable: domain variables (inputs, initial state) and states
  const std::map<std::string, int>& all_variables() const { return all_variables_; }
  const std::map<std::string, int>& domain() const { return domain_; }
  const std::string& state_vector_name() const { return state_vector_name_; }
  const std::vector<std::string>& axes() const { return axes_; }
  bool time_variant() const { return time_variant_; }
  long horizon() const { return static_cast<long>(S_.dimension(0)); }

 protected:
  void set_sizes() {  // dynamics.py:260-267
    const int n_inputs = static_cast<int>(input_names_.size());
    domain_.clear();
    all_variables_.clear();
    for (const auto& kv : domain_ID_)
      domain_[kv.first] = static_cast<int>(kv.second < n_inputs ? U_.dimension(2) : S_.dimension(1));
    all_variables_ = domain_;
    for (const auto& kv : state_ID_) all_variables_[kv.first] = static_cast<int>(S_.dimension(0));
  }
  void update_sizes() {
    if (time_variant_) set_sizes();
  }

  std::vector<std::string> input_names_, state_names_;
  std::string state_vector_name_;
  Tensor3 S_;
  Tensor4 U_;
  std::vector<std::string> axes_;
  bool time_variant_;
  Extender how_to_update_ext_matrices_;
  std::map<std::string, int> domain_ID_, state_ID_, domain_, all_variables_;
};

}  // namespace tools
}  // namespace gecko

#endif  // MPCI_GECKO_HH
