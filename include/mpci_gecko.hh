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
#ifndef MPCI_GECKO_HH
#define MPCI_GECKO_HH

#include <cstddef>
#include <stdexcept>
#include <string>
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

}  // namespace tools
}  // namespace gecko

#endif  // MPCI_GECKO_HH
