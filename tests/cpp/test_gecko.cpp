// The reference's C++ test cases (cpp/tests/tools.cpp: test_extend_matrices_body,
// test_extend_matrices_dynamics) against gecko::tools::extend_matrices of
// include/mpci_gecko.hh.  Eigen is not in this image, so the containers are small
// column-major stand-ins with the part of Eigen's interface the routine uses; the
// expected values are recomputed here with plain loops (test code, not product).
//
//   g++ -std=c++17 -I include tests/cpp/test_gecko.cpp -L mpc_interface_b200/csrc -lmpci
//   ./a.out          runs on the GPU      ./a.out --link-only   checks the build only
#include <cmath>
#include <cstdio>
#include <cstring>
#include <vector>

#include "mpci_gecko.hh"

struct MatrixXd {  // column-major, like Eigen::MatrixXd
  long r = 0, c = 0;
  std::vector<double> v;
  MatrixXd(long rows, long cols) : r(rows), c(cols), v(rows * cols, 0.0) {}
  long rows() const { return r; }
  long cols() const { return c; }
  double& operator()(long i, long j) { return v[i + j * r]; }
  double operator()(long i, long j) const { return v[i + j * r]; }
};

template <int Rank>
struct Tensor {  // column-major, like Eigen::Tensor<double, Rank>
  long d[4] = {0, 0, 0, 0};
  std::vector<double> v;
  void resize(long a, long b, long c) { d[0] = a; d[1] = b; d[2] = c; d[3] = 1; v.assign(a * b * c, -7.0); }
  void resize(long a, long b, long c, long e) { d[0] = a; d[1] = b; d[2] = c; d[3] = e; v.assign(a * b * c * e, -7.0); }
  long dimension(int i) const { return d[i]; }
  double& operator()(long a, long b, long c) { return v[a + d[0] * (b + d[1] * c)]; }
  double& operator()(long a, long b, long c, long e) { return v[a + d[0] * (b + d[1] * (c + d[2] * e))]; }
};

static MatrixXd matmul(const MatrixXd& X, const MatrixXd& Y) {
  MatrixXd Z(X.rows(), Y.cols());
  for (long i = 0; i < X.rows(); ++i)
    for (long j = 0; j < Y.cols(); ++j) {
      double acc = 0.0;
      for (long l = 0; l < X.cols(); ++l) acc += X(i, l) * Y(l, j);
      Z(i, j) = acc;
    }
  return Z;
}

static int check(const char* name, unsigned N, const MatrixXd& A, const MatrixXd& B) {
  Tensor<3> S;
  Tensor<4> U;
  gecko::tools::extend_matrices(S, U, N, A, B);
  const long n = B.rows(), m = B.cols();
  if (S.dimension(0) != N || S.dimension(1) != n || S.dimension(2) != n || U.dimension(0) != m ||
      U.dimension(1) != N || U.dimension(2) != N || U.dimension(3) != n) {
    std::printf("%s: wrong dimensions\n", name);
    return 1;
  }
  // powers of A and A^d B by repeated left multiplication (tools.cc:100-121)
  std::vector<MatrixXd> AdB;
  AdB.push_back(B);
  for (unsigned d = 1; d < N; ++d) AdB.push_back(matmul(A, AdB.back()));
  MatrixXd Ak = A;
  double worst = 0.0, scale = 0.0;
  for (unsigned k = 0; k < N; ++k) {
    if (k) Ak = matmul(A, Ak);
    for (long c = 0; c < n; ++c)
      for (long i = 0; i < n; ++i) {
        worst = std::fmax(worst, std::fabs(S(k, c, i) - Ak(i, c)));
        scale = std::fmax(scale, std::fabs(Ak(i, c)));
      }
    for (long j = 0; j < m; ++j)
      for (unsigned p = 0; p < N; ++p)
        for (long i = 0; i < n; ++i) {
          const double want = (p <= k) ? AdB[k - p](i, j) : 0.0;
          if (p > k && U(j, k, p, i) != 0.0) {
            std::printf("%s: U above the diagonal is not exactly zero\n", name);
            return 1;
          }
          worst = std::fmax(worst, std::fabs(U(j, k, p, i) - want));
          scale = std::fmax(scale, std::fabs(want));
        }
  }
  const double rel = worst / scale;
  std::printf("%s: N=%u n=%ld m=%ld  relative error %.3e\n", name, N, n, m, rel);
  return rel <= 1e-10 ? 0 : 1;  // north_star tolerance
}

int main(int argc, char** argv) {
  if (argc > 1 && std::strcmp(argv[1], "--link-only") == 0) {
    std::printf("mpci version %d, devices %d\n", mpci_version(), mpci_device_count());
    return mpci_version() >= 100 ? 0 : 1;
  }
  int bad = 0;
  {  // test_extend_matrices_body (cpp/tests/tools.cpp:13-31)
    MatrixXd A(3, 3), B(3, 1);
    const double a[3][3] = {{1., 0.1, 0.005}, {0., 1., 0.1}, {0., 0., 1.}};
    const double b[3] = {0.00016667, 0.005, 0.1};
    for (int i = 0; i < 3; ++i) {
      B(i, 0) = b[i];
      for (int j = 0; j < 3; ++j) A(i, j) = a[i][j];
    }
    bad += check("body", 9, A, B);
  }
  {  // test_extend_matrices_dynamics (cpp/tests/tools.cpp:33-44): 8 x 8 shift + ones row
    MatrixXd A(8, 8), B(8, 6);
    for (int i = 0; i < 7; ++i) A(i, i + 1) = 1.0;
    for (int j = 0; j < 8; ++j) A(7, j) = 1.0;
    for (int i = 0; i < 8; ++i)
      for (int j = 0; j < 6; ++j) B(i, j) = 1.0;
    bad += check("dynamics", 20, A, B);
  }
  try {  // misuse: A not n x n
    MatrixXd A(2, 3), B(3, 1);
    Tensor<3> S;
    Tensor<4> U;
    gecko::tools::extend_matrices(S, U, 4, A, B);
    std::printf("shape check did not throw\n");
    ++bad;
  } catch (const std::invalid_argument&) {
  }
  std::printf(bad ? "FAILED\n" : "gecko::tools::extend_matrices ok\n");
  return bad;
}
