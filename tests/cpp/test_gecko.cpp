// The reference's C++ test cases (cpp/tests/tools.cpp: test_extend_matrices_body,
// test_extend_matrices_dynamics) against gecko::tools::extend_matrices of
// include/mpci_gecko.hh.  Eigen is not in this image, so the containers are small
// column-major stand-ins with the part of Eigen's interface the routine uses; the
// expected values are recomputed here with plain loops (test code, not product).
//
// It also exercises the two bookkeeping classes of dynamics.hh (check_classes).
//
//   g++ -std=c++17 -I include tests/cpp/test_gecko.cpp -L mpc_interface_b200/csrc -lmpci
//   ./a.out          runs on the GPU      ./a.out --link-only   checks the build only
#include <cmath>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
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

// ControlSystem / ExtendedSystem (cpp/include/mpc-interface/dynamics.hh:34-146): the
// reference has no test for them and leaves from_control_system as a TODO; the expected
// maps are those of the Python classes (dynamics.py:248-267).
static int check_classes() {
  using CS = gecko::tools::ControlSystem<Tensor<3>, Tensor<4>, MatrixXd>;
  using ES = gecko::tools::ExtendedSystem<Tensor<3>, Tensor<4>, MatrixXd>;
  MatrixXd A(3, 3), B(3, 1);
  const double a[3][3] = {{1., 0.1, 0.005}, {0., 1., 0.1}, {0., 0., 1.}};
  const double b[3] = {0.00016667, 0.005, 0.1};
  for (int i = 0; i < 3; ++i) {
    B(i, 0) = b[i];
    for (int j = 0; j < 3; ++j) A(i, j) = a[i][j];
  }
  const std::vector<std::string> inputs = {"CoM_dddot"}, states = {"CoM", "CoM_dot", "CoM_ddot"};
  const std::vector<std::string> axes = {"_x", "_y"};
  int bad = 0;
  CS lip(inputs, states, A, B, axes, CS::Updater(), /*time_variant=*/true);
  ES ext = ES::from_control_system(lip, "x", 9);
  auto expect = [&](const std::map<std::string, int>& map, const char* key, int want) {
    auto it = map.find(key);
    if (it == map.end() || it->second != want) {
      std::printf("classes: %s -> %d expected\n", key, want);
      ++bad;
    }
  };
  expect(ext.domain_ID(), "CoM_dddot_x", 0);
  expect(ext.domain_ID(), "x0_y", 1);
  expect(ext.state_ID(), "CoM_ddot_y", 2);
  expect(ext.domain(), "CoM_dddot_y", 9);
  expect(ext.domain(), "x0_x", 3);
  expect(ext.all_variables(), "CoM_dot_x", 9);
  if (ext.domain_ID().size() != 4 || ext.state_ID().size() != 6 || ext.all_variables().size() != 10 ||
      ext.horizon() != 9) {
    std::printf("classes: wrong map sizes\n");
    ++bad;
  }
  // S(0) = A, and a time-variant system follows new matrices: doubling B doubles U
  Tensor<3> S0 = ext.S();
  Tensor<4> U0 = ext.U();
  double worst = 0.0;
  for (int c = 0; c < 3; ++c)
    for (int i = 0; i < 3; ++i) worst = std::fmax(worst, std::fabs(S0(0, c, i) - A(i, c)));
  MatrixXd B2 = B;
  for (int i = 0; i < 3; ++i) B2(i, 0) = 2.0 * B(i, 0);
  ext.update(A, B2);
  Tensor<4> U1 = ext.U();
  for (std::size_t e = 0; e < U0.v.size(); ++e) worst = std::fmax(worst, std::fabs(U1.v[e] - 2.0 * U0.v[e]));
  if (worst > 1e-12) {
    std::printf("classes: S / U wrong by %.3e\n", worst);
    ++bad;
  }
  try {  // misuse: B does not match the names
    MatrixXd Bw(3, 2);
    CS wrong(inputs, states, A, Bw, axes);
    std::printf("classes: shape check did not throw\n");
    ++bad;
  } catch (const std::invalid_argument&) {
  }
  if (!bad) std::printf("ControlSystem / ExtendedSystem ok\n");
  return bad;
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
  bad += check_classes();
  std::printf(bad ? "FAILED\n" : "gecko::tools::extend_matrices ok\n");
  return bad;
}
