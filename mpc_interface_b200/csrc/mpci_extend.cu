// K1: prediction matrices of x+ = A x + B u over a horizon of N samples,
// for a batch of independent systems.
//
// Replaces tools.extend_matrices (reference python/mpc_interface/tools.py:14-33)
// and gecko::tools::extend_matrices (cpp/src/tools.cc:83-144).  One thread
// block per system keeps the running block [A^(k+1) | A^k B] in shared memory
// and left-multiplies it by A once per horizon step -- the association order
// of the reference's recursion (tools.py:23-29) -- writing the COMPACT impulse
// response (Phi, Gam).  A second, purely streaming kernel expands it into the
// dense reference layouts S (N,n,n) and U (m,N,N,n): the lower-block-Toeplitz
// structure is index arithmetic, so that kernel is bound by the HBM write.
#include "mpci_common.cuh"

namespace {

constexpr int kMaxN = 32;   // states per system
constexpr int kMaxM = 32;   // inputs per system

// Phi[b][k][i][c] = (A^(k+1))[i][c];  Gam[b][d][i][j] = (A^d B)[i][j]
__global__ void __launch_bounds__(128)
extend_compact_kernel(const double* __restrict__ A, const double* __restrict__ B,
                      double* __restrict__ Phi, double* __restrict__ Gam, int n, int m, int N) {
  extern __shared__ double sm[];
  const int w = n + m;              // width of the running block [Phi_k | Gam_k]
  double* As = sm;                  // n x n
  double* cur = As + n * n;         // n x w
  double* nxt = cur + n * w;        // n x w
  const long b = blockIdx.x;
  const double* Ab = A + b * n * n;
  const double* Bb = B + b * n * m;
  for (int e = threadIdx.x; e < n * n; e += blockDim.x) As[e] = Ab[e];
  // step 0: [A | B]
  for (int e = threadIdx.x; e < n * w; e += blockDim.x) {
    const int i = e / w, c = e % w;
    cur[e] = (c < n) ? Ab[i * n + c] : Bb[i * m + (c - n)];
  }
  __syncthreads();
  double* phi = Phi ? Phi + b * (long)N * n * n : nullptr;
  double* gam = Gam ? Gam + b * (long)N * n * m : nullptr;
  for (int k = 0; k < N; ++k) {
    for (int e = threadIdx.x; e < n * w; e += blockDim.x) {
      const int i = e / w, c = e % w;
      const double v = cur[e];
      if (c < n) {
        if (phi) phi[((long)k * n + i) * n + c] = v;
      } else if (gam) {
        gam[((long)k * n + i) * m + (c - n)] = v;
      }
      if (k + 1 < N) {  // next = A @ cur  (tools.py:24-29)
        double acc = 0.0;
        for (int l = 0; l < n; ++l) acc += As[i * n + l] * cur[l * w + c];
        nxt[e] = acc;
      }
    }
    __syncthreads();
    double* t = cur; cur = nxt; nxt = t;
  }
}

// S[b][k][c][i] = Phi[b][k][i][c]    (tools.py:31)
__global__ void __launch_bounds__(256)
expand_S_kernel(const double* __restrict__ Phi, double* __restrict__ S, long total, int n) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long e = (long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride) {
    const int i = e % n;
    const int c = (e / n) % n;
    const long bk = e / ((long)n * n);
    S[e] = Phi[(bk * n + i) * n + c];
  }
}

// U[b][j][k][p][i] = (p <= k) ? Gam[b][k-p][i][j] : 0     (tools.py:32)
// grid: x over rows (b, j, k); each row is N*n contiguous doubles.
__global__ void __launch_bounds__(256)
expand_U_kernel(const double* __restrict__ Gam, double* __restrict__ U, long rows, int n, int m,
                int N) {
  const int row_len = N * n;
  for (long row = blockIdx.x; row < rows; row += gridDim.x) {
    const int k = row % N;
    const int j = (row / N) % m;
    const long b = row / ((long)N * m);
    const double* g = Gam + b * (long)N * n * m;
    double* out = U + row * row_len;
    for (int e = threadIdx.x; e < row_len; e += blockDim.x) {
      const int p = e / n, i = e % n;
      out[e] = (p <= k) ? g[((long)(k - p) * n + i) * m + j] : 0.0;
    }
  }
}

int check_sizes(int batch, int n, int m, int N) {
  if (batch < 0 || n < 1 || m < 1 || N < 1)
    return mpci::fail(-1, "extend_matrices: sizes must be positive (batch=%d n=%d m=%d N=%d)",
                      batch, n, m, N);
  if (n > kMaxN || m > kMaxM)
    return mpci::fail(-2, "extend_matrices: n=%d, m=%d exceed the limits %d, %d", n, m, kMaxN,
                      kMaxM);
  return 0;
}

int launch_compact(const double* A, const double* B, double* Phi, double* Gam, int batch, int n,
                   int m, int N, cudaStream_t st) {
  if (batch == 0) return 0;
  const size_t smem = sizeof(double) * (n * n + 2 * n * (n + m));
  extend_compact_kernel<<<batch, 128, smem, st>>>(A, B, Phi, Gam, n, m, N);
  MPCI_LAUNCH_CHECK("extend_compact_kernel");
  return 0;
}

}  // namespace

extern "C" int mpci_extend_compact(const double* A, const double* B, double* Phi, double* Gam,
                                   int batch, int n, int m, int N, void* stream) {
  if (int rc = check_sizes(batch, n, m, N)) return rc;
  if (!A || !B) return mpci::fail(-3, "extend_compact: A and B must not be NULL");
  return launch_compact(A, B, Phi, Gam, batch, n, m, N, mpci::as_stream(stream));
}

extern "C" int mpci_extend_matrices(const double* A, const double* B, double* S, double* U,
                                    int batch, int n, int m, int N, void* stream) {
  if (int rc = check_sizes(batch, n, m, N)) return rc;
  if (!A || !B) return mpci::fail(-3, "extend_matrices: A and B must not be NULL");
  if (batch == 0) return 0;
  cudaStream_t st = mpci::as_stream(stream);
  double* scratch = nullptr;
  const size_t phi_elems = (size_t)batch * N * n * n, gam_elems = (size_t)batch * N * n * m;
  MPCI_CUDA(cudaMallocAsync(&scratch, sizeof(double) * (phi_elems + gam_elems), st));
  double* Phi = scratch;
  double* Gam = scratch + phi_elems;
  int rc = launch_compact(A, B, S ? Phi : nullptr, U ? Gam : nullptr, batch, n, m, N, st);
  if (rc == 0 && S) {
    const long total = (long)phi_elems;
    const int blocks = (int)((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
    expand_S_kernel<<<blocks, 256, 0, st>>>(Phi, S, total, n);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) rc = mpci::cuda_fail(e, "expand_S_kernel");
  }
  if (rc == 0 && U) {
    const long rows = (long)batch * m * N;
    const int blocks = (int)(rows < 148L * 16 ? rows : 148L * 16);
    expand_U_kernel<<<blocks, 256, 0, st>>>(Gam, U, rows, n, m, N);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) rc = mpci::cuda_fail(e, "expand_U_kernel");
  }
  cudaFreeAsync(scratch, st);
  return rc;
}

extern "C" int mpci_extend_matrices_host(const double* A, const double* B, double* S, double* U,
                                         int batch, int n, int m, int N) {
  if (int rc = check_sizes(batch, n, m, N)) return rc;
  if (!A || !B) return mpci::fail(-3, "extend_matrices_host: A and B must not be NULL");
  if (batch == 0) return 0;
  const size_t a_b = sizeof(double) * batch * n * n, b_b = sizeof(double) * batch * n * m;
  const size_t s_b = sizeof(double) * batch * N * n * n;
  const size_t u_b = sizeof(double) * batch * m * N * N * n;
  double *dA = nullptr, *dB = nullptr, *dS = nullptr, *dU = nullptr;
  int rc = 0;
  cudaError_t e;
  if ((e = cudaMalloc(&dA, a_b)) != cudaSuccess || (e = cudaMalloc(&dB, b_b)) != cudaSuccess ||
      (S && (e = cudaMalloc(&dS, s_b)) != cudaSuccess) ||
      (U && (e = cudaMalloc(&dU, u_b)) != cudaSuccess)) {
    rc = mpci::cuda_fail(e, "cudaMalloc");
  }
  if (rc == 0 && ((e = cudaMemcpy(dA, A, a_b, cudaMemcpyHostToDevice)) != cudaSuccess ||
                  (e = cudaMemcpy(dB, B, b_b, cudaMemcpyHostToDevice)) != cudaSuccess))
    rc = mpci::cuda_fail(e, "cudaMemcpy H2D");
  if (rc == 0) rc = mpci_extend_matrices(dA, dB, dS, dU, batch, n, m, N, nullptr);
  if (rc == 0 && S && (e = cudaMemcpy(S, dS, s_b, cudaMemcpyDeviceToHost)) != cudaSuccess)
    rc = mpci::cuda_fail(e, "cudaMemcpy D2H S");
  if (rc == 0 && U && (e = cudaMemcpy(U, dU, u_b, cudaMemcpyDeviceToHost)) != cudaSuccess)
    rc = mpci::cuda_fail(e, "cudaMemcpy D2H U");
  cudaFree(dA); cudaFree(dB); cudaFree(dS); cudaFree(dU);
  return rc;
}
