// Dense FP64 building blocks used while an assembly plan is built:
//   * mpci_gemm   -- C = alpha * op(A) @ B + beta * C on the FP64 tensor cores
//                    (mma.sync.aligned.m8n8k4 f64 -> DMMA.8x8x4 in SASS)
//   * mpci_axpy2d -- dst += alpha * rowscale * src with row broadcasting
// They stand where the reference calls NumPy dots: body.py:184-191 (derived
// definitions), :254-255 (limits with L) and :279-292 (V^T C cost products).
#include <cmath>

#include "mpci_common.cuh"

namespace {

constexpr int BM = 64, BN = 64, BK = 16;
constexpr int LDS = 68;  // 68 mod 16 = 4: the 4 x 8 fragment reads hit distinct banks

__device__ __forceinline__ void dmma8x8x4(double& d0, double& d1, double a, double b) {
  asm volatile(
      "mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
      : "+d"(d0), "+d"(d1)
      : "d"(a), "d"(b));
}

// 256 threads = 8 warps laid out 4 (M) x 2 (N); a warp owns 16 x 32 of the
// 64 x 64 block tile = 2 x 4 DMMA tiles.
template <bool TRANS_A>
__global__ void __launch_bounds__(256)
gemm_dmma_kernel(int M, int N, int K, double alpha, const double* __restrict__ A, int lda,
                 long strideA, const double* __restrict__ B, int ldb, long strideB, double beta,
                 double* __restrict__ C, int ldc, long strideC) {
  __shared__ double As[BK][LDS];  // As[k][m] = op(A)[m0 + m][k0 + k]
  __shared__ double Bs[BK][LDS];  // Bs[k][n] = B[k0 + k][n0 + n]
  A += blockIdx.z * strideA;
  B += blockIdx.z * strideB;
  C += blockIdx.z * strideC;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int wm = (warp >> 1) * 16, wn = (warp & 1) * 32;
  const int g = lane >> 2, t = lane & 3;

  double acc[2][4][2];
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

  for (int k0 = 0; k0 < K; k0 += BK) {
    // stage the tiles (zero padded at the borders)
    for (int e = tid; e < BK * BM; e += 256) {
      int k, m;
      if (TRANS_A) { k = e / BM; m = e % BM; } else { m = e / BK; k = e % BK; }
      const int gm = m0 + m, gk = k0 + k;
      double v = 0.0;
      if (gm < M && gk < K) v = TRANS_A ? A[(long)gk * lda + gm] : A[(long)gm * lda + gk];
      As[k][m] = v;
    }
    for (int e = tid; e < BK * BN; e += 256) {
      const int k = e / BN, n = e % BN;
      const int gk = k0 + k, gn = n0 + n;
      Bs[k][n] = (gk < K && gn < N) ? B[(long)gk * ldb + gn] : 0.0;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; kk += 4) {
      double a[2], b[4];
#pragma unroll
      for (int i = 0; i < 2; ++i) a[i] = As[kk + t][wm + 8 * i + g];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = Bs[kk + t][wn + 8 * j + g];
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) dmma8x8x4(acc[i][j][0], acc[i][j][1], a[i], b[j]);
    }
    __syncthreads();
  }
  // epilogue: lane holds C[g][2t], C[g][2t+1] of every 8 x 8 tile
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int row = m0 + wm + 8 * i + g;
      const int col = n0 + wn + 8 * j + 2 * t;
      if (row < M) {
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          if (col + c < N) {
            double* p = C + (long)row * ldc + col + c;
            const double v = alpha * acc[i][j][c];
            *p = (beta == 0.0) ? v : v + beta * *p;
          }
        }
      }
    }
}

__global__ void __launch_bounds__(256)
axpy2d_kernel(int rows, int cols, double alpha, const double* __restrict__ rowscale,
              int n_rowscale, const double* __restrict__ src, int ld_src, int src_rows,
              double* __restrict__ dst, int ld_dst) {
  const long total = (long)rows * cols;
  const long stride = (long)gridDim.x * blockDim.x;
  for (long e = (long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride) {
    const int i = e / cols, j = e % cols;
    double v = alpha * src[(long)(src_rows == 1 ? 0 : i) * ld_src + j];
    if (n_rowscale) v *= rowscale[n_rowscale == 1 ? 0 : i];
    dst[(long)i * ld_dst + j] += v;
  }
}

// out[b] = sum_e (V[b][e] - aim[b * aim_stride + e / rows])^2 : one warp per instance
__global__ void __launch_bounds__(256)
goal_distance_kernel(const double* __restrict__ V, const double* __restrict__ aim, long aim_stride,
                     int rows, int n_axes, double* __restrict__ out, long batch) {
  const int lane = threadIdx.x & 31;
  const long b = (long)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (b >= batch) return;
  const double* v = V + b * (long)rows * n_axes;
  const double* t = aim + b * aim_stride;
  double acc = 0.0;
  for (int a = 0; a < n_axes; ++a) {
    const double ta = t[a];
    for (int r = lane; r < rows; r += 32) {
      const double gap = v[(long)a * rows + r] - ta;
      acc = fma(gap, gap, acc);
    }
  }
  for (int off = 16; off; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
  if (lane == 0) out[b] = acc;
}

}  // namespace

extern "C" int mpci_goal_distance(const double* V, const double* aim, long aim_stride, int rows,
                                  int n_axes, double* out, long batch, void* stream) {
  if (rows < 0 || n_axes < 0 || batch < 0 || aim_stride < 0)
    return mpci::fail(-1, "goal_distance: negative size");
  if (batch == 0) return 0;
  if (!V || !aim || !out) return mpci::fail(-3, "goal_distance: NULL pointer");
  if (aim_stride != 0 && aim_stride < n_axes)
    return mpci::fail(-4, "goal_distance: aim stride %ld smaller than %d axes", aim_stride, n_axes);
  goal_distance_kernel<<<(unsigned)((batch + 7) / 8), 256, 0, mpci::as_stream(stream)>>>(
      V, aim, aim_stride, rows, n_axes, out, batch);
  MPCI_LAUNCH_CHECK("goal_distance_kernel");
  return 0;
}

namespace {
// lb / ub of the variables from the rows of G that bound ONE variable with a coefficient that
// is the same for every instance: thread (b, c) walks the rows of column c.
__global__ void __launch_bounds__(256)
box_bounds_kernel(const double* __restrict__ h, long ld_h, int nv, const int* __restrict__ col_ptr,
                  const int* __restrict__ row, const double* __restrict__ coef,
                  double* __restrict__ lb, double* __restrict__ ub, long batch) {
  const long e = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= batch * nv) return;
  const long b = e / nv;
  const int c = (int)(e - b * nv);
  double lo = -INFINITY, hi = INFINITY;
  for (int x = col_ptr[c]; x < col_ptr[c + 1]; ++x) {
    const double a = coef[x], v = h[b * ld_h + row[x]] / a;  // a z_c <= h
    if (a > 0.0) hi = fmin(hi, v);
    else lo = fmax(lo, v);
  }
  lb[e] = lo;
  ub[e] = hi;
}
}  // namespace

extern "C" int mpci_box_bounds(const double* h, long ld_h, int nv, const int* col_ptr, const int* row,
                               const double* coef, double* lb, double* ub, long batch, void* stream) {
  if (nv < 0 || batch < 0 || ld_h < 0) return mpci::fail(-1, "box_bounds: negative size");
  if (batch == 0 || nv == 0) return 0;
  if (!h || !col_ptr || !lb || !ub) return mpci::fail(-3, "box_bounds: NULL pointer");
  const long total = batch * nv;
  box_bounds_kernel<<<(unsigned)((total + 255) / 256), 256, 0, mpci::as_stream(stream)>>>(
      h, ld_h, nv, col_ptr, row, coef, lb, ub, batch);
  MPCI_LAUNCH_CHECK("box_bounds_kernel");
  return 0;
}

extern "C" int mpci_gemm(int transA, int M, int N, int K, double alpha, const double* A, int lda,
                         long strideA, const double* B, int ldb, long strideB, double beta,
                         double* C, int ldc, long strideC, int batch, void* stream) {
  if (M < 0 || N < 0 || K < 0 || batch < 0)
    return mpci::fail(-1, "gemm: negative size (M=%d N=%d K=%d batch=%d)", M, N, K, batch);
  if (M == 0 || N == 0 || batch == 0) return 0;
  if (!C || (K > 0 && (!A || !B))) return mpci::fail(-3, "gemm: NULL matrix pointer");
  if (ldc < N || ldb < N || lda < (transA ? M : K))
    return mpci::fail(-4, "gemm: leading dimension too small (lda=%d ldb=%d ldc=%d)", lda, ldb,
                      ldc);
  if (batch > 65535) return mpci::fail(-5, "gemm: batch %d exceeds 65535", batch);
  dim3 grid((N + BN - 1) / BN, (M + BM - 1) / BM, batch);
  cudaStream_t st = mpci::as_stream(stream);
  if (transA)
    gemm_dmma_kernel<true><<<grid, 256, 0, st>>>(M, N, K, alpha, A, lda, strideA, B, ldb, strideB,
                                                 beta, C, ldc, strideC);
  else
    gemm_dmma_kernel<false><<<grid, 256, 0, st>>>(M, N, K, alpha, A, lda, strideA, B, ldb,
                                                  strideB, beta, C, ldc, strideC);
  MPCI_LAUNCH_CHECK("gemm_dmma_kernel");
  return 0;
}

extern "C" int mpci_axpy2d(int rows, int cols, double alpha, const double* rowscale,
                           int n_rowscale, const double* src, int ld_src, int src_rows,
                           double* dst, int ld_dst, void* stream) {
  if (rows < 0 || cols < 0) return mpci::fail(-1, "axpy2d: negative size");
  if (rows == 0 || cols == 0) return 0;
  if (!src || !dst) return mpci::fail(-3, "axpy2d: NULL matrix pointer");
  if (src_rows != 1 && src_rows != rows)
    return mpci::fail(-4, "axpy2d: src has %d rows, expected 1 or %d", src_rows, rows);
  if (n_rowscale != 0 && n_rowscale != 1 && n_rowscale != rows)
    return mpci::fail(-4, "axpy2d: rowscale has %d entries, expected 0, 1 or %d", n_rowscale,
                      rows);
  if (n_rowscale && !rowscale) return mpci::fail(-3, "axpy2d: NULL rowscale");
  const long total = (long)rows * cols;
  const int blocks = (int)((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
  axpy2d_kernel<<<blocks, 256, 0, mpci::as_stream(stream)>>>(rows, cols, alpha, rowscale,
                                                             n_rowscale, src, ld_src, src_rows,
                                                             dst, ld_dst);
  MPCI_LAUNCH_CHECK("axpy2d_kernel");
  return 0;
}
