// K1: prediction matrices of x+ = A x + B u over a horizon of N samples,
// for a batch of independent systems.
//
// Replaces tools.extend_matrices (reference python/mpc_interface/tools.py:14-33)
// and gecko::tools::extend_matrices (cpp/src/tools.cc:83-144).  One thread
// block per system keeps the running block [A^(k+1) | A^k B] in shared memory
// and left-multiplies it by A once per horizon step -- the association order
// of the reference's recursion (tools.py:23-29).  The dense reference layouts
// S (N,n,n) and U (m,N,N,n) are written by the same block: every row of U is a
// window of a shared-memory array and leaves as one TMA bulk copy
// (extend_fused_kernel).  The compact impulse response (Phi, Gam) and a streaming
// expansion of it remain for the plan compiler and for shapes the bulk copies do
// not fit.
#include <cstdint>
#include <mutex>

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

// ---------------------------------------------------------------------------
// The dense layouts in one pass.  Row (j, k) of U -- N*n contiguous doubles --
// is  [Gam_k[:, j], Gam_{k-1}[:, j], ..., Gam_0[:, j], 0, ..., 0]:  with
//     Z_j[y*n + i] = Gam_{N-1-y}[i][j]  (y < N),   0  (N <= y < 2N)
// in shared memory it is the window Z_j[(N-1-k)*n ...] of length N*n, so the block
// that ran the recursion hands every row to the TMA engine as ONE bulk copy
// (cp.async.bulk shared -> global); a second copy of Z shifted by one element keeps
// the source 16-byte aligned when (N-1-k)*n is odd.  S leaves with plain stores while
// the recursion runs (it is n / (m N) of the bytes).  Needs N*n even and a 16-byte
// aligned U; other shapes take the three-kernel path above.
// ---------------------------------------------------------------------------
constexpr int kFusedThreads = 128;

__device__ __forceinline__ void bulk_row(double* dst, const double* src_smem, unsigned bytes) {
  const unsigned src = (unsigned)__cvta_generic_to_shared(src_smem);
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;\n"
               :: "l"(dst), "r"(src), "r"(bytes) : "memory");
}

__host__ __device__ inline size_t fused_smem_doubles(int n, int m, int N, bool shifted) {
  const size_t mats = ((size_t)n * n + 2 * (size_t)n * (n + m) + 1) & ~(size_t)1;
  return mats + (size_t)m * 2 * N * n * (shifted ? 2 : 1);
}

__global__ void __launch_bounds__(kFusedThreads)
extend_fused_kernel(const double* __restrict__ A, const double* __restrict__ B,
                    double* __restrict__ S, double* __restrict__ U, long batch, int n, int m, int N,
                    int shifted) {
  extern __shared__ double sm[];
  const int w = n + m, tid = threadIdx.x;
  const int zld = 2 * N * n, row_len = N * n;
  double* As = sm;
  double* cur = As + n * n;
  double* nxt = cur + n * w;
  double* Z = sm + (((size_t)n * n + 2 * (size_t)n * w + 1) & ~(size_t)1);  // [m][zld]
  double* Z1 = Z + (size_t)m * zld;                                        // Z1[x] = Z[x + 1]
  // the zero halves are written once; the recursion only touches [0, N*n)
  for (int e = tid; e < m * zld; e += blockDim.x) {
    const int x = e % zld;
    if (x >= row_len) Z[e] = 0.0;
    if (shifted && x >= row_len - 1) Z1[e] = 0.0;
  }
  for (long b = blockIdx.x; b < batch; b += gridDim.x) {
    // the rows of the previous system have left Z (bulk groups are per thread)
    asm volatile("cp.async.bulk.wait_group.read 0;\n" ::: "memory");
    __syncthreads();
    const double* Ab = A + b * n * n;
    const double* Bb = B + b * n * m;
    for (int e = tid; e < n * n; e += blockDim.x) As[e] = Ab[e];
    for (int e = tid; e < n * w; e += blockDim.x) {
      const int i = e / w, c = e - i * w;
      cur[e] = (c < n) ? Ab[i * n + c] : Bb[i * m + (c - n)];
    }
    __syncthreads();
    double* Sb = S ? S + b * (long)N * n * n : nullptr;
    for (int k = 0; k < N; ++k) {
      for (int e = tid; e < n * w; e += blockDim.x) {
        const int i = e / w, c = e - i * w;
        const double v = cur[e];
        if (c < n) {
          if (Sb) Sb[((long)k * n + c) * n + i] = v;  // S[k][c][i] = (A^(k+1))[i][c]
        } else {
          const int x = (c - n) * zld + (N - 1 - k) * n + i;
          Z[x] = v;
          if (shifted && ((N - 1 - k) * n + i) > 0) Z1[x - 1] = v;
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
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
    __syncthreads();
    double* Ub = U + b * (long)m * N * row_len;
    for (int r = tid; r < m * N; r += blockDim.x) {
      const int j = r / N, k = r - j * N;
      const int start = (N - 1 - k) * n;
      const double* src = (start & 1) ? Z1 + (size_t)j * zld + (start - 1) : Z + (size_t)j * zld + start;
      bulk_row(Ub + (long)r * row_len, src, (unsigned)row_len * 8u);
    }
    asm volatile("cp.async.bulk.commit_group;\n" ::: "memory");
  }
  asm volatile("cp.async.bulk.wait_group 0;\n" ::: "memory");
}

// ---------------------------------------------------------------------------
// Small systems (the LIP: n = 3, m = 1).  At the reference's default horizon N = 24 the rows
// are 576 bytes and the TMA engine takes a bulk copy every ~20 ns whatever its size: 3.5 TB/s
// (0.53 of the copy peak, profiles/r01_tma_rate.txt).  Measured here: 4.74 TB/s (0.72) at
// N = 24 and 5.7 TB/s (0.87, from 5.3 with bulk rows) at N = 100; taken whenever eight systems
// fit a block's shared memory.  (The same plain stores in the block-per-system kernel gave
// nothing at n = 12, m = 4, N = 50 or n = 6, m = 2, N = 400: 4.95 / 5.28 against 5.19 / 5.35 TB/s.)
// Here a WARP owns a system -- no block-wide barrier anywhere: one lane per element of the
// running block [A^(k+1) | A^k B] (n (n + m) <= 32), S collected in shared memory, and both
// outputs leave as flat streams of 16-byte stores (U[j] is N rows of N n doubles, each a window
// of Z_j; the rows are contiguous in memory, so the warp walks the whole of U[j] 512 bytes at
// a time).
// ---------------------------------------------------------------------------
constexpr int kSmallWarps = 8;
constexpr int kSmallMaxN = 5;  // n (n + m) <= 32 with m >= 1

__host__ __device__ inline size_t small_warp_doubles(int n, int m, int N) {
  return (size_t)m * 2 * N * n + (((size_t)N * n * n + 1) & ~(size_t)1);
}

#ifndef MPCI_SMALL_MINBLOCKS
#define MPCI_SMALL_MINBLOCKS 4
#endif
__global__ void __launch_bounds__(32 * kSmallWarps, MPCI_SMALL_MINBLOCKS)
extend_small_kernel(const double* __restrict__ A, const double* __restrict__ B, double* __restrict__ S,
                    double* __restrict__ U, long batch, int n, int m, int N) {
  extern __shared__ double sm[];
  const int w = n + m, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int zld = 2 * N * n, row_len = N * n;
  double* Z = sm + (size_t)warp * small_warp_doubles(n, m, N);  // [m][zld]
  double* Sb = Z + (size_t)m * zld;                             // [N][n][n]
  for (int e = lane; e < m * zld; e += 32)
    if (e % zld >= row_len) Z[e] = 0.0;  // the zero halves, once
  const int i = lane / w, c = lane - i * w;  // this lane's element of the running block
  const bool mine = lane < n * w;
  for (long b = (long)blockIdx.x * kSmallWarps + warp; b < batch; b += (long)gridDim.x * kSmallWarps) {
    __syncwarp();  // the previous system has left shared memory
    const double* Ab = A + b * n * n;
    const double* Bb = B + b * n * m;
    // the running block lives in registers, one element per lane: row i of A beside it, the
    // column of the block through shuffles (n <= 5 here) -- the recursion touches shared memory
    // only to park its results
    double arow[kSmallMaxN];
#pragma unroll
    for (int l = 0; l < kSmallMaxN; ++l) arow[l] = (mine && l < n) ? Ab[i * n + l] : 0.0;
    double v = mine ? ((c < n) ? Ab[i * n + c] : Bb[i * m + (c - n)]) : 0.0;
    const int col = mine ? c : 0;
    for (int k = 0; k < N; ++k) {
      if (mine) {
        if (c < n) Sb[(k * n + c) * n + i] = v;                   // S[k][c][i] = (A^(k+1))[i][c]
        else Z[(c - n) * zld + (N - 1 - k) * n + i] = v;
      }
      double acc = 0.0;  // next = A @ cur  (tools.py:24-29; the last one is not used)
#pragma unroll
      for (int l = 0; l < kSmallMaxN; ++l) {
        const double below = __shfl_sync(0xffffffffu, v, l * w + col);
        if (l < n) acc += arow[l] * below;
      }
      v = acc;
    }
    __syncwarp();
    if (S) {
      double* Sg = S + b * (long)N * n * n;  // N n n doubles; 16-byte aligned when N n n is even
      const int total = N * n * n;
      if ((total & 1) == 0 && (reinterpret_cast<uintptr_t>(S) & 15) == 0) {
        for (int e = 2 * lane; e < total; e += 64)
          __stcs(reinterpret_cast<double2*>(Sg + e), make_double2(Sb[e], Sb[e + 1]));
      } else {
        for (int e = lane; e < total; e += 32) __stcs(Sg + e, Sb[e]);
      }
    }
    for (int j = 0; j < m; ++j) {  // U[j]: flat index f = k row_len + e  <-  Z_j[(N-1-k) n + e]
      double* Ug = U + (b * m + j) * (long)N * row_len;
      const double* Zj = Z + (size_t)j * zld;
      int e = 2 * lane, k = 0;
      while (e >= row_len) { e -= row_len; ++k; }
      for (; k < N;) {
        const double* src = Zj + (N - 1 - k) * n + e;
        __stcs(reinterpret_cast<double2*>(Ug + (long)k * row_len + e), make_double2(src[0], src[1]));
        e += 64;
        while (e >= row_len) { e -= row_len; ++k; }
      }
    }
  }
}

// *launched stays false when the shape is not eligible (the caller then takes the
// three-kernel path); the return value is the usual status.
int launch_fused(const double* A, const double* B, double* S, double* U, int batch, int n, int m,
                 int N, cudaStream_t st, bool* launched) {
  *launched = false;
  if (!U || ((long)N * n) % 2 != 0 || (reinterpret_cast<uintptr_t>(U) & 15) != 0) return 0;
  if (n * (n + m) <= 32) {  // small systems whose data fits eight to a block: a warp per system, plain stores
    const size_t bytes = sizeof(double) * small_warp_doubles(n, m, N) * kSmallWarps;
    if (bytes <= 96 * 1024) {
      if (bytes > 48 * 1024)
        MPCI_CUDA(cudaFuncSetAttribute(extend_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
      int dev = 0, sms = 148, per_sm = 1;
      MPCI_CUDA(cudaGetDevice(&dev));
      MPCI_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
      MPCI_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, extend_small_kernel, 32 * kSmallWarps, bytes));
#ifdef MPCI_SMALL_PER_SM
      if (per_sm > MPCI_SMALL_PER_SM) per_sm = MPCI_SMALL_PER_SM;
#endif
      if (per_sm >= 1) {
        const long want = ((long)batch + kSmallWarps - 1) / kSmallWarps, resident = (long)sms * per_sm;
        extend_small_kernel<<<(unsigned)(want < resident ? want : resident), 32 * kSmallWarps, bytes, st>>>(
            A, B, S, U, batch, n, m, N);
        MPCI_LAUNCH_CHECK("extend_small_kernel");
        *launched = true;
        return 0;
      }
    }
  }
  const bool shifted = (n % 2) != 0;
  const size_t smem = sizeof(double) * fused_smem_doubles(n, m, N, shifted);
  if (smem > 227 * 1024) return 0;
  if (smem > 48 * 1024)
    MPCI_CUDA(cudaFuncSetAttribute(extend_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)smem));
  int dev = 0, sms = 148, per_sm = 1;
  MPCI_CUDA(cudaGetDevice(&dev));
  MPCI_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  // small systems: one warp per block, so that up to 32 systems are in flight on an SM
  // (measured at n=3, m=1: N=100 74 -> 81 %, N=24 36 -> 53 % of the HBM peak); the larger
  // ones need the threads for the recursion and the row copies (n=12, m=4, N=50: 79 %
  // with 128 threads, 62 % with 32)
  const int threads = (n * (n + m) <= 32 && m * N <= 128) ? 32 : kFusedThreads;
  MPCI_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, extend_fused_kernel, threads, smem));
  if (per_sm < 1) return 0;
  const long resident = (long)sms * per_sm;
  const unsigned blocks = (unsigned)(batch < resident ? batch : resident);
  extend_fused_kernel<<<blocks, threads, smem, st>>>(A, B, S, U, batch, n, m, N, shifted ? 1 : 0);
  MPCI_LAUNCH_CHECK("extend_fused_kernel");
  *launched = true;
  return 0;
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
  {
    bool launched = false;
    const int rc = launch_fused(A, B, S, U, batch, n, m, N, st, &launched);
    if (rc != 0 || launched) return rc;
  }
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

// ---------------------------------------------------------------------------
// K1 without any allocation (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA graph capture, tight serving
// loops, harnesses that forbid it).  The three-kernel path above -- compact recursion into
// (Phi, Gam), then the two expansions -- with (Phi, Gam) in a file-scope device array instead
// of a cudaMallocAsync'd buffer.  Every GPU that loads the library holds the array until the
// process exits; every call on a device shares it, so calls that overlap on that GPU race.
// Nothing in it has to start at zero: the recursion writes every element the expansions read.
// ---------------------------------------------------------------------------
namespace {

constexpr size_t kStaticScratchDoubles = MPCI_EXTEND_STATIC_MAX_DOUBLES;  // 4 MiB
static __device__ __align__(256) double extend_static_scratch[kStaticScratchDoubles];

constexpr int kMaxDevices = 64;
std::mutex static_scratch_lock;
double* static_scratch_addr[kMaxDevices];  // device address of the array, per device (set-up once)

// The array's name is a host-side stand-in here, not a device address: the kernels get what
// cudaGetSymbolAddress returns for the CURRENT device.
int static_scratch(double** out) {
  int dev = 0;
  MPCI_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return mpci::fail(-2, "extend_matrices_static: device %d out of range", dev);
  std::lock_guard<std::mutex> hold(static_scratch_lock);
  if (!static_scratch_addr[dev]) {
    double* ptr = nullptr;
    MPCI_CUDA(cudaGetSymbolAddress((void**)&ptr, extend_static_scratch));
    static_scratch_addr[dev] = ptr;
  }
  *out = static_scratch_addr[dev];
  return 0;
}

}  // namespace

extern "C" long mpci_extend_static_capacity(void) { return (long)kStaticScratchDoubles; }

extern "C" int mpci_extend_matrices_static(const double* A, const double* B, double* S, double* U,
                                           int batch, int n, int m, int N, void* stream) {
  if (int rc = check_sizes(batch, n, m, N)) return rc;
  if (!A || !B) return mpci::fail(-3, "extend_matrices_static: A and B must not be NULL");
  if (batch == 0) return 0;
  const size_t phi_elems = S ? (size_t)batch * N * n * n : 0, gam_elems = U ? (size_t)batch * N * n * m : 0;
  if (phi_elems + gam_elems > kStaticScratchDoubles)
    return mpci::fail(-8, "extend_matrices_static: batch=%d n=%d m=%d N=%d needs %zu doubles of scratch, the "
                      "static array holds %zu (use mpci_extend_matrices, or split the batch)", batch, n, m,
                      N, phi_elems + gam_elems, kStaticScratchDoubles);
  double* scratch = nullptr;
  if (int rc = static_scratch(&scratch)) return rc;
  cudaStream_t st = mpci::as_stream(stream);
  double* Phi = scratch;
  double* Gam = scratch + phi_elems;
  if (int rc = launch_compact(A, B, S ? Phi : nullptr, U ? Gam : nullptr, batch, n, m, N, st)) return rc;
  if (S) {
    const long total = (long)phi_elems;
    const int blocks = (int)((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
    expand_S_kernel<<<blocks, 256, 0, st>>>(Phi, S, total, n);
    MPCI_LAUNCH_CHECK("expand_S_kernel");
  }
  if (U) {
    const long rows = (long)batch * m * N;
    const int blocks = (int)(rows < 148L * 16 ? rows : 148L * 16);
    expand_U_kernel<<<blocks, 256, 0, st>>>(Gam, U, rows, n, m, N);
    MPCI_LAUNCH_CHECK("expand_U_kernel");
  }
  return 0;
}
