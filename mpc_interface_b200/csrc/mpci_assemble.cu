// The batched QP assembly kernels (the hot path).
//
// For every instance b of the batch (reference: Formulation.
// generate_all_qp_matrices, python/mpc_interface/body.py:333-348, and under it
// generate_qp_cost :266-302 / generate_qp_constraint :236-264):
//
//   c_m(b)    = prod_f scalars[b][mono[m][f]]
//   P[b]      = sum_t c_{Pt}(b) * DP[t]          streamed, nv*nv doubles
//   G[b]      = sum_t c_{Gt}(b) * DG[t]          streamed, nc*nv doubles
//   [q;h][b]  = sum_t c_{Yt}(b) * DY[t] @ [given[b]; 1]
//
// stream_terms_kernel is bound by the HBM write of P and G (99.5 % of the
// output bytes): constants stay in registers across a group of instances, every
// thread issues 16-byte streaming stores, a warp covers 512 contiguous bytes.
// qh_kernel is a small FP64 product that skips the all-zero blocks of DY.
#include <cstdint>

#include "mpci_common.cuh"
#include "mpci_plan.cuh"

namespace mpci {

// ---------------------------------------------------------------------------
// P / G streams
// ---------------------------------------------------------------------------
constexpr int kStreamThreads = 256;
constexpr int kStreamGroup = 32;  // instances that share one load of the constants

template <int K, typename VecT>
__global__ void __launch_bounds__(kStreamThreads)
stream_terms_kernel(StreamArgs args, const double* __restrict__ scalars, int n_scalars,
                    double* __restrict__ out, long batch) {
  constexpr int V = sizeof(VecT) / sizeof(double);
  __shared__ double coef[kStreamGroup][K];
  const long b0 = (long)blockIdx.y * kStreamGroup;
  const int group = (int)min((long)kStreamGroup, batch - b0);

  for (int e = threadIdx.x; e < group * K; e += blockDim.x) {
    const int t = e / K, k = e % K;
    const double* s = scalars + (b0 + t) * n_scalars;
    double c = 1.0;
#pragma unroll
    for (int f = 0; f < MPCI_MAX_FACTORS; ++f) {
      const int idx = args.factors[k][f];
      if (idx >= 0) c *= s[idx];
    }
    coef[t][k] = c;
  }
  __syncthreads();

  const long n_vec = args.n_elems / V;
  const long v = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= n_vec) return;

  VecT d[K];
#pragma unroll
  for (int k = 0; k < K; ++k)
    d[k] = reinterpret_cast<const VecT*>(args.consts + (long)k * args.n_elems)[v];

  VecT* dst = reinterpret_cast<VecT*>(out + b0 * args.n_elems) + v;
  const long step = args.n_elems / V;
  for (int t = 0; t < group; ++t) {
    VecT r;
    if constexpr (V == 2) {
      r.x = coef[t][0] * d[0].x;
      r.y = coef[t][0] * d[0].y;
#pragma unroll
      for (int k = 1; k < K; ++k) {
        r.x = fma(coef[t][k], d[k].x, r.x);
        r.y = fma(coef[t][k], d[k].y, r.y);
      }
    } else {
      r = coef[t][0] * d[0];
#pragma unroll
      for (int k = 1; k < K; ++k) r = fma(coef[t][k], d[k], r);
    }
    __stcs(dst, r);  // streaming store: the outputs are not read again here
    dst += step;
  }
}

template <int K>
static int launch_stream_k(const StreamArgs& a, const double* scalars, int n_scalars, double* out,
                           long batch, cudaStream_t st) {
  const bool vec2 = (a.n_elems % 2 == 0) && ((reinterpret_cast<uintptr_t>(out) & 15) == 0) &&
                    ((reinterpret_cast<uintptr_t>(a.consts) & 15) == 0);
  const long n_vec = vec2 ? a.n_elems / 2 : a.n_elems;
  dim3 grid((unsigned)((n_vec + kStreamThreads - 1) / kStreamThreads),
            (unsigned)((batch + kStreamGroup - 1) / kStreamGroup));
  if (grid.y > 65535) return fail(-6, "assemble: batch %ld too large for one launch", batch);
  if (vec2)
    stream_terms_kernel<K, double2><<<grid, kStreamThreads, 0, st>>>(a, scalars, n_scalars, out,
                                                                    batch);
  else
    stream_terms_kernel<K, double><<<grid, kStreamThreads, 0, st>>>(a, scalars, n_scalars, out,
                                                                   batch);
  MPCI_LAUNCH_CHECK("stream_terms_kernel");
  return 0;
}

int launch_stream(const StreamArgs& a, const double* scalars, int n_scalars, double* out,
                  long batch, cudaStream_t st) {
  if (a.n_elems == 0 || batch == 0) return 0;
  switch (a.n_terms) {
    case 1: return launch_stream_k<1>(a, scalars, n_scalars, out, batch, st);
    case 2: return launch_stream_k<2>(a, scalars, n_scalars, out, batch, st);
    case 3: return launch_stream_k<3>(a, scalars, n_scalars, out, batch, st);
    case 4: return launch_stream_k<4>(a, scalars, n_scalars, out, batch, st);
    case 5: return launch_stream_k<5>(a, scalars, n_scalars, out, batch, st);
    case 6: return launch_stream_k<6>(a, scalars, n_scalars, out, batch, st);
    case 7: return launch_stream_k<7>(a, scalars, n_scalars, out, batch, st);
    case 8: return launch_stream_k<8>(a, scalars, n_scalars, out, batch, st);
  }
  return fail(-7, "assemble: %d terms in one stream (limit %d)", a.n_terms, MPCI_MAX_TERMS);
}

// ---------------------------------------------------------------------------
// [q; h] = sum_t c_t * DY_t @ [given; 1]
// ---------------------------------------------------------------------------
constexpr int kQhGroup = 16;  // instances per block

// DYt is the transposed copy built at plan creation: DYt[t][k][row], so that the
// threads of a block (one row each) read consecutive addresses.
__global__ void __launch_bounds__(MPCI_QH_ROW_TILE)
qh_kernel(QhArgs a, const double* __restrict__ scalars, const double* __restrict__ given,
          double* __restrict__ q, double* __restrict__ h, long batch) {
  extern __shared__ double sm[];
  const int ngx = a.ng + 1;
  double* gs = sm;                       // [ngx][kQhGroup]
  double* cs = sm + (long)ngx * kQhGroup;  // [n_terms][kQhGroup]
  const long b0 = (long)blockIdx.y * kQhGroup;
  const int group = (int)min((long)kQhGroup, batch - b0);
  const int rows = a.nv + a.nc;

  for (int e = threadIdx.x; e < ngx * kQhGroup; e += blockDim.x) {
    const int k = e / kQhGroup, t = e % kQhGroup;  // conflict-free shared stores
    double v = 0.0;
    if (t < group) v = (k < a.ng) ? given[(b0 + t) * a.ng + k] : 1.0;
    gs[e] = v;
  }
  for (int e = threadIdx.x; e < a.n_terms * kQhGroup; e += blockDim.x) {
    const int j = e / kQhGroup, t = e % kQhGroup;
    double c = 0.0;
    if (t < group) {
      c = 1.0;
      const double* s = scalars + (b0 + t) * a.n_scalars;
      for (int f = 0; f < MPCI_MAX_FACTORS; ++f) {
        const int idx = a.factors[j * MPCI_MAX_FACTORS + f];
        if (idx >= 0) c *= s[idx];
      }
    }
    cs[j * kQhGroup + t] = c;
  }
  __syncthreads();

  const int row = blockIdx.x * MPCI_QH_ROW_TILE + threadIdx.x;
  if (row >= rows) return;
  double y[kQhGroup];
#pragma unroll
  for (int t = 0; t < kQhGroup; ++t) y[t] = 0.0;

  const int r_begin = a.tile_ptr[blockIdx.x], r_end = a.tile_ptr[blockIdx.x + 1];
  for (int r = r_begin; r < r_end; ++r) {
    const int4 range = reinterpret_cast<const int4*>(a.ranges)[r];  // (tile, term, k0, k1)
    const double* d = a.DYt + ((long)range.y * ngx + range.z) * rows + row;
    double acc[kQhGroup];
#pragma unroll
    for (int t = 0; t < kQhGroup; ++t) acc[t] = 0.0;
    for (int k = range.z; k < range.w; ++k, d += rows) {
      const double dv = *d;
      const double2* g2 = reinterpret_cast<const double2*>(gs + k * kQhGroup);
#pragma unroll
      for (int t2 = 0; t2 < kQhGroup / 2; ++t2) {
        const double2 g = g2[t2];
        acc[2 * t2] = fma(dv, g.x, acc[2 * t2]);
        acc[2 * t2 + 1] = fma(dv, g.y, acc[2 * t2 + 1]);
      }
    }
    const double* c = cs + range.y * kQhGroup;
#pragma unroll
    for (int t = 0; t < kQhGroup; ++t) y[t] = fma(c[t], acc[t], y[t]);
  }
#pragma unroll
  for (int t = 0; t < kQhGroup; ++t) {
    if (t < group) {
      if (row < a.nv) {
        if (q) q[(b0 + t) * a.nv + row] = y[t];
      } else if (h) {
        h[(b0 + t) * a.nc + (row - a.nv)] = y[t];
      }
    }
  }
}

int launch_qh(const QhArgs& a, const double* scalars, const double* given, double* q, double* h,
              long batch, cudaStream_t st) {
  const int rows = a.nv + a.nc;
  if (rows == 0 || batch == 0 || (!q && !h)) return 0;
  const size_t smem = sizeof(double) * ((size_t)(a.ng + 1) + a.n_terms) * kQhGroup;
  if (smem > 200 * 1024)
    return fail(-8, "assemble: given_len %d too large for the [q;h] kernel", a.ng);
  if (smem > 48 * 1024)
    MPCI_CUDA(cudaFuncSetAttribute(qh_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)smem));
  dim3 grid((rows + MPCI_QH_ROW_TILE - 1) / MPCI_QH_ROW_TILE,
            (unsigned)((batch + kQhGroup - 1) / kQhGroup));
  if (grid.y > 65535) return fail(-6, "assemble: batch %ld too large for one launch", batch);
  qh_kernel<<<grid, MPCI_QH_ROW_TILE, smem, st>>>(a, scalars, given, q, h, batch);
  MPCI_LAUNCH_CHECK("qh_kernel");
  return 0;
}

// DYt[t][k][row] = DY[t][row][k]
__global__ void __launch_bounds__(256)
transpose_kernel(const double* __restrict__ in, double* __restrict__ out, int n_terms, int rows,
                 int cols) {
  const long total = (long)n_terms * rows * cols;
  const long stride = (long)gridDim.x * blockDim.x;
  for (long e = (long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride) {
    const int row = e % rows;
    const int k = (e / rows) % cols;
    const long t = e / ((long)rows * cols);
    out[e] = in[(t * rows + row) * cols + k];
  }
}

int launch_transpose(const double* in, double* out, int n_terms, int rows, int cols,
                     cudaStream_t st) {
  const long total = (long)n_terms * rows * cols;
  if (total == 0) return 0;
  const int blocks = (int)((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
  transpose_kernel<<<blocks, 256, 0, st>>>(in, out, n_terms, rows, cols);
  MPCI_LAUNCH_CHECK("transpose_kernel");
  return 0;
}

}  // namespace mpci
