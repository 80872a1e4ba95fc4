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
// qh_kernel is a sparse FP64 product (one instance per thread, the shared
// sparsity pattern walked in lock-step).
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
// DY is sparse (axis blocks, triangular Toeplitz blocks, shifted diagonals:
// 51 k non-zeros of 825 k entries at the biped N=100), and the sparsity is the
// same for every instance.  So a THREAD owns an instance and a warp walks the
// non-zeros of its rows in lock-step: the constant (value, column) is one
// 16-byte load at a warp-uniform address, the instance's [given; 1] sits in
// shared memory (column-major, padded: conflict free), one DFMA per non-zero.
// Rows come in chunks of 8 that a warp stages in shared memory and writes out
// as one 64-byte run per instance.
constexpr int kQhInst = 32;                 // instances per block (one per lane)
constexpr int kQhLd = kQhInst + 1;          // padded leading dimension of gs / cs
constexpr int kQhTileLd = kQhChunk + 1;

__global__ void __launch_bounds__(kQhWarps * 32)
qh_kernel(QhArgs a, const double* __restrict__ scalars, const double* __restrict__ given,
          double* __restrict__ q, double* __restrict__ h, long batch) {
  extern __shared__ double sm[];
  const int ngx = a.ng + 1;
  QhNz* ring = reinterpret_cast<QhNz*>(sm) + (threadIdx.x >> 5) * 32;  // [kQhWarps][32], 16-byte aligned
  double* gs = sm + kQhWarps * 32 * (sizeof(QhNz) / sizeof(double));  // [ngx][kQhLd]
  double* cs = gs + (size_t)ngx * kQhLd;             // [n_terms][kQhLd]
  double* tiles = cs + (size_t)a.n_terms * kQhLd;    // [kQhWarps][kQhInst][kQhTileLd]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const long b0 = (long)blockIdx.x * kQhInst;
  const int group = (int)min((long)kQhInst, batch - b0);

  // stage [given; 1] of the block's instances: coalesced reads, conflict-free writes
  for (int t = warp; t < kQhInst; t += kQhWarps) {
    const double* g = given + (b0 + t) * a.ng;
    for (int k = lane; k < ngx; k += 32) {
      double v = 0.0;
      if (t < group) v = (k < a.ng) ? g[k] : 1.0;
      gs[k * kQhLd + t] = v;
    }
  }
  for (int e = tid; e < a.n_terms * kQhInst; e += blockDim.x) {
    const int j = e / kQhInst, t = e % kQhInst;
    double c = 0.0;
    if (t < group) {
      c = 1.0;
      const double* s = scalars + (b0 + t) * a.n_scalars;
      for (int f = 0; f < MPCI_MAX_FACTORS; ++f) {
        const int idx = a.factors[j * MPCI_MAX_FACTORS + f];
        if (idx >= 0) c *= s[idx];
      }
    }
    cs[j * kQhLd + t] = c;
  }
  __syncthreads();

  double* tile = tiles + (size_t)warp * kQhInst * kQhTileLd;
  const double* gl = gs + lane;
  const double* cl = cs + lane;
  const int e_begin = a.warp_nz_ptr[warp], e_end = a.warp_nz_ptr[warp + 1];
  int ci = a.warp_chunk_ptr[warp];
  double acc = 0.0, part = 0.0;
  int r = 0;
  // the warp's record stream comes in batches of 32: one coalesced 16-byte load
  // per lane (the next batch is in flight while this one is consumed), parked in
  // a 512-byte ring and read back as warp-uniform LDS.128
  QhNz next = {0.0, 0, 0};
  if (e_begin + lane < e_end) next = a.nz[e_begin + lane];
  for (int base = e_begin; base < e_end; base += 32) {
    ring[lane] = next;
    if (base + 32 + lane < e_end) next = a.nz[base + 32 + lane];
    __syncwarp();
    const int cnt = min(32, e_end - base);
#pragma unroll 4
    for (int i = 0; i < cnt; ++i) {
      const QhNz rec = ring[i];
      part = fma(rec.val, gl[rec.col * kQhLd], part);
      if (rec.meta & ~kQhTermMask) {  // warp-uniform
        acc = fma(cl[(rec.meta & kQhTermMask) * kQhLd], part, acc);
        part = 0.0;
        if (rec.meta & kQhEndRow) {
          tile[lane * kQhTileLd + r] = acc;
          acc = 0.0;
          ++r;
        }
        if (rec.meta & kQhEndChunk) {
          const int2 chunk = a.chunks[ci++];  // (first row, rows)
          __syncwarp();
          // 8 lanes cover the chunk of one instance (64 contiguous bytes), 4 instances per pass
          const bool is_q = chunk.x < a.nv;
          double* out = is_q ? q : h;
          if (out) {
            const int rr = lane & (kQhChunk - 1), sub = lane / kQhChunk;
            const long stride = is_q ? a.nv : a.nc;
            const int row = is_q ? chunk.x : chunk.x - a.nv;
            if (rr < chunk.y) {
#pragma unroll
              for (int it = 0; it < kQhInst / (32 / kQhChunk); ++it) {
                const int t = it * (32 / kQhChunk) + sub;
                if (t < group) out[(b0 + t) * stride + row + rr] = tile[t * kQhTileLd + rr];
              }
            }
          }
          __syncwarp();
          r = 0;
        }
      }
    }
    __syncwarp();
  }
}

size_t qh_smem_bytes(const QhArgs& a) {
  return sizeof(double) * ((size_t)(a.ng + 1 + a.n_terms) * kQhLd +
                           (size_t)kQhWarps * kQhInst * kQhTileLd) +
         sizeof(QhNz) * kQhWarps * 32;
}

int launch_qh(const QhArgs& a, const double* scalars, const double* given, double* q, double* h,
              long batch, cudaStream_t st) {
  const int rows = a.nv + a.nc;
  if (rows == 0 || batch == 0 || (!q && !h)) return 0;
  const size_t smem = qh_smem_bytes(a);
  if (smem > 227 * 1024)
    return fail(-8, "assemble: given_len %d too large for the [q;h] kernel", a.ng);
  if (smem > 48 * 1024)
    MPCI_CUDA(cudaFuncSetAttribute(qh_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)smem));
  const long blocks = (batch + kQhInst - 1) / kQhInst;
  if (blocks > 2147483647L) return fail(-6, "assemble: batch %ld too large for one launch", batch);
  qh_kernel<<<(unsigned)blocks, kQhWarps * 32, smem, st>>>(a, scalars, given, q, h, batch);
  MPCI_LAUNCH_CHECK("qh_kernel");
  return 0;
}

}  // namespace mpci
