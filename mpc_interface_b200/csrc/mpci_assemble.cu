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

// More than 8 monomials (e.g. several boxes turned by different per-instance angles):
// the constants do not fit the register file next to 32 instances, so the thread works on
// eight instances at a time -- eight accumulators, every constant vector fetched once
// per eight outputs through the read-only path (L1 / L2), the eight coefficients of a
// term as two warp-uniform 32-byte shared-memory reads.  Same grid, same store pattern,
// no extra HBM bytes.
constexpr int kGenericSub = 8;
template <typename VecT>
__global__ void __launch_bounds__(kStreamThreads)
stream_terms_generic(StreamArgs args, const double* __restrict__ scalars, int n_scalars,
                     double* __restrict__ out, long batch) {
  constexpr int V = sizeof(VecT) / sizeof(double);
  __shared__ __align__(16) double coef[MPCI_MAX_TERMS][kStreamGroup];  // [term][instance]
  const int K = args.n_terms;
  const long b0 = (long)blockIdx.y * kStreamGroup;
  const int group = (int)min((long)kStreamGroup, batch - b0);
  for (int e = threadIdx.x; e < kStreamGroup * K; e += blockDim.x) {
    const int k = e / kStreamGroup, t = e - k * kStreamGroup;
    double c = 0.0;  // instances past the batch: zero coefficients, nothing stored
    if (t < group) {
      const double* s = scalars + (b0 + t) * n_scalars;
      c = 1.0;
#pragma unroll
      for (int f = 0; f < MPCI_MAX_FACTORS; ++f) {
        const int idx = args.factors[k][f];
        if (idx >= 0) c *= s[idx];
      }
    }
    coef[k][t] = c;
  }
  __syncthreads();
  const long n_vec = args.n_elems / V;
  const long v = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= n_vec) return;
  const VecT* src = reinterpret_cast<const VecT*>(args.consts) + v;
  VecT* dst = reinterpret_cast<VecT*>(out + b0 * args.n_elems) + v;
  for (int t0 = 0; t0 < group; t0 += kGenericSub, dst += kGenericSub * n_vec) {
    VecT acc[kGenericSub];
#pragma unroll
    for (int i = 0; i < kGenericSub; ++i) {
      if constexpr (V == 2) acc[i].x = acc[i].y = 0.0;
      else acc[i] = 0.0;
    }
    for (int k = 0; k < K; ++k) {
      const VecT d = __ldg(src + (long)k * n_vec);
      const double4 ca = *reinterpret_cast<const double4*>(&coef[k][t0]);
      const double4 cb = *reinterpret_cast<const double4*>(&coef[k][t0 + 4]);
      const double c[kGenericSub] = {ca.x, ca.y, ca.z, ca.w, cb.x, cb.y, cb.z, cb.w};
#pragma unroll
      for (int i = 0; i < kGenericSub; ++i) {
        if constexpr (V == 2) {
          acc[i].x = fma(c[i], d.x, acc[i].x);
          acc[i].y = fma(c[i], d.y, acc[i].y);
        } else {
          acc[i] = fma(c[i], d, acc[i]);
        }
      }
    }
#pragma unroll
    for (int i = 0; i < kGenericSub; ++i)
      if (t0 + i < group) __stcs(dst + (long)i * n_vec, acc[i]);
  }
}

static int launch_stream_generic(const StreamArgs& a, const double* scalars, int n_scalars,
                                 double* out, long batch, cudaStream_t st) {
  const bool vec2 = (a.n_elems % 2 == 0) && ((reinterpret_cast<uintptr_t>(out) & 15) == 0) &&
                    ((reinterpret_cast<uintptr_t>(a.consts) & 15) == 0);
  const long n_vec = vec2 ? a.n_elems / 2 : a.n_elems;
  dim3 grid((unsigned)((n_vec + kStreamThreads - 1) / kStreamThreads),
            (unsigned)((batch + kStreamGroup - 1) / kStreamGroup));
  if (grid.y > 65535) return fail(-6, "assemble: batch %ld too large for one launch", batch);
  if (vec2)
    stream_terms_generic<double2><<<grid, kStreamThreads, 0, st>>>(a, scalars, n_scalars, out, batch);
  else
    stream_terms_generic<double><<<grid, kStreamThreads, 0, st>>>(a, scalars, n_scalars, out, batch);
  MPCI_LAUNCH_CHECK("stream_terms_generic");
  return 0;
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
  if (a.n_terms > 8 && a.n_terms <= MPCI_MAX_TERMS)
    return launch_stream_generic(a, scalars, n_scalars, out, batch, st);
  return fail(-7, "assemble: %d terms in one stream (limit %d)", a.n_terms, MPCI_MAX_TERMS);
}

// ---------------------------------------------------------------------------
// [q; h] = sum_t c_t * DY_t @ [given; 1]
// ---------------------------------------------------------------------------
// DY is sparse (axis blocks, triangular Toeplitz blocks, shifted diagonals:
// 25.6 k non-zeros of 825 k entries at the biped N=100), and the sparsity is the
// same for every instance.  So a THREAD owns two instances and a warp walks the
// non-zeros of its rows in lock-step: the constant (value, column) record is a
// warp-uniform 16-byte shared-memory read, the two instances' [given; 1] entries
// are one more (column-major double2, padded: conflict free), two DFMAs per
// record.  Rows come in chunks of 4 that a warp stages in shared memory and
// writes out as one 32-byte run per instance.
constexpr int kQhInst = 64;                 // instances per block (two per lane)
constexpr int kQhTileLd = kQhChunk + 1;

__global__ void __launch_bounds__(kQhWarps * 32)
qh_kernel(QhArgs a, const double* __restrict__ scalars, const double* __restrict__ given,
          double* __restrict__ q, double* __restrict__ h, long batch) {
  extern __shared__ double sm[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  QhNz* ring = reinterpret_cast<QhNz*>(sm) + warp * 32;                // [kQhWarps][32]
  double* gs = sm + kQhWarps * 32 * (sizeof(QhNz) / sizeof(double));   // [max_cols][33] double2
  double* cs = gs + (size_t)a.max_cols * 2 * kQhColStride;                        // [n_terms][33] double2
  double* tiles = cs + (size_t)a.n_terms * 2 * kQhColStride;              // [kQhWarps][kQhInst][kQhTileLd]
  // persistent: at most one block per SM, so that every block is resident at once
  // and the P / G streams launched beside this kernel can use the other thread slots
  for (long b0 = (long)blockIdx.x * kQhInst; b0 < batch; b0 += (long)gridDim.x * kQhInst) {
  const int group = (int)min((long)kQhInst, batch - b0);
  __syncthreads();  // the previous group is done with gs / cs

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
    cs[j * 66 + (t & 31) * 2 + (t >> 5)] = c;
  }

  // one pass per block of columns of [given; 1] (a single pass unless the columns do
  // not fit in shared memory): pass 0 writes the rows, the later ones add to them
  for (int pass = 0; pass < a.n_pass; ++pass) {
  const int c0 = a.pass_col[pass], ncol = a.pass_col[pass + 1] - c0;
  if (pass) __syncthreads();  // the previous pass is done with gs
  // stage the columns: instance t sits in lane t % 32, half t / 32 of every column
  for (int t = warp; t < kQhInst; t += kQhWarps) {
    const double* g = given + (b0 + t) * a.ng;
    double* dst = gs + (t & 31) * 2 + (t >> 5);
    for (int k = lane; k < ncol; k += 32) {
      double v = 0.0;
      if (t < group) v = (c0 + k < a.ng) ? g[c0 + k] : 1.0;
      dst[k * 66] = v;
    }
  }
  __syncthreads();
  const int* nz_ptr = a.warp_nz_ptr + pass * (kQhLists + 1);
  const int* chunk_ptr = a.warp_chunk_ptr + pass * (kQhLists + 1);

  double* tile = tiles + (size_t)warp * kQhInst * kQhTileLd;
  const double2* gl = reinterpret_cast<const double2*>(gs) + lane;
  const double2* cl = reinterpret_cast<const double2*>(cs) + lane;
  for (int list = warp + kQhWarps * (int)blockIdx.y; list < kQhLists; list += kQhWarps * (int)gridDim.y) {
  const int e_begin = nz_ptr[list], e_end = nz_ptr[list + 1];
  int ci = chunk_ptr[list];
  double acc0 = 0.0, acc1 = 0.0, p0 = 0.0, p1 = 0.0;
  int r = 0;
  // The list's record stream comes in batches of 32: one coalesced 16-byte load
  // per lane (the next batch is in flight while this one is consumed), parked in
  // a 512-byte ring and read back as warp-uniform LDS.128.  A ballot marks the
  // records that end a segment, so the runs between them are a branch-free loop.
  QhNz next = {0.0, 0, 0};
  if (e_begin + lane < e_end) next = a.nz[e_begin + lane];
  for (int base = e_begin; base < e_end; base += 32) {
    const int cnt = min(32, e_end - base);
    ring[lane] = next;
    const unsigned flagged = __ballot_sync(0xffffffffu, lane < cnt && (next.meta & ~kQhTermMask));
    if (base + 32 + lane < e_end) next = a.nz[base + 32 + lane];
    __syncwarp();
    int i = 0;
    while (i < cnt) {
      const unsigned rest = flagged >> i;
      const int stop = rest ? i + __ffs(rest) : cnt;  // one past the next flagged record
#pragma unroll 4
      for (; i < stop; ++i) {
        const QhNz rec = ring[i];
        const double2 g2 = gl[rec.col];
        p0 = fma(rec.val, g2.x, p0);
        p1 = fma(rec.val, g2.y, p1);
      }
      if (!rest) break;
      const int meta = ring[i - 1].meta;  // the flagged record just consumed
      const double2 c2 = cl[(meta & kQhTermMask) * 33];
      acc0 = fma(c2.x, p0, acc0);
      acc1 = fma(c2.y, p1, acc1);
      p0 = p1 = 0.0;
      if (meta & kQhEndRow) {
        tile[lane * kQhTileLd + r] = acc0;
        tile[(lane + 32) * kQhTileLd + r] = acc1;
        acc0 = acc1 = 0.0;
        ++r;
      }
      if (meta & kQhEndChunk) {
        const int2 chunk = a.chunks[ci++];  // (first row, rows)
        __syncwarp();
        // 4 lanes cover the chunk of one instance (32 contiguous bytes), 8 instances per pass
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
              if (t < group) {
                double* o = out + (b0 + t) * stride + row + rr;
                *o = pass ? *o + tile[t * kQhTileLd + rr] : tile[t * kQhTileLd + rr];
              }
            }
          }
        }
        __syncwarp();
        r = 0;
      }
    }
    __syncwarp();
  }
  }  // lists of this warp
  }  // column passes
  }  // groups of instances
}

size_t qh_smem_bytes(const QhArgs& a) {
  return sizeof(double) * ((size_t)(a.max_cols + a.n_terms) * 66 +
                           (size_t)kQhWarps * kQhInst * kQhTileLd) +
         sizeof(QhNz) * kQhWarps * 32;
}

int launch_qh(const QhArgs& a, const double* scalars, const double* given, double* q, double* h,
              long batch, cudaStream_t st) {
  const int rows = a.nv + a.nc;
  if (rows == 0 || batch == 0 || (!q && !h)) return 0;
  const size_t smem = qh_smem_bytes(a);
  if (smem > 227 * 1024)
    return fail(-8, "assemble: %d monomials of [q;h] do not fit the kernel's shared memory", a.n_terms);
  if (smem > 48 * 1024)
    MPCI_CUDA(cudaFuncSetAttribute(qh_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)smem));
  int dev = 0, sms = 148, per_sm = 1;
  MPCI_CUDA(cudaGetDevice(&dev));
  MPCI_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  MPCI_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, qh_kernel, kQhWarps * 32, smem));
  const long groups = (batch + kQhInst - 1) / kQhInst;
  const long resident = (long)sms * (per_sm < 1 ? 1 : per_sm);
  const unsigned blocks = (unsigned)(groups < resident ? groups : resident);
  int split = 1;  // blocks per group of instances: as many as leave no SM without one
  while (split < kQhSplit && (long)blocks * split * 2 <= resident) split *= 2;
  qh_kernel<<<dim3(blocks, split), kQhWarps * 32, smem, st>>>(a, scalars, given, q, h, batch);
  MPCI_LAUNCH_CHECK("qh_kernel");
  return 0;
}

}  // namespace mpci
