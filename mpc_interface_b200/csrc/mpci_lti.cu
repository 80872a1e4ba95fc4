// Per-instance dynamics: the whole condensed LTI QP of one instance in one CTA.
//
// Mode B of DESIGN.md.  Reference path, per instance: tools.extend_matrices
// (python/mpc_interface/tools.py:14-33), Formulation.make_preview_matrices
// (body.py:149-193), generate_qp_cost (body.py:266-302) for every cost and
// generate_qp_constraint (body.py:236-264) for every limit -- for a formulation
// whose variables are point-wise linear forms of the plant's states and inputs
// (lowered to stage tables by mpc_interface_b200/lti.py, see include/mpci.h).
//
// Phases of one instance (everything below lives in shared memory):
//   1. Gam'_d = A^d [B | x0], d = 0..N, by parallel prefix doubling
//   2. gz/wz: zero-padded Toeplitz generators Gam and W = Qxx Gam;  r_k
//   3. P   -- FP64 tensor cores (DMMA.8x8x4) with block-triangular skipping, or
//             the displacement recursion (O(N^2) flops)
//   4. q   -- correlation of Gam with r
//   5. G,h -- generated row by row, 16-byte coalesced stores
#include <algorithm>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "mpci_common.cuh"
#include "mpci_couple.cuh"

namespace {

#ifndef MPCI_LTI_THREADS
#define MPCI_LTI_THREADS 256
#endif
#ifndef MPCI_LTI_MINBLOCKS
#define MPCI_LTI_MINBLOCKS 3  /* resident CTAs per SM the recursion kernel is compiled for */
#endif
constexpr int kThreads = MPCI_LTI_THREADS;
constexpr int kWarps = kThreads / 32;
constexpr int kWideThreads = 384;
constexpr int kPadL = 36;  // zeros before d = 0 of every generator (>= 8*max(TM,TN)+3)
constexpr int TM = 2, TN = 4;  // 8x8 DMMA tiles per warp tile (rows x cols)

struct LtiArgs {
  int n, m, N, S, n_groups;
  int pad;  // zeros before d = 0 of every generator row: kPadL (DMMA mode) or 0
  int LD;   // row length of gz: pad + (N + 1 rounded up to a multiple of 4), plus the
            // skew that keeps the strided DMMA operand loads of the scan conflict-free
  int LDw;  // row length of wz, likewise
  int per_sm;  // resident CTAs per SM of this launch
  int sms;  // SMs of the device (CTA i shares an SM with i + sms, i + 2 sms, ...)
  int scan; // recursion mode: 1 = DMMA row bands + diagonal scan (needs tb / prev), 0 = plain,
            // 2 = whole rows per warp (short horizons, hessian_rows)
  int ldp;  // scan == 2: row length (doubles) of a warp's T patch
  int scan_warps;  // scan == 2: warps that own a T patch (the others do not take Hessian items)
  int aux;  // doubles of the aux region (row windows of the constraint phase / T patches)
  const double *Qxx, *Quu, *Tx, *Tu, *cx, *cu, *arrow, *bound;  // device tables
  // instance strides of the tables (doubles; 0 = one table for the batch): per-instance
  // cost weights / form coefficients / bounds (mpci_lti_assemble_ex)
  long sQ, sR, sTx, sTu, scx, scu, sbd;
  const int* grp_state;  // bit 0: the group has a non-zero cx; bit 1: one arrow for all rows
  const int* grp_k0;     // first horizon sample of the group's rows (a limit's schedule)
  const int* grp_cnt;    // number of rows of the group
  const int* grp_row0;   // first row of the group in G / h
  int nc;                // rows of G / h: sum of grp_cnt
  const int2* tiles;     // DMMA warp tiles (row tile, column tile) of one (j, j') block, longest first
  int n_tiles;
  const double *A, *B, *x0, *scalars;
  long strideA, strideB;
  double *P, *q, *G, *h;
  long batch;
  int hessian_mode;
  // embedded layout (EMB kernels only): the plant's inputs are SOME of the QP's variables
  // and its limits SOME of the rows -- input j's samples live at columns voff[j] .. + N of
  // rows of ldv doubles, group g's rows start at grp_row0[g] of nct rows per instance
  int ldv, nct;
  int emb_even;  // ldv and every voff[j] even: 16-byte stores / bulk copies stay aligned
  const int* voff;
  // the rows' constant entries under the auxiliary variables (mpci_lti_embed::Gz, z_col): written
  // with the rows, so that every 32-byte sector of G is complete when it leaves the L2
  const double* Gz;
  const int* zcol;
  int nz;
  // the blocks of P that involve the auxiliary variables (coupled_hessian_emb): the coupled
  // costs' plant forms (instance strides sccx / sccu, 0 = the plan's), weights, the columns of
  // their Mz as CSR differences over (cost, z), and the constant block Pzz
  int nC;
  const double *ccx, *ccu, *ccw, *d_val, *Pzz;
  long sccx, sccu, sccw, sPzz, sGz;  // (instance strides, 0 = the plan's table)
  const int *d_ptr, *d_tau;
};

// Where the kernel's variables live in the QP.  EMB = false: the plant's inputs ARE the
// variables (row length m N, input j at column j N) -- the arithmetic of the plain kernels.
// (int there, as the plain kernels always computed it; long in the embedded layout)
template <bool EMB> __device__ __forceinline__ auto row_len(const LtiArgs& a) {
  if constexpr (EMB) return (long)a.ldv;
  else return a.m * a.N;
}
template <bool EMB> __device__ __forceinline__ auto var_off(const LtiArgs& a, int j) {
  if constexpr (EMB) return (long)a.voff[j];
  else return j * a.N;
}

struct Smem {
  double *A, *M0, *M1, *Q, *R, *gz, *wz, *r, *tx, *tu, *cg, *tb;
  int* next_item;  // work counter of the Hessian scan (warps fetch items as they get free)
};

// Shared memory of one instance (doubles):
//   A, M0, M1, Q   4 n^2  (A for the adjoint recursion of q, A^have ping-pong, Qxx)
//   gz  [n][m+1][LD]  gz[i][c][pad + d] = (A^d [B | x0])[i][c], d = 0..N: the doubling
//                     writes straight into it; column c = m is the free response,
//                     f_k[i] = gz[i][m][pad + 1 + k]
//   wz  [n][m][LDw]   (Qxx Gam_d)[i][j]
//   r   [n][N],  tx [n],  tu [m]
//   aux = tb [warps][8][40] (the warps' T patches, scan only) / cg (row windows and
//         group tables of the constraint phase)
__host__ __device__ inline int lti_row_len(int N, int pad) { return pad + (N + 1 + 3) / 4 * 4; }
__host__ __device__ inline size_t lti_wz_doubles(int n, int m, int LDw) {
  return ((size_t)n * m * LDw + 1) & ~(size_t)1;
}
__host__ __device__ inline size_t lti_r_doubles(int n, int N) { return ((size_t)n * N + 1) & ~(size_t)1; }
// aux: the constraint phase's row windows and group tables (cg) and, in the scan, the
// warps' T patches (tb) share one region -- the two phases never run at the same time
__host__ __device__ inline size_t lti_aux_doubles(int m, int N, size_t patches) {
  const size_t cg = (size_t)2 * 4 * m * N;  // two window sets; `patches` = doubles of the T patches
  return ((cg > patches ? cg : patches) + 1) & ~(size_t)1;
}
// hessian_rows: a T patch is 8 rows of ldp doubles, ldp = 8 * ceil(N / 8) bumped to 8 (mod 16)
// so that the double2 stores of the DMMA fragments (rows g, g+1 per quarter-warp) do not
// collide on a bank
__host__ __device__ inline int lti_rows_ldp(int N) {
  int x = (N + 2 + 7) / 8 * 8;  // whole 8 x 8 tiles, at least two (zero) columns past the block
  while (x % 16 != 8) x += 8;
  return x;
}
__host__ __device__ inline size_t lti_mats_doubles(int n, int m) {  // A, M0, M1, Qxx | Quu
  return (size_t)4 * n * n + (((size_t)m * m + 1) & ~(size_t)1);
}
__host__ __device__ inline size_t lti_gz_doubles(int n, int m, int LD) {
  return ((size_t)n * (m + 1) * LD + 1) & ~(size_t)1;
}
// The scan reads its DMMA operands with lane (g, t) at word  -g + t * rows * LD: the 16
// lanes of a half-warp hit 16 different 8-byte banks iff rows * LD = 4 or 12 (mod 16).
__host__ __device__ inline int lti_skew(int LD, int rows) {
  for (int x = LD; x < LD + 16; ++x)
    if ((rows * x) % 16 == 4 || (rows * x) % 16 == 12) return x;
  return LD;
}
__host__ __device__ inline size_t lti_smem_doubles(int n, int m, int N, int LD, int LDw, size_t aux) {
  return lti_mats_doubles(n, m) + lti_gz_doubles(n, m, LD) + lti_wz_doubles(n, m, LDw) +
         lti_r_doubles(n, N) + aux + n + m + 2;
}

__device__ __forceinline__ Smem carve(double* base, const LtiArgs& a) {
  Smem s;
  const int n = a.n, m = a.m, N = a.N;
  s.A = base;
  s.M0 = s.A + n * n;
  s.M1 = s.M0 + n * n;
  s.Q = s.M1 + n * n;
  s.R = s.Q + n * n;
  s.gz = base + lti_mats_doubles(n, m);
  s.wz = s.gz + lti_gz_doubles(n, m, a.LD);
  s.r = s.wz + lti_wz_doubles(n, m, a.LDw);
  s.tb = s.r + lti_r_doubles(n, N);  // 16-byte aligned: double2 stores of the T tiles,
  s.cg = s.tb;                       // source of the bulk copies
  s.tx = s.tb + a.aux;
  s.tu = s.tx + n;
  s.next_item = reinterpret_cast<int*>(s.tu + m);
  return s;
}

__device__ __forceinline__ void dmma8x8x4(double& d0, double& d1, double a, double b) {
  asm volatile(
      "mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
      : "+d"(d0), "+d"(d1)
      : "d"(a), "d"(b));
}

// One 8x8 tile of D = X Y on the FP64 tensor cores: X is n x n, row-major in shared
// memory (SYM: its symmetric part), Y(l, x) = Y[l * ys + x].  Lane (g, t) gets
// D[i0 + g][x0 + 2t] and D[i0 + g][x0 + 2t + 1].  Row g of D depends on row g of X only,
// so rows i >= n are simply read from row n - 1 (in bounds) and come out as garbage the
// caller does not store; likewise columns whose Y entries are not valid data yet.  Only
// the padding of K to a multiple of 4 must contribute exact zeros (KFULL: there is none).
template <bool SYM, bool KFULL>
__device__ __forceinline__ void small_tile(const double* X, int n, const double* Y, int ys, int i0,
                                           int x0, int g, int t, double& d0, double& d1) {
  d0 = d1 = 0.0;
  const int ic = min(i0 + g, n - 1);
  const double* Xi = X + ic * n + t;
  const double* Xt = X + ic + t * n;  // column ic, for the transposed read of SYM
  const double* Yg = Y + x0 + g + t * ys;
  const int full = n >> 2;
  for (int ks = 0; ks < full; ++ks, Xi += 4, Xt += 4 * n, Yg += 4 * ys) {
    const double av = SYM ? 0.5 * (Xi[0] + Xt[0]) : Xi[0];
    dmma8x8x4(d0, d1, av, Yg[0]);
  }
  if (!KFULL) {
    const bool k_ok = 4 * full + t < n;
    double av = 0.0, bv = 0.0;
    if (k_ok) {
      av = SYM ? 0.5 * (Xi[0] + Xt[0]) : Xi[0];
      bv = Yg[0];
    }
    dmma8x8x4(d0, d1, av, bv);
  }
}

// q = e / d for 0 <= e < 2^16 by a multiplication (magic = 0 stands for d = 1)
__device__ __forceinline__ unsigned magic_of(int d) { return d == 1 ? 0u : 0xFFFFFFFFu / (unsigned)d + 1u; }
__device__ __forceinline__ int magic_div(int e, unsigned magic) {
  return magic ? (int)__umulhi((unsigned)e, magic) : e;
}

// ---------------------------------------------------------------------------
// phase 1 + 2: prediction data of instance b
// ---------------------------------------------------------------------------
template <bool KFULL>
__device__ void build_prediction(const LtiArgs& a, const Smem& s, long b) {
  const int n = a.n, m = a.m, N = a.N, w = a.m + 1, LD = a.LD, pad = a.pad, tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  const double* Ab = a.A + b * a.strideA;
  const double* Bb = a.B + b * a.strideB;
  const double* xb = a.x0 + b * n;
  for (int e = tid; e < n * n; e += (int)blockDim.x) {
    const double v = Ab[e];
    s.A[e] = v;
    s.M0[e] = v;
    s.Q[e] = a.Qxx[b * a.sQ + e];
  }
  for (int e = tid; e < m * m; e += (int)blockDim.x) s.R[e] = a.Quu[b * a.sR + e];
  if (tid == 0) *s.next_item = 0;
  // d = 0 of every row, and the zeros before it (causality of the DMMA operands)
  for (int e = tid; e < n * w; e += (int)blockDim.x) {
    const int i = e / w, c = e - i * w;
    s.gz[(size_t)e * LD + pad] = (c < m) ? Bb[i * m + c] : xb[i];
  }
  if (pad) {
    for (int row = warp; row < n * w + n * m; row += ((int)blockDim.x >> 5)) {
      double* dst = row < n * w ? s.gz + (size_t)row * LD : s.wz + (size_t)(row - n * w) * a.LDw;
      for (int x = lane; x < pad; x += 32) dst[x] = 0.0;
    }
  }
  // per-instance offsets of the gradient: t = T[:, S] + T[:, :S] @ scalars_b
  for (int e = tid; e < n + m; e += (int)blockDim.x) {
    const double* T = (e < n) ? a.Tx + b * a.sTx + (long)e * (a.S + 1)
                              : a.Tu + b * a.sTu + (long)(e - n) * (a.S + 1);
    double acc = T[a.S];
#pragma unroll 4
    for (int k = 0; k < a.S; ++k) acc = fma(T[k], a.scalars[b * a.S + k], acc);
    if (e < n) s.tx[e] = acc; else s.tu[e - n] = acc;
  }
  __syncthreads();

  // prefix doubling: Gam'_{have+d} = A^have Gam'_d,  A^(2 have) = (A^have)^2, d = 0..N.
  // Every round is a handful of small matrix products with A^have on the left; they run
  // as 8x8 DMMA tiles (states x 8 consecutive d of one column c), a warp per tile, the
  // tiles of the doubling and of the squaring in one list between two barriers.
  double* M0 = s.M0;
  double* M1 = s.M1;
  const int g = lane >> 2, t = lane & 3, warps = (int)blockDim.x >> 5;
  const int it_n = (n + 7) >> 3;  // row tiles of an n x n left factor
  const int wLD = w * LD;
  const unsigned magic_w = magic_of(w), magic_m = magic_of(m), magic_it = magic_of(it_n);
  for (int have = 1; have < N + 1;) {
    const int cnt = min(have, N + 1 - have);
    const int n_dt = (cnt + 7) >> 3;
    const unsigned magic_dt = magic_of(n_dt);
    const int n_dbl = it_n * w * n_dt;
    const int total = n_dbl + ((have + cnt < N + 1) ? it_n * it_n : 0);  // + the next power, if needed
    for (int idx = warp; idx < total; idx += warps) {
      double d0, d1;
      if (idx < n_dbl) {
        const int rc = magic_div(idx, magic_dt), dt = idx - rc * n_dt;
        const int it = magic_div(rc, magic_w), c = rc - it * w;
        small_tile<false, KFULL>(M0, n, s.gz + c * LD + pad, wLD, it * 8, dt * 8, g, t, d0, d1);
        const int i = it * 8 + g, x = dt * 8 + 2 * t;
        double* o = s.gz + (i * w + c) * LD + pad + have + x;
        if (i < n && x < cnt) o[0] = d0;
        if (i < n && x + 1 < cnt) o[1] = d1;
      } else {
        const int e = idx - n_dbl, it = magic_div(e, magic_it), ct = e - it * it_n;
        small_tile<false, KFULL>(M0, n, M0, n, it * 8, ct * 8, g, t, d0, d1);
        const int i = it * 8 + g, x = ct * 8 + 2 * t;
        if (i < n && x < n) M1[i * n + x] = d0;
        if (i < n && x + 1 < n) M1[i * n + x + 1] = d1;
      }
    }
    __syncthreads();
    double* tmp = M0; M0 = M1; M1 = tmp;
    have += cnt;
  }

  // wz[i][j][pad + d] = (Qxx Gam_d)[i][j], d < N, and
  // r[i][k] = sum_l sym(Qxx)[i][l] f_k[l] - tx[i],  f_k = Gam'_{k+1}[:, m]: the same tiles
  const int n_dt = (N + 7) >> 3;
  const unsigned magic_dt = magic_of(n_dt);
  const int n_wz = it_n * m * n_dt;
  for (int idx = warp; idx < n_wz + it_n * n_dt; idx += warps) {
    double d0, d1;
    if (idx < n_wz) {
      const int rc = magic_div(idx, magic_dt), dt = idx - rc * n_dt;
      const int it = magic_div(rc, magic_m), j = rc - it * m;
      small_tile<false, KFULL>(s.Q, n, s.gz + j * LD + pad, wLD, it * 8, dt * 8, g, t, d0, d1);
      const int i = it * 8 + g, x = dt * 8 + 2 * t;
      double* o = s.wz + (i * m + j) * a.LDw + pad + x;
      if (i < n && x < N) o[0] = d0;
      if (i < n && x + 1 < N) o[1] = d1;
    } else {
      const int e = idx - n_wz, it = magic_div(e, magic_dt), dt = e - it * n_dt;
      small_tile<true, KFULL>(s.Q, n, s.gz + m * LD + pad + 1, wLD, it * 8, dt * 8, g, t, d0, d1);
      const int i = it * 8 + g, x = dt * 8 + 2 * t;
      const double txi = s.tx[min(i, n - 1)];
      if (i < n && x < N) s.r[i * N + x] = d0 - txi;
      if (i < n && x + 1 < N) s.r[i * N + x + 1] = d1 - txi;
    }
  }
  __syncthreads();
}

// ---------------------------------------------------------------------------
// phase 3a: P by the dense contraction on the FP64 tensor cores
//   C[(j,p)][(j',p')] = sum_i sum_k gz[i][j][k-p] * wz[i][j'][k-p']
// A warp owns (8 TM) x (8 TN) entries of one (j, j') block.  The DMMA operand
// of lane (g = lane/4, t = lane%4) is one LDS.64: gz[..][k0 + t - p0 - g] --
// 11 consecutive words per warp, so there are no bank conflicts; the zero
// padding before d = 0 encodes causality (k >= p) without a branch.
// ---------------------------------------------------------------------------
// the XX x YY live sub-tiles of a warp tile, one step of four samples, all the states
template <int XX, int YY>
__device__ __forceinline__ void dmma_states(double (&acc)[TM][TN][2], const double* ga, const double* wb,
                                            int n, long gi, long wi) {
#pragma unroll 2
  for (int i = 0; i < n; ++i, ga += gi, wb += wi) {
    double av[XX], bv[YY];
#pragma unroll
    for (int x = 0; x < XX; ++x) av[x] = ga[-8 * x];
#pragma unroll
    for (int y = 0; y < YY; ++y) bv[y] = wb[-8 * y];
#pragma unroll
    for (int x = 0; x < XX; ++x)
#pragma unroll
      for (int y = 0; y < YY; ++y) dmma8x8x4(acc[x][y][0], acc[x][y][1], av[x], bv[y]);
  }
}

template <bool EMB>
__device__ void hessian_dmma(const LtiArgs& a, const Smem& s, double* __restrict__ Pb) {
  const int n = a.n, m = a.m, N = a.N, LD = a.LD;
  const auto nv = row_len<EMB>(a);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int mm = m * m;
  const bool vec_ok = (N % 2 == 0) && ((reinterpret_cast<uintptr_t>(Pb) & 15) == 0) && (!EMB || a.emb_even);

  // tiles are sorted by the length of their k-loop and dealt round-robin, the
  // (j, j') blocks innermost: the eight warps get nearly equal work
  for (int idx = warp; idx < a.n_tiles * mm; idx += ((int)blockDim.x >> 5)) {
    const int tt = idx / mm, jj = idx - tt * mm;
    const int2 rc = a.tiles[tt];
    const int j = jj / m, jp = jj - j * m;
    const int p0 = rc.x * 8 * TM, q0 = rc.y * 8 * TN;
    // k runs downwards from N-1 in steps of 4 (lane t takes k = N-1-s0-t) until
    // it passes max(p, p') of the 8x8 sub-tile: below that every product has a
    // factor with d < 0 (block-triangular skipping, per sub-tile)
    const int span = N - max(p0, q0);
    double acc[TM][TN][2];
#pragma unroll
    for (int x = 0; x < TM; ++x)
#pragma unroll
      for (int y = 0; y < TN; ++y) acc[x][y][0] = acc[x][y][1] = 0.0;

    // K runs over (step of four samples, state), the steps outermost.  Which sub-tiles are
    // still inside their span depends on the step only, and they always form a rectangle
    // XX x YY of the warp tile (sub-tile (x, y) is live while s0 < N - max(p0 + 8x, q0 + 8y)):
    // the step picks the loop over the states that issues exactly those DMMAs -- none is
    // issued predicated-off (at N = 50 the spans are a dozen steps and most of them are in
    // the triangular tail: predicated-off DMMAs kept the tensor pipe 89 % busy at 55 % use)
    const double* ga0 = s.gz + (size_t)j * LD + kPadL + (N - 1 - t - p0 - g);
    const double* wb0 = s.wz + (size_t)jp * a.LDw + kPadL + (N - 1 - t - q0 - g);
    const long gi = (long)(m + 1) * LD, wi = (long)m * a.LDw;
    for (int s0 = 0; s0 < span; s0 += 4) {
      const int XX = min(TM, (N - s0 - p0 + 7) >> 3), YY = min(TN, (N - s0 - q0 + 7) >> 3);
      const double* ga = ga0 - s0;
      const double* wb = wb0 - s0;
      switch (XX * 8 + YY) {
        case 2 * 8 + 4: dmma_states<2, 4>(acc, ga, wb, n, gi, wi); break;
        case 2 * 8 + 3: dmma_states<2, 3>(acc, ga, wb, n, gi, wi); break;
        case 2 * 8 + 2: dmma_states<2, 2>(acc, ga, wb, n, gi, wi); break;
        case 2 * 8 + 1: dmma_states<2, 1>(acc, ga, wb, n, gi, wi); break;
        case 1 * 8 + 4: dmma_states<1, 4>(acc, ga, wb, n, gi, wi); break;
        case 1 * 8 + 3: dmma_states<1, 3>(acc, ga, wb, n, gi, wi); break;
        case 1 * 8 + 2: dmma_states<1, 2>(acc, ga, wb, n, gi, wi); break;
        case 1 * 8 + 1: dmma_states<1, 1>(acc, ga, wb, n, gi, wi); break;
        default: break;
      }
    }
    // epilogue: + Quu[j][j'] on the diagonal of the block, store
    const double ruu = s.R[j * m + jp];
#pragma unroll
    for (int x = 0; x < TM; ++x) {
      const int p = p0 + 8 * x + g;
      if (p >= N) continue;
      double* row = Pb + (size_t)(var_off<EMB>(a, j) + p) * nv + (size_t)var_off<EMB>(a, jp);
#pragma unroll
      for (int y = 0; y < TN; ++y) {
        const int c = q0 + 8 * y + 2 * t;
        if (c >= N) continue;
        double v0 = acc[x][y][0], v1 = acc[x][y][1];
        if (c == p) v0 += ruu;
        if (c + 1 == p) v1 += ruu;
        if (vec_ok && c + 1 < N) {
          __stcs(reinterpret_cast<double2*>(row + c), make_double2(v0, v1));
        } else {
          __stcs(row + c, v0);
          if (c + 1 < N) __stcs(row + c + 1, v1);
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------
// phase 3b: P by the displacement recursion
//   H[p][p'] = H[p+1][p'+1] + Gam_{N-1-p}^T W_{N-1-p'}    (per (j, j') block)
// A thread owns one diagonal delta = p' - p of all m x m blocks and walks the
// rows from p = N-1 up to 0, so that a warp writes consecutive p' of one row.
// ---------------------------------------------------------------------------
template <int M, bool EMB>
__device__ void hessian_recursion(const LtiArgs& a, const Smem& s, double* __restrict__ Pb) {
  const int n = a.n, N = a.N, LD = a.LD;
  const auto nv = row_len<EMB>(a);
  for (int dl0 = 0; dl0 < 2 * N - 1; dl0 += (int)blockDim.x) {
    const int dl = dl0 + threadIdx.x;
    const int delta = dl - (N - 1);
    const bool has = dl < 2 * N - 1;
    // rows of this diagonal, and of the warp's 32 diagonals together: the lanes
    // of a warp walk the SAME row p in every step (coalesced stores)
    const int p_hi = has ? min(N - 1, N - 1 - delta) : -1;
    const int p_lo = has ? max(0, -delta) : N;
    const int w_hi = __reduce_max_sync(0xffffffffu, p_hi);
    const int w_lo = __reduce_min_sync(0xffffffffu, p_lo);
    double acc[M][M];
#pragma unroll
    for (int x = 0; x < M; ++x)
#pragma unroll
      for (int y = 0; y < M; ++y) acc[x][y] = 0.0;
    for (int p = w_hi; p >= w_lo; --p) {
      if (p > p_hi || p < p_lo) continue;  // this diagonal has no entry in row p
      const int c = p + delta;
      const double* ga = s.gz + a.pad + (N - 1 - p);
      const double* wb = s.wz + a.pad + (N - 1 - c);
      for (int i = 0; i < n; ++i) {
        double gv[M], wv[M];
#pragma unroll
        for (int x = 0; x < M; ++x) gv[x] = ga[(i * (M + 1) + x) * LD];
#pragma unroll
        for (int y = 0; y < M; ++y) wv[y] = wb[(i * M + y) * a.LDw];
#pragma unroll
        for (int x = 0; x < M; ++x)
#pragma unroll
          for (int y = 0; y < M; ++y) acc[x][y] = fma(gv[x], wv[y], acc[x][y]);
      }
      double* dst = Pb + (size_t)p * nv + c;
#pragma unroll
      for (int x = 0; x < M; ++x)
#pragma unroll
        for (int y = 0; y < M; ++y)
          __stcs(dst + (size_t)var_off<EMB>(a, x) * nv + var_off<EMB>(a, y),
                 delta == 0 ? acc[x][y] + s.R[x * M + y] : acc[x][y]);
    }
  }
}

// general m: one (j, j') block at a time
template <bool EMB>
__device__ void hessian_recursion_any(const LtiArgs& a, const Smem& s, double* __restrict__ Pb) {
  const int n = a.n, m = a.m, N = a.N, LD = a.LD;
  const auto nv = row_len<EMB>(a);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int dchunks = (2 * N - 1 + 31) / 32;  // a warp takes 32 diagonals of one block
  for (int item = warp; item < m * m * dchunks; item += ((int)blockDim.x >> 5)) {
    const int jj = item / dchunks, dc = item - jj * dchunks, j = jj / m, jp = jj - j * m;
    const int dl = dc * 32 + lane;
    const int delta = dl - (N - 1);
    const bool has = dl < 2 * N - 1;
    const int p_hi = has ? min(N - 1, N - 1 - delta) : -1;
    const int p_lo = has ? max(0, -delta) : N;
    const int w_hi = __reduce_max_sync(0xffffffffu, p_hi);
    const int w_lo = __reduce_min_sync(0xffffffffu, p_lo);
    const double ruu = (delta == 0) ? s.R[j * m + jp] : 0.0;
    double acc = 0.0;
    for (int p = w_hi; p >= w_lo; --p) {
      if (p > p_hi || p < p_lo) continue;
      const int c = p + delta;
      for (int i = 0; i < n; ++i)
        acc = fma(s.gz[(size_t)(i * (m + 1) + j) * LD + a.pad + (N - 1 - p)],
                  s.wz[(size_t)(i * m + jp) * a.LDw + a.pad + (N - 1 - c)], acc);
      __stcs(Pb + (size_t)(var_off<EMB>(a, j) + p) * nv + (size_t)var_off<EMB>(a, jp) + c, acc + ruu);
    }
  }
}

// ---------------------------------------------------------------------------
// phase 3c: the displacement recursion, products on the tensor cores
//   T[(j,p)][(j',c)] = Gam_{N-1-p}[:,j] . W_{N-1-c}[:,j']      (a rank-n product)
//   P[(j,p)][(j',c)] = T + P[(j,p+1)][(j',c+1)]  (+ Quu[j][j'] on p == c)
// A warp owns 32 adjacent diagonals of one (j, j') block, lane = diagonal, and walks
// them bottom-up in bands of 8 rows.  Per band the warp forms the 8 x 40 patch of T
// that its diagonals cross with five DMMA.8x8x4 per 4 states (one LDS per operand,
// the tiles need no alignment because the operands are Toeplitz), parks it in its own
// 2.5 kB of shared memory, and every lane then adds its 8 entries to the running sum
// it keeps in a register: one LDS, one DADD and one coalesced store per entry of P
// instead of 2 n LDS + n DFMA.  No block-wide barrier anywhere.
// ---------------------------------------------------------------------------
constexpr int kScanLd = 40;  // 32 diagonals + 8 rows
// Entries of the patch that no active (p, c) consumes are never read back, so the
// operand loads are NOT bounds-checked: a lane may fetch up to 47 doubles before wz or
// N + 8 past a row end -- still inside this CTA's shared memory (mpci_lti_plan_create
// only enables the scan when the layout guarantees it) -- and garbage there only
// reaches its own row / column of T.  Only the padding of K to a multiple of 4 has to
// contribute exact zeros.
// KFULL: n is a multiple of 4, no K padding to mask.
template <bool KFULL, bool EMB>
__device__ void hessian_scan(const LtiArgs& a, const Smem& s, double* __restrict__ Pb) {
  const int n = a.n, m = a.m, N = a.N, LD = a.LD, w = a.m + 1;
  const auto nv = row_len<EMB>(a);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int ksteps = (n + 3) / 4;
  const int dchunks = (2 * N - 1 + 31) / 32;
  double* tw = s.tb + (size_t)warp * 8 * kScanLd;
  double2* tst = reinterpret_cast<double2*>(tw + g * kScanLd + 2 * t);  // this lane's C slots
  const double* twl = tw + lane;                                         // this lane's diagonal
  const long ga_step = (long)4 * w * LD, wb_step = (long)4 * m * a.LDw;
  // items are fetched from a shared counter: warp 0 joins late (it runs the adjoint
  // recursion of q first) and the items differ in length
  for (;;) {
    int item = 0;
    if (lane == 0) item = atomicAdd(s.next_item, 1);
    item = __shfl_sync(0xffffffffu, item, 0);
    if (item >= m * m * dchunks) break;
    const int jj = item / dchunks, dc = item - jj * dchunks, j = jj / m, jp = jj - j * m;
    const int d0 = dc * 32 - (N - 1);             // diagonal of lane 0 (delta = c - p)
    const int delta = d0 + lane;
    const int p_first = min(N - 1, N - 1 - d0);   // bottom-most row any lane touches
    const int p_last = max(0, -min(d0 + 31, N - 1));  // top-most
    // rows of this lane's diagonal: p_lo .. p_lo + p_rng (none when delta > N - 1)
    const bool has = delta <= N - 1;
    const int p_lo = has ? max(0, -delta) : N;
    const unsigned p_rng = has ? (unsigned)(min(N - 1, N - 1 - delta) - p_lo) : 0u;
    int p0 = N - 8 * ((N - 1 - p_first) / 8 + 1);  // rows p0 .. p0+7, bottom band first
    const double* ga = s.gz + (size_t)(t * w + j) * LD + a.pad + (N - 1 - p0 - g);
    const double* wb = s.wz + (size_t)(t * m + jp) * a.LDw + a.pad + (N - 1 - (p0 + d0) - g);
    double* qd = Pb + (size_t)(var_off<EMB>(a, j) + p0 + 7) * nv + (size_t)var_off<EMB>(a, jp) + (p0 + 7 + delta);
    const long qstep = (long)nv + 1;
    // the running sum of the main diagonal starts at Quu[j][j']: every entry of that
    // diagonal then carries the + Quu delta_pc term, no separate add per entry
    double carry = (delta == 0) ? s.R[j * m + jp] : 0.0;
    for (; p0 + 7 >= p_last; p0 -= 8, ga += 8, wb += 8) {
      double acc[5][2];
#pragma unroll
      for (int y = 0; y < 5; ++y) acc[y][0] = acc[y][1] = 0.0;
      const double* gk = ga;
      const double* wk = wb;
      for (int ks = 0; ks < ksteps; ++ks, gk += ga_step, wk += wb_step) {
        const bool k_ok = KFULL || 4 * ks + t < n;
        const double av = k_ok ? gk[0] : 0.0;
        double bv[5];
#pragma unroll
        for (int y = 0; y < 5; ++y) bv[y] = k_ok ? wk[-8 * y] : 0.0;
#pragma unroll
        for (int y = 0; y < 5; ++y) dmma8x8x4(acc[y][0], acc[y][1], av, bv[y]);
      }
#pragma unroll
      for (int y = 0; y < 5; ++y) tst[4 * y] = make_double2(acc[y][0], acc[y][1]);
      __syncwarp();
#pragma unroll
      for (int r = 7; r >= 0; --r) {
        if ((unsigned)(p0 + r - p_lo) <= p_rng) {
          carry += twl[r * (kScanLd + 1)];
          __stcs(qd, carry);
        }
        qd -= qstep;
      }
      __syncwarp();
    }
  }
}

// ---------------------------------------------------------------------------
// phase 3d: the displacement recursion for short horizons (N <= 60)
// The band scan above gives a lane one DIAGONAL; at N = 50 its 8 x 40 patches are 41-57 %
// full (2N - 1 = 99 diagonals in chunks of 32, triangles at both ends) and the warp spends
// its issue slots on empty slots.  Here a warp owns a whole (j, j') block and a lane two
// adjacent COLUMNS (2 lane, 2 lane + 1) of it: the band's T patch is 8 rows x N columns --
// every slot is used -- and per row of the block the lane does one LDS.128, one shuffle
// (column 2 lane + 2 of the row below lives in lane + 1), two DADDs and one 16-byte store:
//   P[p][c] = T[p][c] + P[p+1][c+1],   T[p][c] = Gam_{N-1-p}[:,j] . W_{N-1-c}[:,j']
// The + Quu[j][j'] of the main diagonal enters as the virtual entry P[N][N] = Quu.  Only
// the first a.scan_warps warps own a patch (shared memory is what limits the CTAs per SM);
// the last warp runs the adjoint recursion of q meanwhile.  Operand loads of rows p < 0
// and columns >= N are not bounds-checked: they stay inside the CTA's shared memory and
// only reach T entries nobody reads (as in hessian_scan).
// ---------------------------------------------------------------------------
// NT column tiles (8 x 8 each) of the band's T patch: K = n in steps of 4, operands straight
// from the Toeplitz generators (one LDS each), results parked in the warp's patch.
// MASKED: the group's last two tiles may reach past column N - 1; there the lanes whose
// column does not exist (ok0 / ok1 false) feed zeros, so those patch columns hold exact
// zeros (the row scan relies on that).
template <bool KFULL, int NT, bool MASKED>
__device__ __forceinline__ void rows_tiles(const double* gk, const double* wk, long ga_step, long wb_step,
                                           int n, int t, double2* dst, bool ok0, bool ok1) {
  double acc[NT][2];
#pragma unroll
  for (int y = 0; y < NT; ++y) acc[y][0] = acc[y][1] = 0.0;
  auto step = [&](bool k_ok) {
    const double av = k_ok ? gk[0] : 0.0;
    double bv[NT];
#pragma unroll
    for (int y = 0; y < NT; ++y) bv[y] = k_ok ? wk[-8 * y] : 0.0;
    if (MASKED) {
      if (NT >= 2 && !ok0) bv[NT >= 2 ? NT - 2 : 0] = 0.0;
      if (!ok1) bv[NT - 1] = 0.0;
    }
#pragma unroll
    for (int y = 0; y < NT; ++y) dmma8x8x4(acc[y][0], acc[y][1], av, bv[y]);
  };
  for (int ks = 0; ks < (n >> 2); ++ks, gk += ga_step, wk += wb_step) step(true);
  if (!KFULL) step((n & ~3) + t < n);  // the last, partial step of K
#pragma unroll
  for (int y = 0; y < NT; ++y) dst[4 * y] = make_double2(acc[y][0], acc[y][1]);
}

__device__ __forceinline__ double2 lds128(unsigned addr) {
  double2 v;
  asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];\n" : "=d"(v.x), "=d"(v.y) : "r"(addr));
  return v;
}

// CT column tiles: 8 CT - 2 >= N, so the patch keeps two zero columns after the block
template <bool KFULL, int CT, bool EMB>
__device__ void hessian_rows(const LtiArgs& a, const Smem& s, double* __restrict__ Pb) {
  constexpr int LDP = (8 * CT) % 16 == 8 ? 8 * CT : 8 * CT + 8;  // = lti_rows_ldp
  constexpr int T0 = CT > 4 ? CT / 2 : CT, T1 = CT - T0;          // two groups (register budget)
  const int n = a.n, m = a.m, N = a.N, LD = a.LD, w = a.m + 1;
  const long nv = row_len<EMB>(a);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (warp >= a.scan_warps) return;
  const int g = lane >> 2, t = lane & 3;
  double* tw = s.tb + (size_t)warp * 8 * LDP;
  double2* tst = reinterpret_cast<double2*>(tw + g * LDP + 2 * t);  // this lane's C slots of tile 0
  const long ga_step = (long)4 * w * LD, wb_step = (long)4 * m * a.LDw;
  const int col = 2 * lane;
  const bool act0 = col < N, act1 = col + 1 < N;
  // the lane's two columns; lanes past the tiles read the tiles' last two columns, which
  // are zero (8 CT - 2 >= N)
  const unsigned trow = (unsigned)__cvta_generic_to_shared(tw) + 8u * (unsigned)min(col, 8 * CT - 2);
  const bool vec = (N % 2 == 0) && ((reinterpret_cast<uintptr_t>(Pb) & 15) == 0) && (!EMB || a.emb_even);
  // columns of the last two tiles that exist
  const bool ok0 = 8 * (CT - 2) + g < N, ok1 = 8 * (CT - 1) + g < N;

  for (;;) {
    int item = 0;
    if (lane == 0) item = atomicAdd(s.next_item, 1);
    item = __shfl_sync(0xffffffffu, item, 0);
    if (item >= m * m) break;
    const int j = item / m, jp = item - j * m;
    const double ruu = s.R[j * m + jp];
    // carries = row p + 1 of the block in columns 2 lane, 2 lane + 1; the virtual P[N][N] = Quu
    double c0 = (col == N) ? ruu : 0.0, c1 = (col + 1 == N) ? ruu : 0.0;
    const double* ga = s.gz + (size_t)(t * w + j) * LD + a.pad + (7 - g);           // band p0 = N - 8
    const double* wb = s.wz + (size_t)(t * m + jp) * a.LDw + a.pad + (N - 1 - g);
    double* dst = Pb + (size_t)(var_off<EMB>(a, j) + N - 1) * nv + (size_t)var_off<EMB>(a, jp) + col;  // row N - 1
    for (int p0 = N - 8; p0 > -8; p0 -= 8, ga += 8) {  // rows p0 .. p0 + 7, bottom band first
      if (T1 > 0) {
        rows_tiles<KFULL, T0, false>(ga, wb, ga_step, wb_step, n, t, tst, true, true);
        rows_tiles<KFULL, (T1 > 0 ? T1 : 1), true>(ga, wb - 8 * T0, ga_step, wb_step, n, t, tst + 4 * T0, ok0, ok1);
      } else {
        rows_tiles<KFULL, T0, true>(ga, wb, ga_step, wb_step, n, t, tst, ok0, ok1);
      }
      __syncwarp();
      // one row of the block per step: the lane's two entries of the patch row plus the
      // carries of the row below (column 2 lane + 2 lives in lane + 1), one 16-byte store
      const int r_min = p0 < 0 ? -p0 : 0;  // the top band overhangs row 0
      if (vec && r_min == 0) {
#pragma unroll
        for (int r = 7; r >= 0; --r, dst -= nv) {
          const double2 tv = lds128(trow + 8u * (unsigned)(r * LDP));
          const double up = __shfl_down_sync(0xffffffffu, c0, 1);
          c0 = tv.x + c1;
          c1 = tv.y + up;
          if (act0) __stcs(reinterpret_cast<double2*>(dst), make_double2(c0, c1));
        }
      } else {
        for (int r = 7; r >= r_min; --r, dst -= nv) {
          const double2 tv = lds128(trow + 8u * (unsigned)(r * LDP));
          const double up = __shfl_down_sync(0xffffffffu, c0, 1);
          c0 = tv.x + c1;
          c1 = tv.y + up;
          if (vec) {
            if (act0) __stcs(reinterpret_cast<double2*>(dst), make_double2(c0, c1));
          } else {
            if (act0) __stcs(dst, c0);
            if (act1) __stcs(dst + 1, c1);
          }
        }
      }
      __syncwarp();
    }
  }
}

template <bool KFULL, bool EMB>
__device__ void hessian_rows_any(const LtiArgs& a, const Smem& s, double* __restrict__ Pb) {
  switch ((a.N + 2 + 7) >> 3) {  // column tiles: two zero columns follow the block (N <= 60)
    case 1: return hessian_rows<KFULL, 1, EMB>(a, s, Pb);
    case 2: return hessian_rows<KFULL, 2, EMB>(a, s, Pb);
    case 3: return hessian_rows<KFULL, 3, EMB>(a, s, Pb);
    case 4: return hessian_rows<KFULL, 4, EMB>(a, s, Pb);
    case 5: return hessian_rows<KFULL, 5, EMB>(a, s, Pb);
    case 6: return hessian_rows<KFULL, 6, EMB>(a, s, Pb);
    case 7: return hessian_rows<KFULL, 7, EMB>(a, s, Pb);
    default: return hessian_rows<KFULL, 8, EMB>(a, s, Pb);
  }
}

// ---------------------------------------------------------------------------
// phase 4: q[(j,p)] = sum_{d>=0} Gam_d[:,j] . r_{p+d} - tu[j]
// ---------------------------------------------------------------------------
template <bool EMB>
__device__ void gradient(const LtiArgs& a, const Smem& s, double* __restrict__ qb) {
  const int n = a.n, m = a.m, N = a.N, LD = a.LD;
  for (int e = threadIdx.x; e < m * N; e += (int)blockDim.x) {
    const int j = e / N, p = e - j * N;
    const int len = N - p;
    double acc0 = 0.0, acc1 = 0.0, acc2 = 0.0, acc3 = 0.0;  // four chains hide the DFMA latency
    for (int i = 0; i < n; ++i) {
      const double* gam = s.gz + (size_t)(i * (m + 1) + j) * LD + a.pad;  // warp-uniform reads
      const double* r = s.r + (size_t)i * N + p;
      int d = 0;
      for (; d + 3 < len; d += 4) {
        acc0 = fma(gam[d], r[d], acc0);
        acc1 = fma(gam[d + 1], r[d + 1], acc1);
        acc2 = fma(gam[d + 2], r[d + 2], acc2);
        acc3 = fma(gam[d + 3], r[d + 3], acc3);
      }
      for (; d < len; ++d) acc0 = fma(gam[d], r[d], acc0);
    }
    qb[EMB ? var_off<EMB>(a, j) + p : (long)e] = ((acc0 + acc1) + (acc2 + acc3)) - s.tu[j];
  }
}

// ---------------------------------------------------------------------------
// phase 4b: q by the adjoint (costate) recursion, one warp
//   s_p = r_p + A^T s_{p+1},   q[(j,p)] = B[:,j] . s_p - tu[j]
// which is the same sum  sum_{k>=p} Gam_{k-p}[:,j] . r_k  in O(N n^2) instead of
// O(N^2 n m) operations.  Lane i carries s[i]; the N steps are sequential but short, and
// the warp runs them while the other warps are already in the Hessian scan.  s_p
// overwrites r_p in shared memory (nothing else reads r).  Needs n <= 32.
// ---------------------------------------------------------------------------
template <bool EMB>
__device__ void gradient_costate(const LtiArgs& a, const Smem& s, double* __restrict__ qb) {
  const int n = a.n, m = a.m, N = a.N, w = a.m + 1, LD = a.LD, lane = threadIdx.x & 31;
  const bool mine = lane < n;
  double sv = 0.0;  // s_{p+1}[lane]
  for (int p = N - 1; p >= 0; --p) {
    double acc0 = mine ? s.r[lane * N + p] : 0.0, acc1 = 0.0;
    int l = 0;
    for (; l + 1 < n; l += 2) {
      const double s0 = __shfl_sync(0xffffffffu, sv, l), s1 = __shfl_sync(0xffffffffu, sv, l + 1);
      if (mine) {
        acc0 = fma(s.A[l * n + lane], s0, acc0);
        acc1 = fma(s.A[(l + 1) * n + lane], s1, acc1);
      }
    }
    if (l < n) {
      const double s0 = __shfl_sync(0xffffffffu, sv, l);
      if (mine) acc0 = fma(s.A[l * n + lane], s0, acc0);
    }
    sv = acc0 + acc1;
    if (mine) s.r[lane * N + p] = sv;
  }
  __syncwarp();
  for (int e = lane; e < m * N; e += 32) {
    const int j = e / N, p = e - j * N;
    double acc = -s.tu[j];
    for (int i = 0; i < n; ++i) acc = fma(s.gz[(i * w + j) * LD + a.pad], s.r[i * N + p], acc);
    qb[EMB ? var_off<EMB>(a, j) + p : (long)e] = acc;
  }
}

// ---------------------------------------------------------------------------
// phase 5: constraint rows.  Group g, row k, column (j, p):
//   G = arrow[k] * ((p <= k ? (cx . Gam_{k-p})[j] : 0) + (p == k ? cu[j] : 0))
//   h = bound[k] - arrow[k] * cx . f_k
// With cz[j][y] = (cx . Gam_{N-1-y})[j] for y < N-1, that plus cu[j] at y = N-1 and
// 0 for N <= y < 2N, row k of block j is arrow[k] times the window
// cz[j][N-1-k .. 2N-2-k]: a sliding copy.  When the arrow is the same for every
// row (the usual case) cz is pre-scaled and the rows are written by the TMA engine:
// one cp.async.bulk (shared -> global, N*8 bytes) per (row, block), issued by one
// lane per warp; a second copy of cz shifted by one element keeps every source
// 16-byte aligned.  Otherwise the warps scale and store the windows themselves.
// ---------------------------------------------------------------------------
__device__ __forceinline__ void bulk_store(double* dst, const double* src_smem, unsigned bytes) {
  const unsigned src = (unsigned)__cvta_generic_to_shared(src_smem);
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;\n"
               :: "l"(dst), "r"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;\n" ::: "memory"); }
template <int PENDING>
__device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;\n" :: "n"(PENDING) : "memory");
}
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;\n" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory"); }

// windows and h entries of one group, by the threads tid = 0 .. nthreads-1
__device__ __forceinline__ void group_windows(const LtiArgs& a, const Smem& s, long b, int grp, double* cz,
                                              double* cz1, bool bulk, double* __restrict__ hb, int tid,
                                              int nthreads, unsigned magic_2N) {
  const int n = a.n, m = a.m, N = a.N, czld = 2 * N;
  const int gstride = (m + 1) * a.LD;
  const bool state = (a.grp_state[grp] & 1) != 0;
  const double* arrow = a.arrow + (size_t)grp * N;
  const double* cx = a.cx + b * a.scx + (size_t)grp * n;  // warp-uniform global reads (L1 hits)
  const double* cu = a.cu + b * a.scu + (size_t)grp * m;
  const double* bound = a.bound + b * a.sbd + (size_t)grp * N;
  const int k0 = a.grp_k0[grp], k1 = k0 + a.grp_cnt[grp];
  const double scale = bulk ? __ldg(arrow + k0) : 1.0;
  for (int e = tid; e < m * 2 * N; e += nthreads) {  // y fastest
    const int j = magic_div(e, magic_2N), y = e - j * 2 * N;
    double acc = (y == N - 1) ? __ldg(cu + j) : 0.0;
    if (state && y < N) {
      const double* src = s.gz + j * a.LD + a.pad + (N - 1 - y);
#pragma unroll 4
      for (int i = 0; i < n; ++i) acc = fma(__ldg(cx + i), src[i * gstride], acc);
    }
    acc *= scale;
    cz[j * czld + y] = acc;
    if (y > 0) cz1[j * czld + y - 1] = acc;
    if (y == 2 * N - 1) cz1[j * czld + y] = 0.0;
  }
  if (hb) {
    const double* fk = s.gz + m * a.LD + a.pad + 1;  // f_k[i] = fk[i (m+1) LD + k]
    double* hg = hb + a.grp_row0[grp] - k0;  // row of sample k: grp_row0 + (k - k0)
    for (int k = k0 + tid; k < k1; k += nthreads) {
      double acc = 0.0;
      if (state) {
#pragma unroll 4
        for (int i = 0; i < n; ++i) acc = fma(__ldg(cx + i), fk[i * gstride + k], acc);
      }
      hg[k] = __ldg(bound + k) - __ldg(arrow + k) * acc;
    }
  }
}

// A group without a state part (cx = 0: input boxes and the like) whose arrow is the same on
// every row: row k is  scale * cu[j]  at column j N + k and zero elsewhere, i.e. the window
// L[N-1-k .. N-1-k + m N) of ONE line  L[y] = scale * cu[j] at y = N - 1 + j N  -- a single
// bulk copy of the whole row (m N doubles) instead of one per input block, and no window
// arithmetic at all.  L1 is the copy shifted by one element (16-byte aligned sources).
__device__ __forceinline__ void group_line(const LtiArgs& a, long b, int grp, double* L, double* L1,
                                           double* __restrict__ hb, int tid, int nthreads) {
  const int m = a.m, N = a.N, len = (m + 1) * N;
  const double* arrow = a.arrow + (size_t)grp * N;
  const double* cu = a.cu + b * a.scu + (size_t)grp * m;
  const double* bound = a.bound + b * a.sbd + (size_t)grp * N;
  const int k0 = a.grp_k0[grp], k1 = k0 + a.grp_cnt[grp];
  const double scale = __ldg(arrow + k0);
  for (int y = tid; y < len; y += nthreads) {
    double v = 0.0, v1 = 0.0;  // L[y], L[y + 1]
    for (int j = 0; j < m; ++j) {
      if (y == N - 1 + j * N) v = scale * __ldg(cu + j);
      if (y + 1 == N - 1 + j * N) v1 = scale * __ldg(cu + j);
    }
    L[y] = v;
    L1[y] = v1;
  }
  if (hb) {
    double* hg = hb + a.grp_row0[grp] - k0;
    for (int k = k0 + tid; k < k1; k += nthreads) hg[k] = __ldg(bound + k);
  }
}

// rows of a group whose arrow changes from row to row (or that are not 16-byte aligned):
// warps warp0, warp0 + nwarps, ... scale and store the windows themselves
// N scaled doubles to dst (8-byte aligned, either parity) by one warp: 16-byte stores over the
// aligned middle, single ones at an odd head / tail
__device__ __forceinline__ void warp_store_scaled(double* __restrict__ dst, const double* src, int N,
                                                  double ak, int lane) {
  const int head = (int)((reinterpret_cast<uintptr_t>(dst) >> 3) & 1);
  if (head && lane == 0) __stcs(dst, ak * src[0]);
  const int pairs = (N - head) >> 1;
  for (int i = lane; i < pairs; i += 32) {
    const int c = head + 2 * i;
    __stcs(reinterpret_cast<double2*>(dst + c), make_double2(ak * src[c], ak * src[c + 1]));
  }
  if (((N - head) & 1) && lane == 31) __stcs(dst + N - 1, ak * src[N - 1]);
}

template <bool EMB>
__device__ __forceinline__ void group_rows_by_warps(const LtiArgs& a, int grp, const double* cz,
                                                    double* __restrict__ rows, bool aligned, int warp,
                                                    int nwarps, int lane) {
  const int m = a.m, N = a.N;
  const auto nv = row_len<EMB>(a);
  const double* arrow = a.arrow + (size_t)grp * N;
  const int k0 = a.grp_k0[grp], k1 = k0 + a.grp_cnt[grp];
  for (int k = k0 + warp; k < k1; k += nwarps) {
    const double ak = __ldg(arrow + k);
    double* row = rows + (size_t)(k - k0) * nv;
    const double* win = cz + (N - 1 - k);  // column c = j N + p reads win[c + j N]
    if (EMB) {  // block j at its own column offset, whatever its parity
      for (int j = 0; j < m; ++j) warp_store_scaled(row + var_off<EMB>(a, j), win + 2 * j * N, N, ak, lane);
      continue;
    }
    if (aligned) {
      for (int c = 2 * lane; c < nv; c += 64) {  // c and c+1 share j because N is even
        int off = c;
        for (int jj = 1; jj < m; ++jj) off += (c >= jj * N) ? N : 0;
        __stcs(reinterpret_cast<double2*>(row + c), make_double2(ak * win[off], ak * win[off + 1]));
      }
    } else {
      for (int c = lane; c < nv; c += 32) {
        int off = c;
        for (int jj = 1; jj < m; ++jj) off += (c >= jj * N) ? N : 0;
        __stcs(row + c, ak * win[off]);
      }
    }
  }
}

template <bool EMB>
__device__ void constraint_rows(const LtiArgs& a, const Smem& s, long b, double* __restrict__ Gb,
                                double* __restrict__ hb) {
  const int m = a.m, N = a.N, tid = threadIdx.x;
  const auto nv = row_len<EMB>(a);
  const int lane = tid & 31, warp = tid >> 5, nwarps = (int)blockDim.x >> 5;
  const int czld = 2 * N, set_doubles = 4 * m * N;  // one set of windows: cz | cz1
  const bool aligned = (N % 2 == 0) && Gb && ((reinterpret_cast<uintptr_t>(Gb) & 15) == 0) &&
                       (!EMB || a.emb_even);
  const unsigned magic_2N = magic_of(2 * N);
  // Two sets of windows alternate, so the TMA engine drains the rows of one group while
  // the next group's windows are built (a third set was measured: no gain); whatever the
  // previous instance left in flight is waited for once, here.
  bulk_wait_read<0>();  // (bulk groups are per thread: every thread waits for its own)
  __syncthreads();

  for (int grp = 0; grp < a.n_groups; ++grp) {
    const bool bulk = aligned && (a.grp_state[grp] & 2);  // bit 1: the arrow is constant over the rows
    double* cz = s.cg + (grp & 1) * set_doubles;   // [m][2N]
    double* cz1 = cz + m * czld;                   // [m][2N]   cz1[j][y] = cz[j][y + 1]
    if (grp >= 2) {  // the copies of group grp - 2 have read this set
      bulk_wait_read<1>();
      __syncthreads();
    }
    const bool line = !EMB && bulk && !(a.grp_state[grp] & 1) && m <= 8;  // no state part: one copy per row
    if (line) group_line(a, b, grp, cz, cz + (((m + 1) * N + 1) & ~1), hb, tid, (int)blockDim.x);
    else group_windows(a, s, b, grp, cz, cz1, bulk, hb, tid, (int)blockDim.x, magic_2N);
    fence_async_smem();  // the windows must be visible to the async proxy
    __syncthreads();
    if (!Gb) continue;
    double* rows = Gb + (size_t)a.grp_row0[grp] * nv;
    const int k0 = a.grp_k0[grp], k1 = k0 + a.grp_cnt[grp];
    if (line) {
      const double* L = cz;
      const double* L1 = cz + (((m + 1) * N + 1) & ~1);
      // copies of 800 bytes and more go out fastest from one lane per warp, shorter ones
      // from all lanes (profiles/r01_tma_rate.txt)
      const int issue_lanes = (nv < 100) ? 32 : 1;
      for (int k = k0 + warp + nwarps * lane; lane < issue_lanes && k < k1; k += nwarps * issue_lanes) {
        const int start = N - 1 - k;
        bulk_store(rows + (size_t)(k - k0) * nv, (start & 1) ? L1 + (start - 1) : L + start, (unsigned)nv * 8u);
      }
      bulk_commit();
      continue;
    }
    if (bulk) {
      // Short rows (< 2 kB per copy): every lane issues, the cost is the issue itself
      // (a single issuing warp per CTA fed through named barriers was measured at C4:
      // G, h 3.0 -> 4.3 ms).  Long rows: one lane per warp, so the rows in flight stay
      // close together in memory (C4 5.87 -> 5.53 ms with 32 lanes, C5 4.58 -> 4.43 ms with 1).
      const int issue_lanes = (N < 256) ? 32 : 1;
      for (int k = k0 + warp + nwarps * lane; lane < issue_lanes && k < k1; k += nwarps * issue_lanes) {
        const int start = N - 1 - k;
        const double* win = (start & 1) ? cz1 + (start - 1) : cz + start;
        double* row = rows + (size_t)(k - k0) * nv;
        for (int j = 0; j < m; ++j) bulk_store(row + var_off<EMB>(a, j), win + j * czld, (unsigned)N * 8u);
      }
      bulk_commit();  // every lane commits (an empty group for the idle ones)
      continue;
    }
    group_rows_by_warps<EMB>(a, grp, cz, rows, aligned, warp, nwarps, lane);
    bulk_commit();  // keeps the group count in step with the bulk groups
  }
}

// ---------------------------------------------------------------------------
// phase 5, embedded layout.  The biped-sized problems that run here (N = 100, n = 6, m = 2,
// 1 MB per instance) are latency-bound, not store-bound, so this version removes the round
// trips of the generic one: the instance's cx / cu (staged in shared memory with the plant,
// one DRAM round trip for everything) and the group tables (staged once per CTA) are read
// from shared memory; groups of a row or two (the terminal box: 4 groups of ONE row) are
// written entry by entry without windows, fences or barriers.
// ---------------------------------------------------------------------------
constexpr int kShortRows = 2;

struct EmbStage {
  const double *cx, *cu;  // [G][n], [G][m] of this instance (shared memory)
  const int* tab;         // state | k0 | cnt | row0, Gs entries each (shared memory)
  int Gs;
  unsigned magic_2N;
};
__host__ __device__ inline size_t lti_emb_doubles(int n, int m, int G) {
  const size_t Gs = G > 0 ? G : 1;
  return (((size_t)G * (n + m) + (4 * Gs + 1) / 2) + 1) & ~(size_t)1;
}

__device__ void constraint_rows_emb(const LtiArgs& a, const Smem& s, const EmbStage& st, long b,
                                    double* __restrict__ Gb, double* __restrict__ hb) {
  const int n = a.n, m = a.m, N = a.N, tid = threadIdx.x, nthr = (int)blockDim.x;
  const int lane = tid & 31, warp = tid >> 5, nwarps = nthr >> 5;
  const long nv = a.ldv;
  const int czld = 2 * N, set_doubles = 4 * m * N, gstride = (m + 1) * a.LD, nu = m * N;
  const int *g_state = st.tab, *g_k0 = st.tab + st.Gs, *g_cnt = st.tab + 2 * st.Gs, *g_row0 = st.tab + 3 * st.Gs;
  const bool can_bulk = Gb != nullptr && N >= 4;
  const double* fk = s.gz + m * a.LD + a.pad + 1;  // f_k[i] = fk[i (m+1) LD + k]
  bulk_wait_read<0>();  // whatever the previous instance left in flight has read the windows
  __syncthreads();

  // groups of one or two rows: entry by entry
  for (int grp = 0; grp < a.n_groups; ++grp) {
    const int cnt = g_cnt[grp];
    if (cnt > kShortRows) continue;
    const int k0 = g_k0[grp], row0 = g_row0[grp];
    const bool state = (g_state[grp] & 1) != 0;
    const double* cx = st.cx + grp * n;
    const double* cu = st.cu + grp * m;
    const double* arrow = a.arrow + (size_t)grp * N;
    if (Gb) {
      for (int e = tid; e < cnt * nu; e += nthr) {
        const int kk = e / nu, r = e - kk * nu, j = r / N, p = r - j * N, k = k0 + kk;
        double acc = (p == k) ? cu[j] : 0.0;
        if (state && p <= k) {
          const double* src = s.gz + j * a.LD + a.pad + (k - p);
#pragma unroll 4
          for (int i = 0; i < n; ++i) acc = fma(cx[i], src[i * gstride], acc);
        }
        __stcs(Gb + (size_t)(row0 + kk) * nv + a.voff[j] + p, __ldg(arrow + k) * acc);
      }
      for (int e = tid; e < cnt * a.nz; e += nthr) {
        const int kk = e / a.nz, z = e - kk * a.nz;
        __stcs(Gb + (size_t)(row0 + kk) * nv + a.zcol[z], __ldg(a.Gz + b * a.sGz + (size_t)(row0 + kk) * a.nz + z));
      }
    }
    if (hb && tid < cnt) {
      const int k = k0 + tid;
      double acc = 0.0;
      if (state)
        for (int i = 0; i < n; ++i) acc = fma(cx[i], fk[i * gstride + k], acc);
      hb[row0 + tid] = __ldg(a.bound + b * a.sbd + (size_t)grp * N + k) - __ldg(arrow + k) * acc;
    }
  }

  int it = 0;  // window sets alternate over the groups that use them
  for (int grp = 0; grp < a.n_groups; ++grp) {
    const int cnt = g_cnt[grp];
    if (cnt <= kShortRows) continue;
    const int k0 = g_k0[grp], k1 = k0 + cnt, row0 = g_row0[grp];
    const bool state = (g_state[grp] & 1) != 0;
    const bool bulk = can_bulk && (g_state[grp] & 2);  // bit 1: one arrow for all rows
    double* cz = s.cg + (it & 1) * set_doubles;  // [m][2N]
    double* cz1 = cz + m * czld;                 // cz1[j][y] = cz[j][y + 1]
    if (it >= 2) {  // the copies of two groups ago have read this set
      bulk_wait_read<1>();
      __syncthreads();
    }
    ++it;
    const double* cx = st.cx + grp * n;
    const double* cu = st.cu + grp * m;
    const double* arrow = a.arrow + (size_t)grp * N;
    const double scale = bulk ? __ldg(arrow + k0) : 1.0;
    for (int e = tid; e < m * 2 * N; e += nthr) {  // y fastest
      const int j = magic_div(e, st.magic_2N), y = e - j * 2 * N;
      double acc = (y == N - 1) ? cu[j] : 0.0;
      if (state && y < N) {
        const double* src = s.gz + j * a.LD + a.pad + (N - 1 - y);
#pragma unroll 6
        for (int i = 0; i < n; ++i) acc = fma(cx[i], src[i * gstride], acc);
      }
      acc *= scale;
      cz[j * czld + y] = acc;
      if (y > 0) cz1[j * czld + y - 1] = acc;
      if (y == 2 * N - 1) cz1[j * czld + y] = 0.0;
    }
    if (hb) {
      const double* bound = a.bound + b * a.sbd + (size_t)grp * N;
      double* hg = hb + row0 - k0;  // row of sample k: row0 + (k - k0)
      for (int k = k0 + tid; k < k1; k += nthr) {
        double acc = 0.0;
        if (state) {
#pragma unroll 6
          for (int i = 0; i < n; ++i) acc = fma(cx[i], fk[i * gstride + k], acc);
        }
        hg[k] = __ldg(bound + k) - __ldg(arrow + k) * acc;
      }
    }
    fence_async_smem();  // the windows must be visible to the async proxy
    __syncthreads();
    if (!Gb) continue;
    double* rows = Gb + (size_t)row0 * nv;
    if (bulk) {
      // A block may start at an odd column (the biped's y block comes after N + 9 columns of
      // x): the bulk copy takes the 16-byte aligned middle of the block -- from the copy of
      // the window whose parity matches -- and the lane stores the odd head / tail itself.
      const int issue_lanes = (N < 256) ? 32 : 1;
      for (int k = k0 + warp + nwarps * lane; lane < issue_lanes && k < k1; k += nwarps * issue_lanes) {
        const int start = N - 1 - k;
        double* row = rows + (size_t)(k - k0) * nv;
        for (int j = 0; j < m; ++j) {
          double* dst = row + a.voff[j];
          const int head = (int)((reinterpret_cast<uintptr_t>(dst) >> 3) & 1);
          const int len = (N - head) & ~1, src0 = start + head;
          const double* wj = cz + j * czld;
          bulk_store(dst + head, (src0 & 1) ? cz1 + j * czld + (src0 - 1) : wj + src0, (unsigned)len * 8u);
          if (head) __stcs(dst, wj[start]);
          if (head + len < N) __stcs(dst + N - 1, wj[start + N - 1]);
        }
      }
    } else {
      group_rows_by_warps<true>(a, grp, cz, rows, false, warp, nwarps, lane);
    }
    bulk_commit();  // every thread, every group: keeps the group counts in step
    // the rows' entries under the auxiliary variables (constants of the formulation)
    for (int e = tid; e < cnt * a.nz; e += nthr) {
      const int kk = e / a.nz, z = e - kk * a.nz;
      __stcs(rows + (size_t)kk * nv + a.zcol[z], __ldg(a.Gz + b * a.sGz + (size_t)(row0 + kk) * a.nz + z));
    }
  }
}

// ---------------------------------------------------------------------------
// phase 3e, embedded layout: the blocks of P that involve the auxiliary variables z.
//   P[u,z] = sum_c w_c T_c^T Mz_c  (and its transpose P[z,u]),   P[z,z] = Pzz
// with T_c the Toeplitz operator of cost c's plant signal hs_c[j][d] = cx_c . Gam_d[:,j]
// (+ cu_c[j] at d = 0).  The columns of Mz are stored as their differences d[tau] = Mz[tau] -
// Mz[tau-1] (CSR; a footstep's column is a unit step: ONE entry), so with the running sums
// Hs[i] = hs[0] + .. + hs[i]
//   sum_{k>=p} hs[k-p] Mz[k][z] = sum_tau d[tau] (Hs[N-1-p] - Hs[max(tau,p)-p-1]):
// one term per difference instead of N - p products.  Written here, next to the plant blocks
// of the same rows, so that every 32-byte sector of P is complete when it leaves the L2 (a
// second kernel filling the 72-byte gaps later cost a DRAM read-modify-write per gap).
// hs / Hs live in the aux region (free between the Hessian scan and the constraint rows).
// ---------------------------------------------------------------------------
__device__ void coupled_hessian_emb(const LtiArgs& a, const Smem& s, long b, double* __restrict__ Pb) {
  const int n = a.n, m = a.m, N = a.N, nz = a.nz, nC = a.nC, nu = a.m * a.N;
  const int tid = threadIdx.x, nthr = (int)blockDim.x, gstride = (m + 1) * a.LD;
  const long nv = a.ldv;
  double* hs = s.cg;                  // [nC][m][N]
  double* Hs = hs + (size_t)nC * nu;  // running sums
  bulk_wait_read<0>();  // (row windows of this CTA still being read by the TMA engine)
  __syncthreads();      // the T patches / windows in aux are done with
  for (int e = tid; e < nC * nu; e += nthr) {
    const int c = e / nu, r = e - c * nu, j = r / N, d = r - j * N;
    const double* cx = a.ccx + b * a.sccx + (size_t)c * n;
    double acc = (d == 0) ? a.ccu[b * a.sccu + (size_t)c * m + j] : 0.0;
    const double* src = s.gz + j * a.LD + a.pad + d;
    for (int i = 0; i < n; ++i) acc = fma(__ldg(cx + i), src[i * gstride], acc);
    hs[e] = acc;
  }
  __syncthreads();
  {  // running sums of every (cost, input) signal: a warp per signal, shuffle scan in chunks of 32
    const int lane = tid & 31, warp = tid >> 5, nwarps = nthr >> 5;
    for (int e = warp; e < nC * m; e += nwarps) {
      const double* src = hs + (size_t)e * N;
      double* dst = Hs + (size_t)e * N;
      double carry = 0.0;
      for (int t0 = 0; t0 < N; t0 += 32) {
        double v = (t0 + lane < N) ? src[t0 + lane] : 0.0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const double up = __shfl_up_sync(0xffffffffu, v, o);
          if (lane >= o) v += up;
        }
        v += carry;
        if (t0 + lane < N) dst[t0 + lane] = v;
        carry = __shfl_sync(0xffffffffu, v, 31);
      }
    }
  }
  __syncthreads();
  auto entry = [&](int z, int j, int p) {
    double acc = 0.0;
    for (int c = 0; c < nC; ++c) {
      const double* H = Hs + (size_t)(c * m + j) * N;
      const double top = H[N - 1 - p];
      double sum = 0.0;
      for (int x = a.d_ptr[c * nz + z]; x < a.d_ptr[c * nz + z + 1]; ++x) {
        const int lo = max(a.d_tau[x], p) - p - 1;
        sum = fma(a.d_val[x], lo >= 0 ? top - H[lo] : top, sum);
      }
      acc = fma(a.ccw[b * a.sccw + c], sum, acc);
    }
    return acc;
  };
  // the rows of z (consecutive p: coalesced), then the columns of z in the plant's rows with z
  // fastest (the auxiliary columns come in runs: 72-byte pieces instead of single doubles);
  // the entry is cheap enough to be computed once per store
  for (int e = tid; e < nz * nu; e += nthr) {
    const int z = e / nu, r = e - z * nu, j = r / N, p = r - j * N;
    __stcs(Pb + (long)a.zcol[z] * nv + a.voff[j] + p, entry(z, j, p));
  }
  for (int e = tid; e < nz * nu; e += nthr) {
    const int r = e / nz, z = e - r * nz, j = r / N, p = r - j * N;
    __stcs(Pb + (long)(a.voff[j] + p) * nv + a.zcol[z], entry(z, j, p));
  }
  for (int e = tid; e < nz * nz; e += nthr) {
    const int z = e / nz, y = e - z * nz;
    __stcs(Pb + (long)a.zcol[z] * nv + a.zcol[y], __ldg(a.Pzz + b * a.sPzz + e));
  }
  __syncthreads();  // aux is free again
}

// One kernel per Hessian mode: the recursion needs neither the DMMA accumulators nor
// the zero padding, so it is compiled for more resident CTAs per SM (the phases of
// different instances -- latency-bound prediction, store-bound P / G -- then overlap).
// WIDE: plans whose shared memory allows one CTA per SM anyway run it with kWideThreads
// threads (more warps to hide the latency-bound phases; measured at C5: 4.47 -> 4.23 ms).
// EMB: the embedded layout (LtiArgs::ldv / voff / nct) -- separate instantiations, so the
// plain kernels keep their code.
template <int MODE, int WIDE, bool EMB>
__global__ void __launch_bounds__(WIDE ? kWideThreads : kThreads,
                                  WIDE ? 1 : (MODE == MPCI_LTI_HESSIAN_RECURSION ? MPCI_LTI_MINBLOCKS : 2))
lti_assemble_kernel(LtiArgs a) {
  extern __shared__ double lti_smem[];
  const Smem s = carve(lti_smem, a);
  const size_t nv = EMB ? (size_t)a.ldv : (size_t)a.m * a.N, nc = EMB ? (size_t)a.nct : (size_t)a.nc;
  [[maybe_unused]] EmbStage stage;
  if constexpr (EMB) {  // group tables: once per CTA (visible after the first barrier below)
    const int Gs = a.n_groups > 0 ? a.n_groups : 1;
    double* ecx = s.tu + a.m + 2;
    double* ecu = ecx + a.n_groups * a.n;
    int* etab = reinterpret_cast<int*>(ecu + a.n_groups * a.m);
    for (int e = threadIdx.x; e < 4 * Gs; e += (int)blockDim.x) etab[e] = a.grp_state[e];
    stage.cx = ecx; stage.cu = ecu; stage.tab = etab; stage.Gs = Gs;
    stage.magic_2N = magic_of(2 * a.N);
  }
  auto limit_rows = [&](double* Gb, double* hb, long b) {
    if constexpr (EMB) constraint_rows_emb(a, s, stage, b, Gb, hb);
    else constraint_rows<false>(a, s, b, Gb, hb);
  };
  for (long b = blockIdx.x; b < a.batch; b += gridDim.x) {
    __syncthreads();  // the previous instance is done with shared memory
    if constexpr (EMB) {  // this instance's limit forms, fetched together with the plant
      double* ecx = const_cast<double*>(stage.cx);
      double* ecu = const_cast<double*>(stage.cu);
      for (int e = threadIdx.x; e < a.n_groups * a.n; e += (int)blockDim.x) ecx[e] = a.cx[b * a.scx + e];
      for (int e = threadIdx.x; e < a.n_groups * a.m; e += (int)blockDim.x) ecu[e] = a.cu[b * a.scu + e];
    }
    if (a.n % 4 == 0) build_prediction<true>(a, s, b);
    else build_prediction<false>(a, s, b);
    // P, q and G only read the prediction data, so their order is free: every other
    // resident CTA of an SM starts with the (store-bound, TMA-driven) constraint rows
    // while its neighbours are in the (issue-bound) Hessian, instead of all of them
    // hitting the same unit in lock-step.
    const bool limits = (a.G || a.h) && a.n_groups;
    // (with one CTA per SM the alternation is between neighbouring SMs: half of them stream
    // constraint rows while the other half computes -- measured 3% at C5)
    // (three resident CTAs: two of them start with the rows -- measured at C4, 16,384 instances:
    // 4.17 ms with one of three, 4.06 with two, 4.7-4.8 with none or all; scripts/gpu_lti_sweep.sh)
    const int slot = a.per_sm == 1 ? (int)blockIdx.x : (int)(blockIdx.x / a.sms);  // resident slot
    const bool g_first = MODE == MPCI_LTI_HESSIAN_RECURSION &&
                         (a.per_sm == 3 ? (slot % 3 != 0) : (slot & 1));
    if (limits && g_first) {
      limit_rows(a.G ? a.G + (size_t)b * nc * nv : nullptr, a.h ? a.h + (size_t)b * nc : nullptr, b);
    }
    if (a.P && a.scan && limits) {  // aux: the windows must have been read before the T patches
      bulk_wait_read<0>();
      __syncthreads();
    }
    // scan mode: warp 0 produces q by the adjoint recursion, then joins the scan (short
    // problems keep the all-threads correlation: their N sequential steps would not hide
    // under the few scan items)
    const bool adjoint = MODE == MPCI_LTI_HESSIAN_RECURSION && a.scan && a.n <= 32 && a.m * a.N >= 128;
    // (rows mode: the last warp, which owns no T patch, runs it beside the Hessian warps)
    const int adjoint_warp = (a.scan == 2 && a.scan_warps < ((int)blockDim.x >> 5)) ? ((int)blockDim.x >> 5) - 1 : 0;
    if (adjoint && a.q && (threadIdx.x >> 5) == adjoint_warp) gradient_costate<EMB>(a, s, a.q + (size_t)b * nv);
    if (a.P) {
      double* Pb = a.P + (size_t)b * nv * nv;
      if (MODE == MPCI_LTI_HESSIAN_DMMA) hessian_dmma<EMB>(a, s, Pb);
      else if (a.scan == 2 && a.n % 4 == 0) hessian_rows_any<true, EMB>(a, s, Pb);
      else if (a.scan == 2) hessian_rows_any<false, EMB>(a, s, Pb);
      else if (a.scan && a.n % 4 == 0) hessian_scan<true, EMB>(a, s, Pb);
      else if (a.scan) hessian_scan<false, EMB>(a, s, Pb);
      else if (a.m == 1) hessian_recursion<1, EMB>(a, s, Pb);
      else if (a.m == 2) hessian_recursion<2, EMB>(a, s, Pb);
      else if (a.m == 3) hessian_recursion<3, EMB>(a, s, Pb);
      else if (a.m == 4) hessian_recursion<4, EMB>(a, s, Pb);
      else hessian_recursion_any<EMB>(a, s, Pb);
      if constexpr (EMB) {
        if (a.nz > 0) coupled_hessian_emb(a, s, b, Pb);
      }
    }
    if (a.q && !adjoint) gradient<EMB>(a, s, a.q + (size_t)b * nv);
    if (limits && !g_first)
      limit_rows(a.G ? a.G + (size_t)b * nc * nv : nullptr, a.h ? a.h + (size_t)b * nc : nullptr, b);
  }
  bulk_wait_all();  // bulk stores of the last rows
}

}  // namespace

// ---------------------------------------------------------------------------
// Closed forms of the named LIP systems (the reference derives them
// with SymPy, tools.py:240-285 / :328-358): one thread per instance.
// ---------------------------------------------------------------------------
namespace {
__global__ void named_system_kernel(int system, const double* __restrict__ tau, long st,
                                    const double* __restrict__ omega, long sw, double* __restrict__ A,
                                    double* __restrict__ B, long batch) {
  const long b = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= batch) return;
  const double T = tau[b * st];
  const double w = omega ? omega[b * sw] : 0.0;
  const double c = cosh(w * T), s = sinh(w * T);
  if (system == 0) {  // J->CCC: jerk -> CoM, CoM_dot, CoM_ddot
    double* a = A + b * 9;
    a[0] = 1.0; a[1] = T; a[2] = T * T / 2; a[3] = 0.0; a[4] = 1.0; a[5] = T; a[6] = 0.0; a[7] = 0.0; a[8] = 1.0;
    B[b * 3] = T * T * T / 6; B[b * 3 + 1] = T * T / 2; B[b * 3 + 2] = T;
  } else if (system == 1) {  // P->CC: CoP -> CoM, CoM_dot
    double* a = A + b * 4;
    a[0] = c; a[1] = s / w; a[2] = w * s; a[3] = c;
    B[b * 2] = 1.0 - c; B[b * 2 + 1] = -w * s;
  } else if (system == 2) {  // P->X: CoP -> DCM
    A[b] = exp(w * T);
    B[b] = 1.0 - exp(w * T);
  } else if (system == 3) {  // dP->CCC: CoP rate -> CoM, CoM_dot, CoM_ddot
    double* a = A + b * 9;
    a[0] = 1.0; a[1] = s / w; a[2] = (c - 1.0) / (w * w);
    a[3] = 0.0; a[4] = c; a[5] = s / w;
    a[6] = 0.0; a[7] = w * s; a[8] = c;
    B[b * 3] = T - s / w; B[b * 3 + 1] = 1.0 - c; B[b * 3 + 2] = -w * s;
  } else {  // dP->CCP: CoP rate -> CoM, CoM_dot, CoP
    double* a = A + b * 9;
    a[0] = c; a[1] = s / w; a[2] = 1.0 - c;
    a[3] = w * s; a[4] = c; a[5] = -w * s;
    a[6] = 0.0; a[7] = 0.0; a[8] = 1.0;
    B[b * 3] = T - s / w; B[b * 3 + 1] = 1.0 - c; B[b * 3 + 2] = T;
  }
}
}  // namespace

extern "C" int mpci_named_system(int system, const double* tau, long stride_tau, const double* omega,
                                 long stride_omega, double* A, double* B, long batch, void* stream) {
  if (system < 0 || system > 4) return mpci::fail(-1, "named_system: unknown system %d", system);
  if (batch < 0) return mpci::fail(-1, "named_system: negative batch");
  if (batch == 0) return 0;
  if (!tau || !A || !B || (system != 0 && !omega))
    return mpci::fail(-3, "named_system: NULL argument");
  named_system_kernel<<<(unsigned)((batch + 127) / 128), 128, 0, mpci::as_stream(stream)>>>(
      system, tau, stride_tau, omega, stride_omega, A, B, batch);
  MPCI_LAUNCH_CHECK("named_system_kernel");
  return 0;
}

typedef void (*lti_kernel_t)(LtiArgs);
static lti_kernel_t lti_kernel(int mode, int wide, int emb = 0) {
  if (emb) {
    if (mode == MPCI_LTI_HESSIAN_DMMA)
      return wide ? lti_assemble_kernel<MPCI_LTI_HESSIAN_DMMA, 1, true>
                  : lti_assemble_kernel<MPCI_LTI_HESSIAN_DMMA, 0, true>;
    return wide ? lti_assemble_kernel<MPCI_LTI_HESSIAN_RECURSION, 1, true>
                : lti_assemble_kernel<MPCI_LTI_HESSIAN_RECURSION, 0, true>;
  }
  if (mode == MPCI_LTI_HESSIAN_DMMA)
    return wide ? lti_assemble_kernel<MPCI_LTI_HESSIAN_DMMA, 1, false>
                : lti_assemble_kernel<MPCI_LTI_HESSIAN_DMMA, 0, false>;
  return wide ? lti_assemble_kernel<MPCI_LTI_HESSIAN_RECURSION, 1, false>
              : lti_assemble_kernel<MPCI_LTI_HESSIAN_RECURSION, 0, false>;
}

static cudaError_t lti_opt_in_smem() {
  cudaError_t e = cudaSuccess;
  for (int emb = 0; emb < 2; ++emb)
    for (int mode = 0; mode < 2; ++mode)
      for (int wide = 0; wide < 2 && e == cudaSuccess; ++wide)
        e = cudaFuncSetAttribute(lti_kernel(mode, wide, emb), cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 227 * 1024);
  return e;
}

struct mpci_lti_plan {
  int n, m, N, S, n_groups, nc;
  void* dev;  // one allocation holding every table
  LtiArgs args;  // table pointers and sizes filled in, per-call fields zero
  // per Hessian mode (index = MPCI_LTI_HESSIAN_*): the DMMA operands need kPadL zeros
  // before every generator row, the recursion does not and so fits more CTAs per SM
  int pad[2], LD[2], LDw[2], scan;  // scan: 0 plain, 1 band scan, 2 whole rows (N <= 64)
  int aux[2], ldp, scan_warps;
  int threads[2];  // CTA size per mode: kThreads, or kWideThreads when one CTA per SM fits anyway
  size_t smem_bytes[2];
  int blocks_per_sm[2];
  // staging of mpci_lti_assemble_host (two slots), allocated on first use
  long slot_capacity;
  double* slot_in[2];   // A | B | x0 | scalars
  double* slot_out[2];  // P | q | G | h
  cudaStream_t streams[2];
  cudaEvent_t slot_done[2];
  // embedded layout (mpci_lti_plan_create_embedded): total sizes and the coupling stage
  int embedded, nvt, nct;
  mpci::CouplePlan* couple;
};

static int lti_plan_create(const mpci_lti_desc* d, const mpci_lti_embed* emb, mpci_lti_plan** out);

extern "C" int mpci_lti_plan_create(const mpci_lti_desc* d, mpci_lti_plan** out) {
  return lti_plan_create(d, nullptr, out);
}

extern "C" int mpci_lti_plan_create_embedded(const mpci_lti_desc* d, const mpci_lti_embed* emb,
                                             mpci_lti_plan** out) {
  if (!emb) return mpci::fail(-3, "lti_plan_create_embedded: NULL embed");
  return lti_plan_create(d, emb, out);
}

static int lti_plan_create(const mpci_lti_desc* d, const mpci_lti_embed* emb, mpci_lti_plan** out) {
  if (!d || !out) return mpci::fail(-3, "lti_plan_create: NULL argument");
  *out = nullptr;
  if (d->n < 1 || d->m < 1 || d->N < 1 || d->n_scalars < 0 || d->n_groups < 0)
    return mpci::fail(-1, "lti_plan_create: bad sizes n=%d m=%d N=%d scalars=%d groups=%d", d->n,
                      d->m, d->N, d->n_scalars, d->n_groups);
  if (!d->Qxx || !d->Quu || !d->Tx || !d->Tu ||
      (d->n_groups && (!d->grp_cx || !d->grp_cu || !d->grp_arrow || !d->grp_bound)))
    return mpci::fail(-3, "lti_plan_create: NULL table");
  const int n = d->n, m = d->m, N = d->N, S = d->n_scalars, G = d->n_groups;
  // (coupled_hessian_emb keeps 2 n_costs m N doubles in the aux region)
  const size_t aux_floor = emb ? (size_t)2 * (emb->n_costs > 0 ? emb->n_costs : 0) * m * N : 0;
  const int pads[2] = {kPadL, 0};
  int LDs[2], LDws[2];
  size_t smems[2], auxs[2];
  for (int mode = 0; mode < 2; ++mode) {
    LDs[mode] = LDws[mode] = lti_row_len(N, pads[mode]);
    auxs[mode] = lti_aux_doubles(m, N, std::max<size_t>(0, aux_floor));
    smems[mode] = sizeof(double) * lti_smem_doubles(n, m, N, LDs[mode], LDws[mode], auxs[mode]);
  }
  // The DMMA scans of the recursion mode: skewed rows, the warps' T patches, and their
  // unchecked operand loads reach 47 doubles before wz (hessian_scan / hessian_rows).
  //   N <= 60: whole rows per warp (scan = 2); as many warps own a patch as fit without
  //            costing a resident CTA (registers allow three CTAs per SM at most)
  //   else:    32 diagonals per warp in bands of 8 rows (scan = 1)
  const int LDg_scan = lti_skew(LDs[1], m + 1), LDw_scan = lti_skew(LDs[1], m);
  int scan = 0, ldp = 0, scan_warps = 0;
  if (lti_mats_doubles(n, m) + lti_gz_doubles(n, m, LDg_scan) >= 48) {
    auto bytes_with = [&](size_t patches) {
      return sizeof(double) * lti_smem_doubles(n, m, N, LDg_scan, LDw_scan, lti_aux_doubles(m, N, std::max<size_t>(patches, aux_floor)));
    };
    if (N <= 60) {  // (lane 31 must neither own a column nor hold the virtual P[N][N])
      ldp = lti_rows_ldp(N);
      const size_t patch = (size_t)8 * ldp;
      auto ctas = [&](int sw) {
        const size_t bytes = bytes_with(sw * patch);
        return bytes > 227 * 1024 ? 0 : std::min(MPCI_LTI_MINBLOCKS, (int)(233472 / (bytes + 1024)));
      };
      const int target = ctas(4);
      if (target >= 1) {
        scan = 2;
        scan_warps = 4;
        while (scan_warps < kWarps && ctas(scan_warps + 1) >= target) ++scan_warps;
        auxs[1] = lti_aux_doubles(m, N, std::max<size_t>(scan_warps * patch, aux_floor));
      }
    }
    if (!scan && bytes_with((size_t)kWarps * 8 * 40) <= 227 * 1024) {
      scan = 1;
      auxs[1] = lti_aux_doubles(m, N, std::max<size_t>((size_t)kWarps * 8 * 40, aux_floor));
    }
    if (scan) {
      LDs[1] = LDg_scan;
      LDws[1] = LDw_scan;
      smems[1] = sizeof(double) * lti_smem_doubles(n, m, N, LDs[1], LDws[1], auxs[1]);
    }
  }
  const size_t emb_bytes = emb ? sizeof(double) * lti_emb_doubles(n, m, G) : 0;  // staged limit tables
  for (int mode = 0; mode < 2; ++mode) smems[mode] += emb_bytes;
  if (smems[MPCI_LTI_HESSIAN_RECURSION] > 227 * 1024)
    return mpci::fail(-8, "lti_plan_create: n=%d m=%d N=%d needs %zu bytes of shared memory "
                      "(limit 232448)", n, m, N, smems[MPCI_LTI_HESSIAN_RECURSION]);
  // pack the tables: Qxx | Quu | Tx | Tu | cx | cu | arrow | bound | flags(int)
  std::vector<double> host;
  auto push = [&](const double* p, size_t count) {
    const size_t off = host.size();
    host.insert(host.end(), p, p + count);
    return off;
  };
  const size_t oQ = push(d->Qxx, (size_t)n * n), oR = push(d->Quu, (size_t)m * m);
  const size_t oTx = push(d->Tx, (size_t)n * (S + 1)), oTu = push(d->Tu, (size_t)m * (S + 1));
  size_t ocx = 0, ocu = 0, oar = 0, obd = 0;
  if (G) {
    ocx = push(d->grp_cx, (size_t)G * n);
    ocu = push(d->grp_cu, (size_t)G * m);
    oar = push(d->grp_arrow, (size_t)G * N);
    obd = push(d->grp_bound, (size_t)G * N);
  }
  // DMMA warp tiles of one (j, j') block, sorted by the length of their k-loop
  const int np8 = (N + 7) / 8, RT = (np8 + TM - 1) / TM, CT = (np8 + TN - 1) / TN;
  std::vector<int> tile_rc;
  for (int span = N; span > 0; --span)
    for (int rt = 0; rt < RT; ++rt)
      for (int ct = 0; ct < CT; ++ct)
        if (N - std::max(rt * 8 * TM, ct * 8 * TN) == span) {
          tile_rc.push_back(rt);
          tile_rc.push_back(ct);
        }
  // int tables: flags | k0 | rows | row0 (G entries each, at least one), then the DMMA tiles
  const int Gs = G > 0 ? G : 1;
  std::vector<int> flags(4 * (size_t)Gs, 0);
  int nc = 0;
  for (int g = 0; g < G; ++g) {
    const int k0 = d->grp_k0 ? d->grp_k0[g] : 0, cnt = d->grp_rows ? d->grp_rows[g] : N;
    if (k0 < 0 || cnt < 0 || k0 + cnt > N)
      return mpci::fail(-1, "lti_plan_create: group %d has rows %d .. %d outside the horizon %d", g, k0,
                        k0 + cnt - 1, N);
    for (int i = 0; i < n; ++i)
      if (d->grp_cx[(size_t)g * n + i] != 0.0) flags[g] |= 1;
    bool same = true;  // one arrow for all rows: the rows can be written as bulk copies
    for (int k = k0 + 1; k < k0 + cnt; ++k)
      same = same && d->grp_arrow[(size_t)g * N + k] == d->grp_arrow[(size_t)g * N + k0];
    if (same) flags[g] |= 2;
    flags[Gs + g] = k0;
    flags[2 * Gs + g] = cnt;
    flags[3 * Gs + g] = nc;
    if (emb) {  // the group's rows inside the larger G / h
      if (!emb->grp_row0) return mpci::fail(-3, "lti_plan_create_embedded: NULL grp_row0");
      const int r0 = emb->grp_row0[g];
      if (r0 < 0 || r0 + cnt > emb->nc_total)
        return mpci::fail(-1, "lti_plan_create_embedded: group %d at rows %d .. %d of %d", g, r0,
                          r0 + cnt - 1, emb->nc_total);
      flags[3 * Gs + g] = r0;
    }
    nc += cnt;
  }
  mpci::CouplePlan* couple = nullptr;
  if (emb) {
    const int rc = mpci::couple_create(n, m, N, emb, &couple);  // validates the layout as well
    if (rc) return rc;
  }
  mpci_lti_plan* p = new mpci_lti_plan();  // value-initialised: every member zero
  p->couple = couple;
  if (emb) {  // voff goes in front of the group tables' tail: flags | voff(m) | tiles
    flags.insert(flags.end(), emb->var_off, emb->var_off + m);
  }
  const size_t n_flags = flags.size();
  flags.insert(flags.end(), tile_rc.begin(), tile_rc.end());  // flags | tiles (int2), 8-byte aligned
  if (n_flags % 2) flags.insert(flags.begin() + n_flags, 0);
  const size_t tiles_off = (n_flags + 1) / 2 * 2;
  const size_t bytes_d = host.size() * sizeof(double), bytes_i = flags.size() * sizeof(int);
  cudaError_t e = cudaMalloc(&p->dev, bytes_d + bytes_i);
  if (e == cudaSuccess) e = cudaMemcpy(p->dev, host.data(), bytes_d, cudaMemcpyHostToDevice);
  if (e == cudaSuccess)
    e = cudaMemcpy(static_cast<char*>(p->dev) + bytes_d, flags.data(), bytes_i,
                   cudaMemcpyHostToDevice);
  if (e == cudaSuccess)  // the largest opt-in size: plans of different sizes share the kernels
    e = lti_opt_in_smem();
  int per_sm[2] = {0, 0}, threads[2] = {kThreads, kThreads};
  for (int mode = 0; mode < 2 && e == cudaSuccess; ++mode) {
    if (smems[mode] > 227 * 1024) continue;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm[mode], lti_kernel(mode, 0, emb ? 1 : 0), kThreads,
                                                      smems[mode]);
    if (e != cudaSuccess || per_sm[mode] > 1) continue;
    // one CTA per SM: give it more warps if its shared memory (larger T-patch area) still fits
    size_t aux_wide = auxs[mode];  // the band scan keeps one T patch per warp
    if (mode == MPCI_LTI_HESSIAN_RECURSION && scan == 1)
      aux_wide = lti_aux_doubles(m, N, std::max<size_t>((size_t)(kWideThreads / 32) * 8 * 40, aux_floor));
    const size_t wide = sizeof(double) * lti_smem_doubles(n, m, N, LDs[mode], LDws[mode], aux_wide) + emb_bytes;
    int fit = 0;
    if (wide <= 227 * 1024)
      e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&fit, lti_kernel(mode, 1, emb ? 1 : 0), kWideThreads, wide);
    if (e == cudaSuccess && fit >= 1) {
      threads[mode] = kWideThreads;
      smems[mode] = wide;
      auxs[mode] = aux_wide;
      per_sm[mode] = fit;
    }
  }
  if (e != cudaSuccess) {
    int rc = mpci::cuda_fail(e, "lti_plan_create");
    if (p->dev) cudaFree(p->dev);
    mpci::couple_destroy(p->couple);
    delete p;
    return rc;
  }
  const double* base = static_cast<const double*>(p->dev);
  p->n = n; p->m = m; p->N = N; p->S = S; p->n_groups = G; p->nc = nc;
  LtiArgs& a = p->args;
  a.n = n; a.m = m; a.N = N; a.S = S; a.n_groups = G; a.nc = nc;
  a.Qxx = base + oQ; a.Quu = base + oR; a.Tx = base + oTx; a.Tu = base + oTu;
  a.cx = base + ocx; a.cu = base + ocu; a.arrow = base + oar; a.bound = base + obd;
  a.grp_state = reinterpret_cast<const int*>(static_cast<const char*>(p->dev) + bytes_d);
  a.grp_k0 = a.grp_state + Gs;
  a.grp_cnt = a.grp_state + 2 * Gs;
  a.grp_row0 = a.grp_state + 3 * Gs;
  a.tiles = reinterpret_cast<const int2*>(a.grp_state + tiles_off);
  a.n_tiles = (int)tile_rc.size() / 2;
  if (emb) {
    p->embedded = 1;
    p->nvt = emb->nv_total;
    p->nct = emb->nc_total;
    a.ldv = emb->nv_total;
    a.nct = emb->nc_total;
    a.voff = a.grp_state + 4 * Gs;
    mpci::CoupleShare sh;
    mpci::couple_share(p->couple, &sh);
    a.Gz = sh.Gz; a.zcol = sh.z_col; a.nz = sh.nz;
    a.nC = sh.n_costs; a.ccx = sh.cost_cx; a.ccu = sh.cost_cu; a.ccw = sh.cost_w;
    a.d_ptr = sh.d_ptr; a.d_tau = sh.d_tau; a.d_val = sh.d_val; a.Pzz = sh.Pzz;
    a.sccx = a.sccu = a.sccw = a.sPzz = a.sGz = 0;
    a.emb_even = (emb->nv_total % 2 == 0);
    for (int j = 0; j < m; ++j) a.emb_even = a.emb_even && (emb->var_off[j] % 2 == 0);
  }
  p->scan = scan;
  p->ldp = ldp;
  p->scan_warps = scan_warps;
  for (int mode = 0; mode < 2; ++mode) {
    p->aux[mode] = (int)auxs[mode];
    p->pad[mode] = pads[mode];
    p->LD[mode] = LDs[mode];
    p->threads[mode] = threads[mode];
    p->LDw[mode] = LDws[mode];
    p->smem_bytes[mode] = smems[mode];
    p->blocks_per_sm[mode] = per_sm[mode] < 1 ? 1 : per_sm[mode];
  }
  if (emb)  // coupled_hessian_emb keeps 2 n_costs m N doubles in the aux region
    for (int mode = 0; mode < 2; ++mode)
      if ((size_t)2 * a.nC * m * N > (size_t)p->aux[mode]) {
        mpci_lti_plan_destroy(p);
        return mpci::fail(-8, "lti_plan_create_embedded: %d coupled costs need %zu doubles of work space, the "
                          "plan has %d", a.nC, (size_t)2 * a.nC * m * N, p->aux[mode]);
      }
  *out = p;
  return 0;
}

extern "C" int mpci_lti_plan_destroy(mpci_lti_plan* p) {
  if (!p) return 0;
  for (int s = 0; s < 2; ++s) {
    if (p->slot_in[s]) cudaFree(p->slot_in[s]);
    if (p->slot_out[s]) cudaFree(p->slot_out[s]);
    if (p->streams[s]) cudaStreamDestroy(p->streams[s]);
    if (p->slot_done[s]) cudaEventDestroy(p->slot_done[s]);
  }
  if (p->dev) cudaFree(p->dev);
  mpci::couple_destroy(p->couple);
  delete p;
  return 0;
}

extern "C" double mpci_lti_hessian_flops(const mpci_lti_plan* p, int mode) {
  if (!p) return 0.0;
  const double n = p->n, m = p->m, N = p->N;
  if (mode == MPCI_LTI_HESSIAN_RECURSION) return 2.0 * n * m * m * N * N + 2.0 * n * n * m * N;
  // executed by hessian_dmma: per warp tile n * (Nk - kstart)/4 steps of TM*TN DMMAs (512 flop)
  const int np8 = (p->N + 7) / 8, RT = (np8 + TM - 1) / TM, CT = (np8 + TN - 1) / TN;
  double dmmas = 0.0;  // per (i, j, j'): one DMMA per 8x8 sub-tile and step of 4 inside its span
  for (int rt = 0; rt < RT; ++rt)
    for (int ct = 0; ct < CT; ++ct)
      for (int x = 0; x < TM; ++x)
        for (int y = 0; y < TN; ++y) {
          const int ps = rt * 8 * TM + 8 * x, qs = ct * 8 * TN + 8 * y;
          const int span = p->N - (ps > qs ? ps : qs);
          if (span > 0) dmmas += (span + 3) / 4;
        }
  return dmmas * n * m * m * 512.0 + 2.0 * n * n * m * N;
}

extern "C" int mpci_lti_plan_info(const mpci_lti_plan* p, int mode, long info[6]) {
  if (!p || !info) return mpci::fail(-3, "lti_plan_info: NULL argument");
  if (mode != MPCI_LTI_HESSIAN_DMMA && mode != MPCI_LTI_HESSIAN_RECURSION)
    return mpci::fail(-1, "lti_plan_info: unknown hessian_mode %d", mode);
  info[0] = (long)p->smem_bytes[mode];
  info[1] = p->blocks_per_sm[mode];
  info[2] = mode == MPCI_LTI_HESSIAN_RECURSION ? p->scan : 0;
  info[3] = p->LD[mode];
  info[4] = p->LDw[mode];
  info[5] = p->threads[mode];
  return 0;
}

static int lti_launch(const mpci_lti_plan* p, const double* A, long strideA, const double* B,
                      long strideB, const double* x0, const double* scalars,
                      const mpci_lti_tables* tab, double* P, double* q, double* G, double* h,
                      long batch, int mode, void* stream, bool coupled = false,
                      const mpci_lti_coupled_tables* ct = nullptr);

extern "C" int mpci_lti_assemble(const mpci_lti_plan* p, const double* A, long strideA,
                                 const double* B, long strideB, const double* x0,
                                 const double* scalars, double* P, double* q, double* G,
                                 double* h, long batch, int mode, void* stream) {
  return lti_launch(p, A, strideA, B, strideB, x0, scalars, nullptr, P, q, G, h, batch, mode, stream);
}

extern "C" int mpci_lti_assemble_ex(const mpci_lti_plan* p, const double* A, long strideA,
                                    const double* B, long strideB, const double* x0,
                                    const double* scalars, const mpci_lti_tables* tab, double* P,
                                    double* q, double* G, double* h, long batch, int mode,
                                    void* stream) {
  if (p && tab && (tab->tx || tab->tu) && p->S != 0)
    return mpci::fail(-1, "lti_assemble_ex: per-instance tx / tu need a plan with n_scalars = 0");
  return lti_launch(p, A, strideA, B, strideB, x0, scalars, tab, P, q, G, h, batch, mode, stream);
}

extern "C" int mpci_lti_assemble_coupled(const mpci_lti_plan* p, const double* A, long strideA,
                                         const double* B, long strideB, const double* x0,
                                         const double* given, const double* scalars,
                                         const mpci_lti_tables* tab, const double* cost_cx,
                                         const double* cost_cu, double* P, double* q, double* G,
                                         double* h, long batch, int mode, void* stream) {
  mpci_lti_coupled_tables ct{};
  ct.cost_cx = cost_cx;
  ct.cost_cu = cost_cu;
  return mpci_lti_assemble_coupled_ex(p, A, strideA, B, strideB, x0, given, scalars, tab, &ct, P, q, G, h, batch,
                                      mode, stream);
}

extern "C" int mpci_lti_assemble_coupled_ex(const mpci_lti_plan* p, const double* A, long strideA,
                                            const double* B, long strideB, const double* x0,
                                            const double* given, const double* scalars,
                                            const mpci_lti_tables* tab, const mpci_lti_coupled_tables* ct,
                                            double* P, double* q, double* G, double* h, long batch,
                                            int mode, void* stream) {
  if (!p) return mpci::fail(-3, "lti_assemble_coupled: NULL plan");
  if (!p->embedded)
    return mpci::fail(-1, "lti_assemble_coupled: the plan was not created by mpci_lti_plan_create_embedded");
  if (tab && (tab->tx || tab->tu) && p->S != 0)
    return mpci::fail(-1, "lti_assemble_coupled: per-instance tx / tu need a plan with n_scalars = 0");
  if (ct && ct->cost_w && !ct->Pzz) {
    mpci::CoupleShare sh;
    mpci::couple_share(p->couple, &sh);
    if (sh.nz > 0)
      return mpci::fail(-1, "lti_assemble_coupled: per-instance cost_w needs the per-instance Pzz of those weights");
  }
  int rc = lti_launch(p, A, strideA, B, strideB, x0, scalars, tab, P, q, G, h, batch, mode, stream, true, ct);
  if (rc) return rc;
  return mpci::couple_launch(p->couple, A, strideA, B, strideB, x0, given, ct, q, G, h, batch,
                             mpci::as_stream(stream));
}

static int lti_launch(const mpci_lti_plan* p, const double* A, long strideA, const double* B,
                      long strideB, const double* x0, const double* scalars,
                      const mpci_lti_tables* tab, double* P, double* q, double* G, double* h,
                      long batch, int mode, void* stream, bool coupled,
                      const mpci_lti_coupled_tables* ct) {
  if (!p) return mpci::fail(-3, "lti_assemble: NULL plan");
  if (p->embedded && !coupled)
    return mpci::fail(-1, "lti_assemble: a plan with the embedded layout runs through mpci_lti_assemble_coupled");
  if (batch < 0) return mpci::fail(-1, "lti_assemble: negative batch");
  if (batch == 0) return 0;
  if (!A || !B || !x0) return mpci::fail(-3, "lti_assemble: NULL A, B or x0");
  if (p->S > 0 && !scalars) return mpci::fail(-3, "lti_assemble: NULL scalars");
  if (mode != MPCI_LTI_HESSIAN_DMMA && mode != MPCI_LTI_HESSIAN_RECURSION)
    return mpci::fail(-1, "lti_assemble: unknown hessian_mode %d", mode);
  if ((strideA != 0 && strideA < (long)p->n * p->n) || (strideB != 0 && strideB < (long)p->n * p->m))
    return mpci::fail(-4, "lti_assemble: instance strides smaller than the matrices");
  if (p->smem_bytes[mode] > 227 * 1024)
    return mpci::fail(-8, "lti_assemble: n=%d m=%d N=%d needs %zu bytes of shared memory in the "
                      "DMMA mode (limit 232448); use MPCI_LTI_HESSIAN_RECURSION", p->n, p->m, p->N,
                      p->smem_bytes[mode]);
  LtiArgs a = p->args;
  a.pad = p->pad[mode]; a.LD = p->LD[mode]; a.LDw = p->LDw[mode];
  a.scan = (mode == MPCI_LTI_HESSIAN_RECURSION) ? p->scan : 0;
  a.aux = p->aux[mode]; a.ldp = p->ldp; a.scan_warps = p->scan_warps;
  a.A = A; a.B = B; a.x0 = x0; a.scalars = scalars; a.strideA = strideA; a.strideB = strideB;
  a.P = P; a.q = q; a.G = G; a.h = h; a.batch = batch; a.hessian_mode = mode;
  a.sQ = a.sR = a.sTx = a.sTu = a.scx = a.scu = a.sbd = 0;
  if (ct && ct->cost_cx) { a.ccx = ct->cost_cx; a.sccx = (long)a.nC * p->n; }
  if (ct && ct->cost_cu) { a.ccu = ct->cost_cu; a.sccu = (long)a.nC * p->m; }
  if (ct && ct->cost_w) { a.ccw = ct->cost_w; a.sccw = a.nC; }
  if (ct && ct->Pzz) { a.Pzz = ct->Pzz; a.sPzz = (long)a.nz * a.nz; }
  if (ct && ct->Gz) { a.Gz = ct->Gz; a.sGz = (long)a.nct * a.nz; }
  if (tab) {  // per-instance tables replace the plan's
    const long n = p->n, m = p->m, Gn = p->n_groups, N = p->N;
    if (tab->Qxx) { a.Qxx = tab->Qxx; a.sQ = n * n; }
    if (tab->Quu) { a.Quu = tab->Quu; a.sR = m * m; }
    if (tab->tx) { a.Tx = tab->tx; a.sTx = n; }
    if (tab->tu) { a.Tu = tab->tu; a.sTu = m; }
    if (tab->cx) { a.cx = tab->cx; a.scx = Gn * n; }
    if (tab->cu) { a.cu = tab->cu; a.scu = Gn * m; }
    if (tab->bound) { a.bound = tab->bound; a.sbd = Gn * N; }
  }
  int dev = 0, sms = 148;
  MPCI_CUDA(cudaGetDevice(&dev));
  MPCI_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  MPCI_CUDA(lti_opt_in_smem());  // per device: the plan may be used on another GPU
  a.sms = sms;
  a.per_sm = p->blocks_per_sm[mode];
  const long resident = (long)sms * a.per_sm;
  const unsigned grid = (unsigned)(batch < resident ? batch : resident);
  lti_kernel(mode, p->threads[mode] == kWideThreads, p->embedded)
      <<<grid, p->threads[mode], p->smem_bytes[mode], mpci::as_stream(stream)>>>(a);
  MPCI_LAUNCH_CHECK("lti_assemble_kernel");
  return 0;
}

// Host buffers: chunks of instances alternate between two slots / streams, so the
// upload of one chunk and the download of the previous one overlap the kernel.
// ring == 0: the host arrays hold the whole batch; ring >= 1: `ring` chunks, chunk c in
// slot c % ring, `on_chunk` when a chunk is complete (see mpci_assemble_host_stream).
static int lti_host(mpci_lti_plan* p, const double* A, long strideA, const double* B, long strideB,
                    const double* x0, const double* scalars, double* P, double* q, double* G,
                    double* h, long batch, int mode, long chunk, long ring, mpci_chunk_fn on_chunk,
                    void* user) {
  if (batch < 0 || chunk < 1) return mpci::fail(-1, "lti_assemble_host: bad batch / chunk");
  if (p->embedded)
    return mpci::fail(-1, "lti_assemble_host: plans with the embedded layout have no host-buffer entry point");
  if (batch == 0) return 0;
  if (!A || !B || !x0 || (p->S > 0 && !scalars))
    return mpci::fail(-3, "lti_assemble_host: NULL input");
  const long n = p->n, m = p->m, S = p->S;
  if ((strideA != 0 && strideA < n * n) || (strideB != 0 && strideB < n * m))
    return mpci::fail(-4, "lti_assemble_host: instance strides smaller than the matrices");
  if (chunk > batch) chunk = batch;
  const long n_chunks = (batch + chunk - 1) / chunk;
  if (ring == 0) ring = n_chunks;
  if (ring < 2 && n_chunks > 1)
    return mpci::fail(-1, "lti_assemble_host: a ring of one slot cannot overlap two chunks");
  const long nv = m * p->N, nc = p->nc;
  const long in_per = n * n + n * m + n + S, out_per = nv * nv + nv + nc * nv + nc;
  if (p->slot_capacity < chunk) {
    for (int s = 0; s < 2; ++s) {
      if (p->slot_in[s]) cudaFree(p->slot_in[s]);
      if (p->slot_out[s]) cudaFree(p->slot_out[s]);
      p->slot_in[s] = p->slot_out[s] = nullptr;
    }
    p->slot_capacity = 0;
    for (int s = 0; s < 2; ++s) {
      MPCI_CUDA(cudaMalloc(&p->slot_in[s], sizeof(double) * in_per * chunk));
      MPCI_CUDA(cudaMalloc(&p->slot_out[s], sizeof(double) * out_per * chunk));
      if (!p->streams[s])
        MPCI_CUDA(cudaStreamCreateWithFlags(&p->streams[s], cudaStreamNonBlocking));
      if (!p->slot_done[s])
        MPCI_CUDA(cudaEventCreateWithFlags(&p->slot_done[s], cudaEventDisableTiming));
    }
    p->slot_capacity = chunk;
  }
  for (long c = 0; c <= n_chunks; ++c) {
    if (c < n_chunks) {
      const int slot = (int)(c & 1);
      const long b0 = c * chunk, nb = (batch - b0 < chunk) ? batch - b0 : chunk;
      const long at = (c % ring) * chunk;
      cudaStream_t st = p->streams[slot];
      double* d_A = p->slot_in[slot];
      double* d_B = d_A + n * n * chunk;
      double* d_x = d_B + n * m * chunk;
      double* d_s = d_x + n * chunk;
      double* d_P = p->slot_out[slot];
      double* d_q = d_P + nv * nv * chunk;
      double* d_G = d_q + nv * chunk;
      double* d_h = d_G + nc * nv * chunk;
      // shared dynamics (stride 0): one matrix per chunk; per-instance: a strided copy
      if (strideA == 0)
        MPCI_CUDA(cudaMemcpyAsync(d_A, A, sizeof(double) * n * n, cudaMemcpyHostToDevice, st));
      else
        MPCI_CUDA(cudaMemcpy2DAsync(d_A, sizeof(double) * n * n, A + b0 * strideA,
                                    sizeof(double) * strideA, sizeof(double) * n * n, nb,
                                    cudaMemcpyHostToDevice, st));
      if (strideB == 0)
        MPCI_CUDA(cudaMemcpyAsync(d_B, B, sizeof(double) * n * m, cudaMemcpyHostToDevice, st));
      else
        MPCI_CUDA(cudaMemcpy2DAsync(d_B, sizeof(double) * n * m, B + b0 * strideB,
                                    sizeof(double) * strideB, sizeof(double) * n * m, nb,
                                    cudaMemcpyHostToDevice, st));
      MPCI_CUDA(cudaMemcpyAsync(d_x, x0 + b0 * n, sizeof(double) * nb * n, cudaMemcpyHostToDevice, st));
      if (S)
        MPCI_CUDA(cudaMemcpyAsync(d_s, scalars + b0 * S, sizeof(double) * nb * S,
                                  cudaMemcpyHostToDevice, st));
      int rc = mpci_lti_assemble(p, d_A, strideA ? n * n : 0, d_B, strideB ? n * m : 0, d_x,
                                 S ? d_s : nullptr, P ? d_P : nullptr, q ? d_q : nullptr,
                                 G ? d_G : nullptr, h ? d_h : nullptr, nb, mode, st);
      if (rc) return rc;
      if (P) MPCI_CUDA(cudaMemcpyAsync(P + at * nv * nv, d_P, sizeof(double) * nb * nv * nv, cudaMemcpyDeviceToHost, st));
      if (q) MPCI_CUDA(cudaMemcpyAsync(q + at * nv, d_q, sizeof(double) * nb * nv, cudaMemcpyDeviceToHost, st));
      if (G) MPCI_CUDA(cudaMemcpyAsync(G + at * nc * nv, d_G, sizeof(double) * nb * nc * nv, cudaMemcpyDeviceToHost, st));
      if (h) MPCI_CUDA(cudaMemcpyAsync(h + at * nc, d_h, sizeof(double) * nb * nc, cudaMemcpyDeviceToHost, st));
      MPCI_CUDA(cudaEventRecord(p->slot_done[slot], st));
    }
    if (c >= 1) {
      const long done = c - 1, b0 = done * chunk;
      MPCI_CUDA(cudaEventSynchronize(p->slot_done[done & 1]));
      if (on_chunk) on_chunk(user, b0, (batch - b0 < chunk) ? batch - b0 : chunk, done % ring);
    }
  }
  return 0;
}

extern "C" int mpci_lti_assemble_host(mpci_lti_plan* p, const double* A, long strideA,
                                      const double* B, long strideB, const double* x0,
                                      const double* scalars, double* P, double* q, double* G,
                                      double* h, long batch, int mode, long chunk) {
  if (!p) return mpci::fail(-3, "lti_assemble_host: NULL plan");
  return lti_host(p, A, strideA, B, strideB, x0, scalars, P, q, G, h, batch, mode, chunk, 0,
                  nullptr, nullptr);
}

extern "C" int mpci_lti_assemble_host_stream(mpci_lti_plan* p, const double* A, long strideA,
                                             const double* B, long strideB, const double* x0,
                                             const double* scalars, double* P, double* q,
                                             double* G, double* h, long batch, int mode,
                                             long chunk, long ring, mpci_chunk_fn on_chunk,
                                             void* user) {
  if (!p) return mpci::fail(-3, "lti_assemble_host_stream: NULL plan");
  if (ring < 1) return mpci::fail(-1, "lti_assemble_host_stream: ring must hold at least one chunk");
  return lti_host(p, A, strideA, B, strideB, x0, scalars, P, q, G, h, batch, mode, chunk, ring,
                  on_chunk, user);
}
