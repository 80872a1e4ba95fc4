// Coupling stage of an embedded per-instance plan (include/mpci.h: mpci_lti_embed).
//
// The fused kernel of mpci_lti.cu writes everything of one instance's QP that involves only
// the plant, ALL of P (the blocks that involve the auxiliary variables included: it has the
// prediction in shared memory and writes whole 32-byte sectors while the rows are in the L2)
// and the rows of G that have a plant part.  What is left of q, G, h -- the parts that involve
// the auxiliary optimisation variables z (the biped's footstep displacements Ds) or the
// auxiliary given variables (s0, the bias n) -- is written here, one CTA per instance:
//   1. V_d = A^d [B | x0], d = 0..N, by prefix doubling (plain FP64 FMAs: n <= a dozen)
//   2. per coupled cost c: hs_c[j][d] = cx_c . Gam_d[:,j] (+ cu_c[j] at d = 0),
//      fr_c[k] = cx_c . A^{k+1} x0,  e_c = E_c [given; 1]
//   3. q[z] = sum_c w_c Mz_c[:,z] . (fr_c + e_c),  q[u] += sum_c w_c T_c^T e_c,  the rows of G
//      without a plant part (whole rows),  h (+)= Hg [given; 1]
// Reference: generate_qp_cost / generate_qp_constraint (python/mpc_interface/body.py:236-302)
// on preview matrices whose optimisation block is [T_c | Mz_c] and whose given block is
// [S_c | Mg_c] (make_preview_matrices, body.py:149-193).
#include <algorithm>
#include <cstdint>
#include <vector>

#include "mpci_couple.cuh"

namespace {

constexpr int kThreads = 256;

struct CoupleArgs {
  int n, m, N, nz, nC, glen, nvt, nct, n_aux_rows;

  const int *voff, *zcol;
  const double *ccx, *ccu;  // [nC][n], [nC][m] (+ instance strides)
  long sccx, sccu, scw, sGz, shv;     // (scw: stride of the per-instance weights and aims, 0 = the plan's / none)
  const double* caim;       // [batch][nC] per-instance aims, or nullptr
  const double *cw, *MzT;   // [nC], [nC][nz][N]
  const unsigned char* act; // [nC][nz]: column z of Mz_c is not identically zero
  const int *d_ptr, *d_tau; // CSR over (c, z) of the column differences Mz[tau] - Mz[tau-1]
  const double* d_val;
  const int *e_ptr, *e_col;
  const double* e_val;
  const double *Pzz, *Gz;
  const int *h_ptr, *h_col;
  const double* h_val;
  const int *row_kind, *aux_rows;
  const double *A, *B, *x0, *given;
  long sA, sB;
  double *q, *G, *h;
  long batch;
};

// shared memory (doubles): Mp, Mq [n][n] | V [N+1][n][m+1] | hs [nC][m][N] | eps, ee [nC][N]
// | gv [glen + 1] ; then ints: colmap [nvt] | hsact [nC m]
__host__ __device__ inline size_t couple_doubles(int n, int m, int N, int nz, int nC, int glen) {
  return (size_t)2 * n * n + (size_t)(N + 1) * n * (m + 1) + (size_t)nC * m * N + (size_t)2 * nC * N +
         (size_t)glen + 1;
}

__global__ void __launch_bounds__(kThreads) lti_couple_kernel(CoupleArgs a) {
  extern __shared__ double couple_smem[];
  const int n = a.n, m = a.m, N = a.N, nz = a.nz, nC = a.nC, w = a.m + 1, nu = a.m * a.N;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nthr = (int)blockDim.x, nwarps = nthr >> 5;
  double* Mp = couple_smem;
  double* Mq = Mp + n * n;
  double* V = Mq + n * n;                 // V[d][i][c]
  double* hs = V + (size_t)(N + 1) * n * w;  // hs[c][j][d]
  double* eps = hs + (size_t)nC * nu;     // fr + e
  double* ee = eps + (size_t)nC * N;      // e
  double* gv = ee + (size_t)nC * N;       // [given; 1]
  int* colmap = reinterpret_cast<int*>(gv + a.glen + 1);
  int* hsact = colmap + a.nvt;

  // column -> what lives there: j N + p for the plant's inputs, -1 - z for an auxiliary variable
  for (int c = tid; c < a.nvt; c += nthr) colmap[c] = INT32_MIN;
  __syncthreads();
  for (int e = tid; e < nu; e += nthr) colmap[a.voff[e / N] + e % N] = e;
  for (int z = tid; z < nz; z += nthr) colmap[a.zcol[z]] = -1 - z;
  __syncthreads();

  for (long b = blockIdx.x; b < a.batch; b += gridDim.x) {
    const double* Ab = a.A + b * a.sA;
    const double* Bb = a.B + b * a.sB;
    // ---- 1. prediction ---------------------------------------------------------------
    for (int e = tid; e < n * n; e += nthr) Mp[e] = Ab[e];
    for (int e = tid; e < n * w; e += nthr) {
      const int i = e / w, c = e - i * w;
      V[e] = (c < m) ? Bb[i * m + c] : a.x0[b * n + i];
    }
    for (int e = tid; e < a.glen; e += nthr) gv[e] = a.given[b * a.glen + e];
    if (tid == 0) gv[a.glen] = 1.0;
    for (int e = tid; e < nC * m; e += nthr) hsact[e] = 0;
    __syncthreads();
    double* P0 = Mp;
    double* P1 = Mq;
    const int slab = n * w;
    for (int have = 1; have < N + 1;) {
      const int cnt = min(have, N + 1 - have);
      for (int e = tid; e < cnt * slab; e += nthr) {
        const int d = e / slab, r = e - d * slab, i = r / w, c = r - i * w;
        const double* src = V + (size_t)d * slab + c;
        double acc = 0.0;
        for (int l = 0; l < n; ++l) acc = fma(P0[i * n + l], src[l * w], acc);
        V[(size_t)(have + d) * slab + r] = acc;
      }
      if (have + cnt < N + 1) {
        for (int e = tid; e < n * n; e += nthr) {
          const int i = e / n, c = e - i * n;
          double acc = 0.0;
          for (int l = 0; l < n; ++l) acc = fma(P0[i * n + l], P0[l * n + c], acc);
          P1[e] = acc;
        }
      }
      __syncthreads();
      double* t = P0; P0 = P1; P1 = t;
      have += cnt;
    }
    // ---- 2. signals of the coupled costs -----------------------------------------------
    for (int e = tid; e < nC * w * N; e += nthr) {
      const int c = e / (w * N), r = e - c * w * N, j = r / N, d = r - j * N;
      const double* cx = a.ccx + b * a.sccx + (size_t)c * n;
      if (j < m) {
        double acc = (d == 0) ? a.ccu[b * a.sccu + (size_t)c * m + j] : 0.0;
        const double* src = V + (size_t)d * slab + j;
        for (int i = 0; i < n; ++i) acc = fma(cx[i], src[i * w], acc);
        hs[(size_t)c * nu + r] = acc;
        if (acc != 0.0) hsact[c * m + j] = 1;
      } else {  // k = d: free response plus the given part
        double fr = 0.0;
        const double* src = V + (size_t)(d + 1) * slab + m;
        for (int i = 0; i < n; ++i) fr = fma(cx[i], src[i * w], fr);
        double acc = 0.0;
        for (int z = a.e_ptr[c * N + d]; z < a.e_ptr[c * N + d + 1]; ++z)
          acc = fma(a.e_val[z], gv[a.e_col[z]], acc);
        if (a.caim) acc -= a.caim[b * (long)nC + c];
        ee[c * N + d] = acc;
        eps[c * N + d] = fr + acc;
      }
    }
    __syncthreads();
    // ---- 4. outputs ---------------------------------------------------------------------------
    if (a.q) {
      double* qb = a.q + (size_t)b * a.nvt;
      for (int z = warp; z < nz; z += nwarps) {  // q[z] = sum_c w_c Mz_c[:,z] . (fr_c + e_c)
        double acc = 0.0;
        for (int c = 0; c < nC; ++c) {
          if (!a.act[c * nz + z]) continue;
          const double* mz = a.MzT + ((size_t)c * nz + z) * N;
          double s = 0.0;
          for (int k = lane; k < N; k += 32) s = fma(__ldg(mz + k), eps[c * N + k], s);
          acc = fma(a.cw[b * a.scw + c], s, acc);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (lane == 0) qb[a.zcol[z]] = acc;
      }
      for (int r = tid; r < nu; r += nthr) {  // q[u] += sum_c w_c T_c^T e_c
        const int j = r / N, p = r - j * N;
        double acc = 0.0;
        for (int c = 0; c < nC; ++c) {
          if (!hsact[c * m + j]) continue;
          const double* h_ = hs + (size_t)c * nu + j * N;
          const double* ev = ee + c * N + p;
          double s = 0.0;
          for (int t = 0; t < N - p; ++t) s = fma(h_[t], ev[t], s);
          acc = fma(a.cw[b * a.scw + c], s, acc);
        }
        qb[a.voff[j] + p] += acc;
      }
    }
    if (a.G) {
      // rows without a plant part, whole rows (the fused kernel writes the rows of its groups,
      // the auxiliary variables' columns included): zeros under the plant's inputs, constants
      // under the auxiliary variables
      double* Gb = a.G + (size_t)b * a.nct * a.nvt;
      for (int e = tid; e < a.n_aux_rows * a.nvt; e += nthr) {
        const int i = e / a.nvt, col = e - i * a.nvt, what = colmap[col], r = a.aux_rows[i];
        if (what == INT32_MIN) continue;
        __stcs(Gb + (size_t)r * a.nvt + col, what >= 0 ? 0.0 : __ldg(a.Gz + b * a.sGz + r * nz + (-1 - what)));
      }
    }
    if (a.h) {
      double* hb = a.h + (size_t)b * a.nct;
      for (int r = tid; r < a.nct; r += nthr) {
        const int z0 = a.h_ptr[r], z1 = a.h_ptr[r + 1], kind = a.row_kind[r];
        if (z0 == z1 && !kind) continue;
        double acc = 0.0;
        for (int z = z0; z < z1; ++z) acc = fma(a.h_val[b * a.shv + z], gv[a.h_col[z]], acc);
        hb[r] = kind ? acc : hb[r] + acc;
      }
    }
    __syncthreads();  // shared memory is reused by the next instance
  }
}

}  // namespace

namespace mpci {

struct CouplePlan {
  CoupleArgs args;
  void* dev;
  size_t smem_bytes;
  int blocks_per_sm;
  int h_nnz;  // entries of the h map (instance stride of per-instance values)
};

int couple_create(int n, int m, int N, const mpci_lti_embed* e, CouplePlan** out) {
  *out = nullptr;
  const int nz = e->nz, nC = e->n_costs, nvt = e->nv_total, nct = e->nc_total, glen = e->given_len;
  if (nz < 0 || nC < 0 || glen < 0 || nvt < m * N + nz || nct < 0)
    return fail(-1, "lti_plan_create_embedded: bad sizes nz=%d costs=%d given_len=%d nv_total=%d nc_total=%d",
                nz, nC, glen, nvt, nct);
  if (!e->var_off || (nz && (!e->z_col || !e->Pzz)) || (nct && (!e->h_ptr || !e->row_kind)) ||
      (nct && nz && !e->Gz) ||
      (nC && (!e->cost_cx || !e->cost_cu || !e->cost_w || !e->cost_e_ptr || (nz && !e->cost_MzT))))
    return fail(-3, "lti_plan_create_embedded: NULL table");
  // every column is used at most once
  std::vector<char> used((size_t)nvt, 0);
  for (int j = 0; j < m; ++j) {
    if (e->var_off[j] < 0 || e->var_off[j] + N > nvt)
      return fail(-1, "lti_plan_create_embedded: var_off[%d] = %d outside the %d variables", j, e->var_off[j], nvt);
    for (int p = 0; p < N; ++p) {
      if (used[e->var_off[j] + p]) return fail(-1, "lti_plan_create_embedded: overlapping variable blocks");
      used[e->var_off[j] + p] = 1;
    }
  }
  for (int z = 0; z < nz; ++z) {
    if (e->z_col[z] < 0 || e->z_col[z] >= nvt || used[e->z_col[z]])
      return fail(-1, "lti_plan_create_embedded: z_col[%d] = %d is not a free column", z, e->z_col[z]);
    used[e->z_col[z]] = 1;
  }
  const int e_nnz = nC ? e->cost_e_ptr[(size_t)nC * N] : 0, h_nnz = nct ? e->h_ptr[nct] : 0;
  for (int z = 0; z < e_nnz; ++z)
    if (e->cost_e_col[z] < 0 || e->cost_e_col[z] > glen)
      return fail(-1, "lti_plan_create_embedded: cost_e_col out of range");
  for (int z = 0; z < h_nnz; ++z)
    if (e->h_col[z] < 0 || e->h_col[z] > glen) return fail(-1, "lti_plan_create_embedded: h_col out of range");

  std::vector<int> aux_rows;
  for (int r = 0; r < nct; ++r)
    if (e->row_kind[r]) aux_rows.push_back(r);
  std::vector<unsigned char> act((size_t)std::max(1, nC * nz), 0);
  std::vector<int> d_ptr(1, 0), d_tau;
  std::vector<double> d_val;
  for (int c = 0; c < nC; ++c)
    for (int z = 0; z < nz; ++z) {
      const double* col = e->cost_MzT + ((size_t)c * nz + z) * N;
      double before = 0.0;
      for (int k = 0; k < N; ++k) {
        if (col[k] != 0.0) act[(size_t)c * nz + z] = 1;
        if (col[k] != before) {
          d_tau.push_back(k);
          d_val.push_back(col[k] - before);
        }
        before = col[k];
      }
      d_ptr.push_back((int)d_tau.size());
    }

  // one allocation: doubles first, then ints, then bytes
  std::vector<double> dbl;
  std::vector<int> ints;
  auto pushd = [&](const double* p, size_t count) { const size_t o = dbl.size(); if (count) dbl.insert(dbl.end(), p, p + count); return o; };
  auto pushi = [&](const int* p, size_t count) { const size_t o = ints.size(); if (count) ints.insert(ints.end(), p, p + count); return o; };
  const size_t o_ccx = pushd(e->cost_cx, (size_t)nC * n), o_ccu = pushd(e->cost_cu, (size_t)nC * m);
  const size_t o_cw = pushd(e->cost_w, nC), o_mz = pushd(e->cost_MzT, (size_t)nC * nz * N);
  const size_t o_dv = pushd(d_val.data(), d_val.size());
  const size_t o_ev = pushd(e->cost_e_val, e_nnz), o_pzz = pushd(e->Pzz, (size_t)nz * nz);
  const size_t o_gz = pushd(e->Gz, (size_t)nct * nz), o_hv = pushd(e->h_val, h_nnz);
  const size_t o_voff = pushi(e->var_off, m), o_zcol = pushi(e->z_col, nz);
  const int zero = 0;
  const size_t o_eptr = nC ? pushi(e->cost_e_ptr, (size_t)nC * N + 1) : pushi(&zero, 1);
  const size_t o_ecol = pushi(e->cost_e_col, e_nnz);
  const size_t o_hptr = nct ? pushi(e->h_ptr, (size_t)nct + 1) : pushi(&zero, 1);
  const size_t o_hcol = pushi(e->h_col, h_nnz), o_kind = pushi(e->row_kind, nct);
  const size_t o_aux = pushi(aux_rows.data(), aux_rows.size());
  const size_t o_dp = pushi(d_ptr.data(), d_ptr.size()), o_dt = pushi(d_tau.data(), d_tau.size());
  const size_t bytes_d = dbl.size() * sizeof(double), bytes_i = ints.size() * sizeof(int);

  CouplePlan* p = new CouplePlan();
  cudaError_t err = cudaMalloc(&p->dev, bytes_d + bytes_i + act.size() + 16);
  char* base = static_cast<char*>(p->dev);
  if (err == cudaSuccess && bytes_d) err = cudaMemcpy(base, dbl.data(), bytes_d, cudaMemcpyHostToDevice);
  if (err == cudaSuccess && bytes_i) err = cudaMemcpy(base + bytes_d, ints.data(), bytes_i, cudaMemcpyHostToDevice);
  if (err == cudaSuccess) err = cudaMemcpy(base + bytes_d + bytes_i, act.data(), act.size(), cudaMemcpyHostToDevice);
  const size_t smem = sizeof(double) * couple_doubles(n, m, N, nz, nC, glen) + sizeof(int) * ((size_t)nvt + (size_t)nC * m + 2);
  if (err == cudaSuccess && smem > 227 * 1024) {
    cudaFree(p->dev);
    delete p;
    return fail(-8, "lti_plan_create_embedded: the coupling stage needs %zu bytes of shared memory (limit 232448)", smem);
  }
  if (err == cudaSuccess)
    err = cudaFuncSetAttribute(lti_couple_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  int per_sm = 1;
  if (err == cudaSuccess)
    err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, lti_couple_kernel, kThreads, smem);
  if (err != cudaSuccess) {
    if (p->dev) cudaFree(p->dev);
    delete p;
    return cuda_fail(err, "lti_plan_create_embedded");
  }
  const double* d = reinterpret_cast<const double*>(base);
  const int* i = reinterpret_cast<const int*>(base + bytes_d);
  CoupleArgs& a = p->args;
  a.n = n; a.m = m; a.N = N; a.nz = nz; a.nC = nC; a.glen = glen; a.nvt = nvt; a.nct = nct;
  a.n_aux_rows = (int)aux_rows.size();
  a.ccx = d + o_ccx; a.ccu = d + o_ccu; a.cw = d + o_cw; a.MzT = d + o_mz; a.e_val = d + o_ev;
  a.Pzz = d + o_pzz; a.Gz = d + o_gz; a.h_val = d + o_hv;
  a.voff = i + o_voff; a.zcol = i + o_zcol; a.e_ptr = i + o_eptr; a.e_col = i + o_ecol;
  a.h_ptr = i + o_hptr; a.h_col = i + o_hcol; a.row_kind = i + o_kind; a.aux_rows = i + o_aux;
  a.act = reinterpret_cast<const unsigned char*>(base + bytes_d + bytes_i);
  a.d_ptr = i + o_dp; a.d_tau = i + o_dt; a.d_val = d + o_dv;
  p->smem_bytes = smem;
  p->h_nnz = h_nnz;
  p->blocks_per_sm = per_sm < 1 ? 1 : per_sm;
  *out = p;
  return 0;
}

void couple_share(const CouplePlan* p, CoupleShare* out) {
  const CoupleArgs& a = p->args;
  out->Gz = a.Gz; out->Pzz = a.Pzz; out->cost_cx = a.ccx; out->cost_cu = a.ccu; out->cost_w = a.cw;
  out->d_val = a.d_val; out->z_col = a.zcol; out->d_ptr = a.d_ptr; out->d_tau = a.d_tau;
  out->nz = a.nz; out->n_costs = a.nC;
}

void couple_destroy(CouplePlan* p) {
  if (!p) return;
  if (p->dev) cudaFree(p->dev);
  delete p;
}

int couple_launch(const CouplePlan* p, const double* A, long strideA, const double* B, long strideB,
                  const double* x0, const double* given, const mpci_lti_coupled_tables* ct,
                  double* q, double* G, double* h, long batch, cudaStream_t stream) {
  if (batch == 0 || (!q && !G && !h)) return 0;
  CoupleArgs a = p->args;
  if (a.glen && !given) return fail(-3, "lti_assemble_coupled: NULL given");
  a.A = A; a.B = B; a.x0 = x0; a.given = given; a.sA = strideA; a.sB = strideB;
  a.sccx = a.sccu = a.scw = a.sGz = a.shv = 0;
  a.caim = nullptr;
  if (ct && ct->cost_cx) { a.ccx = ct->cost_cx; a.sccx = (long)a.nC * a.n; }
  if (ct && ct->cost_cu) { a.ccu = ct->cost_cu; a.sccu = (long)a.nC * a.m; }
  if (ct && ct->cost_w) { a.cw = ct->cost_w; a.scw = a.nC; }
  if (ct && ct->cost_aim) a.caim = ct->cost_aim;
  if (ct && ct->Gz) { a.Gz = ct->Gz; a.sGz = (long)a.nct * a.nz; }
  if (ct && ct->h_val) { a.h_val = ct->h_val; a.shv = p->h_nnz; }
  a.q = q; a.G = G; a.h = h; a.batch = batch;
  int dev = 0, sms = 148;
  MPCI_CUDA(cudaGetDevice(&dev));
  MPCI_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  MPCI_CUDA(cudaFuncSetAttribute(lti_couple_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  const long resident = (long)sms * p->blocks_per_sm;
  const unsigned grid = (unsigned)(batch < resident ? batch : resident);
  lti_couple_kernel<<<grid, kThreads, p->smem_bytes, stream>>>(a);
  MPCI_LAUNCH_CHECK("lti_couple_kernel");
  return 0;
}

}  // namespace mpci
