// Costs with a `schedule` on the per-instance path (include/mpci.h: mpci_lti_sched_desc).
//
// The fused kernel of mpci_lti.cu builds P from ONE stage weight (its Hessian variants all rest
// on  P[p,p'] = sum_k Gam_{k-p}^T Qxx Gam_{k-p'}  with the same Qxx at every sample).  A cost that
// acts on some samples only (goal.py:46-60 `schedule`; the rows sigma of body.py:273-283 -- the
// terminal cost of an MPC is schedule = range(N - 1, N)) is a sum of a few outer products
// instead, and is ADDED here to the P and q the assembly wrote, one CTA per instance:
//   1. V_d = A^d [B | x0] (d <= K, K = 1 + the last scheduled sample) by prefix doubling, as the
//      coupling kernel does (what extend_matrices builds by recursion, tools.py:23-29)
//   2. per scheduled row r = (term t, sample k), a chunk of rows at a time in shared memory:
//      v_r [nv] = row k of the variable's preview matrix over the optimisation variables,
//      c_r the crossed variable's, and the given parts fv_r, fc_r
//   3. P[i][j] += sum_r w_r v_r[i] c_r[j]                                        (body.py:292)
//      q[i]    += sum_r w_r / 2 (v_r[i] (fc_r - caim_r) + c_r[i] (fv_r - aim_r))  (body.py:293-300)
// P is read and written once per chunk of rows: a terminal cost costs one more pass over P.
#include <algorithm>
#include <vector>

#include "mpci_common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr size_t kRowBudget = 64 * 1024;    // bytes of shared memory for the chunk of rows ...
constexpr size_t kSmemLimit = 200 * 1024;   // ... unless one row pair needs more

struct SchedArgs {
  int n, m, N, nv, nz, glen, T, R, K, chunk;
  const int *voff, *zcol, *row_term, *row_out;
  const int *ent_ptr, *ent_k, *cent_ptr, *cent_k;  // row r = sum over its entries of val * (sample k's row)
  const double *ent_val, *cent_val;                // (cent_*: the crossed variable's, nullptr = the same)
  const double *vMz, *vMg, *cMz, *cMg;
  const double *A, *B, *x0, *given, *terms;
  long sA, sB, sT;
  double *P, *q;
  double *G, *h;  // rows mode (mpci_lti_sched_rows): the rows are ADDED to rows row_out[r] of G / taken from h
  long nc;
  long batch;
};

// shared memory (doubles): Mp, Mq [n][n] | S [K + 1][n][m + 1] | V, C [chunk][nv] | epsV, epsC [chunk]
__host__ __device__ inline size_t fixed_doubles(int n, int m, int K) {
  const size_t d = (size_t)2 * n * n + (size_t)(K + 1) * n * (m + 1);
  return (d + 1) & ~(size_t)1;  // (the rows are read 16 bytes at a time)
}

// P[i .. i + ROWS)[2 j, 2 j + 1] += sum_r (w v_r)[i ..] c_r[2 j ..]: ROWS 16-byte loads in flight per lane
template <int ROWS>
__device__ __forceinline__ void update_rows(double* __restrict__ Pb, const double* V, const double* C, int nv,
                                            int rows, int i, int lane) {
  const int half = nv >> 1;
  for (int j = lane; j < half; j += 32) {
    double2 p[ROWS];
#pragma unroll
    for (int u = 0; u < ROWS; ++u) p[u] = reinterpret_cast<const double2*>(Pb + (long)(i + u) * nv)[j];
    for (int rr = 0; rr < rows; ++rr) {
      const double2 c = reinterpret_cast<const double2*>(C + (size_t)rr * nv)[j];
#pragma unroll
      for (int u = 0; u < ROWS; u += 2) {  // (i is even and so is nv: 16-byte broadcasts)
        const double2 v = *reinterpret_cast<const double2*>(V + (size_t)rr * nv + i + u);
        p[u].x = fma(v.x, c.x, p[u].x), p[u].y = fma(v.x, c.y, p[u].y);
        p[u + 1].x = fma(v.y, c.x, p[u + 1].x), p[u + 1].y = fma(v.y, c.y, p[u + 1].y);
      }
    }
#pragma unroll
    for (int u = 0; u < ROWS; ++u) reinterpret_cast<double2*>(Pb + (long)(i + u) * nv)[j] = p[u];
  }
}

__global__ void __launch_bounds__(kThreads, 3) lti_sched_kernel(SchedArgs a) {
  extern __shared__ double sched_smem[];
  const int n = a.n, m = a.m, nv = a.nv, K = a.K, tl = 3 + 2 * a.n + 2 * a.m, w1 = a.m + 1, slab = a.n * w1;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nthr = (int)blockDim.x, nwarps = nthr >> 5;
  double* Mp = sched_smem;
  double* Mq = Mp + n * n;
  double* S = Mq + n * n;  // S[d][i][c]: c < m: (A^d B)[i][c];  c == m: (A^d x0)[i]
  double* V = sched_smem + fixed_doubles(n, m, K);
  double* C = V + (size_t)a.chunk * nv;
  double* epsV = C + (size_t)a.chunk * nv;
  double* epsC = epsV + a.chunk;

  for (long b = blockIdx.x; b < a.batch; b += gridDim.x) {
    const double* Ab = a.A + b * a.sA;
    const double* Bb = a.B + b * a.sB;
    const double* xb = a.x0 + b * (long)n;
    const double* gb = a.given ? a.given + b * (long)a.glen : nullptr;
    const double* tb = a.terms + b * a.sT;
    __syncthreads();  // (the previous instance's rows are still being read)
    for (int e = tid; e < n * n; e += nthr) Mp[e] = Ab[e];
    for (int e = tid; e < slab; e += nthr) {
      const int i = e / w1, c = e - i * w1;
      S[e] = (c < m) ? Bb[i * m + c] : xb[i];
    }
    __syncthreads();
    double* P0 = Mp;
    double* P1 = Mq;
    for (int have = 1; have < K + 1;) {  // S[have .. have + cnt) = A^have S[0 .. cnt)
      const int cnt = min(have, K + 1 - have);
      for (int e = tid; e < cnt * slab; e += nthr) {
        const int d = e / slab, r = e - d * slab, i = r / w1, c = r - i * w1;
        const double* src = S + (size_t)d * slab + c;
        double acc = 0.0;
        for (int l = 0; l < n; ++l) acc = fma(P0[i * n + l], src[l * w1], acc);
        S[(size_t)(have + d) * slab + r] = acc;
      }
      if (have + cnt < K + 1) {
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

    for (int r0 = 0; r0 < a.R; r0 += a.chunk) {
      const int rows = min(a.chunk, a.R - r0);
      if (r0) __syncthreads();
      for (int e = tid; e < rows * nv; e += nthr) V[e] = 0.0, C[e] = 0.0;
      __syncthreads();
      for (int rr = warp; rr < rows; rr += nwarps) {  // a warp per row; V is kept scaled by the weight
        const int r = r0 + rr, t = a.row_term[r];
        const double* tt = tb + (long)t * tl;
        const double *vcx = tt + 3, *vcu = vcx + n, *ccx = vcu + m, *ccu = ccx + n;
        const double w = tt[0];
        double* Vr = V + (size_t)rr * nv;
        double* Cr = C + (size_t)rr * nv;
        const int e0 = a.ent_ptr[r], e1 = a.ent_ptr[r + 1];
        const bool own = a.cent_ptr == nullptr;  // the crossed variable combines the same samples
        const int c0 = own ? e0 : a.cent_ptr[r], c1 = own ? e1 : a.cent_ptr[r + 1];
        for (int j = 0; j < m; ++j) {
          double* Vj = Vr + a.voff[j];
          double* Cj = Cr + a.voff[j];
          for (int p = lane; p < K; p += 32) {
            double sv = 0.0, sc = 0.0;
            for (int e = e0; e < e1; ++e) {
              const int k = a.ent_k[e];
              if (p > k) continue;
              const double* Wd = S + (size_t)(k - p) * slab + j;
              double av = p == k ? vcu[j] : 0.0, ac = p == k ? ccu[j] : 0.0;
              for (int i = 0; i < n; ++i) {
                const double wv = Wd[i * w1];
                av = fma(vcx[i], wv, av);
                ac = fma(ccx[i], wv, ac);
              }
              sv = fma(a.ent_val[e], av, sv);
              if (own) sc = fma(a.ent_val[e], ac, sc);
            }
            for (int e = own ? c1 : c0; e < c1; ++e) {
              const int k = a.cent_k[e];
              if (p > k) continue;
              const double* Wd = S + (size_t)(k - p) * slab + j;
              double ac = p == k ? ccu[j] : 0.0;
              for (int i = 0; i < n; ++i) ac = fma(ccx[i], Wd[i * w1], ac);
              sc = fma(a.cent_val[e], ac, sc);
            }
            Vj[p] = w * sv;
            Cj[p] = sc;
          }
        }
        const double* cMz = a.cMz ? a.cMz : a.vMz;
        for (int i = lane; i < a.nz; i += 32) {
          if (a.vMz) Vr[a.zcol[i]] = w * a.vMz[(size_t)r * a.nz + i];
          if (cMz) Cr[a.zcol[i]] = cMz[(size_t)r * a.nz + i];
        }
        // the given parts: cx . A^{k+1} x0 + Mg[r,:] . given
        const double* cMg = a.cMg ? a.cMg : a.vMg;
        double fv = 0.0, fc = 0.0;
        for (int i = lane; i < n; i += 32) {
          double xv = 0.0, xc = 0.0;  // the combined free responses sum_e val_e (A^{k_e + 1} x0)[i]
          for (int e = e0; e < e1; ++e) xv = fma(a.ent_val[e], S[(size_t)(a.ent_k[e] + 1) * slab + i * w1 + m], xv);
          if (own) xc = xv;
          for (int e = own ? c1 : c0; e < c1; ++e)
            xc = fma(a.cent_val[e], S[(size_t)(a.cent_k[e] + 1) * slab + i * w1 + m], xc);
          fv = fma(vcx[i], xv, fv);
          fc = fma(ccx[i], xc, fc);
        }
        if (gb) {
          for (int g = lane; g < a.glen; g += 32) {
            const double gv = gb[g];
            if (a.vMg) fv = fma(a.vMg[(size_t)r * a.glen + g], gv, fv);
            if (cMg) fc = fma(cMg[(size_t)r * a.glen + g], gv, fc);
          }
        }
        for (int o = 16; o; o >>= 1) {
          fv += __shfl_xor_sync(0xffffffffu, fv, o);
          fc += __shfl_xor_sync(0xffffffffu, fc, o);
        }
        if (lane == 0) {
          epsV[rr] = a.G || a.h ? fv : 0.5 * w * (fv - tt[1]);  // times c_r  (rows mode: the given part itself)
          epsC[rr] = 0.5 * (fc - tt[2]);                        // times w v_r
        }
      }
      __syncthreads();
      if (a.G || a.h) {  // rows mode: G[row_out[r]][plant columns] += w v_r,  h[row_out[r]] -= w fv_r
        for (int rr = warp; rr < rows; rr += nwarps) {
          const int r = r0 + rr, t = a.row_term[r];
          const long row = (long)b * a.nc + a.row_out[r];
          if (a.G) {
            double* Gr = a.G + row * nv;
            for (int e = lane; e < m * a.N; e += 32) {
              const int j = e / a.N, p = e - j * a.N, col = a.voff[j] + p;
              const double v = V[(size_t)rr * nv + col];
              if (v != 0.0) atomicAdd(Gr + col, v);  // (the axes of one limit are terms of their own on the same rows)
            }
          }
          if (a.h && lane == 0) atomicAdd(a.h + row, -tb[(long)t * tl] * epsV[rr]);
        }
        continue;
      }
      if (a.P) {
        double* Pb = a.P + b * (long)nv * nv;
        if ((nv & 1) == 0 && (reinterpret_cast<size_t>(Pb) & 15) == 0) {
          // four rows of P per warp pass, 16 bytes per lane and row
          for (int i = 4 * warp; i < nv; i += 4 * nwarps) {
            if (i + 4 <= nv)
              update_rows<4>(Pb, V, C, nv, rows, i, lane);
            else
              update_rows<2>(Pb, V, C, nv, rows, i, lane);
          }
        } else {
          for (int i = warp; i < nv; i += nwarps) {
            double* Pi = Pb + (long)i * nv;
            for (int j = lane; j < nv; j += 32) {
              double acc = Pi[j];
              for (int rr = 0; rr < rows; ++rr) acc = fma(V[(size_t)rr * nv + i], C[(size_t)rr * nv + j], acc);
              Pi[j] = acc;
            }
          }
        }
      }
      if (a.q) {
        double* qb = a.q + b * (long)nv;
        for (int i = tid; i < nv; i += nthr) {
          double acc = 0.0;
          for (int rr = 0; rr < rows; ++rr)
            acc += V[(size_t)rr * nv + i] * epsC[rr] + C[(size_t)rr * nv + i] * epsV[rr];
          qb[i] += acc;
        }
      }
    }
  }
}

template <typename T>
int upload(T** dst, const T* src, size_t count) {
  *dst = nullptr;
  if (!src || !count) return 0;
  MPCI_CUDA(cudaMalloc(dst, count * sizeof(T)));
  MPCI_CUDA(cudaMemcpy(*dst, src, count * sizeof(T), cudaMemcpyHostToDevice));
  return 0;
}

}  // namespace

struct mpci_lti_sched {
  int n = 0, m = 0, N = 0, nv = 0, nz = 0, glen = 0, T = 0, R = 0, K = 0, chunk = 0, device = 0;
  size_t smem = 0;
  int *voff = nullptr, *zcol = nullptr, *row_term = nullptr, *row_out = nullptr;
  int *ent_ptr = nullptr, *ent_k = nullptr, *cent_ptr = nullptr, *cent_k = nullptr;
  double *ent_val = nullptr, *cent_val = nullptr;
  double *vMz = nullptr, *vMg = nullptr, *cMz = nullptr, *cMg = nullptr;
};

extern "C" int mpci_lti_sched_destroy(mpci_lti_sched* s) {
  if (!s) return 0;
  cudaFree(s->voff), cudaFree(s->zcol), cudaFree(s->row_term), cudaFree(s->row_out);
  cudaFree(s->ent_ptr), cudaFree(s->ent_k), cudaFree(s->ent_val);
  cudaFree(s->cent_ptr), cudaFree(s->cent_k), cudaFree(s->cent_val);
  cudaFree(s->vMz), cudaFree(s->vMg), cudaFree(s->cMz), cudaFree(s->cMg);
  delete s;
  return 0;
}

extern "C" int mpci_lti_sched_create(const mpci_lti_sched_desc* d, mpci_lti_sched** out) {
  using mpci::fail;
  if (!d || !out) return fail(-1, "mpci_lti_sched_create: null argument");
  *out = nullptr;
  if (d->n <= 0 || d->m <= 0 || d->N <= 0 || d->nv <= 0 || d->nz < 0 || d->given_len < 0 || d->n_terms <= 0)
    return fail(-2, "mpci_lti_sched_create: bad sizes (n %d, m %d, N %d, nv %d, nz %d, given_len %d, terms %d)",
                d->n, d->m, d->N, d->nv, d->nz, d->given_len, d->n_terms);
  if (!d->row_ptr || (!d->row_k && !d->ent_ptr))
    return fail(-1, "mpci_lti_sched_create: row_ptr / row_k (or ent_ptr) missing");
  if (d->ent_ptr && (!d->ent_k || !d->ent_val)) return fail(-1, "mpci_lti_sched_create: ent_k / ent_val missing");
  if (d->c_ent_ptr && (!d->ent_ptr || !d->c_ent_k || !d->c_ent_val))
    return fail(-1, "mpci_lti_sched_create: c_ent_* needs ent_* and its own c_ent_k / c_ent_val");
  if (d->nz > 0 && !d->z_col) return fail(-1, "mpci_lti_sched_create: z_col missing");
  if (d->row_ptr[0] != 0) return fail(-2, "mpci_lti_sched_create: row_ptr[0] must be 0");
  const int R = d->row_ptr[d->n_terms];
  if (R <= 0) return fail(-2, "mpci_lti_sched_create: no scheduled rows");
  std::vector<int> row_term(R), voff(d->m), ent_ptr(R + 1), ent_k;
  std::vector<double> ent_val;
  for (int t = 0; t < d->n_terms; ++t) {
    if (d->row_ptr[t + 1] < d->row_ptr[t]) return fail(-2, "mpci_lti_sched_create: row_ptr decreases at term %d", t);
    for (int r = d->row_ptr[t]; r < d->row_ptr[t + 1]; ++r) row_term[r] = t;
  }
  if (d->ent_ptr) {  // rows that combine samples (`L`)
    if (d->ent_ptr[0] != 0) return fail(-2, "mpci_lti_sched_create: ent_ptr[0] must be 0");
    for (int r = 0; r < R; ++r)
      if (d->ent_ptr[r + 1] < d->ent_ptr[r]) return fail(-2, "mpci_lti_sched_create: ent_ptr decreases at row %d", r);
    ent_ptr.assign(d->ent_ptr, d->ent_ptr + R + 1);
    ent_k.assign(d->ent_k, d->ent_k + ent_ptr[R]);
    ent_val.assign(d->ent_val, d->ent_val + ent_ptr[R]);
  } else {  // one sample per row
    for (int r = 0; r <= R; ++r) ent_ptr[r] = r;
    ent_k.assign(d->row_k, d->row_k + R);
    ent_val.assign(R, 1.0);
  }
  int K = 0;
  auto check = [&](const int* k, int count) {
    for (int e = 0; e < count; ++e) {
      if (k[e] < 0 || k[e] >= d->N) return k[e];
      K = std::max(K, k[e] + 1);
    }
    return 0;
  };
  int bad = check(ent_k.data(), (int)ent_k.size());
  if (!bad && d->c_ent_ptr) {
    if (d->c_ent_ptr[0] != 0) return fail(-2, "mpci_lti_sched_create: c_ent_ptr[0] must be 0");
    for (int r = 0; r < R; ++r)
      if (d->c_ent_ptr[r + 1] < d->c_ent_ptr[r])
        return fail(-2, "mpci_lti_sched_create: c_ent_ptr decreases at row %d", r);
    bad = check(d->c_ent_k, d->c_ent_ptr[R]);
  }
  if (bad || (ent_k.size() && K == 0))
    return fail(-3, "mpci_lti_sched_create: a term schedules sample %d outside the horizon of %d", bad, d->N);
  if (K == 0) K = 1;
  if (d->row_out)
    for (int r = 0; r < R; ++r)
      if (d->row_out[r] < 0) return fail(-3, "mpci_lti_sched_create: row %d goes to row %d of G", r, d->row_out[r]);
  for (int j = 0; j < d->m; ++j) {
    voff[j] = d->var_off ? d->var_off[j] : j * d->N;
    if (voff[j] < 0 || voff[j] + d->N > d->nv)
      return fail(-3, "mpci_lti_sched_create: input %d at columns %d..%d of %d", j, voff[j], voff[j] + d->N, d->nv);
  }
  for (int i = 0; i < d->nz; ++i) {
    if (d->z_col[i] < 0 || d->z_col[i] >= d->nv)
      return fail(-3, "mpci_lti_sched_create: auxiliary variable %d at column %d of %d", i, d->z_col[i], d->nv);
    for (int j = 0; j < d->m; ++j)
      if (d->z_col[i] >= voff[j] && d->z_col[i] < voff[j] + d->N)
        return fail(-3, "mpci_lti_sched_create: auxiliary variable %d at column %d, inside input %d's samples", i,
                    d->z_col[i], j);
  }
  if ((d->c_Mz && !d->v_Mz) || (d->c_Mg && !d->v_Mg))
    return fail(-2, "mpci_lti_sched_create: a crossed auxiliary table needs the variable's own");

  auto* s = new mpci_lti_sched;
  s->n = d->n, s->m = d->m, s->N = d->N, s->nv = d->nv, s->nz = d->nz, s->glen = d->given_len;
  s->T = d->n_terms, s->R = R, s->K = K;
  const size_t fixed = fixed_doubles(d->n, d->m, K) * sizeof(double);
  const size_t per_row = ((size_t)2 * d->nv + 2) * sizeof(double);
  size_t chunk = std::max<size_t>(1, kRowBudget / per_row);
  chunk = std::min<size_t>(chunk, (size_t)R);
  if (fixed + per_row > kSmemLimit) {
    delete s;
    return fail(-4, "mpci_lti_sched_create: %zu bytes of shared memory needed (n %d, m %d, %d samples, nv %d), "
                "%zu available", fixed + per_row, d->n, d->m, K, d->nv, kSmemLimit);
  }
  while (fixed + chunk * per_row > kSmemLimit) --chunk;
  s->chunk = (int)chunk;
  s->smem = fixed + chunk * per_row;
  cudaGetDevice(&s->device);
  int rc = 0;
  if (!rc) rc = upload(&s->voff, voff.data(), (size_t)d->m);
  if (!rc) rc = upload(&s->zcol, d->z_col, (size_t)d->nz);
  if (!rc) rc = upload(&s->row_term, row_term.data(), (size_t)R);
  if (!rc) rc = upload(&s->row_out, d->row_out, (size_t)R);
  if (!rc) rc = upload(&s->ent_ptr, ent_ptr.data(), (size_t)R + 1);
  if (!rc) rc = upload(&s->ent_k, ent_k.data(), ent_k.size());
  if (!rc) rc = upload(&s->ent_val, ent_val.data(), ent_val.size());
  if (!rc && d->c_ent_ptr) {
    rc = upload(&s->cent_ptr, d->c_ent_ptr, (size_t)R + 1);
    if (!rc) rc = upload(&s->cent_k, d->c_ent_k, (size_t)d->c_ent_ptr[R]);
    if (!rc) rc = upload(&s->cent_val, d->c_ent_val, (size_t)d->c_ent_ptr[R]);
  }
  if (!rc) rc = upload(&s->vMz, d->v_Mz, (size_t)R * d->nz);
  if (!rc) rc = upload(&s->vMg, d->v_Mg, (size_t)R * d->given_len);
  if (!rc) rc = upload(&s->cMz, d->c_Mz, (size_t)R * d->nz);
  if (!rc) rc = upload(&s->cMg, d->c_Mg, (size_t)R * d->given_len);
  if (!rc && s->smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(lti_sched_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)kSmemLimit);
    if (e != cudaSuccess) rc = mpci::cuda_fail(e, "cudaFuncSetAttribute(lti_sched_kernel)");
  }
  if (rc) {
    mpci_lti_sched_destroy(s);
    return rc;
  }
  *out = s;
  return 0;
}

static int sched_launch(const mpci_lti_sched* s, const char* who, const double* A, long strideA, const double* B,
                        long strideB, const double* x0, const double* given, const double* terms,
                        long stride_terms, double* P, double* q, double* G, double* h, long nc, long batch,
                        void* stream) {
  using mpci::fail;
  if (!s || !A || !B || !x0 || !terms) return fail(-1, "%s: null argument", who);
  if (batch < 0) return fail(-2, "%s: negative batch", who);
  if (s->vMg && !given) return fail(-1, "%s: the terms read `given`, which is NULL", who);
  const long tl = (long)s->T * (3 + 2 * s->n + 2 * s->m);
  if (stride_terms != 0 && stride_terms != tl)
    return fail(-2, "%s: stride_terms %ld (0 or %ld expected)", who, stride_terms, tl);
  if (batch == 0 || (!P && !q && !G && !h)) return 0;
  SchedArgs a{};
  a.n = s->n, a.m = s->m, a.N = s->N, a.nv = s->nv, a.nz = s->nz, a.glen = s->glen, a.T = s->T, a.R = s->R;
  a.K = s->K, a.chunk = s->chunk;
  a.voff = s->voff, a.zcol = s->zcol, a.row_term = s->row_term, a.row_out = s->row_out;
  a.ent_ptr = s->ent_ptr, a.ent_k = s->ent_k, a.ent_val = s->ent_val;
  a.cent_ptr = s->cent_ptr, a.cent_k = s->cent_k, a.cent_val = s->cent_val;
  a.vMz = s->vMz, a.vMg = s->vMg, a.cMz = s->cMz, a.cMg = s->cMg;
  a.A = A, a.B = B, a.x0 = x0, a.given = given, a.terms = terms;
  a.sA = strideA, a.sB = strideB, a.sT = stride_terms;
  a.P = P, a.q = q, a.G = G, a.h = h, a.nc = nc, a.batch = batch;
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, s->device);
  const long grid = std::min<long>(batch, (long)sms * 8);
  lti_sched_kernel<<<(unsigned)grid, kThreads, s->smem, mpci::as_stream(stream)>>>(a);
  MPCI_LAUNCH_CHECK("lti_sched_kernel");
  return 0;
}

extern "C" int mpci_lti_sched_add(const mpci_lti_sched* s, const double* A, long strideA, const double* B,
                                  long strideB, const double* x0, const double* given, const double* terms,
                                  long stride_terms, double* P, double* q, long batch, void* stream) {
  if (s && s->row_out) return mpci::fail(-1, "mpci_lti_sched_add: the description holds limit rows (row_out)");
  return sched_launch(s, "mpci_lti_sched_add", A, strideA, B, strideB, x0, given, terms, stride_terms, P, q,
                      nullptr, nullptr, 0, batch, stream);
}

extern "C" int mpci_lti_sched_rows(const mpci_lti_sched* s, const double* A, long strideA, const double* B,
                                   long strideB, const double* x0, const double* terms, long stride_terms,
                                   double* G, double* h, long nc_total, long batch, void* stream) {
  if (s && !s->row_out) return mpci::fail(-1, "mpci_lti_sched_rows: the description has no row_out");
  if (s && (s->vMz || s->vMg)) return mpci::fail(-1, "mpci_lti_sched_rows: limit rows carry no auxiliary tables");
  return sched_launch(s, "mpci_lti_sched_rows", A, strideA, B, strideB, x0, nullptr, terms, stride_terms, nullptr,
                      nullptr, G, h, nc_total, batch, stream);
}
