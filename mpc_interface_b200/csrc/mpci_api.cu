// C ABI: plan life cycle, device-resident assembly, host-buffer assembly.
#include <cstring>

#include "mpci_common.cuh"
#include "mpci_plan.cuh"

namespace mpci {
char* error_buffer() {
  static thread_local char buf[512] = "";
  return buf;
}
}  // namespace mpci

extern "C" int mpci_version(void) { return MPCI_VERSION; }

extern "C" const char* mpci_last_error(void) { return mpci::error_buffer(); }

extern "C" int mpci_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

namespace {

int fill_stream(mpci::StreamArgs& s, int n_terms, const int* terms, const int* monomials,
                int n_monomials, const double* consts, long n_elems, const char* what) {
  if (n_terms < 1 || n_terms > MPCI_MAX_TERMS)
    return mpci::fail(-7, "plan: %s has %d terms (1..%d allowed)", what, n_terms, MPCI_MAX_TERMS);
  if (n_elems > 0 && !consts) return mpci::fail(-3, "plan: %s constants are NULL", what);
  s.n_terms = n_terms;
  s.consts = consts;
  s.n_elems = n_elems;
  for (int k = 0; k < n_terms; ++k) {
    if (terms[k] < 0 || terms[k] >= n_monomials)
      return mpci::fail(-9, "plan: %s term %d names monomial %d of %d", what, k, terms[k],
                        n_monomials);
    for (int f = 0; f < MPCI_MAX_FACTORS; ++f)
      s.factors[k][f] = monomials[terms[k] * MPCI_MAX_FACTORS + f];
  }
  return 0;
}

}  // namespace

extern "C" int mpci_plan_create(const mpci_plan_desc* d, mpci_plan** out) {
  if (!d || !out) return mpci::fail(-3, "plan_create: NULL argument");
  *out = nullptr;
  if (d->nv < 0 || d->ng < 0 || d->nc < 0 || d->n_scalars < 1 || d->n_monomials < 1)
    return mpci::fail(-1, "plan_create: bad sizes nv=%d ng=%d nc=%d scalars=%d monomials=%d",
                      d->nv, d->ng, d->nc, d->n_scalars, d->n_monomials);
  for (int i = 0; i < d->n_monomials * MPCI_MAX_FACTORS; ++i)
    if (d->monomials[i] >= d->n_scalars)
      return mpci::fail(-9, "plan_create: monomial factor %d out of %d scalars", d->monomials[i],
                        d->n_scalars);
  mpci_plan* p = new mpci_plan();
  std::memset(p, 0, sizeof(*p));
  p->nv = d->nv; p->ng = d->ng; p->nc = d->nc; p->n_scalars = d->n_scalars;
  int rc = fill_stream(p->P, d->n_P_terms, d->P_terms, d->monomials, d->n_monomials, d->DP_dev,
                       (long)d->nv * d->nv, "P");
  if (!rc)
    rc = fill_stream(p->G, d->n_G_terms, d->G_terms, d->monomials, d->n_monomials, d->DG_dev,
                     (long)d->nc * d->nv, "G");
  const int rows = d->nv + d->nc, ngx = d->ng + 1;
  const int n_tiles = (rows + MPCI_QH_ROW_TILE - 1) / MPCI_QH_ROW_TILE;
  if (!rc && (d->n_Y_terms < 1 || !d->Y_terms || (rows > 0 && !d->DY_dev)))
    rc = mpci::fail(-3, "plan_create: the [q;h] terms are missing");
  if (!rc && d->n_qh_ranges > 0 && (!d->qh_ranges || !d->qh_tile_ptr))
    rc = mpci::fail(-3, "plan_create: qh range tables are NULL");
  if (!rc) {
    for (int r = 0; r < d->n_qh_ranges; ++r) {
      const int* q = d->qh_ranges + 4 * r;
      if (q[0] < 0 || q[0] >= n_tiles || q[1] < 0 || q[1] >= d->n_Y_terms || q[2] < 0 ||
          q[3] > ngx || q[2] > q[3]) {
        rc = mpci::fail(-9, "plan_create: bad qh range %d (%d %d %d %d)", r, q[0], q[1], q[2],
                        q[3]);
        break;
      }
    }
  }
  if (rc) { delete p; return rc; }

  // device tables: factors of the Y terms | ranges | tile_ptr
  std::vector<int> host;
  for (int j = 0; j < d->n_Y_terms; ++j) {
    const int m = d->Y_terms[j];
    if (m < 0 || m >= d->n_monomials) {
      delete p;
      return mpci::fail(-9, "plan_create: Y term %d names monomial %d", j, m);
    }
    for (int f = 0; f < MPCI_MAX_FACTORS; ++f)
      host.push_back(d->monomials[m * MPCI_MAX_FACTORS + f]);
  }
  const size_t off_ranges = host.size();
  for (int i = 0; i < 4 * d->n_qh_ranges; ++i) host.push_back(d->qh_ranges[i]);
  while (host.size() % 4) host.push_back(0);  // keep ranges int4-aligned relative to base
  const size_t off_tiles = host.size();
  for (int i = 0; i <= n_tiles; ++i) host.push_back(d->n_qh_ranges ? d->qh_tile_ptr[i] : 0);

  // ranges must start on a 16-byte boundary: pad the factor block
  std::vector<int> packed;
  const size_t factors_len = off_ranges;
  const size_t pad = (4 - factors_len % 4) % 4;
  packed.insert(packed.end(), host.begin(), host.begin() + factors_len);
  packed.insert(packed.end(), pad, 0);
  packed.insert(packed.end(), host.begin() + off_ranges, host.end());
  const size_t p_ranges = factors_len + pad, p_tiles = off_tiles + pad;

  cudaError_t e = cudaMalloc(&p->dev_tables, packed.size() * sizeof(int));
  if (e == cudaSuccess)
    e = cudaMemcpy(p->dev_tables, packed.data(), packed.size() * sizeof(int),
                   cudaMemcpyHostToDevice);
  const size_t dyt_elems = (size_t)d->n_Y_terms * rows * ngx;
  if (e == cudaSuccess && dyt_elems) e = cudaMalloc(&p->DYt, dyt_elems * sizeof(double));
  if (e != cudaSuccess) {
    rc = mpci::cuda_fail(e, "plan_create: device tables");
    mpci_plan_destroy(p);
    return rc;
  }
  rc = mpci::launch_transpose(d->DY_dev, p->DYt, d->n_Y_terms, rows, ngx, nullptr);
  if (!rc && (e = cudaDeviceSynchronize()) != cudaSuccess)
    rc = mpci::cuda_fail(e, "plan_create: transpose");
  if (rc) { mpci_plan_destroy(p); return rc; }

  int* base = static_cast<int*>(p->dev_tables);
  p->Y.nv = d->nv; p->Y.ng = d->ng; p->Y.nc = d->nc;
  p->Y.n_scalars = d->n_scalars; p->Y.n_terms = d->n_Y_terms;
  p->Y.factors = base;
  p->Y.ranges = base + p_ranges;
  p->Y.tile_ptr = base + p_tiles;
  p->Y.DYt = p->DYt;
  *out = p;
  return 0;
}

extern "C" int mpci_plan_destroy(mpci_plan* p) {
  if (!p) return 0;
  for (int s = 0; s < 2; ++s) {
    if (p->slot_in[s]) cudaFree(p->slot_in[s]);
    if (p->slot_out[s]) cudaFree(p->slot_out[s]);
    if (p->slot_free[s]) cudaEventDestroy(p->slot_free[s]);
    if (p->streams[s]) cudaStreamDestroy(p->streams[s]);
  }
  if (p->DYt) cudaFree(p->DYt);
  if (p->dev_tables) cudaFree(p->dev_tables);
  delete p;
  return 0;
}

extern "C" int mpci_assemble_launches(const mpci_plan* p) {
  if (!p) return 0;
  return (p->P.n_elems ? 1 : 0) + (p->G.n_elems ? 1 : 0) + ((p->nv + p->nc) ? 1 : 0);
}

extern "C" int mpci_assemble(const mpci_plan* p, const double* scalars, const double* given,
                             double* P, double* q, double* G, double* h, long batch,
                             void* stream) {
  if (!p) return mpci::fail(-3, "assemble: NULL plan");
  if (batch < 0) return mpci::fail(-1, "assemble: negative batch");
  if (batch == 0) return 0;
  if (!scalars) return mpci::fail(-3, "assemble: NULL scalars");
  if (p->ng > 0 && !given && (q || h)) return mpci::fail(-3, "assemble: NULL given");
  cudaStream_t st = mpci::as_stream(stream);
  // grid.y limits: split very large batches
  const long piece = 65535L * 16;
  for (long b0 = 0; b0 < batch; b0 += piece) {
    const long nb = (batch - b0 < piece) ? batch - b0 : piece;
    const double* s = scalars + b0 * p->n_scalars;
    int rc = 0;
    if (P) rc = mpci::launch_stream(p->P, s, p->n_scalars, P + b0 * p->P.n_elems, nb, st);
    if (!rc && G) rc = mpci::launch_stream(p->G, s, p->n_scalars, G + b0 * p->G.n_elems, nb, st);
    if (!rc)
      rc = mpci::launch_qh(p->Y, s, given ? given + b0 * p->ng : nullptr,
                           q ? q + b0 * p->nv : nullptr, h ? h + b0 * p->nc : nullptr, nb, st);
    if (rc) return rc;
  }
  return 0;
}

extern "C" int mpci_assemble_host(mpci_plan* p, const double* scalars, const double* given,
                                  double* P, double* q, double* G, double* h, long batch,
                                  long chunk) {
  if (!p) return mpci::fail(-3, "assemble_host: NULL plan");
  if (batch < 0 || chunk < 1) return mpci::fail(-1, "assemble_host: bad batch / chunk");
  if (batch == 0) return 0;
  if (!scalars || (p->ng > 0 && !given)) return mpci::fail(-3, "assemble_host: NULL input");
  if (chunk > batch) chunk = batch;
  const long in_per = p->n_scalars + p->ng;
  const long nP = p->P.n_elems, nG = p->G.n_elems;
  const long out_per = nP + p->nv + nG + p->nc;
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
      if (!p->streams[s]) MPCI_CUDA(cudaStreamCreateWithFlags(&p->streams[s], cudaStreamNonBlocking));
      if (!p->slot_free[s]) MPCI_CUDA(cudaEventCreateWithFlags(&p->slot_free[s], cudaEventDisableTiming));
    }
    p->slot_capacity = chunk;
  }
  int slot = 0;
  for (long b0 = 0; b0 < batch; b0 += chunk, slot ^= 1) {
    const long nb = (batch - b0 < chunk) ? batch - b0 : chunk;
    cudaStream_t st = p->streams[slot];
    double* d_s = p->slot_in[slot];
    double* d_g = d_s + (long)p->n_scalars * chunk;
    double* d_P = p->slot_out[slot];
    double* d_q = d_P + nP * chunk;
    double* d_G = d_q + (long)p->nv * chunk;
    double* d_h = d_G + nG * chunk;
    MPCI_CUDA(cudaMemcpyAsync(d_s, scalars + b0 * p->n_scalars, sizeof(double) * nb * p->n_scalars,
                              cudaMemcpyHostToDevice, st));
    if (p->ng)
      MPCI_CUDA(cudaMemcpyAsync(d_g, given + b0 * p->ng, sizeof(double) * nb * p->ng,
                                cudaMemcpyHostToDevice, st));
    int rc = mpci_assemble(p, d_s, d_g, P ? d_P : nullptr, q ? d_q : nullptr, G ? d_G : nullptr,
                           h ? d_h : nullptr, nb, st);
    if (rc) return rc;
    if (P) MPCI_CUDA(cudaMemcpyAsync(P + b0 * nP, d_P, sizeof(double) * nb * nP, cudaMemcpyDeviceToHost, st));
    if (q) MPCI_CUDA(cudaMemcpyAsync(q + b0 * p->nv, d_q, sizeof(double) * nb * p->nv, cudaMemcpyDeviceToHost, st));
    if (G) MPCI_CUDA(cudaMemcpyAsync(G + b0 * nG, d_G, sizeof(double) * nb * nG, cudaMemcpyDeviceToHost, st));
    if (h) MPCI_CUDA(cudaMemcpyAsync(h + b0 * p->nc, d_h, sizeof(double) * nb * p->nc, cudaMemcpyDeviceToHost, st));
  }
  MPCI_CUDA(cudaStreamSynchronize(p->streams[0]));
  MPCI_CUDA(cudaStreamSynchronize(p->streams[1]));
  return 0;
}
