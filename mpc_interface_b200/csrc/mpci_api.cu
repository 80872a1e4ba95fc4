// C ABI: plan life cycle, device-resident assembly, host-buffer assembly.
#include <algorithm>
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
  s.n_terms = n_terms;
  s.consts = consts;
  s.n_elems = consts ? n_elems : 0;  // without constants the plan cannot write this array
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
  if (!d->monomials || !d->P_terms || !d->G_terms)
    return mpci::fail(-3, "plan_create: NULL monomial / term table");
  if (d->n_Y_terms > mpci::kQhTermMask + 1)  // the records carry the term index in 8 bits
    return mpci::fail(-7, "plan_create: [q;h] has %d monomials (limit %d)", d->n_Y_terms,
                      mpci::kQhTermMask + 1);
  for (int i = 0; i < d->n_monomials * MPCI_MAX_FACTORS; ++i)
    if (d->monomials[i] >= d->n_scalars)
      return mpci::fail(-9, "plan_create: monomial factor %d out of %d scalars", d->monomials[i],
                        d->n_scalars);
  mpci_plan* p = new mpci_plan();  // members carry their own zero initialisers
  p->nv = d->nv; p->ng = d->ng; p->nc = d->nc; p->n_scalars = d->n_scalars;
  int rc = fill_stream(p->P, d->n_P_terms, d->P_terms, d->monomials, d->n_monomials, d->DP_dev,
                       (long)d->nv * d->nv, "P");
  if (!rc)
    rc = fill_stream(p->G, d->n_G_terms, d->G_terms, d->monomials, d->n_monomials, d->DG_dev,
                     (long)d->nc * d->nv, "G");
  const int rows = d->nv + d->nc, ngx = d->ng + 1;
  if (!rc && (d->n_Y_terms < 1 || !d->Y_terms))
    rc = mpci::fail(-3, "plan_create: the [q;h] terms are missing");
  if (!rc && (d->n_Y_segs < 0 || d->n_Y_nz < 0 ||
              (d->n_Y_segs > 0 && (!d->Y_segs || !d->Y_val || !d->Y_col))))
    rc = mpci::fail(-3, "plan_create: the sparse [q;h] rows are missing");
  if (!rc) {
    int last_row = 0, last_end = 0;
    for (int s = 0; s < d->n_Y_segs; ++s) {
      const int* g = d->Y_segs + 4 * s;  // (row, term, nz_begin, nz_end)
      if (g[0] < last_row || g[0] >= rows || g[1] < 0 || g[1] >= d->n_Y_terms ||
          g[2] != last_end || g[3] <= g[2] || g[3] > d->n_Y_nz) {
        rc = mpci::fail(-9, "plan_create: bad [q;h] segment %d (%d %d %d %d)", s, g[0], g[1],
                        g[2], g[3]);
        break;
      }
      last_row = g[0];
      last_end = g[3];
    }
    for (int e = 0; !rc && e < d->n_Y_nz; ++e)
      if (d->Y_col[e] < 0 || d->Y_col[e] >= ngx)
        rc = mpci::fail(-9, "plan_create: [q;h] column %d out of %d", d->Y_col[e], ngx);
  }
  for (int j = 0; !rc && j < d->n_Y_terms; ++j)
    if (d->Y_terms[j] < 0 || d->Y_terms[j] >= d->n_monomials)
      rc = mpci::fail(-9, "plan_create: Y term %d names monomial %d", j, d->Y_terms[j]);
  if (rc) { delete p; return rc; }

  // Row chunks: at most kQhChunk consecutive rows, never across the q / h border
  // (so a chunk is one contiguous run of q or of h).  Chunks are dealt to the
  // kQhLists lists longest-first (greedy balance of the non-zero count).
  struct Chunk { int row0, nrows, seg_begin, seg_end; long work; };
  std::vector<Chunk> chunks;
  {
    int s = 0;
    for (int part = 0; part < 2; ++part) {
      const int lo = part == 0 ? 0 : d->nv, hi = part == 0 ? d->nv : rows;
      for (int r0 = lo; r0 < hi; r0 += mpci::kQhChunk) {
        Chunk c{r0, std::min(mpci::kQhChunk, hi - r0), s, s, 0};
        while (s < d->n_Y_segs && d->Y_segs[4 * s] < r0 + c.nrows) {
          c.work += d->Y_segs[4 * s + 3] - d->Y_segs[4 * s + 2] + 2;
          ++s;
        }
        c.seg_end = s;
        c.work += c.nrows;
        chunks.push_back(c);
      }
    }
  }
  std::vector<int> order(chunks.size());
  for (size_t i = 0; i < order.size(); ++i) order[i] = (int)i;
  std::stable_sort(order.begin(), order.end(),
                   [&](int x, int y) { return chunks[x].work > chunks[y].work; });
  std::vector<std::vector<int>> per_warp(mpci::kQhLists);
  std::vector<long> load(mpci::kQhLists, 0);
  for (int idx : order) {
    int w = 0;
    for (int k = 1; k < mpci::kQhLists; ++k) if (load[k] < load[w]) w = k;
    per_warp[w].push_back(idx);
    load[w] += chunks[idx].work;
  }
  // column blocks of [given; 1]: as few passes as the shared memory of one block allows
  // (ring + row tiles are fixed, every staged column and every monomial costs 528 bytes)
  const long fixed_bytes = (long)sizeof(mpci::QhNz) * mpci::kQhWarps * 32 +
                           (long)sizeof(double) * mpci::kQhWarps * 64 * (mpci::kQhChunk + 1);
  const long col_budget = (227L * 1024 - fixed_bytes) / (66 * (long)sizeof(double)) - d->n_Y_terms;
  if (col_budget < 1) {
    delete p;
    return mpci::fail(-8, "plan_create: %d [q;h] monomials do not fit the kernel's shared memory",
                      d->n_Y_terms);
  }
  const int n_pass = (int)((ngx + col_budget - 1) / col_budget);
  if (n_pass > mpci::kQhMaxPass) {
    delete p;
    return mpci::fail(-8, "plan_create: given_len %d needs %d column passes (limit %d)", d->ng,
                      n_pass, mpci::kQhMaxPass);
  }
  const int cols_per = (ngx + n_pass - 1) / n_pass;
  for (int k = 0; k <= mpci::kQhMaxPass; ++k) p->Y.pass_col[k] = std::min(k * cols_per, ngx);
  p->Y.n_pass = n_pass;
  p->Y.max_cols = cols_per;
  // per-pass, per-list record streams: chunks -> rows -> (row, term) segments -> the
  // non-zeros whose column falls into the pass's block.  Pass 0 carries every chunk
  // (an all-zero row still has to be written); a later pass only the chunks it adds to.
  std::vector<mpci::QhNz> h_nz;
  std::vector<int> h_chunks, h_nz_ptr, h_chunk_ptr;
  for (int pass = 0; pass < n_pass; ++pass) {
    const int c_lo = p->Y.pass_col[pass], c_hi = p->Y.pass_col[pass + 1];
    h_nz_ptr.push_back((int)h_nz.size());
    h_chunk_ptr.push_back((int)h_chunks.size() / 2);
    for (int w = 0; w < mpci::kQhLists; ++w) {
      if (pass == 0) std::sort(per_warp[w].begin(), per_warp[w].end());
      for (int idx : per_warp[w]) {
        const Chunk& c = chunks[idx];
        if (pass > 0) {  // anything to add in this block of columns?
          bool any = false;
          for (int sg = c.seg_begin; sg < c.seg_end && !any; ++sg)
            for (int e = d->Y_segs[4 * sg + 2]; e < d->Y_segs[4 * sg + 3] && !any; ++e)
              any = d->Y_col[e] >= c_lo && d->Y_col[e] < c_hi;
          if (!any) continue;
        }
        h_chunks.push_back(c.row0);
        h_chunks.push_back(c.nrows);
        int sg = c.seg_begin;
        for (int r = 0; r < c.nrows; ++r) {
          const size_t row_start = h_nz.size();
          while (sg < c.seg_end && d->Y_segs[4 * sg] == c.row0 + r) {
            const int* g = d->Y_segs + 4 * sg;
            const size_t seg_start = h_nz.size();
            for (int e = g[2]; e < g[3]; ++e)
              if (d->Y_col[e] >= c_lo && d->Y_col[e] < c_hi)
                h_nz.push_back(mpci::QhNz{d->Y_val[e], (d->Y_col[e] - c_lo) * mpci::kQhColStride, g[1]});
            if (h_nz.size() > seg_start) h_nz.back().meta |= mpci::kQhEndSeg;
            ++sg;
          }
          if (h_nz.size() == row_start)  // a row without entries here: 0 (written / added)
            h_nz.push_back(mpci::QhNz{0.0, 0, mpci::kQhEndSeg});
          h_nz.back().meta |= mpci::kQhEndRow;
        }
        h_nz.back().meta |= mpci::kQhEndChunk;
      }
      h_nz_ptr.push_back((int)h_nz.size());
      h_chunk_ptr.push_back((int)h_chunks.size() / 2);
    }
  }
  std::vector<int> h_factors;
  for (int j = 0; j < d->n_Y_terms; ++j)
    for (int f = 0; f < MPCI_MAX_FACTORS; ++f)
      h_factors.push_back(d->monomials[d->Y_terms[j] * MPCI_MAX_FACTORS + f]);

  // one device allocation: nz | chunks | factors | nz_ptr | chunk_ptr  (16-byte aligned parts)
  auto align16 = [](size_t x) { return (x + 15) & ~size_t(15); };
  const size_t o_nz = 0;
  const size_t o_chunks = align16(o_nz + h_nz.size() * sizeof(mpci::QhNz));
  const size_t o_factors = align16(o_chunks + h_chunks.size() * sizeof(int));
  const size_t o_nzp = align16(o_factors + h_factors.size() * sizeof(int));
  const size_t o_chp = align16(o_nzp + h_nz_ptr.size() * sizeof(int));
  const size_t total = o_chp + h_chunk_ptr.size() * sizeof(int);
  std::vector<char> blob(total, 0);
  if (!h_nz.empty()) std::memcpy(&blob[o_nz], h_nz.data(), h_nz.size() * sizeof(mpci::QhNz));
  if (!h_chunks.empty()) std::memcpy(&blob[o_chunks], h_chunks.data(), h_chunks.size() * sizeof(int));
  std::memcpy(&blob[o_factors], h_factors.data(), h_factors.size() * sizeof(int));
  std::memcpy(&blob[o_nzp], h_nz_ptr.data(), h_nz_ptr.size() * sizeof(int));
  std::memcpy(&blob[o_chp], h_chunk_ptr.data(), h_chunk_ptr.size() * sizeof(int));
  cudaError_t e = cudaMalloc(&p->dev_tables, total);
  if (e == cudaSuccess) e = cudaMemcpy(p->dev_tables, blob.data(), total, cudaMemcpyHostToDevice);
  if (e != cudaSuccess) {
    rc = mpci::cuda_fail(e, "plan_create: device tables");
    mpci_plan_destroy(p);
    return rc;
  }
  const char* base = static_cast<const char*>(p->dev_tables);
  p->Y.nv = d->nv; p->Y.ng = d->ng; p->Y.nc = d->nc;
  p->Y.n_scalars = d->n_scalars; p->Y.n_terms = d->n_Y_terms;
  p->Y.nz = reinterpret_cast<const mpci::QhNz*>(base + o_nz);
  p->Y.chunks = reinterpret_cast<const int2*>(base + o_chunks);
  p->Y.factors = reinterpret_cast<const int*>(base + o_factors);
  p->Y.warp_nz_ptr = reinterpret_cast<const int*>(base + o_nzp);
  p->Y.warp_chunk_ptr = reinterpret_cast<const int*>(base + o_chp);
  *out = p;
  return 0;
}

extern "C" int mpci_plan_destroy(mpci_plan* p) {
  if (!p) return 0;
  for (int s = 0; s < 2; ++s) {
    if (p->slot_in[s]) cudaFree(p->slot_in[s]);
    if (p->slot_out[s]) cudaFree(p->slot_out[s]);
    if (p->slot_done[s]) cudaEventDestroy(p->slot_done[s]);
    if (p->streams[s]) cudaStreamDestroy(p->streams[s]);
  }
  if (p->side) cudaStreamDestroy(p->side);
  if (p->fork) cudaEventDestroy(p->fork);
  if (p->join) cudaEventDestroy(p->join);
  if (p->dev_tables) cudaFree(p->dev_tables);
  for (cudaEvent_t e : p->timers) cudaEventDestroy(e);
  delete p;
  return 0;
}

extern "C" int mpci_assemble_launches(const mpci_plan* p) {
  if (!p) return 0;
  return (p->P.n_elems ? 1 : 0) + (p->G.n_elems ? 1 : 0) + ((p->nv + p->nc) ? 1 : 0);
}

extern "C" int mpci_plan_set_timing(mpci_plan* p, int on) {
  if (!p) return mpci::fail(-3, "plan_set_timing: NULL plan");
  for (cudaEvent_t e : p->timers) cudaEventDestroy(e);
  p->timers.clear();
  p->timing = on != 0;
  return 0;
}

extern "C" int mpci_plan_read_timing(mpci_plan* p, double ms[3], long* pieces) {
  if (!p || !ms) return mpci::fail(-3, "plan_read_timing: NULL plan or output");
  ms[0] = ms[1] = ms[2] = 0.0;
  const long n = static_cast<long>(p->timers.size() / 6);
  for (long c = 0; c < n; ++c) {
    for (int k = 0; k < 3; ++k) {
      cudaEvent_t a = p->timers[6 * c + 2 * k], b = p->timers[6 * c + 2 * k + 1];
      MPCI_CUDA(cudaEventSynchronize(b));
      float t = 0.f;
      MPCI_CUDA(cudaEventElapsedTime(&t, a, b));
      ms[k] += t;
    }
  }
  if (pieces) *pieces = n;
  return 0;
}

namespace {
int assemble_with(const mpci_plan* p, const mpci::StreamArgs& Pa, const mpci::StreamArgs& Ga,
                  const double* scalars, const double* given, double* P, double* q, double* G,
                  double* h, long batch, void* stream);
}

extern "C" int mpci_assemble(const mpci_plan* p, const double* scalars, const double* given,
                             double* P, double* q, double* G, double* h, long batch,
                             void* stream) {
  if (!p) return mpci::fail(-3, "assemble: NULL plan");
  return assemble_with(p, p->P, p->G, scalars, given, P, q, G, h, batch, stream);
}

extern "C" int mpci_plan_set_compact(mpci_plan* p, const double* DPc, long nP, const double* DGc,
                                     long nG) {
  if (!p) return mpci::fail(-3, "plan_set_compact: NULL plan");
  if (nP < 0 || nG < 0 || (nP > 0 && !DPc) || (nG > 0 && !DGc))
    return mpci::fail(-1, "plan_set_compact: bad sizes or NULL constants");
  p->Pc = p->P; p->Pc.consts = DPc; p->Pc.n_elems = nP;
  p->Gc = p->G; p->Gc.consts = DGc; p->Gc.n_elems = nG;
  return 0;
}

extern "C" int mpci_assemble_compact(const mpci_plan* p, const double* scalars,
                                     const double* given, double* Pc, double* q, double* Gc,
                                     double* h, long batch, void* stream) {
  if (!p) return mpci::fail(-3, "assemble_compact: NULL plan");
  if ((Pc && !p->Pc.consts) || (Gc && !p->Gc.consts))
    return mpci::fail(-3, "assemble_compact: mpci_plan_set_compact was not called");
  return assemble_with(p, p->Pc, p->Gc, scalars, given, Pc, q, Gc, h, batch, stream);
}

namespace {
int assemble_with(const mpci_plan* p, const mpci::StreamArgs& Pa, const mpci::StreamArgs& Ga,
                  const double* scalars, const double* given, double* P, double* q, double* G,
                  double* h, long batch, void* stream) {
  if (batch < 0) return mpci::fail(-1, "assemble: negative batch");
  if (batch == 0) return 0;
  if (!scalars) return mpci::fail(-3, "assemble: NULL scalars");
  if (p->ng > 0 && !given && (q || h)) return mpci::fail(-3, "assemble: NULL given");
  if ((P && !Pa.consts) || (G && !Ga.consts))
    return mpci::fail(-3, "assemble: the plan was created without the constants of P / G");
  cudaStream_t st = mpci::as_stream(stream);
  // grid.y limits: split very large batches
  const long piece = 65535L * 16;
  // [q;h] first, on the side stream, so that its blocks (one per SM) are resident
  // while the P / G streams saturate HBM with the remaining thread slots
  const bool overlap = (P || G) && (q || h) && (p->nv + p->nc) > 0;
  cudaStream_t qh_stream = st;
  if (overlap) {
    if (!p->side) {
      MPCI_CUDA(cudaStreamCreateWithFlags(&p->side, cudaStreamNonBlocking));
      MPCI_CUDA(cudaEventCreateWithFlags(&p->fork, cudaEventDisableTiming));
      MPCI_CUDA(cudaEventCreateWithFlags(&p->join, cudaEventDisableTiming));
    }
    MPCI_CUDA(cudaEventRecord(p->fork, st));
    MPCI_CUDA(cudaStreamWaitEvent(p->side, p->fork, 0));
    qh_stream = p->side;
  }
  for (long b0 = 0; b0 < batch; b0 += piece) {
    const long nb = (batch - b0 < piece) ? batch - b0 : piece;
    const double* s = scalars + b0 * p->n_scalars;
    cudaEvent_t* ev = nullptr;  // P begin/end, G begin/end, [q;h] begin/end
    if (p->timing) {
      const size_t at = p->timers.size();
      p->timers.resize(at + 6);
      for (int i = 0; i < 6; ++i) MPCI_CUDA(cudaEventCreate(&p->timers[at + i]));
      ev = &p->timers[at];
    }
    if (ev) MPCI_CUDA(cudaEventRecord(ev[4], qh_stream));
    int rc = mpci::launch_qh(p->Y, s, given ? given + b0 * p->ng : nullptr,
                             q ? q + b0 * p->nv : nullptr, h ? h + b0 * p->nc : nullptr, nb,
                             qh_stream);
    if (ev) MPCI_CUDA(cudaEventRecord(ev[5], qh_stream));
    if (ev) MPCI_CUDA(cudaEventRecord(ev[0], st));
    if (!rc && P) rc = mpci::launch_stream(Pa, s, p->n_scalars, P + b0 * Pa.n_elems, nb, st);
    if (ev) MPCI_CUDA(cudaEventRecord(ev[1], st));
    if (ev) MPCI_CUDA(cudaEventRecord(ev[2], st));
    if (!rc && G) rc = mpci::launch_stream(Ga, s, p->n_scalars, G + b0 * Ga.n_elems, nb, st);
    if (ev) MPCI_CUDA(cudaEventRecord(ev[3], st));
    if (rc) return rc;
  }
  if (overlap) {
    MPCI_CUDA(cudaEventRecord(p->join, p->side));
    MPCI_CUDA(cudaStreamWaitEvent(st, p->join, 0));
  }
  return 0;
}
}  // namespace

namespace {
int assemble_host_with(mpci_plan* p, const mpci::StreamArgs& Pa, const mpci::StreamArgs& Ga,
                       const double* scalars, const double* given, double* P, double* q,
                       double* G, double* h, long batch, long chunk, long ring,
                       mpci_chunk_fn on_chunk, void* user);
}

extern "C" int mpci_assemble_host(mpci_plan* p, const double* scalars, const double* given,
                                  double* P, double* q, double* G, double* h, long batch,
                                  long chunk) {
  if (!p) return mpci::fail(-3, "assemble_host: NULL plan");
  return assemble_host_with(p, p->P, p->G, scalars, given, P, q, G, h, batch, chunk, 0, nullptr,
                            nullptr);
}

extern "C" int mpci_assemble_host_stream(mpci_plan* p, const double* scalars, const double* given,
                                         double* P, double* q, double* G, double* h, long batch,
                                         long chunk, long ring, mpci_chunk_fn on_chunk,
                                         void* user) {
  if (!p) return mpci::fail(-3, "assemble_host_stream: NULL plan");
  if (ring < 1) return mpci::fail(-1, "assemble_host_stream: ring must hold at least one chunk");
  return assemble_host_with(p, p->P, p->G, scalars, given, P, q, G, h, batch, chunk, ring,
                            on_chunk, user);
}

extern "C" int mpci_assemble_compact_host(mpci_plan* p, const double* scalars,
                                          const double* given, double* Pc, double* q, double* Gc,
                                          double* h, long batch, long chunk) {
  if (!p) return mpci::fail(-3, "assemble_compact_host: NULL plan");
  if ((Pc && !p->Pc.consts) || (Gc && !p->Gc.consts))
    return mpci::fail(-3, "assemble_compact_host: mpci_plan_set_compact was not called");
  return assemble_host_with(p, p->Pc, p->Gc, scalars, given, Pc, q, Gc, h, batch, chunk, 0,
                            nullptr, nullptr);
}

extern "C" int mpci_assemble_compact_host_stream(mpci_plan* p, const double* scalars,
                                                 const double* given, double* Pc, double* q,
                                                 double* Gc, double* h, long batch, long chunk,
                                                 long ring, mpci_chunk_fn on_chunk, void* user) {
  if (!p) return mpci::fail(-3, "assemble_compact_host_stream: NULL plan");
  if (ring < 1) return mpci::fail(-1, "assemble_compact_host_stream: ring must hold at least one chunk");
  if ((Pc && !p->Pc.consts) || (Gc && !p->Gc.consts))
    return mpci::fail(-3, "assemble_compact_host_stream: mpci_plan_set_compact was not called");
  return assemble_host_with(p, p->Pc, p->Gc, scalars, given, Pc, q, Gc, h, batch, chunk, ring,
                            on_chunk, user);
}

namespace {
// Chunks of instances alternate between two device slots / streams (upload, the three
// kernels, download), so the copies of one chunk overlap the kernels of the other.
// ring == 0: the host arrays hold the whole batch (chunk c lands at instance c * chunk);
// ring >= 1: they hold `ring` chunks and chunk c lands in slot c % ring -- logical batches
// larger than host memory -- with `on_chunk` called (on this thread) once a chunk is
// complete in host memory and before its slot is written again.
int assemble_host_with(mpci_plan* p, const mpci::StreamArgs& Pa, const mpci::StreamArgs& Ga,
                       const double* scalars, const double* given, double* P, double* q,
                       double* G, double* h, long batch, long chunk, long ring,
                       mpci_chunk_fn on_chunk, void* user) {
  if (batch < 0 || chunk < 1) return mpci::fail(-1, "assemble_host: bad batch / chunk");
  if (batch == 0) return 0;
  if (!scalars || (p->ng > 0 && !given)) return mpci::fail(-3, "assemble_host: NULL input");
  if (chunk > batch) chunk = batch;
  const long n_chunks = (batch + chunk - 1) / chunk;
  if (ring == 0) ring = n_chunks;
  if (ring < 2 && n_chunks > 1)
    return mpci::fail(-1, "assemble_host: a ring of one slot cannot overlap two chunks");
  const long in_per = p->n_scalars + p->ng;
  const long nP = Pa.n_elems, nG = Ga.n_elems;
  const long out_per = nP + p->nv + nG + p->nc;
  if (p->slot_capacity < chunk || p->slot_out_doubles < out_per * chunk) {
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
      if (!p->slot_done[s]) MPCI_CUDA(cudaEventCreateWithFlags(&p->slot_done[s], cudaEventDisableTiming));
    }
    p->slot_capacity = chunk;
    p->slot_out_doubles = out_per * chunk;
  }
  for (long c = 0; c <= n_chunks; ++c) {
    if (c < n_chunks) {
      const int slot = (int)(c & 1);
      const long b0 = c * chunk, nb = (batch - b0 < chunk) ? batch - b0 : chunk;
      const long at = (c % ring) * chunk;  // first instance of the host slot
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
      int rc = assemble_with(p, Pa, Ga, d_s, d_g, P ? d_P : nullptr, q ? d_q : nullptr,
                             G ? d_G : nullptr, h ? d_h : nullptr, nb, st);
      if (rc) return rc;
      if (P) MPCI_CUDA(cudaMemcpyAsync(P + at * nP, d_P, sizeof(double) * nb * nP, cudaMemcpyDeviceToHost, st));
      if (q) MPCI_CUDA(cudaMemcpyAsync(q + at * p->nv, d_q, sizeof(double) * nb * p->nv, cudaMemcpyDeviceToHost, st));
      if (G) MPCI_CUDA(cudaMemcpyAsync(G + at * nG, d_G, sizeof(double) * nb * nG, cudaMemcpyDeviceToHost, st));
      if (h) MPCI_CUDA(cudaMemcpyAsync(h + at * p->nc, d_h, sizeof(double) * nb * p->nc, cudaMemcpyDeviceToHost, st));
      MPCI_CUDA(cudaEventRecord(p->slot_done[slot], st));
    }
    if (c >= 1) {  // chunk c - 1 is complete (chunk c keeps the device busy meanwhile)
      const long done = c - 1, b0 = done * chunk;
      MPCI_CUDA(cudaEventSynchronize(p->slot_done[done & 1]));
      if (on_chunk) on_chunk(user, b0, (batch - b0 < chunk) ? batch - b0 : chunk, done % ring);
    }
  }
  return 0;
}
}  // namespace
