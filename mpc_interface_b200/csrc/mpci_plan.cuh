// Internal: the plan object behind the opaque mpci_plan handle and the
// argument blocks of the assembly kernels.
#pragma once
#include <cuda_runtime.h>

#include <vector>

#include "mpci.h"

namespace mpci {

// One streamed array (P or G): out[b][e] = sum_k c_k(b) * consts[k][e]
struct StreamArgs {
  int n_terms;
  int factors[MPCI_MAX_TERMS][MPCI_MAX_FACTORS];  // scalar indices of each monomial
  const double* consts;                            // dev [n_terms][n_elems]
  long n_elems;
};

struct QhArgs {
  int nv, ng, nc, n_scalars, n_terms;
  const int* factors;   // dev [n_terms][MPCI_MAX_FACTORS]
  const double* DYt;    // dev [n_terms][ng + 1][nv + nc]
  const int* ranges;    // dev [n_ranges][4]
  const int* tile_ptr;  // dev [n_tiles + 1]
};

int launch_stream(const StreamArgs& a, const double* scalars, int n_scalars, double* out,
                  long batch, cudaStream_t st);
int launch_qh(const QhArgs& a, const double* scalars, const double* given, double* q, double* h,
              long batch, cudaStream_t st);
int launch_transpose(const double* in, double* out, int n_terms, int rows, int cols,
                     cudaStream_t st);

}  // namespace mpci

struct mpci_plan {
  int nv, ng, nc, n_scalars;
  mpci::StreamArgs P, G;
  mpci::QhArgs Y;
  void* dev_tables;  // one allocation: factors | ranges | tile_ptr
  double* DYt;
  // staging of mpci_assemble_host (two slots), allocated on first use
  long slot_capacity;
  double* slot_in[2];   // scalars | given
  double* slot_out[2];  // P | q | G | h
  cudaStream_t streams[2];
  cudaEvent_t slot_free[2];
};
