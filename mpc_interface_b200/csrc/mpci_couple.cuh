// The coupling stage of an embedded per-instance plan (include/mpci.h: mpci_lti_embed):
// everything of P, q, G, h that involves the auxiliary optimisation variables or the
// auxiliary given variables.  Implemented in mpci_couple.cu, owned by an mpci_lti_plan.
#pragma once
#include "mpci_common.cuh"

namespace mpci {

struct CouplePlan;

// n, m, N: the plant as the fused kernel sees it.  Copies every table of `e` to the device.
int couple_create(int n, int m, int N, const mpci_lti_embed* e, CouplePlan** out);
void couple_destroy(CouplePlan* p);

// Device tables the fused kernel reads in the embedded layout (formulation constants): Gz
// [nc_total][nz] (the rows' entries under the auxiliary variables' columns), z_col [nz], and what
// it needs for the blocks of P that involve the auxiliary variables -- the coupled costs' plant
// forms cost_cx [nC][n] / cost_cu [nC][m] and weights, the columns of their Mz as CSR differences
// (d_ptr over (cost, z); d_tau, d_val), Pzz [nz][nz].
struct CoupleShare {
  const double *Gz, *Pzz, *cost_cx, *cost_cu, *cost_w, *d_val;
  const int *z_col, *d_ptr, *d_tau;
  int nz, n_costs;
};
void couple_share(const CouplePlan* p, CoupleShare* out);

// Runs after the fused kernel on the same stream (it adds to the q and h that kernel wrote; all
// of P is the fused kernel's).  ct: per-instance tables of the coupled costs (signal forms, weights,
// aims; include/mpci.h), nullptr or NULL members for the plan's.  Any output may be nullptr.
int couple_launch(const CouplePlan* p, const double* A, long strideA, const double* B, long strideB,
                  const double* x0, const double* given, const mpci_lti_coupled_tables* ct,
                  double* q, double* G, double* h, long batch, cudaStream_t stream);

}  // namespace mpci
