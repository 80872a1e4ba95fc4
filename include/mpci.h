/* mpci.h -- C ABI of the B200 QP-assembly library (libmpci.so).
 *
 * Drop-in boundary for the matrix-generation path of Gepetto/mpc-interface.
 * The reference has no FFI layer: its boundary is the Python object API
 * (python/mpc_interface/body.py:333-348 `generate_all_qp_matrices`,
 * python/mpc_interface/tools.py:14-33 `extend_matrices`) and one C++ free
 * function, gecko::tools::extend_matrices (cpp/include/mpc-interface/tools.hh:24-25,
 * cpp/src/tools.cc:83-144).  Every entry point below names the reference
 * routine it replaces.  Plain pointers and sizes only; all matrices are
 * row-major (C order) float64, exactly the layout of the reference's NumPy
 * arrays.  "dev" pointers are CUDA device pointers, "host" pointers are host
 * memory (pinned memory gives full PCIe speed).  `stream` is a cudaStream_t
 * passed as void* (NULL = default stream); launches are asynchronous.
 *
 * Return value: 0 = ok, negative = invalid argument (see mpci_last_error()),
 * positive = CUDA error code.  Nothing throws across this boundary.
 */
#ifndef MPCI_H
#define MPCI_H

#ifdef __cplusplus
extern "C" {
#endif

#define MPCI_VERSION 100
#define MPCI_MAX_FACTORS 4   /* per-instance scalar factors of one monomial */
#define MPCI_MAX_TERMS 32    /* monomials of one streamed array (P or G); up to 8 keep the
                              * constants in registers, more are re-read through L1 / L2 */

int mpci_version(void);

/* Thread-local description of the last error returned on this thread. */
const char* mpci_last_error(void);

/* Number of CUDA devices visible (0 without a driver / GPU). */
int mpci_device_count(void);

/* ---- K1: prediction matrices -------------------------------------------
 * Replaces tools.extend_matrices (python/mpc_interface/tools.py:14-33) and
 * gecko::tools::extend_matrices (cpp/src/tools.cc:83-144) for a batch of
 * independent systems x+ = A x + B u.
 *   A_dev [batch][n][n], B_dev [batch][n][m]
 *   S_dev [batch][N][n][n]      S[k][c][i]    = (A^(k+1))[i][c]
 *   U_dev [batch][m][N][N][n]   U[j][k][p][i] = (A^(k-p) B)[i][j], 0 for p > k
 * (U[j] is the reference's list element j; the rank-4 order m,N,N,n is the one
 * of the C++ tensor, cpp/src/tools.cc:139-143.)  S_dev or U_dev may be NULL.
 */
int mpci_extend_matrices(const double* A_dev, const double* B_dev, double* S_dev,
                         double* U_dev, int batch, int n, int m, int N, void* stream);

/* Compact form (what the assembly kernels consume):
 *   Phi_dev [batch][N][n][n]  Phi[k][i][c] = (A^(k+1))[i][c]
 *   Gam_dev [batch][N][n][m]  Gam[d][i][j] = (A^d B)[i][j]                    */
int mpci_extend_compact(const double* A_dev, const double* B_dev, double* Phi_dev,
                        double* Gam_dev, int batch, int n, int m, int N, void* stream);

/* Host-pointer convenience: copies A, B in, runs K1, copies S, U out. */
int mpci_extend_matrices_host(const double* A_host, const double* B_host, double* S_host,
                              double* U_host, int batch, int n, int m, int N);

/* K1 without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid it).  Same arguments and results as
 * mpci_extend_matrices, by its three-kernel path, with the intermediate (Phi, Gam) in a
 * static device array of MPCI_EXTEND_STATIC_MAX_DOUBLES doubles (4 MiB, held by every GPU
 * that loads the library until the process exits):
 *     batch * N * n * (n [if S] + m [if U])  <=  MPCI_EXTEND_STATIC_MAX_DOUBLES,
 * larger problems are refused with code -8 and a message that names both numbers.  Calls
 * neither cudaMalloc nor cudaFree nor any other allocating routine; the caller passes every
 * input and output.  All calls on one GPU share the array: two calls that overlap on that GPU (other
 * streams, other host threads) race -- serialise them.  Nothing in the array needs clearing
 * between calls (every element read was written by the same call).                         */
#define MPCI_EXTEND_STATIC_MAX_DOUBLES 524288
int mpci_extend_matrices_static(const double* A_dev, const double* B_dev, double* S_dev,
                                double* U_dev, int batch, int n, int m, int N, void* stream);
long mpci_extend_static_capacity(void); /* MPCI_EXTEND_STATIC_MAX_DOUBLES of the loaded library */

/* ---- dense building blocks used while a plan is built -------------------
 * The NumPy dots of body.py:184-191 (definitions), :254-255 (L-type limits)
 * and :279-292 (cost products) for the constant coefficient matrices.
 * C = alpha * op(A) @ B + beta * C on FP64 tensor cores (DMMA m8n8k4),
 * op(A) = A (M x K, lda) or A^T (A stored K x M, lda) when transA != 0.
 * Strided batch: matrix i starts at X + i * strideX.                        */
int mpci_gemm(int transA, int M, int N, int K, double alpha, const double* A_dev, int lda,
              long strideA, const double* B_dev, int ldb, long strideB, double beta,
              double* C_dev, int ldc, long strideC, int batch, void* stream);

/* dst[i][j] += alpha * rowscale[i or 0] * src[i or 0][j]   (rows x cols);
 * src has src_rows (1 = broadcast) rows, rowscale n_rowscale (0 = none,
 * 1 = broadcast) entries: the `c * Mo[schedule]` of body.py:258-259.       */
int mpci_axpy2d(int rows, int cols, double alpha, const double* rowscale_dev, int n_rowscale,
                const double* src_dev, int ld_src, int src_rows, double* dst_dev, int ld_dst,
                void* stream);

/* ---- the assembly plan ---------------------------------------------------
 * One plan = one formulation structure (index layout + constant matrices).
 * Per instance b the kernels evaluate
 *     c_m(b)    = prod_f scalars[b][monomials[m][f]]        (index -1 = skip)
 *     P[b]      = sum_t c_{P_terms[t]}(b) * DP[t]            nv x nv
 *     G[b]      = sum_t c_{G_terms[t]}(b) * DG[t]            nc x nv
 *     [q; h][b] = sum_t c_{Y_terms[t]}(b) * DY[t] @ [given[b]; 1]
 * which is Formulation.generate_all_qp_matrices (body.py:333-348) for a
 * formulation whose per-instance numbers are the scalars and `given`.       */
typedef struct {
  int nv, ng, nc;       /* optim_len, given_len, constraint rows             */
  int n_scalars;        /* per-instance scalars (>= 1; column 0 may be dummy) */
  int n_monomials;
  const int* monomials; /* host [n_monomials][MPCI_MAX_FACTORS]               */
  int n_P_terms;        /* <= MPCI_MAX_TERMS                                  */
  const int* P_terms;   /* host [n_P_terms] indices into monomials            */
  const double* DP_dev; /* dev  [n_P_terms][nv][nv]                           */
  int n_G_terms;        /* <= MPCI_MAX_TERMS                                  */
  const int* G_terms;
  const double* DG_dev; /* dev  [n_G_terms][nc][nv]                           */
  int n_Y_terms;
  const int* Y_terms;
  /* the constants DY[t] ([nv + nc] x [ng + 1]) as sparse rows: segment s =
   * (row, term, nz_begin, nz_end), sorted by row; nz e = (Y_val[e], Y_col[e])
   * with Y_col in [0, ng] (ng = the constant column).  All host pointers.   */
  int n_Y_segs;
  const int* Y_segs;    /* host [n_Y_segs][4]                                 */
  int n_Y_nz;
  const double* Y_val;  /* host [n_Y_nz]                                      */
  const int* Y_col;     /* host [n_Y_nz]                                      */
} mpci_plan_desc;

typedef struct mpci_plan mpci_plan;

/* A plan handle is used by one host thread at a time (the calls keep per-plan streams,
 * events and staging slots); different plans may be driven from different threads.
 * Copies the small host tables; keeps the DP/DG/DY device pointers (owned by
 * the caller, must outlive the plan) and a device copy of the sparse Y rows. */
int mpci_plan_create(const mpci_plan_desc* desc, mpci_plan** plan_out);
int mpci_plan_destroy(mpci_plan* plan);

/* Device-resident assembly: replaces generate_all_qp_matrices for `batch`
 * instances.  scalars_dev [batch][n_scalars], given_dev [batch][ng];
 * outputs P_dev [batch][nv][nv], q_dev [batch][nv], G_dev [batch][nc][nv],
 * h_dev [batch][nc], row-major and unpadded.  Any output may be NULL.       */
int mpci_assemble(const mpci_plan* plan, const double* scalars_dev, const double* given_dev,
                  double* P_dev, double* q_dev, double* G_dev, double* h_dev, long batch,
                  void* stream);

/* Same with HOST buffers: inputs are copied to the device and results back in
 * chunks of `chunk` instances on two streams (copy/compute overlap).  The plan
 * owns the device staging buffers (allocated on first use, freed by destroy). */
int mpci_assemble_host(mpci_plan* plan, const double* scalars_host, const double* given_host,
                       double* P_host, double* q_host, double* G_host, double* h_host,
                       long batch, long chunk);

/* Streaming form for logical batches larger than host memory (BASELINE configs[2..4]:
 * 1.2 TB / 337 GB of results): the host output arrays hold `ring` chunks of `chunk`
 * instances and chunk c lands in slot c % ring (instance (c % ring) * chunk of every
 * array).  `on_chunk` (may be NULL) is called on the calling thread when a chunk is
 * complete in host memory and before its slot is written again -- the place where the
 * caller's solver takes the QPs (the reference hands them to the solver one by one,
 * biped_mpc_loop.py:57-60).  ring >= 2 unless the batch is a single chunk.           */
typedef void (*mpci_chunk_fn)(void* user, long first_instance, long count, long slot);
int mpci_assemble_host_stream(mpci_plan* plan, const double* scalars_host,
                              const double* given_host, double* P_ring, double* q_ring,
                              double* G_ring, double* h_ring, long batch, long chunk, long ring,
                              mpci_chunk_fn on_chunk, void* user);

/* ---- after the solve: preview and goal distance ---------------------------
 * Formulation.preview (body.py:209-219), v = Mg given + Mo optim, is the same
 * product as [q; h] with [given; optim] in the place of given: build a plan whose
 * Y constants are the variable's preview matrices (nv = rows, ng = given + optim,
 * nc = 0, DP_dev = DG_dev = NULL) and call mpci_assemble with only q requested.
 * Formulation.goal_distance (body.py:221-229):
 *   out[b] = sum_a sum_r (V[b][a][r] - aim[b * aim_stride + a])^2
 * V_dev [batch][n_axes][rows], aim_dev [batch][>= n_axes] (aim_stride = 0: one aim
 * row for the whole batch), out_dev [batch].                                   */
int mpci_goal_distance(const double* V_dev, const double* aim_dev, long aim_stride, int rows,
                       int n_axes, double* out_dev, long batch, void* stream);

/* ---- solver hand-off formats ---------------------------------------------
 * The example feeds the solver scipy.sparse.csc_matrix(Q) / csc_matrix(A)
 * (python/use_examples/simple_functional_example/biped_mpc_loop.py:57-60).  The
 * sparsity pattern of P and G is the same for every instance of a batch, so the
 * host derives it once (indices / indptr, bit-exact integers) and hands the
 * library the constants gathered in that order; the same streaming kernels then
 * write only the values:  Pc[b][e] = sum_t c_t(b) * DPc[t][e].
 * DPc_dev [n_P_terms][nP], DGc_dev [n_G_terms][nG] (device, caller-owned; nP / nG
 * even for 16-byte stores).  Any entry order works: CSC, upper triangle, ...   */
int mpci_plan_set_compact(mpci_plan* plan, const double* DPc_dev, long nP,
                          const double* DGc_dev, long nG);
int mpci_assemble_compact(const mpci_plan* plan, const double* scalars_dev,
                          const double* given_dev, double* Pc_dev, double* q_dev,
                          double* Gc_dev, double* h_dev, long batch, void* stream);
int mpci_assemble_compact_host(mpci_plan* plan, const double* scalars_host,
                               const double* given_host, double* Pc_host, double* q_host,
                               double* Gc_host, double* h_host, long batch, long chunk);

/* Box bounds for solvers that take lb <= z <= ub beside G z <= h (qpsolvers' `lb`, `ub`; the
 * reference's example passes everything as rows of A, biped_mpc_loop.py:57-60).  The rows of G
 * that bound ONE variable with a coefficient that does not depend on the instance (the stepping
 * area and every `Box` / `Constraint` on an optimisation variable itself, restrictions.py:175-199)
 * are found once per plan on the host; per instance only h differs:
 *   for x in col_ptr[c] .. col_ptr[c+1]:  a = coef[x], v = h[b][row[x]] / a;
 *        a > 0: ub[b][c] = min(ub, v)      a < 0: lb[b][c] = max(lb, v)
 * h_dev [batch][ld_h]; col_ptr_dev [nv+1], row_dev / coef_dev [col_ptr[nv]] (device);
 * lb_dev / ub_dev [batch][nv], -inf / +inf where a variable has no such row.                */
int mpci_box_bounds(const double* h_dev, long ld_h, int nv, const int* col_ptr_dev,
                    const int* row_dev, const double* coef_dev, double* lb_dev, double* ub_dev,
                    long batch, void* stream);

int mpci_assemble_compact_host_stream(mpci_plan* plan, const double* scalars_host,
                                      const double* given_host, double* Pc_ring, double* q_ring,
                                      double* Gc_ring, double* h_ring, long batch, long chunk,
                                      long ring, mpci_chunk_fn on_chunk, void* user);

/* Number of kernel launches one mpci_assemble call issues for this plan.    */
int mpci_assemble_launches(const mpci_plan* plan);

/* Kernel timers (measurement only; the reference has no counterpart).  While
 * on, every mpci_assemble / mpci_assemble_compact call on this plan brackets
 * its launches with CUDA events on the streams they run on, so that a bench
 * reads kernel durations from INSIDE its timed region.  set_timing(on) clears
 * what was collected.  read_timing waits for the recorded events and returns
 * the summed durations in ms: ms[0] P stream, ms[1] G stream, ms[2] [q;h]
 * (side stream, concurrent with the other two); *pieces = number of
 * (call, batch piece) records summed.                                        */
int mpci_plan_set_timing(mpci_plan* plan, int on);
int mpci_plan_read_timing(mpci_plan* plan, double ms[3], long* pieces);

/* ---- per-instance dynamics: the condensed LTI problem ---------------------
 * Mode B of DESIGN.md: every instance b has its own x+ = A_b x + B_b u, so the
 * prediction matrices (tools.py:14-33), the Hessian contraction
 * sum_c w_c Vo^T Co (body.py:292), the gradient (body.py:293-300) and the
 * constraint rows (body.py:236-264) are all computed per instance by one fused
 * kernel.  The formulation is lowered (mpc_interface_b200/lti.py) to stage
 * tables over the n states z_i and m inputs u_j of the plant:
 *   Qxx [n][n], Quu [m][m]   sum over costs of w * (variable form) x (cross form)
 *   Tx  [n][n_scalars + 1]   (w/2)(cv*caim + cc*aim) as an affine map of the
 *   Tu  [m][n_scalars + 1]    per-instance scalars (last column = constant)
 *   groups: limit g has N rows  arrow[k] * (cx . x_{k+1} + cu . u_k) <= bound[k]
 * Optimisation vector: input-major, z = [u_0(0..N-1), u_1(0..N-1), ...] (the
 * reference's optim_ID order for such a formulation); given = x0.
 *   P[(j,p),(j',p')] = sum_{k>=max(p,p')} Gam_{k-p}[:,j]^T Qxx Gam_{k-p'}[:,j'] + Quu[j][j'] d_pp'
 *   q[(j,p)]         = sum_{k>=p} Gam_{k-p}[:,j]^T (sym(Qxx) A^{k+1} x0 - tx) - tu[j]
 *   G[g*N+k,(j,p)]   = arrow[k] * (cx . Gam_{k-p}[:,j] [p<=k] + cu[j] [p==k])
 *   h[g*N+k]         = bound[k] - arrow[k] * cx . A^{k+1} x0
 * with Gam_d = A^d B.  All tables are host pointers, copied at creation.      */
typedef struct {
  int n, m, N;
  int n_scalars;           /* columns of Tx/Tu minus one (may be 0)           */
  int n_groups;            /* nc = n_groups * N                               */
  const double* Qxx;       /* [n][n]                                          */
  const double* Quu;       /* [m][m]                                          */
  const double* Tx;        /* [n][n_scalars + 1]                              */
  const double* Tu;        /* [m][n_scalars + 1]                              */
  const double* grp_cx;    /* [n_groups][n]                                   */
  const double* grp_cu;    /* [n_groups][m]                                   */
  const double* grp_arrow; /* [n_groups][N]                                   */
  const double* grp_bound; /* [n_groups][N]                                   */
  /* rows of every group: horizon samples grp_k0[g] .. grp_k0[g] + grp_rows[g] - 1 (a
   * limit's `schedule`, body.py:246; restrictions.py:147-162).  NULL = all N samples.
   * nc = sum of grp_rows; the tables above stay [n_groups][N], indexed by the sample. */
  const int* grp_k0;       /* [n_groups] or NULL                              */
  const int* grp_rows;     /* [n_groups] or NULL                              */
} mpci_lti_desc;

typedef struct mpci_lti_plan mpci_lti_plan;

int mpci_lti_plan_create(const mpci_lti_desc* desc, mpci_lti_plan** plan_out);
int mpci_lti_plan_destroy(mpci_lti_plan* plan);

/* hessian_mode: 0 = dense contraction on the FP64 tensor cores (DMMA) with
 * block-triangular skipping; 1 = displacement recursion
 * P[p,p'] = P[p+1,p'+1] + Gam_{N-1-p}^T Qxx Gam_{N-1-p'} (O(N^2) flops).     */
#define MPCI_LTI_HESSIAN_DMMA 0
#define MPCI_LTI_HESSIAN_RECURSION 1

/* A_dev [batch][n][n] (strideA = n*n) or one shared matrix (strideA = 0);
 * B_dev likewise ([n][m]); x0_dev [batch][n]; scalars_dev [batch][n_scalars]
 * (may be NULL when n_scalars == 0).  Outputs as in mpci_assemble with
 * nv = m*N, nc = n_groups*N; any output may be NULL.                          */
int mpci_lti_assemble(const mpci_lti_plan* plan, const double* A_dev, long strideA,
                      const double* B_dev, long strideB, const double* x0_dev,
                      const double* scalars_dev, double* P_dev, double* q_dev, double* G_dev,
                      double* h_dev, long batch, int hessian_mode, void* stream);

/* Per-instance tables: cost weights, form coefficients or bounds that differ between the
 * instances (a `Param` inside a LineCombo / Cost / Constraint of a formulation lowered to
 * this path: e.g. the LIP's CoP  b = CoM - CoM_ddot / omega_b^2  with one omega per robot).
 * Every non-NULL pointer replaces the plan's table of the same name for this call; its
 * instance stride is the table size (Qxx [batch][n][n], Quu [batch][m][m], tx [batch][n]
 * and tu [batch][m] -- the offsets of the gradient, replacing Tx / Tu x scalars --,
 * cx [batch][n_groups][n], cu [batch][n_groups][m], bound [batch][n_groups][N]).
 * Device pointers.  The plan must have been created with n_scalars = 0 when tx / tu are
 * given, and with non-zero grp_cx entries wherever an instance's cx may be non-zero.      */
typedef struct {
  const double* Qxx;
  const double* Quu;
  const double* tx;
  const double* tu;
  const double* cx;
  const double* cu;
  const double* bound;
} mpci_lti_tables;
int mpci_lti_assemble_ex(const mpci_lti_plan* plan, const double* A_dev, long strideA,
                         const double* B_dev, long strideB, const double* x0_dev,
                         const double* scalars_dev, const mpci_lti_tables* per_instance,
                         double* P_dev, double* q_dev, double* G_dev, double* h_dev, long batch,
                         int hessian_mode, void* stream);

/* ---- per-instance plant INSIDE a larger formulation --------------------------------
 * The biped of the reference's example (biped_formulation.py:23-191) with one plant per
 * instance: besides the plant's inputs the QP has auxiliary optimisation variables z (the
 * footstep displacements Ds of the shared `steps` system s = F s0 + E Ds, tools.py:79-101)
 * and besides x0 it has auxiliary given variables (s0, the bias n); costs and limits are
 * written on signals   f_k = cx . x_{k+1} + cu . u_k + Mz[k,:] z + Mg[k,:] given + const
 * (`(b+n-s)`, `(DCM-s)`: LineCombos with the matrix coefficient of biped_formulation.py:57,63
 * folded into Mz / Mg on the host).  Everything that involves only the plant is written by
 * the fused kernel at its place in the larger matrices (the embedded layout below); the
 * blocks that involve z or the auxiliary givens come from per-instance signals and formulation
 * constants -- those of P in a phase of the fused kernel, those of q, G, h in a second kernel
 * (csrc/mpci_couple.cu):
 *   P[u,z] = sum_c w_c T_c^T Mz_c  (and its transpose),   P[z,z] = sum_c w_c Mz_c^T Mz_c
 *   q[u]  += sum_c w_c T_c^T e_c,   q[z] = sum_c w_c Mz_c^T (fr_c + e_c)
 *   G[r,z] = Gz[r,z],   h[r] (+)= Hg[r,:] . [given; 1]
 * with T_c the Toeplitz operator of cost c's signal (rows k, columns (j,p): cx . Gam_{k-p}[:,j]
 * + cu[j] [p == k]), fr_c[k] = cx . A^{k+1} x0 and e_c = E_c [given; 1] (body.py:236-302 for
 * such a formulation).  All pointers are host pointers, copied at creation.               */
typedef struct {
  int nv_total;            /* variables of the QP (row length of P and G)                  */
  int nc_total;            /* rows of G / h                                                */
  const int* var_off;      /* [m] column of sample 0 of input j (its N samples contiguous) */
  const int* grp_row0;     /* [n_groups] first row of every group of the mpci_lti_desc     */
  int nz;                  /* auxiliary optimisation variables                             */
  const int* z_col;        /* [nz] their columns                                           */
  int given_len;           /* doubles per instance of `given_dev` (0: no auxiliary givens) */
  int n_costs;             /* costs whose signal involves z or the auxiliary givens        */
  const double* cost_cx;   /* [n_costs][n]   plant part of the signal                      */
  const double* cost_cu;   /* [n_costs][m]                                                 */
  const double* cost_w;    /* [n_costs]                                                    */
  const double* cost_MzT;  /* [n_costs][nz][N]  (Mz transposed)                            */
  const int* cost_e_ptr;   /* CSR over the rows (c, k): e_c[k] = sum val * [given; 1][col] */
  const int* cost_e_col;   /*   [n_costs * N + 1], [nnz], [nnz]                            */
  const double* cost_e_val;
  const double* Pzz;       /* [nz][nz]                                                     */
  const double* Gz;        /* [nc_total][nz]                                               */
  const int* h_ptr;        /* CSR over the rows of h: [nc_total + 1], [nnz], [nnz]         */
  const int* h_col;
  const double* h_val;
  const int* row_kind;     /* [nc_total] 0: a row of a plant group (the CSR part is ADDED  */
                           /* to h); 1: a row without plant part (h = the CSR part, the    */
                           /* plant's columns of G are zeroed)                             */
} mpci_lti_embed;

/* A plan whose outputs have the embedded layout (P [nv_total][nv_total], q [nv_total],
 * G [nc_total][nv_total], h [nc_total] per instance).  Runs through
 * mpci_lti_assemble_coupled only.                                                         */
int mpci_lti_plan_create_embedded(const mpci_lti_desc* desc, const mpci_lti_embed* embed,
                                  mpci_lti_plan** plan_out);

/* As mpci_lti_assemble_ex, plus: given_dev [batch][given_len] (the formulation's whole given
 * vector; x0_dev stays the plant's initial state in the kernel's order), and optional
 * per-instance signal tables of the coupled costs, cost_cx_dev [batch][n_costs][n] /
 * cost_cu_dev [batch][n_costs][m] (NULL = the plan's).  Two launches on `stream`: the fused
 * per-instance kernel, then the coupling kernel.                                           */
int mpci_lti_assemble_coupled(const mpci_lti_plan* plan, const double* A_dev, long strideA,
                              const double* B_dev, long strideB, const double* x0_dev,
                              const double* given_dev, const double* scalars_dev,
                              const mpci_lti_tables* per_instance, const double* cost_cx_dev,
                              const double* cost_cu_dev, double* P_dev, double* q_dev,
                              double* G_dev, double* h_dev, long batch, int hessian_mode,
                              void* stream);

/* mpci_lti_assemble_coupled with every per-instance table of the coupled costs: besides the signal
 * tables, one WEIGHT per instance and cost (cost_w [batch][n_costs], replacing the plan's cost_w --
 * the caller then also passes Pzz [batch][nz][nz] = sum_c w_c Mz_c^T Mz_c for its weights) and one
 * AIM per instance and cost (cost_aim [batch][n_costs], subtracted from every sample of e_c on top
 * of the constant column of the plan's cost_e map).  Any pointer may be NULL (= the plan's / zero). */
typedef struct {
  const double* cost_cx;   /* [batch][n_costs][n] */
  const double* cost_cu;   /* [batch][n_costs][m] */
  const double* cost_w;    /* [batch][n_costs]    */
  const double* cost_aim;  /* [batch][n_costs]    */
  const double* Pzz;       /* [batch][nz][nz]     */
  /* limits turned per instance (a Box rotated by one yaw per robot) on signals with auxiliary
   * parts: the rows' entries under the auxiliary variables and the values of the plan's h map
   * (same pattern h_ptr / h_col) per instance                                              */
  const double* Gz;        /* [batch][nc_total][nz]   */
  const double* h_val;     /* [batch][h_ptr[nc_total]] */
} mpci_lti_coupled_tables;
int mpci_lti_assemble_coupled_ex(const mpci_lti_plan* plan, const double* A_dev, long strideA,
                                 const double* B_dev, long strideB, const double* x0_dev,
                                 const double* given_dev, const double* scalars_dev,
                                 const mpci_lti_tables* per_instance,
                                 const mpci_lti_coupled_tables* coupled, double* P_dev, double* q_dev,
                                 double* G_dev, double* h_dev, long batch, int hessian_mode,
                                 void* stream);

/* ---- costs with a `schedule` on the per-instance path ---------------------------------
 * A Cost whose `schedule` picks some samples of the horizon (goal.py:46-60; the rows `sigma`
 * of body.py:273-283 -- a terminal cost is schedule = range(N - 1, N)) has no time-invariant
 * stage weight, so it is not part of the fused kernel's Hessian: this stage ADDS it to the P
 * and q that mpci_lti_assemble / _ex / _coupled wrote, on the same stream:
 *   P += sum_t w_t sum_{k in sigma_t} v_{t,k} c_{t,k}^T                         (body.py:292)
 *   q += sum_t w_t / 2 sum_k ( v_{t,k} (fc_{t,k} - caim_t) + c_{t,k} (fv_{t,k} - aim_t) )
 *                                                                               (body.py:293-300)
 * v_{t,k} [nv] = row k of the preview matrix of the term's variable over the optimisation
 * variables (plant columns (j, p <= k): cx . A^{k-p} B[:,j] + cu[j] [p == k]; auxiliary columns:
 * Mz[k,:]), fv_{t,k} = cx . A^{k+1} x0 + Mg[k,:] . given its given part; c / fc the same for the
 * crossed variable (equal to v / fv for a plain cost).  Mz / Mg hold one row per SCHEDULED
 * sample, in the order of row_k.  All pointers of the description are host pointers, copied
 * at creation.                                                                              */
typedef struct {
  int n, m, N;            /* the plant, as mpci_lti_desc                                     */
  int nv;                 /* variables of the QP: row length of P (m * N in the plain layout) */
  const int* var_off;     /* [m] column of sample 0 of input j; NULL = j * N                  */
  int nz;                 /* auxiliary optimisation variables (0: none)                       */
  const int* z_col;       /* [nz] their columns                                               */
  int given_len;          /* doubles per instance of given_dev (0: no auxiliary givens)       */
  int n_terms;
  const int* row_ptr;     /* [n_terms + 1] the samples of term t: row_k[row_ptr[t] ..]        */
  const int* row_k;       /* [row_ptr[n_terms]] horizon samples, 0 <= k < N                   */
  const double* v_Mz;     /* [row_ptr[n_terms]][nz] or NULL (= zeros)                         */
  const double* v_Mg;     /* [row_ptr[n_terms]][given_len] or NULL                            */
  const double* c_Mz;     /* the crossed variable's; NULL = the variable's own                */
  const double* c_Mg;
  /* Rows that COMBINE samples -- a Cost's / Constraint's `L` (goal.py:88-104, restrictions.py:75-90;
   * vMo = L @ Mo[sigma], body.py:279-290): row r = sum over e in ent_ptr[r] .. ent_ptr[r + 1] of
   * ent_val[e] * (row ent_k[e] of the preview matrix).  NULL = the single sample row_k[r] (then
   * row_k is required).  c_ent_*: the crossed variable's combination (`cross_L`), NULL = the same.
   * The Mz / Mg tables above hold one row per ROW (the combination already applied).           */
  const int* ent_ptr;     /* [rows + 1] or NULL */
  const int* ent_k;
  const double* ent_val;
  const int* c_ent_ptr;
  const int* c_ent_k;
  const double* c_ent_val;
  /* Limit rows (mpci_lti_sched_rows): the row of G / h every row is added to; NULL for costs.  */
  const int* row_out;     /* [rows] or NULL */
} mpci_lti_sched_desc;

typedef struct mpci_lti_sched mpci_lti_sched;

int mpci_lti_sched_create(const mpci_lti_sched_desc* desc, mpci_lti_sched** out);
int mpci_lti_sched_destroy(mpci_lti_sched* sched);

/* terms_dev: per term (w, aim, caim, v_cx [n], v_cu [m], c_cx [n], c_cu [m]) = 3 + 2n + 2m
 * doubles, [batch][n_terms][3 + 2n + 2m] (stride_terms = n_terms * (3 + 2n + 2m)) or one set
 * for the batch (stride_terms = 0).  A, B, x0 as mpci_lti_assemble; given_dev [batch][given_len]
 * or NULL.  P_dev [batch][nv][nv] / q_dev [batch][nv] may be NULL.
 * One CTA per instance; read-modify-write of P (the one pass over P this stage costs).        */
int mpci_lti_sched_add(const mpci_lti_sched* sched, const double* A_dev, long strideA,
                       const double* B_dev, long strideB, const double* x0_dev,
                       const double* given_dev, const double* terms_dev, long stride_terms,
                       double* P_dev, double* q_dev, long batch, void* stream);

/* Limits with `L` on the per-instance path (restrictions.py:175-179: matrices() = arrow * L;
 * body.py:253-255: cMo += c @ Mo[sigma]): the fused kernel's row generator writes single samples,
 * so these rows are left to it as rows WITHOUT a plant part (row_kind 1 of the embedded layout: the
 * coupling kernel writes their auxiliary entries, zeros under the plant's inputs and h = the bound
 * and given parts) and this call ADDS the plant part afterwards, on the same stream:
 *   G[row_out[r]][plant columns] += w_t v_r      h[row_out[r]] -= w_t (cx_t . sum_e val_e A^{k_e + 1} x0)
 * for a description created with row_out (terms: w = 1 unless the arrow is folded into it; aims
 * unused).  nc_total = rows of G / h per instance.                                              */
int mpci_lti_sched_rows(const mpci_lti_sched* sched, const double* A_dev, long strideA,
                        const double* B_dev, long strideB, const double* x0_dev,
                        const double* terms_dev, long stride_terms, double* G_dev, double* h_dev,
                        long nc_total, long batch, void* stream);

/* Closed forms of the reference's named LIP systems (tools.get_system_matrices,
 * python/mpc_interface/tools.py:240-285: A = exp(G tau), B = int_0^tau exp(G t) dt h, there
 * by SymPy) for a batch with one sampling period and one pendulum frequency per instance:
 * system 0 "J->CCC", 1 "P->CC", 2 "P->X", 3 "dP->CCC", 4 "dP->CCP".
 * tau_dev / omega_dev [batch] (stride 0 = one value for the batch; omega may be NULL for
 * system 0); A_dev [batch][n][n], B_dev [batch][n][1], n = 3, 2, 1, 3, 3.                  */
int mpci_named_system(int system, const double* tau_dev, long stride_tau, const double* omega_dev,
                      long stride_omega, double* A_dev, double* B_dev, long batch, void* stream);

/* Same with HOST buffers (pinned memory gives full PCIe speed): inputs go to the device
 * and results come back in chunks of `chunk` instances on two streams.  The plan owns the
 * device staging buffers (allocated on first use, freed by mpci_lti_plan_destroy).      */
int mpci_lti_assemble_host(mpci_lti_plan* plan, const double* A_host, long strideA,
                           const double* B_host, long strideB, const double* x0_host,
                           const double* scalars_host, double* P_host, double* q_host,
                           double* G_host, double* h_host, long batch, int hessian_mode,
                           long chunk);

/* Streaming form (host rings of `ring` chunks, see mpci_assemble_host_stream).           */
int mpci_lti_assemble_host_stream(mpci_lti_plan* plan, const double* A_host, long strideA,
                                  const double* B_host, long strideB, const double* x0_host,
                                  const double* scalars_host, double* P_ring, double* q_ring,
                                  double* G_ring, double* h_ring, long batch, int hessian_mode,
                                  long chunk, long ring, mpci_chunk_fn on_chunk, void* user);

/* Flops the Hessian phase of one instance executes in `hessian_mode`
 * (for the FP64 roofline; the structural count is SURVEY.md 8(d)).            */
double mpci_lti_hessian_flops(const mpci_lti_plan* plan, int hessian_mode);

/* How the plan maps onto an SM in `hessian_mode` (for benchmarks and ncu cross-checks):
 * info[0] = dynamic shared memory per CTA (bytes), info[1] = resident CTAs per SM,
 * info[2] = 1 when the recursion uses the DMMA patch scan, info[3] / info[4] = row
 * lengths (doubles) of the generator arrays gz / wz in shared memory, info[5] = threads
 * per CTA (256, or 384 when the shared memory allows only one CTA per SM anyway).       */
int mpci_lti_plan_info(const mpci_lti_plan* plan, int hessian_mode, long info[6]);

#ifdef __cplusplus
}
#endif
#endif /* MPCI_H */
