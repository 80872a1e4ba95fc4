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

// [q; h]: sparse rows, one instance per thread (mpci_assemble.cu)
constexpr int kQhWarps = 16;  // warps of a block; each owns a list of row chunks
constexpr int kQhChunk = 4;   // rows per chunk (one 32-byte run per instance)
// The chunks are dealt to kQhWarps * kQhSplit lists; warp w of a block works through the
// lists w, w + kQhWarps, ...  A batch too small to give every SM a block is launched with
// up to kQhSplit blocks per group of instances (gridDim.y), each taking a share of the lists
// of every warp: one instance then runs on 8 x 16 warps instead of 16.
constexpr int kQhSplit = 8;
constexpr int kQhLists = kQhWarps * kQhSplit;

struct QhNz {      // 16 bytes; a warp's records are consecutive, in processing order
  double val;
  int col;         // column of [given; 1] times kQhColStride (index into the staged double2s)
  int meta;        // term | kQhEndSeg | kQhEndRow | kQhEndChunk
};
constexpr int kQhColStride = 33;  // double2 per staged column: 32 lanes x 2 instances, padded
constexpr int kQhTermMask = 0xff;
constexpr int kQhEndSeg = 0x100;    // last non-zero of a (row, term) segment
constexpr int kQhEndRow = 0x200;    // ... of a row: the accumulator is the result
constexpr int kQhEndChunk = 0x400;  // ... of a chunk of rows: write the chunk out

// Columns of [given; 1] are staged in shared memory for the instances of a block; when
// they do not all fit (long horizons: the biped has ng = 8 + 2N; previews stage
// [given; optim; 1]) the columns are cut into n_pass blocks and the kernel makes one pass
// per block: pass 0 writes the rows, the later ones add to them.  A chunk of rows belongs
// to the same list in every pass, so the same thread re-reads what it wrote.
constexpr int kQhMaxPass = 16;
struct QhArgs {
  int nv, ng, nc, n_scalars, n_terms;
  int n_pass;              // column blocks (1 when everything fits)
  int max_cols;            // widest column block (sizes the shared memory)
  int pass_col[kQhMaxPass + 1];  // block p = columns pass_col[p] .. pass_col[p+1]-1 of [given; 1]
  const int* factors;      // dev [n_terms][MPCI_MAX_FACTORS]
  const QhNz* nz;          // dev: the lists' record streams, concatenated (pass-major)
  const int* warp_nz_ptr;  // dev [n_pass][kQhLists + 1]
  const int2* chunks;      // dev: (first row, rows) in the order the lists meet them
  const int* warp_chunk_ptr;  // dev [n_pass][kQhLists + 1]
};

int launch_stream(const StreamArgs& a, const double* scalars, int n_scalars, double* out,
                  long batch, cudaStream_t st);
int launch_qh(const QhArgs& a, const double* scalars, const double* given, double* q, double* h,
              long batch, cudaStream_t st);

}  // namespace mpci

// A plan is used by ONE host thread at a time (mpci_assemble forks a side stream and
// records timers through the handle; the host-buffer entry points own staging slots).
struct mpci_plan {
  int nv = 0, ng = 0, nc = 0, n_scalars = 0;
  mpci::StreamArgs P{}, G{};
  mpci::StreamArgs Pc{}, Gc{};  // compact constant sets (CSC values ...), mpci_plan_set_compact
  mpci::QhArgs Y{};
  void* dev_tables = nullptr;  // one allocation: nz | chunks | factors | warp pointers
  // staging of mpci_assemble_host (two slots), allocated on first use
  long slot_capacity = 0;       // instances
  long slot_out_doubles = 0;    // doubles per output slot
  double* slot_in[2] = {nullptr, nullptr};   // scalars | given
  double* slot_out[2] = {nullptr, nullptr};  // P | q | G | h
  cudaStream_t streams[2] = {nullptr, nullptr};
  cudaEvent_t slot_done[2] = {nullptr, nullptr};  // the slot's results are in host memory
  // mpci_assemble runs the (shared-memory bound) [q;h] kernel beside the
  // (HBM bound) P / G streams: a side stream forked from the caller's stream
  mutable cudaStream_t side = nullptr;
  mutable cudaEvent_t fork = nullptr, join = nullptr;
  // mpci_plan_set_timing: event pairs around the three launches of every call
  // (P stream, G stream, [q;h]) so that a bench reads kernel durations measured
  // INSIDE its timed region; six events per call and piece, in call order
  mutable bool timing = false;
  mutable std::vector<cudaEvent_t> timers;
};
