// Measured ceilings of one B200 for the roofline denominators of this repo (one JSON line):
//   hbm_fill_gbs   write-only stream: 8 GiB of 16-byte streaming stores, one pass, no reads --
//                  what the P / G streams are up against; the best of a grid-stride fill, a
//                  fill that walks 32 rows per thread (the stores of stream_terms_kernel
//                  without its arithmetic) and cudaMemsetAsync
//   hbm_copy_gbs   read + write bytes of a double2 copy of 4 GiB -> 4 GiB (the quantity
//                  MEASURED_PEAKS.json reports from a torch copy)
//   fp64_dmma_tflops / fp64_dfma_tflops   register-resident DMMA.8x8x4 / DFMA loops
//   d2h_pinned_gbs device -> pinned host, one cudaMemcpyAsync stream of 1 GiB pieces
//                  (the ceiling of the host-buffer entry points); --d2h-only N repeats
//                  only this part (run it on all GPUs of a box at once for the N-GPU ceiling)
// best of 10 after one warm-up, CUDA events.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a scripts/peaks.cu -o scripts/peaks.bin
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { \
  fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)

__global__ void fill_kernel(double2* __restrict__ out, long n_vec, double v) {
  const long stride = (long)gridDim.x * blockDim.x;
  const double2 val = make_double2(v, -v);
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n_vec; i += stride) __stcs(out + i, val);
}

// one 16-byte store per thread, consecutive threads on consecutive vectors; ROWS rows of
// `pitch` vectors per thread (the access pattern of stream_terms_kernel: a thread walks
// down a column of a [rows][pitch] array, a warp covers 512 contiguous bytes per row)
template <int ROWS>
__global__ void fill_rows_kernel(double2* __restrict__ out, long pitch, double v) {
  const long col = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (col >= pitch) return;
  double2* dst = out + (long)blockIdx.y * ROWS * pitch + col;
  const double2 val = make_double2(v, -v);
#pragma unroll 4
  for (int r = 0; r < ROWS; ++r) { __stcs(dst, val); dst += pitch; }
}

__global__ void copy_kernel(const double2* __restrict__ in, double2* __restrict__ out, long n_vec) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n_vec; i += stride)
    __stcs(out + i, __ldcs(in + i));
}

__global__ void dfma_loop(double* out, int iters) {
  double a[8], x = 1.0 + threadIdx.x * 1e-9, y = 1e-9;
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = fma(a[i], x, y);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void dmma_loop(double* out, int iters) {
  double c[8][2], a = 1.0 + threadIdx.x * 1e-9, b = 1e-9 * threadIdx.x;
#pragma unroll
  for (int i = 0; i < 8; ++i) c[i][0] = c[i][1] = 0.0;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                   : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main(int argc, char** argv) {
  int d2h_only = 0;
  if (argc > 2 && !strcmp(argv[1], "--d2h-only")) d2h_only = atoi(argv[2]);
  int sms = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  double fill = 0, fill_rows = 0, fill_memset = 0, copy = 0, dmma = 0, dfma = 0, d2h = 0;

  if (!d2h_only) {
    const long bytes = 8L << 30, n_vec = bytes / 16;
    double2* buf;
    CK(cudaMalloc(&buf, bytes));
    for (int blocks_per_sm : {4, 8, 16})
      for (int rep = 0; rep < 11; ++rep) {
        CK(cudaEventRecord(e0));
        fill_kernel<<<sms * blocks_per_sm, 256>>>(buf, n_vec, 1.0 + rep);
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (rep && bytes / ms / 1e6 > fill) fill = bytes / ms / 1e6;
      }
    // the same bytes as [rows][pitch] with 32 rows per thread (pitch = 1 MiB of vectors)
    for (int rep = 0; rep < 11; ++rep) {
      const long pitch = 65536;
      dim3 grid((unsigned)(pitch / 256), (unsigned)(n_vec / pitch / 32));
      CK(cudaEventRecord(e0));
      fill_rows_kernel<32><<<grid, 256>>>(buf, pitch, 1.0 + rep);
      CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
      float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
      if (rep && bytes / ms / 1e6 > fill_rows) fill_rows = bytes / ms / 1e6;
    }
    for (int rep = 0; rep < 11; ++rep) {
      CK(cudaEventRecord(e0));
      CK(cudaMemsetAsync(buf, rep, bytes, 0));
      CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
      float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
      if (rep && bytes / ms / 1e6 > fill_memset) fill_memset = bytes / ms / 1e6;
    }
    for (int blocks_per_sm : {4, 8, 16})
      for (int rep = 0; rep < 11; ++rep) {
        CK(cudaEventRecord(e0));
        copy_kernel<<<sms * blocks_per_sm, 256>>>(buf, buf + n_vec / 2, n_vec / 2);
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (rep && bytes / ms / 1e6 > copy) copy = bytes / ms / 1e6;  // read 4 GiB + write 4 GiB
      }
    CK(cudaFree(buf));
    double* out;
    CK(cudaMalloc(&out, sizeof(double) * sms * 8 * 256));
    const int iters = 20000, threads = 256, per_sm = 4;
    for (int which = 0; which < 2; ++which)
      for (int rep = 0; rep < 6; ++rep) {
        CK(cudaEventRecord(e0));
        if (which == 0) dfma_loop<<<sms * per_sm, threads>>>(out, iters);
        else dmma_loop<<<sms * per_sm, threads>>>(out, iters);
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        const double flops = which == 0 ? 2.0 * 8 * iters * threads * per_sm * sms
                                        : 512.0 * 8 * iters * (threads / 32) * per_sm * sms;
        double& best = which ? dmma : dfma;
        if (rep && flops / ms / 1e9 > best) best = flops / ms / 1e9;
      }
    CK(cudaFree(out));
  }
  {
    const long piece = 1L << 30;
    char *dev, *host;
    CK(cudaMalloc(&dev, piece));
    CK(cudaMemset(dev, 1, piece));
    CK(cudaHostAlloc(&host, piece, cudaHostAllocDefault));
    const int reps = d2h_only ? d2h_only : 10;
    double sum = 0;
    for (int rep = 0; rep <= reps; ++rep) {
      CK(cudaEventRecord(e0));
      CK(cudaMemcpyAsync(host, dev, piece, cudaMemcpyDeviceToHost, 0));
      CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
      float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
      if (rep) { sum += piece / ms / 1e6; if (piece / ms / 1e6 > d2h) d2h = piece / ms / 1e6; }
    }
    if (d2h_only) d2h = sum / reps;  // sustained mean when several GPUs copy at once
    CK(cudaFreeHost(host)); CK(cudaFree(dev));
  }
  const double fill_best = fill > fill_rows ? (fill > fill_memset ? fill : fill_memset)
                                            : (fill_rows > fill_memset ? fill_rows : fill_memset);
  printf("{\"sms\": %d, \"hbm_fill_gbs\": %.1f, \"fill_grid_stride_gbs\": %.1f, \"fill_rows_gbs\": %.1f, "
         "\"fill_memset_gbs\": %.1f, \"hbm_copy_gbs\": %.1f, \"fp64_dmma_tflops\": %.2f, "
         "\"fp64_dfma_tflops\": %.2f, \"d2h_pinned_gbs\": %.2f, \"d2h_only\": %d}\n",
         sms, fill_best, fill, fill_rows, fill_memset, copy, dmma, dfma, d2h, d2h_only);
  return 0;
}
