// FP64 peak of the box: register-resident DFMA and DMMA.8x8x4 loops, all SMs.
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a scripts/fp64_peak.cu -o /tmp/fp64_peak
#include <cstdio>
#include <cuda_runtime.h>

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

int main() {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  double* out;
  cudaMalloc(&out, sizeof(double) * sms * 8 * 256);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int threads : {128, 256}) for (int per_sm : {2, 4, 8}) {
    const int iters = 20000;
    for (int which = 0; which < 2; ++which) {
      float best = 1e30f;
      for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0);
        if (which == 0) dfma_loop<<<sms * per_sm, threads>>>(out, iters);
        else dmma_loop<<<sms * per_sm, threads>>>(out, iters);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (rep && ms < best) best = ms;
      }
      const double flops = which == 0 ? 2.0 * 8 * iters * threads * per_sm * sms
                                      : 512.0 * 8 * iters * (threads / 32) * per_sm * sms;
      printf("%s threads=%d blocks/SM=%d  %.3f ms  %.2f TFLOP/s\n", which ? "DMMA.8x8x4" : "DFMA      ",
             threads, per_sm, best, flops / best / 1e9);
    }
  }
  return cudaDeviceSynchronize() != cudaSuccess;
}
