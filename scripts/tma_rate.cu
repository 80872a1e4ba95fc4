// Issue rate of small shared -> global bulk copies (cp.async.bulk) on one SM, as a function
// of the copy size and of how many warps / lanes issue them.  Backs the choice of issuing
// lanes in the constraint phase of mpci_lti.cu.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tma_rate scripts/tma_rate.cu && ./tma_rate
#include <cstdio>
#include <cuda_runtime.h>

__global__ void issue(double* out, int bytes, int warps, int lanes, int per_thread, long cta_stride) {
  extern __shared__ double buf[];
  for (int i = threadIdx.x; i < 8192; i += blockDim.x) buf[i] = i;
  asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp < warps && lane < lanes) {
    const unsigned src = (unsigned)__cvta_generic_to_shared(buf);
    double* dst = out + blockIdx.x * cta_stride + (size_t)(warp * lanes + lane) * per_thread * (bytes / 8);
    for (int r = 0; r < per_thread; ++r) {
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;\n"
                   :: "l"(dst + (size_t)r * (bytes / 8)), "r"(src + (r & 7) * 16), "r"(bytes) : "memory");
      if ((r & 63) == 63) {
        asm volatile("cp.async.bulk.commit_group;\n" ::: "memory");
        asm volatile("cp.async.bulk.wait_group.read 1;\n" ::: "memory");
      }
    }
    asm volatile("cp.async.bulk.commit_group;\n" ::: "memory");
    asm volatile("cp.async.bulk.wait_group 0;\n" ::: "memory");
  }
}

int main() {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  cudaFuncSetAttribute(issue, cudaFuncAttributeMaxDynamicSharedMemorySize, 72 * 1024);
  double* out;
  const size_t total = (size_t)4 << 30;  // 4 GB
  cudaMalloc(&out, total);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  printf("bytes ctas/SM warps lanes  copies/CTA   ns/copy/SM  cycles@1.9GHz  GB/s(all SMs)\n");
  const int sizes[] = {400, 800, 1600, 3200};
  for (int per_sm = 1; per_sm <= 3; per_sm += 2)
    for (int bytes : sizes)
      for (int warps : {1, 8})
        for (int lanes : {1, 8, 32}) {
          const int grid = sms * per_sm;
          const size_t cta_bytes = total / grid / 16 * 16;
          long copies = (long)(cta_bytes / bytes);
          if (copies > 16384) copies = 16384;
          const int per_thread = (int)(copies / (warps * lanes));
          if (per_thread < 8) continue;
          copies = (long)per_thread * warps * lanes;
          float best = 1e30f;
          for (int rep = 0; rep < 3; ++rep) {
            cudaEventRecord(e0);
            issue<<<grid, 256, 66 * 1024>>>(out, bytes, warps, lanes, per_thread, (long)(cta_bytes / 8));
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            if (ms < best) best = ms;
          }
          const double ns = best * 1e6 / (copies * per_sm);
          printf("%5d %6d %6d %5d %10ld %11.1f %11.1f %12.0f\n", bytes, per_sm, warps, lanes, copies, ns,
                 ns * 1.9, (double)copies * grid * bytes / best / 1e6);
        }
  if (cudaDeviceSynchronize() != cudaSuccess) { printf("CUDA error\n"); return 1; }
  return 0;
}
