// Shared helpers of libmpci: error reporting across the C ABI.
#pragma once
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>

#include "mpci.h"

namespace mpci {

char* error_buffer();  // thread-local, 512 bytes (mpci_api.cu)

inline int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(error_buffer(), 512, fmt, ap);
  va_end(ap);
  return code;
}

inline int cuda_fail(cudaError_t e, const char* where) {
  snprintf(error_buffer(), 512, "%s: %s", where, cudaGetErrorString(e));
  return static_cast<int>(e);
}

#define MPCI_CUDA(call)                                              \
  do {                                                               \
    cudaError_t e_ = (call);                                         \
    if (e_ != cudaSuccess) return ::mpci::cuda_fail(e_, #call);      \
  } while (0)

#define MPCI_LAUNCH_CHECK(name)                                      \
  do {                                                               \
    cudaError_t e_ = cudaGetLastError();                             \
    if (e_ != cudaSuccess) return ::mpci::cuda_fail(e_, name);       \
  } while (0)

inline cudaStream_t as_stream(void* s) { return static_cast<cudaStream_t>(s); }

}  // namespace mpci
