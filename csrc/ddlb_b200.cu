// libddlb_b200.so — unity build of the native layer (one translation unit so the
// device-side error word and the inline-PTX helpers exist once).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -shared -Xcompiler -fPIC \
//        -Iinclude csrc/ddlb_b200.cu -o ddlb_b200/_native/libddlb_b200.so
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include "ddlb_b200.h"

namespace ddlb {

static thread_local char g_last_error[1024] = "";

static int set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_last_error, sizeof(g_last_error), fmt, ap);
  va_end(ap);
  return 1;
}

#define DDLB_CUDA(expr)                                                                   \
  do {                                                                                    \
    cudaError_t _e = (expr);                                                              \
    if (_e != cudaSuccess) {                                                              \
      (void)cudaGetLastError();                                                           \
      return ::ddlb::set_error("%s failed: %s", #expr, cudaGetErrorString(_e));           \
    }                                                                                     \
  } while (0)

}  // namespace ddlb

extern "C" const char* ddlb_last_error(void) { return ddlb::g_last_error; }
extern "C" int ddlb_version(void) { return 200; }   // 0.2.0: config gained pdl, wide_n; info has 12 slots

#include "tp_gemm_sm100.cuh"
#include "tp_gemm_host.inl"
#include "runtime.inl"
#include "vmm_heap.inl"
