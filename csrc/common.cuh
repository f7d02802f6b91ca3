// Shared host-side helpers for libb200ops: error reporting, launch counting, driver entry points.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>

#include <atomic>
#include <utility>
#include <cstdlib>

namespace b200 {

extern thread_local char g_last_error[512];
extern std::atomic<long long> g_launch_count;

inline int set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_last_error, sizeof(g_last_error), fmt, ap);
  va_end(ap);
  return code;
}

inline int check_launch(const char* what) {
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_error(-2, "%s: launch failed: %s", what, cudaGetErrorString(e));
  return 0;
}

#define B200_CUDA_OK(expr)                                                                       \
  do {                                                                                           \
    cudaError_t _e = (expr);                                                                     \
    if (_e != cudaSuccess) return b200::set_error(-3, "%s: %s", #expr, cudaGetErrorString(_e)); \
  } while (0)

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
PFN_encodeTiled get_encode_tiled();

// bf16 tensor map with 128B swizzle and zero OOB fill.  dims/strides innermost first;
// strides[i] is the byte stride of dim i+1 (dim 0 is contiguous).
int make_tmap_bf16(CUtensorMap* m, const void* base, int rank, const unsigned long long* dims,
                   const unsigned long long* strides_bytes, const unsigned int* box);

int num_sms();

// Per-DEVICE one-time set-up (function attributes such as the dynamic shared-memory limit live in the
// device's context, and one process may drive several GPUs: engine/multi_tower.py).  Usage:
//   static std::atomic<unsigned long long> done{0};
//   if (device_needs_setup(done)) { ...cudaFuncSetAttribute...; device_setup_done(done); }
// Launch with the programmatic-dependent-launch attribute (see pdl_wait in ptx_sm100.cuh); B200_PDL=0 turns it off.
bool pdl_enabled();
// B200_DETERMINISTIC=1: every floating-point accumulation that is normally split over CTAs and merged with atomics
// (weight gradients, bias column sums, the stem filter gradient, sum of squares, loss sums) runs in ONE CTA per output
// element instead, so two runs from the same state are bit-identical.  Slow; meant for tests and debugging.
bool deterministic();
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

// same with a thread-block cluster of `cluster_x` CTAs along x (grid.x must be a multiple of it)
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl_cluster(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                      unsigned cluster_x, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  attr[1].id = cudaLaunchAttributeClusterDimension;
  attr[1].val.clusterDim.x = cluster_x;
  attr[1].val.clusterDim.y = 1;
  attr[1].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 2;
  return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

bool device_needs_setup(const std::atomic<unsigned long long>& mask);
void device_setup_done(std::atomic<unsigned long long>& mask);

}  // namespace b200
