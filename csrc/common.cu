#include "common.cuh"

#include "../include/b200_ops.h"

namespace b200 {

thread_local char g_last_error[512] = {0};
std::atomic<long long> g_launch_count{0};

PFN_encodeTiled get_encode_tiled() {
  static PFN_encodeTiled fn = nullptr;
  if (fn) return fn;
  void* p = nullptr;
  cudaDriverEntryPointQueryResult qres;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres);
  if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !p) return nullptr;
  fn = reinterpret_cast<PFN_encodeTiled>(p);
  return fn;
}

int make_tmap_bf16(CUtensorMap* m, const void* base, int rank, const unsigned long long* dims,
                   const unsigned long long* strides_bytes, const unsigned int* box) {
  PFN_encodeTiled enc = get_encode_tiled();
  if (!enc) return set_error(-4, "cuTensorMapEncodeTiled entry point unavailable (no CUDA driver?)");
  cuuint64_t gdim[5];
  cuuint64_t gstr[5];
  cuuint32_t bx[5];
  cuuint32_t es[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    bx[i] = box[i];
    es[i] = 1;
    if (i + 1 < rank) gstr[i] = strides_bytes[i];
  }
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, const_cast<void*>(base), gdim, gstr, bx, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    return set_error(-5,
                     "cuTensorMapEncodeTiled failed (%d): rank %d dims [%llu,%llu,%llu,%llu] strides "
                     "[%llu,%llu,%llu] box [%u,%u,%u,%u] base %p",
                     (int)r, rank, dims[0], rank > 1 ? dims[1] : 0ull, rank > 2 ? dims[2] : 0ull,
                     rank > 3 ? dims[3] : 0ull, strides_bytes[0], rank > 2 ? strides_bytes[1] : 0ull,
                     rank > 3 ? strides_bytes[2] : 0ull, box[0], rank > 1 ? box[1] : 0u, rank > 2 ? box[2] : 0u,
                     rank > 3 ? box[3] : 0u, base);
  }
  return 0;
}

int num_sms() {
  static int n = 0;
  if (n) return n;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
  return n;
}

bool pdl_enabled() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("B200_PDL");
    on = (e && e[0] == '0') ? 0 : 1;
  }
  return on == 1;
}

bool deterministic() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("B200_DETERMINISTIC");
    on = (e && e[0] == '1') ? 1 : 0;
  }
  return on == 1;
}

static unsigned long long device_bit() {
  int dev = 0;
  cudaGetDevice(&dev);
  return 1ull << (dev & 63);
}
bool device_needs_setup(const std::atomic<unsigned long long>& mask) { return !(mask.load() & device_bit()); }
void device_setup_done(std::atomic<unsigned long long>& mask) { mask.fetch_or(device_bit()); }

}  // namespace b200

extern "C" {
const char* b200_last_error(void) { return b200::g_last_error; }
int b200_ops_abi_version(void) { return 2; }
int b200_deterministic(void) { return b200::deterministic() ? 1 : 0; }
long long b200_launch_count(void) { return b200::g_launch_count.load(); }
void b200_reset_launch_count(void) { b200::g_launch_count.store(0); }
/* Kernels executed by replaying a captured CUDA graph never pass through the launch wrappers: the owner of the
 * graph reports them here (count taken while the graph was captured). */
void b200_add_launch_count(long long n) { b200::g_launch_count.fetch_add(n); }
}
