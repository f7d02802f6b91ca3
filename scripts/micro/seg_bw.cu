// Micro-benchmark behind DESIGN.md 15: how much does the per-row segment size of a warp's global access pattern
// matter?  A [rows x 512 B] bf16 tensor (256 channels) is written (or read) by warps that each own a 32-row x SEG-byte
// block and move it with 16-byte accesses per lane: SEG = 32 -> 16 rows x 32 B per instruction, 64 -> 8 rows x 64 B
// (the conv epilogue today), 128 -> 4 rows x 128 B, 512 -> 1 row x 512 B (fully contiguous).
#include <cuda_runtime.h>
#include <stdint.h>
template <int SEG, bool WRITE>
__global__ void seg_kernel(uint4* __restrict__ buf, long long rows, uint4* __restrict__ sink) {
  constexpr int kLanesPerRow = SEG / 16;           // lanes covering one row segment
  constexpr int kRowsPerInstr = 32 / kLanesPerRow;
  constexpr int kInstr = 32 / kRowsPerInstr;       // instructions for 32 rows
  constexpr int kSegsPerRow = 512 / SEG;
  const int lane = threadIdx.x & 31;
  const long long warp = (blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x) >> 5;
  const long long nwarps = (gridDim.x * static_cast<long long>(blockDim.x)) >> 5;
  const long long blocks = (rows / 32) * kSegsPerRow;
  uint4 acc = make_uint4(0, 0, 0, 0);
  for (long long b = warp; b < blocks; b += nwarps) {
    const long long rb = b / kSegsPerRow;
    const int seg = static_cast<int>(b % kSegsPerRow);
#pragma unroll
    for (int i = 0; i < kInstr; ++i) {
      const long long row = rb * 32 + i * kRowsPerInstr + lane / kLanesPerRow;
      uint4* p = buf + row * 32 + seg * kLanesPerRow + (lane % kLanesPerRow);
      if (WRITE) *p = make_uint4(lane, i, seg, 7);
      else { uint4 v = *p; acc.x ^= v.x; acc.y ^= v.y; acc.z ^= v.z; acc.w ^= v.w; }
    }
  }
  if (!WRITE && acc.x == 0x12345678u) *sink = acc;
}
extern "C" int seg_run(int seg, int write, void* buf, long long rows, void* sink, int blocks, void* stream) {
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  uint4* b = static_cast<uint4*>(buf); uint4* k = static_cast<uint4*>(sink);
#define RUN(S) if (seg == S) { if (write) seg_kernel<S, true><<<blocks, 512, 0, s>>>(b, rows, k); else seg_kernel<S, false><<<blocks, 512, 0, s>>>(b, rows, k); }
  RUN(32) RUN(64) RUN(128) RUN(512)
  return static_cast<int>(cudaGetLastError());
}
