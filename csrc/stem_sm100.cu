// RetinaNet/ResNet stem: 7x7 stride-2 convolution, Cin = 3 (stored as 4), Cout = 64, fused with the
// frozen-BN shift and ReLU, on tcgen05 tensor cores.
//
// Cin = 3 cannot feed a 64-wide K tile through TMA, so this kernel builds the implicit-GEMM A tile
// itself: the input is NHWC with 4 channels (8 bytes per pixel), one filter row of an output pixel
// is 7 pixels = 28 contiguous bf16, and K = 7 rows x 28 = 196, zero-padded to 256 (four 64-wide K
// blocks).  128 producer threads (one per output pixel of the 8x16 tile) copy 49 eight-byte pieces
// each into a SWIZZLE_128B K-major tile in shared memory, fence them to the async proxy, and the MMA
// warp issues 16 tcgen05.mma (M=128, N=64, K=16) per tile with the packed weights resident in smem.
// Two smem stages and two TMEM accumulators overlap gather, MMA and the epilogue (shift + ReLU,
// one 128-byte line per pixel).
//
// Replaces cuDNN's conv1 forward; the reference has no CUDA source.
#include "../include/b200_ops.h"
#include "common.cuh"
#include "ptx_sm100.cuh"

namespace b200 {

constexpr int kStemKB = 4;                       // K blocks of 64
constexpr int kStemABytes = kStemKB * 128 * 128; // 64 KB per stage
constexpr int kStemBBytes = kStemKB * 64 * 128;  // 32 KB, resident
constexpr int kStemStages = 3;
constexpr int kStemThreads = 512;                // w0 MMA, w1 TMEM alloc, w4-7 epilogue, w8-15 producers
constexpr int kStemSmem = kStemStages * kStemABytes + kStemBBytes + 1024 + 128;

struct StemParams {
  const __nv_bfloat16* x;   // [N, H, W, 4]
  const __nv_bfloat16* w;   // [64][256] packed, k = r*28 + s*4 + c, BN scale folded in
  const float* shift;       // [64]
  __nv_bfloat16* y;         // [N, OH, OW, 64]
  int N, H, W, OH, OW;
  int tiles_h, tiles_w, total_tiles;
  int relu;                 // 0: raw conv + shift (a trainable batch norm follows: DenseNet)
};

__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__global__ void __launch_bounds__(kStemThreads, 1) stem_conv_kernel(const StemParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t a_smem = smem_base;
  const uint32_t b_smem = smem_base + kStemStages * kStemABytes;
  const uint32_t bar_base = b_smem + kStemBBytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kStemStages + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * kStemStages + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * kStemStages + 2 + a); };
  const uint32_t tmem_ptr_smem = bar_base + 96;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // zero the A stages once (K padding 196..255 is never written again and must not hold NaNs),
  // and copy the packed weights into the swizzled resident B tile
  for (int i = threadIdx.x; i < kStemStages * kStemABytes / 16; i += kStemThreads)
    reinterpret_cast<uint4*>(smem_gen)[i] = make_uint4(0u, 0u, 0u, 0u);
  for (int i = threadIdx.x; i < 64 * 32; i += kStemThreads) {  // 64 rows x 32 sixteen-byte pieces
    const int row = i >> 5, piece = i & 31;
    const int kb = piece >> 3, ch = piece & 7;
    const uint4 v = *reinterpret_cast<const uint4*>(p.w + row * 256 + piece * 8);
    *reinterpret_cast<uint4*>(smem_gen + kStemStages * kStemABytes + kb * 8192 + row * 128 + ((ch ^ (row & 7)) << 4)) = v;
  }
  if (warp == 0 && lane == 0) {
    for (int s = 0; s < kStemStages; ++s) {
      mbar_init(full_bar(s), 256);
      mbar_init(empty_bar(s), 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tfull_bar(s), 1);
      mbar_init(tempty_bar(s), 128);
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_ptr_smem, 128);
    tmem_relinquish();
  }
  fence_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_ptr_smem));

  auto decode = [&](int tile, int& n, int& y0, int& x0) {
    const int per_img = p.tiles_h * p.tiles_w;
    n = tile / per_img;
    const int rem = tile - n * per_img;
    y0 = (rem / p.tiles_w) * 8;
    x0 = (rem % p.tiles_w) * 16;
  };

  if (warp >= 8) {
    // ---------------------------------------------------------------- producers: gather the A tile
    const int t = (warp - 8) * 32 + lane;          // 256 gather threads: (pixel m, half of the filter rows)
    const int m = t & 127, part = t >> 7;
    const int r_lo = part ? 4 : 0, r_hi = part ? 7 : 4;
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
      int n, y0, x0;
      decode(tile, n, y0, x0);
      const int oy = y0 + (m >> 4), ox = x0 + (m & 15);
      mbar_wait(empty_bar(stage), phase ^ 1);
      uint8_t* a_row = smem_gen + stage * kStemABytes + m * 128;
      const int sw = m & 7;
      // all loads of this thread's rows first (28 in flight), then the swizzled stores
      uint2 v[4][7];
#pragma unroll
      for (int rr = 0; rr < 4; ++rr) {
        const int r = r_lo + rr;
        const int iy = 2 * oy + r - 3;
        const bool row_ok = r < r_hi && iy >= 0 && iy < p.H;
        const __nv_bfloat16* src = p.x + (static_cast<long long>(n) * p.H + (row_ok ? iy : 0)) * p.W * 4;
#pragma unroll
        for (int s = 0; s < 7; ++s) {
          const int ix = 2 * ox + s - 3;
          v[rr][s] = make_uint2(0u, 0u);
          if (row_ok && ix >= 0 && ix < p.W) v[rr][s] = __ldg(reinterpret_cast<const uint2*>(src + static_cast<long long>(ix) * 4));
        }
      }
#pragma unroll
      for (int rr = 0; rr < 4; ++rr) {
        const int r = r_lo + rr;
        if (r < r_hi) {
#pragma unroll
          for (int s = 0; s < 7; ++s) {
            const int kbyte = (r * 28 + s * 4) * 2;          // byte offset of this piece inside the K axis
            const int kb = kbyte >> 7, within = kbyte & 127;
            *reinterpret_cast<uint2*>(a_row + kb * 16384 + (((within >> 4) ^ sw) << 4) + (within & 15)) = v[rr][s];
          }
        }
      }
      fence_async_smem();   // generic-proxy stores -> visible to the tensor core (async proxy)
      mbar_arrive(full_bar(stage));
      if (++stage == kStemStages) {
        stage = 0;
        phase ^= 1;
      }
    }
  } else if (warp == 0) {
    // ---------------------------------------------------------------- MMA issuer
    const uint32_t idesc = make_idesc_bf16(128, 64, 0, 0);
    int stage = 0, acc = 0;
    uint32_t phase = 0, acc_phase = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
      mbar_wait(tempty_bar(acc), acc_phase ^ 1);
      mbar_wait(full_bar(stage), phase);
      tc_fence_after();
      if (elect_one()) {
#pragma unroll
        for (int kb = 0; kb < kStemKB; ++kb) {
          const uint64_t adesc = make_smem_desc(a_smem + stage * kStemABytes + kb * 16384, 16, 1024);
          const uint64_t bdesc = make_smem_desc(b_smem + kb * 8192, 16, 1024);
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16(tmem_base + acc * 64, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
        }
        umma_commit(empty_bar(stage));
        umma_commit(tfull_bar(acc));
      }
      __syncwarp();
      if (++stage == kStemStages) {
        stage = 0;
        phase ^= 1;
      }
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
  } else if (warp >= 4 && warp < 8) {
    // ---------------------------------------------------------------- epilogue: + shift, ReLU, bf16
    const int quarter = warp - 4;
    const int m = quarter * 32 + lane;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
      int n, y0, x0;
      decode(tile, n, y0, x0);
      const int oy = y0 + (m >> 4), ox = x0 + (m & 15);
      const bool valid = oy < p.OH && ox < p.OW;
      __nv_bfloat16* dst = p.y + ((static_cast<long long>(n) * p.OH + oy) * p.OW + ox) * 64;
      mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + acc * 64;
#pragma unroll
      for (int c = 0; c < 64; c += 32) {
        uint32_t r[32];
        tmem_ld_32x32(t_row + c, r);
        tmem_ld_wait();
        if (valid) {
#pragma unroll
          for (int j = 0; j < 32; j += 8) {
            const float lo = p.relu ? 0.f : -3.0e38f;      // ReLU as a lower clamp that can be switched off
            const float4 s0 = __ldg(reinterpret_cast<const float4*>(p.shift + c + j));
            const float4 s1 = __ldg(reinterpret_cast<const float4*>(p.shift + c + j + 4));
            uint4 o;
            o.x = pack_bf16x2(fmaxf(__uint_as_float(r[j]) + s0.x, lo), fmaxf(__uint_as_float(r[j + 1]) + s0.y, lo));
            o.y = pack_bf16x2(fmaxf(__uint_as_float(r[j + 2]) + s0.z, lo), fmaxf(__uint_as_float(r[j + 3]) + s0.w, lo));
            o.z = pack_bf16x2(fmaxf(__uint_as_float(r[j + 4]) + s1.x, lo), fmaxf(__uint_as_float(r[j + 5]) + s1.y, lo));
            o.w = pack_bf16x2(fmaxf(__uint_as_float(r[j + 6]) + s1.z, lo), fmaxf(__uint_as_float(r[j + 7]) + s1.w, lo));
            *reinterpret_cast<uint4*>(dst + c + j) = o;
          }
        }
      }
      tc_fence_before();
      mbar_arrive(tempty_bar(acc));
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 128);
}

}  // namespace b200

extern "C" int b200_stem_conv_ex(const void* x, const void* w_packed, const float* shift, void* y, int N, int H, int W,
                              int relu, void* stream) {
  using namespace b200;
  static std::atomic<unsigned long long> attr{0};   // per device
  if (device_needs_setup(attr)) {
    B200_CUDA_OK(cudaFuncSetAttribute(stem_conv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kStemSmem));
    device_setup_done(attr);
  }
  StemParams p;
  p.x = static_cast<const __nv_bfloat16*>(x);
  p.w = static_cast<const __nv_bfloat16*>(w_packed);
  p.shift = shift;
  p.relu = relu;
  p.y = static_cast<__nv_bfloat16*>(y);
  p.N = N;
  p.H = H;
  p.W = W;
  p.OH = (H + 1) / 2;
  p.OW = (W + 1) / 2;
  p.tiles_h = (p.OH + 7) / 8;
  p.tiles_w = (p.OW + 15) / 16;
  p.total_tiles = N * p.tiles_h * p.tiles_w;
  const int grid = p.total_tiles < num_sms() ? p.total_tiles : num_sms();
  stem_conv_kernel<<<grid, kStemThreads, kStemSmem, static_cast<cudaStream_t>(stream)>>>(p);
  return check_launch("stem_conv_kernel");
}

// ---------------------------------------------------------------------------------------------
// Stem weight gradient:  dW[k][co] += scale[co] * sum_pixels dY[p, co] * Xcol[p, k],  k = r*28 + s*4 + c.
// GEMM: M = co (64 valid of 128), N = 256 (196 valid), K = output pixels.  A = dY tiles by TMA
// (MN-major), B = the same gathered im2col tile as the forward kernel, used as an MN-major operand
// (rows = pixels).  One CTA per SM accumulates its share of the pixels in TMEM and flushes once with
// fp32 atomics into the master-layout gradient [49][64][3].
namespace b200 {

constexpr int kSwStages = 3;
constexpr int kSwAStage = 2 * 8192;   // chunk 0 = dY [64 px][64 co]; chunk 1 stays zero (co 64..127)
constexpr int kSwBStage = 4 * 8192;   // [64 px][256 k] as four 64-column chunks
constexpr int kSwThreads = 512;       // w0 TMA+MMA, w1 TMEM alloc, w4-7 epilogue, w8-15 producers
constexpr int kSwSmem = kSwStages * (kSwAStage + kSwBStage) + 1024 + 128;

struct StemWgradParams {
  CUtensorMap dy_map;
  const __nv_bfloat16* x;   // [N, H, W, 4]
  const float* scale;       // [64] or null
  float* dw;                // fp32 [49][64][3], accumulated into
  int N, H, W, OH, OW;
  int tiles_h, tiles_w, total_tiles;
};

__global__ void __launch_bounds__(kSwThreads, 1) stem_wgrad_kernel(const __grid_constant__ StemWgradParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t a_smem = smem_base;
  const uint32_t b_smem = smem_base + kSwStages * kSwAStage;
  const uint32_t bar_base = b_smem + kSwStages * kSwBStage;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kSwStages + s); };
  const uint32_t tfull_bar = bar_base + 8u * (2 * kSwStages);
  const uint32_t tmem_ptr_smem = bar_base + 64;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  for (int i = threadIdx.x; i < kSwStages * (kSwAStage + kSwBStage) / 16; i += kSwThreads)
    reinterpret_cast<uint4*>(smem_gen)[i] = make_uint4(0u, 0u, 0u, 0u);
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.dy_map);
    for (int s = 0; s < kSwStages; ++s) {
      mbar_init(full_bar(s), 257);   // 256 gather threads + the TMA issuer
      mbar_init(empty_bar(s), 1);
    }
    mbar_init(tfull_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_ptr_smem, 256);
    tmem_relinquish();
  }
  fence_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_ptr_smem));

  auto decode = [&](int tile, int& n, int& y0, int& x0) {
    const int per_img = p.tiles_h * p.tiles_w;
    n = tile / per_img;
    const int rem = tile - n * per_img;
    y0 = (rem / p.tiles_w) * 4;
    x0 = (rem % p.tiles_w) * 16;
  };
  const int my_tiles = (p.total_tiles - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);

  if (warp >= 8) {
    // gather: thread -> (pixel m, half of the filter rows)
    const int t = (warp - 8) * 32 + lane;
    const int m = t & 63, part = t >> 6;            // 4 parts: filter rows {0,1} {2,3} {4,5} {6}
    const int r_lo = part * 2, r_hi = part == 3 ? 7 : part * 2 + 2;
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
      int n, y0, x0;
      decode(tile, n, y0, x0);
      const int oy = y0 + (m >> 4), ox = x0 + (m & 15);
      const bool pix_ok = oy < p.OH && ox < p.OW;   // dY is zero-filled there, but keep X finite anyway
      mbar_wait(empty_bar(stage), phase ^ 1);
      uint8_t* b_row = smem_gen + kSwStages * kSwAStage + stage * kSwBStage + m * 128;
      const int sw = m & 7;
      uint2 v[2][7];                                // both rows' loads in flight before the first store
#pragma unroll
      for (int rr = 0; rr < 2; ++rr) {
        const int r = r_lo + rr;
        const int iy = 2 * oy + r - 3;
        const bool row_ok = r < r_hi && pix_ok && iy >= 0 && iy < p.H;
        const __nv_bfloat16* src = p.x + (static_cast<long long>(n) * p.H + (row_ok ? iy : 0)) * p.W * 4;
#pragma unroll
        for (int s = 0; s < 7; ++s) {
          const int ix = 2 * ox + s - 3;
          v[rr][s] = make_uint2(0u, 0u);
          if (row_ok && ix >= 0 && ix < p.W) v[rr][s] = __ldg(reinterpret_cast<const uint2*>(src + static_cast<long long>(ix) * 4));
        }
      }
#pragma unroll
      for (int rr = 0; rr < 2; ++rr) {
        const int r = r_lo + rr;
        if (r < r_hi) {
#pragma unroll
          for (int s = 0; s < 7; ++s) {
            const int kbyte = (r * 28 + s * 4) * 2;
            const int kb = kbyte >> 7, within = kbyte & 127;
            *reinterpret_cast<uint2*>(b_row + kb * 8192 + (((within >> 4) ^ sw) << 4) + (within & 15)) = v[rr][s];
          }
        }
      }
      fence_async_smem();
      mbar_arrive(full_bar(stage));
      if (++stage == kSwStages) {
        stage = 0;
        phase ^= 1;
      }
    }
  } else if (warp == 0) {
    // TMA issue for dY + MMA issue (one lane does both: the loop is short and strictly ordered)
    const uint32_t idesc = make_idesc_bf16(128, 256, 1, 1);
    int stage_p = 0, stage_c = 0;
    uint32_t phase_p = 0, phase_c = 0;
    int issued = 0;
    int tile_p = blockIdx.x;
    for (int it = 0; it < my_tiles; ++it) {
      // keep the dY loads kSwStages ahead
      while (issued < my_tiles && issued < it + kSwStages) {
        mbar_wait(empty_bar(stage_p), phase_p ^ 1);
        if (elect_one()) {
          int n, y0, x0;
          decode(tile_p, n, y0, x0);
          mbar_expect_tx(full_bar(stage_p), 8192);
          tma_load_4d(a_smem + stage_p * kSwAStage, &p.dy_map, full_bar(stage_p), 0, x0, y0, n);
        }
        __syncwarp();
        tile_p += gridDim.x;
        ++issued;
        if (++stage_p == kSwStages) {
          stage_p = 0;
          phase_p ^= 1;
        }
      }
      mbar_wait(full_bar(stage_c), phase_c);
      tc_fence_after();
      if (elect_one()) {
        const uint64_t adesc = make_smem_desc(a_smem + stage_c * kSwAStage, 8192, 1024);
        const uint64_t bdesc = make_smem_desc(b_smem + stage_c * kSwBStage, 8192, 1024);
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_bf16(tmem_base, adesc + 128 * k, bdesc + 128 * k, idesc, (it | k) != 0);
        umma_commit(empty_bar(stage_c));
        if (it == my_tiles - 1) umma_commit(tfull_bar);
      }
      __syncwarp();
      if (++stage_c == kSwStages) {
        stage_c = 0;
        phase_c ^= 1;
      }
    }
  } else if (warp >= 4 && warp < 6 && my_tiles > 0) {
    // epilogue: rows co = 0..63 live in TMEM lanes 0..63 (quarters 0 and 1)
    const int quarter = warp - 4;
    const int co = quarter * 32 + lane;
    const float rs = p.scale ? __ldg(p.scale + co) : 1.f;
    mbar_wait(tfull_bar, 0);
    tc_fence_after();
    const uint32_t t_row = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16);
    for (int c = 0; c < 224; c += 32) {
      uint32_t r[32];
      tmem_ld_32x32(t_row + c, r);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const int k = c + j;
        const int ci = k & 3;
        if (k < 196 && ci < 3) {
          const int tap = (k / 28) * 7 + (k % 28) / 4;
          atomicAdd(p.dw + (tap * 64 + co) * 3 + ci, __uint_as_float(r[j]) * rs);
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 256);
}

}  // namespace b200

extern "C" int b200_stem_wgrad(const void* x, const void* dy, const float* scale, float* dw, int N, int H, int W,
                               void* stream) {
  using namespace b200;
  static std::atomic<unsigned long long> attr{0};   // per device
  if (device_needs_setup(attr)) {
    B200_CUDA_OK(cudaFuncSetAttribute(stem_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSwSmem));
    device_setup_done(attr);
  }
  StemWgradParams p;
  memset(&p, 0, sizeof(p));
  p.x = static_cast<const __nv_bfloat16*>(x);
  p.scale = scale;
  p.dw = dw;
  p.N = N;
  p.H = H;
  p.W = W;
  p.OH = (H + 1) / 2;
  p.OW = (W + 1) / 2;
  p.tiles_h = (p.OH + 3) / 4;
  p.tiles_w = (p.OW + 15) / 16;
  p.total_tiles = N * p.tiles_h * p.tiles_w;
  unsigned long long dims[4] = {64ull, static_cast<unsigned long long>(p.OW), static_cast<unsigned long long>(p.OH),
                                static_cast<unsigned long long>(N)};
  unsigned long long strides[3] = {128ull, static_cast<unsigned long long>(p.OW) * 128ull,
                                   static_cast<unsigned long long>(p.OH) * p.OW * 128ull};
  const unsigned int box[4] = {64u, 16u, 4u, 1u};
  int rc = make_tmap_bf16(&p.dy_map, dy, 4, dims, strides, box);
  if (rc) return rc;
  const int grid = deterministic() ? 1 : (p.total_tiles < num_sms() ? p.total_tiles : num_sms());
  stem_wgrad_kernel<<<grid, kSwThreads, kSwSmem, static_cast<cudaStream_t>(stream)>>>(p);
  return check_launch("stem_wgrad_kernel");
}

extern "C" int b200_stem_conv(const void* x, const void* w_packed, const float* shift, void* y, int N, int H, int W,
                              void* stream) {
  return b200_stem_conv_ex(x, w_packed, shift, y, N, H, W, 1, stream);
}
