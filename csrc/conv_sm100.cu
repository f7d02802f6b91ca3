// Implicit-GEMM convolution (fprop and dgrad) for sm_100a: TMA + tcgen05.mma + TMEM, TMA-fed epilogue.
//
//   out[n, y, x, :] = epilogue( sum_{tap} sum_{k} in[n, y + dh(tap), x + dw(tap), k] * w[tap][:, k] )
//
// GEMM view: M = output pixels (tiles of th x tw = 128 pixels of ONE image, so a tap is a pure
// (dh, dw) shift of the TMA box and the conv halo / padding is TMA's out-of-bounds zero fill),
// N = output channels (BN <= 256 per tile), K = taps x input channels (64 per pipeline stage).
//
// Warp roles (672 threads, 1 CTA per SM, persistent over tiles):
//   warp 0   activation producer (one elected thread): one TMA box per GROUP of the K loop (see ConvParams::mode)
//   warp 20  weight producer (one elected thread): box [64ch, BN rows, 1 tap] per sub-step, own ring; idle after the
//            first tile when all weight tiles of the layer fit its ring (they stay resident)
//   warp 1   MMA issuer: per group one wait, per sub-step 4 x tcgen05.mma (K=16) under elect.sync, commits
//   warp 2   TMEM allocator (512 columns = two 256-column accumulators), then the STORE issuer: one TMA store per
//            finished 128-pixel x 64-channel chunk, then releases the chunk's smem slot.  Optionally sums the
//            chunk's columns first (the bias gradient of the layer that consumes this output as its dY)
//   warp 3   epilogue-operand loader: TMA loads of the residual / ReLU-mask chunk of every tile into smem slots,
//            running ahead of the MMA as far as free slots allow
//   warps 4-19 epilogue math: tcgen05.ld (32 lanes x 32 columns), + shift, + residual, ReLU mask, ReLU, bf16,
//            written IN PLACE over the operand chunk in its slot (row-per-thread, conflict-free under the 128 B swizzle)
//
// The first version of this kernel (git history: round 1) moved the epilogue operands and results through registers
// with LSU loads / stores: the memory-bound bottleneck layers then ran at 3.2-3.8 TB/s, limited by the epilogue's
// instruction issue and by how many bytes its loads kept in flight (18.60 -> 17.67 ms per step at batch 8).  Here the epilogue touches global memory only
// through TMA: nothing but smem <-> register traffic and the math is left in the 16 epilogue warps, and the smem not
// needed by the operand ring (memory-bound layers have short K loops) becomes up to 12 chunk slots of prefetch depth.
//
// Stride-2 convolutions use one tensor map per input parity class (strided views), so every tap is still a plain
// shifted box.  Stride-2 dgrad computes one output parity class per launch through a strided output map.
//
// Replaces cuDNN fprop/dgrad + separate bias/BN/ReLU/add kernels of the reference stack
// (SURVEY.md G2-G4, G6, G8, G12); the reference itself has no CUDA source.
#include <cstdlib>
#include <cstring>

#include "../include/b200_ops.h"
#include "common.cuh"
#include "ptx_sm100.cuh"

#include <type_traits>

namespace b200 {

constexpr int kBM = 128;
constexpr int kBK = 64;
constexpr int kAStageBytes = kBM * kBK * 2;       // 16 KB
constexpr int kMaxAStages = 6;
constexpr int kMaxBStages = 12;
constexpr int kMaxSlots = 12;
constexpr int kSlotBytes = kBM * 64 * 2;          // one [128 pixels][64 channels] bf16 chunk, 128 B rows, SWIZZLE_128B
constexpr int kMaxTaps = 49;
constexpr int kMaxMaps = 8;
constexpr int kTmemCols = 512;
constexpr int kEpiWarps = 16;                     // 4 TMEM lane quarters x 2 column halves x 2 chunk groups
constexpr int kConvThreads = 128 + 32 * kEpiWarps + 32;    // + the weight producer (warp 20)
constexpr int kWeightWarp = kConvThreads / 32 - 1;
constexpr int kShiftBytes = kEpiWarps * 128;      // per epilogue warp: the 32 shift values of its current chunk half
constexpr int kBarBytes = 640;
constexpr int kSmemLimit = 232448;                // 227 KB

struct Tap {
  int8_t dh, dw, map, pad;
  int widx;
};

struct DevProblem {
  int OH, OW;           // GEMM pixel grid
  int a_bytes;          // bytes of one activation box
  int sub_stride;       // byte distance between the sub-tiles of a box
  int tiles_h, tiles_w; // spatial tiles per image
  int tile_begin;       // first m-tile of this problem
  int th_log2, tw_log2;
  int map_base;
};

struct ConvParams {
  CUtensorMap in_maps[kMaxMaps];
  CUtensorMap w_map;
  CUtensorMap out_maps[B200_MAX_PROBLEMS];
  CUtensorMap add_maps[B200_MAX_PROBLEMS];
  CUtensorMap mask_maps[B200_MAX_PROBLEMS];
  DevProblem prob[B200_MAX_PROBLEMS];
  Tap taps[kMaxTaps];
  const float* shift;
  float* colsum[B200_MAX_PROBLEMS];   // optional: += column sums of the problem's output (fp32 [Nrows])
  int num_problems, num_m_tiles, num_n_tiles;
  int Nrows, BN;
  int num_taps, num_kc;
  int relu, add_after_mask;
  // K loop = activation GROUPS (one TMA box each) x up to gsz SUB-STEPS (one weight tile, 4 MMAs each):
  //   mode 0 plain: box [64ch, tw, th, 1], gsz = 1, one group per (tap, k chunk);
  //   mode 1 halo (3x3, stride 1): box [64ch, tw, th + 2, 1] serves the three taps dh = -1, 0, +1 of one filter
  //          column as row-offset views (1024-byte aligned), so activations cross L2 -> smem 3 (th+2)/th times, not 9;
  //   mode 2 kc (K a multiple of 64, >= 2 chunks): 5-D box [64ch, tw, th, 1, gsz chunks] = gsz k chunks per issue.
  // Fewer, larger boxes matter because one thread issues them: ~600 cycles of instruction latency per issue bounded
  // every layer with N <= 128 (ncu: producer and MMA warps busy, tensor pipe 20 % active).
  int mode, gsz, tap_groups, kgroups;
  int a_stages, b_stages, slots, a_stage_bytes, b_stage_bytes, smem_bytes;
  int widx3[3][3];      // halo mode: [dw + 1][dh + 1] -> weight tap index
  // pair mode (streamed weights): the launch runs as clusters of two CTAs that work on neighbouring pixel tiles of the
  // same N tile in lock step and share the weight stream -- each CTA fetches HALF of every weight tile and TMA
  // multicasts it into both shared memories, so a weight tile crosses L2 -> SM once per pair instead of once per CTA
  // (ncu on the res4 1x1 layers: 7.5-10 TB/s of L2 -> SM traffic, two thirds of it identical weight tiles, tensor
  // pipe 30-46 % active).  num_m_tiles is then padded to an even count; real_m_tiles is the true one and the extra
  // tile only keeps the shared ring moving.
  int pair, real_m_tiles;
  long long* trace;     // optional in-kernel timeline: 16 clock64 stamps per CTA (b200_conv_set_trace)
};

__device__ __forceinline__ void decode_tile(const ConvParams& p, int tile, int& pi, int& n, int& y0, int& x0,
                                            int& n_tile) {
  const int m_tile = tile % p.num_m_tiles;
  n_tile = tile / p.num_m_tiles;
  pi = 0;
#pragma unroll
  for (int i = 1; i < B200_MAX_PROBLEMS; ++i)
    if (i < p.num_problems && m_tile >= p.prob[i].tile_begin) pi = i;
  const DevProblem& q = p.prob[pi];
  const int local = m_tile - q.tile_begin;
  const int per_img = q.tiles_h * q.tiles_w;
  n = local / per_img;
  const int rem = local - n * per_img;
  const int ty = rem / q.tiles_w;
  const int tx = rem - ty * q.tiles_w;
  y0 = ty << q.th_log2;
  x0 = tx << q.tw_log2;
}

// 64-column chunks of an N tile that hold at least one (8-padded) valid output channel
__device__ __forceinline__ int tile_chunks(const ConvParams& p, int n_tile) {
  const int col0 = n_tile * p.BN;
  const int col_end = min(p.Nrows, col0 + p.BN);
  return (((col_end + 7) & ~7) - col0 + 63) >> 6;
}

// Ring of chunk slots.  A chunk takes 1 position (operand loaded in place, or no operand) or 2 (residual + mask).
struct SlotCursor {
  int slot;
  uint32_t par;
  __device__ __forceinline__ void advance(int n, int nslots) {
    slot += n;
    if (slot >= nslots) {
      slot -= nslots;
      par ^= 1u;
    }
  }
};

template <bool kAdd, bool kMask>
__global__ void __launch_bounds__(kConvThreads, 1) conv_igemm_kernel(const __grid_constant__ ConvParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  constexpr bool kLoads = kAdd || kMask;
  constexpr int kPPC = (kAdd && kMask) ? 2 : 1;     // slot positions per chunk
  const int a_stages = p.a_stages, b_stages = p.b_stages, nslots = p.slots;
  const uint32_t raw_base = smem_u32(smem_raw);
  const uint32_t smem_base = (raw_base + 1023u) & ~1023u;
  const uint32_t a_smem = smem_base;
  const uint32_t b_smem = a_smem + a_stages * p.a_stage_bytes;
  const uint32_t slot_base = b_smem + b_stages * p.b_stage_bytes;
  const uint32_t shift_base = slot_base + nslots * kSlotBytes;
  const uint32_t bar_base = shift_base + kShiftBytes;
  if (bar_base + kBarBytes - raw_base > static_cast<uint32_t>(p.smem_bytes)) __trap();   // layout does not fit
  // barriers (8 bytes each)
  auto afull_bar = [&](int s) { return bar_base + 8u * s; };
  auto aempty_bar = [&](int s) { return bar_base + 48u + 8u * s; };
  auto bfull_bar = [&](int s) { return bar_base + 96u + 8u * s; };
  auto bempty_bar = [&](int s) { return bar_base + 192u + 8u * s; };
  auto tfull_bar = [&](int a) { return bar_base + 288u + 8u * a; };
  auto tempty_bar = [&](int a) { return bar_base + 304u + 8u * a; };
  auto ldfull_bar = [&](int s) { return bar_base + 320u + 8u * s; };    // operand chunk landed in slot s
  auto free_bar = [&](int s) { return bar_base + 416u + 8u * s; };      // slot s may be overwritten
  auto stfull_bar = [&](int s) { return bar_base + 512u + 8u * s; };    // result chunk complete in slot s
  const uint32_t tmem_ptr_smem = bar_base + 608u;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int total_tiles = p.num_m_tiles * p.num_n_tiles;
  const bool pair = p.pair != 0;
  const uint32_t crank = pair ? cluster_ctarank() : 0u;
  // pair mode: the odd CTA's last tile may lie past the end (every role but the weight ring skips it)
  auto is_dummy = [&](int tile) { return pair && (tile % p.num_m_tiles) >= p.real_m_tiles; };
  // in-kernel timeline (build with -DB200_CONV_TRACE: B200_CONV_TRACE=1 python build_native.py); compiled out by
  // default, the pointer alone costs the epilogue warps two registers
#ifdef B200_CONV_TRACE
  long long* const tr = p.trace ? p.trace + static_cast<size_t>(blockIdx.x) * 16 : nullptr;
#define B200_TR(i) do { if (tr) tr[i] = clock64(); } while (0)
  if (tr && threadIdx.x == 0) {
    unsigned long long gt;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
    tr[0] = static_cast<long long>(gt);
    tr[1] = clock64();
  }
#else
#define B200_TR(i) do { } while (0)
#endif
  const int gsz = p.gsz, kgroups = p.kgroups;
  const int num_groups = p.tap_groups * kgroups;          // activation boxes per tile
  const int num_substeps = p.num_taps * p.num_kc;         // weight tiles per tile
  const uint32_t b_bytes = static_cast<uint32_t>(p.BN) * kBK * 2;
  // Weights stay resident: when every weight tile of the K loop has its own ring stage and there is a single N
  // tile, stage s always holds the same tile, so it is fetched once (first pass) and never released.
  const bool weights_resident = (p.num_n_tiles == 1) && (b_stages == num_substeps);
  // sub-steps of k group kg (the last group of a tap may be short in kc mode)
  auto nsub = [&](int kg) { return p.mode == 2 ? min(gsz, p.num_kc - kg * gsz) : gsz; };

  // Set-up is on the critical path of every launch (one CTA per SM with ~227 KB of shared memory: the next kernel's
  // CTAs cannot become resident before this one's exit, so programmatic launch overlaps little of it).  The in-kernel
  // timeline showed 1.8k cycles to the first __syncthreads and 2.2k more before the first TMA issue, most of it one
  // thread initialising 76 barriers and dependent constant-bank misses in decode_tile: spread both over lanes / warps.
  if (warp == 0) {
    // one descriptor per lane: in_maps[0..7], w_map, then out / add / mask per problem
    const void* d = nullptr;
    if (lane < kMaxMaps) d = &p.in_maps[lane];
    else if (lane == kMaxMaps) d = &p.w_map;
    else if (lane < kMaxMaps + 1 + 3 * B200_MAX_PROBLEMS) {
      const int k = lane - kMaxMaps - 1, i = k / 3, which = k - 3 * i;
      if (i < p.num_problems) {
        if (which == 0) d = &p.out_maps[i];
        else if (which == 1 && kAdd) d = &p.add_maps[i];
        else if (which == 2 && kMask) d = &p.mask_maps[i];
      }
    }
    if (d) tma_prefetch_desc(d);
  }
  if (warp == 1) {
    // barrier k sits at bar_base + 8 k: [0,12) activation full / empty, [12,24) weight full, [24,36) weight empty,
    // 36-37 accumulator full, 38-39 accumulator empty, [40,52) operand landed, [52,64) slot free, [64,76) chunk complete
    for (int k = lane; k < 76; k += 32) {
      uint32_t count = 1;
      if (k >= 24 && k < 36) count = pair ? 2 : 1;      // pair mode: both CTAs' MMA warps release a weight stage
      else if (k == 38 || k == 39) count = 32 * kEpiWarps;
      else if (k >= 64) count = 256;
      mbar_init(bar_base + 8u * k, count);
    }
    fence_barrier_init();
  }
  // first tile of this CTA, decoded by the producer warp while the other warps set up (warms the constant bank)
  // (results parked in the spare bytes after the barriers, not in registers every warp would have to carry)
  const uint32_t t0_smem = bar_base + 616u;
  if (warp == 0 && lane == 0 && static_cast<int>(blockIdx.x) < total_tiles && !is_dummy(blockIdx.x)) {
    int pi, n, y0, x0, nt;
    decode_tile(p, blockIdx.x, pi, n, y0, x0, nt);
    asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(t0_smem), "r"(pi), "r"(n) : "memory");
    asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(t0_smem + 8u), "r"(y0), "r"(x0) : "memory");
    asm volatile("st.shared.b32 [%0], %1;" ::"r"(t0_smem + 16u), "r"(nt) : "memory");
  }
  if (warp == 2) {
    tmem_alloc(tmem_ptr_smem, kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  if (pair) cluster_sync_all();          // the partner's barriers exist before anything is multicast into them
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_ptr_smem));
  // set-up done: let the next kernel start its own set-up, then wait for the previous kernel's results
  pdl_launch_dependents();
  pdl_wait();
  if (threadIdx.x == 0) B200_TR(2);

  if (warp == 0) {
    // ------------------------------------------------------------------ activation producer (one elected thread)
    // elect.sync, not `lane == 0`: ptxas then knows the region is single-lane and keeps the TMA operands in uniform
    // registers instead of wrapping every instruction in a leader-election loop.
    if (elect_one()) {
      int sa = 0;
      uint32_t pa = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        if (is_dummy(tile)) continue;
        int pi, n, y0, x0, n_tile;
        if (tile == static_cast<int>(blockIdx.x)) {
          asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(pi), "=r"(n) : "r"(t0_smem) : "memory");
          asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(y0), "=r"(x0) : "r"(t0_smem + 8u) : "memory");
          asm volatile("ld.shared.b32 %0, [%1];" : "=r"(n_tile) : "r"(t0_smem + 16u) : "memory");
        } else {
          decode_tile(p, tile, pi, n, y0, x0, n_tile);
        }
        const DevProblem& q = p.prob[pi];
        const uint32_t a_bytes = static_cast<uint32_t>(q.a_bytes);
        for (int t = 0; t < p.tap_groups; ++t) {
          int x = x0, y = y0, map = q.map_base;
          if (p.mode == 1) {            // filter column t - 1, rows y0 - 1 .. y0 + th
            x += t - 1;
            y -= 1;
          } else {
            const Tap tap = p.taps[t];
            x += tap.dw;
            y += tap.dh;
            map += tap.map;
          }
          const void* amap = &p.in_maps[map];
          for (int kg = 0; kg < kgroups; ++kg) {
            mbar_wait(aempty_bar(sa), pa ^ 1);
            mbar_expect_tx(afull_bar(sa), a_bytes);
            const uint32_t dst = a_smem + sa * p.a_stage_bytes;
            if (p.mode == 2) tma_load_5d(dst, amap, afull_bar(sa), 0, x, y, n, kg * gsz);
            else tma_load_4d(dst, amap, afull_bar(sa), kg * kBK, x, y, n);
            if (tile == static_cast<int>(blockIdx.x) && t == 0 && kg == 0) B200_TR(3);
            if (++sa == a_stages) {
              sa = 0;
              pa ^= 1;
            }
          }
        }
      }
    }
  } else if (warp == kWeightWarp) {
    // ------------------------------------------------------------------ weight producer (one elected thread)
    if (elect_one()) {
      int sb = 0;
      uint32_t pb = 0;
      const uint32_t half_off = crank * (b_bytes >> 1);
      const int half_row = static_cast<int>(crank) * (p.BN >> 1);
      auto produce = [&](auto pair_c) {          // `pair` resolved once, outside the issue loop (see the MMA warp)
        constexpr bool kPair = decltype(pair_c)::value;
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
          const int row0 = (tile / p.num_m_tiles) * p.BN;
          for (int t = 0; t < p.tap_groups; ++t) {
            for (int kg = 0; kg < kgroups; ++kg) {
              const int ns = nsub(kg);
              for (int j = 0; j < ns; ++j) {
                const int widx = p.mode == 1 ? p.widx3[t][j] : p.taps[t].widx;
                const int kc = p.mode == 2 ? kg * gsz + j : kg;
                mbar_wait(bempty_bar(sb), pb ^ 1);
                mbar_expect_tx(bfull_bar(sb), b_bytes);     // pair mode: both halves, mine and the partner's
                if constexpr (kPair)
                  tma_load_3d_mc(b_smem + sb * p.b_stage_bytes + half_off, &p.w_map, bfull_bar(sb), kc * kBK,
                                 row0 + half_row, widx, static_cast<uint16_t>(3));
                else
                  tma_load_3d(b_smem + sb * p.b_stage_bytes, &p.w_map, bfull_bar(sb), kc * kBK, row0, widx);
                if (++sb == b_stages) {
                  sb = 0;
                  pb ^= 1;
                }
              }
            }
          }
          if (!kPair && weights_resident) break;      // one pass filled every stage for good
        }
      };
      if (pair) produce(std::true_type{});
      else produce(std::false_type{});
      if (pair) {
        // every release the partner still owes my barriers has landed before either CTA may exit
        for (int i = 0; i < b_stages; ++i) {
          mbar_wait(bempty_bar(sb), pb ^ 1);
          if (++sb == b_stages) {
            sb = 0;
            pb ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (one elected thread)
    // This thread's instruction latency IS the kernel's speed for every layer with N <= 128 (a sub-step is 128 cycles
    // of tensor work at N = 64; ncu showed ~600 cycles of address arithmetic, elections and reconvergence around
    // it).  So: one elect.sync for the whole loop (operands stay in uniform registers, no per-step ELECT / BSSY /
    // BSYNC), descriptors advanced incrementally in 16-byte units (the 14-bit address field never carries: smem
    // addresses are < 256 KB), no divisions in the K loop.
    if (elect_one()) {
      const uint64_t desc_hi = make_smem_desc(0, 16, 1024);        // everything but the start address
      const uint32_t a_first = a_smem >> 4, b_first = b_smem >> 4;
      const uint32_t a_step = static_cast<uint32_t>(p.a_stage_bytes) >> 4, b_step = static_cast<uint32_t>(p.b_stage_bytes) >> 4;
      uint32_t a_cur = a_first, b_cur = b_first;
      uint32_t afull = afull_bar(0), aempty = aempty_bar(0), bfull = bfull_bar(0), bempty = bempty_bar(0);
      int sa = 0, sb = 0;
      uint32_t pa = 0, pb = 0;
      bool b_wait = true;                 // false once resident weights have all landed
      int acc = 0;
      uint32_t acc_phase = 0;
      const bool one_problem = p.num_problems == 1;
      uint32_t sub_step = static_cast<uint32_t>(p.prob[0].sub_stride) >> 4;
      // The loop exists twice, selected ONCE outside it: with `pair` tested at every sub-step the compiler reloaded it
      // (and the ring size) from the constant bank inside the issue loop and every N <= 128 layer lost 30 %.
      auto issue_loop = [&](auto pair_c) {
        constexpr bool kPair = decltype(pair_c)::value;
        auto next_b = [&]() {
          b_cur += b_step;
          bfull += 8;
          bempty += 8;
          if (++sb == b_stages) {
            sb = 0;
            pb ^= 1;
            b_cur = b_first;
            bfull = bfull_bar(0);
            bempty = bempty_bar(0);
            if (!kPair && weights_resident) b_wait = false;
          }
        };
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
          const int n_tile = tile / p.num_m_tiles;
          if (kPair && (tile - n_tile * p.num_m_tiles) >= p.real_m_tiles) {
            // no pixels here: receive and release the shared weight tiles only
            for (int j = 0; j < num_substeps; ++j) {
              mbar_wait(bfull, pb);
              umma_commit_mc(bempty, static_cast<uint16_t>(3));
              next_b();
            }
            continue;
          }
          if (!one_problem) {               // halo tiles differ per pyramid level
            int pi, n, y0, x0, nt2;
            decode_tile(p, tile, pi, n, y0, x0, nt2);
            sub_step = static_cast<uint32_t>(p.prob[pi].sub_stride) >> 4;
          }
          // the last N tile may be narrower: do not multiply zero-filled weight rows
          const int cols = min(p.Nrows - n_tile * p.BN, p.BN);
          const uint32_t idesc = make_idesc_bf16(kBM, max(16, (cols + 15) & ~15), 0, 0);
          mbar_wait(tempty_bar(acc), acc_phase ^ 1);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + acc * 256;
          uint32_t accumulate = 0;
          int kg = 0;
          for (int g = 0; g < num_groups; ++g) {
            const int ns = p.mode == 2 ? min(gsz, p.num_kc - kg * gsz) : gsz;
            if (++kg == kgroups) kg = 0;
            mbar_wait(afull, pa);
            tc_fence_after();
            if (tile == static_cast<int>(blockIdx.x) && g == 0) B200_TR(4);
            if (tile == static_cast<int>(blockIdx.x) && g == num_groups - 1) B200_TR(13);
            uint32_t a_sub = a_cur;
            for (int j = 0; j < ns; ++j) {
              if (kPair || b_wait) {
                mbar_wait(bfull, pb);
                tc_fence_after();
                if (tile == static_cast<int>(blockIdx.x) && g == 0 && j == 0) B200_TR(5);
              }
              const uint64_t adesc = desc_hi | a_sub, bdesc = desc_hi | b_cur;
#pragma unroll
              for (int k = 0; k < kBK / 16; ++k) {
                // advance 16 bf16 = 32 bytes along K inside the 128B swizzle row: +2 in (addr >> 4) units
                umma_bf16(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, accumulate | k);
              }
              if constexpr (kPair) {
                umma_commit_mc(bempty, static_cast<uint16_t>(3));
              } else {
                if (!weights_resident) umma_commit(bempty);
              }
              accumulate = 1;
              a_sub += sub_step;
              next_b();
            }
            umma_commit(aempty);
            if (g == num_groups - 1) {
              umma_commit(tfull_bar(acc));
              B200_TR(tile == static_cast<int>(blockIdx.x) ? 6 : 9);
            }
            a_cur += a_step;
            afull += 8;
            aempty += 8;
            if (++sa == a_stages) {
              sa = 0;
              pa ^= 1;
              a_cur = a_first;
              afull = afull_bar(0);
              aempty = aempty_bar(0);
            }
          }
          acc ^= 1;
          if (acc == 0) acc_phase ^= 1;
        }
      };
      if (pair) issue_loop(std::true_type{});
      else issue_loop(std::false_type{});
    }
  } else if (warp == 2) {
    // ------------------------------------------------------------------ store issuer (one thread)
    // Waits until all 256 threads of a chunk have written (and proxy-fenced) their results, stores the chunk with
    // one TMA box -- pixels / channels outside the tensor are clipped by the hardware -- and hands the slot(s)
    // back once the TMA engine has read them.  One store is in flight per CTA; a 16 KB smem read is short next
    // to the time a tile takes.
    //
    // Column sums (bias gradients): when the output of this launch is the dY of a biased layer, its bias gradient
    // is the sum of this tensor over pixels.  The finished chunk sits in smem anyway, so the warp adds up its
    // columns (lane -> 2 adjacent channels, 128 conflict-free 4-byte reads) before the store, keeps the sums in
    // registers over all tiles of one (problem, N tile) and flushes them with one atomic per column.  Rows of a
    // tile that lie outside the pixel grid are skipped: their accumulators are not zero (a tap can reach back
    // into the image), only their store is clipped.
    {
      SlotCursor cur{0, 0u};
      float cs[4][2];
#pragma unroll
      for (int c = 0; c < 4; ++c) cs[c][0] = cs[c][1] = 0.f;
      int cs_pi = -1, cs_nt = -1;
      auto flush = [&]() {
        if (cs_pi >= 0) {
          float* dst = p.colsum[cs_pi];
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const int col = cs_nt * p.BN + c * 64 + 2 * lane;
            if (col < p.Nrows && cs[c][0] != 0.f) atomicAdd(dst + col, cs[c][0]);
            if (col + 1 < p.Nrows && cs[c][1] != 0.f) atomicAdd(dst + col + 1, cs[c][1]);
            cs[c][0] = cs[c][1] = 0.f;
          }
        }
      };
      const uint32_t cs_off = (static_cast<uint32_t>(lane & 3) << 2);
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        if (is_dummy(tile)) continue;
        int pi, n, y0, x0, n_tile;
        decode_tile(p, tile, pi, n, y0, x0, n_tile);
        const int nch = tile_chunks(p, n_tile);
        const bool do_cs = p.colsum[pi] != nullptr;
        if (do_cs && (pi != cs_pi || n_tile != cs_nt)) {
          flush();
          cs_pi = pi;
          cs_nt = n_tile;
        }
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          if (c < nch) {
            mbar_wait(stfull_bar(cur.slot), cur.par);
            const uint32_t sA = slot_base + cur.slot * kSlotBytes;
            if (do_cs) {
              const DevProblem& q = p.prob[pi];
              const int ny = q.OH - y0, nx = q.OW - x0, twm = (1 << q.tw_log2) - 1;
              float s0 = 0.f, s1 = 0.f;
#pragma unroll 16
              for (int row = 0; row < kBM; ++row) {
                // unconditional load + select (a branch here would serialise the 128 smem round trips)
                uint32_t w;
                asm volatile("ld.shared.b32 %0, [%1];" : "=r"(w) : "r"(sA + row * 128 + (((lane >> 2) ^ (row & 7)) << 4) + cs_off));
                if ((row >> q.tw_log2) >= ny || (row & twm) >= nx) w = 0u;
                s0 += bf16_lo(w);
                s1 += bf16_hi(w);
              }
              cs[c][0] += s0;
              cs[c][1] += s1;
              __syncwarp();
            }
            if (elect_one()) {
              if constexpr (kPPC == 2) {
                int sb = cur.slot + 1;
                if (sb >= nslots) sb -= nslots;
                mbar_arrive(free_bar(sb));              // the mask chunk has been consumed
              }
              tma_store_4d(&p.out_maps[pi], sA, n_tile * p.BN + c * 64, x0, y0, n);
              tma_store_commit();
              tma_store_wait_read();
              mbar_arrive(free_bar(cur.slot));
            }
            __syncwarp();
            cur.advance(kPPC, nslots);
          }
        }
      }
      flush();
      if (lane == 0) B200_TR(10);
    }
  } else if (warp == 3) {
    // ------------------------------------------------------------------ epilogue-operand loader (one thread)
    if constexpr (kLoads) {
      if (elect_one()) {
        SlotCursor cur{0, 0u};
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
          if (is_dummy(tile)) continue;
          int pi, n, y0, x0, n_tile;
          decode_tile(p, tile, pi, n, y0, x0, n_tile);
          const int nch = tile_chunks(p, n_tile);
          for (int c = 0; c < nch; ++c) {
            const int col = n_tile * p.BN + c * 64;
            mbar_wait(free_bar(cur.slot), cur.par ^ 1);
            mbar_expect_tx(ldfull_bar(cur.slot), kSlotBytes);
            tma_load_4d(slot_base + cur.slot * kSlotBytes, kAdd ? &p.add_maps[pi] : &p.mask_maps[pi],
                        ldfull_bar(cur.slot), col, x0, y0, n);
            if constexpr (kPPC == 2) {
              SlotCursor cb = cur;
              cb.advance(1, nslots);
              mbar_wait(free_bar(cb.slot), cb.par ^ 1);
              mbar_expect_tx(ldfull_bar(cb.slot), kSlotBytes);
              tma_load_4d(slot_base + cb.slot * kSlotBytes, &p.mask_maps[pi], ldfull_bar(cb.slot), col, x0, y0, n);
            }
            cur.advance(kPPC, nslots);
          }
        }
      }
    }
  } else if (warp < 4 + kEpiWarps) {
    // ------------------------------------------------------------------ epilogue math
    // Warp (quarter q, half hh, group gi): TMEM lanes 32q..32q+31 (the rows a warp may read are fixed by
    // warp_id % 4), columns 32hh..32hh+31 of a 64-column chunk, chunks with (running chunk index) % 2 == gi.
    // A thread owns one pixel row of the chunk: its 4 x 16 B pieces sit at (piece ^ (row & 7)) inside the row's
    // 128 bytes (SWIZZLE_128B, the layout TMA reads and writes), so a quarter-warp's 16-byte accesses cover all
    // 32 banks once.
    const int ew = warp - 4;
    const int quarter = ew & 3;
    const int hh = (ew >> 2) & 1;
    const int gi = ew >> 3;
    const int m = quarter * 32 + lane;
    uint32_t poff[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) poff[j] = static_cast<uint32_t>(m) * 128u + (static_cast<uint32_t>((hh * 4 + j) ^ (m & 7)) << 4);
    const uint32_t shift_s = shift_base + ew * 128;
    const float* __restrict__ shiftp = p.shift;
    const bool relu = p.relu != 0, add_post = p.add_after_mask != 0;
    const int BN = p.BN, Nrows = p.Nrows;
    SlotCursor cur{0, 0u};
    int g = 0;                 // running chunk index of this CTA
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      if (is_dummy(tile)) continue;
      const int n_tile = tile / p.num_m_tiles;
      const int col0 = n_tile * BN;
      const int col_end = min(Nrows, col0 + BN);
      const int nch = tile_chunks(p, n_tile);
      // my chunks of this tile (at most 2 of the <= 4) and their shift values, requested before the accumulator wait
      int my_last = -1;
      float shv[2] = {0.f, 0.f};
      {
        int k = 0;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          if (c < nch && ((g + c) & 1) == gi) {
            my_last = c;
            const int col = col0 + c * 64 + hh * 32 + lane;
            if (shiftp && col < Nrows) shv[k] = __ldg(shiftp + col);
            ++k;
          }
        }
      }
      mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      if (ew == 0 && lane == 0) B200_TR(tile == static_cast<int>(blockIdx.x) ? 7 : 14);
      if (my_last < 0) {
        tc_fence_before();
        mbar_arrive(tempty_bar(acc));
      }
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + acc * 256;
      int k = 0;
      for (int c = 0; c < nch; ++c) {
        if (((g + c) & 1) == gi) {
          const int col = col0 + c * 64 + hh * 32;
          const uint32_t sA = slot_base + cur.slot * kSlotBytes;
          uint32_t sB = sA;
          if constexpr (kLoads) {
            mbar_wait(ldfull_bar(cur.slot), cur.par);
            if constexpr (kPPC == 2) {
              SlotCursor cb = cur;
              cb.advance(1, nslots);
              mbar_wait(ldfull_bar(cb.slot), cb.par);
              sB = slot_base + cb.slot * kSlotBytes;
            }
          } else {
            mbar_wait(free_bar(cur.slot), cur.par ^ 1);
          }
          uint32_t r[32];
          tmem_ld_32x32(t_row + c * 64 + hh * 32, r);
          uint4 ka[4], km[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            if constexpr (kAdd)
              asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(ka[j].x), "=r"(ka[j].y), "=r"(ka[j].z), "=r"(ka[j].w) : "r"(sA + poff[j]) : "memory");
            if constexpr (kMask)
              asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(km[j].x), "=r"(km[j].y), "=r"(km[j].z), "=r"(km[j].w) : "r"(sB + poff[j]) : "memory");
          }
          if (shiftp) {
            __syncwarp();
            asm volatile("st.shared.f32 [%0], %1;" ::"r"(shift_s + lane * 4), "f"(k == 0 ? shv[0] : shv[1]) : "memory");
            __syncwarp();
          }
          tmem_ld_wait();
          if (c == my_last) {          // the accumulator is in registers: the MMA of the tile after next may start
            tc_fence_before();
            mbar_arrive(tempty_bar(acc));
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int cj = col + j * 8;
            float v[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(r[j * 8 + e]);
            if (shiftp) {
              float4 s0, s1;
              asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(s0.x), "=f"(s0.y), "=f"(s0.z), "=f"(s0.w) : "r"(shift_s + j * 32) : "memory");
              asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(s1.x), "=f"(s1.y), "=f"(s1.z), "=f"(s1.w) : "r"(shift_s + j * 32 + 16) : "memory");
              v[0] += s0.x; v[1] += s0.y; v[2] += s0.z; v[3] += s0.w;
              v[4] += s1.x; v[5] += s1.y; v[6] += s1.z; v[7] += s1.w;
            }
            if constexpr (kAdd) {
              if (!add_post) {
                v[0] += bf16_lo(ka[j].x); v[1] += bf16_hi(ka[j].x); v[2] += bf16_lo(ka[j].y); v[3] += bf16_hi(ka[j].y);
                v[4] += bf16_lo(ka[j].z); v[5] += bf16_hi(ka[j].z); v[6] += bf16_lo(ka[j].w); v[7] += bf16_hi(ka[j].w);
              }
            }
            if constexpr (kMask) {
              // bf16 > 0  <=>  its 16-bit pattern, read as a signed integer, is > 0
              const uint32_t w4[4] = {km[j].x, km[j].y, km[j].z, km[j].w};
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                if (static_cast<int>(w4[e] << 16) <= 0) v[2 * e] = 0.f;
                if (static_cast<int>(w4[e] & 0xffff0000u) <= 0) v[2 * e + 1] = 0.f;
              }
            }
            if constexpr (kAdd) {
              if (add_post) {
                v[0] += bf16_lo(ka[j].x); v[1] += bf16_hi(ka[j].x); v[2] += bf16_lo(ka[j].y); v[3] += bf16_hi(ka[j].y);
                v[4] += bf16_lo(ka[j].z); v[5] += bf16_hi(ka[j].z); v[6] += bf16_lo(ka[j].w); v[7] += bf16_hi(ka[j].w);
              }
            }
            if (relu) {
#pragma unroll
              for (int e = 0; e < 8; ++e) v[e] = fmaxf(v[e], 0.f);
            }
            if (cj + 8 > col_end) {  // channel tail (Nrows not a multiple of 8): keep the padding zero
#pragma unroll
              for (int e = 0; e < 8; ++e)
                if (cj + e >= col_end) v[e] = 0.f;
            }
            const uint32_t o0 = pack_bf16x2(v[0], v[1]), o1 = pack_bf16x2(v[2], v[3]);
            const uint32_t o2 = pack_bf16x2(v[4], v[5]), o3 = pack_bf16x2(v[6], v[7]);
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(sA + poff[j]), "r"(o0), "r"(o1), "r"(o2), "r"(o3) : "memory");
          }
          fence_proxy_async();
          mbar_arrive(stfull_bar(cur.slot));
          ++k;
        }
        cur.advance(kPPC, nslots);
      }
      g += nch;
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
      if (ew == 0 && lane == 0 && tile == static_cast<int>(blockIdx.x)) B200_TR(8);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (pair) cluster_sync_all();          // my shared memory stays alive until the partner has stopped writing into it
  if (warp == 2) tmem_dealloc(tmem_base, kTmemCols);
#ifdef B200_CONV_TRACE
  if (tr && threadIdx.x == 0) {
    unsigned long long gt;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
    tr[11] = clock64();
    tr[12] = static_cast<long long>(gt);
  }
#endif
#undef B200_TR
}

// ---------------------------------------------------------------------------------------------
// host side

static int floordiv(int a, int b) { return (a >= 0) ? a / b : -((-a + b - 1) / b); }
static int posmod(int a, int b) { return ((a % b) + b) % b; }

static void pick_tile(int OH, int OW, int& th_log2, int& tw_log2, bool halo) {
  // 128 pixels as (128/tw) x tw; minimise padded tiles, prefer wide tiles on ties.  Halo mode loads th + 2 rows per
  // box: only 8 x 16 and 16 x 8 tiles keep the box small (<= 20 KB) and every row offset 1024-byte aligned.
  long long best = -1;
  for (int tl = halo ? 4 : 7; tl >= (halo ? 3 : 0); --tl) {
    const int tw = 1 << tl, th = 128 >> tl;
    const long long tiles = static_cast<long long>((OH + th - 1) / th) * ((OW + tw - 1) / tw);
    if (best < 0 || tiles < best) {
      best = tiles;
      tw_log2 = tl;
      th_log2 = 7 - tl;
    }
  }
}

// One N tile when the output channels fit 256 columns; otherwise 256-column tiles (a multiple of the 64-column
// store chunk, so a chunk never straddles two N tiles; the last tile's MMA is narrowed to its valid columns).
static int pick_bn(int nrows) { return nrows <= 256 ? (nrows + 15) / 16 * 16 : 256; }

template <bool kAdd, bool kMask>
static int launch_variant(const ConvParams& p, int grid, cudaStream_t stream) {
  static std::atomic<unsigned long long> attr_set{0};   // per device
  if (device_needs_setup(attr_set)) {
    B200_CUDA_OK(cudaFuncSetAttribute(conv_igemm_kernel<kAdd, kMask>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit));
    device_setup_done(attr_set);
  }
  if (p.pair) {
    // clusters of two need both SMs free in one GPC: ask once how many fit at the same time (every layout of this
    // kernel is one CTA per SM, so the answer does not depend on the shared-memory size)
    static std::atomic<int> fit{0};
    int clusters = fit.load();
    if (clusters == 0) {
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(static_cast<unsigned>(num_sms() & ~1));
      cfg.blockDim = dim3(kConvThreads);
      cfg.dynamicSmemBytes = static_cast<size_t>(p.smem_bytes);
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeClusterDimension;
      attr[0].val.clusterDim.x = 2;
      attr[0].val.clusterDim.y = 1;
      attr[0].val.clusterDim.z = 1;
      cfg.attrs = attr;
      cfg.numAttrs = 1;
      if (cudaOccupancyMaxActiveClusters(&clusters, conv_igemm_kernel<kAdd, kMask>, &cfg) != cudaSuccess || clusters < 1) {
        cudaGetLastError();
        clusters = num_sms() / 2 - 2;
      }
      fit.store(clusters);
    }
    if (grid > 2 * clusters) grid = 2 * clusters;
    launch_pdl_cluster(conv_igemm_kernel<kAdd, kMask>, grid, kConvThreads, static_cast<size_t>(p.smem_bytes), stream, 2u, p);
    return check_launch("conv_igemm_kernel (pairs)");
  }
  launch_pdl(conv_igemm_kernel<kAdd, kMask>, grid, kConvThreads, static_cast<size_t>(p.smem_bytes), stream, p);
  return check_launch("conv_igemm_kernel");
}

}  // namespace b200

static std::atomic<long long*> g_conv_trace{nullptr};

extern "C" void b200_conv_set_trace(void* buf) { g_conv_trace.store(static_cast<long long*>(buf)); }

extern "C" int b200_conv(const B200ConvArgs* a, void* stream_) {
  using namespace b200;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (!a || a->num_problems < 1 || a->num_problems > B200_MAX_PROBLEMS) return set_error(-1, "b200_conv: bad num_problems");
  if (a->stride != 1 && a->stride != 2) return set_error(-1, "b200_conv: stride must be 1 or 2");
  if (a->stride == 2 && a->num_problems != 1) return set_error(-1, "b200_conv: grouped launch needs stride 1");
  if (a->KH * a->KW > kMaxTaps) return set_error(-1, "b200_conv: too many taps");
  if ((a->ldi % 8) || (a->ldw % 8) || (a->ldo % 8)) return set_error(-1, "b200_conv: ld* must be multiples of 8");
  if (a->out_fp32) return set_error(-1, "b200_conv: fp32 output is not supported (bf16 activations only)");
  const bool has_add = a->prob[0].add != nullptr, has_mask = a->prob[0].mask != nullptr;
  for (int i = 1; i < a->num_problems; ++i)
    if ((a->prob[i].add != nullptr) != has_add || (a->prob[i].mask != nullptr) != has_mask)
      return set_error(-1, "b200_conv: all problems of a grouped launch must agree on add/mask presence");

  ConvParams p;
  memset(&p, 0, sizeof(p));
  p.trace = g_conv_trace.load();
  p.num_problems = a->num_problems;
  p.Nrows = a->Nrows;
  p.BN = pick_bn(a->Nrows);
  {
    // Small problems (batch 1: res4 has 33 pixel tiles, res5 nine) leave most SMs idle and the K loop of one wide tile
    // is the whole kernel: split the output channels over more CTAs (N tiles of 128 or 64 columns; activation tiles
    // are re-read from L2, which is cheap at this size).  B200_CONV_SPLIT_N=0 keeps 256-column tiles.
    const char* e = getenv("B200_CONV_SPLIT_N");
    if (!(e && e[0] == '0') && a->Nrows >= 128) {
      long long m_tiles = 0;
      for (int i = 0; i < a->num_problems; ++i) {
        const B200ConvProblem& q = a->prob[i];
        long long px = static_cast<long long>(q.N) * q.OHo * q.OWo;
        if (a->mode == 1 && a->stride == 2) px = (px + 3) / 4;          // one parity class per launch
        m_tiles += (px + kBM - 1) / kBM;
      }
      while (p.BN > 64 && m_tiles * ((a->Nrows + p.BN - 1) / p.BN) * 2 < num_sms()) p.BN = p.BN > 128 ? 128 : 64;
    }
  }
  p.num_n_tiles = (a->Nrows + p.BN - 1) / p.BN;
  p.num_kc = (a->K + kBK - 1) / kBK;
  p.shift = a->shift;
  for (int i = 0; i < a->num_problems; ++i) p.colsum[i] = a->prob[i].colsum;
  p.relu = a->relu;
  p.add_after_mask = a->add_after_mask;
  int os = 1, oy = 0, ox = 0;   // output scatter: (y*os + oy, x*os + ox)

  // ---- tap table
  const int s = a->stride;
  int nt = 0;
  if (a->mode == 0) {
    for (int r = 0; r < a->KH; ++r)
      for (int c = 0; c < a->KW; ++c) {
        const int eh = r - a->pad_h, ew = c - a->pad_w;
        Tap t;
        const int py = posmod(eh, s), px = posmod(ew, s);
        t.dh = static_cast<int8_t>(floordiv(eh, s));
        t.dw = static_cast<int8_t>(floordiv(ew, s));
        t.map = static_cast<int8_t>(py * s + px);
        t.pad = 0;
        t.widx = r * a->KW + c;
        p.taps[nt++] = t;
      }
  } else {
    if (s == 2) {
      os = 2;
      oy = a->par_y;
      ox = a->par_x;
    }
    for (int r = 0; r < a->KH; ++r)
      for (int c = 0; c < a->KW; ++c) {
        const int eh = (s == 2 ? a->par_y : 0) + a->pad_h - r;
        const int ew = (s == 2 ? a->par_x : 0) + a->pad_w - c;
        if (posmod(eh, s) != 0 || posmod(ew, s) != 0) continue;
        Tap t;
        t.dh = static_cast<int8_t>(floordiv(eh, s));
        t.dw = static_cast<int8_t>(floordiv(ew, s));
        t.map = 0;
        t.pad = 0;
        t.widx = r * a->KW + c;
        p.taps[nt++] = t;
      }
  }
  p.num_taps = nt;
  if (nt == 0) return set_error(-6, "b200_conv: empty tap set (caller should zero-fill this parity class)");
  // ---- K-loop mode (see ConvParams)
  static int mode_env = -1;
  if (mode_env < 0) {
    const char* e = getenv("B200_CONV_GROUPS");      // 0: plain boxes only (A/B arm for profiles)
    mode_env = (e && e[0] == '0') ? 0 : 1;
  }
  p.mode = 0;
  p.gsz = 1;
  if (mode_env && a->KH == 3 && a->KW == 3 && s == 1 && nt == 9) {
    p.mode = 1;
    for (int t = 0; t < 9; ++t) {
      const Tap& tp = p.taps[t];
      if (tp.dh < -1 || tp.dh > 1 || tp.dw < -1 || tp.dw > 1 || tp.map != 0) p.mode = 0;
    }
    if (p.mode == 1) {
      p.gsz = 3;
      for (int t = 0; t < 9; ++t) p.widx3[p.taps[t].dw + 1][p.taps[t].dh + 1] = p.taps[t].widx;
    }
  }
  if (mode_env && p.mode == 0 && (a->K % kBK) == 0 && p.num_kc >= 2) {
    p.mode = 2;
    p.gsz = p.num_kc < 4 ? p.num_kc : 4;
  }

  // ---- shared-memory split: activation ring | weight ring | epilogue chunk slots.
  // The weight ring holds every weight tile of the layer when they fit in 96 KB and there is one N tile (then they
  // are fetched once); otherwise 4 stages.  The chunk slots the epilogue needs come next (2 for plain outputs, 4 / 6
  // when residual / residual + mask chunks are loaded), and the activation ring gets the rest: it must cover the
  // ~1.5 us of a box that misses L2 with stage times as short as 130 cycles when N is small.
  const int ppc = (has_add && has_mask) ? 2 : 1;
  const int num_substeps = nt * p.num_kc;
  p.b_stage_bytes = p.BN * kBK * 2;
  const int fixed = kShiftBytes + kBarBytes;
  const int want_slots = (has_add || has_mask) ? 2 + 2 * ppc : 2;
  int b_st = 4;
  const bool resident = p.num_n_tiles == 1 && num_substeps <= kMaxBStages && num_substeps * p.b_stage_bytes <= 98304;
  if (resident) b_st = num_substeps;
  // kc boxes pay off where the step time is short (small layers, resident weights); with N = 256 and streamed
  // weights, 4 fine-grained activation stages hide the TMA latency better than 2 coarse ones (measured)
  if (p.mode == 2 && !resident) {
    p.mode = 0;
    p.gsz = 1;
  }
  int a_st = 0;
  for (;;) {
    p.a_stage_bytes = p.mode == 1 ? 20480 : p.gsz * kAStageBytes;
    const int room = kSmemLimit - fixed - b_st * p.b_stage_bytes - want_slots * kSlotBytes;
    a_st = room / p.a_stage_bytes;
    if (a_st > kMaxAStages) a_st = kMaxAStages;
    if (a_st >= (p.mode == 2 ? 3 : 2)) break;
    if (p.mode == 2) {                          // smaller activation boxes first
      p.gsz = p.gsz > 2 ? 2 : 1;
      if (p.gsz == 1) p.mode = 0;
      continue;
    }
    if (b_st > 2) {                             // then a shallower weight ring (gives up residency)
      b_st = b_st > 4 ? 4 : b_st - 1;
      continue;
    }
    return set_error(-1, "b200_conv: shared-memory layout does not fit");
  }
  p.a_stages = a_st;
  p.b_stages = b_st;
  {
    int sl = (kSmemLimit - fixed - a_st * p.a_stage_bytes - b_st * p.b_stage_bytes) / kSlotBytes;
    if (sl > kMaxSlots) sl = kMaxSlots;
    if (ppc == 2) sl &= ~1;
    p.slots = sl;
    if (sl < 2 * ppc) return set_error(-1, "b200_conv: shared-memory layout does not fit (slots)");
    const int need = a_st * p.a_stage_bytes + b_st * p.b_stage_bytes + sl * kSlotBytes + fixed;
    p.smem_bytes = need + 1024 <= kSmemLimit ? need + 1024 : kSmemLimit;   // the base is 1024-aligned in practice
  }
  p.tap_groups = p.mode == 1 ? 3 : nt;
  p.kgroups = p.mode == 2 ? (p.num_kc + p.gsz - 1) / p.gsz : p.num_kc;

  // ---- problems + tensor maps
  int tile_begin = 0;
  int map_count = 0;
  const unsigned long long c8 = static_cast<unsigned long long>((a->Nrows + 7) / 8 * 8);   // stored channels
  for (int i = 0; i < a->num_problems; ++i) {
    B200ConvProblem q = a->prob[i];
    // 1x1 / stride 1: pixels are independent, so the problem is a flat list of N*H*W pixels and a tile is any 128
    // consecutive ones -- no partly filled tiles at row ends or image ends (8x50x84: 263 tiles instead of 312, and two
    // waves on 148 SMs instead of three; 25x42 maps were 46 % padding)
    if (a->KH == 1 && a->KW == 1 && s == 1 && a->pad_h == 0 && a->pad_w == 0 && q.OHo == q.H && q.OWo == q.W) {
      const long long px = static_cast<long long>(q.N) * q.H * q.W;
      if (px < (1LL << 31)) {
        q.N = 1;
        q.H = q.OHo = 1;
        q.W = q.OWo = static_cast<int>(px);
      }
    }
    DevProblem& d = p.prob[i];
    int OH, OW;   // GEMM pixel grid
    if (a->mode == 1 && s == 2) {
      OH = (q.OHo - a->par_y + 1) / 2;
      OW = (q.OWo - a->par_x + 1) / 2;
    } else {
      OH = q.OHo;
      OW = q.OWo;
    }
    if (OH <= 0 || OW <= 0) return set_error(-1, "b200_conv: empty problem");
    d.OH = OH;
    d.OW = OW;
    pick_tile(OH, OW, d.th_log2, d.tw_log2, p.mode == 1);
    const int th = 1 << d.th_log2, tw = 1 << d.tw_log2;
    d.a_bytes = p.mode == 1 ? (th + 2) * tw * kBK * 2 : p.gsz * kAStageBytes;
    d.sub_stride = p.mode == 1 ? tw * kBK * 2 : kAStageBytes;
    d.tiles_h = (OH + th - 1) / th;
    d.tiles_w = (OW + tw - 1) / tw;
    d.tile_begin = tile_begin;
    tile_begin += q.N * d.tiles_h * d.tiles_w;
    d.map_base = map_count;
    const unsigned int box[5] = {static_cast<unsigned>(kBK), static_cast<unsigned>(tw),
                                 static_cast<unsigned>(p.mode == 1 ? th + 2 : th), 1u, static_cast<unsigned>(p.gsz)};
    const int nmaps = (a->mode == 0 && s == 2) ? 4 : 1;
    if (map_count + nmaps > kMaxMaps) return set_error(-1, "b200_conv: too many tensor maps");
    for (int mi = 0; mi < nmaps; ++mi) {
      const int py = mi / 2, px = mi % 2;
      const int step = (nmaps == 4) ? 2 : 1;
      const long long Wd = (q.W - px + step - 1) / step;
      const long long Hd = (q.H - py + step - 1) / step;
      const __nv_bfloat16* base = static_cast<const __nv_bfloat16*>(q.in) + (static_cast<long long>(py) * q.W + px) * a->ldi;
      // kc mode: the channel axis is split [k chunk][64] so that one 5-D box carries gsz chunks
      const bool kcm = p.mode == 2;
      unsigned long long dims[5] = {static_cast<unsigned long long>(kcm ? kBK : a->K),
                                    static_cast<unsigned long long>(Wd > 0 ? Wd : 1),
                                    static_cast<unsigned long long>(Hd > 0 ? Hd : 1), static_cast<unsigned long long>(q.N),
                                    static_cast<unsigned long long>(p.num_kc)};
      unsigned long long strides[4] = {static_cast<unsigned long long>(step) * a->ldi * 2ull,
                                       static_cast<unsigned long long>(step) * q.W * a->ldi * 2ull,
                                       static_cast<unsigned long long>(q.H) * q.W * a->ldi * 2ull, kBK * 2ull};
      int rc = make_tmap_bf16(&p.in_maps[map_count + mi], base, kcm ? 5 : 4, dims, strides, box);
      if (rc) return rc;
    }
    map_count += nmaps;
    // output-geometry maps (out / residual / mask): 64-channel x tw x th boxes over the GEMM pixel grid, which for a
    // stride-2 dgrad parity class is a strided view of the output tensor
    {
      const long long off = (static_cast<long long>(oy) * q.OWo + ox) * a->ldo;
      unsigned long long dims[4] = {c8, static_cast<unsigned long long>(OW), static_cast<unsigned long long>(OH),
                                    static_cast<unsigned long long>(q.N)};
      unsigned long long strides[3] = {static_cast<unsigned long long>(os) * a->ldo * 2ull,
                                       static_cast<unsigned long long>(os) * q.OWo * a->ldo * 2ull,
                                       static_cast<unsigned long long>(q.OHo) * q.OWo * a->ldo * 2ull};
      const unsigned int obox[4] = {64u, static_cast<unsigned>(tw), static_cast<unsigned>(th), 1u};
      int rc = make_tmap_bf16(&p.out_maps[i], static_cast<const __nv_bfloat16*>(q.out) + off, 4, dims, strides, obox);
      if (rc) return rc;
      if (has_add) {
        rc = make_tmap_bf16(&p.add_maps[i], static_cast<const __nv_bfloat16*>(q.add) + off, 4, dims, strides, obox);
        if (rc) return rc;
      }
      if (has_mask) {
        rc = make_tmap_bf16(&p.mask_maps[i], static_cast<const __nv_bfloat16*>(q.mask) + off, 4, dims, strides, obox);
        if (rc) return rc;
      }
    }
  }
  p.num_m_tiles = p.real_m_tiles = tile_begin;
  static int pair_env = -1;
  if (pair_env < 0) {
    const char* e = getenv("B200_CONV_PAIRS");       // 0: every CTA streams its own weights (A/B arm for profiles)
    pair_env = (e && e[0] == '0') ? 0 : 1;
  }
  const bool streamed = !(p.num_n_tiles == 1 && b_st == num_substeps);
  // The two CTAs of a pair advance through the weight ring in lock step, so a stall of one stalls both; that coupling
  // costs ~8 % on launches of two or three waves (measured on the res4 3x3 layers) and nothing on long ones.
  static int pair_min_waves = -1;
  if (pair_min_waves < 0) {
    const char* e = getenv("B200_CONV_PAIR_MIN_WAVES");
    pair_min_waves = e ? atoi(e) : 4;
  }
  const long long tiles_all = static_cast<long long>(p.real_m_tiles) * p.num_n_tiles;
  p.pair = (pair_env && streamed && (p.BN % 16) == 0 && num_sms() >= 4 &&
            tiles_all >= static_cast<long long>(pair_min_waves) * num_sms()) ? 1 : 0;
  if (p.pair) {
    // TMA multicast loads here read only this GPU's memory: the weights must be a device allocation of the current
    // device (a peer mapping, an IPC mapping or managed memory falls back to per-CTA weight streams)
    cudaPointerAttributes at;
    int cur = -1;
    if (cudaPointerGetAttributes(&at, a->weight) != cudaSuccess || cudaGetDevice(&cur) != cudaSuccess ||
        at.type != cudaMemoryTypeDevice || at.device != cur) {
      cudaGetLastError();
      p.pair = 0;
    }
  }
  if (p.pair) p.num_m_tiles = (p.real_m_tiles + 1) & ~1;
  {
    unsigned long long dims[3] = {static_cast<unsigned long long>(a->K), static_cast<unsigned long long>(a->Nrows),
                                  static_cast<unsigned long long>(a->KH * a->KW)};
    unsigned long long strides[2] = {static_cast<unsigned long long>(a->ldw) * 2ull,
                                     static_cast<unsigned long long>(a->Nrows) * a->ldw * 2ull};
    const unsigned int box[3] = {static_cast<unsigned>(kBK), static_cast<unsigned>(p.pair ? p.BN / 2 : p.BN), 1u};
    int rc = make_tmap_bf16(&p.w_map, a->weight, 3, dims, strides, box);
    if (rc) return rc;
  }
  // unused map slots must still be valid descriptors for the prefetch: replicate slot 0
  for (int i = map_count; i < kMaxMaps; ++i) p.in_maps[i] = p.in_maps[0];

  // ---- shared-memory split between the operand ring and the epilogue chunk slots.
  // Short K loops (the memory-bound 1x1 layers) need few ring stages -- the weights stay resident and one stage is one
  // whole tile -- and give the rest to chunk slots = bytes of residual / mask / result in flight per SM.
  const long long total = static_cast<long long>(p.num_m_tiles) * p.num_n_tiles;      // even in pair mode
  const int sms = p.pair ? (num_sms() & ~1) : num_sms();
  const int grid = static_cast<int>(total < sms ? total : sms);
  if (has_add && has_mask) return launch_variant<true, true>(p, grid, stream);
  if (has_add) return launch_variant<true, false>(p, grid, stream);
  if (has_mask) return launch_variant<false, true>(p, grid, stream);
  return launch_variant<false, false>(p, grid, stream);
}
