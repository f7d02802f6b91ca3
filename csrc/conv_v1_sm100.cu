// FIRST-GENERATION conv kernel (register/LSU epilogue), kept behind B200_CONV_IMPL=1 as the A/B arm for the
// TMA-epilogue kernel in conv_sm100.cu.
// Implicit-GEMM convolution (fprop and dgrad) for sm_100a: TMA + tcgen05.mma + TMEM.
//
//   out[n, y, x, :] = epilogue( sum_{tap} sum_{k} in[n, y + dh(tap), x + dw(tap), k] * w[tap][:, k] )
//
// GEMM view: M = output pixels (tiles of th x tw = 128 pixels of ONE image, so a tap is a pure
// (dh, dw) shift of the TMA box and the conv halo / padding is TMA's out-of-bounds zero fill),
// N = output channels (BN <= 256 per tile), K = taps x input channels (64 per pipeline stage).
//
// Warp roles (640 threads, 1 CTA per SM, persistent over tiles):
//   warp 0   TMA producer  (one lane): A box [64ch, tw, th, 1] + B box [64ch, BN rows, 1 tap]
//   warp 1   MMA issuer    (one lane): 4 x tcgen05.mma (K=16) per stage, commit -> empty barrier
//   warp 2   TMEM allocator (512 columns = two 256-column accumulators, double-buffered)
//   warps 4-19 epilogue (4 per TMEM lane quarter): tcgen05.ld 32 lanes x 32 columns, fused
//            shift/add/mask/ReLU, swizzled smem transpose, coalesced bf16 loads/stores
//
// Stride-2 convolutions use one tensor map per input parity class (strided views), so every
// tap is still a plain shifted box.  Stride-2 dgrad computes one output parity class per launch.
//
// Replaces cuDNN fprop/dgrad + separate bias/BN/ReLU/add kernels of the reference stack
// (SURVEY.md G2-G4, G6, G8, G12); the reference itself has no CUDA source.
#include <cstdlib>

#include "../include/b200_ops.h"
#include "common.cuh"
#include "ptx_sm100.cuh"

namespace b200 {
namespace v1 {

constexpr int kStages = 4;
constexpr int kBM = 128;
constexpr int kBK = 64;
constexpr int kAStageBytes = kBM * kBK * 2;       // 16 KB
constexpr int kBStageBytes = 256 * kBK * 2;       // 32 KB (BN <= 256)
constexpr int kMaxTaps = 49;
constexpr int kMaxMaps = 8;
constexpr int kTmemCols = 512;
constexpr int kEpiWarps = 16;                        // 4 warps per TMEM lane quarter (column slots)
constexpr int kConvThreads = 128 + 32 * kEpiWarps;
constexpr int kStagingBytes = kEpiWarps * 2048;      // per epilogue warp: one 32 rows x 64 B buffer
constexpr int kSmemBytes = kStages * (kAStageBytes + kBStageBytes) + kStagingBytes + 1024 /*align*/ + 256 /*barriers*/;

struct Tap {
  int8_t dh, dw, map, pad;
  int widx;
};

struct DevProblem {
  void* out;
  const void* add;
  const void* mask;
  int N, OH, OW;        // GEMM pixel grid
  int OHo, OWo;         // output tensor spatial dims
  int tiles_h, tiles_w; // spatial tiles per image
  int tile_begin;       // first m-tile of this problem
  int th_log2, tw_log2;
  int map_base;
};

struct ConvParams {
  CUtensorMap in_maps[kMaxMaps];
  CUtensorMap w_map;
  DevProblem prob[B200_MAX_PROBLEMS];
  Tap taps[kMaxTaps];
  const float* shift;
  int num_problems, num_m_tiles, num_n_tiles;
  int Nrows, ldo, BN;
  int num_taps, num_kc;
  int os, oy, ox;       // output scatter: (y*os + oy, x*os + ox)
  int relu, out_fp32, add_after_mask;
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

template <bool kAdd, bool kMask>
__global__ void __launch_bounds__(kConvThreads, 1) conv_igemm_kernel(const __grid_constant__ ConvParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t a_smem = smem_base;
  const uint32_t b_smem = smem_base + kStages * kAStageBytes;
  const uint32_t stg_base = b_smem + kStages * kBStageBytes;
  const uint32_t bar_base = stg_base + kStagingBytes;
  // barrier layout (8 bytes each): full[kStages], empty[kStages], tmem_full[2], tmem_empty[2], tmem_ptr
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kStages + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * kStages + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * kStages + 2 + a); };
  const uint32_t tmem_ptr_smem = bar_base + 8u * (2 * kStages + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int total_tiles = p.num_m_tiles * p.num_n_tiles;
  const int num_k = p.num_taps * p.num_kc;
  const uint32_t b_bytes = static_cast<uint32_t>(p.BN) * kBK * 2;

  if (warp == 0 && lane == 0) {
    for (int i = 0; i < kMaxMaps; ++i) tma_prefetch_desc(&p.in_maps[i]);
    tma_prefetch_desc(&p.w_map);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull_bar(a), 1);
      mbar_init(tempty_bar(a), 32 * kEpiWarps);
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_ptr_smem, kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_ptr_smem));
  // set-up done: let the next kernel start its own set-up, then wait for the previous kernel's results
  pdl_launch_dependents();
  pdl_wait();

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    // The whole warp runs the loop; lane 0 arms the barrier and issues the activation box, lane 1 issues
    // the weight box, so the two TMA issues of a stage overlap (a single issuing thread was measurably
    // the bottleneck in the weight-gradient kernel).
    int stage = 0;
    uint32_t phase = 0;
    bool first_pass = true;
    // Weights stay resident: when the whole K loop fits the ring (num_k divides kStages) and there is a
    // single N tile, stage s always holds the same weight tile, so it is fetched once.
    const bool weights_resident = (p.num_n_tiles == 1) && (kStages % num_k == 0);
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      int pi, n, y0, x0, n_tile;
      decode_tile(p, tile, pi, n, y0, x0, n_tile);
      const int map_base = p.prob[pi].map_base;
      for (int t = 0; t < p.num_taps; ++t) {
        const Tap tap = p.taps[t];
        const void* amap = &p.in_maps[map_base + tap.map];
        for (int kc = 0; kc < p.num_kc; ++kc) {
          mbar_wait(empty_bar(stage), phase ^ 1);
          const bool load_b = !weights_resident || first_pass;
          if (lane == 0) {
            mbar_expect_tx(full_bar(stage), kAStageBytes + (load_b ? b_bytes : 0u));
            tma_load_4d(a_smem + stage * kAStageBytes, amap, full_bar(stage), kc * kBK, x0 + tap.dw, y0 + tap.dh, n);
          }
          __syncwarp();
          if (lane == 1 && load_b)
            tma_load_3d(b_smem + stage * kBStageBytes, &p.w_map, full_bar(stage), kc * kBK, n_tile * p.BN, tap.widx);
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
            first_pass = false;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    const uint32_t idesc = make_idesc_bf16(kBM, p.BN, 0, 0);
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      mbar_wait(tempty_bar(acc), acc_phase ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + acc * 256;
      for (int kb = 0; kb < num_k; ++kb) {
        mbar_wait(full_bar(stage), phase);
        tc_fence_after();
        if (lane == 0) {
          const uint64_t adesc = make_smem_desc(a_smem + stage * kAStageBytes, 16, 1024);
          const uint64_t bdesc = make_smem_desc(b_smem + stage * kBStageBytes, 16, 1024);
#pragma unroll
          for (int k = 0; k < kBK / 16; ++k) {
            // advance 16 bf16 = 32 bytes along K inside the 128B swizzle row: +2 in (addr >> 4) units
            umma_bf16(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
          }
          umma_commit(empty_bar(stage));
          if (kb == num_k - 1) umma_commit(tfull_bar(acc));
        }
        __syncwarp();
        if (++stage == kStages) {
          stage = 0;
          phase ^= 1;
        }
      }
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
  } else if (warp >= 4) {
    // ------------------------------------------------------------------ epilogue
    // 16 warps: warp (quarter q, column slot cq) owns TMEM lanes 32q..32q+31 and the 32-column chunks
    // cq, cq+4, ... of the tile.  A thread holds one accumulator row, but row-per-thread global
    // accesses would touch 32 lines per instruction with partial sectors, so every chunk goes through a
    // warp-private XOR-swizzled 2 KB smem buffer: global <-> smem moves are 8 rows x 64 B per
    // instruction (sector-complete), smem <-> registers are row-per-thread.  The kernel is latency
    // bound here for small K (ncu: long_scoreboard), hence 4 warps per scheduler and the mask/residual
    // loads of a chunk issued back to back before anything waits on them.
    const int ew = warp - 4;
    const int quarter = ew & 3;   // TMEM lanes this warp may read: 32 * (warp_id % 4)
    const int cq = ew >> 2;       // column slot
    const int m = quarter * 32 + lane;
    const uint32_t buf = stg_base + ew * 2048;
    const int crow = lane >> 2, cpiece = lane & 3;
    const uint32_t own_sw = (lane >> 1) & 3;
    const uint32_t own_base = buf + lane * 64;
    uint32_t own_off[4], so_r[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) own_off[j] = own_base + ((static_cast<uint32_t>(j) ^ own_sw) << 4);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int rr = i * 8 + crow;
      so_r[i] = buf + rr * 64 + ((cpiece ^ ((rr >> 1) & 3)) << 4);
    }
    const float* __restrict__ shiftp = p.shift;
    const bool relu = p.relu != 0, add_post = p.add_after_mask != 0;
    const int ldo = p.ldo, BN = p.BN, Nrows = p.Nrows;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      int pi, n, y0, x0, n_tile;
      decode_tile(p, tile, pi, n, y0, x0, n_tile);
      const DevProblem& q = p.prob[pi];
      const int y = y0 + (m >> q.tw_log2);
      const int x = x0 + (m & ((1 << q.tw_log2) - 1));
      const bool valid = (y < q.OH) && (x < q.OW);
      const int pix = (n * q.OHo + (y * p.os + p.oy)) * q.OWo + (x * p.os + p.ox);  // < 2^31 pixels
      const uint32_t valid_mask = __ballot_sync(0xffffffffu, valid);
      const int col0 = n_tile * BN;
      const int col_end = min(Nrows, col0 + BN);
      const int col_end_pad = (col_end + 7) & ~7;
      __nv_bfloat16* __restrict__ outp = static_cast<__nv_bfloat16*>(q.out);
      const __nv_bfloat16* __restrict__ addp = static_cast<const __nv_bfloat16*>(q.add);
      const __nv_bfloat16* __restrict__ maskp = static_cast<const __nv_bfloat16*>(q.mask);
      long long roff_r[4];  // element offset of (row handled in the coalesced pattern, this lane's 16-byte piece)
      bool ok_r[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int rr = i * 8 + crow;
        roff_r[i] = static_cast<long long>(__shfl_sync(0xffffffffu, pix, rr)) * ldo + col0 + cpiece * 8;
        ok_r[i] = (valid_mask >> rr) & 1u;
      }

      // A warp owns at most two 32-column chunks of a tile (c0 and c0 + 128).  Their epilogue operands
      // (residual / ReLU mask: the 64 contiguous bytes of this thread's row, 4 x 16 B, every sector fully
      // consumed) and the shift row (one coalesced float per lane, broadcast later with shuffles) are all
      // requested BEFORE waiting for the accumulator: the epilogue is bound by DRAM latency x bytes in
      // flight, so everything a tile needs is put in flight at once, behind the MMA.
      constexpr int kChunkStep = 32 * (kEpiWarps / 4);
      const int c0 = cq * 32;
      // Operands are fetched in the same COALESCED pattern as the stores (8 rows x 64 B per instruction, lane ->
      // row i*8 + lane/4, 16-byte piece lane%4) and transposed to row-per-thread through the staging buffer.
      // Row-per-thread loads touch 32 different 128-byte lines per instruction; the L1 tag stage then caps the
      // operand stream at ~3.2 TB/s (scripts/epilogue_probe.py), well below what the TMA-fed operands reach.
      uint4 ka[2][4], km[2][4];
      float sh[2] = {0.f, 0.f};
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int c = c0 + h * kChunkStep;
        const bool chunk_ok = c < BN && col0 + c < col_end;
        const bool pc_ok = chunk_ok && (col0 + c + cpiece * 8 < col_end_pad);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const bool ok = pc_ok && ok_r[i];
          if constexpr (kMask && !kAdd) km[h][i] = ok ? *reinterpret_cast<const uint4*>(maskp + roff_r[i] + c) : make_uint4(0u, 0u, 0u, 0u);
          if constexpr (kAdd && !kMask) ka[h][i] = ok ? *reinterpret_cast<const uint4*>(addp + roff_r[i] + c) : make_uint4(0u, 0u, 0u, 0u);
          if constexpr (kAdd && kMask) {
            if (h == 0) {  // both operands: only the first chunk up front (register budget)
              km[0][i] = ok ? *reinterpret_cast<const uint4*>(maskp + roff_r[i] + c) : make_uint4(0u, 0u, 0u, 0u);
              ka[0][i] = ok ? *reinterpret_cast<const uint4*>(addp + roff_r[i] + c) : make_uint4(0u, 0u, 0u, 0u);
            }
          }
        }
        if (shiftp && chunk_ok && col0 + c + lane < col_end_pad) sh[h] = __ldg(shiftp + col0 + c + lane);
      }
      // coalesced layout -> row-per-thread layout (in place in the registers), through the warp's staging buffer
      auto to_rows = [&](uint4(&v)[4]) {
#pragma unroll
        for (int i = 0; i < 4; ++i)
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(so_r[i]), "r"(v[i].x), "r"(v[i].y), "r"(v[i].z), "r"(v[i].w) : "memory");
        __syncwarp();
#pragma unroll
        for (int j = 0; j < 4; ++j)
          asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v[j].x), "=r"(v[j].y), "=r"(v[j].z), "=r"(v[j].w) : "r"(own_off[j]) : "memory");
        __syncwarp();
      };
      mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + acc * 256;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int c = c0 + h * kChunkStep;
        const int col = col0 + c;
        if (c >= BN || col >= col_end) break;
        const bool piece_ok = col + cpiece * 8 < col_end_pad;
        if constexpr (kAdd && kMask) {
          if (h == 1) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const bool ok = piece_ok && ok_r[i];
              km[1][i] = ok ? *reinterpret_cast<const uint4*>(maskp + roff_r[i] + c) : make_uint4(0u, 0u, 0u, 0u);
              ka[1][i] = ok ? *reinterpret_cast<const uint4*>(addp + roff_r[i] + c) : make_uint4(0u, 0u, 0u, 0u);
            }
          }
        }
        uint32_t r[32];
        tmem_ld_32x32(t_row + c, r);
        if constexpr (kAdd) to_rows(ka[h]);     // the TMEM load is in flight meanwhile
        if constexpr (kMask) to_rows(km[h]);
        tmem_ld_wait();
        // ---- row-per-thread math, result into the staging buffer (issue-bound: keep it short)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int cj = col + j * 8;
          float v[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(r[j * 8 + e]);
          if (shiftp) {
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] += __shfl_sync(0xffffffffu, sh[h], j * 8 + e);
          }
          if (valid && cj < col_end) {
            if constexpr (kAdd) {
              if (!add_post) {
                v[0] += bf16_lo(ka[h][j].x); v[1] += bf16_hi(ka[h][j].x); v[2] += bf16_lo(ka[h][j].y); v[3] += bf16_hi(ka[h][j].y);
                v[4] += bf16_lo(ka[h][j].z); v[5] += bf16_hi(ka[h][j].z); v[6] += bf16_lo(ka[h][j].w); v[7] += bf16_hi(ka[h][j].w);
              }
            }
            if constexpr (kMask) {
              // bf16 > 0  <=>  its 16-bit pattern, read as a signed integer, is > 0
              const uint32_t w4[4] = {km[h][j].x, km[h][j].y, km[h][j].z, km[h][j].w};
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                if (static_cast<int>(w4[e] << 16) <= 0) v[2 * e] = 0.f;
                if (static_cast<int>(w4[e] & 0xffff0000u) <= 0) v[2 * e + 1] = 0.f;
              }
            }
            if constexpr (kAdd) {
              if (add_post) {
                v[0] += bf16_lo(ka[h][j].x); v[1] += bf16_hi(ka[h][j].x); v[2] += bf16_lo(ka[h][j].y); v[3] += bf16_hi(ka[h][j].y);
                v[4] += bf16_lo(ka[h][j].z); v[5] += bf16_hi(ka[h][j].z); v[6] += bf16_lo(ka[h][j].w); v[7] += bf16_hi(ka[h][j].w);
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
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(own_off[j]), "r"(o0), "r"(o1), "r"(o2), "r"(o3) : "memory");
          }
        }
        __syncwarp();
        // ---- coalesced smem -> global store
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          if (ok_r[i] && piece_ok) {
            uint4 t;
            asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(t.x), "=r"(t.y), "=r"(t.z), "=r"(t.w) : "r"(so_r[i]));
            *reinterpret_cast<uint4*>(outp + roff_r[i] + c) = t;
          }
        }
        __syncwarp();
      }
      tc_fence_before();
      mbar_arrive(tempty_bar(acc));
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, kTmemCols);
}

// ---------------------------------------------------------------------------------------------
// host side

static int floordiv(int a, int b) { return (a >= 0) ? a / b : -((-a + b - 1) / b); }
static int posmod(int a, int b) { return ((a % b) + b) % b; }

static void pick_tile(int OH, int OW, int& th_log2, int& tw_log2) {
  // 128 pixels as (128/tw) x tw; minimise padded tiles, prefer wide tiles on ties.
  long long best = -1;
  for (int tl = 7; tl >= 0; --tl) {
    const int tw = 1 << tl, th = 128 >> tl;
    const long long tiles = static_cast<long long>((OH + th - 1) / th) * ((OW + tw - 1) / tw);
    if (best < 0 || tiles < best) {
      best = tiles;
      tw_log2 = tl;
      th_log2 = 7 - tl;
    }
  }
}

static int pick_bn(int nrows) {
  if (nrows <= 256) return (nrows + 15) / 16 * 16;
  // largest multiple of 16 <= 256 that minimises padded work
  int best = 256;
  long long best_cost = -1;
  for (int bn = 256; bn >= 128; bn -= 16) {
    const long long tiles = (nrows + bn - 1) / bn;
    const long long cost = tiles * bn;
    if (best_cost < 0 || cost < best_cost) {
      best_cost = cost;
      best = bn;
    }
  }
  return best;
}

}  // namespace v1
}  // namespace b200

extern "C" int b200_conv_v1(const B200ConvArgs* a, void* stream_) {
  using namespace b200;
  using namespace b200::v1;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (!a || a->num_problems < 1 || a->num_problems > B200_MAX_PROBLEMS) return set_error(-1, "b200_conv: bad num_problems");
  if (a->stride != 1 && a->stride != 2) return set_error(-1, "b200_conv: stride must be 1 or 2");
  if (a->stride == 2 && a->num_problems != 1) return set_error(-1, "b200_conv: grouped launch needs stride 1");
  if (a->KH * a->KW > kMaxTaps) return set_error(-1, "b200_conv: too many taps");
  if ((a->ldi % 8) || (a->ldw % 8) || (a->ldo % 8)) return set_error(-1, "b200_conv: ld* must be multiples of 8");

  if (a->out_fp32) return set_error(-1, "b200_conv: fp32 output is not supported (bf16 activations only)");
  static std::atomic<unsigned long long> attr_set{0};   // per device
  if (device_needs_setup(attr_set)) {
    B200_CUDA_OK(cudaFuncSetAttribute(conv_igemm_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    B200_CUDA_OK(cudaFuncSetAttribute(conv_igemm_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    B200_CUDA_OK(cudaFuncSetAttribute(conv_igemm_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    B200_CUDA_OK(cudaFuncSetAttribute(conv_igemm_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    device_setup_done(attr_set);
  }
  const bool has_add = a->prob[0].add != nullptr, has_mask = a->prob[0].mask != nullptr;
  for (int i = 1; i < a->num_problems; ++i)
    if ((a->prob[i].add != nullptr) != has_add || (a->prob[i].mask != nullptr) != has_mask)
      return set_error(-1, "b200_conv: all problems of a grouped launch must agree on add/mask presence");

  ConvParams p;
  memset(&p, 0, sizeof(p));
  p.num_problems = a->num_problems;
  p.Nrows = a->Nrows;
  p.ldo = a->ldo;
  p.BN = pick_bn(a->Nrows);
  p.num_n_tiles = (a->Nrows + p.BN - 1) / p.BN;
  p.num_kc = (a->K + kBK - 1) / kBK;
  p.shift = a->shift;
  p.relu = a->relu;
  p.out_fp32 = a->out_fp32;
  p.add_after_mask = a->add_after_mask;
  p.os = 1;
  p.oy = p.ox = 0;

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
      p.os = 2;
      p.oy = a->par_y;
      p.ox = a->par_x;
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

  // ---- problems + tensor maps
  int tile_begin = 0;
  int map_count = 0;
  for (int i = 0; i < a->num_problems; ++i) {
    const B200ConvProblem& q = a->prob[i];
    DevProblem& d = p.prob[i];
    d.out = q.out;
    d.add = q.add;
    d.mask = q.mask;
    d.N = q.N;
    d.OHo = q.OHo;
    d.OWo = q.OWo;
    if (a->mode == 1 && s == 2) {
      d.OH = (q.OHo - a->par_y + 1) / 2;
      d.OW = (q.OWo - a->par_x + 1) / 2;
    } else {
      d.OH = q.OHo;
      d.OW = q.OWo;
    }
    if (d.OH <= 0 || d.OW <= 0) return set_error(-1, "b200_conv: empty problem");
    pick_tile(d.OH, d.OW, d.th_log2, d.tw_log2);
    const int th = 1 << d.th_log2, tw = 1 << d.tw_log2;
    d.tiles_h = (d.OH + th - 1) / th;
    d.tiles_w = (d.OW + tw - 1) / tw;
    d.tile_begin = tile_begin;
    tile_begin += q.N * d.tiles_h * d.tiles_w;
    d.map_base = map_count;
    const unsigned int box[4] = {static_cast<unsigned>(kBK), static_cast<unsigned>(tw), static_cast<unsigned>(th), 1u};
    const int nmaps = (a->mode == 0 && s == 2) ? 4 : 1;
    if (map_count + nmaps > kMaxMaps) return set_error(-1, "b200_conv: too many tensor maps");
    for (int mi = 0; mi < nmaps; ++mi) {
      const int py = mi / 2, px = mi % 2;
      const int step = (nmaps == 4) ? 2 : 1;
      const long long Wd = (q.W - px + step - 1) / step;
      const long long Hd = (q.H - py + step - 1) / step;
      const __nv_bfloat16* base = static_cast<const __nv_bfloat16*>(q.in) + (static_cast<long long>(py) * q.W + px) * a->ldi;
      unsigned long long dims[4] = {static_cast<unsigned long long>(a->K), static_cast<unsigned long long>(Wd > 0 ? Wd : 1),
                                    static_cast<unsigned long long>(Hd > 0 ? Hd : 1), static_cast<unsigned long long>(q.N)};
      unsigned long long strides[3] = {static_cast<unsigned long long>(step) * a->ldi * 2ull,
                                       static_cast<unsigned long long>(step) * q.W * a->ldi * 2ull,
                                       static_cast<unsigned long long>(q.H) * q.W * a->ldi * 2ull};
      int rc = make_tmap_bf16(&p.in_maps[map_count + mi], base, 4, dims, strides, box);
      if (rc) return rc;
    }
    map_count += nmaps;
  }
  p.num_m_tiles = tile_begin;
  {
    unsigned long long dims[3] = {static_cast<unsigned long long>(a->K), static_cast<unsigned long long>(a->Nrows),
                                  static_cast<unsigned long long>(a->KH * a->KW)};
    unsigned long long strides[2] = {static_cast<unsigned long long>(a->ldw) * 2ull,
                                     static_cast<unsigned long long>(a->Nrows) * a->ldw * 2ull};
    const unsigned int box[3] = {static_cast<unsigned>(kBK), static_cast<unsigned>(p.BN), 1u};
    int rc = make_tmap_bf16(&p.w_map, a->weight, 3, dims, strides, box);
    if (rc) return rc;
  }
  // unused map slots must still be valid descriptors for the prefetch: replicate slot 0
  for (int i = map_count; i < kMaxMaps; ++i) p.in_maps[i] = p.in_maps[0];

  const long long total = static_cast<long long>(p.num_m_tiles) * p.num_n_tiles;
  const int grid = static_cast<int>(total < num_sms() ? total : num_sms());
  if (has_add && has_mask) launch_pdl(conv_igemm_kernel<true, true>, grid, kConvThreads, kSmemBytes, stream, p);
  else if (has_add) launch_pdl(conv_igemm_kernel<true, false>, grid, kConvThreads, kSmemBytes, stream, p);
  else if (has_mask) launch_pdl(conv_igemm_kernel<false, true>, grid, kConvThreads, kSmemBytes, stream, p);
  else launch_pdl(conv_igemm_kernel<false, false>, grid, kConvThreads, kSmemBytes, stream, p);
  return check_launch("conv_igemm_kernel");
}
