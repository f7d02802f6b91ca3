// Convolution weight gradient for sm_100a: TMA + tcgen05.mma with MN-major operands.
//
//   dW[tap][co][ci] += rowscale[co] * sum_{n,y,x} dY[n, y, x, co] * X[n, y*s + dh(tap), x*s + dw(tap), ci]
//
// GEMM view: M = co (128 per CTA), N = ci (<= 256 per CTA), K = output pixels.  Both operands
// come straight out of NHWC activation tensors: a TMA box [64 ch, tw, th, 1] lands in smem as
// 64 pixel rows of 128 bytes (64 channels), which is exactly UMMA's MN-major SWIZZLE_128B
// layout (channels contiguous, rows step along K).  So no transposes are ever materialised.
//
// Work split: (tap, co-tile, ci-tile) output tiles x a split of the pixel range; partial sums
// are added with vectorised fp32 reductions (red.global.add.v4.f32) directly into the gradient
// bucket, which is what the allreduce kernel later reads.  Shared heads accumulate their five
// pyramid levels in the same launch.
//
// Replaces cuDNN backward-filter; no counterpart source in the reference.
#include "../include/b200_ops.h"
#include "common.cuh"
#include "ptx_sm100.cuh"

namespace b200 {

constexpr int kWStages = 3;
constexpr int kKT = 64;                           // pixels per pipeline stage
constexpr int kChunkBytes = kKT * 128;            // one [64 px][64 ch] chunk = 8 KB
constexpr int kWAStage = 4 * kChunkBytes;         // up to two 128-row M tiles -> 4 chunks
constexpr int kWBStage = 4 * kChunkBytes;         // N <= 256 -> 4 chunks
constexpr int kWThreads = 128 + 8 * 32;           // TMA, MMA, TMEM alloc, spare + 8 epilogue warps
constexpr int kWSmemBytes = kWStages * (kWAStage + kWBStage) + 1024 + 256;
constexpr int kWMaxTaps = 49;

struct WTap {
  int8_t dh, dw, map, pad;
};

struct WProblem {
  int N, OH, OW;
  int tiles_h, tiles_w;
  int tile_begin;  // first pixel tile of this problem
  int th_log2, tw_log2;
  int xmap_base;
};

// One CTA = one output block [MT x 128 rows of co] x [up to 4 packed 64-column chunks], over a range
// of pixel tiles.  A packed chunk is (tap, 64-channel slice of ci): for Cin >= 256 the four chunks are
// four ci slices of one tap; for Cin = 128 two taps x two slices; for Cin = 64 four taps.  The B tile
// (shifted X) is shared by both M tiles, and for small Cin the A tile (dY) is shared by several taps.
struct WgradParams {
  CUtensorMap x_maps[8];
  CUtensorMap dy_maps[B200_MAX_PROBLEMS];
  WProblem prob[B200_MAX_PROBLEMS];
  WTap taps[kWMaxTaps];
  float* dw;
  const float* rowscale;
  int num_problems, num_pixel_tiles;
  int num_taps, taps_per_unit, chunks_per_tap, num_tap_groups, num_m_groups, num_n_tiles, mt, ksplit, chunk;
  int Cin, Cout;
};

__device__ __forceinline__ void w_decode(const WgradParams& p, int pt, int& pi, int& n, int& y0, int& x0) {
  pi = 0;
#pragma unroll
  for (int i = 1; i < B200_MAX_PROBLEMS; ++i)
    if (i < p.num_problems && pt >= p.prob[i].tile_begin) pi = i;
  const WProblem& q = p.prob[pi];
  const int local = pt - q.tile_begin;
  const int per_img = q.tiles_h * q.tiles_w;
  n = local / per_img;
  const int rem = local - n * per_img;
  const int ty = rem / q.tiles_w;
  const int tx = rem - ty * q.tiles_w;
  y0 = ty << q.th_log2;
  x0 = tx << q.tw_log2;
}

__global__ void __launch_bounds__(kWThreads, 1) wgrad_kernel(const __grid_constant__ WgradParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t a_smem = smem_base;
  const uint32_t b_smem = smem_base + kWStages * kWAStage;
  const uint32_t bar_base = b_smem + kWStages * kWBStage;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kWStages + s); };
  const uint32_t tfull_bar = bar_base + 8u * (2 * kWStages);
  const uint32_t tmem_ptr_smem = bar_base + 8u * (2 * kWStages + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  // unit decode: blockIdx.x = ((tap_group * num_m_groups + m_group) * num_n_tiles + n_tile) * ksplit + split
  int u = blockIdx.x;
  const int split = u % p.ksplit;
  u /= p.ksplit;
  const int n_tile = u % p.num_n_tiles;
  u /= p.num_n_tiles;
  const int m_group = u % p.num_m_groups;
  const int tap_group = u / p.num_m_groups;
  const int pt_begin = split * p.chunk;
  const int pt_end = min(p.num_pixel_tiles, pt_begin + p.chunk);
  const int tap0 = tap_group * p.taps_per_unit;
  const int ntaps_here = min(p.taps_per_unit, p.num_taps - tap0);
  // packed B chunks of this unit
  const int n_chunks = (p.taps_per_unit > 1) ? ntaps_here * p.chunks_per_tap : min(4, (p.Cin - n_tile * 256 + 63) >> 6);
  const int N = n_chunks * 64;
  // A chunks: MT tiles x 2 halves of 64 output channels, loaded only where co < Cout
  const int co_base = m_group * p.mt * 128;
  int a_chunks = 0;
  for (int c = 0; c < 2 * p.mt; ++c) a_chunks += (co_base + c * 64 < p.Cout) ? 1 : 0;
  const int mt_valid = (a_chunks + 1) >> 1;
  const uint32_t stage_bytes = static_cast<uint32_t>(a_chunks + n_chunks) * kChunkBytes;

  if (warp == 0 && lane == 0) {
    for (int i = 0; i < 8; ++i) tma_prefetch_desc(&p.x_maps[i]);
    for (int i = 0; i < B200_MAX_PROBLEMS; ++i) tma_prefetch_desc(&p.dy_maps[i]);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kWStages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    mbar_init(tfull_bar, 1);
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_ptr_smem, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_ptr_smem));
  pdl_launch_dependents();   // set-up done: the next kernel may start its own; then wait for our inputs
  pdl_wait();

  if (pt_begin < pt_end) {
    if (warp == 0) {
      // Producer warp: lane 0 arms the barrier, then lanes 0..(a_chunks + n_chunks - 1) each issue ONE of the
      // stage's TMA boxes, so the (up to 8) box issues of a stage overlap instead of serialising in one thread.
      int stage = 0;
      uint32_t phase = 0;
      int t = tap0, ci = 0;
      bool is_a = lane < a_chunks;
      const int bc = lane - a_chunks;   // packed B chunk of this lane
      if (!is_a && bc < n_chunks) {
        if (p.taps_per_unit > 1) {
          t = tap0 + bc / p.chunks_per_tap;
          ci = (bc % p.chunks_per_tap) * 64;
        } else {
          ci = n_tile * 256 + bc * 64;
        }
      }
      const bool active = lane < a_chunks + n_chunks;
      const WTap tap = p.taps[t];
      for (int pt = pt_begin; pt < pt_end; ++pt) {
        int pi, n, y0, x0;
        w_decode(p, pt, pi, n, y0, x0);
        mbar_wait(empty_bar(stage), phase ^ 1);
        if (lane == 0) mbar_expect_tx(full_bar(stage), stage_bytes);
        __syncwarp();
        if (active) {
          if (is_a)
            tma_load_4d(a_smem + stage * kWAStage + lane * kChunkBytes, &p.dy_maps[pi], full_bar(stage), co_base + lane * 64,
                        x0, y0, n);
          else
            tma_load_4d(b_smem + stage * kWBStage + bc * kChunkBytes, &p.x_maps[p.prob[pi].xmap_base + tap.map],
                        full_bar(stage), ci, x0 + tap.dw, y0 + tap.dh, n);
        }
        if (++stage == kWStages) {
          stage = 0;
          phase ^= 1;
        }
      }
    } else if (warp == 1) {
      const uint32_t idesc = make_idesc_bf16(128, N, 1, 1);
      int stage = 0;
      uint32_t phase = 0;
      for (int pt = pt_begin; pt < pt_end; ++pt) {
        mbar_wait(full_bar(stage), phase);
        tc_fence_after();
        if (elect_one()) {   // elect.sync, not lane == 0: keeps the MMA operands in uniform registers (no election loops)
          // MN-major: LBO = distance between 64-channel chunks, SBO = 8 pixel rows = 1024 B
          const uint64_t bdesc = make_smem_desc(b_smem + stage * kWBStage, kChunkBytes, 1024);
          for (int mt = 0; mt < mt_valid; ++mt) {
            const uint64_t adesc = make_smem_desc(a_smem + stage * kWAStage + mt * 2 * kChunkBytes, kChunkBytes, 1024);
#pragma unroll
            for (int k = 0; k < kKT / 16; ++k) {
              // 16 pixel rows = 2048 bytes -> +128 in (addr >> 4) units
              umma_bf16(tmem_base + mt * 256, adesc + 128 * k, bdesc + 128 * k, idesc, (pt != pt_begin) || (k != 0));
            }
          }
          umma_commit(empty_bar(stage));
          if (pt == pt_end - 1) umma_commit(tfull_bar);
        }
        __syncwarp();
        if (++stage == kWStages) {
          stage = 0;
          phase ^= 1;
        }
      }
    } else if (warp >= 4) {
      const int ew = warp - 4;
      const int quarter = ew & 3;
      const int mt = ew >> 2;
      if (mt < mt_valid) {
        const int co = co_base + mt * 128 + quarter * 32 + lane;
        const bool valid = co < p.Cout;
        const float rs = (valid && p.rowscale) ? __ldg(p.rowscale + co) : 1.f;
        mbar_wait(tfull_bar, 0);
        tc_fence_after();
        const uint32_t t_row = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + mt * 256;
        for (int c = 0; c < N; c += 32) {
          uint32_t r[32];
          tmem_ld_32x32(t_row + c, r);
          tmem_ld_wait();
          const int pc = c >> 6;  // packed chunk
          int t, ci;
          if (p.taps_per_unit > 1) {
            t = tap0 + pc / p.chunks_per_tap;
            ci = (pc % p.chunks_per_tap) * 64;
          } else {
            t = tap0;
            ci = n_tile * 256 + pc * 64;
          }
          if (valid) {
            float* row = p.dw + (static_cast<long long>(t) * p.Cout + co) * p.Cin + ci + (c & 63);
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              if (ci + (c & 63) + j >= p.Cin) break;      // Cin need not fill its last 64-channel chunk (zero-filled by TMA)
              asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(row + j),
                           "f"(__uint_as_float(r[j]) * rs), "f"(__uint_as_float(r[j + 1]) * rs),
                           "f"(__uint_as_float(r[j + 2]) * rs), "f"(__uint_as_float(r[j + 3]) * rs)
                           : "memory");
            }
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, 512);
}

static int floordiv_w(int a, int b) { return (a >= 0) ? a / b : -((-a + b - 1) / b); }
static int posmod_w(int a, int b) { return ((a % b) + b) % b; }

static void pick_tile64(int OH, int OW, int& th_log2, int& tw_log2) {
  long long best = -1;
  for (int tl = 6; tl >= 0; --tl) {
    const int tw = 1 << tl, th = 64 >> tl;
    const long long tiles = static_cast<long long>((OH + th - 1) / th) * ((OW + tw - 1) / tw);
    if (best < 0 || tiles < best) {
      best = tiles;
      tw_log2 = tl;
      th_log2 = 6 - tl;
    }
  }
}

}  // namespace b200

extern "C" int b200_colsum_multi(const void* const* xs, const long long* rows, int n, int cols, int ld, float* out_accum,
                                 void* stream);

extern "C" int b200_wgrad(const B200WgradArgs* a, void* stream_) {
  using namespace b200;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (!a || a->num_problems < 1 || a->num_problems > B200_MAX_PROBLEMS) return set_error(-1, "b200_wgrad: bad num_problems");
  if (a->stride != 1 && a->stride != 2) return set_error(-1, "b200_wgrad: stride must be 1 or 2");
  if (a->stride == 2 && a->num_problems > 2) return set_error(-1, "b200_wgrad: at most 2 problems with stride 2");
  if (a->KH * a->KW > kWMaxTaps) return set_error(-1, "b200_wgrad: too many taps");
  if ((a->ldx % 8) || (a->ldy % 8)) return set_error(-1, "b200_wgrad: ld* must be multiples of 8");
  if (a->Cin % 8) return set_error(-1, "b200_wgrad: Cin must be a multiple of 8");

  static std::atomic<unsigned long long> attr_set{0};   // per device
  if (device_needs_setup(attr_set)) {
    B200_CUDA_OK(cudaFuncSetAttribute(wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kWSmemBytes));
    device_setup_done(attr_set);
  }

  WgradParams p;
  memset(&p, 0, sizeof(p));
  p.dw = a->dw;
  p.rowscale = a->rowscale;
  p.num_problems = a->num_problems;
  p.Cin = a->Cin;
  p.Cout = a->Cout;
  p.chunks_per_tap = a->Cin >= 256 ? 4 : (a->Cin + 63) / 64;
  p.taps_per_unit = a->Cin >= 256 ? 1 : 4 / p.chunks_per_tap;   // pack taps into N when Cin is small
  p.num_n_tiles = a->Cin >= 256 ? (a->Cin + 255) / 256 : 1;
  p.mt = a->Cout > 128 ? 2 : 1;
  p.num_m_groups = (a->Cout + 128 * p.mt - 1) / (128 * p.mt);
  const int s = a->stride;
  int nt = 0;
  for (int r = 0; r < a->KH; ++r)
    for (int c = 0; c < a->KW; ++c) {
      const int eh = r - a->pad_h, ew = c - a->pad_w;
      WTap t;
      t.dh = static_cast<int8_t>(floordiv_w(eh, s));
      t.dw = static_cast<int8_t>(floordiv_w(ew, s));
      t.map = static_cast<int8_t>(posmod_w(eh, s) * s + posmod_w(ew, s));
      t.pad = 0;
      p.taps[nt++] = t;
    }
  p.num_taps = nt;
  if (p.taps_per_unit > nt) p.taps_per_unit = nt;
  p.num_tap_groups = (nt + p.taps_per_unit - 1) / p.taps_per_unit;

  int tile_begin = 0, map_count = 0;
  for (int i = 0; i < a->num_problems; ++i) {
    B200WgradProblem q = a->prob[i];
    // 1x1 / stride 1: a flat list of pixels, tiles of any 64 consecutive ones (no partly filled tiles at row / image ends)
    if (a->KH == 1 && a->KW == 1 && s == 1 && a->pad_h == 0 && a->pad_w == 0 && q.OH == q.H && q.OW == q.W) {
      const long long px = static_cast<long long>(q.N) * q.H * q.W;
      if (px < (1LL << 31)) {
        q.N = 1;
        q.H = q.OH = 1;
        q.W = q.OW = static_cast<int>(px);
      }
    }
    WProblem& d = p.prob[i];
    d.N = q.N;
    d.OH = q.OH;
    d.OW = q.OW;
    pick_tile64(q.OH, q.OW, d.th_log2, d.tw_log2);
    const int th = 1 << d.th_log2, tw = 1 << d.tw_log2;
    d.tiles_h = (q.OH + th - 1) / th;
    d.tiles_w = (q.OW + tw - 1) / tw;
    d.tile_begin = tile_begin;
    tile_begin += q.N * d.tiles_h * d.tiles_w;
    d.xmap_base = map_count;
    const unsigned int box[4] = {64u, static_cast<unsigned>(tw), static_cast<unsigned>(th), 1u};
    const int nmaps = (s == 2) ? 4 : 1;
    if (map_count + nmaps > 8) return set_error(-1, "b200_wgrad: too many tensor maps");
    for (int mi = 0; mi < nmaps; ++mi) {
      const int py = mi / 2, px = mi % 2;
      const int step = (nmaps == 4) ? 2 : 1;
      const long long Wd = (q.W - px + step - 1) / step;
      const long long Hd = (q.H - py + step - 1) / step;
      const __nv_bfloat16* base = static_cast<const __nv_bfloat16*>(q.x) + (static_cast<long long>(py) * q.W + px) * a->ldx;
      unsigned long long dims[4] = {static_cast<unsigned long long>(a->Cin), static_cast<unsigned long long>(Wd > 0 ? Wd : 1),
                                    static_cast<unsigned long long>(Hd > 0 ? Hd : 1), static_cast<unsigned long long>(q.N)};
      unsigned long long strides[3] = {static_cast<unsigned long long>(step) * a->ldx * 2ull,
                                       static_cast<unsigned long long>(step) * q.W * a->ldx * 2ull,
                                       static_cast<unsigned long long>(q.H) * q.W * a->ldx * 2ull};
      int rc = make_tmap_bf16(&p.x_maps[map_count + mi], base, 4, dims, strides, box);
      if (rc) return rc;
    }
    map_count += nmaps;
    {
      unsigned long long dims[4] = {static_cast<unsigned long long>(a->Cout), static_cast<unsigned long long>(q.OW),
                                    static_cast<unsigned long long>(q.OH), static_cast<unsigned long long>(q.N)};
      unsigned long long strides[3] = {static_cast<unsigned long long>(a->ldy) * 2ull,
                                       static_cast<unsigned long long>(q.OW) * a->ldy * 2ull,
                                       static_cast<unsigned long long>(q.OH) * q.OW * a->ldy * 2ull};
      int rc = make_tmap_bf16(&p.dy_maps[i], q.dy, 4, dims, strides, box);
      if (rc) return rc;
    }
  }
  for (int i = map_count; i < 8; ++i) p.x_maps[i] = p.x_maps[0];
  for (int i = a->num_problems; i < B200_MAX_PROBLEMS; ++i) p.dy_maps[i] = p.dy_maps[0];
  p.num_pixel_tiles = tile_begin;

  const long long units = static_cast<long long>(p.num_tap_groups) * p.num_m_groups * p.num_n_tiles;
  // ~3 waves of CTAs, but at least 8 pixel tiles (512 pixels) per CTA to amortise the epilogue
  // One wave: every CTA does the same amount of MMA work, so <= #SM CTAs is balanced, and each extra
  // split only adds another atomic flush of the whole dW block (the dominant cost at 3 waves).
  long long ks = static_cast<long long>(num_sms()) / units;
  const long long max_ks = (p.num_pixel_tiles + 7) / 8;
  if (ks > max_ks) ks = max_ks;
  if (ks < 1) ks = 1;
  if (deterministic()) ks = 1;          // one CTA per output block: no cross-CTA atomic order
  p.chunk = static_cast<int>((p.num_pixel_tiles + ks - 1) / ks);
  p.ksplit = (p.num_pixel_tiles + p.chunk - 1) / p.chunk;
  const long long grid = units * p.ksplit;
  if (grid > 0x7fffffffLL) return set_error(-1, "b200_wgrad: grid too large");

  launch_pdl(wgrad_kernel, static_cast<unsigned>(grid), kWThreads, kWSmemBytes, stream, p);
  int rc = check_launch("wgrad_kernel");
  if (rc) return rc;
  if (a->dbias) {
    const void* ptrs[B200_MAX_PROBLEMS];
    long long rows[B200_MAX_PROBLEMS];
    for (int i = 0; i < a->num_problems; ++i) {
      ptrs[i] = a->prob[i].dy;
      rows[i] = static_cast<long long>(a->prob[i].N) * a->prob[i].OH * a->prob[i].OW;
    }
    rc = b200_colsum_multi(ptrs, rows, a->num_problems, a->Cout, a->ldy, a->dbias, stream_);
    if (rc) return rc;
  }
  return 0;
}
