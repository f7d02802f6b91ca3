// Bandwidth-bound fused kernels: weight preparation (BN-scale fold + layout transform),
// stem BN+ReLU+max-pool (fwd/bwd), FPN nearest-upsample-to-shape + add (fwd/bwd), ReLU,
// column sums (bias gradients), image normalisation.  All bf16 NHWC, 16-byte vector accesses.
// Replaces the separate TF elementwise/pool/resize kernels of the reference stack
// (SURVEY.md G4, G5, G7, G19); no counterpart source in the reference.
#include "../include/b200_ops.h"
#include "common.cuh"
#include "ptx_sm100.cuh"

namespace b200 {

// ------------------------------------------------------------------ weight preparation
__global__ void prepare_weights_kernel(const B200WeightEntry* __restrict__ table, int n) {
  __shared__ float tile[32][33];
  // locate the entry that owns this block (entries sorted by block_begin)
  int lo = 0, hi = n - 1;
  const long long b = blockIdx.x;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (table[mid].block_begin <= b) lo = mid; else hi = mid - 1;
  }
  const B200WeightEntry e = table[lo];
  long long local = b - e.block_begin;
  const int tiles_ci = (e.Cin + 31) / 32;
  const int tiles_co = (e.Cout + 31) / 32;
  const int tci = static_cast<int>(local % tiles_ci);
  local /= tiles_ci;
  const int tco = static_cast<int>(local % tiles_co);
  const int tap = static_cast<int>(local / tiles_co);
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  __nv_bfloat16* df = static_cast<__nv_bfloat16*>(e.dst_f);
  __nv_bfloat16* dd = static_cast<__nv_bfloat16*>(e.dst_d);
#pragma unroll
  for (int r = ty; r < 32; r += 8) {
    const int co = tco * 32 + r, ci = tci * 32 + tx;
    float v = 0.f;
    if (co < e.Cout && ci < e.Cin) {
      v = e.src[(static_cast<long long>(tap) * e.Cout + co) * e.Cin + ci];
      if (e.scale) v *= e.scale[co];
      if (df) df[(static_cast<long long>(tap) * e.Cout + co) * e.ldf + ci] = __float2bfloat16(v);
    }
    tile[r][tx] = v;
  }
  __syncthreads();
  if (dd) {
#pragma unroll
    for (int r = ty; r < 32; r += 8) {
      const int ci = tci * 32 + r, co = tco * 32 + tx;
      if (co < e.Cout && ci < e.Cin)
        dd[(static_cast<long long>(tap) * e.Cin + ci) * e.ldd + co] = __float2bfloat16(tile[tx][r]);
    }
  }
}

// ------------------------------------------------------------------ stem BN + ReLU + maxpool
__device__ __forceinline__ void load8(const __nv_bfloat16* p, float (&v)[8]) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  v[0] = bf16_lo(u.x); v[1] = bf16_hi(u.x); v[2] = bf16_lo(u.y); v[3] = bf16_hi(u.y);
  v[4] = bf16_lo(u.z); v[5] = bf16_hi(u.z); v[6] = bf16_lo(u.w); v[7] = bf16_hi(u.w);
}
__device__ __forceinline__ void store8(__nv_bfloat16* p, const float (&v)[8]) {
  uint4 u;
  u.x = pack_bf16x2(v[0], v[1]); u.y = pack_bf16x2(v[2], v[3]);
  u.z = pack_bf16x2(v[4], v[5]); u.w = pack_bf16x2(v[6], v[7]);
  *reinterpret_cast<uint4*>(p) = u;
}
__device__ __forceinline__ float round_bf16(float x) { return __bfloat162float(__float2bfloat16(x)); }

// Forward also records, per output element, which window position (r*3+s) won (first maximum in
// scan order; 255 when the maximum is <= 0, i.e. no gradient flows), so that backward is a pure
// gather over <= 4 windows per input position without re-reading the conv output.
// 32-bit index arithmetic (the host checks the element count) and per-channel constants reloaded only when the
// thread's channel group changes (with C = 64 and 256-thread blocks it never does).
__global__ void __launch_bounds__(256) bn_relu_maxpool_fwd_kernel(const __nv_bfloat16* __restrict__ x,
                                                                  const float* __restrict__ scale,
                                                                  const float* __restrict__ shift,
                                                                  __nv_bfloat16* __restrict__ y, uint8_t* __restrict__ idx,
                                                                  int N, int H, int W, int C, int OH, int OW, int pad_h,
                                                                  int pad_w) {
  const unsigned cg = static_cast<unsigned>(C) >> 3;
  const unsigned total = static_cast<unsigned>(N) * OH * OW * cg;
  float sc[8], sh[8];
  int c_have = -1;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int c = static_cast<int>(i % cg) * 8;
    unsigned t = i / cg;
    const int ow = static_cast<int>(t % static_cast<unsigned>(OW));
    t /= static_cast<unsigned>(OW);
    const int oh = static_cast<int>(t % static_cast<unsigned>(OH));
    const int n = static_cast<int>(t / static_cast<unsigned>(OH));
    if (c != c_have) {
#pragma unroll
      for (int e = 0; e < 8; ++e) { sc[e] = scale[c + e]; sh[e] = shift[c + e]; }
      c_have = c;
    }
    float m[8];
    uint32_t am[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) { m[e] = 0.f; am[e] = 255u; }
    const int h0 = oh * 2 - pad_h, w0 = ow * 2 - pad_w;
    const __nv_bfloat16* xb = x + (static_cast<size_t>(n) * H * W) * C + c;
    // all nine window loads are issued before the first compare
    uint4 v[9];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int q = 0; q < 3; ++q) {
        const int h = h0 + r, w = w0 + q;
        const bool ok = h >= 0 && h < H && w >= 0 && w < W;
        v[r * 3 + q] = ok ? *reinterpret_cast<const uint4*>(xb + (static_cast<size_t>(h) * W + w) * C) : make_uint4(0u, 0u, 0u, 0u);
      }
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      const int h = h0 + k / 3, w = w0 + k % 3;
      if (h < 0 || h >= H || w < 0 || w >= W) continue;
      const float f[8] = {bf16_lo(v[k].x), bf16_hi(v[k].x), bf16_lo(v[k].y), bf16_hi(v[k].y),
                          bf16_lo(v[k].z), bf16_hi(v[k].z), bf16_lo(v[k].w), bf16_hi(v[k].w)};
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float z = fmaf(f[e], sc[e], sh[e]);
        if (z > m[e]) { m[e] = z; am[e] = k; }
      }
    }
    const size_t o = static_cast<size_t>(i) * 8;          // output index == loop index (same decomposition)
    store8(y + o, m);
    if (idx) {
      uint2 pk;
      pk.x = am[0] | (am[1] << 8) | (am[2] << 16) | (am[3] << 24);
      pk.y = am[4] | (am[5] << 8) | (am[6] << 16) | (am[7] << 24);
      *reinterpret_cast<uint2*>(idx + o) = pk;
    }
  }
}

// dx[n,h,w,c] = scale[c] * sum over windows whose recorded arg-max is (h,w) of dy.
// An input position lies in at most 2 x 2 windows (3x3 / stride 2).  The four index words are loaded together, then the
// dy vectors of the windows that match, so a thread waits for two memory round trips instead of up to eight.
__global__ void __launch_bounds__(256) bn_relu_maxpool_bwd_kernel(const uint8_t* __restrict__ idx,
                                                                  const float* __restrict__ scale,
                                                                  const __nv_bfloat16* __restrict__ dy,
                                                                  __nv_bfloat16* __restrict__ dx, int N, int H, int W, int C,
                                                                  int OH, int OW, int pad_h, int pad_w) {
  const unsigned cg = static_cast<unsigned>(C) >> 3;
  const unsigned total = static_cast<unsigned>(N) * H * W * cg;
  float sc[8];
  int c_have = -1;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int c = static_cast<int>(i % cg) * 8;
    unsigned t = i / cg;
    const int w = static_cast<int>(t % static_cast<unsigned>(W));
    t /= static_cast<unsigned>(W);
    const int h = static_cast<int>(t % static_cast<unsigned>(H));
    const int n = static_cast<int>(t / static_cast<unsigned>(H));
    if (c != c_have) {
#pragma unroll
      for (int e = 0; e < 8; ++e) sc[e] = scale[c + e];
      c_have = c;
    }
    // candidate windows: oh = (h + pad_h) >> 1 always, and the one above it when h + pad_h is even (same along w)
    const int th = h + pad_h, tw = w + pad_w;
    int ohs[2], ows[2];
    bool vh[2], vw[2];
    ohs[1] = th >> 1;
    ohs[0] = ohs[1] - 1;
    vh[1] = ohs[1] < OH;
    vh[0] = !(th & 1) && ohs[0] >= 0 && ohs[0] < OH;
    ows[1] = tw >> 1;
    ows[0] = ows[1] - 1;
    vw[1] = ows[1] < OW;
    vw[0] = !(tw & 1) && ows[0] >= 0 && ows[0] < OW;
    uint2 pk[4];
    size_t off[4];
    uint32_t me[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int a = k >> 1, b = k & 1;
      const bool ok = vh[a] && vw[b];
      off[k] = ((static_cast<size_t>(n) * OH + (ok ? ohs[a] : 0)) * OW + (ok ? ows[b] : 0)) * C + c;
      me[k] = static_cast<uint32_t>((th - 2 * ohs[a]) * 3 + (tw - 2 * ows[b]));
      pk[k] = ok ? *reinterpret_cast<const uint2*>(idx + off[k]) : make_uint2(0xffffffffu, 0xffffffffu);
    }
    uint4 gy[4];
    bool hit[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const uint32_t me4 = me[k] * 0x01010101u;
      const uint32_t x0 = pk[k].x ^ me4, x1 = pk[k].y ^ me4;          // byte-wise equality test
      hit[k] = (((x0 - 0x01010101u) & ~x0 & 0x80808080u) | ((x1 - 0x01010101u) & ~x1 & 0x80808080u)) != 0u;
      gy[k] = hit[k] ? *reinterpret_cast<const uint4*>(dy + off[k]) : make_uint4(0u, 0u, 0u, 0u);
    }
    float g[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) g[e] = 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (!hit[k]) continue;
      const float f[8] = {bf16_lo(gy[k].x), bf16_hi(gy[k].x), bf16_lo(gy[k].y), bf16_hi(gy[k].y),
                          bf16_lo(gy[k].z), bf16_hi(gy[k].z), bf16_lo(gy[k].w), bf16_hi(gy[k].w)};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        if (((pk[k].x >> (8 * e)) & 0xffu) == me[k]) g[e] += f[e];
        if (((pk[k].y >> (8 * e)) & 0xffu) == me[k]) g[4 + e] += f[4 + e];
      }
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) g[e] *= sc[e];
    store8(dx + static_cast<size_t>(i) * 8, g);
  }
}

// ------------------------------------------------------------------ FPN upsample-to-shape + add
__global__ void upsample_add_kernel(const __nv_bfloat16* __restrict__ lateral, const __nv_bfloat16* __restrict__ top,
                                    __nv_bfloat16* __restrict__ out, int N, int H, int W, int C, int TH, int TW) {
  const int cg = C >> 3;
  const long long total = static_cast<long long>(N) * H * W * cg;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % cg) * 8;
    long long t = i / cg;
    const int w = static_cast<int>(t % W);
    t /= W;
    const int h = static_cast<int>(t % H);
    const int n = static_cast<int>(t / H);
    // nearest (legacy): src = floor(dst * in / out)
    const int th = min(TH - 1, static_cast<int>((static_cast<long long>(h) * TH) / H));
    const int tw = min(TW - 1, static_cast<int>((static_cast<long long>(w) * TW) / W));
    float a[8], b[8];
    load8(lateral + ((static_cast<long long>(n) * H + h) * W + w) * C + c, a);
    load8(top + ((static_cast<long long>(n) * TH + th) * TW + tw) * C + c, b);
#pragma unroll
    for (int e = 0; e < 8; ++e) a[e] += b[e];
    store8(out + ((static_cast<long long>(n) * H + h) * W + w) * C + c, a);
  }
}

__global__ void upsample_bwd_add_kernel(const __nv_bfloat16* __restrict__ dmerged, __nv_bfloat16* __restrict__ dtop,
                                        int N, int H, int W, int C, int TH, int TW) {
  const int cg = C >> 3;
  const long long total = static_cast<long long>(N) * TH * TW * cg;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % cg) * 8;
    long long t = i / cg;
    const int tw = static_cast<int>(t % TW);
    t /= TW;
    const int th = static_cast<int>(t % TH);
    const int n = static_cast<int>(t / TH);
    // destination rows h with floor(h*TH/H) == th  <=>  ceil(th*H/TH) <= h < ceil((th+1)*H/TH)
    const int h0 = static_cast<int>((static_cast<long long>(th) * H + TH - 1) / TH);
    const int h1 = static_cast<int>((static_cast<long long>(th + 1) * H + TH - 1) / TH);
    const int w0 = static_cast<int>((static_cast<long long>(tw) * W + TW - 1) / TW);
    const int w1 = static_cast<int>((static_cast<long long>(tw + 1) * W + TW - 1) / TW);
    float acc[8];
    load8(dtop + ((static_cast<long long>(n) * TH + th) * TW + tw) * C + c, acc);
    for (int h = h0; h < min(h1, H); ++h)
      for (int w = w0; w < min(w1, W); ++w) {
        float g[8];
        load8(dmerged + ((static_cast<long long>(n) * H + h) * W + w) * C + c, g);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] += g[e];
      }
    store8(dtop + ((static_cast<long long>(n) * TH + th) * TW + tw) * C + c, acc);
  }
}

__global__ void relu_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, long long n8) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n8;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float v[8];
    load8(x + i * 8, v);
#pragma unroll
    for (int e = 0; e < 8; ++e) v[e] = fmaxf(v[e], 0.f);
    store8(y + i * 8, v);
  }
}
__global__ void relu_bwd_kernel(const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ dy,
                                __nv_bfloat16* __restrict__ dx, long long n8) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n8;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float v[8], g[8];
    load8(y + i * 8, v);
    load8(dy + i * 8, g);
#pragma unroll
    for (int e = 0; e < 8; ++e) g[e] = v[e] > 0.f ? g[e] : 0.f;
    store8(dx + i * 8, g);
  }
}

__global__ void add_mask_kernel(const __nv_bfloat16* __restrict__ a, const __nv_bfloat16* __restrict__ b,
                                const __nv_bfloat16* __restrict__ mask, __nv_bfloat16* __restrict__ out, long long n8) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n8;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float v[8], m[8];
    load8(a + i * 8, v);
    if (b) {
      float w[8];
      load8(b + i * 8, w);
#pragma unroll
      for (int e = 0; e < 8; ++e) v[e] += w[e];
    }
    if (mask) {
      load8(mask + i * 8, m);
#pragma unroll
      for (int e = 0; e < 8; ++e) v[e] = m[e] > 0.f ? v[e] : 0.f;
    }
    store8(out + i * 8, v);
  }
}

// ------------------------------------------------------------------ column sums (bias grads)
// Block = 8 row-lanes x 32 column-groups (8 channels each); 64 rows per block so that a thread only
// chains 8 loads (the kernel is latency bound, not bandwidth bound, on the small pyramid levels);
// up to 5 matrices (pyramid levels) per launch; one atomicAdd per column per block.
constexpr int kColsumRows = 64;
struct ColsumArgs {
  const __nv_bfloat16* x[5];
  long long rows[5];
  long long block_begin[6];
  int n;
};
__global__ void __launch_bounds__(256) colsum_kernel(ColsumArgs a, int cols, int ld, float* __restrict__ out) {
  __shared__ float part[8][32][8];
  pdl_launch_dependents();   // sits between wgrad and dgrad of every biased layer: keep the dependent-launch chain
  pdl_wait();
  int pi = 0;
#pragma unroll
  for (int i = 1; i < 5; ++i)
    if (i < a.n && static_cast<long long>(blockIdx.x) >= a.block_begin[i]) pi = i;
  const __nv_bfloat16* __restrict__ x = a.x[pi];
  const long long rows = a.rows[pi];
  const int cgx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int c = (blockIdx.y * 32 + cgx) * 8;
  const long long r0 = (static_cast<long long>(blockIdx.x) - a.block_begin[pi]) * kColsumRows;
  float acc[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) acc[e] = 0.f;
  if (c < cols) {
    float v[8][8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const long long r = r0 + ry + 8 * k;
      if (r < rows) {
        load8(x + r * ld + c, v[k]);
      } else {
#pragma unroll
        for (int e = 0; e < 8; ++e) v[k][e] = 0.f;
      }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k)
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] += v[k][e];
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) part[ry][cgx][e] = acc[e];
  __syncthreads();
  const int g2 = threadIdx.x >> 3, e2 = threadIdx.x & 7;
  const int cc = (blockIdx.y * 32 + g2) * 8 + e2;
  if (cc < cols) {
    float sum = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) sum += part[k][g2][e2];
    atomicAdd(out + cc, sum);
  }
}

// deterministic variant: one thread per column walks every row of every matrix in order
__global__ void colsum_det_kernel(ColsumArgs a, int cols, int ld, float* __restrict__ out) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cols) return;
  float acc = 0.f;
  for (int pi = 0; pi < a.n; ++pi)
    for (long long r = 0; r < a.rows[pi]; ++r) acc += __bfloat162float(a.x[pi][r * ld + c]);
  out[c] += acc;
}

// ------------------------------------------------------------------ image normalise + pad
__global__ void preprocess_kernel(const uint8_t* __restrict__ src, __nv_bfloat16* __restrict__ dst, long long pixels,
                                  int H, int W, int C, int ldc, float m0, float m1, float m2, float m3,
                                  const int* __restrict__ img_hw, float scale) {
  const float mean[4] = {m0, m1, m2, m3};
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < pixels;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    bool inside = true;
    if (img_hw) {
      const int x = static_cast<int>(i % W);
      const long long t = i / W;
      const int y = static_cast<int>(t % H), n = static_cast<int>(t / H);
      inside = y < img_hw[2 * n] && x < img_hw[2 * n + 1];
    }
    for (int c = 0; c < ldc; ++c) {
      const float v = (c < C && inside) ? (static_cast<float>(src[i * C + c]) - mean[c & 3]) * scale : 0.f;
      dst[i * ldc + c] = __float2bfloat16(v);
    }
  }
}

static int grid_for(long long work, int block) {
  long long g = (work + block - 1) / block;
  const long long cap = static_cast<long long>(num_sms()) * 16;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return static_cast<int>(g);
}


// ---------------------------------------------------------------------------------------------
// 2x2 / 2 max-pool with "valid" padding (VGG: sizes floor) over NHWC bf16.  The pooled tensor x is a ReLU
// output, so backward folds that ReLU's mask in: the result is the gradient w.r.t. the PRE-activation of the
// conv that produced x (first maximum in scan order wins, as in the library kernels; a maximum of 0 passes
// nothing).  Windows hanging over an odd last row / column only zero-fill their in-range inputs.
__global__ void maxpool2_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, int N, int H,
                                    int W, int C, int OH, int OW) {
  const int cg = C >> 3;
  const long long total = static_cast<long long>(N) * OH * OW * cg;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % cg) * 8;
    long long t = i / cg;
    const int ow = static_cast<int>(t % OW);
    t /= OW;
    const int oh = static_cast<int>(t % OH);
    const int n = static_cast<int>(t / OH);
    const __nv_bfloat16* base = x + ((static_cast<long long>(n) * H + 2 * oh) * W + 2 * ow) * C + c;
    float a[8], b[8], m[8];
    load8(base, a);
    load8(base + C, b);
#pragma unroll
    for (int e = 0; e < 8; ++e) m[e] = fmaxf(a[e], b[e]);
    load8(base + static_cast<long long>(W) * C, a);
    load8(base + static_cast<long long>(W) * C + C, b);
#pragma unroll
    for (int e = 0; e < 8; ++e) m[e] = fmaxf(m[e], fmaxf(a[e], b[e]));
    store8(y + ((static_cast<long long>(n) * OH + oh) * OW + ow) * C + c, m);
  }
}

__global__ void maxpool2_bwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                                    __nv_bfloat16* __restrict__ dx, int N, int H, int W, int C, int OH, int OW) {
  const int cg = C >> 3;
  const int WH = (H + 1) >> 1, WW = (W + 1) >> 1;   // windows including the ones hanging over an odd edge
  const long long total = static_cast<long long>(N) * WH * WW * cg;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % cg) * 8;
    long long t = i / cg;
    const int ow = static_cast<int>(t % WW);
    t /= WW;
    const int oh = static_cast<int>(t % WH);
    const int n = static_cast<int>(t / WH);
    const bool full = oh < OH && ow < OW;
    float v[4][8], g[8], best[8];
    int arg[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) { best[e] = 0.f; arg[e] = -1; g[e] = 0.f; }
    if (full) {
      load8(dy + ((static_cast<long long>(n) * OH + oh) * OW + ow) * C + c, g);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        load8(x + ((static_cast<long long>(n) * H + 2 * oh + (q >> 1)) * W + 2 * ow + (q & 1)) * C + c, v[q]);
#pragma unroll
        for (int e = 0; e < 8; ++e)
          if (v[q][e] > best[e]) { best[e] = v[q][e]; arg[e] = q; }
      }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int h = 2 * oh + (q >> 1), w = 2 * ow + (q & 1);
      if (h >= H || w >= W) continue;
      float o[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] = (arg[e] == q) ? g[e] : 0.f;
      store8(dx + ((static_cast<long long>(n) * H + h) * W + w) * C + c, o);
    }
  }
}

}  // namespace b200

using namespace b200;

extern "C" {

int b200_prepare_weights(const B200WeightEntry* host_entries, int n, void* dev_table, void* stream_) {
  // host_entries: block_begin already filled (see b200_weight_table_blocks); dev_table: device copy.
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (n <= 0) return 0;
  const B200WeightEntry& last = host_entries[n - 1];
  const long long blocks = last.block_begin + static_cast<long long>(last.taps) * ((last.Cout + 31) / 32) * ((last.Cin + 31) / 32);
  prepare_weights_kernel<<<static_cast<unsigned>(blocks), 256, 0, stream>>>(static_cast<const B200WeightEntry*>(dev_table), n);
  return check_launch("prepare_weights_kernel");
}

/* Fill block_begin of every entry; returns total blocks. */
long long b200_weight_table_blocks(B200WeightEntry* host_entries, int n) {
  long long b = 0;
  for (int i = 0; i < n; ++i) {
    host_entries[i].block_begin = b;
    b += static_cast<long long>(host_entries[i].taps) * ((host_entries[i].Cout + 31) / 32) * ((host_entries[i].Cin + 31) / 32);
  }
  return b;
}

int b200_bn_relu_maxpool_fwd(const void* x, const float* scale, const float* shift, void* y, uint8_t* idx, int N,
                             int H, int W, int C, int OH, int OW, int pad_h, int pad_w, void* stream) {
  if (C % 8) return set_error(-1, "maxpool: C must be a multiple of 8");
  const long long total = static_cast<long long>(N) * OH * OW * (C / 8);
  if (static_cast<long long>(N) * H * W * (C / 8) >= (1LL << 31)) return set_error(-1, "maxpool: tensor too large for 32-bit indexing");
  bn_relu_maxpool_fwd_kernel<<<grid_for(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), scale, shift, static_cast<__nv_bfloat16*>(y), idx, N, H, W, C, OH, OW,
      pad_h, pad_w);
  return check_launch("bn_relu_maxpool_fwd");
}

int b200_bn_relu_maxpool_bwd(const uint8_t* idx, const float* scale, const void* dy, void* dx, int N, int H, int W,
                             int C, int OH, int OW, int pad_h, int pad_w, void* stream) {
  if (C % 8) return set_error(-1, "maxpool: C must be a multiple of 8");
  const long long total = static_cast<long long>(N) * H * W * (C / 8);
  if (total >= (1LL << 31)) return set_error(-1, "maxpool: tensor too large for 32-bit indexing");
  bn_relu_maxpool_bwd_kernel<<<grid_for(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      idx, scale, static_cast<const __nv_bfloat16*>(dy), static_cast<__nv_bfloat16*>(dx), N, H, W, C, OH, OW, pad_h,
      pad_w);
  return check_launch("bn_relu_maxpool_bwd");
}

int b200_upsample_add(const void* lateral, const void* top, void* out, int N, int H, int W, int C, int TH, int TW,
                      void* stream) {
  if (C % 8) return set_error(-1, "upsample_add: C must be a multiple of 8");
  const long long total = static_cast<long long>(N) * H * W * (C / 8);
  upsample_add_kernel<<<grid_for(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(lateral), static_cast<const __nv_bfloat16*>(top),
      static_cast<__nv_bfloat16*>(out), N, H, W, C, TH, TW);
  return check_launch("upsample_add");
}

int b200_upsample_bwd_add(const void* dmerged, void* dtop, int N, int H, int W, int C, int TH, int TW, void* stream) {
  if (C % 8) return set_error(-1, "upsample_bwd_add: C must be a multiple of 8");
  const long long total = static_cast<long long>(N) * TH * TW * (C / 8);
  upsample_bwd_add_kernel<<<grid_for(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(dmerged), static_cast<__nv_bfloat16*>(dtop), N, H, W, C, TH, TW);
  return check_launch("upsample_bwd_add");
}

int b200_maxpool2_fwd(const void* x, void* y, int N, int H, int W, int C, void* stream) {
  if (C % 8) return set_error(-1, "maxpool2: C must be a multiple of 8");
  const int OH = H / 2, OW = W / 2;
  const long long total = static_cast<long long>(N) * OH * OW * (C / 8);
  if (total == 0) return 0;
  maxpool2_fwd_kernel<<<grid_for(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), N, H, W, C, OH, OW);
  return check_launch("maxpool2_fwd");
}

int b200_maxpool2_relu_bwd(const void* x, const void* dy, void* dx, int N, int H, int W, int C, void* stream) {
  if (C % 8) return set_error(-1, "maxpool2: C must be a multiple of 8");
  const long long total = static_cast<long long>(N) * ((H + 1) / 2) * ((W + 1) / 2) * (C / 8);
  if (total == 0) return 0;
  maxpool2_bwd_kernel<<<grid_for(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), static_cast<const __nv_bfloat16*>(dy), static_cast<__nv_bfloat16*>(dx), N, H,
      W, C, H / 2, W / 2);
  return check_launch("maxpool2_relu_bwd");
}

int b200_relu(const void* x, void* y, long long n, void* stream) {
  if (n % 8) return set_error(-1, "relu: n must be a multiple of 8");
  relu_kernel<<<grid_for(n / 8, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), n / 8);
  return check_launch("relu");
}

int b200_relu_bwd(const void* y, const void* dy, void* dx, long long n, void* stream) {
  if (n % 8) return set_error(-1, "relu_bwd: n must be a multiple of 8");
  relu_bwd_kernel<<<grid_for(n / 8, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(y), static_cast<const __nv_bfloat16*>(dy), static_cast<__nv_bfloat16*>(dx), n / 8);
  return check_launch("relu_bwd");
}

int b200_add_mask(const void* a, const void* b, const void* mask, void* out, long long n, void* stream) {
  if (n % 8) return set_error(-1, "add_mask: n must be a multiple of 8");
  add_mask_kernel<<<grid_for(n / 8, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(a), static_cast<const __nv_bfloat16*>(b),
      static_cast<const __nv_bfloat16*>(mask), static_cast<__nv_bfloat16*>(out), n / 8);
  return check_launch("add_mask");
}

int b200_colsum_multi(const void* const* xs, const long long* rows, int n, int cols, int ld, float* out_accum,
                      void* stream) {
  if (ld % 8) return set_error(-1, "colsum: ld must be a multiple of 8");
  if (n < 1 || n > 5) return set_error(-1, "colsum: 1..5 matrices");
  ColsumArgs a;
  long long blocks = 0;
  for (int i = 0; i < 5; ++i) {
    a.x[i] = i < n ? static_cast<const __nv_bfloat16*>(xs[i]) : nullptr;
    a.rows[i] = i < n ? rows[i] : 0;
    a.block_begin[i] = blocks;
    if (i < n) blocks += (rows[i] + kColsumRows - 1) / kColsumRows;
  }
  a.block_begin[5] = blocks;
  a.n = n;
  if (blocks == 0) return 0;
  if (deterministic()) {
    colsum_det_kernel<<<(cols + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(a, cols, ld, out_accum);
    return check_launch("colsum_det");
  }
  dim3 grid(static_cast<unsigned>(blocks), (cols + 255) / 256);
  launch_pdl(colsum_kernel, grid, dim3(256), 0, static_cast<cudaStream_t>(stream), a, cols, ld, out_accum);
  return check_launch("colsum");
}

int b200_colsum(const void* x, long long rows, int cols, int ld, float* out_accum, void* stream) {
  return b200_colsum_multi(&x, &rows, 1, cols, ld, out_accum, stream);
}

int b200_preprocess_scaled(const uint8_t* src, void* dst, int N, int H, int W, int C, int ldc, const float* mean,
                           const int* img_hw, float scale, void* stream) {
  const long long pixels = static_cast<long long>(N) * H * W;
  preprocess_kernel<<<grid_for(pixels, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      src, static_cast<__nv_bfloat16*>(dst), pixels, H, W, C, ldc, mean[0], C > 1 ? mean[1] : 0.f, C > 2 ? mean[2] : 0.f,
      C > 3 ? mean[3] : 0.f, img_hw, scale);
  return check_launch("preprocess");
}

int b200_preprocess(const uint8_t* src, void* dst, int N, int H, int W, int C, int ldc, const float* mean,
                    const int* img_hw, void* stream) {
  return b200_preprocess_scaled(src, dst, N, H, W, C, ldc, mean, img_hw, 1.f, stream);
}

}  // extern "C"
