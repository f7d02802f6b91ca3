// Layer types of the MobileNet / DenseNet backbones that the ResNet flagship does not have, as bandwidth-bound sm_100a
// kernels on NHWC bf16 activations (the 1x1 / 3x3 dense convolutions of these backbones run on the tcgen05 kernels):
//   * depthwise 3x3 convolution, stride 1 / 2, TensorFlow `same` padding: forward, input gradient, filter gradient;
//   * trainable batch normalisation with BATCH statistics (keras.applications BN, momentum 0.99): per-channel sums,
//     finalise (+ moving-statistics update), normalise + affine + ReLU / ReLU6, and the backward pair (reductions for
//     dgamma / dbeta and the two projection terms, then dx);
//   * 2x2 average pooling (DenseNet transitions) forward / backward;
//   * filter gradient of a dense convolution with very few input channels (Cin <= 8: the first layer of MobileNet and
//     VGG, where a 64-channel MN-major tensor-core operand would be 95 % padding).
// Reachable from the reference through --backbone (train.py:323-324,390); the reference runs these
// through cuDNN / TF elementwise ops and ships no CUDA source.
#include "../include/b200_ops.h"
#include "common.cuh"
#include "ptx_sm100.cuh"

namespace b200 {

__device__ __forceinline__ void ld8(const __nv_bfloat16* p, float (&v)[8]) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  v[0] = bf16_lo(u.x); v[1] = bf16_hi(u.x); v[2] = bf16_lo(u.y); v[3] = bf16_hi(u.y);
  v[4] = bf16_lo(u.z); v[5] = bf16_hi(u.z); v[6] = bf16_lo(u.w); v[7] = bf16_hi(u.w);
}
__device__ __forceinline__ void st8(__nv_bfloat16* p, const float (&v)[8]) {
  uint4 u;
  u.x = pack_bf16x2(v[0], v[1]); u.y = pack_bf16x2(v[2], v[3]);
  u.z = pack_bf16x2(v[4], v[5]); u.w = pack_bf16x2(v[6], v[7]);
  *reinterpret_cast<uint4*>(p) = u;
}

// ------------------------------------------------------------------ thread mapping
// Every kernel below gives a thread ONE group of 8 channels for its whole life (per-channel constants - filter taps,
// BN scale / shift - sit in registers) and lets it walk pixels.  A block is `cgc` channel groups x `pl` pixel lanes with
// cgc = min(C / 8, 32): narrow layers (MobileNet's first blocks have 32 channels = 4 groups) still fill every lane of
// a warp.  blockIdx.y picks the block of cgc * 8 channels, blockIdx.x a contiguous range of `rpb` pixels.
struct Tile {
  int cgc, pl;
  unsigned gx, gy, rpb;
};

static Tile tile_for(long long rows, int C, int threads, int lanes_per_group = 1) {
  Tile t;
  const int cg = C / 8;
  t.cgc = cg < 32 ? cg : 32;
  t.pl = threads / (t.cgc * lanes_per_group);
  t.gy = static_cast<unsigned>((cg + t.cgc - 1) / t.cgc);
  long long want = static_cast<long long>(num_sms()) * 8 / t.gy;
  if (want < 1) want = 1;
  long long rpb = (rows + want - 1) / want;
  if (rpb < t.pl * 4LL) rpb = t.pl * 4LL;          // a thread sees at least four pixels before its tail reduction
  t.rpb = static_cast<unsigned>(rpb);
  t.gx = static_cast<unsigned>((rows + rpb - 1) / rpb);
  if (t.gx < 1) t.gx = 1;
  return t;
}

// (n, h, w) of a pixel index, advanced without divisions inside the pixel loop
struct PixIter {
  int n, h, w;
  __device__ __forceinline__ PixIter(unsigned p, int H, int W) {
    w = static_cast<int>(p % static_cast<unsigned>(W));
    const unsigned t = p / static_cast<unsigned>(W);
    h = static_cast<int>(t % static_cast<unsigned>(H));
    n = static_cast<int>(t / static_cast<unsigned>(H));
  }
  __device__ __forceinline__ void advance(int step, int H, int W) {
    w += step;
    while (w >= W) {
      w -= W;
      if (++h == H) {
        h = 0;
        ++n;
      }
    }
  }
};

// the 72 filter taps of 8 consecutive channels ([C][9] fp32, 288 contiguous bytes) into registers: wr[e * 9 + tap]
__device__ __forceinline__ void load_taps(const float* __restrict__ w, int c, float (&wr)[72]) {
  const float4* p = reinterpret_cast<const float4*>(w + static_cast<size_t>(c) * 9);
#pragma unroll
  for (int k = 0; k < 18; ++k) {
    const float4 v = __ldg(p + k);
    wr[4 * k] = v.x; wr[4 * k + 1] = v.y; wr[4 * k + 2] = v.z; wr[4 * k + 3] = v.w;
  }
}

// ------------------------------------------------------------------ depthwise 3x3
// w fp32 [C][9] (PyTorch (C,1,3,3) layout, read from the fp32 masters: 9 C floats, no working copy needed)
__global__ void __launch_bounds__(256) dw3x3_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ w,
                                                        __nv_bfloat16* __restrict__ y, int N, int H, int W, int C, int OH,
                                                        int OW, int stride, int pad_h, int pad_w, int cgc, unsigned rpb) {
  const int pl = blockDim.x / cgc, cgx = threadIdx.x % cgc, lp = threadIdx.x / cgc;
  const int c = (blockIdx.y * cgc + cgx) * 8;
  const unsigned npix = static_cast<unsigned>(N) * OH * OW;
  const unsigned p0 = blockIdx.x * rpb, pend = min(npix, p0 + rpb);
  unsigned p = p0 + lp;
  if (c >= C || lp >= pl || p >= pend) return;
  float wr[72];
  load_taps(w, c, wr);
  PixIter it(p, OH, OW);
  for (; p < pend; p += pl, it.advance(pl, OH, OW)) {
    float acc[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = 0.f;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int h = it.h * stride - pad_h + r;
      if (h < 0 || h >= H) continue;
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const int ww = it.w * stride - pad_w + s;
        if (ww < 0 || ww >= W) continue;
        float v[8];
        ld8(x + (static_cast<size_t>(it.n * H + h) * W + ww) * C + c, v);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = fmaf(v[e], wr[e * 9 + r * 3 + s], acc[e]);
      }
    }
    st8(y + static_cast<size_t>(p) * C + c, acc);
  }
}

// dx[n,h,w,c] = sum over taps (r, s) with (h + pad_h - r, w + pad_w - s) on the stride grid of
//               dy[n, (h + pad_h - r) / stride, (w + pad_w - s) / stride, c] * w[c][r][s]
__global__ void __launch_bounds__(256) dw3x3_dgrad_kernel(const __nv_bfloat16* __restrict__ dy, const float* __restrict__ w,
                                                          __nv_bfloat16* __restrict__ dx, int N, int H, int W, int C, int OH,
                                                          int OW, int stride, int pad_h, int pad_w, int cgc, unsigned rpb) {
  const int pl = blockDim.x / cgc, cgx = threadIdx.x % cgc, lp = threadIdx.x / cgc;
  const int c = (blockIdx.y * cgc + cgx) * 8;
  const unsigned npix = static_cast<unsigned>(N) * H * W;
  const unsigned p0 = blockIdx.x * rpb, pend = min(npix, p0 + rpb);
  unsigned p = p0 + lp;
  if (c >= C || lp >= pl || p >= pend) return;
  float wr[72];
  load_taps(w, c, wr);
  PixIter it(p, H, W);
  for (; p < pend; p += pl, it.advance(pl, H, W)) {
    float acc[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = 0.f;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int th = it.h + pad_h - r;
      if (th < 0 || (stride == 2 && (th & 1))) continue;
      const int oh = stride == 2 ? th >> 1 : th;
      if (oh >= OH) continue;
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const int tw = it.w + pad_w - s;
        if (tw < 0 || (stride == 2 && (tw & 1))) continue;
        const int ow = stride == 2 ? tw >> 1 : tw;
        if (ow >= OW) continue;
        float g[8];
        ld8(dy + (static_cast<size_t>(it.n * OH + oh) * OW + ow) * C + c, g);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = fmaf(g[e], wr[e * 9 + r * 3 + s], acc[e]);
      }
    }
    st8(dx + static_cast<size_t>(p) * C + c, acc);
  }
}

// dw[c][r][s] += sum over output pixels of dy[pix, c] * x[pix * stride - pad + (r, s), c].
// A thread owns (8 channels, one filter ROW r, one pixel lane): 24 register accumulators, one dy and three x vectors
// per pixel.  The pixel lanes are reduced through shared memory, then one atomic per (channel, tap) per block.
constexpr int kDwWgThreads = 384;
__global__ void __launch_bounds__(kDwWgThreads) dw3x3_wgrad_kernel(const __nv_bfloat16* __restrict__ x,
                                                                   const __nv_bfloat16* __restrict__ dy,
                                                                   float* __restrict__ dw, int N, int H, int W, int C, int OH,
                                                                   int OW, int stride, int pad_h, int pad_w, int cgc,
                                                                   unsigned rpb) {
  __shared__ float red[kDwWgThreads * 24];
  const int pl = kDwWgThreads / (cgc * 3);
  const int cgx = threadIdx.x % cgc, r = (threadIdx.x / cgc) % 3, lp = threadIdx.x / (cgc * 3);
  const int c = (blockIdx.y * cgc + cgx) * 8;
  const unsigned npix = static_cast<unsigned>(N) * OH * OW;
  const unsigned p0 = blockIdx.x * rpb, pend = min(npix, p0 + rpb);
  float acc[3][8];
#pragma unroll
  for (int s = 0; s < 3; ++s)
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[s][e] = 0.f;
  unsigned p = p0 + lp;
  if (c < C && lp < pl && p < pend) {
    PixIter it(p, OH, OW);
    for (; p < pend; p += pl, it.advance(pl, OH, OW)) {
      const int h = it.h * stride - pad_h + r;
      if (h < 0 || h >= H) continue;
      float g[8];
      ld8(dy + static_cast<size_t>(p) * C + c, g);
      const __nv_bfloat16* xr = x + static_cast<size_t>(it.n * H + h) * W * C + c;
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const int ww = it.w * stride - pad_w + s;
        if (ww < 0 || ww >= W) continue;
        float v[8];
        ld8(xr + static_cast<size_t>(ww) * C, v);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[s][e] = fmaf(g[e], v[e], acc[s][e]);
      }
    }
  }
  const int nout = cgc * 72;          // (channel group, row, tap, channel) results of this block
  if (lp < pl) {
#pragma unroll
    for (int s = 0; s < 3; ++s)
#pragma unroll
      for (int e = 0; e < 8; ++e) red[lp * nout + (cgx * 3 + r) * 24 + s * 8 + e] = acc[s][e];
  }
  __syncthreads();
  for (int o = threadIdx.x; o < nout; o += kDwWgThreads) {
    float sum = 0.f;
    for (int k = 0; k < pl; ++k) sum += red[k * nout + o];
    const int g2 = o / 72, rem = o % 72;
    const int cc = (blockIdx.y * cgc + g2) * 8 + (rem & 7);
    if (cc < C && sum != 0.f) atomicAdd(dw + static_cast<size_t>(cc) * 9 + (rem / 24) * 3 + ((rem % 24) >> 3), sum);
  }
}

// ------------------------------------------------------------------ batch normalisation with batch statistics
// Tail of the two reduction kernels: per-thread partial sums of 8 channels -> one atomic per channel per block.
__device__ __forceinline__ void bn_block_reduce(const float (&s1)[8], const float (&s2)[8], int cgc, int pl, int C,
                                                float* __restrict__ sums) {
  __shared__ float part[2][256][8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    part[0][threadIdx.x][e] = s1[e];
    part[1][threadIdx.x][e] = s2[e];
  }
  __syncthreads();
  if (threadIdx.x < cgc * 8) {          // thread index = lane * cgc + group
    const int g2 = threadIdx.x >> 3, e2 = threadIdx.x & 7;
    const int cc = (blockIdx.y * cgc + g2) * 8 + e2;
    if (cc < C) {
      float a = 0.f, b = 0.f;
      for (int k = 0; k < pl; ++k) {
        a += part[0][k * cgc + g2][e2];
        b += part[1][k * cgc + g2][e2];
      }
      atomicAdd(sums + cc, a);
      atomicAdd(sums + C + cc, b);
    }
  }
}

// sums[0..C) += sum x, sums[C..2C) += sum x^2 over a range of rows (fp32; the finalise kernel turns them into mean / var)
__global__ void __launch_bounds__(256) bn_sums_kernel(const __nv_bfloat16* __restrict__ x, long long rows, int C, int ld,
                                                      float* __restrict__ sums, int cgc, unsigned rpb) {
  const int pl = blockDim.x / cgc, cgx = threadIdx.x % cgc, lp = threadIdx.x / cgc;
  const int c = (blockIdx.y * cgc + cgx) * 8;
  const long long r0 = static_cast<long long>(blockIdx.x) * rpb, rend = min(rows, r0 + rpb);
  float s1[8], s2[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) s1[e] = s2[e] = 0.f;
  if (c < C && lp < pl) {
    for (long long r = r0 + lp; r < rend; r += pl) {
      float v[8];
      ld8(x + r * ld + c, v);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        s1[e] += v[e];
        s2[e] = fmaf(v[e], v[e], s2[e]);
      }
    }
  }
  bn_block_reduce(s1, s2, cgc, pl, C, sums);
}

// mean / biased variance from the sums; Keras moving statistics: moving = moving * m + batch * (1 - m)
// (the moving variance takes the biased batch variance, as tf.keras does); clears the sums for the next use.
__global__ void bn_finalize_kernel(float* __restrict__ sums, long long rows, int C, float* __restrict__ mean,
                                   float* __restrict__ var, float* __restrict__ moving_mean,
                                   float* __restrict__ moving_var, float momentum) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float inv = 1.f / static_cast<float>(rows);
  const float m = sums[c] * inv;
  const float v = fmaxf(sums[C + c] * inv - m * m, 0.f);
  mean[c] = m;
  var[c] = v;
  if (moving_mean) {
    moving_mean[c] = moving_mean[c] * momentum + m * (1.f - momentum);
    moving_var[c] = moving_var[c] * momentum + v * (1.f - momentum);
  }
  sums[c] = 0.f;
  sums[C + c] = 0.f;
}

// y = act(gamma * (x - mean) * rsqrt(var + eps) + beta); act: 0 none, 1 ReLU, 2 ReLU6.  ldx / ldy: channel strides
// (y may be a slice of a wider concatenation buffer or x may be one).
__global__ void __launch_bounds__(256) bn_act_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ mean,
                                                         const float* __restrict__ var, const float* __restrict__ gamma,
                                                         const float* __restrict__ beta, float eps,
                                                         __nv_bfloat16* __restrict__ y, long long rows, int C, int ldx,
                                                         int ldy, int act, int cgc, unsigned rpb) {
  const int pl = blockDim.x / cgc, cgx = threadIdx.x % cgc, lp = threadIdx.x / cgc;
  const int c = (blockIdx.y * cgc + cgx) * 8;
  const long long r0 = static_cast<long long>(blockIdx.x) * rpb, rend = min(rows, r0 + rpb);
  if (c >= C || lp >= pl) return;
  float sc[8], sh[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    sc[e] = gamma[c + e] * rsqrtf(var[c + e] + eps);
    sh[e] = beta[c + e];
  }
  float mu[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) mu[e] = mean[c + e];
  for (long long r = r0 + lp; r < rend; r += pl) {
    float v[8];
    ld8(x + r * ldx + c, v);
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      float z = fmaf(v[e] - mu[e], sc[e], sh[e]);
      if (act >= 1) z = fmaxf(z, 0.f);
      if (act == 2) z = fminf(z, 6.f);
      v[e] = z;
    }
    st8(y + r * ldy + c, v);
  }
}

// Backward, pass 1: with gz = dy * act'(z):  sums[0..C) += sum gz (= dbeta),  sums[C..2C) += sum gz * xhat (= dgamma)
__global__ void __launch_bounds__(256) bn_act_bwd_sums_kernel(const __nv_bfloat16* __restrict__ x,
                                                              const __nv_bfloat16* __restrict__ dy,
                                                              const float* __restrict__ mean, const float* __restrict__ var,
                                                              const float* __restrict__ gamma, const float* __restrict__ beta,
                                                              float eps, long long rows, int C, int ldx, int lddy, int act,
                                                              float* __restrict__ sums, int cgc, unsigned rpb) {
  const int pl = blockDim.x / cgc, cgx = threadIdx.x % cgc, lp = threadIdx.x / cgc;
  const int c = (blockIdx.y * cgc + cgx) * 8;
  const long long r0 = static_cast<long long>(blockIdx.x) * rpb, rend = min(rows, r0 + rpb);
  float s1[8], s2[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) s1[e] = s2[e] = 0.f;
  if (c < C && lp < pl) {
    float mu[8], is[8], ga[8], be[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      mu[e] = mean[c + e];
      is[e] = rsqrtf(var[c + e] + eps);
      ga[e] = gamma[c + e];
      be[e] = beta[c + e];
    }
    for (long long r = r0 + lp; r < rend; r += pl) {
      float v[8], g[8];
      ld8(x + r * ldx + c, v);
      ld8(dy + r * lddy + c, g);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float xh = (v[e] - mu[e]) * is[e];
        const float z = fmaf(xh, ga[e], be[e]);
        const bool pass = act == 0 || (z > 0.f && (act == 1 || z < 6.f));
        const float gz = pass ? g[e] : 0.f;
        s1[e] += gz;
        s2[e] = fmaf(gz, xh, s2[e]);
      }
    }
  }
  bn_block_reduce(s1, s2, cgc, pl, C, sums);
}

// pass 2: dx = gamma * invstd * (gz - mean(gz) - xhat * mean(gz * xhat)); with `accumulate`, dx += (DenseNet: the
// gradient of a concatenation buffer collects the contributions of every later layer)
__global__ void __launch_bounds__(256) bn_act_bwd_apply_kernel(const __nv_bfloat16* __restrict__ x,
                                                               const __nv_bfloat16* __restrict__ dy,
                                                               const float* __restrict__ mean, const float* __restrict__ var,
                                                               const float* __restrict__ gamma, const float* __restrict__ beta,
                                                               float eps, const float* __restrict__ sums, long long rows,
                                                               int C, int ldx, int lddy, int lddx, int act, int accumulate,
                                                               __nv_bfloat16* __restrict__ dx, int cgc, unsigned rpb) {
  const int pl = blockDim.x / cgc, cgx = threadIdx.x % cgc, lp = threadIdx.x / cgc;
  const int c = (blockIdx.y * cgc + cgx) * 8;
  const long long r0 = static_cast<long long>(blockIdx.x) * rpb, rend = min(rows, r0 + rpb);
  if (c >= C || lp >= pl) return;
  const float invn = 1.f / static_cast<float>(rows);
  float mu[8], is[8], ga[8], be[8], k1[8], k2[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    mu[e] = mean[c + e];
    is[e] = rsqrtf(var[c + e] + eps);
    ga[e] = gamma[c + e];
    be[e] = beta[c + e];
    k1[e] = sums[c + e] * invn;
    k2[e] = sums[C + c + e] * invn;
  }
  for (long long r = r0 + lp; r < rend; r += pl) {
    float v[8], g[8], o[8];
    ld8(x + r * ldx + c, v);
    ld8(dy + r * lddy + c, g);
    if (accumulate) ld8(dx + r * lddx + c, o);
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float xh = (v[e] - mu[e]) * is[e];
      const float z = fmaf(xh, ga[e], be[e]);
      const bool pass = act == 0 || (z > 0.f && (act == 1 || z < 6.f));
      const float gz = pass ? g[e] : 0.f;
      const float d = ga[e] * is[e] * (gz - k1[e] - xh * k2[e]);
      o[e] = accumulate ? o[e] + d : d;
    }
    st8(dx + r * lddx + c, o);
  }
}

// dgamma += sums[C..2C), dbeta += sums[0..C); clears the sums
__global__ void bn_param_grads_kernel(float* __restrict__ sums, int C, float* __restrict__ dgamma, float* __restrict__ dbeta) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  if (dbeta) dbeta[c] += sums[c];
  if (dgamma) dgamma[c] += sums[C + c];
  sums[c] = 0.f;
  sums[C + c] = 0.f;
}

// ------------------------------------------------------------------ 2x2 average pool (floor sizes)
__global__ void avgpool2_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, int N, int H, int W,
                                    int C, int OH, int OW, int ldx, int ldy) {
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
    const __nv_bfloat16* b = x + ((static_cast<long long>(n) * H + 2 * oh) * W + 2 * ow) * ldx + c;
    float a[8], v[8];
    ld8(b, a);
    ld8(b + ldx, v);
#pragma unroll
    for (int e = 0; e < 8; ++e) a[e] += v[e];
    ld8(b + static_cast<long long>(W) * ldx, v);
#pragma unroll
    for (int e = 0; e < 8; ++e) a[e] += v[e];
    ld8(b + static_cast<long long>(W) * ldx + ldx, v);
#pragma unroll
    for (int e = 0; e < 8; ++e) a[e] = (a[e] + v[e]) * 0.25f;
    st8(y + ((static_cast<long long>(n) * OH + oh) * OW + ow) * ldy + c, a);
  }
}

__global__ void avgpool2_bwd_kernel(const __nv_bfloat16* __restrict__ dy, __nv_bfloat16* __restrict__ dx, int N, int H, int W,
                                    int C, int OH, int OW, int lddy, int lddx) {
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
    float g[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) g[e] = 0.f;
    if ((h >> 1) < OH && (w >> 1) < OW) {
      ld8(dy + ((static_cast<long long>(n) * OH + (h >> 1)) * OW + (w >> 1)) * lddy + c, g);
#pragma unroll
      for (int e = 0; e < 8; ++e) g[e] *= 0.25f;
    }
    st8(dx + ((static_cast<long long>(n) * H + h) * W + w) * lddx + c, g);
  }
}

// ------------------------------------------------------------------ dense 3x3 conv filter gradient, Cin <= 8
// dw fp32 [9][Cout][Cin] += sum_pix dy[pix, co] * x[pix * stride - pad + tap, ci]; x has ldx >= Cin channels.
// A thread owns one output channel and a contiguous run of one output row: the 3x3 input window lives in registers and
// slides along the row (one new column per pixel at stride 1, two at stride 2; the x loads are warp-uniform broadcasts),
// so a pixel costs one dy load + 3 (6) x loads for 9 * kCi FMAs.  Blocks walk rows persistently; the pixel lanes are
// reduced through shared memory and every block ends with one atomic per filter element.
constexpr int kWsThreads = 256;

template <int kCi>
__device__ __forceinline__ void ws_load_col(float (&dst)[kCi], const __nv_bfloat16* rowp, int col, int W, int ldx, int ci0) {
  uint2 u = make_uint2(0u, 0u);
  if (rowp != nullptr && col >= 0 && col < W) u = *reinterpret_cast<const uint2*>(rowp + static_cast<size_t>(col) * ldx + ci0);
  dst[0] = bf16_lo(u.x);
  if (kCi > 1) dst[1] = bf16_hi(u.x);
  if (kCi > 2) dst[2] = bf16_lo(u.y);
  if (kCi > 3) dst[3] = bf16_hi(u.y);
}

// Step k of a run (k % 3 == kPhase): window column j = k * stride + s lives in slot j % 3.
template <int kStride, int kPhase, int kCi>
__device__ __forceinline__ void ws_step(float (&win)[3][3][kCi], float (&acc)[9][kCi], const __nv_bfloat16* const (&rowp)[3],
                                        int col0, int W, int ldx, int ci0, float g) {
  constexpr int kNew = kStride == 1 ? 1 : 2;
#pragma unroll
  for (int jj = 3 - kNew; jj < 3; ++jj) {
    constexpr int base = kPhase * kStride;
#pragma unroll
    for (int r = 0; r < 3; ++r) ws_load_col<kCi>(win[r][(base + jj) % 3], rowp[r], col0 + jj, W, ldx, ci0);
  }
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int s = 0; s < 3; ++s)
#pragma unroll
      for (int e = 0; e < kCi; ++e)
        acc[r * 3 + s][e] = fmaf(g, win[r][(kPhase * kStride + s) % 3][e], acc[r * 3 + s][e]);
}

template <int kStride, int kCi>
__global__ void __launch_bounds__(kWsThreads, 2) wgrad_small_kernel(const __nv_bfloat16* __restrict__ x,
                                                                    const __nv_bfloat16* __restrict__ dy,
                                                                    float* __restrict__ dw, int N, int H, int W, int Cin,
                                                                    int ldx, int OH, int OW, int Cout, int ldy, int pad_h,
                                                                    int pad_w, int cl, int ci0) {
  __shared__ float red[kWsThreads * 9 * kCi];
  const int pl = kWsThreads / cl;                          // pixel lanes; cl = output-channel lanes (multiple of 32)
  const int lane = threadIdx.x % cl, lp = threadIdx.x / cl;
  const int co = blockIdx.y * cl + lane;
  const int seg = (OW + pl - 1) / pl;
  const int ow0 = lp * seg, len = min(OW, ow0 + seg) - ow0;
  float acc[9][kCi];
#pragma unroll
  for (int t = 0; t < 9; ++t)
#pragma unroll
    for (int e = 0; e < kCi; ++e) acc[t][e] = 0.f;
  if (co < Cout && len > 0) {
    for (int row = blockIdx.x; row < N * OH; row += gridDim.x) {
      const int n = row / OH, oh = row - n * OH;
      const __nv_bfloat16* rowp[3];
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const int h = oh * kStride - pad_h + r;
        rowp[r] = (h >= 0 && h < H) ? x + static_cast<size_t>(n * H + h) * W * ldx : nullptr;
      }
      const __nv_bfloat16* g = dy + (static_cast<size_t>(row) * OW + ow0) * ldy + co;
      const int col0 = ow0 * kStride - pad_w;
      float win[3][3][kCi];
#pragma unroll
      for (int jj = 0; jj < (kStride == 1 ? 2 : 1); ++jj)          // columns the first step does not load itself
#pragma unroll
        for (int r = 0; r < 3; ++r) ws_load_col<kCi>(win[r][jj], rowp[r], col0 + jj, W, ldx, ci0);
      int k = 0;
      for (; k + 3 <= len; k += 3) {
        const float g0 = __bfloat162float(g[static_cast<size_t>(k) * ldy]);
        const float g1 = __bfloat162float(g[static_cast<size_t>(k + 1) * ldy]);
        const float g2 = __bfloat162float(g[static_cast<size_t>(k + 2) * ldy]);
        ws_step<kStride, 0, kCi>(win, acc, rowp, col0 + k * kStride, W, ldx, ci0, g0);
        ws_step<kStride, 1, kCi>(win, acc, rowp, col0 + (k + 1) * kStride, W, ldx, ci0, g1);
        ws_step<kStride, 2, kCi>(win, acc, rowp, col0 + (k + 2) * kStride, W, ldx, ci0, g2);
      }
      if (k < len) ws_step<kStride, 0, kCi>(win, acc, rowp, col0 + k * kStride, W, ldx, ci0, __bfloat162float(g[static_cast<size_t>(k) * ldy]));
      if (k + 1 < len)
        ws_step<kStride, 1, kCi>(win, acc, rowp, col0 + (k + 1) * kStride, W, ldx, ci0, __bfloat162float(g[static_cast<size_t>(k + 1) * ldy]));
    }
  }
  constexpr int kPer = 9 * kCi;
#pragma unroll
  for (int t = 0; t < 9; ++t)
#pragma unroll
    for (int e = 0; e < kCi; ++e) red[(lp * kPer + t * kCi + e) * cl + lane] = acc[t][e];
  __syncthreads();
  for (int o = threadIdx.x; o < kPer * cl; o += kWsThreads) {
    float sum = 0.f;
    for (int k2 = 0; k2 < pl; ++k2) sum += red[k2 * kPer * cl + o];
    const int idx = o / cl, co2 = blockIdx.y * cl + o % cl;
    const int ci = ci0 + idx % kCi;
    if (co2 < Cout && ci < Cin && sum != 0.f) atomicAdd(dw + (static_cast<size_t>(idx / kCi) * Cout + co2) * Cin + ci, sum);
  }
}

static int grid1d(long long work, int block) {
  long long g = (work + block - 1) / block;
  const long long cap = static_cast<long long>(num_sms()) * 16;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return static_cast<int>(g);
}

}  // namespace b200

using namespace b200;
#define BF(p) static_cast<const __nv_bfloat16*>(p)
#define BFM(p) static_cast<__nv_bfloat16*>(p)

extern "C" {

int b200_dwconv3x3(int mode, const void* a, const void* b, void* out, int N, int H, int W, int C, int OH, int OW, int stride,
                   int pad_h, int pad_w, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (C % 8) return set_error(-1, "dwconv3x3: C must be a multiple of 8");
  if (stride != 1 && stride != 2) return set_error(-1, "dwconv3x3: stride must be 1 or 2");
  if (static_cast<long long>(N) * H * W >= (1LL << 31)) return set_error(-1, "dwconv3x3: more than 2^31 pixels");
  if (mode == 0) {          // a = x, b = w fp32 [C][9], out = y
    if (reinterpret_cast<uintptr_t>(b) % 16) return set_error(-1, "dwconv3x3: filter must be 16-byte aligned");
    const Tile t = tile_for(static_cast<long long>(N) * OH * OW, C, 256);
    dw3x3_fwd_kernel<<<dim3(t.gx, t.gy), 256, 0, stream>>>(BF(a), static_cast<const float*>(b), BFM(out), N, H, W, C, OH, OW,
                                                          stride, pad_h, pad_w, t.cgc, t.rpb);
    return check_launch("dw3x3_fwd");
  }
  if (mode == 1) {          // a = dy, b = w, out = dx
    if (reinterpret_cast<uintptr_t>(b) % 16) return set_error(-1, "dwconv3x3: filter must be 16-byte aligned");
    const Tile t = tile_for(static_cast<long long>(N) * H * W, C, 256);
    dw3x3_dgrad_kernel<<<dim3(t.gx, t.gy), 256, 0, stream>>>(BF(a), static_cast<const float*>(b), BFM(out), N, H, W, C, OH, OW,
                                                            stride, pad_h, pad_w, t.cgc, t.rpb);
    return check_launch("dw3x3_dgrad");
  }
  if (mode == 2) {          // a = x, b = dy, out = dw fp32 [C][9] (accumulated)
    if (deterministic()) return set_error(-1, "dwconv3x3 wgrad: no deterministic variant");
    const Tile t = tile_for(static_cast<long long>(N) * OH * OW, C, kDwWgThreads, 3);
    dw3x3_wgrad_kernel<<<dim3(t.gx, t.gy), kDwWgThreads, 0, stream>>>(BF(a), BF(b), static_cast<float*>(out), N, H, W, C, OH, OW,
                                                                     stride, pad_h, pad_w, t.cgc, t.rpb);
    return check_launch("dw3x3_wgrad");
  }
  return set_error(-1, "dwconv3x3: mode 0 / 1 / 2");
}

int b200_bn_stats(const void* x, long long rows, int C, int ld, float* sums2C, float* mean, float* var, float* moving_mean,
                  float* moving_var, float momentum, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if ((C % 8) || (ld % 8)) return set_error(-1, "bn_stats: C and ld must be multiples of 8");
  const Tile t = tile_for(rows, C, 256);
  bn_sums_kernel<<<dim3(t.gx, t.gy), 256, 0, stream>>>(BF(x), rows, C, ld, sums2C, t.cgc, t.rpb);
  int rc = check_launch("bn_sums");
  if (rc) return rc;
  bn_finalize_kernel<<<(C + 127) / 128, 128, 0, stream>>>(sums2C, rows, C, mean, var, moving_mean, moving_var, momentum);
  return check_launch("bn_finalize");
}

int b200_bn_act_fwd(const void* x, const float* mean, const float* var, const float* gamma, const float* beta, float eps,
                    void* y, long long rows, int C, int ldx, int ldy, int act, void* stream_) {
  if ((C % 8) || (ldx % 8) || (ldy % 8)) return set_error(-1, "bn_act_fwd: C / ld must be multiples of 8");
  const Tile t = tile_for(rows, C, 256);
  bn_act_fwd_kernel<<<dim3(t.gx, t.gy), 256, 0, static_cast<cudaStream_t>(stream_)>>>(BF(x), mean, var, gamma, beta, eps, BFM(y),
                                                                                     rows, C, ldx, ldy, act, t.cgc, t.rpb);
  return check_launch("bn_act_fwd");
}

int b200_bn_act_bwd(const void* x, const void* dy, const float* mean, const float* var, const float* gamma, const float* beta,
                    float eps, float* sums2C, long long rows, int C, int ldx, int lddy, int lddx, int act, int accumulate,
                    void* dx, float* dgamma, float* dbeta, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if ((C % 8) || (ldx % 8) || (lddy % 8) || (lddx % 8)) return set_error(-1, "bn_act_bwd: C / ld must be multiples of 8");
  const Tile t = tile_for(rows, C, 256);
  bn_act_bwd_sums_kernel<<<dim3(t.gx, t.gy), 256, 0, stream>>>(BF(x), BF(dy), mean, var, gamma, beta, eps, rows, C, ldx, lddy, act,
                                                              sums2C, t.cgc, t.rpb);
  int rc = check_launch("bn_act_bwd_sums");
  if (rc) return rc;
  bn_act_bwd_apply_kernel<<<dim3(t.gx, t.gy), 256, 0, stream>>>(BF(x), BF(dy), mean, var, gamma, beta, eps, sums2C, rows, C, ldx,
                                                               lddy, lddx, act, accumulate, BFM(dx), t.cgc, t.rpb);
  rc = check_launch("bn_act_bwd_apply");
  if (rc) return rc;
  bn_param_grads_kernel<<<(C + 127) / 128, 128, 0, stream>>>(sums2C, C, dgamma, dbeta);
  return check_launch("bn_param_grads");
}

int b200_avgpool2(int backward, const void* in, void* out, int N, int H, int W, int C, int ld_in, int ld_out, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if ((C % 8) || (ld_in % 8) || (ld_out % 8)) return set_error(-1, "avgpool2: C / ld must be multiples of 8");
  const int OH = H / 2, OW = W / 2;
  if (!backward) {
    if (OH * OW == 0) return 0;
    avgpool2_fwd_kernel<<<grid1d(static_cast<long long>(N) * OH * OW * (C / 8), 256), 256, 0, stream>>>(BF(in), BFM(out), N, H, W, C, OH, OW, ld_in, ld_out);
    return check_launch("avgpool2_fwd");
  }
  avgpool2_bwd_kernel<<<grid1d(static_cast<long long>(N) * H * W * (C / 8), 256), 256, 0, stream>>>(BF(in), BFM(out), N, H, W, C, OH, OW, ld_in, ld_out);
  return check_launch("avgpool2_bwd");
}

int b200_wgrad_small(const void* x, const void* dy, float* dw, int N, int H, int W, int Cin, int ldx, int OH, int OW, int Cout,
                     int ldy, int KH, int KW, int stride, int pad_h, int pad_w, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (Cin > 8 || (ldx % 4) || ldx < ((Cin + 3) & ~3)) return set_error(-1, "wgrad_small: Cin <= 8, ldx a multiple of 4 and >= Cin rounded up to 4");
  if (KH != 3 || KW != 3 || (stride != 1 && stride != 2)) return set_error(-1, "wgrad_small: 3x3 filters, stride 1 or 2");
  if (deterministic()) return set_error(-1, "wgrad_small: no deterministic variant");
  int cl = (Cout + 31) / 32 * 32;
  if (cl > kWsThreads) cl = kWsThreads;
  while (kWsThreads % cl) cl += 32;          // 32, 64, 128 or 256 lanes
  const long long rows = static_cast<long long>(N) * OH;
  long long gx = static_cast<long long>(num_sms()) * 2;
  if (gx > rows) gx = rows;
  dim3 grid(static_cast<unsigned>(gx), (Cout + cl - 1) / cl);
  for (int ci0 = 0; ci0 < Cin; ci0 += 4) {
    const bool three = Cin - ci0 <= 3;
#define WS_LAUNCH(S, CI) \
  wgrad_small_kernel<S, CI><<<grid, kWsThreads, 0, stream>>>(BF(x), BF(dy), dw, N, H, W, Cin, ldx, OH, OW, Cout, ldy, pad_h, pad_w, cl, ci0)
    if (stride == 1) {
      if (three) WS_LAUNCH(1, 3); else WS_LAUNCH(1, 4);
    } else {
      if (three) WS_LAUNCH(2, 3); else WS_LAUNCH(2, 4);
    }
#undef WS_LAUNCH
    const int rc = check_launch("wgrad_small");
    if (rc) return rc;
  }
  return 0;
}

}  // extern "C"
