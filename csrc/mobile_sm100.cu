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

// ------------------------------------------------------------------ depthwise 3x3
// w fp32 [C][9] (PyTorch (C,1,3,3) layout, read from the fp32 masters: 9 C floats, no working copy needed)
__global__ void dw3x3_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ w,
                                 __nv_bfloat16* __restrict__ y, int N, int H, int W, int C, int OH, int OW, int stride,
                                 int pad_h, int pad_w) {
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
    float acc[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = 0.f;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int h = oh * stride - pad_h + r;
      if (h < 0 || h >= H) continue;
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const int ww = ow * stride - pad_w + s;
        if (ww < 0 || ww >= W) continue;
        float v[8];
        ld8(x + ((static_cast<long long>(n) * H + h) * W + ww) * C + c, v);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = fmaf(v[e], __ldg(w + (c + e) * 9 + r * 3 + s), acc[e]);
      }
    }
    st8(y + ((static_cast<long long>(n) * OH + oh) * OW + ow) * C + c, acc);
  }
}

// dx[n,h,w,c] = sum_{r,s : (h + pad_h - r) % stride == 0 ...} dy[n, (h+pad_h-r)/stride, (w+pad_w-s)/stride, c] * w[c][r][s]
__global__ void dw3x3_dgrad_kernel(const __nv_bfloat16* __restrict__ dy, const float* __restrict__ w,
                                   __nv_bfloat16* __restrict__ dx, int N, int H, int W, int C, int OH, int OW, int stride,
                                   int pad_h, int pad_w) {
  const int cg = C >> 3;
  const long long total = static_cast<long long>(N) * H * W * cg;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % cg) * 8;
    long long t = i / cg;
    const int ww = static_cast<int>(t % W);
    t /= W;
    const int h = static_cast<int>(t % H);
    const int n = static_cast<int>(t / H);
    float acc[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = 0.f;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int th = h + pad_h - r;
      if (th < 0 || th % stride) continue;
      const int oh = th / stride;
      if (oh >= OH) continue;
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const int tw = ww + pad_w - s;
        if (tw < 0 || tw % stride) continue;
        const int ow = tw / stride;
        if (ow >= OW) continue;
        float g[8];
        ld8(dy + ((static_cast<long long>(n) * OH + oh) * OW + ow) * C + c, g);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = fmaf(g[e], __ldg(w + (c + e) * 9 + r * 3 + s), acc[e]);
      }
    }
    st8(dx + ((static_cast<long long>(n) * H + h) * W + ww) * C + c, acc);
  }
}

// dw[c][r][s] += sum over output pixels of dy[pix, c] * x[pix * stride - pad + (r, s), c].
// Block = 32 channel groups (of 8) x 8 pixel lanes over a chunk of output pixels; 72 register accumulators per thread.
constexpr int kDwPixPerBlock = 512;
__global__ void __launch_bounds__(256) dw3x3_wgrad_kernel(const __nv_bfloat16* __restrict__ x,
                                                          const __nv_bfloat16* __restrict__ dy, float* __restrict__ dw,
                                                          int N, int H, int W, int C, int OH, int OW, int stride,
                                                          int pad_h, int pad_w) {
  const int cgx = threadIdx.x & 31, lane_p = threadIdx.x >> 5;
  const int c = (blockIdx.y * 32 + cgx) * 8;
  const long long npix = static_cast<long long>(N) * OH * OW;
  const long long p0 = static_cast<long long>(blockIdx.x) * kDwPixPerBlock;
  float acc[9][8];
#pragma unroll
  for (int t = 0; t < 9; ++t)
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[t][e] = 0.f;
  if (c < C) {
    for (long long p = p0 + lane_p; p < min(npix, p0 + kDwPixPerBlock); p += 8) {
      const int ow = static_cast<int>(p % OW);
      long long t2 = p / OW;
      const int oh = static_cast<int>(t2 % OH);
      const int n = static_cast<int>(t2 / OH);
      float g[8];
      ld8(dy + p * C + c, g);
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const int h = oh * stride - pad_h + r;
        if (h < 0 || h >= H) continue;
#pragma unroll
        for (int s = 0; s < 3; ++s) {
          const int ww = ow * stride - pad_w + s;
          if (ww < 0 || ww >= W) continue;
          float v[8];
          ld8(x + ((static_cast<long long>(n) * H + h) * W + ww) * C + c, v);
#pragma unroll
          for (int e = 0; e < 8; ++e) acc[r * 3 + s][e] = fmaf(g[e], v[e], acc[r * 3 + s][e]);
        }
      }
    }
  }
  // reduce the 8 pixel lanes through shared memory, then one atomic per (channel, tap) per block
  __shared__ float red[8][32][9];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    __syncthreads();
#pragma unroll
    for (int t = 0; t < 9; ++t) red[lane_p][cgx][t] = acc[t][e];
    __syncthreads();
    if (lane_p == 0 && c + e < C) {
#pragma unroll
      for (int t = 0; t < 9; ++t) {
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) s += red[k][cgx][t];
        if (s != 0.f) atomicAdd(dw + (c + e) * 9 + t, s);
      }
    }
  }
}

// ------------------------------------------------------------------ batch normalisation with batch statistics
// sums[0..C) += sum x, sums[C..2C) += sum x^2 over a chunk of rows (fp32; the finalise kernel turns them into mean / var)
constexpr int kBnRows = 256;
__global__ void __launch_bounds__(256) bn_sums_kernel(const __nv_bfloat16* __restrict__ x, long long rows, int C, int ld,
                                                      float* __restrict__ sums) {
  __shared__ float part[2][8][32][8];
  const int cgx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int c = (blockIdx.y * 32 + cgx) * 8;
  const long long r0 = static_cast<long long>(blockIdx.x) * kBnRows;
  float s1[8], s2[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) s1[e] = s2[e] = 0.f;
  if (c < C) {
    for (long long r = r0 + ry; r < min(rows, r0 + kBnRows); r += 8) {
      float v[8];
      ld8(x + r * ld + c, v);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        s1[e] += v[e];
        s2[e] = fmaf(v[e], v[e], s2[e]);
      }
    }
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    part[0][ry][cgx][e] = s1[e];
    part[1][ry][cgx][e] = s2[e];
  }
  __syncthreads();
  const int g2 = threadIdx.x >> 3, e2 = threadIdx.x & 7;
  const int cc = (blockIdx.y * 32 + g2) * 8 + e2;
  if (cc < C) {
    float a = 0.f, b = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      a += part[0][k][g2][e2];
      b += part[1][k][g2][e2];
    }
    atomicAdd(sums + cc, a);
    atomicAdd(sums + C + cc, b);
  }
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
__global__ void bn_act_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ mean,
                                  const float* __restrict__ var, const float* __restrict__ gamma,
                                  const float* __restrict__ beta, float eps, __nv_bfloat16* __restrict__ y,
                                  long long rows, int C, int ldx, int ldy, int act) {
  const int cg = C >> 3;
  const long long total = rows * cg;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % cg) * 8;
    const long long r = i / cg;
    float v[8];
    ld8(x + r * ldx + c, v);
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float sc = gamma[c + e] * rsqrtf(var[c + e] + eps);
      float z = fmaf(v[e] - mean[c + e], sc, beta[c + e]);
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
                                                              float* __restrict__ sums) {
  __shared__ float part[2][8][32][8];
  const int cgx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int c = (blockIdx.y * 32 + cgx) * 8;
  const long long r0 = static_cast<long long>(blockIdx.x) * kBnRows;
  float s1[8], s2[8], mu[8], is[8], ga[8], be[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    s1[e] = s2[e] = 0.f;
    const bool ok = c + e < C;
    mu[e] = ok ? mean[c + e] : 0.f;
    is[e] = ok ? rsqrtf(var[c + e] + eps) : 0.f;
    ga[e] = ok ? gamma[c + e] : 0.f;
    be[e] = ok ? beta[c + e] : 0.f;
  }
  if (c < C) {
    for (long long r = r0 + ry; r < min(rows, r0 + kBnRows); r += 8) {
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
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    part[0][ry][cgx][e] = s1[e];
    part[1][ry][cgx][e] = s2[e];
  }
  __syncthreads();
  const int g2 = threadIdx.x >> 3, e2 = threadIdx.x & 7;
  const int cc = (blockIdx.y * 32 + g2) * 8 + e2;
  if (cc < C) {
    float a = 0.f, b = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      a += part[0][k][g2][e2];
      b += part[1][k][g2][e2];
    }
    atomicAdd(sums + cc, a);
    atomicAdd(sums + C + cc, b);
  }
}

// pass 2: dx = gamma * invstd * (gz - mean(gz) - xhat * mean(gz * xhat)); with `accumulate`, dx += (DenseNet: the
// gradient of a concatenation buffer collects the contributions of every later layer)
__global__ void bn_act_bwd_apply_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                                        const float* __restrict__ mean, const float* __restrict__ var,
                                        const float* __restrict__ gamma, const float* __restrict__ beta, float eps,
                                        const float* __restrict__ sums, long long rows, int C, int ldx, int lddy,
                                        int lddx, int act, int accumulate, __nv_bfloat16* __restrict__ dx) {
  const int cg = C >> 3;
  const long long total = rows * cg;
  const float invn = 1.f / static_cast<float>(rows);
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % cg) * 8;
    const long long r = i / cg;
    float v[8], g[8], o[8];
    ld8(x + r * ldx + c, v);
    ld8(dy + r * lddy + c, g);
    if (accumulate) ld8(dx + r * lddx + c, o);
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float is = rsqrtf(var[c + e] + eps);
      const float xh = (v[e] - mean[c + e]) * is;
      const float z = fmaf(xh, gamma[c + e], beta[c + e]);
      const bool pass = act == 0 || (z > 0.f && (act == 1 || z < 6.f));
      const float gz = pass ? g[e] : 0.f;
      const float d = gamma[c + e] * is * (gz - sums[c + e] * invn - xh * sums[C + c + e] * invn);
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
                                    int C, int OH, int OW) {
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
    const __nv_bfloat16* b = x + ((static_cast<long long>(n) * H + 2 * oh) * W + 2 * ow) * C + c;
    float a[8], v[8];
    ld8(b, a);
    ld8(b + C, v);
#pragma unroll
    for (int e = 0; e < 8; ++e) a[e] += v[e];
    ld8(b + static_cast<long long>(W) * C, v);
#pragma unroll
    for (int e = 0; e < 8; ++e) a[e] += v[e];
    ld8(b + static_cast<long long>(W) * C + C, v);
#pragma unroll
    for (int e = 0; e < 8; ++e) a[e] = (a[e] + v[e]) * 0.25f;
    st8(y + ((static_cast<long long>(n) * OH + oh) * OW + ow) * C + c, a);
  }
}

__global__ void avgpool2_bwd_kernel(const __nv_bfloat16* __restrict__ dy, __nv_bfloat16* __restrict__ dx, int N, int H, int W,
                                    int C, int OH, int OW) {
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
      ld8(dy + ((static_cast<long long>(n) * OH + (h >> 1)) * OW + (w >> 1)) * C + c, g);
#pragma unroll
      for (int e = 0; e < 8; ++e) g[e] *= 0.25f;
    }
    st8(dx + ((static_cast<long long>(n) * H + h) * W + w) * C + c, g);
  }
}

// ------------------------------------------------------------------ dense conv filter gradient, Cin <= 8
// dw fp32 [KH*KW][Cout][Cin] += sum_pix dy[pix, co] * x[pix * stride - pad + tap, ci]; x has ldx >= Cin channels.
// One thread per output channel; a block walks a chunk of output pixels (x values are warp-uniform broadcasts).
constexpr int kSmallPix = 256;
template <int kTaps>
__global__ void __launch_bounds__(128) wgrad_small_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                                                          float* __restrict__ dw, int N, int H, int W, int Cin, int ldx,
                                                          int OH, int OW, int Cout, int ldy, int KW, int stride, int pad_h,
                                                          int pad_w) {
  const int co = blockIdx.y * 128 + threadIdx.x;
  const long long npix = static_cast<long long>(N) * OH * OW;
  const long long p0 = static_cast<long long>(blockIdx.x) * kSmallPix;
  float acc[kTaps][4];
  for (int half = 0; half < (Cin + 3) / 4; ++half) {     // 4 input channels per sweep keeps the accumulators in registers
#pragma unroll
    for (int t = 0; t < kTaps; ++t)
#pragma unroll
      for (int e = 0; e < 4; ++e) acc[t][e] = 0.f;
    for (long long p = p0; p < min(npix, p0 + kSmallPix); ++p) {
      const int ow = static_cast<int>(p % OW);
      long long t2 = p / OW;
      const int oh = static_cast<int>(t2 % OH);
      const int n = static_cast<int>(t2 / OH);
      const float g = co < Cout ? __bfloat162float(dy[p * ldy + co]) : 0.f;
#pragma unroll
      for (int t = 0; t < kTaps; ++t) {
        const int h = oh * stride - pad_h + t / KW, ww = ow * stride - pad_w + t % KW;
        if (h < 0 || h >= H || ww < 0 || ww >= W) continue;
        const uint2 u = *reinterpret_cast<const uint2*>(x + ((static_cast<long long>(n) * H + h) * W + ww) * ldx + half * 4);
        acc[t][0] = fmaf(g, bf16_lo(u.x), acc[t][0]);
        acc[t][1] = fmaf(g, bf16_hi(u.x), acc[t][1]);
        acc[t][2] = fmaf(g, bf16_lo(u.y), acc[t][2]);
        acc[t][3] = fmaf(g, bf16_hi(u.y), acc[t][3]);
      }
    }
    if (co < Cout) {
#pragma unroll
      for (int t = 0; t < kTaps; ++t)
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int ci = half * 4 + e;
          if (ci < Cin && acc[t][e] != 0.f) atomicAdd(dw + (static_cast<long long>(t) * Cout + co) * Cin + ci, acc[t][e]);
        }
    }
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
  if (mode == 0) {          // a = x, b = w fp32 [C][9], out = y
    dw3x3_fwd_kernel<<<grid1d(static_cast<long long>(N) * OH * OW * (C / 8), 256), 256, 0, stream>>>(
        BF(a), static_cast<const float*>(b), BFM(out), N, H, W, C, OH, OW, stride, pad_h, pad_w);
    return check_launch("dw3x3_fwd");
  }
  if (mode == 1) {          // a = dy, b = w, out = dx
    dw3x3_dgrad_kernel<<<grid1d(static_cast<long long>(N) * H * W * (C / 8), 256), 256, 0, stream>>>(
        BF(a), static_cast<const float*>(b), BFM(out), N, H, W, C, OH, OW, stride, pad_h, pad_w);
    return check_launch("dw3x3_dgrad");
  }
  if (mode == 2) {          // a = x, b = dy, out = dw fp32 [C][9] (accumulated)
    const long long npix = static_cast<long long>(N) * OH * OW;
    dim3 grid(static_cast<unsigned>((npix + kDwPixPerBlock - 1) / kDwPixPerBlock), (C / 8 + 31) / 32);
    if (deterministic()) return set_error(-1, "dwconv3x3 wgrad: no deterministic variant");
    dw3x3_wgrad_kernel<<<grid, 256, 0, stream>>>(BF(a), BF(b), static_cast<float*>(out), N, H, W, C, OH, OW, stride, pad_h, pad_w);
    return check_launch("dw3x3_wgrad");
  }
  return set_error(-1, "dwconv3x3: mode 0 / 1 / 2");
}

int b200_bn_stats(const void* x, long long rows, int C, int ld, float* sums2C, float* mean, float* var, float* moving_mean,
                  float* moving_var, float momentum, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if ((C % 8) || (ld % 8)) return set_error(-1, "bn_stats: C and ld must be multiples of 8");
  dim3 grid(static_cast<unsigned>((rows + kBnRows - 1) / kBnRows), (C / 8 + 31) / 32);
  bn_sums_kernel<<<grid, 256, 0, stream>>>(BF(x), rows, C, ld, sums2C);
  int rc = check_launch("bn_sums");
  if (rc) return rc;
  bn_finalize_kernel<<<(C + 127) / 128, 128, 0, stream>>>(sums2C, rows, C, mean, var, moving_mean, moving_var, momentum);
  return check_launch("bn_finalize");
}

int b200_bn_act_fwd(const void* x, const float* mean, const float* var, const float* gamma, const float* beta, float eps,
                    void* y, long long rows, int C, int ldx, int ldy, int act, void* stream_) {
  if ((C % 8) || (ldx % 8) || (ldy % 8)) return set_error(-1, "bn_act_fwd: C / ld must be multiples of 8");
  bn_act_fwd_kernel<<<grid1d(rows * (C / 8), 256), 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      BF(x), mean, var, gamma, beta, eps, BFM(y), rows, C, ldx, ldy, act);
  return check_launch("bn_act_fwd");
}

int b200_bn_act_bwd(const void* x, const void* dy, const float* mean, const float* var, const float* gamma, const float* beta,
                    float eps, float* sums2C, long long rows, int C, int ldx, int lddy, int lddx, int act, int accumulate,
                    void* dx, float* dgamma, float* dbeta, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if ((C % 8) || (ldx % 8) || (lddy % 8) || (lddx % 8)) return set_error(-1, "bn_act_bwd: C / ld must be multiples of 8");
  dim3 grid(static_cast<unsigned>((rows + kBnRows - 1) / kBnRows), (C / 8 + 31) / 32);
  bn_act_bwd_sums_kernel<<<grid, 256, 0, stream>>>(BF(x), BF(dy), mean, var, gamma, beta, eps, rows, C, ldx, lddy, act, sums2C);
  int rc = check_launch("bn_act_bwd_sums");
  if (rc) return rc;
  bn_act_bwd_apply_kernel<<<grid1d(rows * (C / 8), 256), 256, 0, stream>>>(BF(x), BF(dy), mean, var, gamma, beta, eps, sums2C,
                                                                          rows, C, ldx, lddy, lddx, act, accumulate, BFM(dx));
  rc = check_launch("bn_act_bwd_apply");
  if (rc) return rc;
  bn_param_grads_kernel<<<(C + 127) / 128, 128, 0, stream>>>(sums2C, C, dgamma, dbeta);
  return check_launch("bn_param_grads");
}

int b200_avgpool2(int backward, const void* in, void* out, int N, int H, int W, int C, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (C % 8) return set_error(-1, "avgpool2: C must be a multiple of 8");
  const int OH = H / 2, OW = W / 2;
  if (!backward) {
    if (OH * OW == 0) return 0;
    avgpool2_fwd_kernel<<<grid1d(static_cast<long long>(N) * OH * OW * (C / 8), 256), 256, 0, stream>>>(BF(in), BFM(out), N, H, W, C, OH, OW);
    return check_launch("avgpool2_fwd");
  }
  avgpool2_bwd_kernel<<<grid1d(static_cast<long long>(N) * H * W * (C / 8), 256), 256, 0, stream>>>(BF(in), BFM(out), N, H, W, C, OH, OW);
  return check_launch("avgpool2_bwd");
}

int b200_wgrad_small(const void* x, const void* dy, float* dw, int N, int H, int W, int Cin, int ldx, int OH, int OW, int Cout,
                     int ldy, int KH, int KW, int stride, int pad_h, int pad_w, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (Cin > 8 || (ldx % 4)) return set_error(-1, "wgrad_small: Cin <= 8 and ldx a multiple of 4");
  if (!(KH == 3 && KW == 3) && !(KH == 1 && KW == 1)) return set_error(-1, "wgrad_small: 3x3 or 1x1 filters");
  const long long npix = static_cast<long long>(N) * OH * OW;
  dim3 grid(static_cast<unsigned>((npix + kSmallPix - 1) / kSmallPix), (Cout + 127) / 128);
  if (KH == 3)
    wgrad_small_kernel<9><<<grid, 128, 0, stream>>>(BF(x), BF(dy), dw, N, H, W, Cin, ldx, OH, OW, Cout, ldy, KW, stride, pad_h, pad_w);
  else
    wgrad_small_kernel<1><<<grid, 128, 0, stream>>>(BF(x), BF(dy), dw, N, H, W, Cin, ldx, OH, OW, Cout, ldy, KW, stride, pad_h, pad_w);
  return check_launch("wgrad_small");
}

}  // extern "C"
