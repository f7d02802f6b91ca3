// Anchor generation, IoU target assignment, fused sigmoid-focal and smooth-L1 losses (fwd+bwd).
// Spec: batchai_retinanet_horovod_coco_b200/ops/{anchors,losses}.py, which restate
// keras-retinanet's anchor_targets_bbox / focal / smooth_l1 as bound by the reference at
// train.py:51,99-102.  The reference builds dense one-hot
// targets on the CPU (65 MB per image over PCIe); here the raw (G,5) boxes are assigned on the
// GPU and the loss kernels consume the compact (state, class, reg_target) form.
#include "../include/b200_ops.h"
#include "common.cuh"
#include "ptx_sm100.cuh"

namespace b200 {

struct BaseAnchors {
  float v[9][4];
};

__global__ void anchors_kernel(float* __restrict__ out, int H, int W, int stride, BaseAnchors base) {
  const int total = H * W * 9;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int a = i % 9;
    const int px = i / 9;
    const int x = px % W, y = px / W;
    const float cx = (x + 0.5f) * stride, cy = (y + 0.5f) * stride;
    float4 r = make_float4(base.v[a][0] + cx, base.v[a][1] + cy, base.v[a][2] + cx, base.v[a][3] + cy);
    reinterpret_cast<float4*>(out)[i] = r;
  }
}

constexpr int kMaxGT = 256;

__global__ void assign_kernel(const float* __restrict__ anchors, int A, const float* __restrict__ gt,
                              const int* __restrict__ gt_count, const int* __restrict__ img_hw, int Gmax,
                              float neg_thr, float pos_thr, int8_t* __restrict__ state, int16_t* __restrict__ cls,
                              float* __restrict__ reg, int* __restrict__ num_pos) {
  __shared__ float sgt[kMaxGT][5];
  __shared__ int spos;
  const int b = blockIdx.y;
  const int G = min(gt_count[b], min(Gmax, kMaxGT));
  if (threadIdx.x == 0) spos = 0;
  for (int i = threadIdx.x; i < G * 5; i += blockDim.x) sgt[i / 5][i % 5] = gt[(static_cast<long long>(b) * Gmax) * 5 + i];
  __syncthreads();
  const int a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a < A) {
    const float4 an = reinterpret_cast<const float4*>(anchors)[a];
    // float64 IoU with the +1 pixel convention, like the Cython compute_overlap
    const double ax1 = an.x, ay1 = an.y, ax2 = an.z, ay2 = an.w;
    const double a_area = (ax2 - ax1 + 1.0) * (ay2 - ay1 + 1.0);
    double best = -1.0;
    int bi = 0;
    for (int g = 0; g < G; ++g) {
      const double gx1 = sgt[g][0], gy1 = sgt[g][1], gx2 = sgt[g][2], gy2 = sgt[g][3];
      double iw = fmin(ax2, gx2) - fmax(ax1, gx1) + 1.0;
      double ih = fmin(ay2, gy2) - fmax(ay1, gy1) + 1.0;
      double iou = 0.0;
      if (iw > 0.0 && ih > 0.0) {
        const double ua = a_area + (gx2 - gx1 + 1.0) * (gy2 - gy1 + 1.0) - iw * ih;
        iou = iw * ih / fmax(ua, 1e-8);
      }
      if (iou > best) {
        best = iou;
        bi = g;
      }
    }
    int st = 0;
    if (G > 0) st = (best >= pos_thr) ? 1 : ((best < neg_thr) ? 0 : -1);
    const float cx = 0.5f * (an.x + an.z), cy = 0.5f * (an.y + an.w);
    if (cx >= static_cast<float>(img_hw[b * 2 + 1]) || cy >= static_cast<float>(img_hw[b * 2 + 0])) st = -1;
    const long long o = static_cast<long long>(b) * A + a;
    state[o] = static_cast<int8_t>(st);
    float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
    int c = 0;
    if (G > 0) {
      c = static_cast<int>(sgt[bi][4]);
      const float aw = an.z - an.x, ah = an.w - an.y;
      t.x = (sgt[bi][0] - an.x) / aw / 0.2f;
      t.y = (sgt[bi][1] - an.y) / ah / 0.2f;
      t.z = (sgt[bi][2] - an.z) / aw / 0.2f;
      t.w = (sgt[bi][3] - an.w) / ah / 0.2f;
    }
    cls[o] = static_cast<int16_t>(c);
    reinterpret_cast<float4*>(reg)[o] = t;
    if (st == 1) atomicAdd(&spos, 1);
  }
  __syncthreads();
  if (threadIdx.x == 0 && spos) atomicAdd(num_pos, spos);
}

__device__ __forceinline__ float softplus(float x) { return fmaxf(x, 0.f) + log1pf(__expf(-fabsf(x))); }

__device__ __forceinline__ float block_sum(float v) {
  __shared__ float red[32];
#pragma unroll
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) red[w] = v;
  __syncthreads();
  v = (threadIdx.x < (blockDim.x >> 5)) ? red[threadIdx.x] : 0.f;
  if (w == 0) {
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  }
  return v;
}

// One pass over the logits: loss (block-reduced, one atomic per block) and dlogits.
// Layout: per pixel a row of `ld` bf16 holding 9*K logits (anchor-major, class fastest).
__global__ void focal_kernel(const __nv_bfloat16* __restrict__ logits, __nv_bfloat16* __restrict__ dlogits,
                             long long pixels_total, int HW, int K, int ld, const int8_t* __restrict__ state,
                             const int16_t* __restrict__ cls, int A_total, int a_off, const int* __restrict__ num_pos,
                             float alpha, float gamma, float loss_scale, float* __restrict__ loss_accum) {
  const int groups = (9 * K + 7) >> 3;
  const long long total = pixels_total * groups;
  const float inv_norm = 1.f / fmaxf(1.f, static_cast<float>(*num_pos));
  const float gscale = loss_scale * inv_norm;
  float lsum = 0.f;
  if ((K & 7) == 0 && total < (1LL << 31)) {
    // Fast path (K a multiple of 8, e.g. COCO's 80): the 8 logits of a thread belong to ONE anchor, so its state and
    // class are read once, ignored anchors cost one store, and there is no per-element anchor / class bookkeeping.
    const unsigned ugroups = static_cast<unsigned>(groups), gpa = static_cast<unsigned>(K) >> 3;   // groups per anchor
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < static_cast<unsigned>(total); i += gridDim.x * blockDim.x) {
      const unsigned g = i % ugroups, pix = i / ugroups;
      const unsigned b = pix / static_cast<unsigned>(HW), p = pix - b * static_cast<unsigned>(HW);
      const unsigned j = g / gpa;
      const int k0 = static_cast<int>(g - j * gpa) * 8;
      const size_t off = static_cast<size_t>(pix) * ld + g * 8;
      const size_t t = static_cast<size_t>(b) * A_total + a_off + p * 9 + j;
      const int st = state[t];
      uint4 o = make_uint4(0u, 0u, 0u, 0u);
      if (st >= 0) {
        const int kpos = (st == 1) ? static_cast<int>(cls[t]) - k0 : -1;     // index of the positive class in this vector
        const uint4 u = *reinterpret_cast<const uint4*>(logits + off);
        const float x[8] = {bf16_lo(u.x), bf16_hi(u.x), bf16_lo(u.y), bf16_hi(u.y),
                            bf16_lo(u.z), bf16_hi(u.z), bf16_lo(u.w), bf16_hi(u.w)};
        float gr[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float xv = x[e];
          const float ex = __expf(-fabsf(xv));
          const float inv = __frcp_rn(1.f + ex);
          const float pr = xv >= 0.f ? inv : ex * inv;
          const float sp_pos = fmaxf(xv, 0.f) + __logf(1.f + ex);   // -log(1-p)
          const float sp_neg = sp_pos - xv;                          // -log(p)
          const bool pos = e == kpos;
          float l, d;
          if (gamma == 2.f) {
            if (pos) {
              const float q = 1.f - pr;
              l = alpha * q * q * sp_neg;
              d = alpha * q * q * (-2.f * pr * sp_neg - q);
            } else {
              l = (1.f - alpha) * pr * pr * sp_pos;
              d = (1.f - alpha) * pr * pr * (pr + 2.f * (1.f - pr) * sp_pos);
            }
          } else {
            if (pos) {
              const float q = 1.f - pr;
              const float m = __powf(fmaxf(q, 1e-12f), gamma);
              l = alpha * m * sp_neg;
              d = alpha * m * (-gamma * pr * sp_neg - q);
            } else {
              const float m = __powf(fmaxf(pr, 1e-12f), gamma);
              l = (1.f - alpha) * m * sp_pos;
              d = (1.f - alpha) * m * (pr + gamma * (1.f - pr) * sp_pos);
            }
          }
          lsum += l;
          gr[e] = d * gscale;
        }
        o.x = pack_bf16x2(gr[0], gr[1]);
        o.y = pack_bf16x2(gr[2], gr[3]);
        o.z = pack_bf16x2(gr[4], gr[5]);
        o.w = pack_bf16x2(gr[6], gr[7]);
      }
      *reinterpret_cast<uint4*>(dlogits + off) = o;
    }
  } else
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int g = static_cast<int>(i % groups);
    const long long pix = i / groups;  // b * HW + p
    const int b = static_cast<int>(pix / HW);
    const int p = static_cast<int>(pix - static_cast<long long>(b) * HW);
    const long long off = pix * ld + g * 8;
    const uint4 u = *reinterpret_cast<const uint4*>(logits + off);
    float x[8] = {bf16_lo(u.x), bf16_hi(u.x), bf16_lo(u.y), bf16_hi(u.y),
                  bf16_lo(u.z), bf16_hi(u.z), bf16_lo(u.w), bf16_hi(u.w)};
    float gr[8];
    int j_cached = -1, st = 0, cid = 0;
    int j = (g * 8) / K, k = g * 8 - j * K - 1;   // one division per thread; (j, k) then advance incrementally
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int ch = g * 8 + e;
      gr[e] = 0.f;
      if (++k == K) {
        k = 0;
        ++j;
      }
      if (ch >= 9 * K) continue;
      if (j != j_cached) {
        const long long t = static_cast<long long>(b) * A_total + a_off + p * 9 + j;
        st = state[t];
        cid = cls[t];
        j_cached = j;
      }
      if (st < 0) continue;
      const bool pos = (st == 1) && (cid == k);
      const float xv = x[e];
      // one exp shared by sigmoid and softplus: e = exp(-|x|)
      const float ex = __expf(-fabsf(xv));
      const float inv = __frcp_rn(1.f + ex);
      const float pr = xv >= 0.f ? inv : ex * inv;
      const float sp_pos = fmaxf(xv, 0.f) + __logf(1.f + ex);   // -log(1-p)
      const float sp_neg = sp_pos - xv;                          // -log(p)
      float l, d;
      if (gamma == 2.f) {
        if (pos) {
          const float q = 1.f - pr;
          l = alpha * q * q * sp_neg;
          d = alpha * q * q * (-2.f * pr * sp_neg - q);
        } else {
          l = (1.f - alpha) * pr * pr * sp_pos;
          d = (1.f - alpha) * pr * pr * (pr + 2.f * (1.f - pr) * sp_pos);
        }
      } else {
        if (pos) {
          const float q = 1.f - pr;
          const float m = __powf(fmaxf(q, 1e-12f), gamma);
          l = alpha * m * sp_neg;
          d = alpha * m * (-gamma * pr * sp_neg - q);
        } else {
          const float m = __powf(fmaxf(pr, 1e-12f), gamma);
          l = (1.f - alpha) * m * sp_pos;
          d = (1.f - alpha) * m * (pr + gamma * (1.f - pr) * sp_pos);
        }
      }
      lsum += l;
      gr[e] = d * gscale;
    }
    uint4 o;
    o.x = pack_bf16x2(gr[0], gr[1]);
    o.y = pack_bf16x2(gr[2], gr[3]);
    o.z = pack_bf16x2(gr[4], gr[5]);
    o.w = pack_bf16x2(gr[6], gr[7]);
    *reinterpret_cast<uint4*>(dlogits + off) = o;
  }
  lsum = block_sum(lsum);
  if (threadIdx.x == 0 && lsum != 0.f) atomicAdd(loss_accum, lsum * inv_norm);
}

// pred row per pixel: ld bf16 holding 9*4 offsets.  One thread per anchor.
__global__ void smooth_l1_kernel(const __nv_bfloat16* __restrict__ pred, __nv_bfloat16* __restrict__ dpred,
                                 long long anchors_total, int HW, int ld, const int8_t* __restrict__ state,
                                 const float* __restrict__ reg_target, int A_total, int a_off,
                                 const int* __restrict__ num_pos, float sigma2, float loss_scale,
                                 float* __restrict__ loss_accum) {
  const float inv_norm = 1.f / fmaxf(1.f, static_cast<float>(*num_pos));
  const float gscale = loss_scale * inv_norm;
  const int A_l = HW * 9;
  float lsum = 0.f;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < anchors_total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int b = static_cast<int>(i / A_l);
    const int a = static_cast<int>(i - static_cast<long long>(b) * A_l);
    const int p = a / 9, j = a - p * 9;
    const long long off = (static_cast<long long>(b) * HW + p) * ld + j * 4;
    const long long t = static_cast<long long>(b) * A_total + a_off + a;
    uint2 o = make_uint2(0u, 0u);
    if (state[t] == 1) {
      const uint2 u = *reinterpret_cast<const uint2*>(pred + off);
      const float4 tg = reinterpret_cast<const float4*>(reg_target)[t];
      const float x[4] = {bf16_lo(u.x), bf16_hi(u.x), bf16_lo(u.y), bf16_hi(u.y)};
      const float tv[4] = {tg.x, tg.y, tg.z, tg.w};
      float g[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float diff = x[e] - tv[e];
        const float ad = fabsf(diff);
        if (ad < 1.f / sigma2) {
          lsum += 0.5f * sigma2 * ad * ad;
          g[e] = sigma2 * diff;
        } else {
          lsum += ad - 0.5f / sigma2;
          g[e] = diff > 0.f ? 1.f : -1.f;
        }
        g[e] *= gscale;
      }
      o.x = pack_bf16x2(g[0], g[1]);
      o.y = pack_bf16x2(g[2], g[3]);
    }
    *reinterpret_cast<uint2*>(dpred + off) = o;
  }
  lsum = block_sum(lsum);
  if (threadIdx.x == 0 && lsum != 0.f) atomicAdd(loss_accum, lsum * inv_norm);
}

static int grid_cap(long long work, int block) {
  long long g = (work + block - 1) / block;
  const long long cap = static_cast<long long>(num_sms()) * 8;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return static_cast<int>(g);
}

}  // namespace b200

using namespace b200;

extern "C" {

int b200_generate_anchors(float* anchors, int H, int W, int stride, float size, void* stream) {
  // ratio-major, scale-fastest; computed in double like the NumPy original
  const double ratios[3] = {0.5, 1.0, 2.0};
  const double scales[3] = {1.0, 1.2599210498948732, 1.5874010519681994};
  BaseAnchors base;
  for (int r = 0; r < 3; ++r)
    for (int s = 0; s < 3; ++s) {
      const double side = size * scales[s];
      const double area = side * side;
      const double w = sqrt(area / ratios[r]);
      const double h = w * ratios[r];
      float* v = base.v[r * 3 + s];
      v[0] = static_cast<float>(-w / 2);
      v[1] = static_cast<float>(-h / 2);
      v[2] = static_cast<float>(w / 2);
      v[3] = static_cast<float>(h / 2);
    }
  const int total = H * W * 9;
  anchors_kernel<<<grid_cap(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(anchors, H, W, stride, base);
  return check_launch("anchors_kernel");
}

int b200_assign_targets(const float* anchors, int A, const float* gt, const int* gt_count, const int* img_hw, int B,
                        int Gmax, float neg_thr, float pos_thr, int8_t* state, int16_t* cls, float* reg_target,
                        int* num_pos, void* stream) {
  if (Gmax > kMaxGT) return set_error(-1, "assign_targets: Gmax %d > %d", Gmax, kMaxGT);
  dim3 grid((A + 255) / 256, B);
  assign_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(anchors, A, gt, gt_count, img_hw, Gmax, neg_thr,
                                                                      pos_thr, state, cls, reg_target, num_pos);
  return check_launch("assign_kernel");
}

int b200_focal_loss(const void* logits, void* dlogits, int B, int A_l, int K, int ld, const int8_t* state,
                    const int16_t* cls, int A_total, int a_off, const int* num_pos, float alpha, float gamma,
                    float loss_scale, float* loss_accum, void* stream) {
  if (A_l % 9 || ld % 8 || ld < 9 * K) return set_error(-1, "focal_loss: bad layout (A_l %d, ld %d, K %d)", A_l, ld, K);
  const int HW = A_l / 9;
  const long long pixels = static_cast<long long>(B) * HW;
  const long long total = pixels * ((9 * K + 7) / 8);
  focal_kernel<<<deterministic() ? 1 : grid_cap(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(logits), static_cast<__nv_bfloat16*>(dlogits), pixels, HW, K, ld, state, cls,
      A_total, a_off, num_pos, alpha, gamma, loss_scale, loss_accum);
  return check_launch("focal_kernel");
}

int b200_smooth_l1(const void* pred, void* dpred, int B, int A_l, int ld, const int8_t* state, const float* reg_target,
                   int A_total, int a_off, const int* num_pos, float sigma, float loss_scale, float* loss_accum,
                   void* stream) {
  if (A_l % 9 || ld % 4 || ld < 36) return set_error(-1, "smooth_l1: bad layout");
  const long long total = static_cast<long long>(B) * A_l;
  smooth_l1_kernel<<<deterministic() ? 1 : grid_cap(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(pred), static_cast<__nv_bfloat16*>(dpred), total, A_l / 9, ld, state,
      reg_target, A_total, a_off, num_pos, sigma * sigma, loss_scale, loss_accum);
  return check_launch("smooth_l1_kernel");
}

}  // extern "C"
