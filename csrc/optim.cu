// Optimizer-side kernels on FLAT fp32 buffers: sum of squares (global-norm clip), clip factor,
// Keras-semantics Adam.  The engine keeps parameters, gradients and Adam slots as single flat
// buffers, so each of these is one launch per step instead of one TF op per variable
// (reference: keras.optimizers.adam(lr=1e-5, clipnorm=0.001), train.py:104).
#include "../include/b200_ops.h"
#include "common.cuh"

namespace b200 {

__global__ void sumsq_kernel(const float* __restrict__ x, long long n, float* __restrict__ out) {
  __shared__ float red[32];
  float acc = 0.f;
  const long long n4 = n >> 2;
  const float4* x4 = reinterpret_cast<const float4*>(x);
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float4 v = x4[i];
    acc += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const float v = x[(n4 << 2) + threadIdx.x];
    acc += v * v;
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    acc = (threadIdx.x < (blockDim.x >> 5)) ? red[threadIdx.x] : 0.f;
#pragma unroll
    for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (threadIdx.x == 0) atomicAdd(out, acc);
  }
}

__global__ void clip_scale_kernel(const float* __restrict__ sumsq, float clipnorm, float* __restrict__ scale) {
  const float norm = sqrtf(*sumsq);
  // Keras clip_norm: g * c / n if n >= c else g
  *scale = (clipnorm > 0.f && norm >= clipnorm) ? clipnorm / norm : 1.f;
}

__global__ void adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                            float* __restrict__ v, long long n, const float* __restrict__ grad_scale, float lr_host,
                            float b1, float b2, float eps, const float* __restrict__ lr_dev) {
  const float gs = grad_scale ? *grad_scale : 1.f;
  const float lr_t = lr_dev ? *lr_dev : lr_host;   // device copy: the value can change between CUDA-graph replays
  const long long n4 = n >> 2;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float4 pp = reinterpret_cast<float4*>(p)[i];
    const float4 gg = reinterpret_cast<const float4*>(g)[i];
    float4 mm = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
#define B200_ADAM1(c)                              \
  {                                                \
    const float gv = gg.c * gs;                    \
    mm.c = b1 * mm.c + (1.f - b1) * gv;            \
    vv.c = b2 * vv.c + (1.f - b2) * gv * gv;       \
    pp.c -= lr_t * mm.c / (sqrtf(vv.c) + eps);     \
  }
    B200_ADAM1(x) B200_ADAM1(y) B200_ADAM1(z) B200_ADAM1(w)
#undef B200_ADAM1
    reinterpret_cast<float4*>(p)[i] = pp;
    reinterpret_cast<float4*>(m)[i] = mm;
    reinterpret_cast<float4*>(v)[i] = vv;
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const long long i = (n4 << 2) + threadIdx.x;
    const float gv = g[i] * gs;
    const float mi = b1 * m[i] + (1.f - b1) * gv;
    const float vi = b2 * v[i] + (1.f - b2) * gv * gv;
    m[i] = mi;
    v[i] = vi;
    p[i] -= lr_t * mi / (sqrtf(vi) + eps);
  }
}

}  // namespace b200

using namespace b200;

extern "C" {

int b200_sumsq(const float* x, long long n, float* out, void* stream) {
  if (reinterpret_cast<uintptr_t>(x) & 15) return set_error(-1, "sumsq: buffer must be 16-byte aligned");
  long long g = (n / 4 + 255) / 256;
  const long long cap = static_cast<long long>(num_sms()) * 8;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  if (deterministic()) g = 1;
  sumsq_kernel<<<static_cast<unsigned>(g), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, n, out);
  return check_launch("sumsq_kernel");
}

int b200_clip_scale(const float* sumsq, float clipnorm, float* scale, void* stream) {
  clip_scale_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(sumsq, clipnorm, scale);
  return check_launch("clip_scale_kernel");
}

static int adam_launch(float* p, const float* g, float* m, float* v, long long n, const float* grad_scale, float lr_t,
                       float beta1, float beta2, float eps, const float* lr_dev, void* stream) {
  if ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
       reinterpret_cast<uintptr_t>(v)) & 15)
    return set_error(-1, "adam: buffers must be 16-byte aligned");
  long long gr = (n / 4 + 255) / 256;
  const long long cap = static_cast<long long>(num_sms()) * 8;
  if (gr > cap) gr = cap;
  if (gr < 1) gr = 1;
  adam_kernel<<<static_cast<unsigned>(gr), 256, 0, static_cast<cudaStream_t>(stream)>>>(p, g, m, v, n, grad_scale, lr_t,
                                                                                       beta1, beta2, eps, lr_dev);
  return check_launch("adam_kernel");
}

int b200_adam(float* p, const float* g, float* m, float* v, long long n, const float* grad_scale, float lr_t,
              float beta1, float beta2, float eps, void* stream) {
  return adam_launch(p, g, m, v, n, grad_scale, lr_t, beta1, beta2, eps, nullptr, stream);
}

/* Same update with the bias-corrected learning rate read from device memory (for CUDA-graph replay). */
int b200_adam_devlr(float* p, const float* g, float* m, float* v, long long n, const float* grad_scale,
                    const float* lr_t_dev, float beta1, float beta2, float eps, void* stream) {
  if (!lr_t_dev) return set_error(-1, "adam: lr_t_dev is null");
  return adam_launch(p, g, m, v, n, grad_scale, 0.f, beta1, beta2, eps, lr_t_dev, stream);
}

}  // extern "C"
