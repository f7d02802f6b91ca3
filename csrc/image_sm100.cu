// Image geometry on the GPU for the real-data path: bilinear resize and affine warp of uint8 BGR images with the
// arithmetic of the OpenCV calls the reference's generator makes on the CPU (keras-retinanet utils/image.py as bound
// at reference train.py:177-214: cv2.warpAffine for --random-transform, cv2.resize to 800/1333), so that a batch
// assembled on the GPU matches the host pipeline to +-1 LSB:
//   * resize (INTER_LINEAR): half-pixel centres, 11-bit fixed-point coefficients, the 8u vertical pass
//     ((b0*(S0>>4))>>16) + ((b1*(S1>>4))>>16) + 2) >> 2;
//   * warpAffine (INTER_LINEAR): inverse map in 10-bit fixed point, source coordinates quantised to 1/32 pixel,
//     15-bit bilinear weights 32(32-fx)(32-fy)..., constant or replicated border.
// The host then only decodes; SURVEY.md G19 / K17.  No counterpart source in the reference (it calls OpenCV).
#include "../include/b200_ops.h"
#include "common.cuh"

namespace b200 {

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

// dst[y, x, c] for x < dw, y < dh; one thread per destination pixel (3 channels)
__global__ void resize_linear_u8_kernel(const uint8_t* __restrict__ src, int sh, int sw, long long src_pitch,
                                        uint8_t* __restrict__ dst, int dh, int dw, long long dst_pitch, double scale_x,
                                        double scale_y) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= dw || y >= dh) return;
  auto coef = [](int d, double scale, int ssize, int& s0, int& s1, int& c0, int& c1) {
    float f = static_cast<float>((d + 0.5) * scale - 0.5);
    int s = static_cast<int>(floorf(f));
    f -= s;
    if (s < 0) { s = 0; f = 0.f; }
    if (s >= ssize - 1) { s = ssize - 1; f = 0.f; }
    s0 = s;
    s1 = min(s + 1, ssize - 1);
    c0 = __float2int_rn((1.f - f) * 2048.f);
    c1 = __float2int_rn(f * 2048.f);
  };
  int x0, x1, a0, a1, y0, y1, b0, b1;
  coef(x, scale_x, sw, x0, x1, a0, a1);
  coef(y, scale_y, sh, y0, y1, b0, b1);
  const uint8_t* r0 = src + y0 * src_pitch;
  const uint8_t* r1 = src + y1 * src_pitch;
  uint8_t* o = dst + y * dst_pitch + x * 3;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const int s0 = r0[x0 * 3 + c] * a0 + r0[x1 * 3 + c] * a1;      // horizontal pass, scaled by 2048
    const int s1 = r1[x0 * 3 + c] * a0 + r1[x1 * 3 + c] * a1;
    const int v = (((b0 * (s0 >> 4)) >> 16) + ((b1 * (s1 >> 4)) >> 16) + 2) >> 2;
    o[c] = static_cast<uint8_t>(clampi(v, 0, 255));
  }
}

struct WarpMatrix {
  double m[6];   // INVERSE map: src = m * (x, y, 1)
};

__global__ void warp_affine_u8_kernel(const uint8_t* __restrict__ src, int sh, int sw, long long src_pitch,
                                      uint8_t* __restrict__ dst, int dh, int dw, long long dst_pitch, WarpMatrix M,
                                      int border_replicate, int cval) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= dw || y >= dh) return;
  // OpenCV: AB_BITS = 10, INTER_BITS = 5, round_delta = 1024 / 32 / 2 = 16
  const int adelta = __double2int_rn(M.m[0] * x * 1024.0), bdelta = __double2int_rn(M.m[3] * x * 1024.0);
  const int X0 = __double2int_rn((M.m[1] * y + M.m[2]) * 1024.0) + 16, Y0 = __double2int_rn((M.m[4] * y + M.m[5]) * 1024.0) + 16;
  const int X = (X0 + adelta) >> 5, Y = (Y0 + bdelta) >> 5;
  const int sx = X >> 5, sy = Y >> 5, fx = X & 31, fy = Y & 31;
  const int w00 = 32 * (32 - fx) * (32 - fy), w01 = 32 * fx * (32 - fy), w10 = 32 * (32 - fx) * fy, w11 = 32 * fx * fy;
  uint8_t* o = dst + y * dst_pitch + x * 3;
  auto px = [&](int yy, int xx, int c) -> int {
    if (border_replicate) return src[clampi(yy, 0, sh - 1) * src_pitch + clampi(xx, 0, sw - 1) * 3 + c];
    return (yy >= 0 && yy < sh && xx >= 0 && xx < sw) ? src[yy * src_pitch + xx * 3 + c] : cval;
  };
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const int v = (px(sy, sx, c) * w00 + px(sy, sx + 1, c) * w01 + px(sy + 1, sx, c) * w10 + px(sy + 1, sx + 1, c) * w11 +
                   16384) >> 15;
    o[c] = static_cast<uint8_t>(clampi(v, 0, 255));
  }
}

}  // namespace b200

using namespace b200;

extern "C" {

int b200_resize_linear_u8(const uint8_t* src, int sh, int sw, long long src_pitch, uint8_t* dst, int dh, int dw,
                          long long dst_pitch, double scale_x, double scale_y, void* stream) {
  if (sh < 1 || sw < 1 || dh < 1 || dw < 1) return set_error(-1, "resize: empty image");
  dim3 block(32, 8), grid((dw + 31) / 32, (dh + 7) / 8);
  resize_linear_u8_kernel<<<grid, block, 0, static_cast<cudaStream_t>(stream)>>>(src, sh, sw, src_pitch, dst, dh, dw,
                                                                                dst_pitch, scale_x, scale_y);
  return check_launch("resize_linear_u8");
}

int b200_warp_affine_u8(const uint8_t* src, int sh, int sw, long long src_pitch, uint8_t* dst, int dh, int dw,
                        long long dst_pitch, const double* inverse_2x3, int border_replicate, int cval, void* stream) {
  if (sh < 1 || sw < 1 || dh < 1 || dw < 1) return set_error(-1, "warp: empty image");
  WarpMatrix M;
  for (int i = 0; i < 6; ++i) M.m[i] = inverse_2x3[i];
  dim3 block(32, 8), grid((dw + 31) / 32, (dh + 7) / 8);
  warp_affine_u8_kernel<<<grid, block, 0, static_cast<cudaStream_t>(stream)>>>(src, sh, sw, src_pitch, dst, dh, dw,
                                                                              dst_pitch, M, border_replicate, cval);
  return check_launch("warp_affine_u8");
}

}  // extern "C"
