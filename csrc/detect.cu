// Inference tail: score threshold + candidate compaction, sort, box decode + clip, class-aware
// bitmask NMS, top-k and -1 padding -- keras-retinanet's RegressBoxes / ClipBoxes /
// FilterDetections as built into the prediction graph at reference train.py:95
// (SURVEY.md B6-B8, G17, G18).  In the reference this is a TF map_fn over images with a CPU NMS
// kernel per class; here it is a handful of launches for the whole batch.
#include "../include/b200_ops.h"
#include "common.cuh"
#include "ptx_sm100.cuh"

namespace b200 {

constexpr int kNmsCap = 8192;  // candidates (best scores first) that enter NMS per image

struct LevelTable {
  const void* cls[5];
  const void* reg[5];
  int anchor_begin[6];
  int num_levels;
};

// Candidate selection.  Every score above the threshold is a candidate (keras-retinanet keeps them all per class);
// the bitmask NMS takes the kNmsCap best of an image.  WHICH ones must not depend on thread timing, so the cut is an
// exact top-k on the 64-bit key (score bits << 32 | ~id, ids are unique): a 4-pass radix select over 16-bit digits
// finds the k-th largest key of every image (pass p histograms digit p of the keys that match the digits chosen so
// far), then one more pass over the logits compacts the keys >= that one -- exactly min(k, #candidates) of them --
// and a bitonic sort in shared memory orders them.  No library sort, no arrival-order truncation.
constexpr int kDigitBins = 1 << 16;

__device__ __forceinline__ unsigned long long det_key(float logit, unsigned int id) {
  const float score = 1.f / (1.f + __expf(-logit));
  return (static_cast<unsigned long long>(__float_as_uint(score)) << 32) | static_cast<unsigned int>(~id);
}

// mode 0: histogram of digit `pass` among keys whose higher digits equal prefix[b];  mode 1: compact keys >= kth[b]
template <int kMode>
__global__ void det_scan_kernel(const __nv_bfloat16* __restrict__ logits, long long pixels_total, int HW, int K, int ld,
                                int a_off, float logit_thr, int pass, const unsigned long long* __restrict__ prefix,
                                int* __restrict__ hist, int* __restrict__ count, unsigned long long* __restrict__ keys) {
  const int groups = (9 * K + 7) >> 3;
  const long long total = pixels_total * groups;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total; i += stride) {
    const int g = static_cast<int>(i % groups);
    const long long pix = i / groups;
    const int b = static_cast<int>(pix / HW);
    const int p = static_cast<int>(pix - static_cast<long long>(b) * HW);
    const uint4 u = *reinterpret_cast<const uint4*>(logits + pix * ld + g * 8);
    const float x[8] = {bf16_lo(u.x), bf16_hi(u.x), bf16_lo(u.y), bf16_hi(u.y),
                        bf16_lo(u.z), bf16_hi(u.z), bf16_lo(u.w), bf16_hi(u.w)};
    const unsigned long long want = prefix[b];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int ch = g * 8 + e;
      if (ch >= 9 * K || !(x[e] > logit_thr)) continue;
      const int j = ch / K, k = ch - j * K;
      const unsigned int id = static_cast<unsigned int>((a_off + p * 9 + j)) * static_cast<unsigned int>(K) + k;
      const unsigned long long key = det_key(x[e], id);
      if (kMode == 0) {
        if (pass == 0 || (key >> (64 - 16 * pass)) == want)
          atomicAdd(hist + static_cast<long long>(b) * kDigitBins + static_cast<int>((key >> (48 - 16 * pass)) & 0xffffull), 1);
      } else {
        if (key >= want) {
          const int slot = atomicAdd(count + b, 1);            // order is fixed by the sort that follows
          if (slot < kNmsCap) keys[static_cast<long long>(b) * kNmsCap + slot] = key;
        }
      }
    }
  }
}

// One block per image: walk the digit histogram from the top until `remaining` keys are covered; extend the prefix.
__global__ void __launch_bounds__(1024) det_select_kernel(int* __restrict__ hist, unsigned long long* __restrict__ prefix,
                                                          int* __restrict__ remaining, int pass) {
  __shared__ int part[1024];
  __shared__ int s_bin, s_above;
  const int b = blockIdx.x, t = threadIdx.x;
  int* h = hist + static_cast<long long>(b) * kDigitBins;
  // thread t owns bins [hi - 63, hi], hi = 65535 - 64 t  (descending order of value)
  const int hi = kDigitBins - 1 - 64 * t;
  int sum = 0;
  for (int q = 0; q < 64; ++q) sum += h[hi - q];
  part[t] = sum;
  __syncthreads();
  if (t == 0) {
    const int need = remaining[b];
    int acc = 0, owner = -1;
    for (int q = 0; q < 1024; ++q) {
      if (acc + part[q] >= need) { owner = q; break; }
      acc += part[q];
    }
    int bin = 0;                 // fewer candidates than needed: everything passes (k-th key = 0)
    if (owner >= 0 && need > 0) {
      const int ohi = kDigitBins - 1 - 64 * owner;
      for (int q = 0; q < 64; ++q) {
        if (acc + h[ohi - q] >= need) { bin = ohi - q; break; }
        acc += h[ohi - q];
      }
    } else {
      acc = 0;
    }
    s_bin = bin;
    s_above = (owner >= 0 && need > 0) ? acc : 0;
  }
  __syncthreads();
  if (t == 0) {
    prefix[b] = (prefix[b] << 16) | static_cast<unsigned long long>(s_bin);
    remaining[b] = remaining[b] - s_above;          // keys still to take inside the chosen bin
  }
  for (int q = t; q < kDigitBins; q += 1024) h[q] = 0;     // ready for the next pass
}

// Descending bitonic sort of one image's <= kNmsCap keys in shared memory (unused slots hold 0 and sink to the end).
__global__ void __launch_bounds__(1024) det_sort_kernel(const unsigned long long* __restrict__ keys, const int* __restrict__ count,
                                                        unsigned long long* __restrict__ sorted) {
  extern __shared__ unsigned long long sk[];
  const int b = blockIdx.x;
  const int n = min(count[b], kNmsCap);
  int m = 64;
  while (m < n) m <<= 1;                                   // power of two >= n
  for (int i = threadIdx.x; i < m; i += blockDim.x) sk[i] = i < n ? keys[static_cast<long long>(b) * kNmsCap + i] : 0ull;
  __syncthreads();
  for (int k = 2; k <= m; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < m; i += blockDim.x) {
        const int l = i ^ j;
        if (l > i) {
          const unsigned long long a = sk[i], c = sk[l];
          const bool desc = (i & k) == 0;
          if (desc ? a < c : a > c) {
            sk[i] = c;
            sk[l] = a;
          }
        }
      }
      __syncthreads();
    }
  }
  for (int i = threadIdx.x; i < n; i += blockDim.x) sorted[static_cast<long long>(b) * kNmsCap + i] = sk[i];
}

__global__ void decode_kernel(const unsigned long long* __restrict__ keys_sorted, const int* __restrict__ count,
                              int max_cand, int K, LevelTable lt, int ld_reg, const float* __restrict__ anchors,
                              int img_h, int img_w, float* __restrict__ boxes, float* __restrict__ scores,
                              int* __restrict__ labels) {
  const int b = blockIdx.y;
  const int n = min(min(count[b], max_cand), kNmsCap);
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const unsigned long long key = keys_sorted[static_cast<long long>(b) * max_cand + i];
  const unsigned int id = ~static_cast<unsigned int>(key & 0xffffffffu);
  const int aidx = id / K, k = id - aidx * K;
  int lv = 0;
  for (int l = 1; l < lt.num_levels; ++l)
    if (aidx >= lt.anchor_begin[l]) lv = l;
  const int la = aidx - lt.anchor_begin[lv];
  const int hw = (lt.anchor_begin[lv + 1] - lt.anchor_begin[lv]) / 9;
  const int p = la / 9, j = la - p * 9;
  const __nv_bfloat16* r = static_cast<const __nv_bfloat16*>(lt.reg[lv]) + (static_cast<long long>(b) * hw + p) * ld_reg + j * 4;
  const float4 an = reinterpret_cast<const float4*>(anchors)[aidx];
  const float w = an.z - an.x, h = an.w - an.y;
  float4 o;
  o.x = fminf(fmaxf(an.x + __bfloat162float(r[0]) * 0.2f * w, 0.f), static_cast<float>(img_w));
  o.y = fminf(fmaxf(an.y + __bfloat162float(r[1]) * 0.2f * h, 0.f), static_cast<float>(img_h));
  o.z = fminf(fmaxf(an.z + __bfloat162float(r[2]) * 0.2f * w, 0.f), static_cast<float>(img_w));
  o.w = fminf(fmaxf(an.w + __bfloat162float(r[3]) * 0.2f * h, 0.f), static_cast<float>(img_h));
  const long long oi = static_cast<long long>(b) * kNmsCap + i;
  reinterpret_cast<float4*>(boxes)[oi] = o;
  scores[oi] = __uint_as_float(static_cast<unsigned int>(key >> 32));
  labels[oi] = k;
}

__device__ __forceinline__ bool iou_gt(const float4& a, const float4& b, float thr) {
  const float iw = fminf(a.z, b.z) - fmaxf(a.x, b.x);
  const float ih = fminf(a.w, b.w) - fmaxf(a.y, b.y);
  if (iw <= 0.f || ih <= 0.f) return false;
  const float inter = iw * ih;
  const float ua = (a.z - a.x) * (a.w - a.y) + (b.z - b.x) * (b.w - b.y) - inter;
  return inter > thr * ua;
}

// mask[i][cb] bit j: box (cb*64+j) comes after i, same class (if labels), IoU > thr.
__global__ void nms_mask_kernel(const float* __restrict__ boxes, const int* __restrict__ labels,
                                const int* __restrict__ count, int fixed_n, int max_cand, long long box_stride,
                                float thr, unsigned long long* __restrict__ mask, int words) {
  const int b = blockIdx.z;
  const int n = count ? min(min(count[b], max_cand), kNmsCap) : fixed_n;
  const int rb = blockIdx.y, cb = blockIdx.x;
  if (rb * 64 >= n || cb * 64 >= n || cb < rb) return;
  __shared__ float4 cbox[64];
  __shared__ int clab[64];
  const float4* bx = reinterpret_cast<const float4*>(boxes) + b * box_stride;
  const int* lb = labels ? labels + b * box_stride : nullptr;
  const int cj = cb * 64 + threadIdx.x;
  if (cj < n) {
    cbox[threadIdx.x] = bx[cj];
    clab[threadIdx.x] = lb ? lb[cj] : 0;
  }
  __syncthreads();
  const int i = rb * 64 + threadIdx.x;
  if (i < n) {
    const float4 me = bx[i];
    const int ml = lb ? lb[i] : 0;
    unsigned long long bits = 0;
    const int jn = min(64, n - cb * 64);
    for (int j = 0; j < jn; ++j) {
      const int gj = cb * 64 + j;
      if (gj > i && clab[j] == ml && iou_gt(me, cbox[j], thr)) bits |= 1ull << j;
    }
    mask[(static_cast<long long>(b) * box_stride + i) * words + cb] = bits;
  }
}

// One block per image: greedy scan in score order.
__global__ void nms_scan_kernel(const unsigned long long* __restrict__ mask, const int* __restrict__ count, int fixed_n,
                                int max_cand, long long box_stride, int words, int max_keep,
                                const float* __restrict__ boxes, const float* __restrict__ scores,
                                const int* __restrict__ labels, float* __restrict__ out_boxes,
                                float* __restrict__ out_scores, int* __restrict__ out_labels,
                                int* __restrict__ keep_flags) {
  extern __shared__ unsigned long long removed[];
  const int b = blockIdx.x;
  const int n = count ? min(min(count[b], max_cand), kNmsCap) : fixed_n;
  for (int w = threadIdx.x; w < words; w += blockDim.x) removed[w] = 0ull;
  __syncthreads();
  int kept = 0;
  for (int i = 0; i < n && kept < max_keep; ++i) {
    const bool dead = (removed[i >> 6] >> (i & 63)) & 1ull;
    if (!dead) {
      const unsigned long long* row = mask + (static_cast<long long>(b) * box_stride + i) * words;
      // words below i>>6 hold no bits (only later boxes are marked)
      for (int w = (i >> 6) + threadIdx.x; w < words; w += blockDim.x) {
        const int cbw = w;
        if (cbw * 64 < n) removed[w] |= row[w];
      }
      if (threadIdx.x == 0) {
        if (keep_flags) keep_flags[b * box_stride + i] = 1;
        if (out_boxes) {
          const long long o = static_cast<long long>(b) * max_keep + kept;
          reinterpret_cast<float4*>(out_boxes)[o] = reinterpret_cast<const float4*>(boxes)[b * box_stride + i];
          out_scores[o] = scores[b * box_stride + i];
          out_labels[o] = labels[b * box_stride + i];
        }
      }
      ++kept;
    }
    __syncthreads();
  }
  if (out_boxes) {
    for (int k = kept + threadIdx.x; k < max_keep; k += blockDim.x) {
      const long long o = static_cast<long long>(b) * max_keep + k;
      reinterpret_cast<float4*>(out_boxes)[o] = make_float4(-1.f, -1.f, -1.f, -1.f);
      out_scores[o] = -1.f;
      out_labels[o] = -1;
    }
  }
}

struct DetWorkspace {
  int* count;
  int* remaining;
  unsigned long long* prefix;
  int* hist;
  unsigned long long* keys;
  unsigned long long* keys_sorted;
  float* boxes;
  float* scores;
  int* labels;
  unsigned long long* mask;
  long long total;
};

static size_t align_up(size_t v) { return (v + 255) & ~static_cast<size_t>(255); }

static DetWorkspace carve(void* base, int B) {
  DetWorkspace w;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    void* p = base ? static_cast<char*>(base) + off : nullptr;
    off += align_up(bytes);
    return p;
  };
  w.count = static_cast<int*>(take(sizeof(int) * B));
  w.remaining = static_cast<int*>(take(sizeof(int) * B));
  w.prefix = static_cast<unsigned long long*>(take(8ull * B));
  w.hist = static_cast<int*>(take(4ull * B * kDigitBins));
  w.keys = static_cast<unsigned long long*>(take(8ull * B * kNmsCap));
  w.keys_sorted = static_cast<unsigned long long*>(take(8ull * B * kNmsCap));
  w.boxes = static_cast<float*>(take(16ull * B * kNmsCap));
  w.scores = static_cast<float*>(take(4ull * B * kNmsCap));
  w.labels = static_cast<int*>(take(4ull * B * kNmsCap));
  w.mask = static_cast<unsigned long long*>(take(8ull * B * kNmsCap * (kNmsCap / 64)));
  w.total = static_cast<long long>(off);
  return w;
}

__global__ void det_init_kernel(int* count, int* remaining, unsigned long long* prefix, int B) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b < B) {
    count[b] = 0;
    remaining[b] = kNmsCap;
    prefix[b] = 0ull;
  }
}

}  // namespace b200

using namespace b200;

extern "C" {

long long b200_filter_detections_workspace(int B, int max_cand) {
  (void)max_cand;       // candidates are no longer capped before the selection (see det_scan_kernel)
  return carve(nullptr, B).total;
}

int b200_filter_detections(const void* const* cls_levels, const void* const* reg_levels, const int* level_anchors,
                           int num_levels, int ld_cls, int ld_reg, int K, const float* anchors, int A_total, int B,
                           int img_h, int img_w, float score_thr, float nms_thr, int max_det, int max_cand,
                           void* workspace, long long workspace_bytes, float* out_boxes, float* out_scores,
                           int* out_labels, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (num_levels < 1 || num_levels > 5) return set_error(-1, "filter_detections: 1..5 levels");
  DetWorkspace w = carve(workspace, B);
  if (w.total > workspace_bytes) return set_error(-1, "filter_detections: workspace too small (%lld < %lld)", workspace_bytes, w.total);
  max_cand = kNmsCap;       // stride of the key arrays
  static std::atomic<unsigned long long> attr{0};   // per device
  if (device_needs_setup(attr)) {
    B200_CUDA_OK(cudaFuncSetAttribute(det_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kNmsCap * 8));
    device_setup_done(attr);
  }
  B200_CUDA_OK(cudaMemsetAsync(w.hist, 0, 4ull * B * kDigitBins, stream));
  det_init_kernel<<<(B + 63) / 64, 64, 0, stream>>>(w.count, w.remaining, w.prefix, B);
  LevelTable lt;
  lt.num_levels = num_levels;
  const float logit_thr = logf(score_thr / (1.f - score_thr));
  int off = 0;
  for (int l = 0; l < num_levels; ++l) {
    lt.cls[l] = cls_levels[l];
    lt.reg[l] = reg_levels[l];
    lt.anchor_begin[l] = off;
    off += level_anchors[l];
  }
  lt.anchor_begin[num_levels] = off;
  if (off != A_total) return set_error(-1, "filter_detections: level anchors sum %d != A_total %d", off, A_total);
  auto scan = [&](int mode, int pass) -> int {
    for (int l = 0; l < num_levels; ++l) {
      const int hw = level_anchors[l] / 9;
      const long long pixels = static_cast<long long>(B) * hw;
      const long long total = pixels * ((9 * K + 7) / 8);
      long long g = (total + 255) / 256;
      if (g > num_sms() * 8LL) g = num_sms() * 8LL;
      const unsigned grid = static_cast<unsigned>(g < 1 ? 1 : g);
      const __nv_bfloat16* lg = static_cast<const __nv_bfloat16*>(cls_levels[l]);
      if (mode == 0)
        det_scan_kernel<0><<<grid, 256, 0, stream>>>(lg, pixels, hw, K, ld_cls, lt.anchor_begin[l], logit_thr, pass, w.prefix,
                                                     w.hist, w.count, w.keys);
      else
        det_scan_kernel<1><<<grid, 256, 0, stream>>>(lg, pixels, hw, K, ld_cls, lt.anchor_begin[l], logit_thr, pass, w.prefix,
                                                     w.hist, w.count, w.keys);
      int rc = check_launch("det_scan_kernel");
      if (rc) return rc;
    }
    return 0;
  };
  int rc = 0;
  for (int pass = 0; pass < 4; ++pass) {
    if ((rc = scan(0, pass))) return rc;
    det_select_kernel<<<B, 1024, 0, stream>>>(w.hist, w.prefix, w.remaining, pass);
    if ((rc = check_launch("det_select_kernel"))) return rc;
  }
  if ((rc = scan(1, 0))) return rc;
  det_sort_kernel<<<B, 1024, kNmsCap * 8, stream>>>(w.keys, w.count, w.keys_sorted);
  if ((rc = check_launch("det_sort_kernel"))) return rc;
  dim3 dgrid((kNmsCap + 255) / 256, B);
  decode_kernel<<<dgrid, 256, 0, stream>>>(w.keys_sorted, w.count, max_cand, K, lt, ld_reg, anchors, img_h, img_w,
                                           w.boxes, w.scores, w.labels);
  rc = check_launch("decode_kernel");
  if (rc) return rc;
  const int words = kNmsCap / 64;
  dim3 mgrid(words, words, B);
  nms_mask_kernel<<<mgrid, 64, 0, stream>>>(w.boxes, w.labels, w.count, 0, max_cand, kNmsCap, nms_thr, w.mask, words);
  rc = check_launch("nms_mask_kernel");
  if (rc) return rc;
  nms_scan_kernel<<<B, 128, words * sizeof(unsigned long long), stream>>>(
      w.mask, w.count, 0, max_cand, kNmsCap, words, max_det, w.boxes, w.scores, w.labels, out_boxes, out_scores,
      out_labels, nullptr);
  return check_launch("nms_scan_kernel");
}

/* boxes fp32 [n,4] already sorted by descending score; cls optional int32 [n];
 * mask_ws: n * ceil(n/64) uint64; keep_flags int32 [n] must be zeroed by the caller. */
int b200_nms_sorted(const float* boxes, const int* cls_or_null, int n, float thr, unsigned long long* mask_ws,
                    int* keep_flags, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (n <= 0) return 0;
  const int words = (n + 63) / 64;
  if (words * 8 > 200 * 1024) return set_error(-1, "nms_sorted: too many boxes (%d)", n);
  dim3 mgrid(words, words, 1);
  nms_mask_kernel<<<mgrid, 64, 0, stream>>>(boxes, cls_or_null, nullptr, n, 0, n, thr, mask_ws, words);
  int rc = check_launch("nms_mask_kernel");
  if (rc) return rc;
  static std::atomic<unsigned long long> attr{0};   // per device
  if (device_needs_setup(attr)) {
    B200_CUDA_OK(cudaFuncSetAttribute(nms_scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    device_setup_done(attr);
  }
  nms_scan_kernel<<<1, 128, words * sizeof(unsigned long long), stream>>>(mask_ws, nullptr, n, 0, n, words, n, boxes,
                                                                          nullptr, nullptr, nullptr, nullptr, nullptr,
                                                                          keep_flags);
  return check_launch("nms_scan_kernel");
}


/* ---------------------------------------------------------------------------------------------
 * b200_nms_sorted_noalloc -- the same greedy NMS as b200_nms_sorted, for callers that are not
 * allowed to allocate device memory.
 *
 * Who that is: code that runs after a deployment has handed out its whole device-memory budget
 * (a serving process whose memory pool belongs to someone else, a plugin loaded into a host
 * application that forbids cudaMalloc from plugins), and code on a latency-critical path where
 * cudaMalloc/cudaFree would stall the device or take a driver lock at an unpredictable moment.
 * Such callers cannot size and pass the n x ceil(n/64) suppression-mask scratch that
 * b200_nms_sorted needs, so this variant keeps that scratch inside the library:
 *
 *   - the mask lives in a file-scope __device__ array of kNmsNoAllocMax x kNmsNoAllocMax/64
 *     64-bit words = 2 MiB.  Every GPU that loads libb200ops.so holds those 2 MiB until the
 *     process exits, whether or not this function is ever called;
 *   - n must be <= kNmsNoAllocMax (4096); larger inputs are rejected with an error, nothing runs;
 *   - the function calls no cudaMalloc / cudaFree / allocating library routine.  The first call
 *     on each device looks the array's device address up once (cudaGetSymbolAddress, with that
 *     device current) and caches it; make that first call before starting any stream capture;
 *   - the caller still owns inputs and outputs: boxes [n,4] fp32 sorted by descending score,
 *     optional int32 classes [n], keep_flags int32 [n].  keep_flags is cleared here (the scan only
 *     writes ones), so unlike b200_nms_sorted the caller need not zero it;
 *   - the scratch keeps whatever the previous call left in it.  That is harmless: every mask word
 *     the scan reads for this n is rewritten by this call's mask kernel first, so nothing has to
 *     start at zero.  But ALL calls share the one array per device: two calls whose kernels may
 *     overlap on the GPU (different streams, or different host threads) race on it.  Serialise
 *     them on one stream.
 * No counterpart in the reference (it has no native code at all).
 */
namespace b200 {
constexpr int kNmsNoAllocMax = 4096;
static __device__ __align__(256) unsigned long long g_nms_noalloc_mask[kNmsNoAllocMax * (kNmsNoAllocMax / 64)];
}  // namespace b200

int b200_nms_sorted_noalloc_max(void) { return b200::kNmsNoAllocMax; }

int b200_nms_sorted_noalloc(const float* boxes, const int* cls_or_null, int n, float thr, int* keep_flags,
                            void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (n < 0) return set_error(-1, "nms_sorted_noalloc: negative n");
  if (n > kNmsNoAllocMax)
    return set_error(-1, "nms_sorted_noalloc: n = %d exceeds the built-in scratch limit of %d boxes; "
                         "use b200_nms_sorted with caller-provided scratch", n, kNmsNoAllocMax);
  if (n == 0) return 0;
  // per-device address of the static scratch (the host-side symbol name is NOT a device pointer)
  constexpr int kMaxDevices = 64;
  static unsigned long long* s_scratch[kMaxDevices] = {nullptr};
  static std::atomic<int> s_lock{0};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return set_error(-1, "nms_sorted_noalloc: device index %d out of range", dev);
  unsigned long long* scratch = s_scratch[dev];
  if (!scratch) {
    while (s_lock.exchange(1, std::memory_order_acquire)) {
    }
    cudaError_t e = cudaSuccess;
    if (!s_scratch[dev]) {
      void* p = nullptr;
      e = cudaGetSymbolAddress(&p, g_nms_noalloc_mask);
      if (e == cudaSuccess) s_scratch[dev] = static_cast<unsigned long long*>(p);
    }
    scratch = s_scratch[dev];
    s_lock.store(0, std::memory_order_release);
    if (e != cudaSuccess || !scratch)
      return set_error(-3, "nms_sorted_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
  }
  B200_CUDA_OK(cudaMemsetAsync(keep_flags, 0, sizeof(int) * static_cast<size_t>(n), stream));
  const int words = (n + 63) / 64;
  dim3 mgrid(words, words, 1);
  nms_mask_kernel<<<mgrid, 64, 0, stream>>>(boxes, cls_or_null, nullptr, n, 0, n, thr, scratch, words);
  int rc = check_launch("nms_mask_kernel(noalloc)");
  if (rc) return rc;
  nms_scan_kernel<<<1, 128, words * sizeof(unsigned long long), stream>>>(scratch, nullptr, n, 0, n, words, n, boxes,
                                                                          nullptr, nullptr, nullptr, nullptr, nullptr,
                                                                          keep_flags);
  return check_launch("nms_scan_kernel(noalloc)");
}

}  // extern "C"
