/* C ABI of libb200ops.so -- hand-written sm_100a kernels for the RetinaNet hot path.
 *
 * Every entry point takes raw device pointers plus the CUDA stream to launch on (as void*),
 * returns 0 on success and a negative code on failure (b200_last_error() has the text).
 * None of them synchronises the stream.  Tensors are NHWC ("channels last"); `ld*` arguments
 * are per-pixel channel strides in elements.
 *
 * The "Replaces" notes cite the reference stack's op (SURVEY.md section 2F ids) so parity can
 * be checked; the reference itself ships no CUDA code (its train.py delegates to TF/cuDNN).
 */
#ifndef B200_OPS_H_
#define B200_OPS_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* b200_last_error(void);
int b200_ops_abi_version(void);
/* 1 when B200_DETERMINISTIC=1 was set at load time: split accumulations run in one CTA per output (bit-reproducible). */
int b200_deterministic(void);
/* Number of kernel launches issued by this library in this process (bench.py gpu_launches). */
long long b200_launch_count(void);
void b200_reset_launch_count(void);
void b200_add_launch_count(long long n); /* kernels run by a CUDA-graph replay (counted at capture time) */

/* ------------------------------------------------------------------ convolution (tcgen05)
 * Implicit-GEMM convolution on the 5th-gen tensor cores: TMA tiled loads (zero-filled halos)
 * -> 128B-swizzled smem ring -> tcgen05.mma (M=128, N<=256, K=16, bf16 in / fp32 acc in TMEM)
 * -> tcgen05.ld epilogue fusing: + shift[c] (bias or frozen-BN shift; the BN scale is folded
 * into the bf16 working weights), + add (residual / gradient accumulation), * (mask > 0)
 * (ReLU backward), ReLU, bf16 store.  The add / mask operands arrive by TMA into smem chunk slots and the
 * results leave by TMA store.  Persistent, warp-specialised (operand TMA / MMA / TMEM alloc + store /
 * epilogue-operand TMA / 16 epilogue warps), double-buffered accumulators.  Up to 5 same-weight problems
 * (the FPN levels of a shared head) are scheduled in one launch.
 * Replaces: cuDNN fprop + bias-add + fused-BN + ReLU + add (G2-G4, G6, G8) and dgrad (G12). */
#define B200_MAX_PROBLEMS 5

typedef struct {
  const void* in;   /* bf16 [N, H, W, ldi] */
  void* out;        /* bf16/fp32 [N, OHo, OWo, ldo] */
  const void* add;  /* optional bf16, geometry of out */
  const void* mask; /* optional bf16, geometry of out */
  int N, H, W;      /* dims of `in` */
  int OHo, OWo;     /* spatial dims of `out` */
  float* colsum;    /* optional fp32 [Nrows]: += sum over pixels of `out` (the bias gradient of the layer whose dY
                       this output is).  Several problems may name the same vector. */
} B200ConvProblem;

typedef struct {
  int mode;         /* 0 fprop: out = conv(in, w); 1 dgrad: out = conv_transpose(in = dY, w) */
  int num_problems;
  B200ConvProblem prob[B200_MAX_PROBLEMS];
  const void* weight; /* bf16 [KH*KW][Nrows][ldw]; fprop: rows = Cout, K = Cin; dgrad: rows = Cin, K = Cout */
  int K;            /* reduction channels (channels of `in`) */
  int ldi;          /* channel stride of `in` */
  int Nrows;        /* output channels of this GEMM */
  int ldw;          /* row stride of weight, elements */
  int ldo;          /* channel stride of out / add / mask */
  int KH, KW, stride, pad_h, pad_w; /* pad_* = padding before (TF 'same' may be asymmetric) */
  int par_y, par_x; /* dgrad with stride 2: output parity class computed by this launch */
  const float* shift; /* optional fp32 [Nrows] */
  int relu;
  int out_fp32;
  int add_after_mask; /* 0: (acc+add)*mask ; 1: acc*mask + add */
} B200ConvArgs;

/* In-kernel timeline of the conv kernel: buf = device array of 16 int64 per CTA (>= 148 * 16), or NULL to switch it
 * off.  Per CTA: [0] globaltimer ns at entry, [1] clock at entry, [2] set-up done, [3] first activation box issued,
 * [4] first activation box landed, [5] first weight tile landed, [6] tile 0: last MMA issued, [13] tile 0: last
 * activation box landed, [7] tile 0: accumulator complete, [8] tile 0: epilogue done, [9] / [14] same as 6 / 7 for the
 * CTA's last tile, [10] last store read out, [11] clock at exit, [12] globaltimer at exit.  Profiling aid. */
void b200_conv_set_trace(void* buf);
int b200_conv(const B200ConvArgs* args, void* stream);

/* Weight gradient: dW[t][co][ci] += rowscale[co] * sum_pixels dY[p, co] * X[p + tap, ci]
 * (fp32 atomics; split over pixels).  Both operands are MN-major UMMA operands straight from
 * NHWC tiles.  Up to 5 problems accumulate into the same dW (shared heads).
 * Replaces: cuDNN backward-filter (G12). */
typedef struct {
  const void* x;    /* bf16 [N, H, W, ldx] forward input */
  const void* dy;   /* bf16 [N, OH, OW, ldy] gradient of conv output */
  int N, H, W, OH, OW;
} B200WgradProblem;

typedef struct {
  int num_problems;
  B200WgradProblem prob[B200_MAX_PROBLEMS];
  float* dw;        /* fp32 [KH*KW][Cout][Cin], accumulated into */
  int Cin, ldx, Cout, ldy;
  int KH, KW, stride, pad_h, pad_w;
  const float* rowscale; /* optional fp32 [Cout] (frozen-BN scale) */
  float* dbias;     /* optional fp32 [Cout]: += sum_pixels dY (computed by the tap-0 CTAs) */
} B200WgradArgs;

int b200_wgrad(const B200WgradArgs* args, void* stream);

/* Stem: 7x7/2 convolution, Cin = 3 stored as 4 channels, Cout = 64, + shift + ReLU on tcgen05.
 * x bf16 [N,H,W,4]; w_packed bf16 [64][256] with k = r*28 + s*4 + c (zero padded, BN scale folded);
 * shift fp32 [64]; y bf16 [N,ceil(H/2),ceil(W/2),64].  Producer warps gather the im2col rows into a
 * swizzled smem tile (Cin=3 cannot feed TMA).  Replaces cuDNN conv1 forward (G1). */
int b200_stem_conv(const void* x, const void* w_packed, const float* shift, void* y, int N, int H, int W,
                   void* stream);
/* Same with the ReLU optional (relu = 0: raw conv + shift, for a stem followed by a trainable batch norm: DenseNet). */
int b200_stem_conv_ex(const void* x, const void* w_packed, const float* shift, void* y, int N, int H, int W, int relu,
                      void* stream);
/* Stem weight gradient: dw fp32 [49][64][3] (tap, co, ci) += scale[co] * dY^T * im2col(x).
 * dy bf16 [N,ceil(H/2),ceil(W/2),64].  Replaces cuDNN backward-filter for conv1 (G12). */
int b200_stem_wgrad(const void* x, const void* dy, const float* scale, float* dw, int N, int H, int W, void* stream);

/* ------------------------------------------------------------------ bandwidth-bound fused ops */

/* Frozen-BN scale fold + layout transform of master weights into the bf16 working copies
 * consumed by b200_conv (table-driven, one launch for every conv of the model).
 * entry: src fp32 [KH*KW][Cout][Cin] (engine master layout), scale fp32 [Cout] or null,
 * dst_f bf16 [KH*KW][Cout][ldf] (fprop), dst_d bf16 [KH*KW][Cin][ldd] (dgrad, transposed). */
typedef struct {
  const float* src;
  const float* scale;
  void* dst_f;
  void* dst_d;
  int taps, Cout, Cin, ldf, ldd;
  long long block_begin; /* filled by the library */
} B200WeightEntry;
long long b200_weight_table_blocks(B200WeightEntry* host_entries, int n); /* fills block_begin */
int b200_prepare_weights(const B200WeightEntry* host_entries, int n, void* dev_table, void* stream);

/* 3x3/2 'same' max-pool fused with the stem's frozen-BN affine + ReLU (G4+G5).  Forward records the
 * winning window position per output (uint8, 255 = no gradient); backward is a gather over it. */
int b200_bn_relu_maxpool_fwd(const void* x, const float* scale, const float* shift, void* y, uint8_t* idx, int N,
                             int H, int W, int C, int OH, int OW, int pad_h, int pad_w, void* stream);
int b200_bn_relu_maxpool_bwd(const uint8_t* idx, const float* scale, const void* dy, void* dx, int N, int H, int W,
                             int C, int OH, int OW, int pad_h, int pad_w, void* stream);
/* Nearest upsample-to-shape + add (FPN top-down merge, G7) and its backward (gather-sum). */
int b200_upsample_add(const void* lateral, const void* top, void* out, int N, int H, int W, int C, int TH, int TW,
                      void* stream);
int b200_upsample_bwd_add(const void* dmerged, void* dtop_inout, int N, int H, int W, int C, int TH, int TW,
                          void* stream);
/* y = relu(x) and dx = dy * (y > 0) on bf16 buffers (P7's input; generic). */
/* 2x2/2 max-pool, "valid" padding (VGG blocks), NHWC bf16 [N,H,W,C] -> [N,H/2,W/2,C].  Backward takes the pooled
 * tensor x (a ReLU output) and also applies that ReLU's mask: dx is the gradient w.r.t. the pre-activation. */
int b200_maxpool2_fwd(const void* x, void* y, int N, int H, int W, int C, void* stream);
int b200_maxpool2_relu_bwd(const void* x, const void* dy, void* dx, int N, int H, int W, int C, void* stream);
int b200_relu(const void* x, void* y, long long n, void* stream);
int b200_relu_bwd(const void* y, const void* dy, void* dx, long long n, void* stream);
/* out = (a + b) * (mask > 0); b optional (merge of gradient branches + ReLU backward). */
int b200_add_mask(const void* a, const void* b, const void* mask, void* out, long long n, void* stream);
/* Column sums of a bf16 [rows, ld] matrix into fp32 (bias gradients). */
int b200_colsum(const void* x, long long rows, int cols, int ld, float* out_accum, void* stream);
int b200_colsum_multi(const void* const* xs, const long long* rows, int n, int cols, int ld, float* out_accum,
                      void* stream); /* up to 5 matrices (pyramid levels) in one launch */

/* Image normalise + pad (G19): uint8 HWC (BGR) -> bf16 NHWC with Caffe mean subtraction. */
int b200_preprocess(const uint8_t* src, void* dst, int N, int H, int W, int C, int ldc, const float* mean,
                    const int* img_hw /* optional device int32 [N,2]: pixels outside (h,w) are zeroed */, void* stream);
/* Same with a scale after the mean: dst = (src - mean) * scale ("tf" mode of MobileNet / DenseNet: mean 127.5, scale 1/127.5). */
int b200_preprocess_scaled(const uint8_t* src, void* dst, int N, int H, int W, int C, int ldc, const float* mean,
                           const int* img_hw, float scale, void* stream);

/* Anchor generation (G15): 9 anchors per location, level l, written at anchors[offset..]. */
int b200_generate_anchors(float* anchors, int H, int W, int stride, float size, void* stream);

/* IoU + arg-max + assignment + target encoding (G16; replaces the Cython compute_overlap path).
 * gt: fp32 [B, Gmax, 5] (x1,y1,x2,y2,label), gt_count int32 [B], img_hw int32 [B,2] un-padded sizes.
 * Outputs: state int8 [B,A], cls int16 [B,A], reg_target fp32 [B,A,4], num_pos int32[1] (+=). */
int b200_assign_targets(const float* anchors, int A, const float* gt, const int* gt_count, const int* img_hw, int B,
                        int Gmax, float neg_thr, float pos_thr, int8_t* state, int16_t* cls, float* reg_target,
                        int* num_pos, void* stream);

/* Sigmoid focal loss fwd+bwd in one pass over the logits (G9+G10).  logits bf16 [B, A_l, ld]
 * (A_l = anchors of this level, K classes used of ld), targets indexed at [b*A_total + a_off + a].
 * loss_accum += sum(loss)/max(1,num_pos); dlogits = dloss/dlogit * loss_scale / max(1,num_pos). */
int b200_focal_loss(const void* logits, void* dlogits, int B, int A_l, int K, int ld, const int8_t* state,
                    const int16_t* cls, int A_total, int a_off, const int* num_pos, float alpha, float gamma,
                    float loss_scale, float* loss_accum, void* stream);
/* Smooth-L1 fwd+bwd on positives only (G11).  pred bf16 [B, A_l, ld] (4 used). */
int b200_smooth_l1(const void* pred, void* dpred, int B, int A_l, int ld, const int8_t* state,
                   const float* reg_target, int A_total, int a_off, const int* num_pos, float sigma,
                   float loss_scale, float* loss_accum, void* stream);

/* Sum of squares of a flat fp32 buffer (G13), accumulated into out[0]. */
int b200_sumsq(const float* x, long long n, float* out, void* stream);
/* Keras-semantics Adam on flat fp32 buffers (G14): grads are first multiplied by
 * grad_scale[0] (device scalar: the global-norm clip factor), then
 * m,v update; p -= lr_t * m / (sqrt(v) + eps) with lr_t = lr*sqrt(1-b2^t)/(1-b1^t). */
int b200_adam(float* p, const float* g, float* m, float* v, long long n, const float* grad_scale, float lr_t,
              float beta1, float beta2, float eps, void* stream);
/* Same update, bias-corrected learning rate read from device memory (a captured CUDA graph replays with new values). */
int b200_adam_devlr(float* p, const float* g, float* m, float* v, long long n, const float* grad_scale,
                    const float* lr_t_dev, float beta1, float beta2, float eps, void* stream);
/* clip factor: scale[0] = (norm >= clip) ? clip / norm : 1 with norm = sqrt(sumsq[0]). */
int b200_clip_scale(const float* sumsq, float clipnorm, float* scale, void* stream);

/* Box decode + clip + score threshold + candidate compaction (G17), then class-aware
 * bitmask NMS + top-k + pad (G18). */
int b200_filter_detections(const void* const* cls_levels, const void* const* reg_levels, const int* level_anchors,
                           int num_levels, int ld_cls, int ld_reg, int K, const float* anchors, int A_total, int B,
                           int img_h, int img_w, float score_thr, float nms_thr, int max_det, int max_cand,
                           void* workspace, long long workspace_bytes, float* out_boxes, float* out_scores,
                           int* out_labels, void* stream);
long long b200_filter_detections_workspace(int B, int max_cand);
/* Stand-alone greedy NMS on fp32 boxes sorted by the caller (descending score); keep mask out. */
int b200_nms_sorted(const float* boxes, const int* cls_or_null, int n, float thr, unsigned long long* mask_ws,
                    int* keep_flags, void* stream);
/* Same NMS for callers that may not allocate device memory (fixed memory budget owned by
 * someone else, plugin hosts that forbid cudaMalloc, latency-critical paths): the suppression-mask
 * scratch is a 2 MiB __device__ array inside the library, held by every GPU that loads it.
 * Limit: n <= b200_nms_sorted_noalloc_max() (4096), larger n is rejected.  keep_flags is cleared
 * by the call.  The first call on a device caches the scratch address (make it before any stream
 * capture); calls that can overlap on the GPU race on the shared scratch.  Has no counterpart in
 * the reference, which ships no native code. */
int b200_nms_sorted_noalloc(const float* boxes, const int* cls_or_null, int n, float thr, int* keep_flags,
                            void* stream);
int b200_nms_sorted_noalloc_max(void);

/* ------------------------------------------------------------------ MobileNet / DenseNet layer types
 * (csrc/mobile_sm100.cu; reachable from the reference through --backbone, train.py:323-324,390).  NHWC bf16.
 * Depthwise 3x3, TF `same` padding: mode 0 forward (a = x, b = w fp32 [C][9], out = y), 1 input gradient (a = dy, b = w,
 * out = dx [N,H,W,C]), 2 filter gradient (a = x, b = dy, out = dw fp32 [C][9], accumulated).  H, W: input; OH, OW: output. */
int b200_dwconv3x3(int mode, const void* a, const void* b, void* out, int N, int H, int W, int C, int OH, int OW, int stride,
                   int pad_h, int pad_w, void* stream);
/* Batch statistics of x [rows, C] (channel stride ld): mean, biased variance; optional Keras moving-statistics update
 * (moving = moving * momentum + batch * (1 - momentum)).  sums2C: zeroed fp32 scratch [2C], left zeroed. */
int b200_bn_stats(const void* x, long long rows, int C, int ld, float* sums2C, float* mean, float* var, float* moving_mean,
                  float* moving_var, float momentum, void* stream);
/* y = act(gamma * (x - mean) / sqrt(var + eps) + beta); act 0 none, 1 ReLU, 2 ReLU6.  Also the inference form (moving stats). */
int b200_bn_act_fwd(const void* x, const float* mean, const float* var, const float* gamma, const float* beta, float eps,
                    void* y, long long rows, int C, int ldx, int ldy, int act, void* stream);
/* Backward of the same through batch statistics: dx (optionally accumulated), dgamma += , dbeta += . */
int b200_bn_act_bwd(const void* x, const void* dy, const float* mean, const float* var, const float* gamma, const float* beta,
                    float eps, float* sums2C, long long rows, int C, int ldx, int lddy, int lddx, int act, int accumulate,
                    void* dx, float* dgamma, float* dbeta, void* stream);
/* 2x2 / 2 average pool (floor sizes): forward in [N,H,W,C] -> out [N,H/2,W/2,C]; backward in = dy -> out = dx [N,H,W,C].
 * ld_in / ld_out: channel strides (either side may be a slice of a wider concatenation buffer). */
int b200_avgpool2(int backward, const void* in, void* out, int N, int H, int W, int C, int ld_in, int ld_out, void* stream);
/* Filter gradient of a dense 3x3 / 1x1 convolution with Cin <= 8 (first layers): dw fp32 [KH*KW][Cout][Cin] += . */
int b200_wgrad_small(const void* x, const void* dy, float* dw, int N, int H, int W, int Cin, int ldx, int OH, int OW, int Cout,
                     int ldy, int KH, int KW, int stride, int pad_h, int pad_w, void* stream);

/* ------------------------------------------------------------------ image geometry (real-data path)
 * uint8 BGR interleaved images, pitches in bytes.  Bilinear resize and affine warp with OpenCV's fixed-point
 * arithmetic (cv2.resize INTER_LINEAR / cv2.warpAffine INTER_LINEAR), so a batch assembled on the GPU matches the
 * reference generator's host pipeline to +-1 LSB.  inverse_2x3: the dst -> src map (cv2 inverts the matrix it is
 * given; pass that inverse).  Replaces: OpenCV resize / warpAffine in keras-retinanet's Generator (G19). */
int b200_resize_linear_u8(const uint8_t* src, int sh, int sw, long long src_pitch, uint8_t* dst, int dh, int dw,
                          long long dst_pitch, double scale_x, double scale_y, void* stream);
int b200_warp_affine_u8(const uint8_t* src, int sh, int sw, long long src_pitch, uint8_t* dst, int dh, int dw,
                        long long dst_pitch, const double* inverse_2x3, int border_replicate, int cval, void* stream);

/* ------------------------------------------------------------------ peer-to-peer collectives
 * Allreduce over NVLink peer mappings: every rank passes the table of all ranks' bucket
 * pointers (mapped with CUDA IPC) and flag pads (host arrays, copied into kernel params).  Fused: per-rank pre-scale (global-norm clip
 * factor read from each peer), 1/N averaging; the staged path can carry the peers' slices as bf16.
 * Replaces: ncclAllReduce / MPI_Allreduce + tf.div + fusion-buffer memcpys (X1-X3). */
typedef struct {
  void* bufs[8];            /* every rank's fp32 bucket base (peer-mapped device pointers) */
  uint32_t* flags[8];       /* every rank's flag pad (b200_flag_pad_bytes(), zero-initialised) */
  const float* prescale[8]; /* every rank's device scalar (its clip factor), if has_prescale */
  void* outs[8];            /* every rank's peer-mapped output buffer (algo 3, push); null entries otherwise */
  float* out;               /* LOCAL result buffer (same offsets as the bucket) */
  int rank, world;          /* world in {1, 2, 4, 8} */
  long long offset, count;  /* element range of the bucket, multiples of 4 */
  float postscale;          /* e.g. 1/world */
  uint32_t epoch;           /* flag base; advance by 4 per call, same on every rank */
  int algo;                 /* 0 auto, 1 one-shot, 2 two-shot (pull), 3 fused reduce + push all-gather,
                               4 NVLS: in-switch reduce (multimem.ld_reduce) + multicast store (needs mc_buf/mc_out) */
  int max_blocks;           /* <= 64; 0 = default */
  int has_prescale;
  unsigned int* epoch_ptr;  /* device counter holding the flag epoch (bumped by 4 after the call); null: use `epoch`.
                               With it a captured CUDA graph of the collective can be replayed. */
  void* mc_buf;             /* multicast (NVLS) address of the bucket, or null */
  void* mc_out;             /* multicast (NVLS) address of the result buffer, or null */
  int external_sync;        /* 1: the caller orders the ranks with events (ranks = threads of one process); the
                               kernel then contains no flag barrier and algo is forced to one-shot */
  void* stage[8];           /* every rank's peer-mapped STAGING area (world x ceil(count/4/world) float4), or null:
                               used by b200_ar_scatter / b200_ar_reduce_staged */
  int wire_bf16;            /* staged path: peers' raw slices travel and are staged as bf16 (half the scatter bytes) */
  unsigned int* err;        /* LOCAL device word pair {code, peer}: a flag barrier that waited longer than
                               timeout_cycles for `peer` stores code = the awaited epoch here and gives up (the
                               results of that call are then garbage); null: never written */
  long long timeout_cycles; /* 0: wait for ever */
} B200ArHostArgs;
int b200_allreduce_host(const B200ArHostArgs* args, void* stream);
/* Allreduce split around the backward pass (Horovod clips by the LOCAL global norm before averaging, reference
 * train.py:103-104, so the weights of the sum are unknown until backward ends -- but the raw gradients are not):
 *   b200_ar_scatter       elements [lo, hi) of this rank's bucket go, raw, to the staging area of the rank that
 *                         owns them (owner = index / ceil(n/world)); call it on a side stream as soon as a range
 *                         of the bucket is final, any number of times per step; no barrier, no flags;
 *   b200_ar_reduce_staged after the clip factors exist: flag barrier, every rank forms
 *                         post * sum_r pre_r * g_r over ITS slice from local memory (own bucket + staged peers) and
 *                         stores it into every rank's result buffer (multimem.st through the switch when mc_out is
 *                         given, else one store per peer), flag barrier.
 * Only the all-gather half stays on the critical path.  `count` must be the same in both calls. */
int b200_ar_scatter(const B200ArHostArgs* args, long long lo, long long hi, void* stream);
int b200_ar_reduce_staged(const B200ArHostArgs* args, void* stream);
int b200_broadcast_host(const B200ArHostArgs* args, int root, void* stream);
long long b200_flag_pad_bytes(void);
/* IPC helpers (cudaIpc*): 64-byte handles exchanged by the Python layer. */
int b200_ipc_alloc(void** ptr, long long bytes);
int b200_ipc_free(void* ptr);
int b200_ipc_get_handle(void* ptr, void* handle64);
int b200_ipc_open_handle(const void* handle64, void** ptr);
int b200_ipc_close_handle(void* ptr);
int b200_enable_peer_access(int device, int world_local);

/* ---- VMM + NVLink multicast set-up (csrc/vmm_multicast.cu): symmetric regions that peers map through a
 * POSIX fd and that are bound to an NVSwitch multicast object for the NVLS allreduce (algo 4).  The
 * reference's stack reaches NVLS only through NCCL, if at all. */
int b200_vmm_supported(int device, int* multicast);
int b200_vmm_granularity(int device, int world, long long* bytes);
int b200_vmm_create(int device, long long bytes, unsigned long long* handle, int* fd);
int b200_vmm_import(int fd, unsigned long long* handle);
int b200_vmm_map(unsigned long long handle, long long bytes, int device, long long alignment, void** ptr);
int b200_vmm_unmap(void* ptr, long long bytes);
int b200_vmm_release(unsigned long long handle);
int b200_mc_create(int world, long long bytes, unsigned long long* handle, int* fd);
int b200_mc_add_device(unsigned long long mc, int device);
int b200_mc_bind(unsigned long long mc, unsigned long long mem, long long bytes);

#ifdef __cplusplus
}
#endif
#endif /* B200_OPS_H_ */
