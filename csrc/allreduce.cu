// Peer-to-peer gradient allreduce / broadcast over NVLink (sm_100a), fused with the per-rank
// global-norm clip factor and the 1/N averaging scale; the staged (overlapped) path can carry bf16.
//
// Every rank maps every peer's bucket and flag pad (CUDA IPC) and runs the same kernel:
//   one-shot : barrier; out[i] = post * sum_r pre_r * peer_r[i]; barrier           (small buckets)
//   two-shot : barrier; reduce my 1/N slice from all peers, write it back to my bucket;
//              barrier; gather every peer's reduced slice into out; barrier        (large buckets)
// Barriers are per-block flag words with release/acquire at system scope and monotonically
// increasing epochs, so no flag is ever reset (no reset race).
//
// Replaces Horovod's fusion-buffer memcpy + ncclAllReduce/MPI_Allreduce + tf.div
// (SURVEY.md C4-C7, X1-X3); the reference has no CUDA source of its own.
#include <cstdlib>

#include "../include/b200_ops.h"
#include <cuda_bf16.h>

#include "common.cuh"

namespace b200 {

constexpr int kMaxWorld = 8;
constexpr int kArThreads = 512;
constexpr int kArMaxBlocks = 296;  // 2 per SM; flag pad = blocks x 8 peers x 4 B

struct ArParams {
  float* bufs[kMaxWorld];
  uint32_t* flags[kMaxWorld];
  const float* prescale[kMaxWorld];
  float* outs[kMaxWorld];  // peer-mapped output buffers (push algorithm), may be null
  float* out;
  int rank, world;
  long long offset, count;
  float postscale;
  uint32_t epoch;   // base flag value of this call; the call uses epoch+1 .. epoch+3
  const uint32_t* epoch_ptr;  // device-resident epoch (CUDA-graph replay), or null
  const float* mc_buf;  // multicast address of the bucket (NVLS), or null
  float* mc_out;        // multicast address of the result buffer (NVLS), or null
  int has_prescale;
  int external_sync;  // no in-kernel flag barriers: the launching streams were ordered by the host
  void* stage[kMaxWorld];   // peer-mapped staging areas (staged path)
  int wire_bf16;
  unsigned int* err;        // local {code, peer} word pair written when a barrier times out
  long long timeout_cycles;
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Block-level barrier across ranks: block b of every rank meets block b of every other rank.
// Flag layout per rank: flags[block][peer].
// Base flag value of this call: a kernel argument, or read from device memory so that a captured graph of the
// collective sees a fresh epoch at every replay (a one-thread kernel bumps the counter after each collective).
__device__ __forceinline__ uint32_t ar_epoch(const ArParams& p) { return p.epoch_ptr ? *p.epoch_ptr : p.epoch; }

// The counter wraps mod 2^32; rank_barrier compares with a signed difference, which is wrap-safe.
__global__ void epoch_bump_kernel(uint32_t* e) { *e = *e + 4u; }

__device__ __forceinline__ void rank_barrier(const ArParams& p, uint32_t value) {
  if (p.external_sync) return;
  __syncthreads();
  if (threadIdx.x < p.world) {
    const int peer = threadIdx.x;
    st_release_sys(p.flags[peer] + blockIdx.x * kMaxWorld + p.rank, value);
    const uint32_t* mine = p.flags[p.rank] + blockIdx.x * kMaxWorld + peer;
    // A peer that died or hangs between launches must not leave this GPU spinning in a kernel for ever (a hung
    // kernel also blocks context teardown): after timeout_cycles the barrier records {awaited epoch, peer} and gives
    // up; the host raises from P2PComm.check().  The results of a timed-out collective are garbage.
    const long long t0 = clock64();
    while (static_cast<int32_t>(ld_acquire_sys(mine) - value) < 0) {
      if (p.timeout_cycles > 0 && clock64() - t0 > p.timeout_cycles) {
        if (p.err && atomicCAS(p.err, 0u, value | 0x80000000u) == 0u) p.err[1] = static_cast<unsigned int>(peer);
        break;
      }
    }
  }
  __syncthreads();
}

// Peer data is read with plain streaming loads: it was published before the flag barrier
// (st.release.sys / ld.acquire.sys), peer memory is never in the local L2, and L1 starts empty at
// kernel launch and is bypassed here, so no stale line can be hit.
__device__ __forceinline__ float4 ld_peer(const float* p) {
  float4 v;
  asm volatile("ld.global.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}

template <int kWorld>
__global__ void __launch_bounds__(kArThreads) allreduce_oneshot_kernel(const ArParams p) {
  rank_barrier(p, ar_epoch(p) + 1);
  float pre[kWorld];
#pragma unroll
  for (int r = 0; r < kWorld; ++r) pre[r] = p.has_prescale ? *p.prescale[r] : 1.f;
  const long long n4 = p.count >> 2;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float4 v[kWorld];
#pragma unroll
    for (int r = 0; r < kWorld; ++r) v[r] = ld_peer(p.bufs[(p.rank + r) % kWorld] + p.offset + i * 4);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int r = 0; r < kWorld; ++r) {
      const float s = pre[(p.rank + r) % kWorld];
      acc.x += s * v[r].x; acc.y += s * v[r].y; acc.z += s * v[r].z; acc.w += s * v[r].w;
    }
    acc.x *= p.postscale; acc.y *= p.postscale; acc.z *= p.postscale; acc.w *= p.postscale;
    reinterpret_cast<float4*>(p.out + p.offset)[i] = acc;
  }
  rank_barrier(p, ar_epoch(p) + 2);
}

template <int kWorld>
__global__ void __launch_bounds__(kArThreads) allreduce_twoshot_kernel(const ArParams p) {
  rank_barrier(p, ar_epoch(p) + 1);
  float pre[kWorld];
#pragma unroll
  for (int r = 0; r < kWorld; ++r) pre[r] = p.has_prescale ? *p.prescale[r] : 1.f;
  const long long n4 = p.count >> 2;
  const long long slice = (n4 + kWorld - 1) / kWorld;
  // phase 1: reduce-scatter my slice, write it back into my own bucket (only I read that slice).
  // kUnroll iterations are batched so that every thread keeps ~8 sixteen-byte peer loads in flight.
  constexpr int kUnroll = kWorld >= 8 ? 1 : 8 / kWorld;
  {
    const long long s0 = slice * p.rank, s1 = min(n4, s0 + slice);
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long i0 = s0 + blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i0 < s1; i0 += stride * kUnroll) {
      float4 v[kUnroll][kWorld];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const long long i = i0 + u * stride;
        if (i < s1) {
#pragma unroll
          for (int r = 0; r < kWorld; ++r) v[u][r] = ld_peer(p.bufs[(p.rank + r) % kWorld] + p.offset + i * 4);
        }
      }
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const long long i = i0 + u * stride;
        if (i < s1) {
          float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
          for (int r = 0; r < kWorld; ++r) {
            const float s = pre[(p.rank + r) % kWorld];
            acc.x += s * v[u][r].x; acc.y += s * v[u][r].y; acc.z += s * v[u][r].z; acc.w += s * v[u][r].w;
          }
          acc.x *= p.postscale; acc.y *= p.postscale; acc.z *= p.postscale; acc.w *= p.postscale;
          reinterpret_cast<float4*>(p.bufs[p.rank] + p.offset)[i] = acc;
        }
      }
    }
  }
  __threadfence_system();
  rank_barrier(p, ar_epoch(p) + 2);
  // phase 2: all-gather the reduced slices (4 loads in flight per thread per peer slice)
#pragma unroll
  for (int r = 0; r < kWorld; ++r) {
    const int src = (p.rank + r) % kWorld;
    const long long s0 = slice * src, s1 = min(n4, s0 + slice);
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long i0 = s0 + blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i0 < s1; i0 += stride * 4) {
      float4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (i0 + u * stride < s1) v[u] = ld_peer(p.bufs[src] + p.offset + (i0 + u * stride) * 4);
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (i0 + u * stride < s1) reinterpret_cast<float4*>(p.out + p.offset)[i0 + u * stride] = v[u];
    }
  }
  rank_barrier(p, ar_epoch(p) + 3);
}

// Fused reduce-scatter + all-gather, push flavour: each rank reduces its 1/N slice from all peers
// (reads) and stores the result straight into EVERY rank's output buffer (writes).  Reads and writes
// travel in opposite NVLink directions, so each direction carries only (N-1)/N * S bytes and there
// is no second phase and no intermediate copy in the bucket.
template <int kWorld>
__global__ void __launch_bounds__(kArThreads) allreduce_push_kernel(const ArParams p) {
  rank_barrier(p, ar_epoch(p) + 1);
  float pre[kWorld];
#pragma unroll
  for (int r = 0; r < kWorld; ++r) pre[r] = p.has_prescale ? *p.prescale[r] : 1.f;
  const long long n4 = p.count >> 2;
  const long long slice = (n4 + kWorld - 1) / kWorld;
  constexpr int kUnroll = kWorld >= 8 ? 1 : 8 / kWorld;
  const long long s0 = slice * p.rank, s1 = min(n4, s0 + slice);
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i0 = s0 + blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i0 < s1; i0 += stride * kUnroll) {
    float4 v[kUnroll][kWorld];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const long long i = i0 + u * stride;
      if (i < s1) {
#pragma unroll
        for (int r = 0; r < kWorld; ++r) v[u][r] = ld_peer(p.bufs[(p.rank + r) % kWorld] + p.offset + i * 4);
      }
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const long long i = i0 + u * stride;
      if (i < s1) {
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int r = 0; r < kWorld; ++r) {
          const float s = pre[(p.rank + r) % kWorld];
          acc.x += s * v[u][r].x; acc.y += s * v[u][r].y; acc.z += s * v[u][r].z; acc.w += s * v[u][r].w;
        }
        acc.x *= p.postscale; acc.y *= p.postscale; acc.z *= p.postscale; acc.w *= p.postscale;
#pragma unroll
        for (int r = 0; r < kWorld; ++r)
          reinterpret_cast<float4*>(p.outs[(p.rank + r) % kWorld] + p.offset)[i] = acc;
      }
    }
  }
  __threadfence_system();
  rank_barrier(p, ar_epoch(p) + 2);
}


// ---- NVLS: the NVSwitch reduces.  multimem.ld_reduce on the multicast address returns the sum of the
// same offset in EVERY rank's bucket (added inside the switch); multimem.st writes one value into every
// rank's result buffer.  Each rank handles 1/N of the bucket, so a GPU sends and receives about S bytes
// instead of 2 (N-1)/N S with unicast reduce-scatter + all-gather.  The switch cannot weight its inputs, so
// the per-rank clip factor is applied to the local bucket in place before the first barrier.
__device__ __forceinline__ float4 multimem_ld_reduce_add(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(mc)
               : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st(float* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}

template <int kWorld>
__global__ void __launch_bounds__(kArThreads) allreduce_nvls_kernel(const ArParams p) {
  const long long n4 = p.count >> 2;
  const long long tid = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  const long long nthreads = static_cast<long long>(gridDim.x) * blockDim.x;
  const long long per = (n4 + kWorld - 1) / kWorld;
  if (p.has_prescale) {
    const float s = *p.prescale[p.rank];
    if (s != 1.f) {
      // The in-place pre-scale uses the SAME index -> block mapping as the reduce below (slice r, then
      // tid + k * nthreads inside it): the elements block b of rank r will ld_reduce were scaled, on every
      // rank, by that rank's block b, and rank_barrier orders exactly block b of all ranks.  (A flat
      // i = tid + k * nthreads mapping hands them to a different block, which nothing orders.)
      float4* mine = reinterpret_cast<float4*>(p.bufs[p.rank] + p.offset);
#pragma unroll 1
      for (int r = 0; r < kWorld; ++r) {
        const long long r_hi = (r + 1) * per < n4 ? (r + 1) * per : n4;
        for (long long i = r * per + tid; i < r_hi; i += nthreads) {
          float4 v = mine[i];
          v.x *= s; v.y *= s; v.z *= s; v.w *= s;
          mine[i] = v;
        }
      }
      __threadfence_system();
    }
  }
  rank_barrier(p, ar_epoch(p) + 1);
  const long long lo = p.rank * per;
  const long long hi = lo + per < n4 ? lo + per : n4;
  const float* in = p.mc_buf + p.offset;
  float* out = p.mc_out + p.offset;
  constexpr int kU = 4;
  long long i = lo + tid;
  for (; i + (kU - 1) * nthreads < hi; i += kU * nthreads) {
    float4 v[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) v[u] = multimem_ld_reduce_add(in + (i + u * nthreads) * 4);
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      v[u].x *= p.postscale; v[u].y *= p.postscale; v[u].z *= p.postscale; v[u].w *= p.postscale;
      multimem_st(out + (i + u * nthreads) * 4, v[u]);
    }
  }
  for (; i < hi; i += nthreads) {
    float4 v = multimem_ld_reduce_add(in + i * 4);
    v.x *= p.postscale; v.y *= p.postscale; v.z *= p.postscale; v.w *= p.postscale;
    multimem_st(out + i * 4, v);
  }
  __threadfence_system();
  rank_barrier(p, ar_epoch(p) + 2);
}

// Broadcast through the switch: the root stores once, every rank's result buffer receives it.
template <int kWorld>
__global__ void __launch_bounds__(kArThreads) broadcast_nvls_kernel(const ArParams p, int root) {
  rank_barrier(p, ar_epoch(p) + 1);
  if (p.rank == root) {
    const long long n4 = p.count >> 2;
    const float4* src = reinterpret_cast<const float4*>(p.bufs[root] + p.offset);
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4;
         i += static_cast<long long>(gridDim.x) * blockDim.x)
      multimem_st(p.mc_out + p.offset + i * 4, src[i]);
    __threadfence_system();
  }
  rank_barrier(p, ar_epoch(p) + 2);
}

template <int kWorld>
__global__ void __launch_bounds__(kArThreads) broadcast_kernel(const ArParams p, int root) {
  rank_barrier(p, ar_epoch(p) + 1);
  if (p.rank != root) {
    const long long n4 = p.count >> 2;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
      const float4 v = ld_peer(p.bufs[root] + p.offset + i * 4);
      reinterpret_cast<float4*>(p.out + p.offset)[i] = v;
    }
  }
  rank_barrier(p, ar_epoch(p) + 2);
}

// ---- staged allreduce: scatter during backward, weighted reduce + all-gather after it ------------------------
// Staging area of owner o: [world][per] float4 (or [world][per] 8-byte bf16x4), slot r = rank r's raw slice o.
__device__ __forceinline__ uint2 pack_bf16x4(float4 v) {
  __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
  return make_uint2(*reinterpret_cast<uint32_t*>(&a), *reinterpret_cast<uint32_t*>(&b));
}
__device__ __forceinline__ float4 unpack_bf16x4(uint2 u) {
  return make_float4(__uint_as_float(u.x << 16), __uint_as_float(u.x & 0xffff0000u), __uint_as_float(u.y << 16),
                     __uint_as_float(u.y & 0xffff0000u));
}

__global__ void __launch_bounds__(256) ar_scatter_kernel(const ArParams p, long long lo4, long long hi4) {
  const long long n4 = p.count >> 2;
  const long long per = (n4 + p.world - 1) / p.world;
  const float4* mine = reinterpret_cast<const float4*>(p.bufs[p.rank] + p.offset);
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = lo4 + blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < hi4; i += stride) {
    const int owner = static_cast<int>(i / per);
    if (owner == p.rank) continue;
    const float4 v = mine[i];
    const long long slot = static_cast<long long>(p.rank) * per + (i - owner * per);
    if (p.wire_bf16) reinterpret_cast<uint2*>(p.stage[owner])[slot] = pack_bf16x4(v);
    else reinterpret_cast<float4*>(p.stage[owner])[slot] = v;
  }
  __threadfence_system();
}

template <int kWorld>
__global__ void __launch_bounds__(kArThreads) ar_reduce_staged_kernel(const ArParams p) {
  rank_barrier(p, ar_epoch(p) + 1);        // every rank's scatters have landed, every clip factor is published
  float pre[kWorld];
#pragma unroll
  for (int r = 0; r < kWorld; ++r) pre[r] = p.has_prescale ? *p.prescale[r] : 1.f;
  const long long n4 = p.count >> 2;
  const long long per = (n4 + kWorld - 1) / kWorld;
  const long long lo = per * p.rank, hi = min(n4, lo + per);
  const float4* mine = reinterpret_cast<const float4*>(p.bufs[p.rank] + p.offset);
  const float4* st32 = reinterpret_cast<const float4*>(p.stage[p.rank]);
  const uint2* st16 = reinterpret_cast<const uint2*>(p.stage[p.rank]);
  float* mc_out = p.mc_out ? p.mc_out + p.offset : nullptr;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = lo + blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < hi; i += stride) {
    // all loads of this element first (local HBM; the staged slices were written by the peers' scatter kernels)
    float4 v[kWorld];
#pragma unroll
    for (int r = 0; r < kWorld; ++r) {
      if (r == p.rank) {
        v[r] = mine[i];
      } else if (p.wire_bf16) {
        uint2 u;
        asm volatile("ld.global.L1::no_allocate.v2.u32 {%0, %1}, [%2];" : "=r"(u.x), "=r"(u.y) : "l"(st16 + r * per + (i - lo)) : "memory");
        v[r] = unpack_bf16x4(u);
      } else {
        v[r] = ld_peer(reinterpret_cast<const float*>(st32 + r * per + (i - lo)));
      }
    }
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int r = 0; r < kWorld; ++r) {
      acc.x += pre[r] * v[r].x; acc.y += pre[r] * v[r].y; acc.z += pre[r] * v[r].z; acc.w += pre[r] * v[r].w;
    }
    acc.x *= p.postscale; acc.y *= p.postscale; acc.z *= p.postscale; acc.w *= p.postscale;
    if (mc_out) {
      multimem_st(mc_out + i * 4, acc);      // one store, delivered to every rank's result buffer by the switch
    } else {
#pragma unroll
      for (int r = 0; r < kWorld; ++r) reinterpret_cast<float4*>(p.outs[(p.rank + r) % kWorld] + p.offset)[i] = acc;
    }
  }
  __threadfence_system();
  rank_barrier(p, ar_epoch(p) + 2);
}

__global__ void scale_copy_kernel(const float* __restrict__ in, float* __restrict__ out, long long n4,
                                  const float* __restrict__ pre, float post) {
  const float s = (pre ? *pre : 1.f) * post;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float4 v = reinterpret_cast<const float4*>(in)[i];
    v.x *= s; v.y *= s; v.z *= s; v.w *= s;
    reinterpret_cast<float4*>(out)[i] = v;
  }
}

}  // namespace b200

using namespace b200;

// Host-side argument block: pointer tables are HOST arrays here (copied by value into the
// kernel parameters), so no device-side table has to be maintained.
extern "C" {



static int fill(ArParams& p, const B200ArHostArgs* a) {
  if (a->world < 1 || a->world > kMaxWorld) return set_error(-1, "allreduce: world %d unsupported", a->world);
  if ((a->count & 3) || (a->offset & 3)) return set_error(-1, "allreduce: offset/count must be multiples of 4 floats");
  for (int r = 0; r < kMaxWorld; ++r) {
    p.bufs[r] = r < a->world ? static_cast<float*>(a->bufs[r]) : nullptr;
    p.flags[r] = r < a->world ? a->flags[r] : nullptr;
    p.prescale[r] = r < a->world ? a->prescale[r] : nullptr;
    p.outs[r] = r < a->world ? static_cast<float*>(a->outs[r]) : nullptr;
  }
  p.out = a->out;
  p.rank = a->rank;
  p.world = a->world;
  p.offset = a->offset;
  p.count = a->count;
  p.postscale = a->postscale;
  p.epoch = a->epoch;
  p.has_prescale = a->has_prescale;
  p.epoch_ptr = a->epoch_ptr;
  p.mc_buf = static_cast<const float*>(a->mc_buf);
  p.mc_out = static_cast<float*>(a->mc_out);
  p.external_sync = a->external_sync;
  for (int r = 0; r < kMaxWorld; ++r) p.stage[r] = r < a->world ? a->stage[r] : nullptr;
  p.wire_bf16 = a->wire_bf16;
  p.err = a->err;
  p.timeout_cycles = a->timeout_cycles;
  return 0;
}

int b200_allreduce_host(const B200ArHostArgs* a, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  ArParams p;
  int rc = fill(p, a);
  if (rc) return rc;
  const long long n4 = p.count >> 2;
  if (n4 == 0) return 0;
  if (p.world == 1) {
    long long g = (n4 + 255) / 256;
    if (g > num_sms() * 4LL) g = num_sms() * 4LL;
    scale_copy_kernel<<<static_cast<unsigned>(g), 256, 0, stream>>>(p.bufs[0] + p.offset, p.out + p.offset, n4,
                                                                    p.has_prescale ? p.prescale[0] : nullptr, p.postscale);
    return check_launch("scale_copy_kernel");
  }
  static int env_blocks = -1;
  if (env_blocks < 0) { const char* e = getenv("B200_AR_BLOCKS"); env_blocks = e ? atoi(e) : 0; }
  // enough blocks to keep NVLink busy for big buckets (measured: 64 -> 0.29, 296 -> 0.68 of the link
  // roofline at 2 GPUs), few for small ones where the flag barriers dominate
  int auto_blocks = static_cast<int>(p.count * 4 / (64 << 10));
  if (auto_blocks < 8) auto_blocks = 8;
  if (auto_blocks > 2 * num_sms()) auto_blocks = 2 * num_sms();
  int blocks = a->max_blocks > 0 ? a->max_blocks : (env_blocks > 0 ? env_blocks : auto_blocks);
  if (blocks > kArMaxBlocks) blocks = kArMaxBlocks;
  const long long bytes = p.count * 4;
  int algo = a->algo;
  const bool can_push = a->outs[0] != nullptr;
  // measured at 8 GPUs: push is never slower than one-shot (23 vs 25 us at 1 KB, 28 vs 61 us at 256 KB)
  const bool can_nvls = p.mc_buf != nullptr && p.mc_out != nullptr && !p.external_sync;
  static long long nvls_min = -1;
  if (nvls_min < 0) { const char* e = getenv("B200_AR_NVLS_MIN"); nvls_min = e ? atoll(e) : (1ll << 20); }
  // NVLS moves S bytes per direction per GPU, unicast push 2(N-1)/N * S / 2: no gain at N = 2 (measured 0.6x)
  if (algo == 0) algo = (can_nvls && p.world >= 4 && bytes >= nvls_min) ? 4 : (can_push ? 3 : ((bytes <= (128 << 10)) ? 1 : 2));
  if (algo == 4 && !can_nvls) return set_error(-1, "allreduce: NVLS algorithm needs multicast addresses");
  if (p.external_sync) algo = 1;   // the multi-phase algorithms need barriers between their phases
  if (algo == 3 && !can_push) return set_error(-1, "allreduce: push algorithm needs peer-mapped output buffers");
  const long long work = (algo == 1) ? n4 : (n4 + p.world - 1) / p.world;
  long long need = (work + kArThreads - 1) / kArThreads;
  if (need < blocks) blocks = static_cast<int>(need < 1 ? 1 : need);
#define B200_AR_LAUNCH(W)                                                                        \
  case W:                                                                                        \
    if (algo == 1)                                                                               \
      allreduce_oneshot_kernel<W><<<blocks, kArThreads, 0, stream>>>(p);                         \
    else if (algo == 3)                                                                          \
      allreduce_push_kernel<W><<<blocks, kArThreads, 0, stream>>>(p);                            \
    else if (algo == 4)                                                                          \
      allreduce_nvls_kernel<W><<<blocks, kArThreads, 0, stream>>>(p);                            \
    else                                                                                         \
      allreduce_twoshot_kernel<W><<<blocks, kArThreads, 0, stream>>>(p);                         \
    break;
  switch (p.world) {
    B200_AR_LAUNCH(2)
    B200_AR_LAUNCH(3)
    B200_AR_LAUNCH(4)
    B200_AR_LAUNCH(5)
    B200_AR_LAUNCH(6)
    B200_AR_LAUNCH(7)
    B200_AR_LAUNCH(8)
    default:
      return set_error(-1, "allreduce: world size %d not instantiated (1..8)", p.world);
  }
#undef B200_AR_LAUNCH
  if (a->epoch_ptr) epoch_bump_kernel<<<1, 1, 0, stream>>>(a->epoch_ptr);
  return check_launch("allreduce_kernel");
}

int b200_ar_scatter(const B200ArHostArgs* a, long long lo, long long hi, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  ArParams p;
  int rc = fill(p, a);
  if (rc) return rc;
  if (p.world == 1 || hi <= lo) return 0;
  if ((lo & 3) || (hi & 3) || hi > p.count) return set_error(-1, "ar_scatter: range must be float4-aligned and inside count");
  if (!p.stage[0]) return set_error(-1, "ar_scatter: no staging areas");
  const long long n4 = (hi - lo) >> 2;
  // few, small blocks: this kernel runs beside the backward pass and only has to keep NVLink stores flowing
  long long blocks = (n4 + 256 * 8 - 1) / (256 * 8);
  if (blocks > 64) blocks = 64;
  if (blocks < 1) blocks = 1;
  ar_scatter_kernel<<<static_cast<unsigned>(blocks), 256, 0, stream>>>(p, lo >> 2, hi >> 2);
  return check_launch("ar_scatter_kernel");
}

int b200_ar_reduce_staged(const B200ArHostArgs* a, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  ArParams p;
  int rc = fill(p, a);
  if (rc) return rc;
  const long long n4 = p.count >> 2;
  if (n4 == 0) return 0;
  if (p.world == 1) return b200_allreduce_host(a, stream_);
  if (p.external_sync) return set_error(-1, "ar_reduce_staged: needs in-kernel flag barriers");
  if (!p.stage[0] || (!p.mc_out && !p.outs[0])) return set_error(-1, "ar_reduce_staged: staging / output areas missing");
  const long long work = (n4 + p.world - 1) / p.world;
  int blocks = a->max_blocks > 0 ? a->max_blocks : 2 * num_sms();
  if (blocks > kArMaxBlocks) blocks = kArMaxBlocks;
  const long long need = (work + kArThreads - 1) / kArThreads;
  if (need < blocks) blocks = static_cast<int>(need < 1 ? 1 : need);
  switch (p.world) {
    case 2: ar_reduce_staged_kernel<2><<<blocks, kArThreads, 0, stream>>>(p); break;
    case 3: ar_reduce_staged_kernel<3><<<blocks, kArThreads, 0, stream>>>(p); break;
    case 4: ar_reduce_staged_kernel<4><<<blocks, kArThreads, 0, stream>>>(p); break;
    case 5: ar_reduce_staged_kernel<5><<<blocks, kArThreads, 0, stream>>>(p); break;
    case 6: ar_reduce_staged_kernel<6><<<blocks, kArThreads, 0, stream>>>(p); break;
    case 7: ar_reduce_staged_kernel<7><<<blocks, kArThreads, 0, stream>>>(p); break;
    case 8: ar_reduce_staged_kernel<8><<<blocks, kArThreads, 0, stream>>>(p); break;
    default: return set_error(-1, "ar_reduce_staged: world size %d not instantiated", p.world);
  }
  if (a->epoch_ptr) epoch_bump_kernel<<<1, 1, 0, stream>>>(a->epoch_ptr);
  return check_launch("ar_reduce_staged_kernel");
}

int b200_broadcast_host(const B200ArHostArgs* a, int root, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  ArParams p;
  int rc = fill(p, a);
  if (rc) return rc;
  const long long n4 = p.count >> 2;
  if (n4 == 0 || p.world == 1) return 0;
  int blocks = a->max_blocks > 0 ? a->max_blocks : kArMaxBlocks;
  if (blocks > kArMaxBlocks) blocks = kArMaxBlocks;
  long long need = (n4 + kArThreads - 1) / kArThreads;
  if (need < blocks) blocks = static_cast<int>(need < 1 ? 1 : need);
  if (a->algo == 4) {
    // through the switch: the result lands in every rank's `result` buffer (mc_out), the root's included
    if (!p.mc_out || p.external_sync) return set_error(-1, "broadcast: NVLS needs a multicast result address");
    switch (p.world) {
      case 2: broadcast_nvls_kernel<2><<<blocks, kArThreads, 0, stream>>>(p, root); break;
      case 3: broadcast_nvls_kernel<3><<<blocks, kArThreads, 0, stream>>>(p, root); break;
      case 4: broadcast_nvls_kernel<4><<<blocks, kArThreads, 0, stream>>>(p, root); break;
      case 5: broadcast_nvls_kernel<5><<<blocks, kArThreads, 0, stream>>>(p, root); break;
      case 6: broadcast_nvls_kernel<6><<<blocks, kArThreads, 0, stream>>>(p, root); break;
      case 7: broadcast_nvls_kernel<7><<<blocks, kArThreads, 0, stream>>>(p, root); break;
      case 8: broadcast_nvls_kernel<8><<<blocks, kArThreads, 0, stream>>>(p, root); break;
      default: return set_error(-1, "broadcast: world size %d not instantiated", p.world);
    }
    if (a->epoch_ptr) epoch_bump_kernel<<<1, 1, 0, stream>>>(a->epoch_ptr);
    return check_launch("broadcast_nvls_kernel");
  }
  switch (p.world) {
    case 2: broadcast_kernel<2><<<blocks, kArThreads, 0, stream>>>(p, root); break;
    case 3: broadcast_kernel<3><<<blocks, kArThreads, 0, stream>>>(p, root); break;
    case 4: broadcast_kernel<4><<<blocks, kArThreads, 0, stream>>>(p, root); break;
    case 5: broadcast_kernel<5><<<blocks, kArThreads, 0, stream>>>(p, root); break;
    case 6: broadcast_kernel<6><<<blocks, kArThreads, 0, stream>>>(p, root); break;
    case 7: broadcast_kernel<7><<<blocks, kArThreads, 0, stream>>>(p, root); break;
    case 8: broadcast_kernel<8><<<blocks, kArThreads, 0, stream>>>(p, root); break;
    default: return set_error(-1, "broadcast: world size %d not instantiated", p.world);
  }
  if (a->epoch_ptr) epoch_bump_kernel<<<1, 1, 0, stream>>>(a->epoch_ptr);
  return check_launch("broadcast_kernel");
}

/* Bytes each rank must allocate (zeroed) for its flag pad. */
long long b200_flag_pad_bytes(void) { return static_cast<long long>(kArMaxBlocks) * kMaxWorld * sizeof(uint32_t); }

int b200_ipc_alloc(void** ptr, long long bytes) {
  B200_CUDA_OK(cudaMalloc(ptr, static_cast<size_t>(bytes)));
  B200_CUDA_OK(cudaMemset(*ptr, 0, static_cast<size_t>(bytes)));
  return 0;
}
int b200_ipc_free(void* ptr) {
  B200_CUDA_OK(cudaFree(ptr));
  return 0;
}
int b200_ipc_get_handle(void* ptr, void* handle64) {
  cudaIpcMemHandle_t h;
  B200_CUDA_OK(cudaIpcGetMemHandle(&h, ptr));
  memcpy(handle64, &h, sizeof(h));
  return 0;
}
int b200_ipc_open_handle(const void* handle64, void** ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, sizeof(h));
  B200_CUDA_OK(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
int b200_ipc_close_handle(void* ptr) {
  B200_CUDA_OK(cudaIpcCloseMemHandle(ptr));
  return 0;
}
int b200_enable_peer_access(int device, int world_local) {
  for (int d = 0; d < world_local; ++d) {
    if (d == device) continue;
    int can = 0;
    B200_CUDA_OK(cudaDeviceCanAccessPeer(&can, device, d));
    if (!can) return set_error(-7, "device %d cannot access peer %d", device, d);
    cudaError_t e = cudaDeviceEnablePeerAccess(d, 0);
    if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return set_error(-3, "enable peer: %s", cudaGetErrorString(e));
    cudaGetLastError();
  }
  return 0;
}

}  // extern "C"
