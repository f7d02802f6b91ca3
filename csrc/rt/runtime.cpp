// Host collective runtime: Horovod-core equivalent over POSIX shared memory (single node).
//
//   user thread:  b200rt_enqueue(name, tensor) -> handle            (returns immediately)
//   cycle thread: every cycle_ms  1. publish this rank's pending names in its shm slot
//                                 2. rank 0 only: coordinator -- a name becomes a *response* once
//                                    every rank has it pending (dtype/count/op validated; a
//                                    mismatch yields an error response delivered to all ranks)
//                                 3. all ranks: execute the response in order; consecutive
//                                    allreduces of one dtype are fused into the shm fusion buffer
//                                    (copy in, barrier, reduce my 1/N slice, barrier, copy out)
//   extras:       stall warnings (names pending on some ranks only), peer-loss detection by
//                 heartbeat, Chrome-trace timeline, fault injection for tests.
//
// Reference behaviour being reproduced: horovod.keras init/rank/allreduce/broadcast as bound by
// train.py:20-21,103,111.  No code is shared with Horovod; the negotiation exists
// for API parity (dynamic, name-keyed collectives) -- the GPU hot path uses a static bucket plan.
#include <fcntl.h>
#include <signal.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

#include <atomic>
#include <condition_variable>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/b200_rt.h"

namespace {

constexpr int kMaxRanks = 64;
constexpr int kMaxPending = 192;
constexpr int kMaxResp = 64;
constexpr int kNameLen = 96;
constexpr uint32_t kMagic = 0xB200C0DEu;

struct ReqDesc {
  char name[kNameLen];
  int64_t count;
  int32_t op, dtype, root, pad;
};

struct RankSlot {
  std::atomic<uint64_t> version;  // seqlock: odd while the owner is writing
  std::atomic<uint64_t> heartbeat_ns;
  uint32_t npending;
  uint32_t pad;
  ReqDesc pending[kMaxPending];
};

struct RespEntry {
  char name[kNameLen];
  int32_t op, dtype, root, status;  // status 0 ok, else error
  int64_t counts[kMaxRanks];
  char error[160];
};

struct Header {
  std::atomic<uint32_t> magic;
  uint32_t world;
  int32_t creator_pid;
  uint32_t pad0;
  uint64_t fusion_bytes;
  std::atomic<uint32_t> bar_count;
  std::atomic<uint32_t> bar_sense;
  std::atomic<uint32_t> abort_flag;
  std::atomic<uint32_t> attached;
  char abort_msg[256];
  std::atomic<uint64_t> resp_seq;
  std::atomic<uint32_t> resp_acks;
  uint32_t resp_n;
  RespEntry resp[kMaxResp];
  RankSlot slots[kMaxRanks];
};

struct Handle {
  std::atomic<int> status{0};  // 0 pending, 1 done, <0 failed
  std::string error;
  long long result_count = 0;
};

struct LocalReq {
  ReqDesc d;
  const void* in;
  void* out;
  double postscale;
  long long handle;
  uint64_t t_enqueue_ns;
};

size_t dtype_size(int dt) {
  switch (dt) {
    case B200RT_F32: return 4;
    case B200RT_F64: return 8;
    case B200RT_I32: return 4;
    case B200RT_I64: return 8;
    case B200RT_U8: return 1;
  }
  return 0;
}

uint64_t now_ns() {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return static_cast<uint64_t>(ts.tv_sec) * 1000000000ull + ts.tv_nsec;
}

thread_local char g_err[512] = {0};
int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

struct Runtime {
  int rank = 0, size = 1, local_rank = 0, local_size = 1;
  double cycle_ms = 5.0;
  uint64_t fusion_bytes = 64ull << 20;
  std::string shm_name;
  size_t shm_bytes = 0;
  Header* hdr = nullptr;
  uint8_t* data = nullptr;  // size regions + 1 result region
  uint32_t my_sense = 0;
  std::thread thread;
  std::atomic<bool> stop{false};
  std::atomic<bool> running{false};
  std::mutex mu;
  std::condition_variable cv;
  std::deque<LocalReq> inbox;
  std::vector<LocalReq> pending;  // in enqueue order
  std::map<long long, std::shared_ptr<Handle>> handles;
  long long next_handle = 1;
  uint64_t last_resp = 0;
  double stall_warn_s = 60.0, peer_timeout_s = 120.0;
  std::map<std::string, uint64_t> first_seen_ns;  // coordinator: name -> first time seen on any rank
  std::map<std::string, uint64_t> last_warn_ns;
  FILE* timeline = nullptr;
  bool timeline_first = true;
  std::mutex tl_mu;
  long long stats[5] = {0, 0, 0, 0, 0};
  // fault injection (tests): B200RT_FAULT="delay:<rank>:<ms>" | "exit:<rank>:<after_n_responses>"
  int fault_kind = 0, fault_rank = -1;
  long long fault_arg = 0;

  uint8_t* region(int r) { return data + static_cast<size_t>(r) * fusion_bytes; }
  uint8_t* result() { return data + static_cast<size_t>(size) * fusion_bytes; }

  void tl_event(const char* phase, const char* tensor, uint64_t t0, uint64_t t1) {
    if (!timeline) return;
    std::lock_guard<std::mutex> g(tl_mu);
    fprintf(timeline, "%s{\"name\":\"%s\",\"cat\":\"%s\",\"ph\":\"X\",\"pid\":%d,\"tid\":1,\"ts\":%.3f,\"dur\":%.3f}",
            timeline_first ? "" : ",\n", phase, tensor, rank, t0 / 1e3, (t1 - t0) / 1e3);
    timeline_first = false;
  }

  void set_abort(const char* msg) {
    uint32_t expect = 0;
    if (hdr->abort_flag.compare_exchange_strong(expect, 1)) {
      snprintf(hdr->abort_msg, sizeof(hdr->abort_msg), "%s", msg);
    }
  }

  // Sense-reversing barrier over the shm header.  Returns false if the job was aborted.
  bool barrier() {
    if (size == 1) return true;
    my_sense ^= 1;
    const uint64_t t0 = now_ns();
    if (hdr->bar_count.fetch_add(1, std::memory_order_acq_rel) + 1 == static_cast<uint32_t>(size)) {
      hdr->bar_count.store(0, std::memory_order_relaxed);
      hdr->bar_sense.store(my_sense, std::memory_order_release);
      return true;
    }
    int spins = 0;
    while (hdr->bar_sense.load(std::memory_order_acquire) != my_sense) {
      if (hdr->abort_flag.load(std::memory_order_acquire)) return false;
      if (++spins > 2000) {
        timespec ts{0, 20000};
        nanosleep(&ts, nullptr);
        hdr->slots[rank].heartbeat_ns.store(now_ns(), std::memory_order_relaxed);
        if ((now_ns() - t0) / 1e9 > peer_timeout_s) {
          set_abort("barrier timeout: a rank stopped participating");
          return false;
        }
      }
    }
    return true;
  }

  void finish(LocalReq& q, int status, const char* err, long long result_count = 0) {
    std::shared_ptr<Handle> h;
    {
      std::lock_guard<std::mutex> g(mu);
      auto it = handles.find(q.handle);
      if (it != handles.end()) h = it->second;
    }
    if (h) {
      if (err) h->error = err;
      h->result_count = result_count;
      h->status.store(status, std::memory_order_release);
    }
    cv.notify_all();
  }

  void fail_all(const char* msg) {
    std::vector<LocalReq> all;
    {
      std::lock_guard<std::mutex> g(mu);
      for (auto& q : inbox) all.push_back(q);
      inbox.clear();
    }
    for (auto& q : pending) all.push_back(q);
    pending.clear();
    for (auto& q : all) finish(q, -9, msg);
  }

  void publish_pending() {
    RankSlot& s = hdr->slots[rank];
    const uint64_t v = s.version.load(std::memory_order_relaxed);
    s.version.store(v + 1, std::memory_order_release);
    std::atomic_thread_fence(std::memory_order_seq_cst);
    const uint32_t n = static_cast<uint32_t>(std::min<size_t>(pending.size(), kMaxPending));
    for (uint32_t i = 0; i < n; ++i) s.pending[i] = pending[i].d;
    s.npending = n;
    std::atomic_thread_fence(std::memory_order_seq_cst);
    s.version.store(v + 2, std::memory_order_release);
    s.heartbeat_ns.store(now_ns(), std::memory_order_relaxed);
  }

  bool read_slot(int r, std::vector<ReqDesc>& out) {
    RankSlot& s = hdr->slots[r];
    for (int attempt = 0; attempt < 64; ++attempt) {
      const uint64_t v0 = s.version.load(std::memory_order_acquire);
      if (v0 & 1) continue;
      std::atomic_thread_fence(std::memory_order_seq_cst);
      const uint32_t n = std::min<uint32_t>(s.npending, kMaxPending);
      out.assign(s.pending, s.pending + n);
      std::atomic_thread_fence(std::memory_order_seq_cst);
      if (s.version.load(std::memory_order_acquire) == v0) return true;
    }
    return false;
  }

  void coordinate();
  bool execute_response();
  void exec_allreduce_group(std::vector<LocalReq*>& group);
  void exec_broadcast(LocalReq& q, const RespEntry& e);
  void exec_allgather(LocalReq& q, const RespEntry& e);
  void loop();
};

Runtime* g_rt = nullptr;

template <typename T>
void reduce_slice(Runtime& rt, size_t e0, size_t e1) {
  T* res = reinterpret_cast<T*>(rt.result());
  const T* r0 = reinterpret_cast<const T*>(rt.region(0));
  for (size_t i = e0; i < e1; ++i) res[i] = r0[i];
  for (int r = 1; r < rt.size; ++r) {
    const T* src = reinterpret_cast<const T*>(rt.region(r));
    for (size_t i = e0; i < e1; ++i) res[i] += src[i];
  }
}

template <typename T>
void copy_scaled(void* out, const void* in, size_t n, double s) {
  T* o = static_cast<T*>(out);
  const T* p = static_cast<const T*>(in);
  if (s == 1.0) {
    memcpy(o, p, n * sizeof(T));
  } else {
    for (size_t i = 0; i < n; ++i) o[i] = static_cast<T>(p[i] * s);
  }
}

}  // namespace

#include "runtime_exec.inc"
