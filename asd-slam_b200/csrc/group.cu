// group.cu -- multi-GPU groups (include/asdmatch.h "multi-GPU groups"): keyframe batches replicated over
// NVLink by the kernel that packs them, batched keyframe-pair searches dealt over the ranks, results stored by the resolve
// kernels straight into host memory (every rank over its own PCIe link) and / or into every rank's GPU, device-side
// barriers. Plain C ABI: no libtorch, no NCCL, no Python on the path. One process may drive all ranks
// (asdm_group_create) or one rank each (asdm_group_join); the code below is the same for both -- a process simply hosts
// `local_count` members of the `world`.
#include <errno.h>
#include <fcntl.h>
#include <sched.h>
#include <stdio.h>
#include <string.h>
#include <sys/mman.h>
#include <time.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <deque>
#include <map>
#include <new>

#include "jobs.cuh"

// ------------------------------------------------------------------------------------------------ window allocator
// Deterministic first fit over [0, size): every rank replays the same alloc / free sequence on its own copy, so a
// symmetric object sits at the same offset of every window without any handle exchange.
struct WinAlloc {
  std::map<size_t, size_t> free_;  // offset -> length
  std::mutex mu;                   // sets (and with them their slab's range) are released from any thread
  void reset(size_t lo, size_t hi) {
    free_.clear();
    free_[lo] = hi - lo;
  }
  bool alloc(size_t bytes, size_t* off) {
    std::lock_guard<std::mutex> lk(mu);
    bytes = align_up(std::max(bytes, (size_t)1), 4096);
    for (auto it = free_.begin(); it != free_.end(); ++it)
      if (it->second >= bytes) {
        *off = it->first;
        const size_t rest = it->second - bytes, at = it->first + bytes;
        free_.erase(it);
        if (rest) free_[at] = rest;
        return true;
      }
    return false;
  }
  void release(size_t off, size_t bytes) {
    std::lock_guard<std::mutex> lk(mu);
    bytes = align_up(std::max(bytes, (size_t)1), 4096);
    auto it = free_.emplace(off, bytes).first;
    auto nx = std::next(it);
    if (nx != free_.end() && it->first + it->second == nx->first) {
      it->second += nx->second;
      free_.erase(nx);
    }
    if (it != free_.begin()) {
      auto pv = std::prev(it);
      if (pv->first + pv->second == it->first) {
        pv->second += it->second;
        free_.erase(it);
      }
    }
  }
};

enum { WIN_HEADER = 4096 };  // barrier flags of the group live in front of the allocator's range

struct GroupMember {
  int rank = 0;
  asdm_ctx* ctx = nullptr;
  bool own_ctx = false;
  char* win = nullptr;               // this member's window
  char* peer[ASDM_MAX_PEERS] = {};   // every rank's window as this member's device addresses it (peer[rank] == win)
  bool opened[ASDM_MAX_PEERS] = {};  // cudaIpc mappings to close
  WinAlloc alloc;
  // Two barrier channels, each used on ONE stream so that its epochs are announced in order: 0 = the main stream (records
  // gathered by jobs), 1 = the preparation stream (keyframe batches replicated into the windows)
  uint32_t epoch[2] = {0, 0};        // barriers queued so far per channel (identical on every rank)
  int32_t* h_err = nullptr;          // page-locked: raised by a barrier that ran out of patience
  // recorded on the main stream whenever a range of this window is released: what is queued there up to that point may
  // still read the range, so whoever takes it next orders itself behind this event (and behind nothing queued later)
  cudaEvent_t rel_evt = nullptr;
  // The barriers that close the exchange of a job's records run on a stream of their own, behind the job's kernels, so
  // that the NEXT job's distance kernel already runs while the last rows of this one are still crossing NVLink
  cudaStream_t bar_stream = nullptr;
  cudaEvent_t ev_main = nullptr, ev_sync = nullptr;
};

struct HostCtl {  // shared by the processes of the group (one page)
  std::atomic<uint32_t> arrived;
  std::atomic<uint32_t> sense;
};

struct HostBlock {
  void* p;
  size_t bytes;
  int fd;         // memfd (multi-process) or -1
  bool shared;
};

struct asdm_group {
  int world = 1;
  int n_proc = 1;
  std::deque<GroupMember> m;  // (a member holds a mutex: never moved)
  asdm_allgather_fn xchg = nullptr;
  void* user = nullptr;
  size_t win_bytes = 0;
  std::recursive_mutex mu;
  std::string err;
  HostCtl* hctl = nullptr;
  int hctl_fd = -1;
  uint32_t host_sense = 0;
  std::vector<HostBlock> host_blocks;
};

static int gfail(asdm_group* g, int code, const char* what, cudaError_t e = cudaSuccess) {
  if (g) {
    char buf[512];
    if (e != cudaSuccess) snprintf(buf, sizeof buf, "%s: %s", what, cudaGetErrorString(e));
    else snprintf(buf, sizeof buf, "%s", what);
    g->err = buf;
  }
  if (e != cudaSuccess) cudaGetLastError();
  return code;
}
#define GCK(call)                                                      \
  do {                                                                 \
    cudaError_t e__ = (call);                                          \
    if (e__ != cudaSuccess) return gfail(g, e__ == cudaErrorMemoryAllocation ? ASDM_ERR_OOM : ASDM_ERR_CUDA, #call, e__); \
  } while (0)

// ------------------------------------------------------------------------------------------------ device barrier
// Rank r announces epoch e by storing it into slot r of every peer's flag array (release, system scope) and then waits
// until every slot of its own array has reached e (acquire). Stream ordered: a signal is queued behind the kernels whose
// peer stores it publishes, so "all flags >= e" means every rank's stores into this window have landed. The wait is
// bounded: a rank that never arrives raises h_err instead of hanging the GPU.
struct BarrierPeers {
  uint32_t* flags[ASDM_MAX_PEERS];
  int n;
};
__global__ void k_group_signal(BarrierPeers P, int rank, uint32_t epoch) {
  const int p = threadIdx.x;
  if (p >= P.n) return;
  __threadfence_system();
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(P.flags[p] + rank), "r"(epoch) : "memory");
}
__global__ void k_group_wait(const uint32_t* mine, int n, uint32_t epoch, int32_t* h_err, unsigned long long budget_ns) {
  const int p = threadIdx.x;
  if (p >= n) return;
  unsigned long long t0;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
  for (unsigned spin = 0;; spin++) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine + p) : "memory");
    if ((int32_t)(v - epoch) >= 0) break;
    if ((spin & 1023u) == 1023u) {
      unsigned long long t1;
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t1));
      if (t1 - t0 > budget_ns) {
        *h_err = 1;
        break;
      }
    }
    __nanosleep(64);
  }
}

// A barrier is queued in two halves so that a process hosting several ranks can queue every member's signal before any
// member's wait: the host never blocks on a stream whose wait kernel needs a signal it has not launched yet.
enum { CH_MAIN = 0, CH_PREP = 1 };
static int group_signal(asdm_group* g, GroupMember& M, int ch) {
  asdm_ctx* ctx = M.ctx;
  ++M.epoch[ch];
  if (g->world == 1) return ASDM_OK;
  GCK(cudaSetDevice(ctx->device));
  BarrierPeers P;
  P.n = g->world;
  for (int r = 0; r < g->world; r++) P.flags[r] = (uint32_t*)M.peer[r] + ch * ASDM_MAX_PEERS;
  k_group_signal<<<1, 32, 0, ch == CH_MAIN ? M.bar_stream : ctx->prep_stream>>>(P, M.rank, M.epoch[ch]);
  ctx->stat_total_launches++;
  GCK(cudaGetLastError());
  return ASDM_OK;
}
static int group_wait(asdm_group* g, GroupMember& M, int ch) {  // for the epoch of the channel's last group_signal
  asdm_ctx* ctx = M.ctx;
  if (g->world == 1) return ASDM_OK;
  GCK(cudaSetDevice(ctx->device));
  k_group_wait<<<1, 32, 0, ch == CH_MAIN ? M.bar_stream : ctx->prep_stream>>>((const uint32_t*)M.win + ch * ASDM_MAX_PEERS, g->world,
                                                                           M.epoch[ch], M.h_err, 10ull * 1000000000ull);
  ctx->stat_total_launches++;
  GCK(cudaGetLastError());
  return ASDM_OK;
}

// ------------------------------------------------------------------------------------------------ host-side plumbing
static int host_ctl_barrier(asdm_group* g) {
  if (g->n_proc <= 1 || !g->hctl) return ASDM_OK;
  const uint32_t want = (g->host_sense ^= 1u);
  if (g->hctl->arrived.fetch_add(1, std::memory_order_acq_rel) + 1 == (uint32_t)g->n_proc) {
    g->hctl->arrived.store(0, std::memory_order_relaxed);
    g->hctl->sense.store(want, std::memory_order_release);
    return ASDM_OK;
  }
  timespec t0;
  clock_gettime(CLOCK_MONOTONIC, &t0);
  for (unsigned spin = 0; g->hctl->sense.load(std::memory_order_acquire) != want; spin++) {
    if ((spin & 0xfff) == 0xfff) {
      timespec t1;
      clock_gettime(CLOCK_MONOTONIC, &t1);
      if (t1.tv_sec - t0.tv_sec > 60) return gfail(g, ASDM_ERR_INTERNAL, "host barrier: a process of the group never arrived");
      sched_yield();
    }
  }
  return ASDM_OK;
}

// shared host mapping: the lowest local rank == 0 creates a memfd, the others open it through /proc/<pid>/fd/<fd>
struct FdBlob {
  int32_t pid, fd;
  int64_t bytes;
};
static int shared_host_map(asdm_group* g, size_t bytes, void** out, int* fd_out) {
  *out = nullptr;
  *fd_out = -1;
  const bool creator = g->m[0].rank == 0;
  FdBlob mine{0, -1, (int64_t)bytes};
  int fd = -1;
  if (creator) {
    fd = memfd_create("asdm_group", 0);
    if (fd < 0 || ftruncate(fd, (off_t)bytes) != 0) {
      if (fd >= 0) close(fd);
      fd = -1;
    }
    mine.pid = (int32_t)getpid();
    mine.fd = fd;
  }
  std::vector<FdBlob> all((size_t)g->n_proc);
  if (g->xchg(g->user, &mine, all.data(), sizeof(FdBlob)) != 0) return gfail(g, ASDM_ERR_INTERNAL, "allgather callback failed");
  if (all[0].fd < 0) return gfail(g, ASDM_ERR_OOM, "memfd_create / ftruncate failed for the shared host block");
  if (!creator) {
    char path[64];
    snprintf(path, sizeof path, "/proc/%d/fd/%d", all[0].pid, all[0].fd);
    fd = open(path, O_RDWR);
  }
  void* p = fd >= 0 ? mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0) : MAP_FAILED;
  int32_t ok = p != MAP_FAILED ? 1 : 0;
  std::vector<int32_t> oks((size_t)g->n_proc);
  if (g->xchg(g->user, &ok, oks.data(), sizeof ok) != 0) return gfail(g, ASDM_ERR_INTERNAL, "allgather callback failed");
  bool all_ok = true;
  for (int32_t v : oks) all_ok &= v != 0;
  if (!all_ok) {
    if (p != MAP_FAILED) munmap(p, bytes);
    if (fd >= 0) close(fd);
    return gfail(g, ASDM_ERR_INTERNAL, "a process of the group could not map the shared host block (/proc/<pid>/fd)");
  }
  *out = p;
  *fd_out = fd;
  return ASDM_OK;
}

// ------------------------------------------------------------------------------------------------ life cycle
static int group_finish_setup(asdm_group* g) {
  for (GroupMember& M : g->m) {
    GCK(cudaSetDevice(M.ctx->device));
    GCK(cudaMemsetAsync(M.win, 0, WIN_HEADER, M.ctx->stream));
    GCK(cudaStreamSynchronize(M.ctx->stream));
    GCK(cudaHostAlloc((void**)&M.h_err, 64, cudaHostAllocMapped | cudaHostAllocPortable));
    *M.h_err = 0;
    GCK(cudaEventCreateWithFlags(&M.rel_evt, cudaEventDisableTiming));
    GCK(cudaEventRecord(M.rel_evt, M.ctx->stream));
    GCK(cudaStreamCreateWithFlags(&M.bar_stream, cudaStreamNonBlocking));
    GCK(cudaEventCreateWithFlags(&M.ev_main, cudaEventDisableTiming));
    GCK(cudaEventCreateWithFlags(&M.ev_sync, cudaEventDisableTiming));
    M.alloc.reset(WIN_HEADER, g->win_bytes);
  }
  return ASDM_OK;
}

extern "C" int asdm_group_create(const int* devices, int n, size_t window_bytes, asdm_group** out) {
  if (!out || !devices || n < 1 || n > ASDM_MAX_PEERS) return ASDM_ERR_INVALID;
  *out = nullptr;
  asdm_group* g = new (std::nothrow) asdm_group();
  if (!g) return ASDM_ERR_OOM;
  g->world = n;
  g->n_proc = 1;
  g->win_bytes = window_bytes ? align_up(window_bytes, 4096) : ((size_t)2 << 30);
  g->m.resize((size_t)n);
  auto bail = [&](int code) {
    asdm_group_destroy(g);
    return code;
  };
  for (int r = 0; r < n; r++) {
    GroupMember& M = g->m[r];
    M.rank = r;
    M.own_ctx = true;
    const int rc = asdm_create(devices[r], &M.ctx);
    if (rc) return bail(rc);
    if (cudaSetDevice(devices[r]) != cudaSuccess || cudaMalloc((void**)&M.win, g->win_bytes) != cudaSuccess) {
      cudaGetLastError();
      return bail(ASDM_ERR_OOM);
    }
  }
  for (int a = 0; a < n; a++) {
    cudaSetDevice(devices[a]);
    for (int b = 0; b < n; b++) {
      g->m[a].peer[b] = g->m[b].win;
      if (devices[a] == devices[b]) continue;
      int can = 0;
      if (cudaDeviceCanAccessPeer(&can, devices[a], devices[b]) != cudaSuccess || !can) {
        cudaGetLastError();
        return bail(ASDM_ERR_CUDA);  // no peer access between these GPUs: a group needs NVLink / PCIe P2P
      }
      const cudaError_t e = cudaDeviceEnablePeerAccess(devices[b], 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
        cudaGetLastError();
        return bail(ASDM_ERR_CUDA);
      }
      cudaGetLastError();
    }
  }
  const int rc = group_finish_setup(g);
  if (rc) return bail(rc);
  *out = g;
  return ASDM_OK;
}

struct JoinBlob {
  cudaIpcMemHandle_t handle;
  int32_t pid, device;
};

extern "C" int asdm_group_join(asdm_ctx* ctx, int rank, int world, size_t window_bytes, asdm_allgather_fn allgather, void* user,
                               asdm_group** out) {
  if (!out || !ctx || world < 1 || world > ASDM_MAX_PEERS || rank < 0 || rank >= world || (world > 1 && !allgather))
    return ASDM_ERR_INVALID;
  *out = nullptr;
  asdm_group* g = new (std::nothrow) asdm_group();
  if (!g) return ASDM_ERR_OOM;
  g->world = world;
  g->n_proc = world;
  g->xchg = allgather;
  g->user = user;
  g->win_bytes = window_bytes ? align_up(window_bytes, 4096) : ((size_t)2 << 30);
  g->m.resize(1);
  GroupMember& M = g->m[0];
  M.rank = rank;
  M.ctx = ctx;
  auto bail = [&](int code) {
    asdm_group_destroy(g);
    return code;
  };
  JoinBlob mine;
  memset(&mine, 0, sizeof mine);
  int32_t ok = 1;
  if (cudaSetDevice(ctx->device) != cudaSuccess || cudaMalloc((void**)&M.win, g->win_bytes) != cudaSuccess ||
      cudaIpcGetMemHandle(&mine.handle, M.win) != cudaSuccess) {
    cudaGetLastError();
    ok = 0;
  }
  mine.pid = (int32_t)getpid();
  mine.device = ok ? ctx->device : -1;
  M.peer[rank] = M.win;
  if (world > 1) {
    std::vector<JoinBlob> all((size_t)world);
    if (allgather(user, &mine, all.data(), sizeof mine) != 0) return bail(gfail(g, ASDM_ERR_INTERNAL, "allgather callback failed"));
    for (int r = 0; r < world; r++) ok &= all[r].device >= 0 ? 1 : 0;
    if (!ok) return bail(ASDM_ERR_OOM);
    for (int r = 0; r < world; r++) {
      if (r == rank) continue;
      void* p = nullptr;
      if (cudaIpcOpenMemHandle(&p, all[r].handle, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
        cudaGetLastError();
        return bail(ASDM_ERR_CUDA);  // no peer access between these GPUs
      }
      M.peer[r] = (char*)p;
      M.opened[r] = true;
    }
    void* hp = nullptr;
    const int rc = shared_host_map(g, 4096, &hp, &g->hctl_fd);
    if (rc) return bail(rc);
    g->hctl = (HostCtl*)hp;  // a fresh memfd is zero filled: arrived = 0, sense = 0
  } else if (!ok) {
    return bail(ASDM_ERR_OOM);
  }
  const int rc = group_finish_setup(g);
  if (rc) return bail(rc);
  if (world > 1) {  // nobody signals into a window that is still being cleared
    const int rb = host_ctl_barrier(g);
    if (rb) return bail(rb);
  }
  *out = g;
  return ASDM_OK;
}

extern "C" int asdm_group_destroy(asdm_group* g) {
  if (!g) return ASDM_ERR_INVALID;
  for (GroupMember& M : g->m) {
    if (!M.ctx) continue;
    cudaSetDevice(M.ctx->device);
    cudaStreamSynchronize(M.ctx->stream);
  }
  if (g->n_proc > 1 && g->hctl) host_ctl_barrier(g);  // nobody still stores into a window that is about to go away
  for (HostBlock& b : g->host_blocks) {
    if (b.shared) {
      cudaHostUnregister(b.p);
      munmap(b.p, b.bytes);
      if (b.fd >= 0) close(b.fd);
    } else {
      cudaFreeHost(b.p);
    }
  }
  for (GroupMember& M : g->m) {
    if (!M.ctx) continue;
    cudaSetDevice(M.ctx->device);
    for (int r = 0; r < g->world; r++)
      if (M.opened[r]) cudaIpcCloseMemHandle(M.peer[r]);
    if (M.win) cudaFree(M.win);
    if (M.h_err) cudaFreeHost(M.h_err);
    if (M.rel_evt) cudaEventDestroy(M.rel_evt);
    if (M.bar_stream) {
      cudaStreamSynchronize(M.bar_stream);
      cudaStreamDestroy(M.bar_stream);
    }
    if (M.ev_main) cudaEventDestroy(M.ev_main);
    if (M.ev_sync) cudaEventDestroy(M.ev_sync);
    if (M.own_ctx) asdm_destroy(M.ctx);
  }
  cudaGetLastError();
  if (g->hctl) munmap(g->hctl, 4096);
  if (g->hctl_fd >= 0) close(g->hctl_fd);
  delete g;
  return ASDM_OK;
}

extern "C" const char* asdm_group_last_error(const asdm_group* g) { return g ? g->err.c_str() : "null group"; }
extern "C" int asdm_group_world(const asdm_group* g) { return g ? g->world : ASDM_ERR_INVALID; }
extern "C" int asdm_group_local_count(const asdm_group* g) { return g ? (int)g->m.size() : ASDM_ERR_INVALID; }
extern "C" int asdm_group_local_rank(const asdm_group* g, int l) {
  return (g && l >= 0 && l < (int)g->m.size()) ? g->m[l].rank : ASDM_ERR_INVALID;
}
extern "C" asdm_ctx* asdm_group_ctx(asdm_group* g, int l) { return (g && l >= 0 && l < (int)g->m.size()) ? g->m[l].ctx : nullptr; }

extern "C" int asdm_group_host_alloc(asdm_group* g, size_t bytes, void** out) {
  if (!g || !out || bytes == 0) return ASDM_ERR_INVALID;
  std::lock_guard<std::recursive_mutex> lk(g->mu);
  *out = nullptr;
  bytes = align_up(bytes, 4096);
  HostBlock b{nullptr, bytes, -1, g->n_proc > 1};
  if (!b.shared) {
    GCK(cudaHostAlloc(&b.p, bytes, cudaHostAllocMapped | cudaHostAllocPortable));
  } else {
    const int rc = shared_host_map(g, bytes, &b.p, &b.fd);
    if (rc) return rc;
    const cudaError_t e = cudaHostRegister(b.p, bytes, cudaHostRegisterMapped | cudaHostRegisterPortable);
    if (e != cudaSuccess) {
      munmap(b.p, bytes);
      if (b.fd >= 0) close(b.fd);
      return gfail(g, ASDM_ERR_CUDA, "cudaHostRegister(shared host block)", e);
    }
  }
  g->host_blocks.push_back(b);
  *out = b.p;
  return ASDM_OK;
}

extern "C" int asdm_group_host_free(asdm_group* g, void* p) {
  if (!g || !p) return ASDM_ERR_INVALID;
  std::lock_guard<std::recursive_mutex> lk(g->mu);
  for (size_t k = 0; k < g->host_blocks.size(); k++) {
    HostBlock& b = g->host_blocks[k];
    if (b.p != p) continue;
    for (GroupMember& M : g->m) {
      cudaSetDevice(M.ctx->device);
      cudaStreamSynchronize(M.ctx->stream);
    }
    if (b.shared) {
      host_ctl_barrier(g);
      cudaHostUnregister(b.p);
      munmap(b.p, b.bytes);
      if (b.fd >= 0) close(b.fd);
    } else {
      cudaFreeHost(b.p);
    }
    cudaGetLastError();
    g->host_blocks.erase(g->host_blocks.begin() + (long)k);
    return ASDM_OK;
  }
  return gfail(g, ASDM_ERR_INVALID, "not a block of asdm_group_host_alloc");
}

extern "C" int asdm_group_host_barrier(asdm_group* g) {
  if (!g) return ASDM_ERR_INVALID;
  std::lock_guard<std::recursive_mutex> lk(g->mu);
  return host_ctl_barrier(g);
}

// ------------------------------------------------------------------------------------------------ replicated sets
static void window_release(GroupMember& M, size_t off, size_t bytes) {
  cudaSetDevice(M.ctx->device);
  if (cudaEventRecord(M.rel_evt, M.ctx->stream) != cudaSuccess) cudaGetLastError();
  M.alloc.release(off, bytes);
}

void group_slab_release(GroupSlab* s) {
  asdm_group* g = s->g;
  GroupMember& M = g->m[(size_t)s->member];
  window_release(M, s->off, s->bytes);
  if (s->h_flags) cudaFreeHost(s->h_flags);
  delete s;
}

struct PushItem {  // one keyframe of the batch; offsets into the (every) window
  unsigned long long o_desc, o_bf, o_norm, o_side;
  uint32_t side16;  // 16-byte pieces of keypoints | node ids | FeatureVector positions
  int32_t n, n_pad, flag_idx;
};
struct PeerTable {
  char* base[ASDM_MAX_PEERS];
  int n, self;
  unsigned long long o_flags;
};

// The owner of a keyframe packs it (bf16 copy, norms, range flag: as k_pack_batch) and replicates it in the same pass:
// every row is read once from the owner's window (where the PCIe upload landed) and stored into every window of the
// group -- fp32 rows to the peers, bf16 / norms to all. grid: (row chunks, owned keyframes).
__global__ void __launch_bounds__(256) k_pack_push(const PushItem* __restrict__ items, PeerTable P) {
  const PushItem it = items[blockIdx.y];
  const int lane = threadIdx.x & 31;
  char* self = P.base[P.self];
  for (int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; row < it.n_pad; row += (gridDim.x * blockDim.x) >> 5) {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    const size_t at = (size_t)row * ASDM_DESC_DIM * 4 + (size_t)lane * 16;
    if (row < it.n) v = *reinterpret_cast<const float4*>(self + it.o_desc + at);
    float s = v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y);
    __nv_bfloat162 hi = __floats2bfloat162_rn(v.z, v.w);
    uint2 packed;
    packed.x = *reinterpret_cast<uint32_t*>(&lo);
    packed.y = *reinterpret_cast<uint32_t*>(&hi);
    const bool ok = (s == 0.f) || (s >= 9.5367431640625e-07f && s <= 1048576.f);  // [2^-20, 2^20], as k_pack_desc
    for (int p = 0; p < P.n; p++) {
      char* b = P.base[p];
      if (p != P.self || row >= it.n) *reinterpret_cast<float4*>(b + it.o_desc + at) = v;
      *reinterpret_cast<uint2*>(b + it.o_bf + (size_t)row * ASDM_DESC_DIM * 2 + (size_t)lane * 8) = packed;
      if (lane == 0) {
        *reinterpret_cast<float*>(b + it.o_norm + (size_t)row * 4) = s;
        if (!ok) reinterpret_cast<int32_t*>(b + P.o_flags)[it.flag_idx] = 1;
      }
    }
  }
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < it.side16; i += gridDim.x * blockDim.x) {
    const uint4 x = *reinterpret_cast<const uint4*>(self + it.o_side + (size_t)i * 16);
    for (int p = 0; p < P.n; p++)
      if (p != P.self) *reinterpret_cast<uint4*>(P.base[p] + it.o_side + (size_t)i * 16) = x;
  }
}

extern "C" int asdm_group_sets_create(asdm_group* g, const asdm_set_desc* descs, int n_sets, asdm_set** out) {
  if (!g || n_sets < 0 || (n_sets > 0 && (!descs || !out))) return ASDM_ERR_INVALID;
  std::lock_guard<std::recursive_mutex> lk(g->mu);
  const int L = (int)g->m.size(), W = g->world;
  for (int k = 0; k < n_sets * L; k++) out[k] = nullptr;
  if (n_sets == 0) return ASDM_OK;
  // ---- layout of the slab (identical in every window): flags | push items | per keyframe: rows | bf16 | norms | side
  struct Lay {
    size_t o_desc, o_bf, o_norm, o_kps, o_fv, side_bytes;
    int n_pad;
  };
  std::vector<Lay> lay((size_t)n_sets);
  size_t off = 0;
  const size_t o_flags = off; off += align_up((size_t)n_sets * 4, 256);
  const size_t o_items = off; off += align_up((size_t)n_sets * sizeof(PushItem), 256);
  for (int k = 0; k < n_sets; k++) {
    const asdm_set_desc& d = descs[k];
    if (d.n < 0) return gfail(g, ASDM_ERR_INVALID, "bad set description");
    if (d.row_stride_bytes != 0 && d.row_stride_bytes < (int)(ASDM_DESC_DIM * sizeof(float)))
      return gfail(g, ASDM_ERR_INVALID, "row stride < 512 B");
    if (d.grid && (!d.kps || d.grid->cols <= 0 || d.grid->rows <= 0 || (int64_t)d.grid->cols * d.grid->rows > (1 << 20)))
      return gfail(g, ASDM_ERR_INVALID, "bad grid");
    Lay& y = lay[(size_t)k];
    y.n_pad = (int)align_up((size_t)std::max(d.n, 1), 256);
    y.o_desc = off; off += align_up((size_t)y.n_pad * ASDM_DESC_DIM * sizeof(float), 1024);
    y.o_bf = off;   off += align_up((size_t)y.n_pad * ASDM_DESC_DIM * sizeof(__nv_bfloat16), 1024);
    y.o_norm = off; off += align_up((size_t)y.n_pad * sizeof(float), 256);
    y.o_kps = off;
    y.side_bytes = (d.kps ? (size_t)y.n_pad * sizeof(asdm_kp) : 0) + (d.featvec ? (size_t)y.n_pad * 8 : 0);
    y.o_fv = y.o_kps + (d.kps ? (size_t)y.n_pad * sizeof(asdm_kp) : 0);
    off += align_up(y.side_bytes, 256);
  }
  const size_t slab_bytes = off;

  std::vector<GroupSlab*> slabs((size_t)L, nullptr);
  auto bail = [&](int code) {
    for (int l = 0; l < L; l++) {
      for (int k = 0; k < n_sets; k++)
        if (out[k * L + l]) {
          asdm_set* s = out[k * L + l];
          s->slab = nullptr;  // the slab is returned below, once
          s->desc = nullptr;
          s->h_flag = nullptr;
          set_free(s);
          out[k * L + l] = nullptr;
        }
      if (slabs[(size_t)l]) group_slab_release(slabs[(size_t)l]);
    }
    return code;
  };
#define GCB(call)                                                                 \
  do {                                                                            \
    cudaError_t e__ = (call);                                                     \
    if (e__ != cudaSuccess) return bail(gfail(g, ASDM_ERR_CUDA, #call, e__));     \
  } while (0)

  for (int l = 0; l < L; l++) {
    GroupMember& M = g->m[(size_t)l];
    asdm_ctx* ctx = M.ctx;
    std::lock_guard<std::recursive_mutex> lkc(ctx->mu);
    GCB(cudaSetDevice(ctx->device));
    size_t base_off = 0;
    if (!M.alloc.alloc(slab_bytes, &base_off)) return bail(gfail(g, ASDM_ERR_OOM, "group window exhausted (asdm_group_create window_bytes)"));
    GroupSlab* slab = new (std::nothrow) GroupSlab();
    if (!slab) {
      M.alloc.release(base_off, slab_bytes);
      return bail(gfail(g, ASDM_ERR_OOM, "host allocation"));
    }
    slab->g = g; slab->member = l; slab->off = base_off; slab->bytes = slab_bytes;
    slabs[(size_t)l] = slab;
    GCB(cudaHostAlloc((void**)&slab->h_flags, align_up((size_t)n_sets * 4, 64), cudaHostAllocDefault));
    for (int k = 0; k < n_sets; k++) slab->h_flags[k] = 1;  // "not proven" until the device flags are back
    char* win = M.win + base_off;
    // The range may have had an occupant: whatever this rank still has queued on it was queued on the main stream before
    // the release (rel_evt). The upload and the preparation streams order themselves behind that -- and behind nothing
    // queued later, so the searches of earlier waves run while this batch crosses PCIe. My own range is mine to clear:
    // peers only store into it after the barrier below.
    GCB(cudaStreamWaitEvent(ctx->upload_stream, M.rel_evt, 0));
    GCB(cudaStreamWaitEvent(ctx->prep_stream, M.rel_evt, 0));
    GCB(cudaMemsetAsync(win + o_flags, 0, align_up((size_t)n_sets * 4, 256), ctx->prep_stream));
    // B1: every rank has left behind whatever it had queued on the previous occupant of this range
    {
      int rc = group_signal(g, M, CH_PREP);
      if (rc) return bail(rc);
    }
  }
  for (int l = 0; l < L; l++) {
    GroupMember& M = g->m[(size_t)l];
    asdm_ctx* ctx = M.ctx;
    std::lock_guard<std::recursive_mutex> lkc(ctx->mu);
    GCB(cudaSetDevice(ctx->device));
    const size_t base_off = slabs[(size_t)l]->off;
    char* win = M.win + base_off;
    {
      int rc = group_wait(g, M, CH_PREP);
      if (rc) return bail(rc);
    }
    // ---- uploads of the keyframes this rank owns (k % W == rank): rows straight from the caller's buffer, the small
    // arrays through the page-locked arena, the push table last
    int n_owned = 0, max_pad = 0;
    for (int k = M.rank; k < n_sets; k += W) n_owned++;
    void* hst = nullptr;
    size_t side_total = 0;
    for (int k = M.rank; k < n_sets; k += W) side_total += align_up(lay[(size_t)k].side_bytes, 256);
    const size_t items_bytes = align_up((size_t)std::max(n_owned, 1) * sizeof(PushItem), 256);
    {
      int rc = stage_alloc(ctx, items_bytes + side_total, &hst);
      if (rc) return bail(rc);
    }
    PushItem* h_items = (PushItem*)hst;
    char* h_side = (char*)hst + items_bytes;
    size_t side_at = 0;
    int oi = 0;
    for (int k = M.rank; k < n_sets; k += W, oi++) {
      const asdm_set_desc& d = descs[k];
      const Lay& y = lay[(size_t)k];
      max_pad = std::max(max_pad, y.n_pad);
      if (d.n > 0) {
        if (!d.desc) return bail(gfail(g, ASDM_ERR_INVALID, "the owner of a keyframe needs its descriptors"));
        const int stride = d.row_stride_bytes ? d.row_stride_bytes : (int)(ASDM_DESC_DIM * sizeof(float));
        if (stride == (int)(ASDM_DESC_DIM * sizeof(float)))
          GCB(cudaMemcpyAsync(win + y.o_desc, d.desc, (size_t)d.n * ASDM_DESC_DIM * sizeof(float), cudaMemcpyDefault, ctx->upload_stream));
        else
          GCB(cudaMemcpy2DAsync(win + y.o_desc, ASDM_DESC_DIM * sizeof(float), d.desc, stride, ASDM_DESC_DIM * sizeof(float), d.n,
                                cudaMemcpyDefault, ctx->upload_stream));
      }
      char* hs = h_side + side_at;
      if (d.kps) {
        memcpy(hs, d.kps, (size_t)d.n * sizeof(asdm_kp));
        memset(hs + (size_t)d.n * sizeof(asdm_kp), 0, (size_t)(y.n_pad - d.n) * sizeof(asdm_kp));
      }
      if (d.featvec) {
        uint32_t* hn = (uint32_t*)(hs + (y.o_fv - y.o_kps));
        int rc = flatten_featvec(ctx, d.featvec, d.n, y.n_pad, hn, (int32_t*)(hn + y.n_pad));
        if (rc) return bail(gfail(g, rc, asdm_last_error(ctx)));
      }
      if (y.side_bytes)
        GCB(cudaMemcpyAsync(win + y.o_kps, hs, y.side_bytes, cudaMemcpyHostToDevice, ctx->upload_stream));
      side_at += align_up(y.side_bytes, 256);
      h_items[oi] = PushItem{base_off + y.o_desc, base_off + y.o_bf, base_off + y.o_norm, base_off + y.o_kps,
                             (uint32_t)(y.side_bytes / 16), d.n, y.n_pad, k};
    }
    if (n_owned)
      GCB(cudaMemcpyAsync(win + o_items, h_items, (size_t)n_owned * sizeof(PushItem), cudaMemcpyHostToDevice, ctx->upload_stream));
    GCB(cudaEventRecord(ctx->upload_done, ctx->upload_stream));
    GCB(cudaStreamWaitEvent(ctx->prep_stream, ctx->upload_done, 0));
    if (n_owned) {
      PeerTable P;
      P.n = W;
      P.self = M.rank;
      P.o_flags = base_off + o_flags;
      for (int r = 0; r < W; r++) P.base[r] = M.peer[r];
      k_pack_push<<<dim3(std::min((max_pad * 32 + 255) / 256, 64), n_owned), 256, 0, ctx->prep_stream>>>((const PushItem*)(win + o_items), P);
      ctx->stat_total_launches++;
      GCB(cudaGetLastError());
    }
    // B2: every owner's rows have landed in this window
    {
      int rc = group_signal(g, M, CH_PREP);
      if (rc) return bail(rc);
    }
  }
  for (int l = 0; l < L; l++) {
    GroupMember& M = g->m[(size_t)l];
    asdm_ctx* ctx = M.ctx;
    std::lock_guard<std::recursive_mutex> lkc(ctx->mu);
    GCB(cudaSetDevice(ctx->device));
    GroupSlab* slab = slabs[(size_t)l];
    char* win = M.win + slab->off;
    {
      int rc = group_wait(g, M, CH_PREP);
      if (rc) return bail(rc);
    }
    GCB(cudaMemcpyAsync(slab->h_flags, win + o_flags, (size_t)n_sets * 4, cudaMemcpyDeviceToHost, ctx->prep_stream));
    GCB(cudaEventRecord(ctx->prep_done, ctx->prep_stream));
    GCB(cudaStreamWaitEvent(ctx->stream, ctx->prep_done, 0));  // whatever is queued on the main stream from here on sees the batch
    // ---- the views
    for (int k = 0; k < n_sets; k++) {
      const asdm_set_desc& d = descs[k];
      const Lay& y = lay[(size_t)k];
      asdm_set* s = new (std::nothrow) asdm_set();
      if (!s) return bail(gfail(g, ASDM_ERR_OOM, "host allocation"));
      out[k * L + l] = s;
      s->ctx = ctx;
      s->slab = slab;
      slab->ref++;
      s->n = d.n;
      s->n_pad = y.n_pad;
      s->desc = (float*)(win + y.o_desc);
      s->desc_bf = (__nv_bfloat16*)(win + y.o_bf);
      s->norm2 = (float*)(win + y.o_norm);
      s->d_flags = (int32_t*)(win + o_flags) + k;
      s->h_flag = slab->h_flags + k;
      if (d.kps) {
        s->kps = (asdm_kp*)(win + y.o_kps);
        s->has_kps = true;
      }
      if (d.grid) {
        s->grid = *d.grid;
        s->has_grid = true;
      }
      if (d.featvec) {
        s->node_id = (uint32_t*)(win + y.o_fv);
        s->fv_pos = (int32_t*)(win + y.o_fv + (size_t)y.n_pad * 4);
        s->has_fv = true;
      }
      if (set_mark_ready(ctx, s, ctx->prep_stream)) return bail(gfail(g, ASDM_ERR_CUDA, "cudaEventRecord(set ready)"));
    }
  }
#undef GCB
  return ASDM_OK;
}

// ------------------------------------------------------------------------------------------------ sharded jobs
enum { GJOB_GENS = 3 };  // generations of the gathered records (see asdm_group_job_run)
struct asdm_gjob {
  asdm_group* g = nullptr;
  int kind = 0, n_pairs = 0, n_units = 0, stride = 1, gather = 0;
  bool split = false;
  std::vector<int32_t> unit_pair;
  struct Local {
    asdm_job* job = nullptr;
    int n_local = 0;
    DevBuf row_map;
    bool has_gather = false;
    cudaEvent_t ev_bar[GJOB_GENS] = {nullptr, nullptr, nullptr};  // barrier behind the run that filled generation g has completed
    bool bar_valid[GJOB_GENS] = {false, false, false};
  };
  std::vector<Local> loc;
  // ASDM_GATHER_DEVICE: GJOB_GENS generations of [rows][stride] match | dist and [units] counts at the same offset of every window
  size_t g_off = 0, g_bytes = 0, o_dist = 0, o_nm = 0, gen_bytes = 0;
  int gen = GJOB_GENS - 1;
  // staging of asdm_group_job_fetch
  int32_t* st_match = nullptr;
  float* st_dist = nullptr;
  int32_t* st_nm = nullptr;
  bool ran = false;
};

// The unit plan of a group job: pure host arithmetic, the same on every rank (no device, no communication).
extern "C" int asdm_group_plan_units(const int32_t* query_rows, int n_pairs, int blocks_per_pair, int world, int32_t* unit_pair,
                                     int32_t* unit_row0, int32_t* unit_row1, int32_t* unit_rank, int cap, int* n_units) {
  if (n_pairs < 0 || (n_pairs > 0 && !query_rows) || blocks_per_pair < 1 || world < 1 || !n_units) return ASDM_ERR_INVALID;
  int nu = 0;
  for (int p = 0; p < n_pairs; p++) {
    const int n = query_rows[p];
    if (n < 0) return ASDM_ERR_INVALID;
    // <= blocks_per_pair blocks on 128-row boundaries: no tile of the tensor-core kernel is computed twice
    const int tiles = (n + 127) / 128, nb = std::max(1, std::min(blocks_per_pair, tiles));
    for (int b = 0; b < nb; b++, nu++) {
      if (nu >= cap) return ASDM_ERR_INVALID;
      if (unit_pair) unit_pair[nu] = p;
      if (unit_row0) unit_row0[nu] = std::min(n, (int)((long)tiles * b / nb) * 128);
      if (unit_row1) unit_row1[nu] = b + 1 == nb ? n : std::min(n, (int)((long)tiles * (b + 1) / nb) * 128);
      if (unit_rank) unit_rank[nu] = nu % world;  // round-robin deal: unit u runs on rank u % world
    }
  }
  *n_units = nu;
  return ASDM_OK;
}

extern "C" int asdm_group_job_prepare(asdm_group* g, const asdm_gjob_desc* gd, asdm_gjob** out) {
  if (!g || !gd || !out) return ASDM_ERR_INVALID;
  *out = nullptr;
  const asdm_job_desc& d = gd->job;
  std::lock_guard<std::recursive_mutex> lk(g->mu);
  const int L = (int)g->m.size(), W = g->world;
  const int bpp = std::max(1, (int)gd->blocks_per_pair);
  if (d.kind != ASDM_JOB_BOW_KF_KF && d.kind != ASDM_JOB_TRIANGULATION) return gfail(g, ASDM_ERR_INVALID, "unknown job kind");
  if (bpp > 1 && (d.kind != ASDM_JOB_TRIANGULATION || d.check_ori))
    return gfail(g, ASDM_ERR_INVALID, "row-block units need a triangulation job without rotation check");
  if (d.row_begin || d.row_end) return gfail(g, ASDM_ERR_INVALID, "group jobs cut their own row blocks (blocks_per_pair)");
  if (d.n_kf <= 0 || d.n_pairs < 0 || !d.sets || !d.masks || (d.n_pairs > 0 && (!d.ia || !d.ib)))
    return gfail(g, ASDM_ERR_INVALID, "bad job description");
  for (int k = 0; k < d.n_kf * L; k++)
    if (!d.sets[k]) return gfail(g, ASDM_ERR_INVALID, "job.sets must hold n_kf * local_count sets (asdm_group_sets_create)");
  asdm_gjob* j = new (std::nothrow) asdm_gjob();
  if (!j) return gfail(g, ASDM_ERR_OOM, "host allocation");
  j->g = g;
  j->kind = d.kind;
  j->n_pairs = d.n_pairs;
  j->gather = gd->gather;
  j->split = bpp > 1;
  j->loc.resize((size_t)L);
  auto bail = [&](int code) {
    asdm_group_job_destroy(j);
    return code;
  };
  // units: pair p as a whole, or its query rows cut into <= bpp blocks on 128-row boundaries (asdm_group_plan_units)
  std::vector<int32_t> u_r0, u_r1, rows_of((size_t)std::max(d.n_pairs, 1));
  int stride = 1;
  for (int p = 0; p < d.n_pairs; p++) {
    if (d.ia[p] < 0 || d.ia[p] >= d.n_kf || d.ib[p] < 0 || d.ib[p] >= d.n_kf) return bail(gfail(g, ASDM_ERR_INVALID, "pair index out of range"));
    rows_of[(size_t)p] = d.sets[(size_t)d.ia[p] * L]->n;
    stride = std::max(stride, rows_of[(size_t)p]);
  }
  {
    const size_t cap = (size_t)std::max(d.n_pairs, 1) * bpp;
    j->unit_pair.resize(cap);
    u_r0.resize(cap);
    u_r1.resize(cap);
    int nu = 0;
    const int rc = asdm_group_plan_units(rows_of.data(), d.n_pairs, bpp, W, j->unit_pair.data(), u_r0.data(), u_r1.data(), nullptr,
                                         (int)cap, &nu);
    if (rc) return bail(gfail(g, rc, "unit plan"));
    j->unit_pair.resize((size_t)nu);
    j->n_units = nu;
  }
  j->stride = stride;
  if (j->gather & ASDM_GATHER_DEVICE) {
    const size_t rows = (size_t)std::max(d.n_pairs, 1);
    j->o_dist = align_up(rows * stride * 4, 256);
    j->o_nm = j->o_dist + align_up(rows * stride * 4, 256);
    j->gen_bytes = j->o_nm + align_up((size_t)std::max(j->n_units, 1) * 4, 256);
    j->g_bytes = GJOB_GENS * j->gen_bytes;
  }
  for (int l = 0; l < L; l++) {
    GroupMember& M = g->m[(size_t)l];
    asdm_ctx* ctx = M.ctx;
    std::vector<asdm_set*> sets((size_t)d.n_kf);
    for (int k = 0; k < d.n_kf; k++) {
      sets[(size_t)k] = d.sets[(size_t)k * L + l];
      if (sets[(size_t)k]->ctx->device != ctx->device || sets[(size_t)k]->ctx->family != ctx->family)
        return bail(gfail(g, ASDM_ERR_INVALID, "job.sets[k * local_count + l] must live on the GPU of local member l"));
    }
    std::vector<int32_t> ia, ib, rb, re, rmap;
    std::vector<float> F12, ep;
    for (int u = M.rank; u < j->n_units; u += W) {
      const int p = j->unit_pair[(size_t)u];
      ia.push_back(d.ia[p]);
      ib.push_back(d.ib[p]);
      rb.push_back(u_r0[(size_t)u]);
      re.push_back(u_r1[(size_t)u]);
      rmap.push_back(p);
      if (d.kind == ASDM_JOB_TRIANGULATION) {
        if (!d.F12 || !d.epipoles) return bail(gfail(g, ASDM_ERR_INVALID, "triangulation job needs F12 and epipoles"));
        F12.insert(F12.end(), d.F12 + 9 * (size_t)p, d.F12 + 9 * (size_t)p + 9);
        ep.insert(ep.end(), d.epipoles + 2 * (size_t)p, d.epipoles + 2 * (size_t)p + 2);
      }
    }
    asdm_job_desc ld = d;
    ld.sets = sets.data();
    ld.n_pairs = (int)ia.size();
    ld.ia = ia.data();
    ld.ib = ib.data();
    ld.F12 = F12.empty() ? d.F12 : F12.data();
    ld.epipoles = ep.empty() ? d.epipoles : ep.data();
    ld.row_begin = j->split ? rb.data() : nullptr;
    ld.row_end = j->split ? re.data() : nullptr;
    asdm_gjob::Local& X = j->loc[(size_t)l];
    X.n_local = ld.n_pairs;
    int rc = asdm_job_prepare(ctx, &ld, &X.job);
    if (rc) return bail(gfail(g, rc, asdm_last_error(ctx)));
    if (j->split && X.n_local) {
      std::lock_guard<std::recursive_mutex> lkc(ctx->mu);
      cudaSetDevice(ctx->device);
      if ((rc = ensure(ctx, X.row_map, rmap.size() * 4)) || (rc = stage_upload(ctx, X.row_map.p, rmap.data(), rmap.size() * 4)))
        return bail(gfail(g, rc, asdm_last_error(ctx)));
    }
    if (j->g_bytes) {
      std::lock_guard<std::recursive_mutex> lkc(ctx->mu);  // the window allocator is guarded by its member's context mutex
      size_t off = 0;
      if (!M.alloc.alloc(j->g_bytes, &off)) return bail(gfail(g, ASDM_ERR_OOM, "group window exhausted (gathered records)"));
      X.has_gather = true;
      if (l == 0) j->g_off = off;
      else if (off != j->g_off) return bail(gfail(g, ASDM_ERR_INTERNAL, "window allocators of the local members diverged"));
    }
  }
  *out = j;
  return ASDM_OK;
}

extern "C" int asdm_group_job_units(const asdm_gjob* j) { return j ? j->n_units : ASDM_ERR_INVALID; }
extern "C" int asdm_group_job_stride(const asdm_gjob* j) { return j ? j->stride : ASDM_ERR_INVALID; }

static bool device_visible(const void* p) {
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return at.type == cudaMemoryTypeHost || at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}

extern "C" int asdm_group_job_run(asdm_gjob* j, int32_t* h_match, float* h_dist, int32_t* h_unit_nm) {
  if (!j) return ASDM_ERR_INVALID;
  asdm_group* g = j->g;
  std::lock_guard<std::recursive_mutex> lk(g->mu);
  const int L = (int)g->m.size(), W = g->world;
  const bool host = h_match || h_dist || h_unit_nm;
  if (host && (!h_match || !h_unit_nm)) return gfail(g, ASDM_ERR_INVALID, "host results need at least the match rows and the unit counts");
  if (host && (!device_visible(h_match) || !device_visible(h_unit_nm) || (h_dist && !device_visible(h_dist))))
    return gfail(g, ASDM_ERR_INVALID, "host result arrays must be page-locked (asdm_group_host_alloc / asdm_host_alloc)");
  j->gen = (j->gen + 1) % GJOB_GENS;
  const bool gather = (j->gather & ASDM_GATHER_DEVICE) != 0;
  if (gather && !j->ran) {
    // The gathered records live in a range of every window that other jobs may have used before: nobody stores into a
    // peer's range until that peer has prepared this job too (and so is done with whatever was there).
    for (int l = 0; l < L; l++) {
      GroupMember& M = g->m[(size_t)l];
      std::lock_guard<std::recursive_mutex> lkc(M.ctx->mu);
      GCK(cudaSetDevice(M.ctx->device));
      GCK(cudaEventRecord(M.ev_main, M.ctx->stream));
      GCK(cudaStreamWaitEvent(M.bar_stream, M.ev_main, 0));
      const int rc = group_signal(g, M, CH_MAIN);
      if (rc) return rc;
    }
    for (int l = 0; l < L; l++) {
      GroupMember& M = g->m[(size_t)l];
      std::lock_guard<std::recursive_mutex> lkc(M.ctx->mu);
      const int rc = group_wait(g, M, CH_MAIN);
      if (rc) return rc;
      GCK(cudaEventRecord(M.ev_sync, M.bar_stream));
      GCK(cudaStreamWaitEvent(M.ctx->stream, M.ev_sync, 0));
    }
  }
  for (int l = 0; l < L; l++) {
    GroupMember& M = g->m[(size_t)l];
    asdm_gjob::Local& X = j->loc[(size_t)l];
    asdm_job* job = X.job;
    std::lock_guard<std::recursive_mutex> lkc(M.ctx->mu);
    PublishArgs& pub = job->pub;
    pub = PublishArgs();
    pub.row0 = M.rank;
    pub.row_step = W;
    pub.row_map = j->split ? (const int32_t*)X.row_map.p : nullptr;
    pub.col_range = j->split ? 1 : 0;
    pub.stride = j->stride;
    int n = 0;
    unsigned dist_mask = 0;  // which destinations take the distance rows
    if (host) {
      if (h_dist) dist_mask |= 1u << n;
      pub.match[n] = h_match;
      pub.dist[n] = h_dist ? h_dist : (float*)h_match;
      pub.nm[n] = h_unit_nm;
      n++;
    }
    if (j->gather & ASDM_GATHER_DEVICE) {
      for (int r = 0; r < W; r++) {
        char* b = M.peer[r] + j->g_off + (size_t)j->gen * j->gen_bytes;
        // ASDM_GATHER_NO_DIST: peers receive what the reference's calls return (index lists + counts); the distances of
        // a unit stay with the rank that computed it (its own window still gets them)
        if (!(j->gather & ASDM_GATHER_NO_DIST) || r == M.rank) dist_mask |= 1u << n;
        pub.match[n] = (int32_t*)b;
        pub.dist[n] = (float*)(b + j->o_dist);
        pub.nm[n] = (int32_t*)(b + j->o_nm);
        n++;
      }
    }
    if (n > ASDM_MAX_PEERS) return gfail(g, ASDM_ERR_INVALID, "too many publish destinations (world + host > 16)");
    pub.n = n;
    pub.dist_mask = dist_mask;
    pub.no_dist = dist_mask ? 0 : 1;
    // This run (s) fills generation `gen`, which run s-3 filled too: a peer may still be reading those rows (device code it
    // queued behind asdm_group_job_flush) until it has started ITS run s-2 -- which it has once the barrier behind run s-2
    // has completed here. With three generations a run therefore waits for the barrier of the run BEFORE the previous
    // one: the exchange of run s-1 (signal, peers' arrival, wait: ~12 us of dependent launches plus the ranks' skew) is
    // still in flight while run s computes, instead of sitting between every two runs.
    const int g_need = (j->gen + 1) % GJOB_GENS;  // generation of run s-2
    if (gather && X.bar_valid[g_need]) {
      GCK(cudaSetDevice(M.ctx->device));
      GCK(cudaStreamWaitEvent(M.ctx->stream, X.ev_bar[g_need], 0));
    }
    int rc = job_run(job);
    if (rc) return gfail(g, rc, asdm_last_error(M.ctx));
    if (gather) {  // the barrier follows the kernels on its own stream: the next run does not queue behind it
      GCK(cudaEventRecord(M.ev_main, M.ctx->stream));
      GCK(cudaStreamWaitEvent(M.bar_stream, M.ev_main, 0));
      if ((rc = group_signal(g, M, CH_MAIN))) return rc;
    }
  }
  if (gather)  // every rank's rows have landed in every window once this has run
    for (int l = 0; l < L; l++) {
      GroupMember& M = g->m[(size_t)l];
      asdm_gjob::Local& X = j->loc[(size_t)l];
      std::lock_guard<std::recursive_mutex> lkc(M.ctx->mu);
      const int rc = group_wait(g, M, CH_MAIN);
      if (rc) return rc;
      if (!X.ev_bar[j->gen]) GCK(cudaEventCreateWithFlags(&X.ev_bar[j->gen], cudaEventDisableTiming));
      GCK(cudaEventRecord(X.ev_bar[j->gen], M.bar_stream));
      X.bar_valid[j->gen] = true;
    }
  j->ran = true;
  return ASDM_OK;
}

// Orders whatever is queued next on the members' streams behind the exchange of the last run (stream-side, no host wait):
// device code that reads asdm_group_job_gathered() goes behind this.
extern "C" int asdm_group_job_flush(asdm_gjob* j) {
  if (!j) return ASDM_ERR_INVALID;
  asdm_group* g = j->g;
  std::lock_guard<std::recursive_mutex> lk(g->mu);
  for (size_t l = 0; l < g->m.size(); l++) {
    GroupMember& M = g->m[l];
    asdm_gjob::Local& X = j->loc[l];
    if (!(j->gather & ASDM_GATHER_DEVICE) || !X.bar_valid[j->gen]) continue;
    std::lock_guard<std::recursive_mutex> lkc(M.ctx->mu);
    GCK(cudaSetDevice(M.ctx->device));
    GCK(cudaStreamWaitEvent(M.ctx->stream, X.ev_bar[j->gen], 0));
  }
  return ASDM_OK;
}

extern "C" int asdm_group_job_wait(asdm_gjob* j) {
  if (!j) return ASDM_ERR_INVALID;
  asdm_group* g = j->g;
  std::lock_guard<std::recursive_mutex> lk(g->mu);
  if (!j->ran) return gfail(g, ASDM_ERR_INVALID, "asdm_group_job_wait before asdm_group_job_run");
  for (size_t l = 0; l < g->m.size(); l++) {
    GroupMember& M = g->m[l];
    std::lock_guard<std::recursive_mutex> lkc(M.ctx->mu);
    int rc = job_settle(j->loc[l].job);
    if (rc) return gfail(g, rc, asdm_last_error(M.ctx));
    GCK(cudaSetDevice(M.ctx->device));
    if (j->loc[l].n_local == 0) GCK(cudaStreamSynchronize(M.ctx->stream));  // nothing ran here, but set-up work may be queued
    if (j->gather & ASDM_GATHER_DEVICE) GCK(cudaStreamSynchronize(M.bar_stream));  // the exchange of the last run
    if (*M.h_err) {
      *M.h_err = 0;
      return gfail(g, ASDM_ERR_INTERNAL, "device barrier timed out: a rank of the group never arrived");
    }
  }
  return host_ctl_barrier(g);
}

extern "C" int asdm_group_job_counts(const asdm_gjob* j, const int32_t* h_unit_nm, int32_t* nmatches) {
  if (!j || !h_unit_nm || !nmatches) return ASDM_ERR_INVALID;
  for (int p = 0; p < j->n_pairs; p++) nmatches[p] = 0;
  for (int u = 0; u < j->n_units; u++) nmatches[j->unit_pair[(size_t)u]] += h_unit_nm[u];
  return ASDM_OK;
}

extern "C" int asdm_group_job_fetch(asdm_gjob* j, int32_t* match12, float* dist12, int32_t* nmatches) {
  if (!j) return ASDM_ERR_INVALID;
  asdm_group* g = j->g;
  std::lock_guard<std::recursive_mutex> lk(g->mu);
  const size_t rows = (size_t)std::max(j->n_pairs, 1), n = rows * (size_t)j->stride;
  if (!j->st_match) {
    void* p = nullptr;
    int rc = asdm_group_host_alloc(g, align_up(n * 4, 256) * 2 + align_up((size_t)std::max(j->n_units, 1) * 4, 256), &p);
    if (rc) return rc;
    j->st_match = (int32_t*)p;
    j->st_dist = (float*)((char*)p + align_up(n * 4, 256));
    j->st_nm = (int32_t*)((char*)p + 2 * align_up(n * 4, 256));
  }
  int rc = asdm_group_job_run(j, j->st_match, j->st_dist, j->st_nm);
  if (!rc) rc = asdm_group_job_wait(j);
  if (rc) return rc;
  const size_t used = (size_t)j->n_pairs * (size_t)j->stride;
  if (match12) memcpy(match12, j->st_match, used * 4);
  if (dist12) memcpy(dist12, j->st_dist, used * 4);
  if (nmatches) asdm_group_job_counts(j, j->st_nm, nmatches);
  return host_ctl_barrier(g);  // nobody re-runs into the staging block while a slower process still copies out of it
}

extern "C" int asdm_group_job_gathered(asdm_gjob* j, int l, void** match, void** dist, void** unit_nm) {
  if (!j || l < 0 || l >= (int)j->g->m.size() || !(j->gather & ASDM_GATHER_DEVICE)) return ASDM_ERR_INVALID;
  char* b = j->g->m[(size_t)l].win + j->g_off + (size_t)j->gen * j->gen_bytes;
  if (match) *match = b;
  if (dist) *dist = b + j->o_dist;
  if (unit_nm) *unit_nm = b + j->o_nm;
  return ASDM_OK;
}

extern "C" int asdm_group_job_destroy(asdm_gjob* j) {
  if (!j) return ASDM_ERR_INVALID;
  asdm_group* g = j->g;
  std::lock_guard<std::recursive_mutex> lk(g->mu);
  for (size_t l = 0; l < j->loc.size(); l++) {
    GroupMember& M = g->m[l];
    asdm_gjob::Local& X = j->loc[l];
    {
      std::lock_guard<std::recursive_mutex> lkc(M.ctx->mu);
      cudaSetDevice(M.ctx->device);
      release(M.ctx, X.row_map);
    }
    if (X.job) asdm_job_destroy(X.job);
    for (cudaEvent_t& e : X.ev_bar)
      if (e) {
        cudaSetDevice(M.ctx->device);
        cudaStreamSynchronize(M.bar_stream);
        cudaEventDestroy(e);
        e = nullptr;
      }
    if (X.has_gather) {
      std::lock_guard<std::recursive_mutex> lkc(M.ctx->mu);
      window_release(M, j->g_off, j->g_bytes);
    }
  }
  if (j->st_match) asdm_group_host_free(g, j->st_match);
  delete j;
  return ASDM_OK;
}
