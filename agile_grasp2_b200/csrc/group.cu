// Multi-GPU host of the path, behind the C ABI (include/ag2.h, ag2_group_*).
//
// The reference is a single process with OpenMP loops over samples (hand_search.cpp:118,194,
// learning.cpp:12); its caller is GraspDetectionNode::detectGraspPosesInFile / InTopic
// (grasp_detection_node.cpp:98-143).  The path shards by grasp sample: every rank holds
// the whole cloud and its cell index, computes 1/world of the normals, evaluates the strided
// shard rank, rank + world, ... of the sample list, and all ranks end with the merged top-k.
//
// One rank = one GPU = one ag2_ctx.  Ranks are host threads of one process (ag2_group_create)
// or one process each (ag2_group_create_rank, e.g. under torchrun).  NCCL bootstraps the group
// (exchange of the window handles) and is the fallback transport; the data path proper runs on
// PEER MEMORY over NVLink, with no host synchronisation between the ranks:
//
//   window   one allocation per rank, mapped by every other rank (cudaIpc handles between
//            processes, direct pointers + cudaDeviceEnablePeerAccess inside a process):
//            flags | result blocks | input cloud (rank 0's is read by all) | sorted normals;
//   flags    flag[kind][source rank] = epoch of the last post; k_post stores the epoch into the
//            flag windows of all ranks with st.release.sys, k_wait spins with ld.acquire.sys
//            (bounded by a timeout: a dead peer turns into an error code, not a hung GPU);
//   inputs   rank 0 stages cloud + samples in its window and posts IN; every rank PULLS them
//            with loads over NVLink (k_copy2) straight into its own ctx buffers;
//   normals  the ctx's sorted normal array IS the window: k_eigen (normals.cu) stores each finished
//            normal of the rank's granules into the arrays of all ranks -- the compute kernel
//            does the all-gather itself, 16 B x N x (world-1)/world per rank over NVLink -- after
//            the GRID flags say that every rank has finished writing its own array's camera
//            words; then NRM flags;
//   results  k_pack_result stores the rank's selected records into block [rank] of every rank's
//            result window (double-buffered by epoch parity), then RES flags; one D2H, host merge
//            with the single-GPU ordering rule (grasp_detector.cpp:238-252).
//
// AG2_GROUP_NCCL=1 (or a failed window mapping) selects the plain-NCCL transport instead:
// ncclBroadcast of the inputs, ncclAllGather of packed normal shards + unpack, ncclAllGather of
// the result blocks.  It is the baseline the peer-store path is measured against.
#include <nccl.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "dev_common.cuh"

extern "C" int ag2_detect(ag2_ctx* c);
extern "C" int ag2_draw_importance_samples(ag2_ctx* c, const ag2_is_params* p, int round, const ag2_hypothesis* hands,
                                           size_t n_hands, double* xyz_out);
extern "C" int ag2_set_sample_points(ag2_ctx* c, const double* xyz, size_t n);
extern "C" int ag2_set_sample_id_map(ag2_ctx* c, int64_t first, int64_t stride);

namespace ag2 {
namespace {

constexpr int kMaxRanks = 8;
enum { F_IN = 0, F_GRID = 1, F_NRM = 2, F_RES = 3, F_IN2 = 5 };
// flag window (uint32 words): [kind * 8 + source rank] epochs (kinds 0..3 and 5); [32] wait timeouts seen here; [33] bad sample indices
constexpr size_t kFlagBytes = 256;
constexpr int kErrTimeout = 32, kErrSample = 33;
constexpr size_t kResHeader = 16;  // uint32 count, uint32 status of the sending rank, 8 bytes unused

struct PeerPtrs {
  unsigned char* base[kMaxRanks];
};

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// Stores `epoch` as flag (kind, rank) into the flag window of every rank.  Everything this rank wrote
// before (own or peer memory, earlier kernels of the stream) is visible to whoever acquires the flag.
__global__ void k_post(PeerPtrs w, int world, int rank, int kind, uint32_t epoch) {
  const int r = threadIdx.x;
  if (r < world) {
    __threadfence_system();
    st_release_sys(reinterpret_cast<uint32_t*>(w.base[r]) + kind * kMaxRanks + rank, epoch);
  }
}

// Spins until every rank of `mask` has posted `epoch` (or a later one) as flag `kind` in this rank's window.
__device__ __forceinline__ void wait_flags(uint32_t* flags, int world, uint32_t mask, int kind, uint32_t epoch,
                                           unsigned long long timeout_ns) {
  const int r = threadIdx.x;
  if (r < world && ((mask >> r) & 1u)) {
    const uint32_t* f = flags + kind * kMaxRanks + r;
    const unsigned long long t0 = global_ns();
    while ((int32_t)(ld_acquire_sys(f) - epoch) < 0) {
      if (global_ns() - t0 > timeout_ns) {
        atomicAdd(flags + kErrTimeout, 1u);
        break;
      }
      __nanosleep(64);
    }
  }
  __syncwarp();
  __threadfence_system();
}
__global__ void k_wait(uint32_t* flags, int world, uint32_t mask, int kind, uint32_t epoch,
                       unsigned long long timeout_ns) {
  wait_flags(flags, world, mask, kind, epoch, timeout_ns);
}
__global__ void k_post_wait(PeerPtrs w, int world, int rank, uint32_t mask, int kind, uint32_t epoch,
                            unsigned long long timeout_ns) {
  const int r = threadIdx.x;
  if (r < world) {
    __threadfence_system();
    st_release_sys(reinterpret_cast<uint32_t*>(w.base[r]) + kind * kMaxRanks + rank, epoch);
  }
  wait_flags(reinterpret_cast<uint32_t*>(w.base[rank]), world, mask, kind, epoch, timeout_ns);
}

// Two byte ranges, source -> destination, 16 bytes per load (both sources are 16-byte aligned; the few tail
// bytes go one by one).  The sources are rank 0's window: loads over NVLink on every other rank.
__global__ void __launch_bounds__(256) k_copy2(const unsigned char* __restrict__ a_src, unsigned char* __restrict__ a_dst,
                                               size_t a_bytes, const unsigned char* __restrict__ b_src,
                                               unsigned char* __restrict__ b_dst, size_t b_bytes) {
  const size_t tid = blockIdx.x * (size_t)blockDim.x + threadIdx.x, nth = (size_t)gridDim.x * blockDim.x;
  const size_t a16 = a_bytes / 16, b16 = b_bytes / 16;
  for (size_t i = tid; i < a16 + b16; i += nth) {
    if (i < a16)
      reinterpret_cast<uint4*>(a_dst)[i] = reinterpret_cast<const uint4*>(a_src)[i];
    else
      reinterpret_cast<uint4*>(b_dst)[i - a16] = reinterpret_cast<const uint4*>(b_src)[i - a16];
  }
  if (tid < 16) {
    if (a16 * 16 + tid < a_bytes) a_dst[a16 * 16 + tid] = a_src[a16 * 16 + tid];
    if (b16 * 16 + tid < b_bytes) b_dst[b16 * 16 + tid] = b_src[b16 * 16 + tid];
  }
}

// Second stage of the input distribution: byte range [0, bytes) of region `off` lives in slices of `slice` bytes,
// slice k in the window of rank k (rank 0's window holds all of them); 16 bytes per load, everything 16-byte aligned
// but the tail of the range.
__device__ __forceinline__ void gather_region(const PeerPtrs& w, int rank, size_t off, size_t bytes, size_t slice,
                                              unsigned char* __restrict__ dst, size_t tid, size_t nth) {
  const size_t n16 = bytes / 16;
  for (size_t i = tid; i < n16; i += nth) {
    const size_t k = (i * 16) / slice;
    const unsigned char* src = w.base[rank == 0 ? 0 : k] + off;
    reinterpret_cast<uint4*>(dst)[i] = reinterpret_cast<const uint4*>(src)[i];
  }
  if (tid < 16 && n16 * 16 + tid < bytes) {
    const size_t b = n16 * 16 + tid;
    dst[b] = (w.base[rank == 0 ? 0 : b / slice] + off)[b];
  }
}
__global__ void __launch_bounds__(256) k_gather_slices(PeerPtrs w, int rank, size_t off_a, size_t a_bytes, size_t a_slice,
                                                       unsigned char* __restrict__ a_dst, size_t off_b, size_t b_bytes,
                                                       size_t b_slice, unsigned char* __restrict__ b_dst) {
  const size_t tid = blockIdx.x * (size_t)blockDim.x + threadIdx.x, nth = (size_t)gridDim.x * blockDim.x;
  gather_region(w, rank, off_a, a_bytes, a_slice, a_dst, tid, nth);
  gather_region(w, rank, off_b, b_bytes, b_slice, b_dst, tid, nth);
}

// Strided shard rank, rank + world, ... of the sample list (range-checked against the cloud).
__global__ void k_take_strided(const int32_t* __restrict__ all, size_t mine, int rank, int world, size_t limit,
                               int32_t* __restrict__ out, uint32_t* __restrict__ bad) {
  const size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (t >= mine) return;
  int32_t v = all[(size_t)rank + t * (size_t)world];
  if (v < 0 || (size_t)v >= limit) {
    atomicAdd(bad, 1u);
    v = 0;
  }
  out[t] = v;
}
// This rank's `cnt` records + header into result block `off_block` of the windows of ranks [r0, r1):
// blockIdx.y = destination rank - r0.
__global__ void __launch_bounds__(256) k_pack_result(const ag2_hypothesis* __restrict__ rec, uint32_t cnt, uint32_t status,
                                                     PeerPtrs w, size_t off_block, int r0) {
  unsigned char* dst = w.base[r0 + blockIdx.y] + off_block;
  const uint4* src = reinterpret_cast<const uint4*>(rec);
  uint4* out = reinterpret_cast<uint4*>(dst + kResHeader);
  const uint32_t words = cnt * (uint32_t)(sizeof(ag2_hypothesis) / 16);
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < words; i += gridDim.x * blockDim.x) out[i] = src[i];
  if (blockIdx.x == 0 && threadIdx.x == 0) *reinterpret_cast<uint4*>(dst) = make_uint4(cnt, status, 0u, 0u);
  __threadfence_system();
}

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// What the ranks tell each other about their windows (all-gathered through NCCL at set-up).
struct WinInfo {
  cudaIpcMemHandle_t handle;
  unsigned long long ptr;
  long long pid;
  unsigned long long host;  // hash of the host name
  int device;
  int ok;
};
static_assert(sizeof(WinInfo) <= 128, "WinInfo travels in 128-byte slots");

// Host barrier of the ranks of one process (threaded groups only).
struct HostBarrier {
  std::mutex m;
  std::condition_variable cv;
  int n = 1, arrived = 0;
  unsigned long long gen = 0;
  bool aborted = false;  // a rank failed: nobody waits for it any more (reset at the start of the next call)
  bool wait() {
    std::unique_lock<std::mutex> lk(m);
    if (aborted) return false;
    const unsigned long long g = gen;
    if (++arrived == n) {
      arrived = 0;
      gen++;
      cv.notify_all();
    } else {
      cv.wait(lk, [&] { return gen != g || aborted; });
    }
    return !aborted;
  }
  void abort() {
    {
      std::lock_guard<std::mutex> lk(m);
      aborted = true;
    }
    cv.notify_all();
  }
  void reset() {
    std::lock_guard<std::mutex> lk(m);
    aborted = false;
    arrived = 0;
  }
};

struct Rank {
  int rank = 0, world = 1, device = 0;
  ag2_ctx* ctx = nullptr;
  ncclComm_t comm = nullptr;
  HostBarrier* hb = nullptr;  // ranks are threads of one process: host barriers guard the flag waits (see detect)
  bool want_peer = true, use_peer = false;
  unsigned char* base = nullptr;
  size_t win_bytes = 0;
  PeerPtrs peers{};
  bool opened[kMaxRanks] = {false, false, false, false, false, false, false, false};
  size_t cap_points = 0, cap_samples = 0, cap_sel = 0;
  size_t off_res = 0, res_stride = 0, off_xyz = 0, off_cam = 0, off_smp = 0, off_nrm = 0;
  uint32_t epoch = 0;
  DevBuf<unsigned char> xchg;  // NCCL staging: window handles; normal shards and unbounded lists on the NCCL path
  unsigned char* h_res = nullptr;  // pinned: result window + flag window after the exchange
  size_t h_res_bytes = 0;
  unsigned long long timeout_ns = 20ull * 1000 * 1000 * 1000;
  std::string err;
  // statistics of the last call (ms, host clock): filled for the bench's per-rank view
  size_t io_h2d = 0, io_d2h = 0;  // host <-> device bytes of the last call
  // segment marks of the last detect call on the ctx stream (per-rank account of a step, ag2_group_last_timing):
  // 0 start | 1 inputs pulled | 2 cell index built | 3 own moments done | 4 own normals stored to all ranks |
  // 5 all normals in, samples taken | 6 local detect done | 7 results of all ranks in
  cudaEvent_t tev[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  bool tev_set[8] = {false, false, false, false, false, false, false, false};
  double host_tail_ms = 0.0, host_total_ms = 0.0;
  void mark(int i) {
    if (!tev[i] && cudaEventCreate(&tev[i]) != cudaSuccess) {
      tev[i] = nullptr;
      return;
    }
    tev_set[i] = cudaEventRecord(tev[i], ctx->stream) == cudaSuccess;
  }

  int fail(int code, const std::string& what) {
    err = what;
    if (ctx && !ctx->err.empty() && what.find(ctx->err) == std::string::npos) err += ": " + ctx->err;
    return code;
  }
#define G_CUDA(expr)                                                                       \
  do {                                                                                     \
    cudaError_t _e = (expr);                                                               \
    if (_e != cudaSuccess) return fail(AG2_E_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e)); \
  } while (0)
#define G_NCCL(expr)                                                                       \
  do {                                                                                     \
    ncclResult_t _r = (expr);                                                              \
    if (_r != ncclSuccess) return fail(AG2_E_CUDA, std::string(#expr) + ": " + ncclGetErrorString(_r)); \
  } while (0)
#define G_TRY(expr)                                              \
  do {                                                           \
    int _rc = (expr);                                            \
    if (_rc != AG2_OK) return err.empty() ? fail(_rc, #expr) : _rc; \
  } while (0)
#define G_LAUNCH()                                                                         \
  do {                                                                                     \
    ctx->launches++;                                                                       \
    cudaError_t _e = cudaGetLastError();                                                   \
    if (_e != cudaSuccess) return fail(AG2_E_CUDA, std::string("kernel launch: ") + cudaGetErrorString(_e)); \
  } while (0)

  uint32_t all_mask() const { return (1u << world) - 1u; }
  uint32_t* flags() const { return reinterpret_cast<uint32_t*>(base); }
  size_t res_block(uint32_t ep, int r) const { return off_res + ((size_t)(ep & 1u) * world + (size_t)r) * res_stride; }

  void close_windows() {
    for (int r = 0; r < kMaxRanks; r++) {
      if (opened[r]) cudaIpcCloseMemHandle(peers.base[r]);
      opened[r] = false;
      peers.base[r] = nullptr;
    }
    if (ctx) ctx->nrm.release();  // adopted storage: nothing is freed
    if (base) cudaFree(base);
    base = nullptr;
    win_bytes = 0;
  }

  // all-gather of one 128-byte slot per rank through NCCL (set-up only)
  int gather_slots(const void* mine, std::vector<unsigned char>& all) {
    cudaStream_t st = ctx->stream;
    G_CUDA(xchg.reserve((size_t)128 * (world + 1)));
    G_CUDA(cudaMemcpyAsync(xchg.p + 128 * world, mine, 128, cudaMemcpyHostToDevice, st));
    G_NCCL(ncclAllGather(xchg.p + 128 * world, xchg.p, 128, ncclChar, comm, st));
    all.resize((size_t)128 * world);
    G_CUDA(cudaMemcpyAsync(all.data(), xchg.p, all.size(), cudaMemcpyDeviceToHost, st));
    G_CUDA(cudaStreamSynchronize(st));
    return AG2_OK;
  }

  // (Re)allocates the window for n points / s samples / sel records per rank and maps the windows of all ranks.
  // Collective: every rank takes the same decision from the same (n, s, sel).
  int ensure_windows(size_t n, size_t s, size_t sel) {
    sel = std::max<size_t>(sel, 1);
    if (base && n <= cap_points && s <= cap_samples && sel <= cap_sel) return AG2_OK;
    cudaStream_t st = ctx->stream;
    G_CUDA(cudaStreamSynchronize(st));
    if (hb) hb->wait();  // nobody of this process still uses a window that is about to go
    close_windows();
    cap_points = std::max(cap_points, n + n / 8 + 1024);
    cap_samples = std::max(cap_samples, s + s / 8 + 1024);
    cap_sel = std::max(cap_sel, sel);
    res_stride = align_up(kResHeader + cap_sel * sizeof(ag2_hypothesis), 256);
    size_t o = kFlagBytes;
    off_res = o, o += 2 * (size_t)world * res_stride;
    off_xyz = o = align_up(o, 256), o += 12 * cap_points;
    off_cam = o = align_up(o, 256), o += cap_points;
    off_smp = o = align_up(o, 256), o += 4 * cap_samples;
    off_nrm = o = align_up(o, 256), o += 16 * cap_points;
    win_bytes = align_up(o, 256);
    G_CUDA(cudaMalloc((void**)&base, win_bytes));
    G_CUDA(cudaMemsetAsync(base, 0, off_xyz, st));  // flags and result headers
    h_res_bytes = 0;
    if (h_res) cudaFreeHost(h_res);
    h_res = nullptr;
    G_CUDA(cudaHostAlloc((void**)&h_res, off_xyz, cudaHostAllocDefault));
    h_res_bytes = off_xyz;
    ctx->nrm.adopt(reinterpret_cast<float4*>(base + off_nrm), cap_points);
    ctx->grid_built = false;
    ctx->have_normals = false;
    peers.base[rank] = base;
    use_peer = false;
    if (world == 1) return AG2_OK;
    // ---- tell the others, map theirs
    WinInfo mine;
    memset(&mine, 0, sizeof(mine));
    mine.ok = want_peer ? 1 : 0;
    if (want_peer && cudaIpcGetMemHandle(&mine.handle, base) != cudaSuccess) {
      cudaGetLastError();
      mine.ok = 0;
    }
    mine.ptr = (unsigned long long)(uintptr_t)base;
    mine.pid = (long long)getpid();
    char hn[256] = {0};
    gethostname(hn, sizeof(hn) - 1);
    unsigned long long hh = 1469598103934665603ull;
    for (const char* q = hn; *q; q++) hh = (hh ^ (unsigned char)*q) * 1099511628211ull;
    mine.host = hh;
    mine.device = device;
    unsigned char slot[128] = {0};
    memcpy(slot, &mine, sizeof(mine));
    std::vector<unsigned char> all;
    G_TRY(gather_slots(slot, all));
    int ok = 1;
    for (int r = 0; r < world && ok; r++) {
      WinInfo wi;
      memcpy(&wi, all.data() + 128 * (size_t)r, sizeof(wi));
      if (!wi.ok || wi.host != mine.host) ok = 0;
    }
    for (int r = 0; r < world && ok; r++) {
      if (r == rank) continue;
      WinInfo wi;
      memcpy(&wi, all.data() + 128 * (size_t)r, sizeof(wi));
      if (wi.pid == mine.pid) {  // a thread of this process: its pointer is ours once peer access is on
        int can = 0;
        if (cudaDeviceCanAccessPeer(&can, device, wi.device) != cudaSuccess || !can) {
          ok = 0;
          break;
        }
        cudaError_t e = cudaDeviceEnablePeerAccess(wi.device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) ok = 0;
        cudaGetLastError();
        peers.base[r] = reinterpret_cast<unsigned char*>((uintptr_t)wi.ptr);
      } else {
        void* p = nullptr;
        if (cudaIpcOpenMemHandle(&p, wi.handle, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
          cudaGetLastError();
          ok = 0;
          break;
        }
        peers.base[r] = static_cast<unsigned char*>(p);
        opened[r] = true;
      }
    }
    // everybody maps everybody, or nobody uses the windows
    unsigned char vote[128] = {0};
    vote[0] = (unsigned char)ok;
    G_TRY(gather_slots(vote, all));
    for (int r = 0; r < world; r++) ok = ok && all[128 * (size_t)r];
    use_peer = ok != 0;
    if (!use_peer) {
      for (int r = 0; r < kMaxRanks; r++) {
        if (opened[r]) cudaIpcCloseMemHandle(peers.base[r]);
        opened[r] = false;
        if (r != rank) peers.base[r] = nullptr;
      }
    }
    if (hb) hb->wait();
    return AG2_OK;
  }

  int post(int kind) {
    k_post<<<1, 32, 0, ctx->stream>>>(peers, world, rank, kind, epoch);
    G_LAUNCH();
    return AG2_OK;
  }
  int wait(int kind, uint32_t mask) {
    // Threads of one process: the posts this wait needs are already executed (host barrier), so the kernel
    // below never spins while another rank's host sits in a device-wide call such as cudaMalloc or cudaFree
    // that may want this device idle.
    if (hb) {
      G_CUDA(cudaStreamSynchronize(ctx->stream));
      if (!hb->wait()) return fail(AG2_E_STATE, "another rank of the group failed");
    }
    k_wait<<<1, 32, 0, ctx->stream>>>(flags(), world, mask, kind, epoch, timeout_ns);
    G_LAUNCH();
    return AG2_OK;
  }
  int post_wait(int kind) {
    if (hb) {
      G_TRY(post(kind));
      return wait(kind, all_mask());
    }
    k_post_wait<<<1, 32, 0, ctx->stream>>>(peers, world, rank, all_mask(), kind, epoch, timeout_ns);
    G_LAUNCH();
    return AG2_OK;
  }

  // ---- inputs: rank 0's cloud and samples -> every rank's ctx; returns with the grid built and the normals complete
  int load_cloud(const float* xyz, size_t n, const uint8_t* cam, int num_cams, const int32_t* smp, size_t s,
                 int on_device) {
    ag2_ctx* c = ctx;
    cudaStream_t st = c->stream;
    const cudaMemcpyKind kind = on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    if (rank == 0) {
      if (!xyz || !smp) return fail(AG2_E_INVALID, "rank 0 needs the cloud and the samples");
      G_CUDA(cudaMemcpyAsync(base + off_xyz, xyz, 12 * n, kind, st));
      if (cam)
        G_CUDA(cudaMemcpyAsync(base + off_cam, cam, n, kind, st));
      else
        G_CUDA(cudaMemsetAsync(base + off_cam, 1, n, st));
      G_CUDA(cudaMemcpyAsync(base + off_smp, smp, 4 * s, kind, st));
      if (!on_device) io_h2d += 12 * n + (cam ? n : 0) + 4 * s;
    }
    const unsigned char* src = base;  // where the inputs are read from
    // Three ranks and more: seven ranks pulling the whole cloud from rank 0 at once are bounded by ONE GPU's NVLink
    // egress.  Two stages instead: rank r copies slice r of cloud and mask from rank 0's window into its own, then
    // every rank collects the slices from the windows of their holders -- every link carries 1/world of the data.
    static const bool force_two_stage = getenv("AG2_GROUP_TWO_STAGE") && atoi(getenv("AG2_GROUP_TWO_STAGE")) != 0;  // test hook
    const bool two_stage = (world > 2 || (force_two_stage && world > 1)) && use_peer;
    size_t xyz_slice = 0, cam_slice = 0;
    if (world > 1 && use_peer) {
      if (rank == 0) G_TRY(post(F_IN));
      G_TRY(wait(F_IN, 1u));
      src = peers.base[0];
      if (two_stage) {
        xyz_slice = align_up((12 * n + world - 1) / world, 16);
        cam_slice = align_up((n + world - 1) / world, 16);
        if (rank != 0) {
          const size_t xb = std::min(12 * n, (size_t)rank * xyz_slice), xe = std::min(12 * n, (size_t)(rank + 1) * xyz_slice);
          const size_t cb = std::min(n, (size_t)rank * cam_slice), ce = std::min(n, (size_t)(rank + 1) * cam_slice);
          k_copy2<<<c->sm_count, 256, 0, st>>>(src + off_xyz + xb, base + off_xyz + xb, xe - xb, src + off_cam + cb,
                                               base + off_cam + cb, ce - cb);
          G_LAUNCH();
        }
        G_TRY(post(F_IN2));
        if (hb || rank != 0) G_TRY(wait(F_IN2, all_mask()));  // rank 0 reads its own window only
      }
    } else if (world > 1) {
      G_NCCL(ncclGroupStart());
      G_NCCL(ncclBroadcast(base + off_xyz, base + off_xyz, 12 * n, ncclChar, 0, comm, st));
      G_NCCL(ncclBroadcast(base + off_cam, base + off_cam, n, ncclChar, 0, comm, st));
      G_NCCL(ncclBroadcast(base + off_smp, base + off_smp, 4 * s, ncclChar, 0, comm, st));
      G_NCCL(ncclGroupEnd());
    }
    G_TRY(cloud_reset(c, n, num_cams));
    c->normals_given = false;
    G_CUDA(c->xyz_in.reserve(3 * n));
    G_CUDA(c->cam_in.reserve(n));
    if (two_stage)
      k_gather_slices<<<c->sm_count * 4, 256, 0, st>>>(peers, rank, off_xyz, 12 * n, xyz_slice,
                                                       reinterpret_cast<unsigned char*>(c->xyz_in.p), off_cam, n, cam_slice,
                                                       c->cam_in.p);
    else
      k_copy2<<<c->sm_count * 4, 256, 0, st>>>(src + off_xyz, reinterpret_cast<unsigned char*>(c->xyz_in.p), 12 * n,
                                               src + off_cam, c->cam_in.p, n);
    G_LAUNCH();
    mark(1);
    G_TRY(grid_build(c));
    mark(2);
    if (world == 1) {
      G_TRY(normals_compute(c));
      mark(5);
      return AG2_OK;
    }
    if (use_peer) {
      // every rank's k_finalize has written the camera words of its own normal array once all GRID flags are in
      G_TRY(post(F_GRID));
      c->nrm_rank = rank, c->nrm_world = world;
      int rc = normals_moments(c);
      mark(3);
      if (rc == AG2_OK) rc = wait(F_GRID, all_mask());
      if (rc == AG2_OK) {
        c->nrm_peers = world;
        for (int r = 0; r < world; r++) c->nrm_peer[r] = reinterpret_cast<float4*>(peers.base[r] + off_nrm);
        rc = normals_finish(c);  // k_eigen: NVLink stores into every rank's array
      }
      c->nrm_rank = 0, c->nrm_world = 1, c->nrm_peers = 0;
      G_TRY(rc);
      mark(4);
      G_TRY(post_wait(F_NRM));
      mark(5);
      return AG2_OK;
    }
    // NCCL transport: packed shard -> all-gather -> unpack
    const size_t shard_cap = normals_shard_points((uint32_t)c->n_finite, 0, world);
    G_CUDA(xchg.reserve(16 * shard_cap * (size_t)(world + 1)));
    float* tx = reinterpret_cast<float*>(xchg.p + 16 * shard_cap * (size_t)world);
    size_t owned = 0;
    G_TRY(normals_shard_compute(c, rank, world, tx, shard_cap, &owned));
    G_NCCL(ncclAllGather(tx, xchg.p, 16 * shard_cap, ncclChar, comm, st));
    G_TRY(normals_shard_assemble(c, reinterpret_cast<const float*>(xchg.p), shard_cap, world));
    mark(5);
    return AG2_OK;
  }

  // ---- this rank's strided shard of the sample list (in rank 0's window, or in the own one after a broadcast)
  int take_samples(size_t s) {
    ag2_ctx* c = ctx;
    const size_t mine = s > (size_t)rank ? (s - rank + world - 1) / world : 0;
    const unsigned char* src = (world > 1 && use_peer) ? peers.base[0] : base;
    if (mine) {
      G_CUDA(c->sample_idx.reserve(mine));
      k_take_strided<<<(unsigned)((mine + 255) / 256), 256, 0, c->stream>>>(
          reinterpret_cast<const int32_t*>(src + off_smp), mine, rank, world, c->n, c->sample_idx.p, flags() + kErrSample);
      G_LAUNCH();
    }
    c->h_sample_idx.clear();
    c->h_sample_idx_stale = mine > 0;
    c->use_points = false;
    c->num_samples = mine;
    c->st.num_samples = (int64_t)mine;
    c->sample_id_offset = rank, c->sample_id_stride = world;
    invalidate_results(c);
    return AG2_OK;
  }

  // ---- per-rank lists -> the single-GPU list on every rank
  int exchange_results(int local_rc, std::vector<ag2_hypothesis>& out) {
    ag2_ctx* c = ctx;
    cudaStream_t st = c->stream;
    out.clear();
    const bool topk = c->p.num_selected > 0 && c->p.antipodal_mode != AG2_MODE_NONE;
    const int which = (world == 1 || topk) ? AG2_LIST_SELECTED : AG2_LIST_ANTIPODAL;
    size_t cnt = local_rc == AG2_OK ? c->list_n[which] : 0;
    if (local_rc == AG2_OK && cnt) {
      int rc = gather_list_records(c, which);
      if (rc != AG2_OK) local_rc = rc, cnt = 0;
    }
    if (world == 1) {
      if (local_rc != AG2_OK) return local_rc;
      out.resize(cnt);
      if (cnt)
        G_CUDA(cudaMemcpyAsync(out.data(), c->list_rec[which].p, cnt * sizeof(ag2_hypothesis), cudaMemcpyDeviceToHost, st));
      io_d2h += cnt * sizeof(ag2_hypothesis);
      G_CUDA(cudaMemcpyAsync(h_res, base, kFlagBytes, cudaMemcpyDeviceToHost, st));
      G_TRY(sync_and_finish(c));
      return check_error_words();
    }
    if (!topk) return exchange_unbounded(local_rc, which, cnt, out);
    if (cnt > cap_sel) return fail(AG2_E_CAPACITY, "selection larger than the result window");
    const ag2_hypothesis* rec = cnt ? c->list_rec[which].p : nullptr;
    const uint32_t status = (uint32_t)(-local_rc);
    const unsigned gx = (unsigned)std::max<size_t>(1, (cnt * 12 + 255) / 256);
    if (use_peer) {
      k_pack_result<<<dim3(gx, world), 256, 0, st>>>(rec, (uint32_t)cnt, status, peers, res_block(epoch, rank), 0);
      G_LAUNCH();
      G_TRY(post_wait(F_RES));
    } else {
      k_pack_result<<<dim3(gx, 1), 256, 0, st>>>(rec, (uint32_t)cnt, status, peers, res_block(epoch, rank), rank);
      G_LAUNCH();
      G_NCCL(ncclAllGather(base + res_block(epoch, rank), base + res_block(epoch, 0), res_stride, ncclChar, comm, st));
    }
    mark(7);
    // flag window (error words) and this epoch's half of the result blocks
    G_CUDA(cudaMemcpyAsync(h_res, base, kFlagBytes, cudaMemcpyDeviceToHost, st));
    G_CUDA(cudaMemcpyAsync(h_res + res_block(epoch, 0), base + res_block(epoch, 0), (size_t)world * res_stride,
                           cudaMemcpyDeviceToHost, st));
    io_d2h += kFlagBytes + (size_t)world * res_stride;
    int rc = sync_and_finish(c);
    if (local_rc != AG2_OK) return local_rc;
    G_TRY(rc);
    G_TRY(check_error_words());
    const ag2_hypothesis* ptrs[kMaxRanks];
    size_t counts[kMaxRanks];
    for (int r = 0; r < world; r++) {
      const unsigned char* blk = h_res + res_block(epoch, r);
      uint32_t hdr[4];
      memcpy(hdr, blk, 16);
      if (hdr[1]) return fail(-(int)hdr[1], "rank " + std::to_string(r) + " of the group failed");
      if (hdr[0] > cap_sel) return fail(AG2_E_CUDA, "corrupt result block");
      counts[r] = hdr[0];
      ptrs[r] = reinterpret_cast<const ag2_hypothesis*>(blk + kResHeader);
    }
    merge_rank_lists(ptrs, counts, world, c->p.num_selected, false, out);
    return AG2_OK;
  }

  // error words of the flag window, as copied to h_res by the caller; a reported error is cleared
  int check_error_words() {
    const uint32_t* hf = reinterpret_cast<const uint32_t*>(h_res);
    const bool timeout = hf[kErrTimeout] != 0, sample = hf[kErrSample] != 0;
    if (timeout || sample) cudaMemsetAsync(flags() + kErrTimeout, 0, 8, ctx->stream);
    if (timeout) return fail(AG2_E_CUDA, "a rank of the group did not answer in time");
    if (sample) return fail(AG2_E_INVALID, "sample index outside the cloud");
    return AG2_OK;
  }

  // lists without a bound (num_selected <= 0, antipodal_mode 0): sizes first, then padded records, over NCCL
  int exchange_unbounded(int local_rc, int which, size_t cnt, std::vector<ag2_hypothesis>& out) {
    ag2_ctx* c = ctx;
    cudaStream_t st = c->stream;
    unsigned char slot[128] = {0};
    uint32_t hdr[2] = {(uint32_t)cnt, (uint32_t)(-local_rc)};
    memcpy(slot, hdr, 8);
    std::vector<unsigned char> all;
    G_TRY(gather_slots(slot, all));
    size_t counts[kMaxRanks], mx = 1;
    for (int r = 0; r < world; r++) {
      uint32_t h2[2];
      memcpy(h2, all.data() + 128 * (size_t)r, 8);
      if (h2[1]) return fail(-(int)h2[1], "rank " + std::to_string(r) + " of the group failed");
      counts[r] = h2[0];
      mx = std::max(mx, counts[r]);
    }
    const size_t blk = mx * sizeof(ag2_hypothesis);
    G_CUDA(xchg.reserve(blk * (size_t)world));
    if (cnt)
      G_CUDA(cudaMemcpyAsync(xchg.p + blk * rank, c->list_rec[which].p, cnt * sizeof(ag2_hypothesis),
                             cudaMemcpyDeviceToDevice, st));
    G_NCCL(ncclAllGather(xchg.p + blk * rank, xchg.p, blk, ncclChar, comm, st));
    std::vector<ag2_hypothesis> host((size_t)world * mx);
    G_CUDA(cudaMemcpyAsync(host.data(), xchg.p, blk * (size_t)world, cudaMemcpyDeviceToHost, st));
    G_CUDA(cudaMemcpyAsync(h_res, base, kFlagBytes, cudaMemcpyDeviceToHost, st));
    G_TRY(sync_and_finish(c));
    G_TRY(check_error_words());
    const ag2_hypothesis* ptrs[kMaxRanks];
    for (int r = 0; r < world; r++) ptrs[r] = host.data() + (size_t)r * mx;
    merge_rank_lists(ptrs, counts, world, c->p.num_selected, c->p.antipodal_mode == AG2_MODE_NONE, out);
    return AG2_OK;
  }

  // Keeps the other ranks from waiting for flags this rank will never post after a failure.
  void release_peers(int rc) {
    if (hb) hb->abort();
    if (world == 1 || !use_peer || !base) return;
    cudaStream_t st = ctx->stream;
    k_pack_result<<<dim3(1, world), 256, 0, st>>>(nullptr, 0u, (uint32_t)(-rc), peers, res_block(epoch, rank), 0);
    for (int kind = F_IN; kind <= F_RES; kind++)
      if (kind != F_IN || rank == 0) k_post<<<1, 32, 0, st>>>(peers, world, rank, kind, epoch);
    cudaStreamSynchronize(st);
    cudaGetLastError();
  }

  int check_args(size_t n, size_t s, int num_cams) {
    if (cudaSetDevice(device) != cudaSuccess) return fail(AG2_E_NO_DEVICE, "cudaSetDevice");
    err.clear();
    ctx->err.clear();
    io_h2d = io_d2h = 0;
    if (n == 0) return fail(AG2_E_EMPTY_CLOUD, "empty cloud");
    if (s == 0) return fail(AG2_E_NO_SAMPLES, "no samples");
    if (n >= 0x7fffffffu || num_cams < 1 || num_cams > 8) return fail(AG2_E_INVALID, "bad sizes");
    return AG2_OK;
  }

  // GraspDetector::detectGraspPoses (grasp_detector.cpp:84-282) over all ranks.
  int detect(const float* xyz, size_t n, const uint8_t* cam, int num_cams, const int32_t* smp, size_t s, int on_device,
             std::vector<ag2_hypothesis>& out) {
    G_TRY(check_args(n, s, num_cams));
    G_TRY(ensure_windows(n, s, ctx->p.num_selected > 0 ? (size_t)ctx->p.num_selected : 1));
    epoch++;
    const auto t_begin = std::chrono::steady_clock::now();
    for (int i = 0; i < 8; i++) tev_set[i] = false;
    mark(0);
    int rc = load_cloud(xyz, n, cam, num_cams, smp, s, on_device);
    if (rc == AG2_OK) rc = take_samples(s);
    if (rc != AG2_OK) {
      release_peers(rc);
      return rc;
    }
    if (ctx->num_samples) rc = ag2_detect(ctx);
    if (rc == AG2_E_EMPTY_CLOUD || rc == AG2_E_NO_SAMPLES) rc = AG2_OK;  // an empty shard contributes nothing
    if (rc != AG2_OK && hb) hb->abort();
    mark(6);
    const auto t_detect = std::chrono::steady_clock::now();
    rc = exchange_results(rc, out);
    const auto t_end = std::chrono::steady_clock::now();
    host_tail_ms = std::chrono::duration<double, std::milli>(t_end - t_detect).count();
    host_total_ms = std::chrono::duration<double, std::milli>(t_end - t_begin).count();
    return rc;
  }

  // ImportanceSampling::detectGraspPoses (importance_sampling.cpp:30-112) over all ranks: the cloud, its index and
  // its normals stay resident; every round's samples are sharded like a plain call; all ranks hold the same
  // hands found so far and draw the same next batch from them (counter-based stream, deviation D5).
  int detect_importance(const float* xyz, size_t n, const uint8_t* cam, int num_cams, const int32_t* smp, size_t s,
                        int on_device, const ag2_is_params* ip, std::vector<ag2_hypothesis>& out) {
    ag2_ctx* c = ctx;
    struct NoCluster {  // the rounds run detectGraspPoses(cloud_cam, false): no clustering inside (:38, :90)
      ag2_ctx* c;
      int saved;
      explicit NoCluster(ag2_ctx* cc) : c(cc), saved(cc->p.min_inliers) { c->p.min_inliers = 0; }
      ~NoCluster() { c->p.min_inliers = saved; }
    } guard(c);
    G_TRY(detect(xyz, n, cam, num_cams, smp, s, on_device, out));
    if (out.empty()) return AG2_OK;  // :40-43
    const int ns = ip->num_samples;
    int32_t offset = (int32_t)s;
    std::vector<double> pts((size_t)3 * ns);
    std::vector<ag2_hypothesis> sel;
    for (int it = 0; it < ip->num_iterations; it++) {
      epoch++;
      int rc = ag2_draw_importance_samples(c, ip, it, out.data(), out.size(), pts.data());
      const size_t mine = (size_t)ns > (size_t)rank ? ((size_t)ns - rank + world - 1) / world : 0;
      if (rc == AG2_OK && mine) {
        std::vector<double> shard(3 * mine);
        for (size_t t = 0; t < mine; t++)
          for (int a = 0; a < 3; a++) shard[3 * t + a] = pts[3 * ((size_t)rank + t * world) + a];
        rc = ag2_set_sample_points(c, shard.data(), mine);
        if (rc == AG2_OK) rc = ag2_set_sample_id_map(c, rank, world);
        if (rc == AG2_OK) rc = ag2_detect(c);
      } else if (rc == AG2_OK) {
        invalidate_results(c);
      }
      G_TRY(exchange_results(rc, sel));
      for (auto& h : sel) h.sample_id += offset;
      offset += ns;
      out.insert(out.end(), sel.begin(), sel.end());
    }
    // :105-109 clusters over the hands of all rounds
    if (guard.saved > 0 && !out.empty()) {
      c->p.min_inliers = guard.saved;
      DevBuf<ag2_hypothesis> d;
      G_CUDA(d.reserve(out.size()));
      G_CUDA(cudaMemcpyAsync(d.p, out.data(), out.size() * sizeof(ag2_hypothesis), cudaMemcpyHostToDevice, c->stream));
      G_TRY(clusters_compute(c, d.p, out.size()));
      const size_t k = c->list_n[AG2_LIST_CLUSTERS];
      out.resize(k);
      if (k)
        G_CUDA(cudaMemcpyAsync(out.data(), c->list_rec[AG2_LIST_CLUSTERS].p, k * sizeof(ag2_hypothesis),
                               cudaMemcpyDeviceToHost, c->stream));
      G_TRY(sync_and_finish(c));
    }
    return AG2_OK;
  }

  void destroy() {
    if (ctx) {
      cudaSetDevice(device);
      cudaStreamSynchronize(ctx->stream);
    }
    if (hb) hb->wait();  // no rank of this process frees a window another one may still touch
    close_windows();
    if (h_res) cudaFreeHost(h_res);
    h_res = nullptr;
    for (int i = 0; i < 8; i++)
      if (tev[i]) cudaEventDestroy(tev[i]), tev[i] = nullptr;
    xchg.release();
    if (comm) ncclCommDestroy(comm);
    comm = nullptr;
    if (ctx) ag2_destroy(ctx);
    ctx = nullptr;
  }
};

// One host thread per rank of a single-process group: runs the jobs the caller's thread hands out.
struct Worker {
  std::thread th;
  std::mutex m;
  std::condition_variable cv;
  std::function<int()> job;
  bool has_job = false, done = false, quit = false;
  int rc = 0;
  void loop() {
    for (;;) {
      std::function<int()> j;
      {
        std::unique_lock<std::mutex> lk(m);
        cv.wait(lk, [&] { return has_job || quit; });
        if (quit && !has_job) return;
        j = job;
      }
      const int r = j();
      {
        std::lock_guard<std::mutex> lk(m);
        rc = r;
        has_job = false;
        done = true;
      }
      cv.notify_all();
    }
  }
  void start() { th = std::thread([this] { loop(); }); }
  void submit(std::function<int()> j) {
    {
      std::lock_guard<std::mutex> lk(m);
      job = std::move(j);
      has_job = true;
      done = false;
    }
    cv.notify_all();
  }
  int join_job() {
    std::unique_lock<std::mutex> lk(m);
    cv.wait(lk, [&] { return done; });
    return rc;
  }
  void stop() {
    {
      std::lock_guard<std::mutex> lk(m);
      quit = true;
    }
    cv.notify_all();
    if (th.joinable()) th.join();
  }
};

}  // namespace
}  // namespace ag2

using namespace ag2;

struct ag2_group {
  std::vector<Rank*> ranks;       // one (a process per rank) or all of them (threads of this process)
  std::vector<Worker*> workers;   // threaded groups only
  HostBarrier hb;
  std::string err;
  std::vector<std::vector<ag2_hypothesis>> results;

  // runs f(local index) on every local rank (in its thread when there are several); first error wins
  int run(const std::function<int(int)>& f) {
    if (workers.empty()) return f(0);
    hb.reset();
    for (size_t i = 0; i < workers.size(); i++) workers[i]->submit([&f, i] { return f((int)i); });
    // the first error wins, but a rank that only gave up because another one failed does not hide the cause
    int rc = AG2_OK;
    bool secondary = true;
    for (size_t i = 0; i < workers.size(); i++) {
      const int r = workers[i]->join_job();
      const bool sec = r == AG2_E_STATE && ranks[i]->err.find("another rank") != std::string::npos;
      if (r != AG2_OK && (rc == AG2_OK || (secondary && !sec))) {
        rc = r;
        secondary = sec;
        err = "rank " + std::to_string(i) + ": " + ranks[i]->err;
      }
    }
    return rc;
  }
};

namespace {
void read_env(Rank* r) {
  if (const char* e = getenv("AG2_GROUP_NCCL")) r->want_peer = atoi(e) == 0;
  if (const char* e = getenv("AG2_GROUP_TIMEOUT_MS")) {
    const long v = atol(e);
    if (v > 0) r->timeout_ns = (unsigned long long)v * 1000000ull;
  }
}

int copy_out(const std::vector<ag2_hypothesis>& res, ag2_hypothesis* out, size_t cap, size_t* n_out) {
  if (n_out) *n_out = res.size();
  if (res.size() > cap) return AG2_E_CAPACITY;
  if (!res.empty()) {
    if (!out) return AG2_E_INVALID;
    memcpy(out, res.data(), res.size() * sizeof(ag2_hypothesis));
  }
  return AG2_OK;
}
}  // namespace

extern "C" {

int ag2_group_unique_id(ag2_group_id* out) {
  static_assert(sizeof(ncclUniqueId) <= sizeof(ag2_group_id), "ag2_group_id holds an ncclUniqueId");
  if (!out) return AG2_E_INVALID;
  memset(out, 0, sizeof(*out));
  ncclUniqueId id;
  if (ncclGetUniqueId(&id) != ncclSuccess) return AG2_E_CUDA;
  memcpy(out, &id, sizeof(id));
  return AG2_OK;
}

int ag2_group_create_rank(const ag2_params* p, const ag2_group_id* id, int rank, int world, int device, ag2_group** out) {
  if (!out || !p || world < 1 || world > kMaxRanks || rank < 0 || rank >= world || (world > 1 && !id)) return AG2_E_INVALID;
  *out = nullptr;
  ag2_ctx* ctx = nullptr;
  int rc = ag2_create(p, device, &ctx);
  if (rc != AG2_OK) return rc;
  ag2_group* g = new (std::nothrow) ag2_group();
  Rank* r = new (std::nothrow) Rank();
  if (!g || !r) {
    delete g;
    delete r;
    ag2_destroy(ctx);
    return AG2_E_NOMEM;
  }
  r->rank = rank, r->world = world, r->device = device, r->ctx = ctx;
  read_env(r);
  g->ranks.push_back(r);
  if (world > 1) {
    ncclUniqueId nid;
    memcpy(&nid, id, sizeof(nid));
    if (ncclCommInitRank(&r->comm, world, nid, rank) != ncclSuccess) {
      r->comm = nullptr;
      ag2_group_destroy(g);
      return AG2_E_CUDA;
    }
  }
  *out = g;
  return AG2_OK;
}

int ag2_group_create(const ag2_params* p, int n_gpus, ag2_group** out) {
  if (!out || !p || n_gpus < 1 || n_gpus > kMaxRanks) return AG2_E_INVALID;
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < n_gpus) {
    cudaGetLastError();
    return AG2_E_NO_DEVICE;
  }
  if (n_gpus == 1) return ag2_group_create_rank(p, nullptr, 0, 1, 0, out);
  ag2_group* g = new (std::nothrow) ag2_group();
  if (!g) return AG2_E_NOMEM;
  g->hb.n = n_gpus;
  std::vector<ncclComm_t> comms((size_t)n_gpus, nullptr);
  std::vector<int> devs((size_t)n_gpus);
  for (int i = 0; i < n_gpus; i++) devs[i] = i;
  int rc = AG2_OK;
  for (int i = 0; i < n_gpus && rc == AG2_OK; i++) {
    ag2_ctx* ctx = nullptr;
    rc = ag2_create(p, i, &ctx);
    if (rc != AG2_OK) break;
    Rank* r = new (std::nothrow) Rank();
    if (!r) {
      ag2_destroy(ctx);
      rc = AG2_E_NOMEM;
      break;
    }
    r->rank = i, r->world = n_gpus, r->device = i, r->ctx = ctx, r->hb = &g->hb;
    read_env(r);
    g->ranks.push_back(r);
  }
  if (rc == AG2_OK && ncclCommInitAll(comms.data(), n_gpus, devs.data()) != ncclSuccess) rc = AG2_E_CUDA;
  if (rc != AG2_OK) {
    g->hb.n = 1;
    for (Rank* r : g->ranks) {
      r->hb = nullptr;
      r->destroy();
      delete r;
    }
    delete g;
    return rc;
  }
  for (int i = 0; i < n_gpus; i++) {
    g->ranks[i]->comm = comms[i];
    Worker* w = new Worker();
    g->workers.push_back(w);
    w->start();
  }
  *out = g;
  return AG2_OK;
}

int ag2_group_destroy(ag2_group* g) {
  if (!g) return AG2_OK;
  g->run([g](int i) {
    g->ranks[i]->destroy();
    return AG2_OK;
  });
  for (Worker* w : g->workers) {
    w->stop();
    delete w;
  }
  for (Rank* r : g->ranks) delete r;
  delete g;
  return AG2_OK;
}

int ag2_group_size(const ag2_group* g, int* world, int* n_local) {
  if (!g || g->ranks.empty()) return AG2_E_INVALID;
  if (world) *world = g->ranks[0]->world;
  if (n_local) *n_local = (int)g->ranks.size();
  return AG2_OK;
}

int ag2_group_local_ctx(ag2_group* g, int i, ag2_ctx** ctx) {
  if (!g || !ctx || i < 0 || (size_t)i >= g->ranks.size()) return AG2_E_INVALID;
  *ctx = g->ranks[i]->ctx;
  return AG2_OK;
}

const char* ag2_group_last_error(const ag2_group* g) {
  if (!g) return "";
  if (!g->err.empty()) return g->err.c_str();
  return g->ranks.empty() ? "" : g->ranks[0]->err.c_str();
}

int ag2_group_transport(const ag2_group* g, int* peer_stores) {
  if (!g || !peer_stores || g->ranks.empty()) return AG2_E_INVALID;
  *peer_stores = g->ranks[0]->use_peer ? 1 : 0;
  return AG2_OK;
}

int ag2_group_last_timing(const ag2_group* g, double* ms) {
  if (!g || g->ranks.empty() || !ms) return AG2_E_INVALID;
  const Rank* r = g->ranks[0];
  for (int i = 0; i < 10; i++) ms[i] = 0.0;
  int prev = 0;
  for (int i = 1; i < 8; i++) {
    if (!r->tev_set[i] || !r->tev_set[prev]) continue;
    float t = 0.f;
    if (cudaEventElapsedTime(&t, r->tev[prev], r->tev[i]) == cudaSuccess) ms[i - 1] = t;
    prev = i;
  }
  ms[7] = r->host_tail_ms;
  ms[8] = r->host_total_ms;
  float t = 0.f;
  if (r->tev_set[0] && r->tev_set[7] && cudaEventElapsedTime(&t, r->tev[0], r->tev[7]) == cudaSuccess) ms[9] = t;
  return AG2_OK;
}

int ag2_group_io_bytes(const ag2_group* g, size_t* h2d, size_t* d2h) {
  if (!g || g->ranks.empty()) return AG2_E_INVALID;
  if (h2d) *h2d = g->ranks[0]->io_h2d;
  if (d2h) *d2h = g->ranks[0]->io_d2h;
  return AG2_OK;
}

int ag2_group_detect(ag2_group* g, const float* xyz, size_t n, const uint8_t* cam_mask, int num_cams,
                     const int32_t* samples, size_t n_samples, int on_device, ag2_hypothesis* out, size_t cap,
                     size_t* n_out) {
  if (!g || g->ranks.empty()) return AG2_E_INVALID;
  g->err.clear();
  g->results.resize(g->ranks.size());
  const int rc = g->run([&](int i) {
    return g->ranks[i]->detect(xyz, n, cam_mask, num_cams, samples, n_samples, on_device, g->results[i]);
  });
  if (rc != AG2_OK) return rc;
  return copy_out(g->results[0], out, cap, n_out);
}

int ag2_group_detect_importance(ag2_group* g, const float* xyz, size_t n, const uint8_t* cam_mask, int num_cams,
                                const int32_t* samples, size_t n_samples, int on_device, const ag2_is_params* ip,
                                ag2_hypothesis* out, size_t cap, size_t* n_out) {
  if (!g || g->ranks.empty() || !ip || ip->num_iterations < 0 || ip->num_samples < 1) return AG2_E_INVALID;
  g->err.clear();
  g->results.resize(g->ranks.size());
  const int rc = g->run([&](int i) {
    return g->ranks[i]->detect_importance(xyz, n, cam_mask, num_cams, samples, n_samples, on_device, ip, g->results[i]);
  });
  if (rc != AG2_OK) return rc;
  return copy_out(g->results[0], out, cap, n_out);
}

}  // extern "C"
