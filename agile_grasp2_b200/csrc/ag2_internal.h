// Internal declarations shared by the translation units of libag2.so.
// Product code: CUDA (sm_100a) only, no CPU fallback, nothing from oracle/.
#ifndef AG2_INTERNAL_H_
#define AG2_INTERNAL_H_

#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/ag2.h"

namespace ag2 {

// ---------------------------------------------------------------- device buffers
template <typename T>
struct DevBuf {
  T* p = nullptr;
  size_t cap = 0;  // elements
  bool external = false;  // storage belongs to someone else (a multi-GPU exchange window, group.cu)
  DevBuf() = default;
  DevBuf(const DevBuf&) = delete;
  DevBuf& operator=(const DevBuf&) = delete;
  ~DevBuf() { release(); }
  void release() {
    if (p && !external) cudaFree(p);
    p = nullptr;
    cap = 0;
    external = false;
  }
  // Uses caller-owned storage of `n` elements instead of an allocation of its own (never freed here).
  void adopt(T* ptr, size_t n) {
    release();
    p = ptr;
    cap = n;
    external = ptr != nullptr;
  }
  // Grows (never shrinks); contents are NOT preserved.
  cudaError_t reserve(size_t n) {
    if (n <= cap) return cudaSuccess;
    if (external) return cudaErrorMemoryAllocation;  // the owner sizes external storage
    release();
    size_t want = n + n / 8 + 64;
    cudaError_t e = cudaMalloc((void**)&p, want * sizeof(T));
    if (e != cudaSuccess) {
      p = nullptr;
      return e;
    }
    cap = want;
    return cudaSuccess;
  }
};

// ---------------------------------------------------------------- uniform grid (K1-K3)
// Cell index (cx + nx*(cy + ny*cz)): x runs are contiguous in the sorted arrays,
// so a fixed-radius query is a loop over (y,z) rows, each ONE contiguous span.
struct GridDev {
  float ox, oy, oz;   // origin (bbox min of the finite points)
  float inv_h;        // 1 / cell size
  float h;            // cell size
  int nx, ny, nz;
  int n;              // finite points (sorted arrays hold exactly these)
  const uint32_t* cell_start;  // [ncells + 1]
  const float4* pts;           // sorted: x, y, z, w = bits of the input-order index
  const float4* nrm;           // sorted: nx, ny, nz, w = bits of the camera mask
};

// Host-built constant tables; values are produced on the host EXACTLY as the
// reference produces them (finger_hand.cpp:7-13,122; hand_search.cpp:179-180,357)
// and never recomputed with device intrinsics.
// Slot-edge lookup: the 40 values fs_lo[i], fs_hi[i] sorted (U); a point collides with
// slot i iff fs_lo[i] < x < fs_hi[i], which depends only on k = #(U[j] < x) and on
// whether x == U[k].  kstart maps a uniform bucket of x to a conservative start for k.
#define AG2_SLOT_EDGES 40
#define AG2_SLOT_BUCKETS 256
struct SlotLut {
  double U[AG2_SLOT_EDGES];
  uint32_t M[AG2_SLOT_EDGES + 1];   // U[k-1] < x < U[k]
  uint32_t Meq[AG2_SLOT_EDGES];     // x == U[k]
  uint8_t kstart[AG2_SLOT_BUCKETS];
  double x0, xscale;                // bucket = (int)((x - x0) * xscale), clamped
};

// Device tables of the FP32 filter (uint32 words, uploaded by build_hand_const; layout shared with hands.cu):
// a bucket entry is the answer itself (depth class / 20-slot mask) when no threshold is near the bucket, else
// 0x80000000 | last << 8 | first for the ONE threshold (or run of thresholds within eps32 of each other) near it
// (certain below lo[first] / above hi[last] = threshold -+ eps32, exact FP64 evaluation in between), or
// 0x80000000 | 0xFFFF when separate thresholds are near (always exact).
constexpr int kYLutBuckets = 256, kXLutBuckets = 1024;
constexpr int kLutY = 0, kLutX = kLutY + kYLutBuckets, kLutYLo = kLutX + kXLutBuckets, kLutYHi = kLutYLo + 64,
              kLutXLo = kLutYHi + 64, kLutXHi = kLutXLo + AG2_SLOT_EDGES, kLutWords = kLutXHi + AG2_SLOT_EDGES;

struct HandConst {
  SlotLut lut;
  double fs_lo[20];   // finger_spacing_(i)
  double fs_hi[20];   // finger_spacing_(i) + finger_width_
  double depth[AG2_MAX_DEEPEN_STEPS + 2];   // t_0 = init_bite, t_j deepen depths
  double bottom[AG2_MAX_DEEPEN_STEPS + 2];  // t_j - hand_depth
  double cos_a[AG2_MAX_ORIENTATIONS];
  double sin_a[AG2_MAX_ORIENTATIONS];
  double finger_width, hand_depth, hand_height, neg_hand_height;
  double cos_friction;        // cos(30 deg), antipodal.cpp:11,23
  double cam_origin[2][3];    // translation columns of cam_tf_left/right
  double x_est_lo, x_est_inv_step;  // slot-index estimator (left half); exact fix-up follows
  int num_depths;             // D + 1 entries in depth[]
  int num_orientations;
  // FP32 filter of the hand sweep (hands.cu): |x32 - x64| <= eps32 for hand-frame coordinates of staged points
  float eps32;
  float ylut_v0, ylut_scale;  // bucket = (int)((v - v0) * scale), clamped to the table
  float xlut_v0, xlut_scale;
  float r2_hands;             // (float)(r*r)
  float rq_hands;             // query half-width with slack
  float r2_taubin, rq_taubin;
  float r2_normals, rq_normals;
  uint64_t rng_seed;
};

// Per (sample, orientation) auxiliary output of the hand kernel.
struct BoxRef {
  uint32_t offset;  // into the in-box pool
  uint32_t count;
};

struct LeNetDev {
  bool loaded = false;
  DevBuf<float> c1w, c1b, c2w, c2b, i1w, i1b, i2w, i2b;
  DevBuf<float> c1w_tc;  // conv1 weights split into tf32 terms, two taps per K step
  DevBuf<float> i1w_tc;  // ip1 weights scaled, split into fp16 terms, per K step and 128-output tile (raw bytes)
  DevBuf<uint16_t> c2w_h;  // conv2 weights scaled, split into fp16 terms and tiled per K step (k_conv2_h)
  float a1_scale = 1.f;      // power of two applied to the pooled conv1 activations before their fp16 split
  float c2_inv_scale = 1.f;  // 1 / (a1_scale * conv2 weight scale)
  float a2_scale = 1.f;      // power of two applied to the pooled conv2 activations before their fp16 split (ip1's operand)
  float i1_inv_scale = 1.f;  // 1 / (a2_scale * ip1 weight scale)
  std::vector<float> host[8];
};

}  // namespace ag2

struct ag2_ctx {
  ag2_params p;
  int device = 0;
  cudaStream_t stream = nullptr;
  std::string err;
  int sm_count = 148;
  int64_t launches = 0;

  // input-order cloud
  size_t n = 0;
  int num_cams = 1;
  ag2::DevBuf<float> xyz_in;       // 3n packed
  ag2::DevBuf<uint8_t> cam_in;     // n
  ag2::DevBuf<float> normals_in;   // 3n when precomputed normals were given
  bool normals_given = false;
  bool have_normals = false;       // sorted normals complete for all points
  int nrm_rank = 0, nrm_world = 1; // normals_compute covers the granules of this shard only (multi-GPU)
  // multi-GPU (group.cu): k_eigen stores every finished normal of this rank's granules into the sorted
  // normals of ALL ranks through peer-mapped pointers; 0 peers = single GPU, store into c->nrm only
  float4* nrm_peer[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  int nrm_peers = 0;
  // deferred read-backs: small device scalars copied to pinned host memory by asynchronous copies and
  // folded into the statistics (and checked for capacity errors) after the next synchronisation of the
  // ctx stream (finish_pending) -- a stage that only needs a number for its statistics does not stall
  uint32_t* h_pin = nullptr;       // kPinWords words of pinned host memory
  bool pend_normals = false;       // h_pin[0..4]: max k_n, -, sum k_n (u64), points over the last tier capacity
  bool pend_select = false;        // h_pin[27..28]: antipodal / selected counts of the single-launch select path
  bool pend_frames = false;        // h_pin[5]: largest frame neighbourhood (capacity check)
  bool frames_attr_set = false;
  bool normals_attr_set = false;   // dynamic shared-memory limits of the normals kernels raised on this device
  uint32_t max_kn_hint = 0;        // densest normals neighbourhood of the previous cloud (sizes speculative launches)
  float min_z = 0.f;               // pcl::getMinMax3D z (grasp_detector.cpp:153)

  // grid
  bool grid_built = false;
  float origin[3] = {0, 0, 0};
  float h = 0.01f, inv_h = 100.f;
  int dim[3] = {1, 1, 1};
  size_t ncells = 0;
  size_t n_finite = 0;
  ag2::DevBuf<uint32_t> cell_of;        // n (input order) cell id or 0xFFFFFFFF
  ag2::DevBuf<uint32_t> cell_start;     // ncells + 1
  ag2::DevBuf<uint32_t> cell_fill;      // ncells
  ag2::DevBuf<uint32_t> tmp_idx;        // n
  ag2::DevBuf<float4> pts;              // sorted
  ag2::DevBuf<float4> nrm;              // sorted
  ag2::DevBuf<uint32_t> sorted_of_orig; // n
  ag2::DevBuf<uint32_t> scan_tmp;
  ag2::DevBuf<uint32_t> kn;             // neighbour counts (normals pass)
  bool pruned_precomputed = false, selected_precomputed = false;  // small lists: built eagerly in one launch
  size_t pruned_count = 0, selected_count = 0;
  ag2::DevBuf<uint32_t> big_list[2];    // points whose neighbourhood exceeds a launch capacity (ping-pong)
  ag2::DevBuf<uint32_t> tier_list;      // normals: points grouped by neighbourhood-size tier, [tier][owned points]
  ag2::DevBuf<uint32_t> tier_cnt;       // normals: per-tier counts + per-launch batch counters
  ag2::DevBuf<float> nrm_mom;           // normals: raw moment sums per point (12 floats), input of k_eigen
  static constexpr int kAux = 12;
  cudaStream_t aux[kAux] = {};  // independent launches (normals tiers, render size classes) run side by side
  cudaEvent_t ev_fork = nullptr, ev_join[kAux] = {};
  ag2::DevBuf<uint32_t> scalars;        // small device scratch (counters, maxima, flags)

  // samples
  bool use_points = false;
  size_t num_samples = 0;
  int64_t sample_id_offset = 0, sample_id_stride = 1;
  ag2::DevBuf<int32_t> sample_idx;
  ag2::DevBuf<double> sample_xyz;
  std::vector<int32_t> h_sample_idx;
  bool h_sample_idx_stale = false;  // indices were set from device memory; host copy fetched on demand
  std::vector<double> h_sample_xyz;

  // hot-path outputs
  ag2::DevBuf<ag2_frame> frames;            // S
  ag2::DevBuf<uint32_t> kt;                 // S neighbour counts at nn_radius_taubin
  ag2::DevBuf<uint32_t> frames_big;         // [2] sizes + two lists: samples whose frame neighbourhood exceeds 1024 / 4096
  ag2::DevBuf<ag2_sweep_record> sweeps;     // S*R
  ag2::DevBuf<ag2_hypothesis> slots;        // S*R
  ag2::DevBuf<ag2::BoxRef> box_ref;         // S*R
  ag2::DevBuf<int32_t> box_pool;            // in-box sorted positions
  ag2::DevBuf<uint32_t> flags;              // 2 * S*R (emitted / pruned)
  ag2::DevBuf<uint32_t> flag_scan;          // 2 * S*R
  ag2::DevBuf<uint32_t> list[4];            // slot ids of AG2_LIST_HYPOTHESES .. AG2_LIST_SELECTED
  size_t list_n[6] = {0, 0, 0, 0, 0, 0};
  // contiguous records per list; lists 4 (importance sampling) and 5 (clusters) exist only as records
  ag2::DevBuf<ag2_hypothesis> list_rec[6];
  bool list_rec_valid[6] = {false, false, false, false, false, false};
  // importance sampling (importance.cu)
  ag2::DevBuf<double> is_samples;           // samples drawn in the rounds, 3 doubles each
  size_t is_samples_n = 0;
  std::vector<size_t> is_round_offset;      // first sample of each round in is_samples
  ag2::DevBuf<double> is_cand;              // candidate staging
  ag2::DevBuf<int32_t> is_cand_idx;
  ag2::DevBuf<double> is_surf;              // grasp surfaces of the hands found so far, packed 3 doubles each
  ag2::DevBuf<uint32_t> is_flag;            // accept flags of a candidate batch, then their exclusive scan
  ag2::DevBuf<ag2_hypothesis> is_hands_in;  // ag2_draw_importance_samples: caller's hands, drawn samples
  ag2::DevBuf<double> is_draw;
  ag2::DevBuf<uint8_t> images;              // pruned * 10800 (HWC u8)
  ag2::DevBuf<uint8_t> render_scratch;      // sort arrays of k_render when a closing region exceeds shared memory
  ag2::DevBuf<float> logits;                // pruned * 2
  ag2::DevBuf<double> scores;               // pruned
  ag2::DevBuf<float> act;                   // LeNet activations
  ag2::DevBuf<ag2_hypothesis> cluster_tmp;  // moved hands before compaction (cluster.cu)
  ag2::DevBuf<uint64_t> sort_keys[2];       // top-k ranking scratch (select.cu)
  ag2::DevBuf<uint32_t> sort_pos[2];
  uint32_t max_in_box = 0;
  bool have_hyps = false, have_pruned = false, have_images = false, have_antipodal = false,
       have_selected = false;

  ag2::LeNetDev net;
  ag2::HandConst hc;
  ag2::DevBuf<uint32_t> hands_lut;          // kLutWords words: tables of the FP32 filter of k_hands
  ag2_stats st;
  cudaEvent_t ev[18];
  bool ev_used[18];
};

namespace ag2 {

// error plumbing -------------------------------------------------------------
int fail(ag2_ctx* c, int code, const char* what, cudaError_t e, const char* file, int line);
#define AG2_CUDA(c, expr)                                                          \
  do {                                                                             \
    cudaError_t _e = (expr);                                                       \
    if (_e != cudaSuccess) return ::ag2::fail((c), AG2_E_CUDA, #expr, _e, __FILE__, __LINE__); \
  } while (0)
#define AG2_TRY(expr)      \
  do {                     \
    int _rc = (expr);      \
    if (_rc != AG2_OK) return _rc; \
  } while (0)
#define AG2_LAUNCH_CHECK(c)                                                        \
  do {                                                                             \
    (c)->launches++;                                                               \
    cudaError_t _e = cudaGetLastError();                                           \
    if (_e != cudaSuccess) return ::ag2::fail((c), AG2_E_CUDA, "kernel launch", _e, __FILE__, __LINE__); \
  } while (0)

// stage entry points (each .cu implements its own) ----------------------------
int build_hand_const(ag2_ctx* c);                   // tables.cpp
int grid_build(ag2_ctx* c);                         // grid.cu
GridDev grid_dev(const ag2_ctx* c);
int exclusive_scan_u32(ag2_ctx* c, uint32_t* data, size_t n);  // in place, total at data[n]
int normals_compute(ag2_ctx* c);                    // normals.cu: normals_moments + normals_finish
int normals_moments(ag2_ctx* c);                    // count pass + one launch per neighbourhood-size tier (raw moment sums)
int normals_finish(ag2_ctx* c);                     // k_eigen over the owned points (stores to all peers when set)
int finish_pending(ag2_ctx* c);                     // api_core.cu: after a stream sync, folds the deferred read-backs
int sync_and_finish(ag2_ctx* c);                    // cudaStreamSynchronize(c->stream) + finish_pending
constexpr int kPinWords = 64;
uint32_t normals_shard_points(uint32_t n, int rank, int world);
int normals_shard_compute(ag2_ctx* c, int rank, int world, float* d_out, size_t cap_points, size_t* n_owned);
int normals_shard_assemble(ag2_ctx* c, const float* d_all, size_t cap_points, int world);
int frames_compute(ag2_ctx* c);                     // frames.cu
int hands_compute(ag2_ctx* c);                      // hands.cu: launch + synchronise + finish, retried when the pool grows
int hands_launch(ag2_ctx* c);                       // k_hands + read-backs to pinned memory (no synchronisation)
int hands_finish(ag2_ctx* c);                       // after a synchronisation: AG2_OK / kHandsRetry / error
constexpr int kHandsRetry = 1;
int lists_build_after_hands(ag2_ctx* c);            // select.cu: launch + synchronise + finish
int lists_launch(ag2_ctx* c);                       // hypothesis list (and, for small runs, the pruned list) enqueued
int lists_finish(ag2_ctx* c);                       // after a synchronisation: counts from pinned memory
int prune_compute(ag2_ctx* c);
int prune_launch(ag2_ctx* c);
int prune_finish(ag2_ctx* c);
int select_mode(ag2_ctx* c);
int select_top(ag2_ctx* c);
int gather_list_records(ag2_ctx* c, int which);
int clusters_compute(ag2_ctx* c, const ag2_hypothesis* d_src, size_t n);  // cluster.cu
int importance_detect(ag2_ctx* c, const ag2_is_params* ip);   // importance.cu
int render_compute(ag2_ctx* c);                     // render.cu
int lenet_upload(ag2_ctx* c);                       // lenet.cu
int lenet_forward(ag2_ctx* c, const uint8_t* d_images, size_t n, float* d_logits);
int classify_compute(ag2_ctx* c);
int radius_search_device(ag2_ctx* c, const float* q, double radius, std::vector<int32_t>& idx,
                         std::vector<float>& d2);   // grid.cu
int load_caffemodel(const char* path, std::vector<float> out[8], std::string& err);  // caffemodel.cpp
void xavier_fill(uint64_t seed, std::vector<float> out[8]);

int cloud_reset(ag2_ctx* c, size_t n, int num_cams);   // api_core.cu: a new cloud of n points (grid, normals, results dropped)
void invalidate_results(ag2_ctx* c);
// host merge of per-rank selections (api_stages.cu): rank r's records are recs[r][0..counts[r]), its samples the
// strided shard r, r + world, ...; result in single-GPU order, see ag2_merge_selected
void merge_rank_lists(const ag2_hypothesis* const* recs, const size_t* counts, int world, int num_selected, bool keep_all,
                      std::vector<ag2_hypothesis>& out);
void phase_begin(ag2_ctx* c, int phase);
void phase_end(ag2_ctx* c, int phase);
int collect_phase_times(ag2_ctx* c);

}  // namespace ag2

#endif
