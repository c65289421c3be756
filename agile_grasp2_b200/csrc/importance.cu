// a17: ImportanceSampling::detectGraspPoses (reference importance_sampling.cpp:30-112) and its
// samplers drawSamplesFromSumOfGaussians / drawSamplesFromMaxOfGaussians (:115-157).
//
// What changes against the reference, and why:
//  * the cell index and the normals are built ONCE; every round is a new sample batch on the
//    resident cloud (the reference rebuilds the kd-tree and recomputes all N normals per round,
//    hand_search.cpp:11-24);
//  * the rounds evaluate the DRAWN xyz samples (calculateLocalFrames(samples), hand_search.cpp:
//    238-317) — the evident intent; as written the reference's rounds re-evaluate the index
//    samples because use_incoming_samples_ is never set on this instance;
//  * rand() / boost::mt19937(time(NULL)) / boost::normal_distribution are replaced by a
//    counter-based stream (deviation D5): candidate c of round i draws its hand index and its
//    three N(0,1) values from splitmix64(seed, i, c, j) with Box-Muller on the exactly
//    reproducible log / cos of include/ag2_detmath.h (the same source is compiled into the
//    oracle).  Everything runs on the device: one thread generates a candidate, one warp runs its
//    O(H) max-of-Gaussians rejection test, a scan of the accept flags picks the first num_gauss
//    accepted candidates in stream order; the host only learns how many were accepted per batch.
//  * the rejection test p(idx) >= max_k p(k) with p = term * exp(c * d2), c = -1/(2 sigma), is
//    evaluated on the exponent arguments c * d2 (exp is monotone); the two differ only when exp
//    rounds two different arguments to the same value (documented tie).
#include <math.h>
#include <string.h>

#include <algorithm>
#include <chrono>
#include <vector>

#include "../../include/ag2_detmath.h"
#include "dev_common.cuh"

namespace ag2 {
namespace {

// AG2_IS_TRACE=1: host wall-clock of the segments of a multi-round run on stderr (profiling aid)
struct IsTrace {
  bool on;
  std::chrono::steady_clock::time_point t0;
  IsTrace() : on(getenv("AG2_IS_TRACE") && atoi(getenv("AG2_IS_TRACE")) != 0), t0(std::chrono::steady_clock::now()) {}
  void lap(const char* what, int round) {
    if (!on) return;
    const auto t1 = std::chrono::steady_clock::now();
    fprintf(stderr, "[ag2 importance] round %d %-10s %8.3f ms\n", round, what,
            std::chrono::duration<double, std::milli>(t1 - t0).count());
    t0 = t1;
  }
};

__host__ __device__ inline uint64_t h_splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
__host__ __device__ inline uint64_t is_hash(uint64_t seed, uint64_t round, uint64_t cand, uint64_t j) {
  uint64_t k = h_splitmix64(seed ^ 0x1234ABCD5678EF01ull);
  k = h_splitmix64(k + round);
  k = h_splitmix64(k + cand);
  return h_splitmix64(k + j);
}
__host__ __device__ inline double is_u01(uint64_t h) { return ((double)(h >> 11) + 0.5) * (1.0 / 9007199254740992.0); }
__host__ __device__ inline double is_normal(uint64_t seed, uint64_t round, uint64_t cand, int comp) {
  const double u1 = is_u01(is_hash(seed, round, cand, 2 * (uint64_t)comp));
  const double u2 = is_u01(is_hash(seed, round, cand, 2 * (uint64_t)comp + 1));
  return ag2_det_normal(u1, u2);
}

// grasp surfaces of the hands found so far, packed 3 doubles each
__global__ void k_is_surf(const ag2_hypothesis* __restrict__ hands, uint32_t H, double* __restrict__ surf) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= H) return;
  surf[3 * k] = hands[k].surface[0], surf[3 * k + 1] = hands[k].surface[1], surf[3 * k + 2] = hands[k].surface[2];
}

// One thread per candidate `first + b` of round `it`: the hand it is drawn around and the point
// surface + N(0, I) * sigma (importance_sampling.cpp:118-121,134-139).
__global__ void k_is_generate(uint64_t seed, uint64_t it, uint64_t first, uint32_t B, const double* __restrict__ surf,
                              uint32_t H, double sigma, double* __restrict__ cand, int32_t* __restrict__ cand_idx) {
  const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const uint64_t cid = first + b;
  const uint32_t idx = (uint32_t)(is_hash(seed, it, cid, 6) % (uint64_t)H);
  cand_idx[b] = (int32_t)idx;
  for (int a = 0; a < 3; a++)
    cand[3 * (size_t)b + a] = ag2_dm_add(surf[3 * (size_t)idx + a], ag2_dm_mul(is_normal(seed, it, cid, a), sigma));
}

// The accepted candidates of a batch whose rank among the accepted ones is below `need` become samples
// got, got + 1, ... (scan = exclusive prefix of the accept flags, total at scan[B]).
__global__ void k_is_take(const double* __restrict__ cand, const uint32_t* __restrict__ scan, uint32_t B, uint32_t need,
                          double* __restrict__ out) {
  const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const uint32_t r = scan[b];
  if (scan[b + 1] == r || r >= need) return;
  out[3 * (size_t)r] = cand[3 * (size_t)b], out[3 * (size_t)r + 1] = cand[3 * (size_t)b + 1];
  out[3 * (size_t)r + 2] = cand[3 * (size_t)b + 2];
}

// rand part of a round: samples.col(j) = cloud->points[r].getVector3fMap().cast<double>(), r drawn per sample
__global__ void k_is_random_points(uint64_t seed, uint64_t it, const float* __restrict__ xyz_in, uint64_t n_points,
                                   uint32_t n, double* __restrict__ out) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  const size_t r = (size_t)(is_hash(seed, it, 0xFFFFFFFFull + (uint64_t)j, 7) % n_points);
  out[3 * j] = (double)xyz_in[3 * r];
  out[3 * j + 1] = (double)xyz_in[3 * r + 1];
  out[3 * j + 2] = (double)xyz_in[3 * r + 2];
}

// One warp per candidate: accept iff c*d2(x, hand[idx]) >= max_k c*d2(x, hand[k]).
__global__ void k_is_accept(const double* __restrict__ cand, const int32_t* __restrict__ cand_idx, uint32_t n_cand,
                            const ag2_hypothesis* __restrict__ hands, const double* __restrict__ surf /* 3H, packed */,
                            uint32_t H, double coef, uint32_t* __restrict__ accept) {
  const uint32_t c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (c >= n_cand) return;
  const double x0 = cand[3 * c], x1 = cand[3 * c + 1], x2 = cand[3 * c + 2];
  // own = coef * d2 to the generating hand; the candidate is rejected as soon as ANY hand scores strictly higher
  // (the same decision as comparing with the maximum over all hands, which nearly every candidate fails early)
  double own;
  {
    const ag2_hypothesis& h = hands[cand_idx[c]];
    const double v0 = x0 - h.surface[0], v1 = x1 - h.surface[1], v2 = x2 - h.surface[2];
    own = coef * (v0 * v0 + (v1 * v1 + v2 * v2));  // Eigen fixed-size inner product order
  }
  bool ok = true;
  for (uint32_t k0 = 0; k0 < H; k0 += 32) {
    const uint32_t k = k0 + lane;
    bool higher = false;
    if (k < H) {
      const double v0 = x0 - surf[3 * k], v1 = x1 - surf[3 * k + 1], v2 = x2 - surf[3 * k + 2];
      higher = coef * (v0 * v0 + (v1 * v1 + v2 * v2)) > own;
    }
    if (__any_sync(AG2_FULL, higher)) {
      ok = false;
      break;
    }
  }
  if (lane == 0) accept[c] = ok ? 1u : 0u;
}

__global__ void k_is_append(const ag2_hypothesis* __restrict__ src, uint32_t n, int32_t sample_offset,
                            ag2_hypothesis* __restrict__ dst) {
  uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t rec = t / 12, part = t % 12;
  if (rec >= n) return;
  uint4 v = reinterpret_cast<const uint4*>(&src[rec])[part];
  if (part == 10) v.x = (uint32_t)((int32_t)v.x + sample_offset);  // byte 160: sample_id
  reinterpret_cast<uint4*>(&dst[rec])[part] = v;
}

int append_selected(ag2_ctx* c, size_t cap, int32_t sample_offset) {
  AG2_TRY(gather_list_records(c, AG2_LIST_SELECTED));
  const size_t n = c->list_n[AG2_LIST_SELECTED];
  if (c->list_n[AG2_LIST_IMPORTANCE] + n > cap) return AG2_E_CAPACITY;
  if (n) {
    k_is_append<<<(unsigned)((n * 12 + 255) / 256), 256, 0, c->stream>>>(
        c->list_rec[AG2_LIST_SELECTED].p, (uint32_t)n, sample_offset,
        c->list_rec[AG2_LIST_IMPORTANCE].p + c->list_n[AG2_LIST_IMPORTANCE]);
    AG2_LAUNCH_CHECK(c);
  }
  c->list_n[AG2_LIST_IMPORTANCE] += n;
  return AG2_OK;
}

// The sample batch of round `it` (importance_sampling.cpp:47-86): n_gauss draws close to the hands found so far
// (d_hands: H records on the device) followed by n_rand random cloud points, written to d_out (3 * num_samples
// doubles on the device).  All on the device; per batch of candidates the host reads back one number (how many
// were accepted) to decide whether another batch is needed.
int draw_round(ag2_ctx* c, const ag2_is_params* ip, int it, const ag2_hypothesis* d_hands, size_t H, double* d_out) {
  cudaStream_t st = c->stream;
  const int ns = ip->num_samples;
  const int n_rand = (int)(ip->prob_rand_samples * ns);  // int num_rand_samples = prob * num_samples
  const int n_gauss = ns - n_rand;
  const double sigma = ip->radius;
  const double coef = -1.0 / (2.0 * sigma);
  if (n_gauss > 0) {
    AG2_CUDA(c, c->is_surf.reserve(3 * H));
    k_is_surf<<<(unsigned)((H + 255) / 256), 256, 0, st>>>(d_hands, (uint32_t)H, c->is_surf.p);
    AG2_LAUNCH_CHECK(c);
  }
  // 3.1 samples close to the grasps found so far
  int got = 0;
  uint64_t next_cand = 0;
  for (int batch_no = 0; got < n_gauss; batch_no++) {
    if (batch_no >= 4096) {
      c->err = "importance sampling: rejection sampler did not terminate";
      return AG2_E_STATE;
    }
    // batch size: the accepted samples are the first n_gauss accepted candidates in id order whatever the
    // batching, so it only trades round trips against wasted candidates: 8 x the missing samples at first, then
    // missing / observed acceptance rate with a quarter to spare
    size_t B = (size_t)(n_gauss - got);
    if (ip->method != 1) {
      const double rate = next_cand ? std::max((double)got / (double)next_cand, 1e-4) : 0.125;
      B = (size_t)std::min(4.0e6, std::max(256.0, 1.25 * (double)(n_gauss - got) / rate));
    }
    AG2_CUDA(c, c->is_cand.reserve(3 * B));
    AG2_CUDA(c, c->is_cand_idx.reserve(B));
    k_is_generate<<<(unsigned)((B + 127) / 128), 128, 0, st>>>(ip->seed, (uint64_t)it, next_cand, (uint32_t)B, c->is_surf.p,
                                                              (uint32_t)H, sigma, c->is_cand.p, c->is_cand_idx.p);
    AG2_LAUNCH_CHECK(c);
    next_cand += B;
    if (ip->method == 1) {  // sum of Gaussians: every draw is a sample (importance_sampling.cpp:115-125)
      AG2_CUDA(c, cudaMemcpyAsync(d_out + 3 * (size_t)got, c->is_cand.p, 24 * B, cudaMemcpyDeviceToDevice, st));
      got += (int)B;
      break;
    }
    AG2_CUDA(c, c->is_flag.reserve(B + 1));
    k_is_accept<<<(unsigned)((B * 32 + 255) / 256), 256, 0, st>>>(c->is_cand.p, c->is_cand_idx.p, (uint32_t)B, d_hands,
                                                                  c->is_surf.p, (uint32_t)H, coef, c->is_flag.p);
    AG2_LAUNCH_CHECK(c);
    AG2_TRY(exclusive_scan_u32(c, c->is_flag.p, B));
    k_is_take<<<(unsigned)((B + 255) / 256), 256, 0, st>>>(c->is_cand.p, c->is_flag.p, (uint32_t)B, (uint32_t)(n_gauss - got),
                                                          d_out + 3 * (size_t)got);
    AG2_LAUNCH_CHECK(c);
    AG2_CUDA(c, cudaMemcpyAsync(c->h_pin + 32, c->is_flag.p + B, 4, cudaMemcpyDeviceToHost, st));
    AG2_TRY(sync_and_finish(c));
    got += (int)std::min<uint32_t>(c->h_pin[32], (uint32_t)(n_gauss - got));
  }
  // 3.2 random cloud points
  if (n_rand) {
    k_is_random_points<<<(n_rand + 255) / 256, 256, 0, st>>>(ip->seed, (uint64_t)it, c->xyz_in.p, (uint64_t)c->n,
                                                           (uint32_t)n_rand, d_out + (size_t)3 * n_gauss);
    AG2_LAUNCH_CHECK(c);
  }
  return AG2_OK;
}

}  // namespace

int importance_detect(ag2_ctx* c, const ag2_is_params* ip) {
  static_assert(offsetof(ag2_hypothesis, sample_id) == 160, "k_is_append patches sample_id at byte 160");
  cudaStream_t st = c->stream;
  const int iters = ip->num_iterations;
  const int ns = ip->num_samples;
  const int R = c->p.num_orientations;
  const size_t per_round_max = (c->p.num_selected > 0 && c->p.antipodal_mode != AG2_MODE_NONE)
                                   ? (size_t)c->p.num_selected
                                   : std::max<size_t>(c->num_samples, (size_t)ns) * (size_t)R;
  const size_t cap = per_round_max * (size_t)(iters + 1);
  AG2_CUDA(c, c->list_rec[AG2_LIST_IMPORTANCE].reserve(std::max<size_t>(cap, 1)));
  AG2_CUDA(c, c->is_samples.reserve(std::max<size_t>((size_t)3 * ns * iters, 3)));
  c->list_n[AG2_LIST_IMPORTANCE] = 0;
  c->is_samples_n = 0;
  c->is_round_offset.clear();

  // the rounds run detectGraspPoses(cloud_cam, false): no clustering inside (:38, :90)
  struct NoCluster {
    ag2_ctx* c;
    int saved;
    explicit NoCluster(ag2_ctx* cc) : c(cc), saved(cc->p.min_inliers) { c->p.min_inliers = 0; }
    ~NoCluster() { c->p.min_inliers = saved; }
  } no_cluster(c);
  // 1. initial grasp hypotheses on the samples the caller has set (importance_sampling.cpp:38)
  IsTrace trace;
  AG2_TRY(ag2_detect(c));
  if (trace.on) cudaStreamSynchronize(st);
  trace.lap("detect", -1);
  int32_t sample_offset = 0;
  AG2_TRY(append_selected(c, cap, sample_offset));
  sample_offset += (int32_t)c->num_samples;
  if (c->list_n[AG2_LIST_IMPORTANCE] == 0) return AG2_OK;  // :40-43

  for (int it = 0; it < iters; it++) {
    AG2_CUDA(c, c->sample_xyz.reserve((size_t)3 * ns));
    AG2_TRY(draw_round(c, ip, it, c->list_rec[AG2_LIST_IMPORTANCE].p, c->list_n[AG2_LIST_IMPORTANCE], c->sample_xyz.p));
    trace.lap("draw", it);
    AG2_CUDA(c, cudaMemcpyAsync(c->is_samples.p + 3 * c->is_samples_n, c->sample_xyz.p, (size_t)24 * ns,
                                cudaMemcpyDeviceToDevice, st));
    c->is_round_offset.push_back(c->is_samples_n);
    c->is_samples_n += (size_t)ns;
    c->use_points = true;
    c->num_samples = (size_t)ns;
    c->st.num_samples = (int64_t)ns;
    c->have_hyps = c->have_pruned = c->have_images = c->have_antipodal = c->have_selected = false;
    for (int i = 0; i < 4; i++) {
      c->list_n[i] = 0;
      c->list_rec_valid[i] = false;
    }
    AG2_TRY(ag2_detect(c));
    if (trace.on) cudaStreamSynchronize(st);
    trace.lap("detect", it);
    AG2_TRY(append_selected(c, cap, sample_offset));
    if (trace.on) cudaStreamSynchronize(st);
    trace.lap("append", it);
    sample_offset += ns;
  }
  // :105-109 clusters over the hands of all rounds -> AG2_LIST_CLUSTERS
  if (no_cluster.saved > 0) {
    c->p.min_inliers = no_cluster.saved;
    AG2_TRY(clusters_compute(c, c->list_rec[AG2_LIST_IMPORTANCE].p, c->list_n[AG2_LIST_IMPORTANCE]));
  }
  return AG2_OK;
}

}  // namespace ag2

using namespace ag2;

extern "C" {

int ag2_default_is_params(ag2_is_params* p) {
  if (!p) return AG2_E_INVALID;
  memset(p, 0, sizeof(*p));
  p->num_iterations = 5;       // importance_sampling.cpp:9-15
  p->num_samples = 50;
  p->num_init_samples = 100;
  p->method = 2;
  p->prob_rand_samples = 0.3;
  p->radius = 0.02;
  p->seed = 0;
  return AG2_OK;
}

int ag2_detect_importance(ag2_ctx* c, const ag2_is_params* p) {
  if (!c || !p || p->num_iterations < 0 || p->num_samples < 1 || !(p->radius > 0) || p->prob_rand_samples < 0 ||
      p->prob_rand_samples > 1 || (p->method != 1 && p->method != 2))
    return AG2_E_INVALID;
  if (cudaSetDevice(c->device) != cudaSuccess) return AG2_E_NO_DEVICE;
  if (c->n == 0) return AG2_E_EMPTY_CLOUD;
  if (c->num_samples == 0) return AG2_E_NO_SAMPLES;
  return importance_detect(c, p);
}

int ag2_draw_importance_samples(ag2_ctx* c, const ag2_is_params* p, int round, const ag2_hypothesis* hands,
                                size_t n_hands, double* xyz_out) {
  if (!c || !p || !hands || !xyz_out || n_hands == 0 || round < 0 || p->num_samples < 1 || !(p->radius > 0) ||
      p->prob_rand_samples < 0 || p->prob_rand_samples > 1 || (p->method != 1 && p->method != 2))
    return AG2_E_INVALID;
  if (cudaSetDevice(c->device) != cudaSuccess) return AG2_E_NO_DEVICE;
  if (c->n == 0) return AG2_E_EMPTY_CLOUD;
  AG2_CUDA(c, c->is_hands_in.reserve(n_hands));
  AG2_CUDA(c, c->is_draw.reserve((size_t)3 * p->num_samples));
  AG2_CUDA(c, cudaMemcpyAsync(c->is_hands_in.p, hands, n_hands * sizeof(ag2_hypothesis), cudaMemcpyHostToDevice,
                              c->stream));
  AG2_TRY(draw_round(c, p, round, c->is_hands_in.p, n_hands, c->is_draw.p));
  AG2_CUDA(c, cudaMemcpyAsync(xyz_out, c->is_draw.p, (size_t)24 * p->num_samples, cudaMemcpyDeviceToHost, c->stream));
  AG2_CUDA(c, cudaStreamSynchronize(c->stream));
  return AG2_OK;
}

int ag2_get_importance_samples(ag2_ctx* c, int round, double* xyz, size_t cap, size_t* n) {
  if (!c || !n || round < 0 || (size_t)round >= c->is_round_offset.size()) return AG2_E_INVALID;
  const size_t b = c->is_round_offset[round];
  const size_t e = (size_t)round + 1 < c->is_round_offset.size() ? c->is_round_offset[round + 1] : c->is_samples_n;
  *n = e - b;
  if (e - b > cap) return AG2_E_CAPACITY;
  if (!xyz) return AG2_E_INVALID;
  cudaSetDevice(c->device);
  AG2_CUDA(c, cudaMemcpyAsync(xyz, c->is_samples.p + 3 * b, (e - b) * 24, cudaMemcpyDeviceToHost, c->stream));
  AG2_CUDA(c, cudaStreamSynchronize(c->stream));
  return AG2_OK;
}

}  // extern "C"
