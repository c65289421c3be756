// f2: grasp clustering, replacing HandleSearch::findClusters (reference handle_search.cpp:4-80,
// called with remove_inliers = false at grasp_detector.cpp:229-236 and
// importance_sampling.cpp:105-109).  All-pairs test over H hands:
//   inlier(i, j) = |axis_i . axis_j| > cos(15 deg)  &&  |bottom_i - bottom_j| <= 0.05
//                  && |(I - axis_i axis_i^T)(bottom_i - bottom_j)| <= 0.005
// A hand with at least min_inliers inliers is kept, moved by (mean inlier bottom - own bottom)
// and scored with the mean inlier score.  One warp per hand i, lanes stride over j; the three
// predicates are evaluated in double with the reference's operation order (bit-exact decisions),
// the position / score sums are lane-partial sums combined in a fixed order (the reference adds
// sequentially over j; difference at rounding level, tolerance 1e-9 in the tests).
#include <math.h>

#include <algorithm>

#include "dev_common.cuh"

namespace ag2 {
namespace {

constexpr int kT = 256;

struct ClusterConst {
  double cos_thresh;   // cos(15 deg), host value
  double max_dist;     // 0.05
  double max_proj;     // 0.005
  int min_inliers;
};

__global__ void k_cluster(const ag2_hypothesis* __restrict__ hands, uint32_t H, ClusterConst cc,
                          ag2_hypothesis* __restrict__ moved /* [H] */, uint32_t* __restrict__ flag /* [H] */) {
  const uint32_t i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (i >= H) return;
  const ag2_hypothesis& hi = hands[i];
  const double a0 = hi.axis[0], a1 = hi.axis[1], a2 = hi.axis[2];
  const double b0 = hi.bottom[0], b1 = hi.bottom[1], b2 = hi.bottom[2];
  // axis_orth_proj = Identity - axis * axis^T (handle_search.cpp:26,49)
  const double P00 = 1.0 - a0 * a0, P01 = 0.0 - a0 * a1, P02 = 0.0 - a0 * a2;
  const double P10 = 0.0 - a1 * a0, P11 = 1.0 - a1 * a1, P12 = 0.0 - a1 * a2;
  const double P20 = 0.0 - a2 * a0, P21 = 0.0 - a2 * a1, P22 = 1.0 - a2 * a2;
  uint32_t cnt = 0;
  double s0 = 0.0, s1 = 0.0, s2 = 0.0, ss = 0.0;
  for (uint32_t j = lane; j < H; j += 32) {
    if (j == i) continue;
    const ag2_hypothesis& hj = hands[j];
    const double al = a0 * hj.axis[0] + (a1 * hj.axis[1] + a2 * hj.axis[2]);
    if (!(fabs(al) > cc.cos_thresh)) continue;
    const double d0 = b0 - hj.bottom[0], d1 = b1 - hj.bottom[1], d2 = b2 - hj.bottom[2];
    const double mag = sqrt(d0 * d0 + (d1 * d1 + d2 * d2));
    if (!(mag <= cc.max_dist)) continue;
    const double p0 = (P00 * d0 + P01 * d1) + P02 * d2;
    const double p1 = (P10 * d0 + P11 * d1) + P12 * d2;
    const double p2 = (P20 * d0 + P21 * d1) + P22 * d2;
    const double pm = sqrt(p0 * p0 + (p1 * p1 + p2 * p2));
    if (!(pm <= cc.max_proj)) continue;
    cnt++;
    s0 += hj.bottom[0], s1 += hj.bottom[1], s2 += hj.bottom[2];
    ss += hj.score;
  }
  for (int o = 16; o > 0; o >>= 1) {
    cnt += __shfl_xor_sync(AG2_FULL, cnt, o);
    s0 += __shfl_xor_sync(AG2_FULL, s0, o);
    s1 += __shfl_xor_sync(AG2_FULL, s1, o);
    s2 += __shfl_xor_sync(AG2_FULL, s2, o);
    ss += __shfl_xor_sync(AG2_FULL, ss, o);
  }
  if (lane != 0) return;
  const bool keep = (int)cnt >= cc.min_inliers;
  flag[i] = keep ? 1u : 0u;
  if (!keep) return;
  ag2_hypothesis g = hi;
  const double n = (double)cnt;
  const double dl[3] = {s0 / n - b0, s1 / n - b1, s2 / n - b2};
  for (int a = 0; a < 3; a++) {
    g.surface[a] = hi.surface[a] + dl[a];
    g.bottom[a] = hi.bottom[a] + dl[a];
    g.top[a] = hi.top[a] + dl[a];
  }
  g.score = ss / n;
  moved[i] = g;
}

__global__ void k_cluster_compact(const ag2_hypothesis* __restrict__ moved, const uint32_t* __restrict__ scan,
                                  uint32_t H, ag2_hypothesis* __restrict__ out) {
  uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t rec = t / 12, part = t % 12;
  if (rec >= H) return;
  uint32_t a = scan[rec], b = scan[rec + 1];
  if (a == b) return;
  reinterpret_cast<uint4*>(&out[a])[part] = reinterpret_cast<const uint4*>(&moved[rec])[part];
}

}  // namespace

// Clusters `n` contiguous device records; result: list_rec[AG2_LIST_CLUSTERS] (source order).
int clusters_compute(ag2_ctx* c, const ag2_hypothesis* d_src, size_t n) {
  cudaStream_t st = c->stream;
  c->list_n[AG2_LIST_CLUSTERS] = 0;
  AG2_CUDA(c, c->list_rec[AG2_LIST_CLUSTERS].reserve(std::max<size_t>(n, 1)));
  if (n == 0) return AG2_OK;
  AG2_CUDA(c, c->cluster_tmp.reserve(n));
  AG2_CUDA(c, c->flags.reserve(std::max<size_t>(n, c->num_samples * (size_t)c->p.num_orientations) + 1));
  ClusterConst cc;
  cc.cos_thresh = cos(15.0 * M_PI / 180.0);
  cc.max_dist = 0.05;
  cc.max_proj = 0.005;
  cc.min_inliers = c->p.min_inliers;
  k_cluster<<<(unsigned)((n * 32 + kT - 1) / kT), kT, 0, st>>>(d_src, (uint32_t)n, cc, c->cluster_tmp.p, c->flags.p);
  AG2_LAUNCH_CHECK(c);
  AG2_TRY(exclusive_scan_u32(c, c->flags.p, n));
  k_cluster_compact<<<(unsigned)((n * 12 + kT - 1) / kT), kT, 0, st>>>(c->cluster_tmp.p, c->flags.p, (uint32_t)n,
                                                                      c->list_rec[AG2_LIST_CLUSTERS].p);
  AG2_LAUNCH_CHECK(c);
  uint32_t total = 0;
  AG2_CUDA(c, cudaMemcpyAsync(&total, c->flags.p + n, 4, cudaMemcpyDeviceToHost, st));
  AG2_CUDA(c, cudaStreamSynchronize(st));
  c->list_n[AG2_LIST_CLUSTERS] = total;
  return AG2_OK;
}

}  // namespace ag2

using namespace ag2;

extern "C" int ag2_find_clusters(ag2_ctx* c, int which_src) {
  if (!c || which_src < 0 || which_src > 4) return AG2_E_INVALID;
  if (cudaSetDevice(c->device) != cudaSuccess) return AG2_E_NO_DEVICE;
  AG2_TRY(gather_list_records(c, which_src));
  return clusters_compute(c, c->list_rec[which_src].p, c->list_n[which_src]);
}

extern "C" int ag2_set_min_inliers(ag2_ctx* c, int min_inliers) {
  if (!c || min_inliers < 0) return AG2_E_INVALID;
  c->p.min_inliers = min_inliers;
  return AG2_OK;
}
