// Host-generated constant tables.  Every value is produced with the same
// operations, in the same order, as the reference produces it on the host, and
// is handed to the kernels as data (never recomputed with device intrinsics):
//   finger_spacing_           finger_hand.cpp:7-13   (Eigen LinSpaced: low + i*step)
//   deepen depth list         finger_hand.cpp:118-122 (repeated double addition)
//   orientation angles        hand_search.cpp:179-180
//   cos/sin of the angles     hand_search.cpp:357
//   cos(friction cone)        antipodal.cpp:11,23
//   (float)(r*r)              PCL KdTreeFLANN::radiusSearch
// Compiled with -ffp-contract=off.
#include <math.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "ag2_internal.h"

namespace ag2 {

namespace {
// One bucket table of the FP32 filter (see ag2_internal.h): `edges` ascending (duplicates allowed), value(k) = the
// answer for a coordinate with exactly k edges below it.
template <class F>
void build_filter_lut(const double* edges, int ne, int nb, double v0, double width, double eps, uint32_t* out, F value) {
  const double delta = 0.01 * width + 2.0 * eps;  // float rounding of the bucket index + the filter's own error
  for (int b = 0; b < nb; b++) {
    const double lo = b == 0 ? -INFINITY : v0 + (double)b * width;
    const double hi = b == nb - 1 ? INFINITY : v0 + (double)(b + 1) * width;
    int first = -1, last = -1;
    for (int k = 0; k < ne; k++)
      if (edges[k] >= lo - delta && edges[k] <= hi + delta) {
        if (first < 0) first = k;
        last = k;
      }
    if (first < 0) {
      const double mid = b == 0 ? hi - 0.5 * width : lo + 0.5 * width;
      int k = 0;
      while (k < ne && edges[k] < mid) k++;
      out[b] = value(k);
    } else if (edges[last] - edges[first] <= eps) {
      // one threshold, or thresholds the FP32 coordinate cannot tell apart (the slot table holds 0 and 5.2e-18):
      // certain below lo[first] / above hi[last], exact in between
      out[b] = 0x80000000u | ((uint32_t)last << 8) | (uint32_t)first;
    } else {
      out[b] = 0x80000000u | 0xFFFFu;
    }
  }
}
inline uint32_t f2u(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  return u;
}
}  // namespace

int build_hand_const(ag2_ctx* c) {
  const ag2_params& p = c->p;
  HandConst& h = c->hc;
  memset(&h, 0, sizeof(h));
  const double fw = p.finger_width, od = p.hand_outer_diameter;
  double step = ((od - fw) - 0.0) / 9.0;
  for (int i = 0; i < 10; i++) {
    double half = 0.0 + (double)i * step;
    h.fs_lo[i] = (half - od) + fw;
    h.fs_lo[10 + i] = half;
  }
  for (int i = 0; i < 20; i++) h.fs_hi[i] = h.fs_lo[i] + fw;
  int nd = 0;
  h.depth[nd++] = p.init_bite;
  for (double depth = p.init_bite + 0.005; depth <= p.hand_depth; depth += 0.005) {
    if (nd >= AG2_MAX_DEEPEN_STEPS + 1) return AG2_E_INVALID;
    h.depth[nd++] = depth;
  }
  h.num_depths = nd;
  for (int j = 0; j < nd; j++) h.bottom[j] = h.depth[j] - p.hand_depth;
  const int R = p.num_orientations;
  double low = -1.0 * M_PI / 2.0, high = M_PI / 2.0;
  double astep = (high - low) / (double)R;
  for (int i = 0; i < R; i++) {
    double a = low + (double)i * astep;
    h.cos_a[i] = cos(a);
    h.sin_a[i] = sin(a);
  }
  h.num_orientations = R;
  h.finger_width = fw;
  h.hand_depth = p.hand_depth;
  h.hand_height = p.hand_height;
  h.neg_hand_height = -1.0 * p.hand_height;
  h.cos_friction = cos(30.0 * M_PI / 180.0);
  for (int a = 0; a < 3; a++) {
    h.cam_origin[0][a] = p.cam_tf_left[4 * a + 3];
    h.cam_origin[1][a] = p.cam_tf_right[4 * a + 3];
  }
  h.x_est_lo = h.fs_lo[0];
  h.x_est_inv_step = step > 0 ? 1.0 / step : 0.0;
  auto rr = [](double r, float& r2, float& rq) {
    r2 = (float)(r * r);
    rq = (float)r * 1.0001f + 1e-6f;
  };
  rr(p.nn_radius_hands, h.r2_hands, h.rq_hands);
  rr(p.nn_radius_taubin, h.r2_taubin, h.rq_taubin);
  rr(p.normals_radius, h.r2_normals, h.rq_normals);
  h.rng_seed = p.rng_seed;
  // slot-edge lookup tables (hands.cu): exact by construction, no arithmetic on the edges
  SlotLut& L = h.lut;
  std::vector<double> u(h.fs_lo, h.fs_lo + 20);
  u.insert(u.end(), h.fs_hi, h.fs_hi + 20);
  std::sort(u.begin(), u.end());
  for (int k = 0; k < AG2_SLOT_EDGES; k++) L.U[k] = u[k];  // duplicates are harmless: their interval is empty
  for (int k = 0; k <= AG2_SLOT_EDGES; k++) {
    uint32_t mask = 0;
    if (k > 0 && k < AG2_SLOT_EDGES)
      for (int i = 0; i < 20; i++)
        if (h.fs_lo[i] <= L.U[k - 1] && h.fs_hi[i] >= L.U[k]) mask |= 1u << i;
    L.M[k] = mask;
  }
  for (int k = 0; k < AG2_SLOT_EDGES; k++) {
    uint32_t mask = 0;
    for (int i = 0; i < 20; i++)
      if (h.fs_lo[i] < L.U[k] && L.U[k] < h.fs_hi[i]) mask |= 1u << i;
    L.Meq[k] = mask;
  }
  const double span = L.U[AG2_SLOT_EDGES - 1] - L.U[0];
  L.x0 = L.U[0];
  L.xscale = span > 0 ? (double)AG2_SLOT_BUCKETS / span : 0.0;
  const double width = span / (double)AG2_SLOT_BUCKETS;
  for (int b = 0; b < AG2_SLOT_BUCKETS; b++) {
    // edges strictly below the bucket's lower edge, with 2 % of a bucket as slack for the
    // rounding of the bucket index on the device; the device scans forward from here
    const double edge = L.x0 + ((double)b - 0.02) * width;
    int k = 0;
    while (k < AG2_SLOT_EDGES && L.U[k] < edge) k++;
    L.kstart[b] = (uint8_t)k;
  }
  // ---- FP32 filter tables.  A hand-frame coordinate is a . p with |a| = 1 and p a float difference shorter than
  // nn_radius_hands; evaluated with three float operations on float-rounded coefficients it is within
  // 4.2e-7 * radius of the FP64 value the reference computes (coefficient rounding 3 x 2^-24 x radius, three
  // roundings of partial sums below sqrt(3) x radius).  eps32 is ten times that.
  const double eps = 4e-6 * std::max(p.nn_radius_hands, 0.01);
  h.eps32 = (float)eps;
  std::vector<uint32_t> lut((size_t)kLutWords, 0u);
  {
    const double y_lo = h.depth[0] - 0.005, y_hi = h.depth[nd - 1] + 0.005;
    const double wy = (y_hi - y_lo) / (double)kYLutBuckets;
    h.ylut_v0 = (float)y_lo, h.ylut_scale = (float)(1.0 / wy);
    // the device computes the bucket with the float-rounded origin and scale; delta above covers the difference
    build_filter_lut(h.depth, nd, kYLutBuckets, (double)h.ylut_v0, 1.0 / (double)h.ylut_scale, eps, lut.data() + kLutY,
                     [](int k) { return (uint32_t)k; });
    const double wx = span > 0 ? span * 1.02 / (double)kXLutBuckets : 1.0;
    const double x_lo = L.U[0] - 0.01 * span;
    h.xlut_v0 = (float)x_lo, h.xlut_scale = (float)(1.0 / wx);
    build_filter_lut(L.U, AG2_SLOT_EDGES, kXLutBuckets, (double)h.xlut_v0, 1.0 / (double)h.xlut_scale, eps,
                     lut.data() + kLutX, [&](int k) { return L.M[k]; });
    for (int k = 0; k < 64; k++) {
      const double t = k < nd ? h.depth[k] : INFINITY;
      lut[kLutYLo + k] = f2u(nextafterf((float)(t - eps), -INFINITY));
      lut[kLutYHi + k] = f2u(nextafterf((float)(t + eps), INFINITY));
    }
    for (int k = 0; k < AG2_SLOT_EDGES; k++) {
      lut[kLutXLo + k] = f2u(nextafterf((float)(L.U[k] - eps), -INFINITY));
      lut[kLutXHi + k] = f2u(nextafterf((float)(L.U[k] + eps), INFINITY));
    }
  }
  if (cudaSetDevice(c->device) != cudaSuccess) return AG2_E_NO_DEVICE;
  if (c->hands_lut.reserve(kLutWords) != cudaSuccess) return AG2_E_NOMEM;
  // synchronous copy from a stack-lifetime vector, after whatever still reads the old tables; parameters change rarely
  if (c->stream) cudaStreamSynchronize(c->stream);
  if (cudaMemcpy(c->hands_lut.p, lut.data(), (size_t)kLutWords * 4, cudaMemcpyHostToDevice) != cudaSuccess)
    return AG2_E_CUDA;
  return AG2_OK;
}

}  // namespace ag2
