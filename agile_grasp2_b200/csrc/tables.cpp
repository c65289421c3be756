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
  return AG2_OK;
}

}  // namespace ag2
