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
  return AG2_OK;
}

}  // namespace ag2
