// K6: hand search, replacing HandSearch::evaluateHands + calculateHand
// (reference hand_search.cpp:173-235, :319-426), FingerHand
// (finger_hand.cpp:17-93, :96-134, :137-214, :313-325) and
// Antipodal::evaluateGrasp (antipodal.cpp:8-84).  One CTA per sample.
//
// Restructuring (results identical, all predicates are order independent):
//  * the r = nn_radius_hands neighbourhood is enumerated from contiguous cell-row
//    spans, tested with the exact float d2 < (float)(r*r), centred in float, cast
//    to double and cropped with -hh < z_f < hh; survivors are staged in shared
//    memory as (dx, dy, dz, sorted position) = 16 B;
//  * per orientation ONE pass over the staged points produces everything
//    evaluateFingers / deepenHand will ever ask: the 20-slot collision mask of a
//    point is a bit range found by exact compares against the host-built slot
//    tables, and it is OR-ed into occ[depth class of y]; the deepen loop is then a
//    prefix-OR over depth classes done by one thread (no FingerHand copies, no
//    re-scans);
//  * pass 2 counts the closing-region points and their x extent, pass 3 writes
//    their ids to the in-box pool and reduces the antipodal statistics.
// Neighbourhoods that exceed the staging buffer are processed in chunks; every
// accumulator is associative (OR / min / max / count), so chunking is exact.
// FP64 throughout, -fmad=false, same association order as the oracle.
#include <math.h>

#include <algorithm>

#include "dev_common.cuh"

namespace ag2 {
namespace {

constexpr int kHandThreads = 256;
constexpr int kMaxRows = 1024;
constexpr int kBatch = 4;  // candidates per thread per fill batch

struct HandsArgs {
  GridDev g;
  const ag2_frame* frames;
  uint32_t cap;                // staging capacity (points)
  ag2_sweep_record* sweeps;    // S*R
  ag2_hypothesis* slots;       // S*R
  BoxRef* box_ref;             // S*R
  int32_t* pool;
  uint32_t pool_cap;
  uint32_t* pool_counter;      // [0] next free, [1] overflow flag, [2] max m
  unsigned long long* stat;    // [0] sum K, [1] sum k', [2] sum m
};

__device__ __forceinline__ long long d_ordered(double v) {
  long long i = __double_as_longlong(v);
  return i >= 0 ? i : i ^ 0x7fffffffffffffffll;
}
__device__ __forceinline__ double ordered_d(long long i) {
  return __longlong_as_double(i >= 0 ? i : i ^ 0x7fffffffffffffffll);
}

// first j in [0, n] with y < thr[j] (n if none); thr ascending, est is only a hint
__device__ __forceinline__ int depth_class(double y, const double* thr, int n, double inv_step) {
  double e = floor((y - thr[0]) * inv_step) + 1.0;
  int j = e < 0.0 ? 0 : (e > (double)n ? n : (int)e);
  while (j > 0 && y < thr[j - 1]) j--;
  while (j < n && !(y < thr[j])) j++;
  return j;
}

// bits i in [0,10) with lo[i] < x && x < hi[i]; lo, hi ascending
__device__ __forceinline__ uint32_t slot_range(double x, const double* lo, const double* hi, double inv_step) {
  double eb = floor((x - lo[0]) * inv_step) + 1.0;
  int b = eb < 0.0 ? 0 : (eb > 10.0 ? 10 : (int)eb);
  while (b > 0 && !(lo[b - 1] < x)) b--;
  while (b < 10 && lo[b] < x) b++;
  double ea = floor((x - hi[0]) * inv_step) + 1.0;
  int a = ea < 0.0 ? 0 : (ea > 10.0 ? 10 : (int)ea);
  while (a > 0 && x < hi[a - 1]) a--;
  while (a < 10 && !(x < hi[a])) a++;
  return b > a ? (((1u << b) - 1u) & ~((1u << a) - 1u)) : 0u;
}

struct OrientState {  // decided by one thread per orientation after pass 1
  int feasible;
  int slot, last;
  double top, bottom, left, right, center;
  double lc, sx, sy, sz;      // unit-box scaling (hand_search.cpp:399-409)
};

struct Shared {
  double F[3][3];
  double sample[3];
  double FRx[AG2_MAX_ORIENTATIONS][3];  // frame_rot column 0 (closing direction)
  double FRy[AG2_MAX_ORIENTATIONS][3];  // frame_rot column 1 (approach)
  uint32_t occ[AG2_MAX_ORIENTATIONS][AG2_MAX_DEEPEN_STEPS + 2];
  int minct[AG2_MAX_ORIENTATIONS];
  int mincb[AG2_MAX_ORIENTATIONS];
  long long ymin[AG2_MAX_ORIENTATIONS];
  OrientState os[AG2_MAX_ORIENTATIONS];
  uint32_t m[AG2_MAX_ORIENTATIONS];
  long long minx[AG2_MAX_ORIENTATIONS], maxx[AG2_MAX_ORIENTATIONS];
  uint32_t off[AG2_MAX_ORIENTATIONS], fill[AG2_MAX_ORIENTATIONS];
  uint32_t anyl[AG2_MAX_ORIENTATIONS], anyr[AG2_MAX_ORIENTATIONS];
  long long ap[AG2_MAX_ORIENTATIONS][8];  // l: maxy miny maxz minz, r: same (ordered)
  uint32_t row_prefix[kMaxRows + 1];
  uint32_t row_begin[kMaxRows];
  uint32_t nkept, total_cand, cursor, pool_ok;
  unsigned long long sumK, sumKc;
};

}  // namespace
}  // namespace ag2

#include "hands_kernel.cuh"
