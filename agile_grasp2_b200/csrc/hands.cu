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
//    evaluateFingers / deepenHand will ever ask:
//      - the 20-slot collision mask of a point depends only on which of the 41
//        intervals between the (sorted) slot edges its x falls into, so it is one
//        table lookup (host-built, exact strict/non-strict handling of x == edge);
//        the mask is OR-ed into occ[depth class of y];
//      - "a point below the hand base" and "no point below the bite" are monotone
//        in y, so both follow from the exact minimum of y;
//    the deepen loop is then a prefix-OR over depth classes done by one thread (no
//    FingerHand copies, no re-scans);
//  * pass 2 counts the closing-region points and their x extent, pass 3 writes
//    their ids to the in-box pool and reduces the antipodal statistics.
// Neighbourhoods that exceed the staging buffer are processed in chunks; every
// accumulator is associative (OR / min / max / count), so chunking is exact.
// FP32 filter, FP64 decision: every predicate of the sweep compares a hand-frame coordinate a . p (|a| = 1,
// p a float difference shorter than the search radius) with a threshold.  The coordinate is first evaluated in
// FP32 (three operations, error <= eps32, tables.cpp); a bucket table turns it into the depth class / slot mask /
// in-box decision whenever it is farther than eps32 from every threshold, and only the rest (a fraction of 1e-3)
// is redone in FP64 with -fmad=false in the oracle's association order.  Minima / maxima that end up in a record
// (min y = grasp surface, x extent of the closing region) are taken exactly: FP32 finds the candidates within
// 2 eps32 of the float extremum, FP64 evaluates those.  Results are bit-identical to the all-FP64 sweep.
#include <math.h>
#include <stdlib.h>

#include <algorithm>

#include "dev_common.cuh"

namespace ag2 {
namespace {

constexpr int kHandThreads = 384;
constexpr int kMaxRows = 1024;
constexpr int kCand = 4;  // consecutive candidates per thread per fill batch
constexpr uint32_t kFillBatch = 32u * kCand;  // candidates per warp and fill batch
constexpr int kBatchRows = 1024;  // batches whose first cell row is tabulated (neighbourhoods up to 131 072 candidates)
constexpr int kMaxDepths = AG2_MAX_DEEPEN_STEPS + 2;

struct HandsArgs {
  GridDev g;
  const uint32_t* lut;         // tables of the FP32 filter (kLutWords words, ag2_internal.h)
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

// Per-orientation state (shared memory, one per orientation of this launch).
struct Orient {
  double FRx[3], FRy[3];  // frame_rot columns 0 (closing direction) and 1 (approach)
  float fx[3], fy[3];     // the same rounded to float: coefficients of the FP32 filter
  // closing-region box in float, shrunk / grown by the filter's error bound: certainly inside / certainly outside
  float top_in, top_out, left_in, left_out, right_in, right_out;
  long long ymin;         // ordered double: points.row(1).minCoeff()
  long long minx, maxx;   // x extent of the closing region
  long long ap[8];        // antipodal extrema, l: maxy miny maxz minz, r: same (ordered)
  // decided by one thread after pass 1
  double top, bottom, left, right, center;
  double lc, sx, sy, sz;  // unit-box scaling (hand_search.cpp:399-409)
  int feasible, slot, last;
  uint32_t m, off, fill, anyl, anyr;
};

struct Fixed {
  uint32_t lut[kLutWords];  // FP32 filter tables (copied from HandsArgs::lut)
  float Fz[3];              // hand axis (frame column 2) in float; crop bounds |z| certainly inside / outside
  float hh_in, hh_out;
  double F[3][3];
  double sample[3];
  // slot-edge lookup (HandConst::lut copied here: divergent indexing of constant memory serialises)
  double U[AG2_SLOT_EDGES];
  uint32_t M[AG2_SLOT_EDGES + 1];
  uint32_t Meq[AG2_SLOT_EDGES];
  uint8_t kstart[AG2_SLOT_BUCKETS];
  double depth[kMaxDepths];
  double bottom[kMaxDepths];
  uint32_t row_prefix[kMaxRows + 1];
  uint32_t row_begin[kMaxRows];
  uint16_t batch_row[kBatchRows];  // cell row that holds the first candidate of fill batch b
  uint32_t nkept, total_cand, cursor, pool_ok, nrows, next_batch;
  unsigned long long sumK, sumKc;
};

__host__ __device__ inline size_t hands_smem_base(int R, int ND) {
  size_t o = (sizeof(Fixed) + 15) & ~(size_t)15;
  o += (size_t)R * sizeof(Orient);
  o = (o + 15) & ~(size_t)15;
  o += (size_t)R * ND * sizeof(uint32_t);
  return (o + 15) & ~(size_t)15;
}

__device__ __forceinline__ int lut_bucket(float v, float v0, float scale, int nb) {
  return min(max((int)((v - v0) * scale), 0), nb - 1);
}
__device__ __forceinline__ float dot3f(const float a0, const float a1, const float a2, const float4& p) {
  return fmaf(a2, p.z, fmaf(a1, p.y, a0 * p.x));
}
__device__ __forceinline__ double dot3d_ref(const double* a, const float4& p) {
  // the reference's FP64 sequence (hand_search.cpp:359: frame_rot^T * points, no contraction)
  return (a[0] * (double)p.x + a[1] * (double)p.y) + a[2] * (double)p.z;
}

}  // namespace
}  // namespace ag2

#include "hands_kernel.cuh"
