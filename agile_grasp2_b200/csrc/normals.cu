// K4: surface normals, replacing pcl::NormalEstimationOMP (reference
// hand_search.cpp:83-94).  A count pass (thread per point) sizes the work; then one warp per point:
//   1. gather the r = normals_radius neighbourhood from the cell rows (float d2,
//      strict <, FLANN L2_Simple order) into shared memory as (d2, id) keys,
//   2. order the keys exactly (bucketed rank, see k_normals): PCL feeds
//      computeMeanAndCovarianceMatrix the neighbours in radiusSearch order (ascending
//      (d2, id)), and the single-pass FLOAT moment sums are order dependent at the
//      1e-3 rad level, so the order is part of the semantics,
//   3. nine lanes run the nine sequential float accumulators in that order,
//   4. one thread per point solves the 3x3 eigenproblem exactly as pcl::eigen33 does and
//      flips the normal toward the viewpoint (0,0,0).
// Multi-GPU: a rank can be asked for the interleaved 32-point granules it owns only
// (normals_shard_compute / normals_shard_assemble).
// Compiled with -fmad=false; every float op below is a separate IEEE operation.
#include <math.h>
#include <stdlib.h>

#include <algorithm>

#include "dev_common.cuh"

namespace ag2 {
namespace {

// thread per point: number of neighbours within the radius (sizes the sort buffers)
constexpr int kTiers = 8;
__constant__ uint32_t c_tier[kTiers] = {64, 96, 128, 192, 256, 384, 512, 1024};

// Multi-GPU: rank r of w owns the 32-point granules g with g % w == r of the cell-sorted cloud
// (interleaved, so that every rank sees every density of the scene).
__host__ __device__ __forceinline__ uint32_t shard_point(uint32_t t, uint32_t rank, uint32_t world) {
  return (((t >> 5) * world + rank) << 5) | (t & 31u);  // t-th owned point -> cell-sorted position
}

__global__ void k_count_neighbors(GridDev g, float r2, float rq, uint32_t* __restrict__ kn,
                                  uint32_t* __restrict__ max_out, unsigned long long* __restrict__ sum_out,
                                  uint32_t* __restrict__ tier_hist /* [kTiers]: points with kn > tier */,
                                  uint32_t shard_rank, uint32_t shard_world) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t cnt = 0;
  if (i < (uint32_t)g.n && (i >> 5) % shard_world == shard_rank) {
    float4 q = g.pts[i];
    QueryWindow w = make_window(g, q.x, q.y, q.z, rq);
    int rows = w.rows();
    for (int row = 0; row < rows; row++) {
      uint32_t b, e;
      row_span_trim(g, w, row, q.x, q.y, q.z, r2 * 1.002f, b, e);
      for (uint32_t j = b; j < e; j++) {
        float4 p = g.pts[j];
        float d2 = dist2_f(q.x - p.x, q.y - p.y, q.z - p.z);
        cnt += (d2 < r2);
      }
    }
    kn[i] = cnt;
  }
  uint32_t m = cnt;
  unsigned long long s = cnt;
  for (int o = 16; o > 0; o >>= 1) {
    m = max(m, __shfl_xor_sync(AG2_FULL, m, o));
    s += __shfl_xor_sync(AG2_FULL, s, o);
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMax(max_out, m);
    atomicAdd(sum_out, s);
  }
#pragma unroll
  for (int k = 0; k < kTiers; k++) {
    uint32_t over = __popc(__ballot_sync(AG2_FULL, cnt > c_tier[k]));
    if ((threadIdx.x & 31) == 0 && over) atomicAdd(&tier_hist[k], over);
  }
}

// pcl::computeRoots2
__device__ __forceinline__ void roots2(float b, float c, float r[3]) {
  r[0] = 0.f;
  float bb = b * b;  // float product, then double arithmetic: Scalar(b*b - 4.0*c)
  float d = (float)((double)bb - 4.0 * (double)c);
  if (d < 0.0f) d = 0.0f;
  float sd = sqrtf(d);
  r[2] = 0.5f * (b + sd);
  r[1] = 0.5f * (b - sd);
}

__device__ __forceinline__ void swapf(float& a, float& b) {
  float t = a;
  a = b;
  b = t;
}

// pcl::computeRoots; trig in double rounded to float (oracle deviation D2)
__device__ void compute_roots(const float m[9], float roots[3]) {
  float m00 = m[0], m01 = m[1], m02 = m[2], m11 = m[4], m12 = m[5], m22 = m[8];
  float c0 = m00 * m11 * m22 + 2.f * m01 * m02 * m12 - m00 * m12 * m12 - m11 * m02 * m02 - m22 * m01 * m01;
  float c1 = m00 * m11 - m01 * m01 + m00 * m22 - m02 * m02 + m11 * m22 - m12 * m12;
  float c2 = m00 + m11 + m22;
  if (fabsf(c0) < 1.1920928955078125e-07f) {
    roots2(c2, c1, roots);
    return;
  }
  const float s_inv3 = (float)(1.0 / 3.0);
  const float s_sqrt3 = sqrtf(3.0f);
  float c2_over_3 = c2 * s_inv3;
  float a_over_3 = (c1 - c2 * c2_over_3) * s_inv3;
  if (a_over_3 > 0.f) a_over_3 = 0.f;
  float half_b = 0.5f * (c0 + c2_over_3 * (2.f * c2_over_3 * c2_over_3 - c1));
  float q = half_b * half_b + a_over_3 * a_over_3 * a_over_3;
  if (q > 0.f) q = 0.f;
  float rho = sqrtf(-a_over_3);
  float theta = (float)atan2((double)sqrtf(-q), (double)half_b) * s_inv3;
  float cos_theta = (float)cos((double)theta);
  float sin_theta = (float)sin((double)theta);
  roots[0] = c2_over_3 + 2.f * rho * cos_theta;
  roots[1] = c2_over_3 - rho * (cos_theta + s_sqrt3 * sin_theta);
  roots[2] = c2_over_3 - rho * (cos_theta - s_sqrt3 * sin_theta);
  if (roots[0] >= roots[1]) swapf(roots[0], roots[1]);
  if (roots[1] >= roots[2]) {
    swapf(roots[1], roots[2]);
    if (roots[0] >= roots[1]) swapf(roots[0], roots[1]);
  }
  if (roots[0] <= 0.f) roots2(c2, c1, roots);
}

__device__ __forceinline__ void cross3f(const float* a, const float* b, float* o) {
  o[0] = a[1] * b[2] - a[2] * b[1];
  o[1] = a[2] * b[0] - a[0] * b[2];
  o[2] = a[0] * b[1] - a[1] * b[0];
}

// moments a[9] (already divided) -> normal flipped toward the origin
__device__ void solve_normal(const float a[9], float px, float py, float pz, float n[3]) {
  float cov[9];
  cov[0] = a[0] - a[6] * a[6];
  cov[1] = a[1] - a[6] * a[7];
  cov[2] = a[2] - a[6] * a[8];
  cov[4] = a[3] - a[7] * a[7];
  cov[5] = a[4] - a[7] * a[8];
  cov[8] = a[5] - a[8] * a[8];
  cov[3] = cov[1];
  cov[6] = cov[2];
  cov[7] = cov[5];
  // pcl::eigen33 (smallest eigenpair)
  float scale = 0.f;
  for (int i = 0; i < 9; i++) scale = fmaxf(scale, fabsf(cov[i]));
  if (scale <= 1.17549435e-38f) scale = 1.0f;
  float m[9];
  for (int i = 0; i < 9; i++) m[i] = cov[i] / scale;
  float roots[3];
  compute_roots(m, roots);
  m[0] -= roots[0];
  m[4] -= roots[0];
  m[8] -= roots[0];
  float v1[3], v2[3], v3[3];
  cross3f(&m[0], &m[3], v1);
  cross3f(&m[0], &m[6], v2);
  cross3f(&m[3], &m[6], v3);
  float l1 = v1[0] * v1[0] + (v1[1] * v1[1] + v1[2] * v1[2]);
  float l2 = v2[0] * v2[0] + (v2[1] * v2[1] + v2[2] * v2[2]);
  float l3 = v3[0] * v3[0] + (v3[1] * v3[1] + v3[2] * v3[2]);
  const float* best;
  float len;
  if (l1 >= l2 && l1 >= l3) {
    best = v1;
    len = l1;
  } else if (l2 >= l1 && l2 >= l3) {
    best = v2;
    len = l2;
  } else {
    best = v3;
    len = l3;
  }
  float s = sqrtf(len);
  n[0] = best[0] / s;
  n[1] = best[1] / s;
  n[2] = best[2] / s;
  float vx = 0.f - px, vy = 0.f - py, vz = 0.f - pz;
  float cos_theta = (vx * n[0] + vy * n[1]) + vz * n[2];
  if (cos_theta < 0) {
    n[0] *= -1;
    n[1] *= -1;
    n[2] *= -1;
  }
}

// Raw moment sums of a point (accu[0..8] of computeMeanAndCovarianceMatrix before the division), padded
// to three float4.  The neighbourhood kernels only produce these; k_eigen finishes every point of
// the cloud in one full-occupancy launch instead of a few lanes at the end of each warp's batch.
constexpr int kMomWords = 12;

// Destination arrays of k_eigen: the sorted normals of every rank (multi-GPU: peer-mapped pointers into
// the other GPUs' memory, group.cu) or just this context's own.
struct NrmDst {
  float4* p[8];
  int n;
};

__global__ void __launch_bounds__(256) k_eigen(GridDev g, const uint32_t* __restrict__ kn,
                                               const float* __restrict__ mom, const float4* nrm_own /* may alias dst.p[] */,
                                               NrmDst dst, uint32_t shard_rank, uint32_t shard_world) {
  const uint32_t i = shard_point(blockIdx.x * blockDim.x + threadIdx.x, shard_rank, shard_world);
  if (i >= (uint32_t)g.n) return;
  const uint32_t count = kn[i];
  const float nanv = __int_as_float(0x7fc00000);
  float n[3] = {nanv, nanv, nanv};  // computePointNormal: fewer than 3 neighbours -> NaN
  if (count >= 3) {
    const float4* m4 = reinterpret_cast<const float4*>(mom + (size_t)i * kMomWords);
    const float4 m0 = m4[0], m1 = m4[1], m2 = m4[2];
    float a[9] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w, m2.x};
    const float inv = 1.0f / (float)count;  // Eigen 3.2 operator/=: multiply by the reciprocal
#pragma unroll
    for (int k = 0; k < 9; k++) a[k] = a[k] * inv;
    const float4 q = g.pts[i];
    solve_normal(a, q.x, q.y, q.z, n);
  }
  // w = camera bits: every rank built the same cell-sorted cloud, so the value is the destination's own
  const float4 v = make_float4(n[0], n[1], n[2], nrm_own[i].w);
  // a warp's 32 points are one granule = 512 contiguous bytes per destination (NVLink stores when remote)
  for (int r = 0; r < dst.n; r++) dst.p[r][i] = v;
}

// ---------------------------------------------------------------------------
// k_normals.  A warp owns `bsz` (<= 32) points at a time, taken from a global counter.
// Per point (warp cooperative):
//   gather   the candidate spans of the <=16 cell rows are flattened into one
//            sequence (prefix over rows by shuffles), each lane tests one candidate
//            per iteration and hits are ballot-compacted into A = (d2 bits, j);
//   order    exact ascending (d2, id) WITHOUT a comparison sort: hits are grouped
//            into kNB monotone d2 buckets (shared atomics + warp prefix), and the
//            rank of an element is bucket start + #smaller (d2, id) inside its
//            bucket (a handful of compares; ids are fetched only on equal d2), then
//            its coordinates are written at `rank` of the staging array S;
//   moments  lanes 0..8 run the nine single-pass float accumulators of
//            computeMeanAndCovarianceMatrix over S in that order (sequential by
//            definition) and store the raw sums; k_eigen (thread per point, one launch
//            for the whole cloud) solves eigen33 and flips the normal.
// Shared memory per warp is 16 B x cap + 1.3 KB, so the launch capacity decides the
// occupancy: the host runs the bulk of the points at a small cap and hands the dense
// ones (appended to big_list here) to further launches with larger caps.
constexpr int kNB = 128;     // d2 buckets (4 per lane in the prefix)

struct NormalsSmem {  // per-warp carve-up (byte offsets, 16-aligned)
  size_t a, s, cnt, start, rowb, rowp, total;
};
__host__ __device__ inline NormalsSmem normals_smem(uint32_t cap) {
  NormalsSmem L;
  size_t o = 0;
  L.a = o, o += (size_t)cap * 8;    // A: (d2 bits, j); later J: positions in radiusSearch order | Z
  L.s = o, o += (size_t)cap * 8;    // G: (d2 bits, j) grouped by bucket; later X | Y
  L.cnt = o, o += kNB * 4;
  L.start = o, o += (kNB + 4) * 4;
  L.rowb = o, o += 32 * 4;
  L.rowp = o, o += 36 * 4;
  L.total = (o + 15) & ~(size_t)15;
  return L;
}

__device__ __forceinline__ int d2_bucket(float d2, float bscale) {
  return min(kNB - 1, (int)__fmul_rn(d2, bscale));  // monotone non-decreasing in d2
}

struct NormalsArgs {
  GridDev g;
  float r2, rq, bscale;
  const uint32_t* kn;
  uint32_t n_items, cap, bsz;
  const uint32_t* n_items_dev;  // when set: the item count lives in device memory (tier sizes of the count pass);
                                // the kernel then derives its own batch size and n_items / bsz above are unused
  uint32_t skip_from;           // with n_items_dev: the launch does nothing when the count is >= skip_from (the
                                // four-lanes-per-point kernel takes such a tier); 0 = never
  uint32_t shard_rank, shard_world;
  const uint32_t* list;       // nullptr: items are the owned cell-sorted points (shard_point), n_items of them
  uint32_t* big_list;         // out: items whose neighbourhood exceeds cap
  uint32_t* big_count;
  uint32_t big_cap;
  uint32_t* batch_counter;    // dynamic batch scheduling
  float* mom_out;             // [n][kMomWords] raw moment sums; k_eigen turns them into normals
  float4* nrm_out;
};

__global__ void __launch_bounds__(256) k_normals(NormalsArgs P) {
  extern __shared__ __align__(16) unsigned char smem[];
  const GridDev& g = P.g;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t n_items = P.n_items, bsz = P.bsz;
  if (P.n_items_dev) {
    // launched at full residency before the count is known on the host: same batch-size rule as the host's
    n_items = *P.n_items_dev;
    if (n_items == 0 || (P.skip_from && n_items >= P.skip_from)) return;
    const uint32_t resident_warps = gridDim.x * (blockDim.x >> 5);
    bsz = 32;
    while (bsz > 1 && (n_items + bsz - 1) / bsz < 4 * resident_warps) bsz >>= 1;
  }
  const uint32_t cap = P.cap;
  const NormalsSmem L = normals_smem(cap);
  unsigned char* base = smem + (size_t)warp * L.total;
  uint2* A = (uint2*)(base + L.a);
  uint2* G = (uint2*)(base + L.s);
  // after the ranking: J | Z reuse A's bytes, X | Y reuse G's (A and G are dead by then)
  uint32_t* J = (uint32_t*)(base + L.a);
  float* Zs = (float*)(base + L.a) + cap;
  float* Xs = (float*)(base + L.s);
  float* Ys = (float*)(base + L.s) + cap;
  uint32_t* cnt = (uint32_t*)(base + L.cnt);
  uint32_t* start = (uint32_t*)(base + L.start);
  uint32_t* rowb = (uint32_t*)(base + L.rowb);
  uint32_t* rowp = (uint32_t*)(base + L.rowp);
  const uint32_t n_batches = (n_items + bsz - 1) / bsz;
  // accumulator role of lanes 0..8: operand selectors 0 = x, 1 = y, 2 = z, 3 = constant 1
  const int sa = lane < 3 ? 0 : (lane < 5 ? 1 : (lane == 5 ? 2 : lane - 6));
  const int sb = lane < 3 ? lane : (lane < 5 ? lane - 2 : (lane == 5 ? 2 : 3));
  __shared__ __align__(16) float ones_s[4];
  if (threadIdx.x < 4) ones_s[threadIdx.x] = 1.0f;
  __syncthreads();

  // Batches of one point (short tier lists) are dealt out without a counter, every warp a contiguous run of the
  // list -- consecutive entries are neighbours in space, so the run re-reads what L1 already holds; larger
  // batches are claimed from the global counter, the following one a batch ahead so that the atomic's round trip
  // hides behind this one.
  const bool dealt = bsz == 1;
  const uint32_t warp_id = blockIdx.x * (blockDim.x >> 5) + warp, n_warps = gridDim.x * (blockDim.x >> 5);
  const uint32_t run = (n_batches + n_warps - 1) / n_warps;
  uint32_t batch = warp_id * run;
  const uint32_t batch_end = dealt ? min(n_batches, batch + run) : n_batches;
  if (!dealt) {
    if (lane == 0) batch = atomicAdd(P.batch_counter, 1u);
    batch = __shfl_sync(AG2_FULL, batch, 0);
  }
  while (batch < batch_end) {
    uint32_t next = batch + 1;
    if (!dealt && lane == 0) next = atomicAdd(P.batch_counter, 1u);
    const uint32_t i0 = batch * bsz;
    const uint32_t nb = min(bsz, n_items - i0);
    const uint32_t my_i =
        lane < nb ? (P.list ? P.list[i0 + lane] : shard_point(i0 + lane, P.shard_rank, P.shard_world)) : 0u;
    const uint32_t my_count = lane < nb ? P.kn[my_i] : 0u;
    const bool too_big = my_count > cap;
    if (too_big && P.big_count) {
      uint32_t slot = atomicAdd(P.big_count, 1u);
      if (slot < P.big_cap) P.big_list[slot] = my_i;
    }
    for (uint32_t k = 0; k < nb; k++) {
      const uint32_t i = __shfl_sync(AG2_FULL, my_i, (int)k);
      const uint32_t count = __shfl_sync(AG2_FULL, my_count, (int)k);
      if (count < 3 || count > cap) continue;  // NaN normal (solve pass) / a later launch
      const float4 q = g.pts[i];
      // ---- gather
      QueryWindow w = make_window(g, q.x, q.y, q.z, P.rq);
      const int rows = w.rows();  // <= 16 whenever the cell size is >= normals_radius
#pragma unroll
      for (int b = 0; b < kNB / 32; b++) cnt[lane + 32 * b] = 0;
      uint32_t filled = 0;
      for (int r0 = 0; r0 < rows; r0 += 32) {
        uint32_t rb = 0, rc = 0;
        if (r0 + lane < rows) {
          uint32_t b, e;
          row_span_trim(g, w, r0 + lane, q.x, q.y, q.z, P.r2 * 1.002f, b, e);
          rb = b, rc = e - b;
        }
        uint32_t incl = rc;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          uint32_t t = __shfl_up_sync(AG2_FULL, incl, o);
          if (lane >= o) incl += t;
        }
        const uint32_t T = __shfl_sync(AG2_FULL, incl, 31);
        __syncwarp();
        rowb[lane] = rb;
        rowp[lane] = incl - rc;
        if (lane == 31) rowp[32] = T;
        __syncwarp();
        int row = 0;
        uint32_t row_lo = 0, row_hi = rowp[1], row_b = rowb[0];
        // the candidate of the NEXT iteration is located and loaded one iteration ahead
        uint32_t j_next = 0;
        float4 p_next = make_float4(0.f, 0.f, 0.f, 0.f);
        if (lane < T) {
          while (lane >= row_hi) {
            row++;
            row_lo = row_hi;
            row_hi = rowp[row + 1];
            row_b = rowb[row];
          }
          j_next = row_b + ((uint32_t)lane - row_lo);
          p_next = g.pts[j_next];
        }
        for (uint32_t t0 = 0; t0 < T; t0 += 32) {
          const uint32_t t = t0 + lane;
          const uint32_t j = j_next;
          const float4 p = p_next;
          const uint32_t tn = t + 32;
          if (tn < T) {
            while (tn >= row_hi) {
              row++;
              row_lo = row_hi;
              row_hi = rowp[row + 1];
              row_b = rowb[row];
            }
            j_next = row_b + (tn - row_lo);
            p_next = g.pts[j_next];
          }
          bool hit = false;
          float d2 = 0.f;
          if (t < T) {
            d2 = dist2_f(q.x - p.x, q.y - p.y, q.z - p.z);
            hit = d2 < P.r2;
          }
          const uint32_t ballot = __ballot_sync(AG2_FULL, hit);
          if (hit) {
            A[filled + __popc(ballot & ((1u << lane) - 1))] = make_uint2(__float_as_uint(d2), j);
            atomicAdd(&cnt[d2_bucket(d2, P.bscale)], 1u);
          }
          filled += __popc(ballot);
        }
      }
      __syncwarp();
      // ---- bucket starts: exclusive prefix over kNB counters, kNB/32 consecutive ones per lane
      {
        constexpr int per = kNB / 32;
        uint32_t cl[per], s = 0;
#pragma unroll
        for (int b = 0; b < per; b++) {
          cl[b] = cnt[per * lane + b];
          s += cl[b];
        }
        uint32_t in2 = s;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          uint32_t t = __shfl_up_sync(AG2_FULL, in2, o);
          if (lane >= o) in2 += t;
        }
        uint32_t ex = in2 - s;
#pragma unroll
        for (int b = 0; b < per; b++) {
          start[per * lane + b] = ex;
          cnt[per * lane + b] = ex;  // cursor
          ex += cl[b];
        }
        if (lane == 31) start[kNB] = in2;
      }
      __syncwarp();
      // ---- group by bucket
      for (uint32_t t = lane; t < count; t += 128) {  // four shared atomics in flight per lane
        uint2 e[4];
        uint32_t slot[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
          e[u] = A[min(t + 32u * u, count - 1)];
          slot[u] = 0;
          if (t + 32u * u < count) slot[u] = atomicAdd(&cnt[d2_bucket(__uint_as_float(e[u].x), P.bscale)], 1u);
        }
#pragma unroll
        for (int u = 0; u < 4; u++)
          if (t + 32u * u < count) G[slot[u]] = e[u];
      }
      __syncwarp();
      // ---- exact rank inside the bucket by (d2, input index); j goes to J[rank]
      for (uint32_t t = lane; t < count; t += 32) {
        const uint2 e = G[t];
        const int b = d2_bucket(__uint_as_float(e.x), P.bscale);
        const uint32_t lo = start[b], hi = start[b + 1];
        uint32_t rank = lo;
        for (uint32_t u = lo; u < hi; u++) {
          const uint2 o = G[u];
          if (o.x < e.x) {
            rank++;
          } else if (o.x == e.x && u != t) {  // equal distances: FLANN orders by index
            if (__float_as_uint(g.pts[o.y].w) < __float_as_uint(g.pts[e.y].w)) rank++;
          }
        }
        J[rank] = e.y;
      }
      __syncwarp();
      for (uint32_t t = lane; t < count; t += 128) {  // four loads in flight per lane
        float4 p[4];
#pragma unroll
        for (int u = 0; u < 4; u++) p[u] = g.pts[J[min(t + 32u * u, count - 1)]];
#pragma unroll
        for (int u = 0; u < 4; u++)
          if (t + 32u * u < count) {
            Xs[t + 32u * u] = p[u].x;
            Ys[t + 32u * u] = p[u].y;
            Zs[t + 32u * u] = p[u].z;
          }
      }
      __syncwarp();
      // ---- sequential single-pass float moments
      if (lane < 9) {
        // accu[6..8] += p.x : the second operand is the constant 1.0f, the product is exact
        const float4* pa = (const float4*)(sa == 0 ? Xs : (sa == 1 ? Ys : Zs));
        const float4* pb = (const float4*)(sb == 3 ? ones_s : (sb == 0 ? Xs : (sb == 1 ? Ys : Zs)));
        const uint32_t step_b = sb == 3 ? 0u : 1u;
        float acc = 0.f;
        const uint32_t n4 = count >> 2;
        for (uint32_t t = 0; t < n4; t++) {
          const float4 va = pa[t], vb = pb[step_b * t];
          acc = __fadd_rn(acc, __fmul_rn(va.x, vb.x));
          acc = __fadd_rn(acc, __fmul_rn(va.y, vb.y));
          acc = __fadd_rn(acc, __fmul_rn(va.z, vb.z));
          acc = __fadd_rn(acc, __fmul_rn(va.w, vb.w));
        }
        const float* ra = (const float*)pa;
        const float* rb_ = (const float*)pb;
        for (uint32_t t = n4 << 2; t < count; t++) acc = __fadd_rn(acc, __fmul_rn(ra[t], rb_[step_b * t]));
        P.mom_out[(size_t)i * kMomWords + lane] = acc;
      }
      __syncwarp();
    }
    // ---- fewer than 3 neighbours -> NaN (computePointNormal); everything else is k_eigen's
    if (lane < nb && my_count < 3) {
      const float nanv = __int_as_float(0x7fc00000);
      P.nrm_out[my_i] = make_float4(nanv, nanv, nanv, P.nrm_out[my_i].w);
    }
    batch = __shfl_sync(AG2_FULL, next, 0);
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------
// k_normals_cta: one CTA (four warps) per point, for the DENSE tiers.  Same algorithm as k_normals
// (flattened candidate sequence, ballot compaction, monotone d2 buckets, exact rank inside the
// bucket, coordinates staged in search order, nine sequential accumulators on lanes 0..8 of warp 0),
// but the 16 B x cap of shared memory now belong to 128 threads instead of 32: a neighbourhood of a
// thousand points no longer leaves the SM with a handful of warps.  Points are claimed from a global
// counter, the next one a point ahead.
constexpr int kCtaThreads = 128;
constexpr int kCtaNB = 512;  // d2 buckets

struct CtaSmem {
  size_t a, s, cnt, start, rowb, rowp, misc, total;
};
__host__ __device__ inline CtaSmem cta_smem(uint32_t cap) {
  CtaSmem L;
  size_t o = 0;
  L.a = o, o += (size_t)cap * 8;   // A: (d2 bits, j); later J | Z
  L.s = o, o += (size_t)cap * 8;   // G: bucket-grouped (d2 bits, j); later X | Y
  L.cnt = o, o += kCtaNB * 4;
  L.start = o, o += (kCtaNB + 4) * 4;
  L.rowb = o, o += 32 * 4;
  L.rowp = o, o += 36 * 4;
  L.misc = o, o += 16 * 4;         // [0] hits so far, [1] T, [2] claimed point, [4..7] warp totals, [8..11] ones
  L.total = (o + 15) & ~(size_t)15;
  return L;
}

__global__ void __launch_bounds__(kCtaThreads) k_normals_cta(NormalsArgs P) {
  extern __shared__ __align__(16) unsigned char smem[];
  const GridDev& g = P.g;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t cap = P.cap;
  const CtaSmem L = cta_smem(cap);
  uint2* A = (uint2*)(smem + L.a);
  uint2* G = (uint2*)(smem + L.s);
  uint32_t* J = (uint32_t*)(smem + L.a);
  float* Zs = (float*)(smem + L.a) + cap;
  float* Xs = (float*)(smem + L.s);
  float* Ys = (float*)(smem + L.s) + cap;
  uint32_t* cnt = (uint32_t*)(smem + L.cnt);
  uint32_t* start = (uint32_t*)(smem + L.start);
  uint32_t* rowb = (uint32_t*)(smem + L.rowb);
  uint32_t* rowp = (uint32_t*)(smem + L.rowp);
  uint32_t* misc = (uint32_t*)(smem + L.misc);
  float* ones_s = (float*)(misc + 8);
  const int sa = lane < 3 ? 0 : (lane < 5 ? 1 : (lane == 5 ? 2 : lane - 6));
  const int sb = lane < 3 ? lane : (lane < 5 ? lane - 2 : (lane == 5 ? 2 : 3));
  const float bscale = (float)kCtaNB / P.r2;
  const uint32_t n_items = P.n_items_dev ? *P.n_items_dev : P.n_items;
  if (n_items == 0) return;
  if (tid < 4) ones_s[tid] = 1.0f;
  if (tid == 0) misc[2] = atomicAdd(P.batch_counter, 1u);
  __syncthreads();
  uint32_t item = misc[2];
  while (item < n_items) {
    __syncthreads();  // everyone has read misc[2]
    if (tid == 0) misc[2] = atomicAdd(P.batch_counter, 1u);  // the next point, read at the end of this one
    const uint32_t i = P.list[item];
    const uint32_t count = P.kn[i];  // 3 <= count <= cap (tier list)
    const float4 q = g.pts[i];
    const QueryWindow w = make_window(g, q.x, q.y, q.z, P.rq);
    const int rows = w.rows();
    for (int b = tid; b < kCtaNB; b += kCtaThreads) cnt[b] = 0;
    if (tid == 0) misc[0] = 0;
    // ---- gather
    for (int r0 = 0; r0 < rows; r0 += 32) {
      __syncthreads();  // rowb / rowp of the previous round are no longer read; counters zeroed
      if (warp == 0) {
        uint32_t rb = 0, rc = 0;
        if (r0 + lane < rows) {
          uint32_t b, e;
          row_span_trim(g, w, r0 + lane, q.x, q.y, q.z, P.r2 * 1.002f, b, e);
          rb = b, rc = e - b;
        }
        uint32_t incl = rc;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          uint32_t t = __shfl_up_sync(AG2_FULL, incl, o);
          if (lane >= o) incl += t;
        }
        rowb[lane] = rb;
        rowp[lane] = incl - rc;
        if (lane == 31) {
          rowp[32] = incl;
          misc[1] = incl;
        }
      }
      __syncthreads();
      const uint32_t T = misc[1];
      int row = 0;
      uint32_t row_lo = 0, row_hi = rowp[1], row_b = rowb[0];
      for (uint32_t t0 = 0; t0 < T; t0 += kCtaThreads) {
        const uint32_t t = t0 + tid;
        bool hit = false;
        float d2 = 0.f;
        uint32_t j = 0;
        if (t < T) {
          while (t >= row_hi) {
            row++;
            row_lo = row_hi;
            row_hi = rowp[row + 1];
            row_b = rowb[row];
          }
          j = row_b + (t - row_lo);
          const float4 p = g.pts[j];
          d2 = dist2_f(q.x - p.x, q.y - p.y, q.z - p.z);
          hit = d2 < P.r2;
        }
        const uint32_t ballot = __ballot_sync(AG2_FULL, hit);
        uint32_t wbase = 0;
        if (lane == 0 && ballot) wbase = atomicAdd(&misc[0], (uint32_t)__popc(ballot));
        wbase = __shfl_sync(AG2_FULL, wbase, 0);
        if (hit) {
          const uint32_t slot = wbase + __popc(ballot & ((1u << lane) - 1));
          if (slot < cap) A[slot] = make_uint2(__float_as_uint(d2), j);
          atomicAdd(&cnt[min(kCtaNB - 1, (int)__fmul_rn(d2, bscale))], 1u);
        }
      }
    }
    __syncthreads();
    // ---- bucket starts: exclusive prefix over kCtaNB counters, four consecutive ones per thread
    {
      constexpr int per = kCtaNB / kCtaThreads;
      uint32_t cl[per], sum = 0;
#pragma unroll
      for (int b = 0; b < per; b++) {
        cl[b] = cnt[per * tid + b];
        sum += cl[b];
      }
      uint32_t in2 = sum;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(AG2_FULL, in2, o);
        if (lane >= o) in2 += t;
      }
      if (lane == 31) misc[4 + warp] = in2;
      __syncthreads();
      uint32_t wofs = 0;
      for (int k = 0; k < warp; k++) wofs += misc[4 + k];
      uint32_t ex = wofs + in2 - sum;
#pragma unroll
      for (int b = 0; b < per; b++) {
        start[per * tid + b] = ex;
        cnt[per * tid + b] = ex;  // cursor
        ex += cl[b];
      }
      if (tid == kCtaThreads - 1) start[kCtaNB] = ex;
    }
    __syncthreads();
    // ---- group by bucket
    for (uint32_t t = tid; t < count; t += kCtaThreads) {
      const uint2 e = A[t];
      G[atomicAdd(&cnt[min(kCtaNB - 1, (int)__fmul_rn(__uint_as_float(e.x), bscale))], 1u)] = e;
    }
    __syncthreads();
    // ---- exact rank inside the bucket by (d2, input index); j goes to J[rank]
    for (uint32_t t = tid; t < count; t += kCtaThreads) {
      const uint2 e = G[t];
      const int b = min(kCtaNB - 1, (int)__fmul_rn(__uint_as_float(e.x), bscale));
      const uint32_t lo = start[b], hi = start[b + 1];
      uint32_t rank = lo;
      for (uint32_t u = lo; u < hi; u++) {
        const uint2 o = G[u];
        if (o.x < e.x) {
          rank++;
        } else if (o.x == e.x && u != t) {  // equal distances: FLANN orders by index
          if (__float_as_uint(g.pts[o.y].w) < __float_as_uint(g.pts[e.y].w)) rank++;
        }
      }
      J[rank] = e.y;
    }
    __syncthreads();
    // ---- coordinates in search order
    for (uint32_t t = tid; t < count; t += kCtaThreads) {
      const float4 p = g.pts[J[t]];
      Xs[t] = p.x;
      Ys[t] = p.y;
      Zs[t] = p.z;
    }
    __syncthreads();
    // ---- sequential single-pass float moments (warp 0, lanes 0..8)
    if (tid < 9) {
      const float4* pa = (const float4*)(sa == 0 ? Xs : (sa == 1 ? Ys : Zs));
      const float4* pb = (const float4*)(sb == 3 ? ones_s : (sb == 0 ? Xs : (sb == 1 ? Ys : Zs)));
      const uint32_t step_b = sb == 3 ? 0u : 1u;
      float acc = 0.f;
      const uint32_t n4 = count >> 2;
      for (uint32_t t = 0; t < n4; t++) {
        const float4 va = pa[t], vb = pb[step_b * t];
        acc = __fadd_rn(acc, __fmul_rn(va.x, vb.x));
        acc = __fadd_rn(acc, __fmul_rn(va.y, vb.y));
        acc = __fadd_rn(acc, __fmul_rn(va.z, vb.z));
        acc = __fadd_rn(acc, __fmul_rn(va.w, vb.w));
      }
      const float* ra = (const float*)pa;
      const float* rb_ = (const float*)pb;
      for (uint32_t t = n4 << 2; t < count; t++) acc = __fadd_rn(acc, __fmul_rn(ra[t], rb_[step_b * t]));
      P.mom_out[(size_t)i * kMomWords + tid] = acc;
    }
    __syncthreads();
    item = misc[2];
  }
}

// ---------------------------------------------------------------------------
// Tiered path.  The count pass groups the points by neighbourhood size into tiers; every
// non-empty tier is ONE launch sized for exactly that capacity, and the launches are
// independent, so they run side by side on the ctx's auxiliary streams.
//   the sparsest tiers (by default 0..3, <= 96 neighbours; any of 0..7 can): k_normals_quad,
//               four lanes per point;
//   the others: k_normals, one warp per point.
// The last capacity is what one CTA's shared memory holds (16 B per neighbour); a denser point is an error
// (AG2_E_CAPACITY, reported at the next synchronisation).
constexpr int kQuadTiers = 8;
constexpr int kAllTiers = 16;
constexpr uint32_t kTierCapHost[kAllTiers] = {48,  64,  80,  96,   112,  128,  160,  208,
                                              256, 384, 512, 1024, 2048, 4096, 8192, 13600};
__constant__ uint32_t c_tier_cap[kAllTiers] = {48,  64,  80,  96,   112,  128,  160,  208,
                                               256, 384, 512, 1024, 2048, 4096, 8192, 13600};
// tier_cnt layout: [0..15] tier sizes, [16..31] batch counters of the tier launches, [32] points over the last capacity
constexpr int kTierCntWords = 48;

// FOUR lanes per point (they share out its cell rows; a rank's share of a cloud is too few points to hide a
// thread-per-point walk's chain of dependent loads): neighbour count, NaN normal below three neighbours, tier
// membership lists (one warp-aggregated atomic per tier present in the warp, so consecutive cell-sorted points
// stay consecutive in their list and keep their cache locality).
template <int LANES>  // lanes per point: 4, or 1 when the launch has points enough to fill the machine by itself
__global__ void __launch_bounds__(128) k_count_tiers(GridDev g, float r2, float rq, uint32_t* __restrict__ kn,
                                                     uint32_t* __restrict__ max_out,
                                                     unsigned long long* __restrict__ sum_out,
                                                     uint32_t* __restrict__ tier_count, uint32_t* __restrict__ tier_list,
                                                     uint32_t stride, float4* __restrict__ nrm_out,
                                                     uint32_t shard_rank, uint32_t shard_world) {
  // group t of the launch = the t-th point this rank owns (whole warps of work at any world size)
  const uint32_t gtid = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t i = shard_point(LANES == 4 ? gtid >> 2 : gtid, shard_rank, shard_world);
  const uint32_t lane = threadIdx.x & 31, sub = LANES == 4 ? lane & 3u : 0u;
  const bool in_range = i < (uint32_t)g.n, owned = in_range && sub == 0;
  uint32_t cnt = 0, cand = 0;
  if (in_range) {
    const float4 q = g.pts[i];
    const QueryWindow w = make_window(g, q.x, q.y, q.z, rq);
    const int rows = w.rows();
    for (int row = (int)sub; row < rows; row += LANES) {
      uint32_t b, e;
      row_span_trim(g, w, row, q.x, q.y, q.z, r2 * 1.002f, b, e);
      cand += e - b;
      for (uint32_t j = b; j < e; j++) {
        const float4 p = g.pts[j];
        cnt += (dist2_f(q.x - p.x, q.y - p.y, q.z - p.z) < r2);
      }
    }
  }
  if (LANES == 4) {
    cnt += __shfl_xor_sync(AG2_FULL, cnt, 1);
    cnt += __shfl_xor_sync(AG2_FULL, cnt, 2);
    cand += __shfl_xor_sync(AG2_FULL, cand, 1);
    cand += __shfl_xor_sync(AG2_FULL, cand, 2);
  }
  if (!owned) cnt = 0;  // one lane per point reports
  if (owned) {
    kn[i] = cnt;
    if (cnt < 3) {  // computePointNormal: fewer than 3 neighbours -> NaN
      const float nanv = __int_as_float(0x7fc00000);
      nrm_out[i] = make_float4(nanv, nanv, nanv, nrm_out[i].w);
    }
  }
  uint32_t m = cnt;
  unsigned long long s = cnt;
  for (int o = 16; o > 0; o >>= 1) {
    m = max(m, __shfl_xor_sync(AG2_FULL, m, o));
    s += __shfl_xor_sync(AG2_FULL, s, o);
  }
  if (lane == 0) {
    atomicMax(max_out, m);
    atomicAdd(sum_out, s);
  }
  int tier = -1;
  if (owned && cnt >= 3) {
    if (cnt > c_tier_cap[kAllTiers - 1]) {
      atomicAdd(&tier_count[32], 1u);  // no launch can hold this neighbourhood: reported as AG2_E_CAPACITY
    } else {
      // A tier fixes the capacity of its launch AND how many lanes walk a point's candidates (4, 32, 128).  A sparse
      // point beside a dense region has few neighbours but thousands of candidates: on four lanes it would outlast
      // its whole launch.  Such a point is filed by its walk (a sixth of the candidates, the usual hit rate being
      // one in five) -- any tier at or above the point's own is a valid home, the normal does not depend on it.
      // Tier 11 (one CTA per point) is as far as the walk alone moves a point: the launches above it are sized
      // by the previous cloud's largest neighbourhood.
      const uint32_t eff = max(cnt, cand / 6u);
      tier = 0;
      while (tier < 11 && eff > c_tier_cap[tier]) tier++;
      while (cnt > c_tier_cap[tier]) tier++;
    }
  }
  const uint32_t peers = __match_any_sync(AG2_FULL, tier);
  if (tier >= 0) {
    const int leader = __ffs(peers) - 1;
    uint32_t base = 0;
    if ((int)lane == leader) base = atomicAdd(&tier_count[tier], (uint32_t)__popc(peers));
    base = __shfl_sync(peers, base, leader);
    tier_list[(size_t)tier * stride + base + __popc(peers & ((1u << lane) - 1u))] = i;
  }
}

// k_normals_quad: FOUR lanes per point, eight points per warp.  The four lanes share one
// shared-memory neighbour list; no data crosses a group.
//   gather   the group's cell rows (trimmed by its four lanes, four rows each) are flattened into one
//            candidate sequence and walked four at a time (float d2, strict <); hits are compacted
//            with the group's four bits of a warp ballot and appended as 32-bit keys (q << jbits) | j,
//            j = cell-sorted position (jbits = bits of n), q = a monotone quantisation of d2 on
//            the remaining bits;
//   order    counting sort on the top five key bits into a byte permutation (shared atomics);
//            each lane then owns eight consecutive buckets = one contiguous segment and insertion-
//            sorts it.  Keys with different q are ordered by q; EQUAL q falls back to the exact
//            (d2, input index) comparison, recomputing d2 with the gather's own float sequence --
//            so the result is exactly FLANN's radiusSearch order;
//   moments  lanes 0..2 of the group run three of the nine sequential float accumulators each and
//            store the raw sums for k_eigen.
// Lists are laid out [entry][group] per warp, so the four consecutive entries a group touches
// fall into distinct banks whatever the other groups do.  Shared memory is 5 B x cap + 128 B per
// point (a quarter of that per thread), which is what gives this kernel its occupancy.
constexpr int kQuadThreads = 128;

struct QuadArgs {
  GridDev g;
  float r2, rq, qscale;
  uint32_t qmax, jbits;
  const uint32_t* kn;
  const uint32_t* list;
  uint32_t n_items, cap;
  const uint32_t* n_items_dev;  // when set: item count in device memory
  uint32_t min_items;           // the launch does nothing below this count (>= 1: an empty list is skipped)
  float* mom_out;
};

__host__ __device__ inline size_t quad_smem_per_warp(uint32_t cap) { return (size_t)cap * 40 + 1024 + 1088; }

__global__ void __launch_bounds__(kQuadThreads) k_normals_quad(QuadArgs P) {
  extern __shared__ __align__(16) unsigned char smem[];
  const GridDev& g = P.g;
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31, grp = lane >> 2, sub = lane & 3;
  const uint32_t gshift = grp * 4;
  const uint32_t cap = P.cap;
  unsigned char* base = smem + (size_t)warp * quad_smem_per_warp(cap);
  uint32_t* L = (uint32_t*)base + grp;                      // L[t * 8]
  uint8_t* ord = base + (size_t)cap * 32 + grp;             // ord[t * 8]
  uint32_t* cur = (uint32_t*)(base + (size_t)cap * 40) + grp;  // cur[b * 8], 32 buckets
  uint32_t* rowb = (uint32_t*)(base + (size_t)cap * 40 + 1024) + grp;  // rowb[r * 8], 16 rows: first position
  uint32_t* rowp = rowb + 16 * 8;                                      // rowp[r * 8], 17 entries: candidate prefix
  // The item count may live in device memory (tier sizes of the count pass): the launch is then sized for the
  // worst case and its CTAs stride over the list, most of them leaving at once.
  const uint32_t n_items = P.n_items_dev ? *P.n_items_dev : P.n_items;
  if (n_items < P.min_items) return;  // a short tier runs on k_normals (launched next to this one)
  for (uint32_t item0 = blockIdx.x * (kQuadThreads / 4); item0 < n_items; item0 += gridDim.x * (kQuadThreads / 4)) {
    if (item0 + warp * 8 >= n_items) continue;  // warp-uniform: nothing below synchronises across warps
    // No lane leaves early, so every warp-wide primitive below runs on the full mask: a sub-warp mask that
    // differs between the groups makes the hardware issue the primitive once per group.  Groups past the end of
    // the list redo its last point (same values written twice).
    const uint32_t item = min(item0 + (threadIdx.x >> 2), n_items - 1);
    const uint32_t i = P.list[item];
    const float4 q = g.pts[i];  // 3 <= kn[i] <= cap: the tier list holds nothing else
    const uint32_t jbits = P.jbits, jmask = (1u << jbits) - 1u;
    const float r2 = P.r2;
    // ---- gather: the group's candidate rows flattened into one sequence, four candidates per step; the step
    //      count is the warp's maximum so that the ballot is warp-uniform
    uint32_t n = 0;
    {
      const QueryWindow w = make_window(g, q.x, q.y, q.z, P.rq);
      const int rows = w.rows();
      int rows_max = rows;
  #pragma unroll
      for (int o = 16; o > 0; o >>= 1) rows_max = max(rows_max, __shfl_xor_sync(AG2_FULL, rows_max, o));
      for (int r0 = 0; r0 < rows_max; r0 += 16) {
        __syncwarp();  // the previous table is no longer read
  #pragma unroll
        for (int k = 0; k < 4; k++) {  // each lane trims four of the sixteen rows
          const int r = (int)sub + 4 * k;
          uint32_t b = 0, e = 0;
          if (r0 + r < rows) row_span_trim(g, w, r0 + r, q.x, q.y, q.z, r2 * 1.002f, b, e);
          rowb[r * 8] = b;
          rowp[r * 8] = e - b;
        }
        __syncwarp();
        if (sub == 0) {
          uint32_t run = 0;
  #pragma unroll
          for (int r = 0; r < 16; r++) {
            const uint32_t c = rowp[r * 8];
            rowp[r * 8] = run;
            run += c;
          }
          rowp[16 * 8] = run;
        }
        __syncwarp();
        const uint32_t T = rowp[16 * 8];
        uint32_t T_max = T;
  #pragma unroll
        for (int o = 16; o > 0; o >>= 1) T_max = max(T_max, __shfl_xor_sync(AG2_FULL, T_max, o));
        int row = 0;
        uint32_t row_lo = 0, row_hi = rowp[8], row_b = rowb[0];
        for (uint32_t t0 = 0; t0 < T_max; t0 += 4) {
          const uint32_t t = t0 + sub;
          bool hit = false;
          float d2 = 0.f;
          uint32_t j = 0;
          if (t < T) {
            while (t >= row_hi) {
              row++;
              row_lo = row_hi;
              row_hi = rowp[(row + 1) * 8];
              row_b = rowb[row * 8];
            }
            j = row_b + (t - row_lo);
            const float4 p = g.pts[j];
            d2 = dist2_f(q.x - p.x, q.y - p.y, q.z - p.z);
            hit = d2 < r2;
          }
          const uint32_t bits = (__ballot_sync(AG2_FULL, hit) >> gshift) & 0xFu;
          if (hit) {
            const uint32_t qv = min(P.qmax, (uint32_t)__fmul_rn(d2, P.qscale));  // monotone in d2
            const uint32_t pos = n + __popc(bits & ((1u << sub) - 1u));
            if (pos < cap) L[pos * 8] = (qv << jbits) | j;
          }
          n += __popc(bits);
        }
      }
      n = min(n, cap);  // == count by construction (same walk, same predicate as the count pass)
    }
    // ---- counting sort on the top five bits -> byte permutation grouped by bucket
  #pragma unroll
    for (int k = 0; k < 8; k++) cur[(sub * 8 + k) * 8] = 0;
    __syncwarp();
    for (uint32_t t = sub; t < n; t += 4) atomicAdd(&cur[(L[t * 8] >> 27) * 8], 1u);
    __syncwarp();
    uint32_t lo, hi;  // this lane's buckets sub*8 .. sub*8+7 occupy ord[lo, hi)
    {
      uint32_t c8[8], tot = 0;
  #pragma unroll
      for (int k = 0; k < 8; k++) {
        c8[k] = cur[(sub * 8 + k) * 8];
        tot += c8[k];
      }
      uint32_t incl = tot;
      uint32_t up = __shfl_up_sync(AG2_FULL, incl, 1, 4);
      if (sub >= 1) incl += up;
      up = __shfl_up_sync(AG2_FULL, incl, 2, 4);
      if (sub >= 2) incl += up;
      lo = incl - tot, hi = incl;
      uint32_t run = lo;
  #pragma unroll
      for (int k = 0; k < 8; k++) {
        cur[(sub * 8 + k) * 8] = run;
        run += c8[k];
      }
    }
    __syncwarp();
    for (uint32_t t = sub; t < n; t += 4) {
      const uint32_t pos = atomicAdd(&cur[(L[t * 8] >> 27) * 8], 1u);
      ord[pos * 8] = (uint8_t)t;
    }
    __syncwarp();
    // ---- insertion sort of the lane's segment (elements never leave their bucket); `top` is the
    //      last key of the sorted prefix, so an element already in place costs one comparison
    auto before = [&](uint32_t ka, uint32_t kb) -> bool {  // does ka precede kb in radiusSearch order?
      const uint32_t qa = ka >> jbits, qb = kb >> jbits;
      if (qa != qb) return qa < qb;
      const float4 pa = g.pts[ka & jmask], pb = g.pts[kb & jmask];
      const float da = dist2_f(q.x - pa.x, q.y - pa.y, q.z - pa.z);
      const float db = dist2_f(q.x - pb.x, q.y - pb.y, q.z - pb.z);
      return da < db || (da == db && __float_as_uint(pa.w) < __float_as_uint(pb.w));
    };
    if (lo < hi) {
      uint32_t top = L[(uint32_t)ord[lo * 8] * 8];
      for (uint32_t t = lo + 1; t < hi; t++) {
        const uint32_t o = ord[t * 8];
        const uint32_t k = L[o * 8];
        if (before(top, k)) {
          top = k;
          continue;
        }
        uint32_t u = t;
        do {
          const uint32_t o2 = ord[(u - 1) * 8];
          if (before(L[o2 * 8], k)) break;
          ord[u * 8] = (uint8_t)o2;
          u--;
        } while (u > lo);
        ord[u * 8] = (uint8_t)o;
      }
    }
    __syncwarp();
    // ---- sequential single-pass float moments (accu[0..8] of computeMeanAndCovarianceMatrix):
    //      lane 0: xx xy xz, lane 1: yy yz zz, lane 2: x y z (times the constant 1, exact)
    float a0 = 0.f, a1 = 0.f, a2 = 0.f;
    if (sub < 3) {
  #pragma unroll 4
      for (uint32_t t = 0; t < n; t++) {
        const float4 p = g.pts[L[(uint32_t)ord[t * 8] * 8] & jmask];
        // m = x | y | 1 by lane; products m*(x,y,z), except lane 1: y*y, y*z, z*z (five selects instead of eight)
        const float m = sub == 0 ? p.x : (sub == 1 ? p.y : 1.0f);
        a0 = __fadd_rn(a0, __fmul_rn(m, sub == 1 ? p.y : p.x));
        a1 = __fadd_rn(a1, __fmul_rn(m, sub == 1 ? p.z : p.y));
        a2 = __fadd_rn(a2, __fmul_rn(sub == 1 ? p.z : m, p.z));
      }
    }
    if (sub < 3) {
      float* m = P.mom_out + (size_t)i * kMomWords + sub * 3;
      m[0] = a0, m[1] = a1, m[2] = a2;
    }
    __syncwarp();  // the lists are reused by the next item
  }
}

}  // namespace

// Launches over a tier list.  `n_dev` (device memory) holds the item count when the host does not know it
// yet: the launch is then sized for the worst case (`n_upper` items / full residency) and reads the count itself.

// Four lanes per point at capacity `cap` (< 256: byte permutation); does nothing below `min_items` items.
static int launch_normals_quad(ag2_ctx* c, cudaStream_t st, uint32_t n_upper, const uint32_t* n_dev, uint32_t min_items,
                               uint32_t cap, const uint32_t* list) {
  const size_t smem = quad_smem_per_warp(cap) * (kQuadThreads / 32);
  uint32_t jbits = 1;
  while (jbits < 27 && (1ull << jbits) < (unsigned long long)c->n_finite) jbits++;
  const uint32_t qbits = 32 - jbits;
  QuadArgs P;
  P.g = grid_dev(c);
  P.r2 = c->hc.r2_normals, P.rq = c->hc.rq_normals;
  P.qscale = ldexpf(1.0f, (int)qbits) / c->hc.r2_normals;
  P.qmax = (uint32_t)((1ull << qbits) - 1ull);
  P.jbits = jbits;
  P.kn = c->kn.p;
  P.list = list;
  P.n_items = n_upper, P.cap = cap;
  P.n_items_dev = n_dev;
  P.min_items = std::max(min_items, 1u);
  P.mom_out = c->nrm_mom.p;
  const uint32_t per_cta = kQuadThreads / 4;
  // CTAs stride over the list; a few waves of resident CTAs is all a launch ever needs
  const uint32_t per_sm = (uint32_t)std::max<size_t>(1, std::min<size_t>(16, (226 * 1024) / (smem + 1024)));
  const uint32_t grid = std::max(1u, std::min((n_upper + per_cta - 1) / per_cta, (uint32_t)c->sm_count * per_sm * 4u));
  k_normals_quad<<<grid, kQuadThreads, smem, st>>>(P);
  AG2_LAUNCH_CHECK(c);
  return AG2_OK;
}

// One warp per point over `n_items` items (all owned points when `list` is null, else a list) at capacity `cap`.
static int launch_normals(ag2_ctx* c, cudaStream_t st, uint32_t n_items, const uint32_t* n_dev, uint32_t skip_from,
                          uint32_t cap, const uint32_t* list, uint32_t* big_list, uint32_t* big_count, uint32_t big_cap,
                          uint32_t* batch_counter) {
  const NormalsSmem L = normals_smem(cap);
  const int warps = (int)std::max<size_t>(1, std::min<size_t>(8, (110 * 1024) / L.total));
  const size_t smem = L.total * warps;
  AG2_CUDA(c, cudaFuncSetAttribute(k_normals, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int blocks_per_sm = (int)std::max<size_t>(1, std::min<size_t>(64 / warps, (226 * 1024) / (smem + 1024)));
  const uint32_t resident_warps = (uint32_t)(c->sm_count * blocks_per_sm * warps);
  // batch size: 32 points per warp when there is plenty of work, fewer when a short list of
  // dense points must still be spread over the whole machine (the kernel applies the same rule to a device-side count)
  uint32_t bsz = 32;
  while (bsz > 1 && (n_items + bsz - 1) / bsz < 4 * resident_warps) bsz >>= 1;
  const uint32_t n_batches = (n_items + bsz - 1) / bsz;
  uint32_t grid = (uint32_t)std::max<size_t>(
      1, std::min<size_t>(((size_t)n_batches + warps - 1) / warps, (size_t)c->sm_count * blocks_per_sm));
  if (n_dev)  // count unknown here: full residency, but never more warps than there can be items
    grid = (uint32_t)std::max<size_t>(1, std::min<size_t>(((size_t)n_items + warps - 1) / warps, (size_t)c->sm_count * blocks_per_sm));
  NormalsArgs P;
  P.g = grid_dev(c);
  P.r2 = c->hc.r2_normals, P.rq = c->hc.rq_normals;
  P.bscale = (float)kNB / c->hc.r2_normals;
  P.kn = c->kn.p;
  P.n_items = n_items, P.cap = cap, P.bsz = bsz;
  P.n_items_dev = n_dev, P.skip_from = skip_from;
  P.shard_rank = (uint32_t)c->nrm_rank, P.shard_world = (uint32_t)c->nrm_world;
  P.list = list;
  P.big_list = big_list, P.big_count = big_count, P.big_cap = big_cap;
  P.batch_counter = batch_counter;
  P.mom_out = c->nrm_mom.p;
  P.nrm_out = c->nrm.p;
  k_normals<<<grid, warps * 32, smem, st>>>(P);
  AG2_LAUNCH_CHECK(c);
  return AG2_OK;
}

// One CTA per point over a tier list at capacity `cap` (dense tiers).
static int launch_normals_cta(ag2_ctx* c, cudaStream_t st, uint32_t n_upper, const uint32_t* n_dev, uint32_t cap,
                              const uint32_t* list, uint32_t* counter, bool unlikely) {
  const CtaSmem L = cta_smem(cap);
  const int per_sm = (int)std::max<size_t>(1, std::min<size_t>(16, (226 * 1024) / (L.total + 1024)));
  NormalsArgs P;
  P.g = grid_dev(c);
  P.r2 = c->hc.r2_normals, P.rq = c->hc.rq_normals;
  P.bscale = 0.f;
  P.kn = c->kn.p;
  P.n_items = n_upper, P.cap = cap, P.bsz = 1;
  P.n_items_dev = n_dev, P.skip_from = 0;
  P.shard_rank = 0, P.shard_world = 1;
  P.list = list;
  P.big_list = nullptr, P.big_count = nullptr, P.big_cap = 0;
  P.batch_counter = counter;
  P.mom_out = c->nrm_mom.p;
  P.nrm_out = c->nrm.p;
  // CTAs claim points from a counter, so any grid is correct.  A tier that the previous cloud's densest
  // neighbourhood says will be empty gets a token grid: its CTAs need most of an SM's shared memory each and
  // would otherwise queue behind the launches that do have work.
  uint32_t grid = (uint32_t)std::max<size_t>(1, std::min<size_t>(n_upper, (size_t)c->sm_count * per_sm));
  if (unlikely) grid = std::min(grid, 4u);
  k_normals_cta<<<grid, kCtaThreads, L.total, st>>>(P);
  AG2_LAUNCH_CHECK(c);
  return AG2_OK;
}

// number of points rank `rank` of `world` owns among n cell-sorted points (see shard_point)
uint32_t normals_shard_points(uint32_t n, int rank, int world) {
  const uint32_t G = (n + 31) / 32;
  if (G <= (uint32_t)rank) return 0;
  const uint32_t owned = (G - (uint32_t)rank + (uint32_t)world - 1) / (uint32_t)world;
  const bool has_tail = (G - 1) % (uint32_t)world == (uint32_t)rank;
  return owned * 32 - (has_tail ? G * 32 - n : 0);
}

namespace {
// owned normals -> compact send buffer, and the gathered buffers of all ranks -> nrm (w = camera bits kept)
__global__ void k_pack_shard(const float4* __restrict__ nrm, uint32_t n_owned, uint32_t rank, uint32_t world,
                             float4* __restrict__ out) {
  uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n_owned) out[t] = nrm[shard_point(t, rank, world)];
}
__global__ void k_unpack_shards(const float4* __restrict__ all, uint32_t cap_points, uint32_t world, uint32_t n,
                                float4* __restrict__ nrm) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t g = i >> 5, r = g % world, t = ((g / world) << 5) | (i & 31u);
  const float4 v = all[(size_t)r * cap_points + t];
  nrm[i] = make_float4(v.x, v.y, v.z, nrm[i].w);
}
}  // namespace

int normals_shard_compute(ag2_ctx* c, int rank, int world, float* d_out, size_t cap_points, size_t* n_owned) {
  AG2_TRY(grid_build(c));
  if (c->normals_given) return AG2_E_STATE;  // normals came with the cloud: nothing to shard
  const uint32_t n = (uint32_t)c->n_finite;
  const uint32_t mine = normals_shard_points(n, rank, world);
  if (n_owned) *n_owned = mine;
  if (mine > cap_points) return AG2_E_CAPACITY;
  c->nrm_rank = rank, c->nrm_world = world;
  c->have_normals = false;
  int rc = normals_compute(c);
  c->nrm_rank = 0, c->nrm_world = 1;
  c->have_normals = false;  // only this rank's granules are filled in
  if (rc != AG2_OK) return rc;
  if (mine) {
    k_pack_shard<<<(mine + 255) / 256, 256, 0, c->stream>>>(c->nrm.p, mine, (uint32_t)rank, (uint32_t)world,
                                                           reinterpret_cast<float4*>(d_out));
    AG2_LAUNCH_CHECK(c);
  }
  AG2_CUDA(c, cudaStreamSynchronize(c->stream));
  return AG2_OK;
}

int normals_shard_assemble(ag2_ctx* c, const float* d_all, size_t cap_points, int world) {
  AG2_TRY(grid_build(c));
  const uint32_t n = (uint32_t)c->n_finite;
  if ((size_t)normals_shard_points(n, 0, world) > cap_points) return AG2_E_CAPACITY;
  if (n) {
    k_unpack_shards<<<(n + 255) / 256, 256, 0, c->stream>>>(reinterpret_cast<const float4*>(d_all), (uint32_t)cap_points,
                                                           (uint32_t)world, n, c->nrm.p);
    AG2_LAUNCH_CHECK(c);
  }
  c->have_normals = true;
  return AG2_OK;
}

// k_eigen over the owned points: normals into this context's sorted array, or into every rank's (multi-GPU).
static int launch_eigen(ag2_ctx* c) {
  const uint32_t own = normals_shard_points((uint32_t)c->n_finite, c->nrm_rank, c->nrm_world);
  NrmDst dst;
  dst.n = 0;
  for (int r = 0; r < c->nrm_peers && r < 8; r++) dst.p[dst.n++] = c->nrm_peer[r];
  if (dst.n == 0) dst.p[dst.n++] = c->nrm.p;
  for (int r = dst.n; r < 8; r++) dst.p[r] = nullptr;
  k_eigen<<<std::max(1u, (((own + 31) & ~31u) + 255) / 256), 256, 0, c->stream>>>(
      grid_dev(c), c->kn.p, c->nrm_mom.p, c->nrm.p, dst, (uint32_t)c->nrm_rank, (uint32_t)c->nrm_world);
  AG2_LAUNCH_CHECK(c);
  return AG2_OK;
}

// Tiered path: count pass with tier lists, then one launch per tier on the auxiliary streams.  The host never
// learns the tier sizes: every launch is sized for the worst case and reads its item count from device memory
// (an empty tier costs a launch that exits at once), so the phase has no host round trip and the launches are
// queued while the count pass still runs.
static int normals_moments_tiered(ag2_ctx* c) {
  cudaStream_t st = c->stream;
  GridDev g = grid_dev(c);
  const uint32_t n = (uint32_t)c->n_finite;
  const uint32_t owned = normals_shard_points(n, c->nrm_rank, c->nrm_world);
  // The four sparsest tiers (<= 96 neighbours) run on k_normals_quad, the rest on k_normals: measured
  // best on c2 and c3.  AG2_NORMALS_QUAD_TIERS = 0..8 moves the boundary (tests).
  uint32_t quad_tiers = 4;
  if (const char* e = getenv("AG2_NORMALS_QUAD_TIERS")) quad_tiers = (uint32_t)std::min(kQuadTiers, std::max(0, atoi(e)));
  // A lone k_normals_quad launch cannot finish faster than one group's chain of dependent accesses
  // (~130 us); a short list is better off on k_normals, whose floor is a tenth of that.
  // dense tiers (capacity 512 and up by default) run one CTA per point; AG2_NORMALS_CTA_FROM = first such tier
  int cta_from = 10;
  if (const char* e = getenv("AG2_NORMALS_CTA_FROM")) cta_from = std::max(0, atoi(e));
  uint32_t quad_min = 8192;
  if (const char* e = getenv("AG2_NORMALS_QUAD_MIN")) quad_min = (uint32_t)std::max(0, atoi(e));
  quad_min = std::max(quad_min, 1u);
  AG2_CUDA(c, c->kn.reserve(n));
  AG2_CUDA(c, c->tier_list.reserve((size_t)kAllTiers * std::max(owned, 1u)));
  AG2_CUDA(c, c->tier_cnt.reserve(kTierCntWords));
  AG2_CUDA(c, c->nrm_mom.reserve((size_t)kMomWords * std::max(n, 1u)));
  AG2_CUDA(c, cudaMemsetAsync(c->scalars.p + 8, 0, 16, st));  // [8] max, [10..11] sum
  AG2_CUDA(c, cudaMemsetAsync(c->tier_cnt.p, 0, kTierCntWords * 4, st));
  {
    const uint32_t owned32 = (owned + 31) & ~31u;
    unsigned long long* sum = (unsigned long long*)(c->scalars.p + 10);
    if (owned32 >= 200000u)
      k_count_tiers<1><<<std::max(1u, (owned32 + 127) / 128), 128, 0, st>>>(g, c->hc.r2_normals, c->hc.rq_normals, c->kn.p,
                                                                           c->scalars.p + 8, sum, c->tier_cnt.p, c->tier_list.p,
                                                                           owned, c->nrm.p, (uint32_t)c->nrm_rank,
                                                                           (uint32_t)c->nrm_world);
    else
      k_count_tiers<4><<<std::max(1u, owned32 / 32), 128, 0, st>>>(g, c->hc.r2_normals, c->hc.rq_normals, c->kn.p,
                                                                  c->scalars.p + 8, sum, c->tier_cnt.p, c->tier_list.p, owned,
                                                                  c->nrm.p, (uint32_t)c->nrm_rank, (uint32_t)c->nrm_world);
  }
  AG2_LAUNCH_CHECK(c);
  // statistics and the capacity check travel to pinned memory behind the count pass; nobody waits for them here
  AG2_CUDA(c, cudaMemcpyAsync(c->h_pin, c->scalars.p + 8, 16, cudaMemcpyDeviceToHost, st));
  AG2_CUDA(c, cudaMemcpyAsync(c->h_pin + 4, c->tier_cnt.p + 32, 4, cudaMemcpyDeviceToHost, st));
  c->pend_normals = true;
  if (!c->normals_attr_set) {  // per device, not per process: a group runs one ctx per GPU in one process
    AG2_CUDA(c, cudaFuncSetAttribute(k_normals_quad, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)(quad_smem_per_warp(kTierCapHost[kQuadTiers - 1]) * (kQuadThreads / 32))));
    AG2_CUDA(c, cudaFuncSetAttribute(k_normals_cta, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)cta_smem(kTierCapHost[kAllTiers - 1]).total));
    c->normals_attr_set = true;
  }
  AG2_CUDA(c, cudaEventRecord(c->ev_fork, st));
  // AG2_NORMALS_STREAMS (1..12): how many launches may run side by side; every launch of a stream waits for the one before
  int n_aux = 4;
  if (const char* e = getenv("AG2_NORMALS_STREAMS")) n_aux = std::min((int)ag2_ctx::kAux, std::max(1, atoi(e)));
  for (int i = 0; i < n_aux; i++) AG2_CUDA(c, cudaStreamWaitEvent(c->aux[i], c->ev_fork, 0));
  int used = 0;
  // dense tiers first: their launches run longest
  for (int k = 0; k < kAllTiers; k++) {
    const int tier = kAllTiers - 1 - k;
    // a tier above c's densest possible neighbourhood cannot hold a point: a cloud of n points has no
    // neighbourhood larger than n (keeps the launch count of small scenes down)
    if (tier > 0 && kTierCapHost[tier - 1] >= n) continue;
    const uint32_t* list = c->tier_list.p + (size_t)tier * owned;
    const uint32_t* n_dev = c->tier_cnt.p + tier;
    const uint32_t cap = kTierCapHost[tier];
    if (tier >= cta_from || tier >= 12) {  // above 1024 neighbours only the CTA kernel has the shared memory
      const bool unlikely = tier >= 12 && kTierCapHost[tier - 1] >= 2u * std::max(c->max_kn_hint, 512u);
      AG2_TRY(launch_normals_cta(c, c->aux[used++ % n_aux], owned, n_dev, cap, list, c->tier_cnt.p + 16 + tier, unlikely));
      continue;
    }
    const bool quad = tier < (int)quad_tiers;
    if (quad) AG2_TRY(launch_normals_quad(c, c->aux[used++ % n_aux], owned, n_dev, quad_min, cap, list));
    // the warp-per-point launch of a quad tier only runs when the tier is too short for k_normals_quad
    AG2_TRY(launch_normals(c, c->aux[used++ % n_aux], owned, n_dev, quad ? quad_min : 0u, cap, list, nullptr, nullptr, 0,
                           c->tier_cnt.p + 16 + tier));
  }
  for (int i = 0; i < std::min(used, n_aux); i++) {
    AG2_CUDA(c, cudaEventRecord(c->ev_join[i], c->aux[i]));
    AG2_CUDA(c, cudaStreamWaitEvent(st, c->ev_join[i], 0));
  }
  return AG2_OK;
}

// Chained capacity levels on the warp-per-point kernel (the path before the tiers existed; test /
// measurement hook AG2_NORMALS_WARP=1, and clouds of more than 2^27 points).
static int normals_moments_chained(ag2_ctx* c) {
  cudaStream_t st = c->stream;
  GridDev g = grid_dev(c);
  const uint32_t n = (uint32_t)c->n_finite;
  AG2_CUDA(c, c->kn.reserve(n));
  AG2_CUDA(c, c->nrm_mom.reserve((size_t)kMomWords * std::max(n, 1u)));
  AG2_CUDA(c, cudaMemsetAsync(c->scalars.p + 8, 0, 16, st));    // [8] max, [10..11] sum
  AG2_CUDA(c, cudaMemsetAsync(c->scalars.p + 32, 0, 32 * 4, st));  // [32..39] tiers, [40..47] list counts, [48..55] batch counters
  k_count_neighbors<<<(n + 127) / 128, 128, 0, st>>>(g, c->hc.r2_normals, c->hc.rq_normals, c->kn.p,
                                                      c->scalars.p + 8,
                                                      (unsigned long long*)(c->scalars.p + 10), c->scalars.p + 32,
                                                      (uint32_t)c->nrm_rank, (uint32_t)c->nrm_world);
  AG2_LAUNCH_CHECK(c);
  uint32_t res[4], hist[kTiers];
  AG2_CUDA(c, cudaMemcpyAsync(res, c->scalars.p + 8, 16, cudaMemcpyDeviceToHost, st));
  AG2_CUDA(c, cudaMemcpyAsync(hist, c->scalars.p + 32, sizeof(hist), cudaMemcpyDeviceToHost, st));
  AG2_CUDA(c, cudaStreamSynchronize(st));
  const uint32_t max_kn = res[0];
  unsigned long long sum;
  memcpy(&sum, &res[2], 8);
  c->st.sum_neighbors_normals = (int64_t)sum;
  const uint32_t cap_max = (std::max(max_kn, 32u) + 31u) & ~31u;
  if (normals_smem(cap_max).total > 220 * 1024) {
    c->err = "normals: a point has more than ~13600 neighbours within normals_radius";
    return AG2_E_CAPACITY;
  }
  // Capacity levels: each launch runs at the smallest tier that leaves at most 25 % of its
  // items for the next one; hist[] gives the exact size of every leftover list.
  static const uint32_t tiers[kTiers] = {64, 96, 128, 192, 256, 384, 512, 1024};
  // test hook: the first level capacity can be forced down so that small scenes exercise the
  // multi-level path (AG2_NORMALS_FIRST_TIER = index into tiers[])
  int forced_tier = -1;
  if (const char* e = getenv("AG2_NORMALS_FIRST_TIER")) forced_tier = atoi(e);
  uint32_t remaining = normals_shard_points(n, c->nrm_rank, c->nrm_world), level = 0;
  const uint32_t* list = nullptr;
  int tier = 0;
  while (remaining > 0 && level < 8) {
    uint32_t cap = cap_max, left = 0;
    for (; tier < kTiers; tier++)
      if (tiers[tier] < cap_max && (hist[tier] <= remaining / 4 || (level == 0 && tier == forced_tier))) {
        cap = tiers[tier];
        left = hist[tier];
        tier++;
        break;
      }
    DevBuf<uint32_t>& out = c->big_list[level & 1];
    if (left) AG2_CUDA(c, out.reserve(left));
    AG2_TRY(launch_normals(c, st, remaining, nullptr, 0, cap, list, out.p, c->scalars.p + 40 + level, left,
                           c->scalars.p + 48 + level));
    list = out.p;
    remaining = left;
    level++;
  }
  return AG2_OK;
}

int normals_moments(ag2_ctx* c) {
  AG2_TRY(grid_build(c));
  if (c->have_normals) return AG2_OK;
  phase_begin(c, 1);
  const char* warp_only = getenv("AG2_NORMALS_WARP");
  if (!(warp_only && atoi(warp_only) != 0) && c->n_finite <= (1u << 27)) return normals_moments_tiered(c);
  return normals_moments_chained(c);
}

int normals_finish(ag2_ctx* c) {
  if (c->have_normals) return AG2_OK;
  AG2_TRY(launch_eigen(c));
  c->have_normals = true;
  phase_end(c, 1);
  return AG2_OK;
}

int normals_compute(ag2_ctx* c) {
  AG2_TRY(normals_moments(c));
  return normals_finish(c);
}

}  // namespace ag2
