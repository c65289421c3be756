// K4: surface normals, replacing pcl::NormalEstimationOMP (reference
// hand_search.cpp:83-94).  One warp per point:
//   1. gather the r = normals_radius neighbourhood from the cell rows (float d2,
//      strict <, FLANN L2_Simple order) into shared memory as (d2, id) keys,
//   2. bitonic-sort the keys: PCL feeds computeMeanAndCovarianceMatrix the
//      neighbours in radiusSearch order (ascending (d2, id)), and the single-pass
//      FLOAT moment sums are order dependent at the 1e-3 rad level, so the order
//      is part of the semantics,
//   3. nine lanes run the nine sequential float accumulators in that order,
//   4. lane 0 solves the 3x3 eigenproblem exactly as pcl::eigen33 does and flips
//      the normal toward the viewpoint (0,0,0).
// Compiled with -fmad=false; every float op below is a separate IEEE operation.
#include <math.h>

#include <algorithm>

#include "dev_common.cuh"

namespace ag2 {
namespace {

// thread per point: number of neighbours within the radius (sizes the sort buffers)
__global__ void k_count_neighbors(GridDev g, float r2, float rq, uint32_t* __restrict__ kn,
                                  uint32_t* __restrict__ max_out, unsigned long long* __restrict__ sum_out) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t cnt = 0;
  if (i < (uint32_t)g.n) {
    float4 q = g.pts[i];
    QueryWindow w = make_window(g, q.x, q.y, q.z, rq);
    int rows = w.rows();
    for (int row = 0; row < rows; row++) {
      uint32_t b, e;
      row_span(g, w, row, b, e);
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

// One warp per point.  smem per warp: cap * (8 + 4) bytes.
__global__ void k_normals(GridDev g, float r2, float rq, const uint32_t* __restrict__ kn,
                          const uint32_t* __restrict__ list, uint32_t n_items, uint32_t cap,
                          float4* __restrict__ nrm_out) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int warps = blockDim.x >> 5;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint64_t* keys = (uint64_t*)smem + (size_t)warp * cap;
  uint32_t* vals = (uint32_t*)((uint64_t*)smem + (size_t)warps * cap) + (size_t)warp * cap;
  for (uint32_t item = blockIdx.x * warps + warp; item < n_items; item += gridDim.x * warps) {
    uint32_t i = list ? list[item] : item;
    float4 q = g.pts[i];
    uint32_t count = kn[i];
    float4 old = nrm_out[i];
    if (count < 3) {  // computePointNormal: fewer than 3 neighbours -> NaN
      if (lane == 0) {
        float nanv = __int_as_float(0x7fc00000);
        nrm_out[i] = make_float4(nanv, nanv, nanv, old.w);
      }
      continue;
    }
    uint32_t n_pad = next_pow2(count);
    // 1. gather (ordered append per warp)
    QueryWindow w = make_window(g, q.x, q.y, q.z, rq);
    int rows = w.rows();
    uint32_t filled = 0;
    for (int row = 0; row < rows; row++) {
      uint32_t b, e;
      row_span(g, w, row, b, e);
      for (uint32_t base = b; base < e; base += 32) {
        uint32_t j = base + lane;
        bool hit = false;
        float d2 = 0.f;
        float4 p = make_float4(0, 0, 0, 0);
        if (j < e) {
          p = g.pts[j];
          d2 = dist2_f(q.x - p.x, q.y - p.y, q.z - p.z);
          hit = d2 < r2;
        }
        uint32_t ballot = __ballot_sync(AG2_FULL, hit);
        if (hit) {
          uint32_t slot = filled + __popc(ballot & ((1u << lane) - 1));
          keys[slot] = nn_key(d2, __float_as_uint(p.w));
          vals[slot] = j;
        }
        filled += __popc(ballot);
      }
    }
    for (uint32_t t = filled + lane; t < n_pad; t += 32) {
      keys[t] = ~0ull;
      vals[t] = 0;
    }
    __syncwarp();
    // 2. sort by (d2, id)
    bitonic_sort_kv(keys, vals, n_pad, (uint32_t)lane, 32u, WarpSync());
    // 3. replace each entry by its coordinates (entry-local, no aliasing)
    for (uint32_t t = lane; t < count; t += 32) {
      float4 p = g.pts[vals[t]];
      float2 xy = make_float2(p.x, p.y);
      *reinterpret_cast<float2*>(&keys[t]) = xy;
      vals[t] = __float_as_uint(p.z);
    }
    __syncwarp();
    // sequential single-pass float moments (computeMeanAndCovarianceMatrix)
    float acc = 0.f;
    if (lane < 9) {
      // operand selectors: 0 = x, 1 = y, 2 = z, 3 = constant 1
      const int sa = lane < 3 ? 0 : (lane < 5 ? 1 : (lane == 5 ? 2 : lane - 6));
      const int sb = lane < 3 ? lane : (lane < 5 ? lane - 2 : (lane == 5 ? 2 : 3));
      for (uint32_t t = 0; t < count; t++) {
        float2 xy = *reinterpret_cast<const float2*>(&keys[t]);
        float z = __uint_as_float(vals[t]);
        float va = sa == 0 ? xy.x : (sa == 1 ? xy.y : z);
        float vb = sb == 0 ? xy.x : (sb == 1 ? xy.y : (sb == 2 ? z : 1.0f));
        // accu[6..8] += p.x : multiplying by 1.0f is exact
        acc = __fadd_rn(acc, __fmul_rn(va, vb));
      }
    }
    float a[9];
#pragma unroll
    for (int k = 0; k < 9; k++) a[k] = __shfl_sync(AG2_FULL, acc, k);
    if (lane == 0) {
      float inv = 1.0f / (float)count;  // Eigen 3.2 operator/=: multiply by the reciprocal
      for (int k = 0; k < 9; k++) a[k] *= inv;
      float n[3];
      solve_normal(a, q.x, q.y, q.z, n);
      nrm_out[i] = make_float4(n[0], n[1], n[2], old.w);
    }
    __syncwarp();
  }
}

}  // namespace

int normals_compute(ag2_ctx* c) {
  AG2_TRY(grid_build(c));
  if (c->have_normals) return AG2_OK;
  phase_begin(c, 1);
  cudaStream_t st = c->stream;
  GridDev g = grid_dev(c);
  const uint32_t n = (uint32_t)c->n_finite;
  AG2_CUDA(c, c->kn.reserve(n));
  AG2_CUDA(c, cudaMemsetAsync(c->scalars.p + 8, 0, 16, st));  // [8] max, [10..11] sum
  k_count_neighbors<<<(n + 127) / 128, 128, 0, st>>>(g, c->hc.r2_normals, c->hc.rq_normals, c->kn.p,
                                                      c->scalars.p + 8,
                                                      (unsigned long long*)(c->scalars.p + 10));
  AG2_LAUNCH_CHECK(c);
  uint32_t res[4];
  AG2_CUDA(c, cudaMemcpyAsync(res, c->scalars.p + 8, 16, cudaMemcpyDeviceToHost, st));
  AG2_CUDA(c, cudaStreamSynchronize(st));
  uint32_t max_kn = res[0];
  unsigned long long sum;
  memcpy(&sum, &res[2], 8);
  c->st.sum_neighbors_normals = (int64_t)sum;
  uint32_t cap = 32;
  while (cap < max_kn) cap <<= 1;
  size_t per_warp = (size_t)cap * 12;
  if (per_warp > 200 * 1024) {
    c->err = "normals: a point has more than 16384 neighbours within normals_radius";
    return AG2_E_CAPACITY;
  }
  int warps = (int)std::max<size_t>(1, std::min<size_t>(8, (96 * 1024) / per_warp));
  size_t smem = per_warp * warps;
  AG2_CUDA(c, cudaFuncSetAttribute(k_normals, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int blocks_per_sm = (int)std::max<size_t>(1, std::min<size_t>(16, (200 * 1024) / smem));
  uint32_t grid = (uint32_t)std::min<size_t>(((size_t)n + warps - 1) / warps,
                                             (size_t)c->sm_count * blocks_per_sm);
  k_normals<<<grid, warps * 32, smem, st>>>(g, c->hc.r2_normals, c->hc.rq_normals, c->kn.p, nullptr, n, cap,
                                            c->nrm.p);
  AG2_LAUNCH_CHECK(c);
  c->have_normals = true;
  phase_end(c, 1);
  return AG2_OK;
}

}  // namespace ag2
