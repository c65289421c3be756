// Device helpers shared by the kernels of libag2.so (sm_100a).
#ifndef AG2_DEV_COMMON_CUH_
#define AG2_DEV_COMMON_CUH_

#include "ag2_internal.h"

namespace ag2 {

#define AG2_FULL 0xffffffffu

__device__ __forceinline__ int cell_coord(float v, float o, float inv_h, int n) {
  // monotone non-decreasing in v; clamped so that any point maps to a valid cell
  int c = (int)floorf(__fmul_rn(__fsub_rn(v, o), inv_h));
  return min(max(c, 0), n - 1);
}

__device__ __forceinline__ uint32_t cell_id(const GridDev& g, int cx, int cy, int cz) {
  return (uint32_t)cx + (uint32_t)g.nx * ((uint32_t)cy + (uint32_t)g.ny * (uint32_t)cz);
}

// FLANN L2_Simple restated: float, sequential over x, y, z, no FMA.
__device__ __forceinline__ float dist2_f(float dx, float dy, float dz) {
  return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
}

// Cell-row query window of a fixed-radius search around q: rows (y,z) in
// [y0,y1]x[z0,z1], each row is the contiguous span of cells x0..x1.
struct QueryWindow {
  int x0, x1, y0, y1, z0, z1;
  __device__ __forceinline__ int rows() const { return (y1 - y0 + 1) * (z1 - z0 + 1); }
};

__device__ __forceinline__ QueryWindow make_window(const GridDev& g, float qx, float qy, float qz,
                                                  float rq) {
  QueryWindow w;
  w.x0 = cell_coord(qx - rq, g.ox, g.inv_h, g.nx);
  w.x1 = cell_coord(qx + rq, g.ox, g.inv_h, g.nx);
  w.y0 = cell_coord(qy - rq, g.oy, g.inv_h, g.ny);
  w.y1 = cell_coord(qy + rq, g.oy, g.inv_h, g.ny);
  w.z0 = cell_coord(qz - rq, g.oz, g.inv_h, g.nz);
  w.z1 = cell_coord(qz + rq, g.oz, g.inv_h, g.nz);
  return w;
}

// Span [begin, end) of sorted positions covered by row `row` of the window.
__device__ __forceinline__ void row_span(const GridDev& g, const QueryWindow& w, int row,
                                         uint32_t& begin, uint32_t& end) {
  int ny = w.y1 - w.y0 + 1;
  int cz = w.z0 + row / ny;
  int cy = w.y0 + row % ny;
  uint32_t c0 = cell_id(g, w.x0, cy, cz);
  begin = __ldg(&g.cell_start[c0]);
  end = __ldg(&g.cell_start[c0 + (uint32_t)(w.x1 - w.x0) + 1]);
}

// Same, with the cells that cannot hold a point within the radius trimmed away: the whole row when
// its (y, z) slab is farther than the radius from q, and the end cells of the row whose near face
// is farther than the remaining chord.  r2m = radius^2 with a small margin (the caller's exact
// test decides membership; the margin only has to cover the float rounding of the cell faces,
// ~1e-7 m).  Cells on the border of the grid are unbounded outwards (cell_coord clamps).
__device__ __forceinline__ void row_span_trim(const GridDev& g, const QueryWindow& w, int row, float qx, float qy,
                                              float qz, float r2m, uint32_t& begin, uint32_t& end) {
  const int ny = w.y1 - w.y0 + 1;
  const int cz = w.z0 + row / ny;
  const int cy = w.y0 + row % ny;
  const float h = g.h;
  const float ylo = g.oy + (float)cy * h, zlo = g.oz + (float)cz * h;
  float dy = 0.f, dz = 0.f;
  if (qy < ylo && cy > 0) dy = ylo - qy;
  if (qy > ylo + h && cy < g.ny - 1) dy = qy - (ylo + h);
  if (qz < zlo && cz > 0) dz = zlo - qz;
  if (qz > zlo + h && cz < g.nz - 1) dz = qz - (zlo + h);
  const float rem = r2m - dy * dy - dz * dz;
  if (rem < 0.f) {
    begin = end = 0;
    return;
  }
  int x0 = w.x0, x1 = w.x1;
  while (x0 < x1) {  // leading cells entirely to the left of the chord
    const float d = qx - (g.ox + (float)(x0 + 1) * h);
    if (d > 0.f && d * d > rem) x0++; else break;
  }
  while (x1 > x0) {  // trailing cells entirely to the right
    const float d = (g.ox + (float)x1 * h) - qx;
    if (d > 0.f && d * d > rem) x1--; else break;
  }
  const uint32_t c0 = cell_id(g, x0, cy, cz);
  begin = __ldg(&g.cell_start[c0]);
  end = __ldg(&g.cell_start[c0 + (uint32_t)(x1 - x0) + 1]);
}

__device__ __forceinline__ uint32_t next_pow2(uint32_t v) {
  uint32_t p = 1;
  while (p < v) p <<= 1;
  return p;
}

// 64-bit sort key: (d2 bits, input-order index) == FLANN DistanceIndex order.
__device__ __forceinline__ uint64_t nn_key(float d2, uint32_t orig) {
  return ((uint64_t)__float_as_uint(d2) << 32) | (uint64_t)orig;
}

// Bitonic sort of n_pad (power of two) key/value pairs in shared (or global) memory
// by `nthreads` cooperating threads; `sync` is __syncwarp or __syncthreads.
template <typename Sync>
__device__ __forceinline__ void bitonic_sort_kv(uint64_t* keys, uint32_t* vals, uint32_t n_pad,
                                                uint32_t tid, uint32_t nthreads, Sync sync) {
  for (uint32_t k = 2; k <= n_pad; k <<= 1) {
    for (uint32_t j = k >> 1; j > 0; j >>= 1) {
      for (uint32_t t = tid; t < (n_pad >> 1); t += nthreads) {
        // index of the lower element of the t-th compare-exchange pair
        uint32_t i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
        uint32_t l = i | j;
        bool up = ((i & k) == 0);
        uint64_t a = keys[i], b = keys[l];
        if ((a > b) == up) {
          keys[i] = b;
          keys[l] = a;
          uint32_t va = vals[i];
          vals[i] = vals[l];
          vals[l] = va;
        }
      }
      sync();
    }
  }
}

template <typename Sync>
__device__ __forceinline__ void bitonic_sort_u32(uint32_t* keys, uint32_t n_pad, uint32_t tid,
                                                 uint32_t nthreads, Sync sync) {
  for (uint32_t k = 2; k <= n_pad; k <<= 1) {
    for (uint32_t j = k >> 1; j > 0; j >>= 1) {
      for (uint32_t t = tid; t < (n_pad >> 1); t += nthreads) {
        uint32_t i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
        uint32_t l = i | j;
        bool up = ((i & k) == 0);
        uint32_t a = keys[i], b = keys[l];
        if ((a > b) == up) {
          keys[i] = b;
          keys[l] = a;
        }
      }
      sync();
    }
  }
}

struct WarpSync {
  __device__ __forceinline__ void operator()() const { __syncwarp(); }
};
struct BlockSync {
  __device__ __forceinline__ void operator()() const { __syncthreads(); }
};

// splitmix64-based replacement of rand() % n (hand_search.cpp:130); documented deviation D1.
__device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
__device__ __forceinline__ uint32_t draw_index(uint64_t seed, int64_t sample_id, int j, uint32_t n) {
  uint64_t h = splitmix64(seed ^ splitmix64(((uint64_t)sample_id << 8) ^ (uint64_t)j));
  return (uint32_t)(h % (uint64_t)n);
}

__device__ __forceinline__ double warp_min_d(double v) {
  for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(AG2_FULL, v, o));
  return v;
}
__device__ __forceinline__ double warp_max_d(double v) {
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(AG2_FULL, v, o));
  return v;
}

// ordered-int encoding of floats for atomicMin/Max
__device__ __forceinline__ int float_ordered(float f) {
  int i = __float_as_int(f);
  return i >= 0 ? i : i ^ 0x7fffffff;
}
__host__ __device__ __forceinline__ float ordered_float(int i) {
  int j = i >= 0 ? i : i ^ 0x7fffffff;
#ifdef __CUDA_ARCH__
  return __int_as_float(j);
#else
  float f;
  memcpy(&f, &j, 4);
  return f;
#endif
}

}  // namespace ag2
#endif
