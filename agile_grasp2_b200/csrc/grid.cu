// K1-K3: uniform-grid cell index replacing pcl::KdTreeFLANN::setInputCloud
// (reference hand_search.cpp:11-12 and the second tree inside NormalEstimationOMP,
// :87-90).  HBM-bound byte work: 50 B/point algorithmic.
//
//   k_bbox      : bounding box of the finite points (+ min z for S13)
//   k_cell_count: cell id per point, histogram
//   k_scan_*    : exclusive scan of the histogram (3 launches)
//   k_scatter   : counting-sort scatter (atomic slot within the cell)
//   k_finalize  : deterministic in-cell order (ascending input index) and the
//                 SoA->float4 repack: pts = (x,y,z,bits(input index)),
//                 nrm.w = camera mask, sorted_of_orig = inverse permutation.
#include <math.h>
#include <string.h>

#include <algorithm>

#include "dev_common.cuh"

namespace ag2 {

namespace {

constexpr int kThreads = 256;

__global__ void k_bbox(const float* __restrict__ xyz, size_t n, int* __restrict__ out /*6 + count*/) {
  float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
  unsigned cnt = 0;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    float x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
    if (isfinite(x) && isfinite(y) && isfinite(z)) {
      mn[0] = fminf(mn[0], x), mx[0] = fmaxf(mx[0], x);
      mn[1] = fminf(mn[1], y), mx[1] = fmaxf(mx[1], y);
      mn[2] = fminf(mn[2], z), mx[2] = fmaxf(mx[2], z);
      cnt++;
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    for (int d = 0; d < 3; d++) {
      mn[d] = fminf(mn[d], __shfl_xor_sync(AG2_FULL, mn[d], o));
      mx[d] = fmaxf(mx[d], __shfl_xor_sync(AG2_FULL, mx[d], o));
    }
    cnt += __shfl_xor_sync(AG2_FULL, cnt, o);
  }
  // one set of atomics per CTA (7 addresses shared by the whole grid)
  __shared__ float s_mn[kThreads / 32][3], s_mx[kThreads / 32][3];
  __shared__ unsigned s_cnt[kThreads / 32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) {
    for (int d = 0; d < 3; d++) s_mn[warp][d] = mn[d], s_mx[warp][d] = mx[d];
    s_cnt[warp] = cnt;
  }
  __syncthreads();
  if (threadIdx.x < 3) {
    const int d = threadIdx.x;
    float a = s_mn[0][d], b = s_mx[0][d];
    for (int k = 1; k < kThreads / 32; k++) a = fminf(a, s_mn[k][d]), b = fmaxf(b, s_mx[k][d]);
    atomicMin(&out[d], float_ordered(a));
    atomicMax(&out[3 + d], float_ordered(b));
  } else if (threadIdx.x == 3) {
    unsigned t = 0;
    for (int k = 0; k < kThreads / 32; k++) t += s_cnt[k];
    atomicAdd((unsigned*)&out[6], t);
  }
}

__global__ void k_cell_count(const float* __restrict__ xyz, size_t n, GridDev g,
                             uint32_t* __restrict__ cell_of, uint32_t* __restrict__ count) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  float x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
  uint32_t c = 0xFFFFFFFFu;
  if (isfinite(x) && isfinite(y) && isfinite(z)) {
    c = cell_id(g, cell_coord(x, g.ox, g.inv_h, g.nx), cell_coord(y, g.oy, g.inv_h, g.ny),
                cell_coord(z, g.oz, g.inv_h, g.nz));
    atomicAdd(&count[c], 1u);
  }
  cell_of[i] = c;
}

// ---- exclusive scan over up to 2^32 counts: per-block scan, scan of block sums, add.
constexpr int kScanItems = 8;  // per thread
constexpr int kScanTile = kThreads * kScanItems;

__global__ void k_scan_tiles(uint32_t* __restrict__ data, size_t n, uint32_t* __restrict__ tile_sums) {
  __shared__ uint32_t warp_sums[kThreads / 32];
  size_t base = (size_t)blockIdx.x * kScanTile + (size_t)threadIdx.x * kScanItems;
  uint32_t v[kScanItems];
  uint32_t sum = 0;
#pragma unroll
  for (int k = 0; k < kScanItems; k++) {
    v[k] = (base + k < n) ? data[base + k] : 0u;
    sum += v[k];
  }
  uint32_t incl = sum;
  int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(AG2_FULL, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) warp_sums[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    uint32_t w = lane < kThreads / 32 ? warp_sums[lane] : 0u;
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t t = __shfl_up_sync(AG2_FULL, w, o);
      if (lane >= o) w += t;
    }
    if (lane < kThreads / 32) warp_sums[lane] = w;  // inclusive
  }
  __syncthreads();
  uint32_t excl = incl - sum + (warp > 0 ? warp_sums[warp - 1] : 0u);
#pragma unroll
  for (int k = 0; k < kScanItems; k++) {
    if (base + k < n) data[base + k] = excl;
    excl += v[k];
  }
  if (threadIdx.x == kThreads - 1) tile_sums[blockIdx.x] = excl;  // tile total
}

// single block: exclusive scan of the tile sums in place, total appended at [ntiles]
__global__ void k_scan_sums(uint32_t* __restrict__ sums, size_t ntiles) {
  __shared__ uint32_t warp_sums[kThreads / 32];
  __shared__ uint32_t carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (size_t base = 0; base < ntiles; base += kThreads) {
    size_t i = base + threadIdx.x;
    uint32_t v = i < ntiles ? sums[i] : 0u;
    uint32_t incl = v;
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t t = __shfl_up_sync(AG2_FULL, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      uint32_t w = lane < kThreads / 32 ? warp_sums[lane] : 0u;
      for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(AG2_FULL, w, o);
        if (lane >= o) w += t;
      }
      if (lane < kThreads / 32) warp_sums[lane] = w;
    }
    __syncthreads();
    uint32_t carry = carry_s;
    uint32_t excl = carry + incl - v + (warp > 0 ? warp_sums[warp - 1] : 0u);
    if (i < ntiles) sums[i] = excl;
    __syncthreads();
    if (threadIdx.x == kThreads - 1) carry_s = excl + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) sums[ntiles] = carry_s;
}

__global__ void k_scan_add(uint32_t* __restrict__ data, size_t n, const uint32_t* __restrict__ tile_sums,
                           size_t ntiles) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i < n) data[i] += tile_sums[i / kScanTile];
  if (i == 0) data[n] = tile_sums[ntiles];  // cell_start[ncells] = total
}

__global__ void k_scatter(const uint32_t* __restrict__ cell_of, size_t n, const uint32_t* __restrict__ cell_start,
                          uint32_t* __restrict__ fill, uint32_t* __restrict__ tmp_idx) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t c = cell_of[i];
  if (c == 0xFFFFFFFFu) return;
  uint32_t slot = cell_start[c] + atomicAdd(&fill[c], 1u);
  tmp_idx[slot] = (uint32_t)i;
}

__global__ void k_finalize(const uint32_t* __restrict__ tmp_idx, uint32_t n_finite,
                           const uint32_t* __restrict__ cell_of, const uint32_t* __restrict__ cell_start,
                           const float* __restrict__ xyz, const uint8_t* __restrict__ cam,
                           const float* __restrict__ normals_in, float4* __restrict__ pts,
                           float4* __restrict__ nrm, uint32_t* __restrict__ sorted_of_orig) {
  uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n_finite) return;
  uint32_t me = tmp_idx[s];
  uint32_t c = cell_of[me];
  uint32_t b = cell_start[c], e = cell_start[c + 1];
  uint32_t rank = 0;
  for (uint32_t j = b; j < e; j++) rank += (tmp_idx[j] < me);
  uint32_t dst = b + rank;
  pts[dst] = make_float4(xyz[3 * (size_t)me], xyz[3 * (size_t)me + 1], xyz[3 * (size_t)me + 2],
                         __uint_as_float(me));
  float4 nv = make_float4(0.f, 0.f, 0.f, __uint_as_float(cam ? (uint32_t)cam[me] : 1u));
  if (normals_in) {
    nv.x = normals_in[3 * (size_t)me];
    nv.y = normals_in[3 * (size_t)me + 1];
    nv.z = normals_in[3 * (size_t)me + 2];
  }
  nrm[dst] = nv;
  sorted_of_orig[me] = dst;
}

// One block: fixed-radius query used by ag2_radius_search (test introspection).
__global__ void k_radius_query(GridDev g, float qx, float qy, float qz, float r2, float rq,
                               uint64_t* __restrict__ keys, uint32_t cap, uint32_t* __restrict__ count) {
  QueryWindow w = make_window(g, qx, qy, qz, rq);
  int rows = w.rows();
  int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  for (int row = warp; row < rows; row += nw) {
    uint32_t b, e;
    row_span(g, w, row, b, e);
    for (uint32_t i = b + lane; i < e; i += 32) {
      float4 p = g.pts[i];
      float d2 = dist2_f(qx - p.x, qy - p.y, qz - p.z);
      if (d2 < r2) {
        uint32_t slot = atomicAdd(count, 1u);
        if (slot < cap) keys[slot] = nn_key(d2, __float_as_uint(p.w));
      }
    }
  }
}

}  // namespace

GridDev grid_dev(const ag2_ctx* c) {
  GridDev g;
  g.ox = c->origin[0], g.oy = c->origin[1], g.oz = c->origin[2];
  g.inv_h = c->inv_h;
  g.h = c->h;
  g.nx = c->dim[0], g.ny = c->dim[1], g.nz = c->dim[2];
  g.n = (int)c->n_finite;
  g.cell_start = c->cell_start.p;
  g.pts = c->pts.p;
  g.nrm = c->nrm.p;
  return g;
}

// In-place exclusive scan of data[0..n) with the total written to data[n].
int exclusive_scan_u32(ag2_ctx* c, uint32_t* data, size_t n) {
  cudaStream_t st = c->stream;
  if (n == 0) {
    AG2_CUDA(c, cudaMemsetAsync(data, 0, sizeof(uint32_t), st));
    return AG2_OK;
  }
  size_t ntiles = (n + kScanTile - 1) / kScanTile;
  AG2_CUDA(c, c->scan_tmp.reserve(ntiles + 1));
  k_scan_tiles<<<(unsigned)ntiles, kThreads, 0, st>>>(data, n, c->scan_tmp.p);
  AG2_LAUNCH_CHECK(c);
  k_scan_sums<<<1, kThreads, 0, st>>>(c->scan_tmp.p, ntiles);
  AG2_LAUNCH_CHECK(c);
  k_scan_add<<<(unsigned)((n + kThreads - 1) / kThreads), kThreads, 0, st>>>(data, n, c->scan_tmp.p, ntiles);
  AG2_LAUNCH_CHECK(c);
  return AG2_OK;
}

int grid_build(ag2_ctx* c) {
  if (c->grid_built) return AG2_OK;
  if (c->n == 0) return AG2_E_EMPTY_CLOUD;
  phase_begin(c, 0);
  cudaStream_t st = c->stream;
  const size_t n = c->n;
  AG2_CUDA(c, c->scalars.reserve(64));
  int init[7];
  {
    float inf = INFINITY;
    int pinf, ninf;
    memcpy(&pinf, &inf, 4);
    float ninff = -INFINITY;
    int raw;
    memcpy(&raw, &ninff, 4);
    ninf = raw ^ 0x7fffffff;
    init[0] = init[1] = init[2] = pinf;
    init[3] = init[4] = init[5] = ninf;
    init[6] = 0;
  }
  AG2_CUDA(c, cudaMemcpyAsync(c->scalars.p, init, sizeof(init), cudaMemcpyHostToDevice, st));
  int blocks = (int)std::min<size_t>((n + kThreads - 1) / kThreads, (size_t)c->sm_count * 4);
  k_bbox<<<blocks, kThreads, 0, st>>>(c->xyz_in.p, n, (int*)c->scalars.p);
  AG2_LAUNCH_CHECK(c);
  int res[7];
  AG2_CUDA(c, cudaMemcpyAsync(res, c->scalars.p, sizeof(res), cudaMemcpyDeviceToHost, st));
  AG2_CUDA(c, cudaStreamSynchronize(st));
  c->n_finite = (size_t)(unsigned)res[6];
  if (c->n_finite == 0) return AG2_E_EMPTY_CLOUD;
  float mn[3], mx[3];
  for (int d = 0; d < 3; d++) {
    mn[d] = ordered_float(res[d]);
    mx[d] = ordered_float(res[3 + d]);
  }
  c->min_z = mn[2];
  // cell size: the normals radius (27-cell normal queries) unless that needs more
  // than 2^26 cells or more than ~26 cell rows per hand query; then grow it.
  double h = c->p.normals_radius;
  h = std::max(h, c->p.nn_radius_hands / 13.0);
  for (;;) {
    double cells = 1;
    for (int d = 0; d < 3; d++) cells *= std::floor((double)(mx[d] - mn[d]) / h) + 1.0;
    if (cells <= (double)(1u << 26)) break;
    h *= 1.25;
  }
  c->h = (float)h;
  c->inv_h = 1.0f / c->h;
  size_t ncells = 1;
  for (int d = 0; d < 3; d++) {
    c->origin[d] = mn[d];
    c->dim[d] = (int)std::floor((double)(mx[d] - mn[d]) / h) + 1;
    ncells *= (size_t)c->dim[d];
  }
  c->ncells = ncells;
  AG2_CUDA(c, c->cell_of.reserve(n));
  AG2_CUDA(c, c->cell_start.reserve(ncells + 1));
  AG2_CUDA(c, c->cell_fill.reserve(ncells));
  AG2_CUDA(c, c->tmp_idx.reserve(n));
  AG2_CUDA(c, c->pts.reserve(n));
  AG2_CUDA(c, c->nrm.reserve(n));
  AG2_CUDA(c, c->sorted_of_orig.reserve(n));
  AG2_CUDA(c, cudaMemsetAsync(c->cell_start.p, 0, (ncells + 1) * sizeof(uint32_t), st));
  AG2_CUDA(c, cudaMemsetAsync(c->cell_fill.p, 0, ncells * sizeof(uint32_t), st));
  AG2_CUDA(c, cudaMemsetAsync(c->sorted_of_orig.p, 0xFF, n * sizeof(uint32_t), st));
  GridDev g = grid_dev(c);
  int nb = (int)((n + kThreads - 1) / kThreads);
  k_cell_count<<<nb, kThreads, 0, st>>>(c->xyz_in.p, n, g, c->cell_of.p, c->cell_start.p);
  AG2_LAUNCH_CHECK(c);
  AG2_TRY(exclusive_scan_u32(c, c->cell_start.p, ncells));
  k_scatter<<<nb, kThreads, 0, st>>>(c->cell_of.p, n, c->cell_start.p, c->cell_fill.p, c->tmp_idx.p);
  AG2_LAUNCH_CHECK(c);
  k_finalize<<<(unsigned)((c->n_finite + kThreads - 1) / kThreads), kThreads, 0, st>>>(
      c->tmp_idx.p, (uint32_t)c->n_finite, c->cell_of.p, c->cell_start.p, c->xyz_in.p, c->cam_in.p,
      c->normals_given ? c->normals_in.p : nullptr, c->pts.p, c->nrm.p, c->sorted_of_orig.p);
  AG2_LAUNCH_CHECK(c);
  c->grid_built = true;
  c->have_normals = c->normals_given;
  phase_end(c, 0);
  return AG2_OK;
}

int radius_search_device(ag2_ctx* c, const float* q, double radius, std::vector<int32_t>& idx,
                         std::vector<float>& d2) {
  AG2_TRY(grid_build(c));
  float r2 = (float)(radius * radius);
  float rq = (float)radius * 1.0001f + 1e-6f;
  uint32_t cap = (uint32_t)c->n_finite;
  DevBuf<uint64_t> keys;
  AG2_CUDA(c, keys.reserve(cap));
  AG2_CUDA(c, cudaMemsetAsync(c->scalars.p + 16, 0, 4, c->stream));
  k_radius_query<<<1, 1024, 0, c->stream>>>(grid_dev(c), q[0], q[1], q[2], r2, rq, keys.p, cap,
                                            c->scalars.p + 16);
  AG2_LAUNCH_CHECK(c);
  uint32_t cnt = 0;
  AG2_CUDA(c, cudaMemcpyAsync(&cnt, c->scalars.p + 16, 4, cudaMemcpyDeviceToHost, c->stream));
  AG2_CUDA(c, cudaStreamSynchronize(c->stream));
  std::vector<uint64_t> hk(cnt);
  if (cnt) AG2_CUDA(c, cudaMemcpy(hk.data(), keys.p, cnt * sizeof(uint64_t), cudaMemcpyDeviceToHost));
  std::sort(hk.begin(), hk.end());
  idx.resize(cnt);
  d2.resize(cnt);
  for (uint32_t i = 0; i < cnt; i++) {
    idx[i] = (int32_t)(uint32_t)(hk[i] & 0xFFFFFFFFu);
    uint32_t bits = (uint32_t)(hk[i] >> 32);
    memcpy(&d2[i], &bits, 4);
  }
  return AG2_OK;
}

}  // namespace ag2
