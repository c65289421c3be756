// f1: pre-processing, replacing CloudCamera::filterWorkspace / voxelizeCloud /
// subsampleUniformly (reference cloud_camera.cpp:89-121, :124-168, :171-178) as driven by
// GraspDetector::preprocessPointCloud (grasp_detector.cpp:285-350).  HBM-bound byte work.
//
//   workspace filter  flag (strict >, < on float coordinates promoted to double; NaN fails)
//                     -> exclusive scan -> ordered compaction of xyz / camera mask / normals;
//   voxelisation      min corner (float), voxel = floor((p - min) / (float)cell) per axis; the
//                     reference inserts the voxels into a std::set ordered by (x, y, z) and pushes
//                     the index of every point that OPENS a voxel onto idx_cam_source, i.e. in scan
//                     order; it then gives the i-th voxel of the SORTED set the camera source of
//                     idx_cam_source[i], the i-th opener in scan order -- not of the voxel's own
//                     first point (cloud_camera.cpp:139-153).  That is what the reference computes,
//                     so it is reproduced: a direct-address table over the voxel box
//                     (id = (vx*ny + vy)*nz + vz, ascending id == the set's order) takes
//                     atomicMin(point index); the openers (table[voxel(i)] == i) are compacted in
//                     point order into first_seen[]; occupied entries are compacted in id order,
//                     emit the voxel corner fl(fl((float)v * cell) + min) exactly as the
//                     reference's float expression, and entry a takes cam[first_seen[a]];
//   subsampling       uniform, without replacement, ascending (Vitter's algorithm A as in
//                     pcl::RandomSample) with the counter-based RNG replacing rand() (D1).
// Documented deviations: the reference's camera_source re-indexing in filterWorkspace uses
// linear indices into a C x N matrix (cloud_camera.cpp:107), which is only right for C = 1; the
// column gather it intends is implemented.  voxelizeCloud leaves normals_ stale (wrong size);
// here precomputed normals are dropped when the cloud is voxelised and recomputed.
#include <math.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "dev_common.cuh"

namespace ag2 {
namespace {

constexpr int kT = 256;

struct Ws {
  double v[6];
};

__global__ void k_ws_flag(const float* __restrict__ xyz, size_t stride_f, size_t n, Ws ws, int enabled,
                          uint32_t* __restrict__ flag) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t keep = 1;
  if (enabled) {
    double x = (double)xyz[i * stride_f], y = (double)xyz[i * stride_f + 1], z = (double)xyz[i * stride_f + 2];
    keep = (x > ws.v[0] && x < ws.v[1] && y > ws.v[2] && y < ws.v[3] && z > ws.v[4] && z < ws.v[5]) ? 1u : 0u;
  }
  flag[i] = keep;
}

__global__ void k_ws_compact(const float* __restrict__ xyz, size_t stride_f, const uint8_t* __restrict__ cam,
                             const float* __restrict__ nrm, size_t n, const uint32_t* __restrict__ scan,
                             float* __restrict__ xyz_out, uint8_t* __restrict__ cam_out, float* __restrict__ nrm_out) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t a = scan[i], b = scan[i + 1];
  if (a == b) return;
  xyz_out[3 * (size_t)a] = xyz[i * stride_f];
  xyz_out[3 * (size_t)a + 1] = xyz[i * stride_f + 1];
  xyz_out[3 * (size_t)a + 2] = xyz[i * stride_f + 2];
  cam_out[a] = cam ? cam[i] : (uint8_t)1;
  if (nrm) {
    nrm_out[3 * (size_t)a] = nrm[3 * i];
    nrm_out[3 * (size_t)a + 1] = nrm[3 * i + 1];
    nrm_out[3 * (size_t)a + 2] = nrm[3 * i + 2];
  }
}

// pts.row(k).minCoeff() / maxCoeff() over ALL points (Eigen min/max: NaN entries never win a <)
__global__ void k_minmax(const float* __restrict__ xyz, size_t n, int* __restrict__ out /*6 ordered ints*/) {
  float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    for (int d = 0; d < 3; d++) {
      float v = xyz[3 * i + d];
      if (v < mn[d]) mn[d] = v;
      if (v > mx[d]) mx[d] = v;
    }
  for (int o = 16; o > 0; o >>= 1)
    for (int d = 0; d < 3; d++) {
      mn[d] = fminf(mn[d], __shfl_xor_sync(AG2_FULL, mn[d], o));
      mx[d] = fmaxf(mx[d], __shfl_xor_sync(AG2_FULL, mx[d], o));
    }
  if ((threadIdx.x & 31) == 0)
    for (int d = 0; d < 3; d++) {
      atomicMin(&out[d], float_ordered(mn[d]));
      atomicMax(&out[3 + d], float_ordered(mx[d]));
    }
}

struct Vox {
  float mn[3];
  float cell;
  int dim[3];
};

__device__ __forceinline__ bool voxel_of(const Vox& v, const float* p, size_t& id) {
  int c[3];
  for (int d = 0; d < 3; d++) {
    float f = floorf(__fdiv_rn(__fsub_rn(p[d], v.mn[d]), v.cell));  // floorVector((p - min_xyz) / cell_size)
    if (!(f >= 0.f && f < (float)v.dim[d])) return false;          // NaN coordinates
    c[d] = (int)f;
  }
  id = ((size_t)c[0] * v.dim[1] + c[1]) * (size_t)v.dim[2] + c[2];
  return true;
}

__global__ void k_vox_mark(const float* __restrict__ xyz, size_t n, Vox v, uint32_t* __restrict__ table) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  size_t id;
  if (voxel_of(v, xyz + 3 * i, id)) atomicMin(&table[id], (uint32_t)i);
}

__global__ void k_vox_flag(const uint32_t* __restrict__ table, size_t ncells, uint32_t* __restrict__ flag) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i < ncells) flag[i] = table[i] != 0xFFFFFFFFu ? 1u : 0u;
}

// idx_cam_source (cloud_camera.cpp:139-141): the points that open a voxel, in scan order
__global__ void k_vox_opener_flag(const float* __restrict__ xyz, size_t n, Vox v, const uint32_t* __restrict__ table,
                                  uint32_t* __restrict__ flag) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  size_t id;
  flag[i] = (voxel_of(v, xyz + 3 * i, id) && table[id] == (uint32_t)i) ? 1u : 0u;
}
__global__ void k_vox_opener_list(const float* __restrict__ xyz, size_t n, Vox v, const uint32_t* __restrict__ table,
                                  const uint32_t* __restrict__ scan, uint32_t* __restrict__ first_seen) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  size_t id;
  if (voxel_of(v, xyz + 3 * i, id) && table[id] == (uint32_t)i) first_seen[scan[i]] = (uint32_t)i;
}

__global__ void k_vox_emit(const uint32_t* __restrict__ table, const uint32_t* __restrict__ scan, size_t ncells,
                           Vox v, const uint8_t* __restrict__ cam, const uint32_t* __restrict__ first_seen,
                           float* __restrict__ xyz_out, uint8_t* __restrict__ cam_out) {
  size_t id = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (id >= ncells) return;
  uint32_t a = scan[id], b = scan[id + 1];
  if (a == b) return;
  int cz = (int)(id % (size_t)v.dim[2]);
  int cy = (int)((id / (size_t)v.dim[2]) % (size_t)v.dim[1]);
  int cx = (int)(id / ((size_t)v.dim[2] * (size_t)v.dim[1]));
  // voxels.row(k) * cell_size + ones * min_xyz(k), float, unfused
  xyz_out[3 * (size_t)a] = __fadd_rn(__fmul_rn((float)cx, v.cell), v.mn[0]);
  xyz_out[3 * (size_t)a + 1] = __fadd_rn(__fmul_rn((float)cy, v.cell), v.mn[1]);
  xyz_out[3 * (size_t)a + 2] = __fadd_rn(__fmul_rn((float)cz, v.cell), v.mn[2]);
  cam_out[a] = cam[first_seen[a]];  // camera_source_(j, idx_cam_source[i]), i = position in the sorted set
}

}  // namespace

int preprocess_cloud(ag2_ctx* c, const float* xyz, size_t stride_bytes, size_t n, const uint8_t* cam, int num_cams,
                     const float* normals, int filter_ws, int voxelize, double voxel_size) {
  cudaStream_t st = c->stream;
  const size_t sf = stride_bytes / 4;
  DevBuf<float> d_in, d_nrm, f_xyz, f_nrm, v_xyz;
  DevBuf<uint8_t> d_cam, f_cam, v_cam;
  DevBuf<uint32_t> flag, table;
  DevBuf<int> mm;
  AG2_CUDA(c, d_in.reserve(n * sf));
  AG2_CUDA(c, cudaMemcpyAsync(d_in.p, xyz, n * stride_bytes, cudaMemcpyHostToDevice, st));
  if (cam) {
    AG2_CUDA(c, d_cam.reserve(n));
    AG2_CUDA(c, cudaMemcpyAsync(d_cam.p, cam, n, cudaMemcpyHostToDevice, st));
  }
  if (normals && !voxelize) {
    AG2_CUDA(c, d_nrm.reserve(3 * n));
    AG2_CUDA(c, cudaMemcpyAsync(d_nrm.p, normals, 12 * n, cudaMemcpyHostToDevice, st));
  }
  // ---- workspace filter (also packs the stride away)
  AG2_CUDA(c, flag.reserve(n + 1));
  Ws ws;
  for (int k = 0; k < 6; k++) ws.v[k] = c->p.workspace[k];
  const unsigned nb = (unsigned)((n + kT - 1) / kT);
  k_ws_flag<<<nb, kT, 0, st>>>(d_in.p, sf, n, ws, filter_ws, flag.p);
  AG2_LAUNCH_CHECK(c);
  AG2_TRY(exclusive_scan_u32(c, flag.p, n));
  uint32_t nf = 0;
  AG2_CUDA(c, cudaMemcpyAsync(&nf, flag.p + n, 4, cudaMemcpyDeviceToHost, st));
  AG2_CUDA(c, cudaStreamSynchronize(st));
  if (nf == 0) {
    AG2_TRY(ag2_set_cloud_device(c, nullptr, 0, nullptr, num_cams, nullptr));
    return AG2_OK;
  }
  AG2_CUDA(c, f_xyz.reserve(3 * (size_t)nf));
  AG2_CUDA(c, f_cam.reserve(nf));
  if (d_nrm.p) AG2_CUDA(c, f_nrm.reserve(3 * (size_t)nf));
  k_ws_compact<<<nb, kT, 0, st>>>(d_in.p, sf, d_cam.p, d_nrm.p, n, flag.p, f_xyz.p, f_cam.p, f_nrm.p);
  AG2_LAUNCH_CHECK(c);
  if (!voxelize) {
    AG2_TRY(ag2_set_cloud_device(c, f_xyz.p, nf, f_cam.p, num_cams, f_nrm.p));
    AG2_CUDA(c, cudaStreamSynchronize(st));
    return AG2_OK;
  }
  // ---- voxelisation
  AG2_CUDA(c, mm.reserve(8));
  int init[6];
  {
    float pinf = INFINITY, ninf = -INFINITY;
    int a, b;
    memcpy(&a, &pinf, 4);
    memcpy(&b, &ninf, 4);
    init[0] = init[1] = init[2] = a;
    init[3] = init[4] = init[5] = b ^ 0x7fffffff;
  }
  AG2_CUDA(c, cudaMemcpyAsync(mm.p, init, sizeof(init), cudaMemcpyHostToDevice, st));
  k_minmax<<<std::min<unsigned>((nf + kT - 1) / kT, (unsigned)c->sm_count * 8), kT, 0, st>>>(f_xyz.p, nf, mm.p);
  AG2_LAUNCH_CHECK(c);
  int res[6];
  AG2_CUDA(c, cudaMemcpyAsync(res, mm.p, sizeof(res), cudaMemcpyDeviceToHost, st));
  AG2_CUDA(c, cudaStreamSynchronize(st));
  Vox v;
  v.cell = (float)voxel_size;
  size_t ncells = 1;
  for (int d = 0; d < 3; d++) {
    v.mn[d] = ordered_float(res[d]);
    float mx = ordered_float(res[3 + d]);
    volatile float q = (mx - v.mn[d]) / v.cell;  // same float expression as the per-point voxel index
    double dim = floor((double)q) + 1.0;
    if (!(dim >= 1.0 && dim < 2097152.0)) return AG2_E_CAPACITY;
    v.dim[d] = (int)dim;
    ncells *= (size_t)v.dim[d];
  }
  if (ncells > ((size_t)1 << 30)) {
    c->err = "voxelize: the voxel box has more than 2^30 cells; filter the workspace first";
    return AG2_E_CAPACITY;
  }
  AG2_CUDA(c, table.reserve(ncells));
  AG2_CUDA(c, cudaMemsetAsync(table.p, 0xFF, ncells * 4, st));
  k_vox_mark<<<(nf + kT - 1) / kT, kT, 0, st>>>(f_xyz.p, nf, v, table.p);
  AG2_LAUNCH_CHECK(c);
  DevBuf<uint32_t> vflag;
  AG2_CUDA(c, vflag.reserve(ncells + 1));
  const unsigned cb = (unsigned)((ncells + kT - 1) / kT);
  k_vox_flag<<<cb, kT, 0, st>>>(table.p, ncells, vflag.p);
  AG2_LAUNCH_CHECK(c);
  AG2_TRY(exclusive_scan_u32(c, vflag.p, ncells));
  uint32_t nv = 0;
  AG2_CUDA(c, cudaMemcpyAsync(&nv, vflag.p + ncells, 4, cudaMemcpyDeviceToHost, st));
  AG2_CUDA(c, cudaStreamSynchronize(st));
  AG2_CUDA(c, v_xyz.reserve(3 * (size_t)std::max(nv, 1u)));
  AG2_CUDA(c, v_cam.reserve(std::max(nv, 1u)));
  // idx_cam_source: openers in point order (flag is free again: at least nf + 1 words)
  DevBuf<uint32_t> first_seen;
  AG2_CUDA(c, first_seen.reserve(std::max(nv, 1u)));
  k_vox_opener_flag<<<(nf + kT - 1) / kT, kT, 0, st>>>(f_xyz.p, nf, v, table.p, flag.p);
  AG2_LAUNCH_CHECK(c);
  AG2_TRY(exclusive_scan_u32(c, flag.p, nf));
  k_vox_opener_list<<<(nf + kT - 1) / kT, kT, 0, st>>>(f_xyz.p, nf, v, table.p, flag.p, first_seen.p);
  AG2_LAUNCH_CHECK(c);
  k_vox_emit<<<cb, kT, 0, st>>>(table.p, vflag.p, ncells, v, f_cam.p, first_seen.p, v_xyz.p, v_cam.p);
  AG2_LAUNCH_CHECK(c);
  AG2_TRY(ag2_set_cloud_device(c, v_xyz.p, nv, v_cam.p, num_cams, nullptr));
  AG2_CUDA(c, cudaStreamSynchronize(st));
  return AG2_OK;
}

}  // namespace ag2

using namespace ag2;

extern "C" {

int ag2_preprocess_cloud(ag2_ctx* c, const float* xyz, size_t stride_bytes, size_t n, const uint8_t* cam_mask,
                         int num_cams, const float* normals, int filter_workspace, int voxelize, double voxel_size) {
  if (!c || (n && !xyz) || stride_bytes < 12 || stride_bytes % 4 || num_cams < 1 || num_cams > 8 ||
      (voxelize && !(voxel_size > 0)))
    return AG2_E_INVALID;
  if (n >= 0x7fffffffu) return AG2_E_CAPACITY;
  if (cudaSetDevice(c->device) != cudaSuccess) return AG2_E_NO_DEVICE;
  if (n == 0) return ag2_set_cloud_device(c, nullptr, 0, nullptr, num_cams, nullptr);
  return preprocess_cloud(c, xyz, stride_bytes, n, cam_mask, num_cams, normals, filter_workspace, voxelize,
                          voxel_size);
}

int ag2_get_cloud(ag2_ctx* c, float* xyz, uint8_t* cam_mask, size_t cap, size_t* n) {
  if (!c || !n) return AG2_E_INVALID;
  *n = c->n;
  if (c->n > cap) return AG2_E_CAPACITY;
  if (c->n == 0) return AG2_OK;
  cudaSetDevice(c->device);
  if (xyz) AG2_CUDA(c, cudaMemcpyAsync(xyz, c->xyz_in.p, 12 * c->n, cudaMemcpyDeviceToHost, c->stream));
  if (cam_mask) AG2_CUDA(c, cudaMemcpyAsync(cam_mask, c->cam_in.p, c->n, cudaMemcpyDeviceToHost, c->stream));
  AG2_CUDA(c, cudaStreamSynchronize(c->stream));
  return AG2_OK;
}

// pcl::RandomSample (Vitter's algorithm A): num_samples distinct indices, ascending.
int ag2_subsample_uniformly(ag2_ctx* c, int num_samples, uint64_t seed) {
  if (!c || num_samples < 0) return AG2_E_INVALID;
  const size_t N0 = c->n;
  if ((size_t)num_samples > N0) return AG2_E_INVALID;
  std::vector<int32_t> out((size_t)num_samples);
  auto unif = [seed](uint64_t k) {
    uint64_t x = seed ^ (0xD1B54A32D192ED03ull * (k + 1));
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    x ^= x >> 31;
    return (float)((double)(x >> 11) * (1.0 / 9007199254740992.0));
  };
  size_t N = N0, top = N0 - (size_t)num_samples, index = 0, i = 0;
  uint64_t draw = 0;
  for (size_t m = (size_t)num_samples; m >= 2; m--) {
    float V = unif(draw++);
    size_t S = 0;
    float quot = (float)top / (float)N;
    while (quot > V) {
      S++;
      top--;
      N--;
      quot = quot * (float)top / (float)N;
    }
    index += S;
    out[i++] = (int32_t)index++;
    N--;
  }
  if (num_samples >= 1) {
    index += (size_t)((float)N * unif(draw++));
    if (index >= N0) index = N0 - 1;
    out[i++] = (int32_t)index;
  }
  return ag2_set_sample_indices(c, out.data(), out.size());
}

int ag2_get_sample_indices(ag2_ctx* c, int32_t* idx, size_t cap, size_t* n) {
  if (!c || !n) return AG2_E_INVALID;
  if (c->use_points) return AG2_E_STATE;
  if (c->h_sample_idx_stale) {
    cudaSetDevice(c->device);
    c->h_sample_idx.resize(c->num_samples);
    AG2_CUDA(c, cudaMemcpyAsync(c->h_sample_idx.data(), c->sample_idx.p, 4 * c->num_samples, cudaMemcpyDeviceToHost,
                                c->stream));
    AG2_CUDA(c, cudaStreamSynchronize(c->stream));
    c->h_sample_idx_stale = false;
  }
  *n = c->h_sample_idx.size();
  if (*n > cap) return AG2_E_CAPACITY;
  if (*n && !idx) return AG2_E_INVALID;
  std::copy(c->h_sample_idx.begin(), c->h_sample_idx.end(), idx);
  return AG2_OK;
}

}  // extern "C"
