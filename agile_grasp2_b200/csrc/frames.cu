// K5: local reference frames, replacing HandSearch::calculateLocalFrames
// (reference hand_search.cpp:97-170 for index samples, :238-317 for xyz samples)
// and LocalFrame::findAverageNormalAxis (local_frame.cpp:26-59).
// One CTA per sample: gather the nn_radius_taubin neighbourhood, sort it by
// (d2, id) (the reference draws `rand() % n` into the distance-sorted list), draw
// min(50, n) normals with the counter-based replacement of rand(), majority
// camera, M = N N^T, cyclic Jacobi (double), Gram^6 column argmax, frame axes.
// All double arithmetic is IEEE and unfused (-fmad=false) in the oracle's order.
#include <math.h>

#include <algorithm>

#include "dev_common.cuh"

namespace ag2 {
namespace {

constexpr int kFrameThreads = 128;
// k_frames sorts a neighbourhood in shared memory (12 B per neighbour, power-of-two capacity).  Three launches cover
// all sizes without the host knowing the largest one: every sample at capacity 1024, and two short persistent
// launches over the (usually empty) lists of denser ones at 4096 and 16384.
constexpr uint32_t kFrameCap0 = 1024, kFrameCap1 = 4096, kFrameCap2 = 16384;
constexpr int kOrdBuckets = 256;                       // d2 buckets of the first class's ordering (multiple of 32)
constexpr size_t kFrameSmem0 = (size_t)kFrameCap0 * 18;  // keys 8 + positions 4 + ordered positions 4 + grouping 2 bytes

__device__ __forceinline__ void sample_query(const float* __restrict__ xyz_in, const int32_t* __restrict__ sidx,
                                             const double* __restrict__ sxyz, uint32_t s, float q[3]) {
  if (sxyz) {
    q[0] = (float)sxyz[3 * (size_t)s];
    q[1] = (float)sxyz[3 * (size_t)s + 1];
    q[2] = (float)sxyz[3 * (size_t)s + 2];
  } else {
    size_t i = (size_t)sidx[s];
    q[0] = xyz_in[3 * i];
    q[1] = xyz_in[3 * i + 1];
    q[2] = xyz_in[3 * i + 2];
  }
}

// warp per sample: neighbour count at the frame radius
__global__ void k_count_sample_neighbors(GridDev g, const float* __restrict__ xyz_in,
                                         const int32_t* __restrict__ sidx, const double* __restrict__ sxyz,
                                         uint32_t S, float r2, float rq, uint32_t* __restrict__ kt,
                                         uint32_t* __restrict__ max_out, uint32_t* __restrict__ big_list /* [2][S] */,
                                         uint32_t* __restrict__ big_n /* [2] */) {
  uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  int lane = threadIdx.x & 31;
  if (s >= S) return;
  float q[3];
  sample_query(xyz_in, sidx, sxyz, s, q);
  QueryWindow w = make_window(g, q[0], q[1], q[2], rq);
  int rows = w.rows();
  uint32_t cnt = 0;
  for (int row = 0; row < rows; row++) {
    uint32_t b, e;
    row_span(g, w, row, b, e);
    for (uint32_t j = b + lane; j < e; j += 32) {
      float4 p = g.pts[j];
      cnt += (dist2_f(q[0] - p.x, q[1] - p.y, q[2] - p.z) < r2);
    }
  }
  for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(AG2_FULL, cnt, o);
  if (lane == 0) {
    kt[s] = cnt;
    atomicMax(max_out, cnt);
    // neighbourhoods beyond the first launch's capacity go to the lists of the larger ones (order is free: every
    // sample writes its own frame)
    if (cnt > kFrameCap0) {
      const int cls = cnt > kFrameCap1 ? 1 : 0;
      big_list[(size_t)cls * S + atomicAdd(&big_n[cls], 1u)] = s;
    }
  }
}

// cyclic Jacobi of a symmetric 3x3 — same sequence of operations as the oracle (D3)
__device__ void jacobi3(double a[3][3], double d[3], double v[3][3]) {
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) v[i][j] = (i == j) ? 1.0 : 0.0;
  for (int sweep = 0; sweep < 32; sweep++) {
    double off = fabs(a[0][1]) + fabs(a[0][2]) + fabs(a[1][2]);
    if (off == 0.0) break;
    for (int p = 0; p < 2; p++)
      for (int q = p + 1; q < 3; q++) {
        double apq = a[p][q];
        if (apq == 0.0) continue;
        double g = 100.0 * fabs(apq);
        if (sweep > 3 && fabs(a[p][p]) + g == fabs(a[p][p]) && fabs(a[q][q]) + g == fabs(a[q][q])) {
          a[p][q] = a[q][p] = 0.0;
          continue;
        }
        double h = a[q][q] - a[p][p];
        double t;
        if (fabs(h) + g == fabs(h)) {
          t = apq / h;
        } else {
          double theta = 0.5 * h / apq;
          t = 1.0 / (fabs(theta) + sqrt(1.0 + theta * theta));
          if (theta < 0.0) t = -t;
        }
        double c = 1.0 / sqrt(1.0 + t * t);
        double s = t * c;
        double app = a[p][p], aqq = a[q][q];
        a[p][p] = app - t * apq;
        a[q][q] = aqq + t * apq;
        a[p][q] = a[q][p] = 0.0;
        int r = 3 - p - q;
        double arp = a[r][p], arq = a[r][q];
        a[r][p] = a[p][r] = c * arp - s * arq;
        a[r][q] = a[q][r] = s * arp + c * arq;
        for (int k = 0; k < 3; k++) {
          double vkp = v[k][p], vkq = v[k][q];
          v[k][p] = c * vkp - s * vkq;
          v[k][q] = s * vkp + c * vkq;
        }
      }
  }
  for (int i = 0; i < 3; i++) d[i] = a[i][i];
}

__device__ __forceinline__ double dot3d(const double* a, const double* b) {
  return a[0] * b[0] + (a[1] * b[1] + a[2] * b[2]);  // Eigen fixed-size reduction order
}

struct FrameArgs {
  GridDev g;
  const float* xyz_in;
  const int32_t* sidx;
  const double* sxyz;
  uint32_t S;
  const uint32_t* kt;
  uint32_t cap, lo;            // this launch handles neighbourhoods of lo < count <= cap (and count == 0 when lo == 0)
  const uint32_t* list;        // null: sample = blockIdx.x; else the CTAs stride over list[0 .. *list_n)
  const uint32_t* list_n;
  float r2, rq;
  int num_cams;
  uint64_t seed;
  int64_t id_offset, id_stride;
  double cam0x, cam0y, cam0z, cam1x, cam1y, cam1z;
  ag2_frame* out;
  uint32_t* num_valid;
};

__device__ void frame_of_sample(const FrameArgs& P, const uint32_t s, unsigned char* smem) {
  const GridDev& g = P.g;
  const float* xyz_in = P.xyz_in;
  const int32_t* sidx = P.sidx;
  const double* sxyz = P.sxyz;
  const uint32_t* kt = P.kt;
  const uint32_t cap = P.cap;
  const float r2 = P.r2, rq = P.rq;
  const int num_cams = P.num_cams;
  const uint64_t seed = P.seed;
  const int64_t id_offset = P.id_offset, id_stride = P.id_stride;
  const double cam0x = P.cam0x, cam0y = P.cam0y, cam0z = P.cam0z, cam1x = P.cam1x, cam1y = P.cam1y, cam1z = P.cam1z;
  ag2_frame* out = P.out;
  uint32_t* num_valid = P.num_valid;
  uint64_t* keys = (uint64_t*)smem;
  uint32_t* vals = (uint32_t*)(keys + cap);
  __shared__ double N[50][3];
  __shared__ double colsum[50];
  __shared__ double M[9];
  __shared__ int num_source[8];
  __shared__ uint32_t filled;
  __shared__ uint32_t bstart[kOrdBuckets + 1], bcursor[kOrdBuckets];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = kFrameThreads / 32;
  float q[3];
  sample_query(xyz_in, sidx, sxyz, s, q);
  const uint32_t count = kt[s];
  if (count > cap || (count <= P.lo && !(count == 0 && P.lo == 0))) return;  // another launch's (uniform per CTA)
  ag2_frame f;
  if (tid == 0) {
    f.sample[0] = (double)q[0], f.sample[1] = (double)q[1], f.sample[2] = (double)q[2];
    for (int a = 0; a < 3; a++) f.normal[a] = f.binormal[a] = f.curvature_axis[a] = 0.0;
    f.valid = 0, f.majority_cam = 0, f.num_neighbors = (int32_t)count, f.reserved = 0;
  }
  if (count == 0) {  // sample dropped (hand_search.cpp:122,157-163)
    if (tid == 0) out[s] = f;
    return;
  }
  if (tid == 0) filled = 0;
  if (tid < 8) num_source[tid] = 0;
  __syncthreads();
  QueryWindow w = make_window(g, q[0], q[1], q[2], rq);
  int rows = w.rows();
  for (int row = warp; row < rows; row += nw) {
    uint32_t b, e;
    row_span(g, w, row, b, e);
    for (uint32_t base = b; base < e; base += 32) {
      uint32_t j = base + lane;
      bool hit = false;
      float d2 = 0.f;
      uint32_t orig = 0;
      if (j < e) {
        float4 p = g.pts[j];
        d2 = dist2_f(q[0] - p.x, q[1] - p.y, q[2] - p.z);
        hit = d2 < r2;
        orig = __float_as_uint(p.w);
      }
      uint32_t ballot = __ballot_sync(AG2_FULL, hit);
      uint32_t basepos = 0;
      if (lane == 0 && ballot) basepos = atomicAdd(&filled, (uint32_t)__popc(ballot));
      basepos = __shfl_sync(AG2_FULL, basepos, 0);
      if (hit) {
        uint32_t slot = basepos + __popc(ballot & ((1u << lane) - 1));
        keys[slot] = nn_key(d2, orig);
        vals[slot] = j;
      }
    }
  }
  __syncthreads();
  const uint32_t* ordered;  // cell-sorted positions of the neighbours in ascending (d2, input index) order
  if (cap == kFrameCap0) {
    // First capacity class (every sample but the few dense ones): the order comes from kOrdBuckets monotone d2
    // buckets and, inside a bucket, rank = number of smaller keys among its handful of members -- five barriers
    // instead of the 55 of a 1024-element bitonic network, which was most of this kernel's dependent chain.
    uint32_t* svals = vals + cap;
    uint16_t* grouped = (uint16_t*)(svals + cap);
    const float bscale = (float)kOrdBuckets / r2;
    auto bucket_of = [&](uint64_t key) -> int {
      return min(kOrdBuckets - 1, (int)(__uint_as_float((uint32_t)(key >> 32)) * bscale));  // monotone in d2
    };
    for (int b = tid; b <= kOrdBuckets; b += kFrameThreads) bstart[b] = 0;
    __syncthreads();
    for (uint32_t t = tid; t < count; t += kFrameThreads) atomicAdd(&bstart[bucket_of(keys[t]) + 1], 1u);
    __syncthreads();
    if (warp == 0) {  // inclusive prefix over the bucket counts, 32 lanes x kOrdBuckets / 32 consecutive buckets
      constexpr int per = kOrdBuckets / 32;
      uint32_t loc[per], run = 0;
#pragma unroll
      for (int k = 0; k < per; k++) run += bstart[1 + lane * per + k], loc[k] = run;
      uint32_t incl = run;
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t up = __shfl_up_sync(AG2_FULL, incl, o);
        if (lane >= o) incl += up;
      }
      const uint32_t before = incl - run;
#pragma unroll
      for (int k = 0; k < per; k++) {
        bstart[1 + lane * per + k] = before + loc[k];
        bcursor[lane * per + k] = before + (k ? loc[k - 1] : 0u);
      }
    }
    __syncthreads();
    for (uint32_t t = tid; t < count; t += kFrameThreads)
      grouped[atomicAdd(&bcursor[bucket_of(keys[t])], 1u)] = (uint16_t)t;
    __syncthreads();
    for (uint32_t p = tid; p < count; p += kFrameThreads) {
      const uint32_t t = grouped[p];
      const uint64_t key = keys[t];
      const int b = bucket_of(key);
      const uint32_t lo = bstart[b], hi = bstart[b + 1];
      uint32_t rank = lo;
      for (uint32_t u = lo; u < hi; u++) rank += keys[grouped[u]] < key;  // keys are distinct: (d2, input index)
      svals[rank] = vals[t];
    }
    __syncthreads();
    ordered = svals;
  } else {
    const uint32_t n_pad = next_pow2(count);
    for (uint32_t t = count + tid; t < n_pad; t += kFrameThreads) {
      keys[t] = ~0ull;
      vals[t] = 0;
    }
    __syncthreads();
    bitonic_sort_kv(keys, vals, n_pad, (uint32_t)tid, (uint32_t)kFrameThreads, BlockSync());
    ordered = vals;
  }
  const int n = (int)min(50u, count);
  if (tid < n) {
    uint32_t r = draw_index(seed, id_offset + id_stride * (int64_t)s, tid, count);
    uint32_t pos = ordered[r];
    float4 nv = g.nrm[pos];
    double nx = (double)nv.x, ny = (double)nv.y, nz = (double)nv.z;
    uint32_t mask = __float_as_uint(nv.w);
    for (int cam = 0; cam < num_cams; cam++)
      if ((mask >> cam) & 1u) atomicAdd(&num_source[cam], 1);
    double mag = sqrt(nx * nx + (ny * ny + nz * nz));
    N[tid][0] = nx / mag;
    N[tid][1] = ny / mag;
    N[tid][2] = nz / mag;
  }
  __syncthreads();
  if (tid < 9) {  // M = sum_j n_j n_j^T, sequential in draw order
    int a = tid / 3, b = tid % 3;
    double acc = 0.0;
    for (int j = 0; j < n; j++) acc += N[j][a] * N[j][b];
    M[tid] = acc;
  }
  if (tid >= 32 && tid < 32 + n) {  // column sums of (N^T N)^.6 (local_frame.cpp:42)
    int j = tid - 32;
    double sacc = 0.0;
    for (int i = 0; i < n; i++) {
      double gij = (N[i][0] * N[j][0] + N[i][1] * N[j][1]) + N[i][2] * N[j][2];
      double g2 = gij * gij;
      sacc += (g2 * g2) * g2;
    }
    colsum[j] = sacc;
  }
  __syncthreads();
  if (tid != 0) return;
  int majority = 0;
  for (int cam = 1; cam < num_cams; cam++)
    if (num_source[cam] > num_source[majority]) majority = cam;
  double A[3][3], d[3], v[3][3];
  for (int a = 0; a < 3; a++)
    for (int b = 0; b < 3; b++) A[a][b] = M[a * 3 + b];
  jacobi3(A, d, v);
  int min_index = 0;
  for (int i = 1; i < 3; i++)
    if (d[i] < d[min_index]) min_index = i;
  double cax[3] = {v[0][min_index], v[1][min_index], v[2][min_index]};
  int max_index = 0;
  double best = -INFINITY;
  for (int j = 0; j < n; j++)
    if (colsum[j] > best) {
      best = colsum[j];
      max_index = j;
    }
  double nm[3] = {N[max_index][0], N[max_index][1], N[max_index][2]};
  double np[3];
  for (int a = 0; a < 3; a++) {
    double P0 = (a == 0 ? 1.0 : 0.0) - cax[a] * cax[0];
    double P1 = (a == 1 ? 1.0 : 0.0) - cax[a] * cax[1];
    double P2 = (a == 2 ? 1.0 : 0.0) - cax[a] * cax[2];
    np[a] = (P0 * nm[0] + P1 * nm[1]) + P2 * nm[2];
  }
  double len = sqrt(dot3d(np, np));
  double nrm[3] = {np[0] / len, np[1] / len, np[2] / len};
  double bin[3] = {cax[1] * nrm[2] - cax[2] * nrm[1], cax[2] * nrm[0] - cax[0] * nrm[2],
                   cax[0] * nrm[1] - cax[1] * nrm[0]};
  int mc = majority > 1 ? 1 : majority;
  double co[3] = {mc ? cam1x : cam0x, mc ? cam1y : cam0y, mc ? cam1z : cam0z};
  double s2s[3] = {f.sample[0] - co[0], f.sample[1] - co[1], f.sample[2] - co[2]};
  if (dot3d(nrm, s2s) > 0)
    for (int a = 0; a < 3; a++) nrm[a] = -nrm[a];
  if (dot3d(bin, s2s) > 0)
    for (int a = 0; a < 3; a++) bin[a] = -bin[a];
  double ax[3] = {nrm[1] * bin[2] - nrm[2] * bin[1], nrm[2] * bin[0] - nrm[0] * bin[2],
                  nrm[0] * bin[1] - nrm[1] * bin[0]};
  for (int a = 0; a < 3; a++) {
    f.normal[a] = nrm[a];
    f.binormal[a] = bin[a];
    f.curvature_axis[a] = ax[a];
  }
  f.valid = 1;
  f.majority_cam = majority;
  out[s] = f;
  atomicAdd(num_valid, 1u);
}

__global__ void __launch_bounds__(kFrameThreads) k_frames(const FrameArgs P) {
  extern __shared__ __align__(16) unsigned char smem[];
  if (!P.list) {
    frame_of_sample(P, blockIdx.x, smem);
    return;
  }
  const uint32_t n = *P.list_n;
  for (uint32_t it = blockIdx.x; it < n; it += gridDim.x) {
    frame_of_sample(P, P.list[it], smem);
    __syncthreads();  // thread 0 finishes the frame alone; the shared arrays are reused
  }
}

}  // namespace

int frames_compute(ag2_ctx* c) {
  phase_begin(c, 2);
  cudaStream_t st = c->stream;
  GridDev g = grid_dev(c);
  const uint32_t S = (uint32_t)c->num_samples;
  AG2_CUDA(c, c->frames.reserve(S));
  AG2_CUDA(c, c->kt.reserve(S));
  const int32_t* sidx = c->use_points ? nullptr : c->sample_idx.p;
  const double* sxyz = c->use_points ? c->sample_xyz.p : nullptr;
  AG2_CUDA(c, c->frames_big.reserve(2 * (size_t)std::max(S, 1u) + 2));
  uint32_t* big_n = c->frames_big.p;  // [2] list sizes, then the two lists
  uint32_t* big_list = c->frames_big.p + 2;
  AG2_CUDA(c, cudaMemsetAsync(c->scalars.p + 12, 0, 8, st));  // [12] max k_t, [13] valid frames
  AG2_CUDA(c, cudaMemsetAsync(big_n, 0, 8, st));
  k_count_sample_neighbors<<<(S * 32 + 127) / 128, 128, 0, st>>>(g, c->xyz_in.p, sidx, sxyz, S, c->hc.r2_taubin,
                                                                c->hc.rq_taubin, c->kt.p, c->scalars.p + 12, big_list, big_n);
  AG2_LAUNCH_CHECK(c);
  // the largest neighbourhood only matters for the capacity check: it travels to pinned memory, nobody waits
  AG2_CUDA(c, cudaMemcpyAsync(c->h_pin + 5, c->scalars.p + 12, 4, cudaMemcpyDeviceToHost, st));
  c->pend_frames = true;
  if (!c->frames_attr_set) {
    AG2_CUDA(c, cudaFuncSetAttribute(k_frames, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kFrameCap2 * 12)));
    c->frames_attr_set = true;
  }
  const HandConst& h = c->hc;
  FrameArgs P;
  P.g = g, P.xyz_in = c->xyz_in.p, P.sidx = sidx, P.sxyz = sxyz, P.S = S, P.kt = c->kt.p;
  P.r2 = h.r2_taubin, P.rq = h.rq_taubin, P.num_cams = c->num_cams, P.seed = h.rng_seed;
  P.id_offset = c->sample_id_offset, P.id_stride = c->sample_id_stride;
  P.cam0x = h.cam_origin[0][0], P.cam0y = h.cam_origin[0][1], P.cam0z = h.cam_origin[0][2];
  P.cam1x = h.cam_origin[1][0], P.cam1y = h.cam_origin[1][1], P.cam1z = h.cam_origin[1][2];
  P.out = c->frames.p, P.num_valid = c->scalars.p + 13;
  P.cap = kFrameCap0, P.lo = 0, P.list = nullptr, P.list_n = nullptr;
  k_frames<<<S, kFrameThreads, kFrameSmem0, st>>>(P);
  AG2_LAUNCH_CHECK(c);
  P.cap = kFrameCap1, P.lo = kFrameCap0, P.list = big_list, P.list_n = big_n;
  k_frames<<<std::min<uint32_t>(S, (uint32_t)c->sm_count * 4), kFrameThreads, (size_t)kFrameCap1 * 12, st>>>(P);
  AG2_LAUNCH_CHECK(c);
  P.cap = kFrameCap2, P.lo = kFrameCap1, P.list = big_list + S, P.list_n = big_n + 1;
  k_frames<<<std::min<uint32_t>(S, (uint32_t)c->sm_count), kFrameThreads, (size_t)kFrameCap2 * 12, st>>>(P);
  AG2_LAUNCH_CHECK(c);
  phase_end(c, 2);
  return AG2_OK;
}

}  // namespace ag2
