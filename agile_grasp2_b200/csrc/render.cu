// K7: grasp images, replacing Learning::createGraspImages / createInstance /
// convertToImageRGB (reference learning.cpp:4-33, :221-249, :143-209) and the
// OpenCV tail dilate(3x3) -> cvtColor(BGR2RGB) -> convertTo(CV_8UC3, 255).
// One CTA per pruned hypothesis:
//   1. min over the scaled y of its in-box points, then the image cell of every point, computed
//      with the hand kernel's exact arithmetic: cell = floor(x'/cs) + 60*floor((y'-min y')/cs)
//      (no clamping), and a histogram over the 3600 cells;
//   2. counting sort by cell (prefix over the histogram, scatter with a shared-memory cursor);
//   3. the reference's per-cell double sums run in radiusSearch order (ascending (d2, id)); only the
//      order inside a cell matters, so each (short) run is ordered by counting smaller keys - no
//      sort of the whole closing region;
//   4. one thread per occupied cell sums its run in that order (the reference scans all m points
//      for each of 3600 cells: O(3600 m));
//   5. normalise, abs, float image in shared memory, 3x3 max, channel swap,
//      round-half-even x255 -> u8 HWC (the cv::Mat layout).
#include <math.h>
#include <stdlib.h>

#include <algorithm>
#include <vector>

#include "dev_common.cuh"

namespace ag2 {
namespace {

constexpr int kRenderThreads = 512;  // measured against 256 and 1024 on c2 / c3
constexpr int IS = AG2_IMAGE_SIZE;
constexpr int IP = IS + 2;                       // image with a one-cell zero border (dilate needs no bounds tests)
constexpr size_t kImgBytes = (size_t)IP * IP * 8;   // per cell: u16x2 (out ch 0, out ch 1) | u16x2 (out ch 2, 0)
constexpr size_t kOutBytes = (size_t)AG2_IMAGE_BYTES;
constexpr uint32_t kSmallCap = 4096;             // size class 1: two CTAs per SM

struct RenderArgs {
  GridDev g;
  const uint32_t* list;        // pruned slot ids
  const ag2_hypothesis* slots;
  const BoxRef* box_ref;
  const int32_t* pool;
  const ag2_frame* frames;
  uint32_t cap;                // capacity (multiple of 8) >= largest m of this launch
  uint32_t m_lo, m_hi;         // size class of this launch: images with m outside [m_lo, m_hi] are skipped
  uint32_t n;                  // images
  uint32_t* claim;             // nullptr: CTA b takes images b, b + grid, ...; else CTAs claim images from this counter
  uint8_t* scratch;            // nullptr: sort arrays in shared memory; else gridDim.x regions of cap*16 bytes
                               // in global memory (closing regions too large for shared memory)
  uint8_t* out;                // n * 10800
  // EXPLICIT mode (Learning::createGraspImage(points, normals, grasp), learning.cpp:69-90): image i is
  // drawn from caller-supplied unit-box points / normals [xoff[i], xoff[i+1]), 3 doubles each
  const double* xpts;
  const double* xnrm;
  const uint32_t* xoff;
  int aligns;                  // subtract the minimum y first (always done on the detection path)
};

// Shared-memory plan.  Region 1 holds the 64-bit keys while the runs are ordered and the image
// afterwards; region 2 holds run_pos / ord / run_cell until the cell sums are done and the output
// bytes afterwards.
struct RenderSmem {
  size_t r1, r2, total;
};
__host__ __device__ inline RenderSmem render_smem(uint32_t cap, bool scratch) {
  RenderSmem L;
  const size_t arr = scratch ? 0 : (size_t)cap * 8;
  L.r1 = 0;
  const size_t s1 = arr > kImgBytes ? arr : kImgBytes;
  L.r2 = (s1 + 15) & ~(size_t)15;
  const size_t s2 = arr > kOutBytes ? arr : kOutBytes;
  L.total = L.r2 + ((s2 + 15) & ~(size_t)15);
  return L;
}

__device__ __forceinline__ long long dord(double v) {
  long long i = __double_as_longlong(v);
  return i >= 0 ? i : i ^ 0x7fffffffffffffffll;
}
__device__ __forceinline__ double ordd(long long i) {
  return __longlong_as_double(i >= 0 ? i : i ^ 0x7fffffffffffffffll);
}
// saturate_cast<uchar>(v * 255): float product, round half to even, clamp
__device__ __forceinline__ uint32_t to_u8(double v) {
  const int iv = __float2int_rn(__fmul_rn((float)fabs(v), 255.0f));
  return (uint32_t)min(max(iv, 0), 255);
}

template <bool EXPLICIT>
__global__ void __launch_bounds__(kRenderThreads)
k_render(RenderArgs A, const __grid_constant__ HandConst hc) {
  extern __shared__ __align__(16) unsigned char smem[];
  const RenderSmem L = render_smem(A.cap, A.scratch != nullptr);
  unsigned char* arr = A.scratch ? A.scratch + (size_t)blockIdx.x * A.cap * 16 : nullptr;
  uint64_t* run_key = (uint64_t*)(arr ? arr : smem + L.r1);        // cap: (d2, id) of the point in run slot s
  unsigned char* a2 = arr ? arr + (size_t)A.cap * 8 : smem + L.r2;
  uint32_t* run_pos = (uint32_t*)a2;                               // cap: its sorted position
  uint16_t* cellv = (uint16_t*)(run_pos + A.cap);                  // cap: cell of in-box point t (0xFFFF = none)
  uint16_t* ord = cellv;                                           //      later: run slots, each run in (d2, id) order
  uint16_t* run_cell = cellv + A.cap;                              // cap: cell of run slot s
  uint2* img = (uint2*)(smem + L.r1);                              // IP x IP cells, after the runs are ordered
  uint8_t* outb = smem + L.r2;                                     // 10800 output bytes, after the cell sums
  // per image cell: count, then (first run slot | slots still to hand out << 16); [3600] = end sentinel
  __shared__ uint32_t cellc[IS * IS + 1];
  __shared__ uint32_t warp_tot[kRenderThreads / 32];
  __shared__ long long miny_s;
  __shared__ double FR[3][3];
  __shared__ uint32_t claim_s;
  const int tid = threadIdx.x;
  // next image of this CTA: a fixed stride, or (second size class: few images, very different sizes) the next
  // unclaimed one, so that no CTA is left with several of the large closing regions
  auto next_image = [&](uint32_t cur) -> uint32_t {
    if (!A.claim) return cur + gridDim.x;
    __syncthreads();
    if (tid == 0) claim_s = atomicAdd(A.claim, 1u);
    __syncthreads();
    return claim_s;
  };
  for (uint32_t image = A.claim ? next_image(0) : blockIdx.x; image < A.n; image = next_image(image)) {
  const uint32_t slot = EXPLICIT ? 0u : A.list[image];
  const ag2_hypothesis& hyp = A.slots[slot];
  BoxRef br;
  if (EXPLICIT) {
    br.offset = A.xoff[image];
    br.count = A.xoff[image + 1] - A.xoff[image];
  } else {
    br = A.box_ref[slot];
  }
  const uint32_t m = br.count;
  if (m < A.m_lo || m > A.m_hi) continue;  // another launch's size class
  float q[3] = {0.f, 0.f, 0.f};
  if (!EXPLICIT) {
    const ag2_frame& fr = A.frames[hyp.sample_id];
    q[0] = (float)fr.sample[0], q[1] = (float)fr.sample[1], q[2] = (float)fr.sample[2];
    if (tid < 9) {
      int a = tid / 3, b = tid % 3;  // frame_rot columns: binormal, approach, axis
      FR[a][b] = b == 0 ? hyp.binormal[a] : (b == 1 ? hyp.approach[a] : hyp.axis[a]);
    }
  }
  if (tid == 0) miny_s = dord(INFINITY);
  for (int i = tid; i <= IS * IS; i += kRenderThreads) cellc[i] = 0;
  __syncthreads();
  // hand-frame constants of this hypothesis (same values the hand kernel used)
  const int e = EXPLICIT ? 0 : hyp.finger_slot;
  const double left = hc.fs_hi[e], right = hc.fs_lo[10 + e];
  const double top = EXPLICIT ? 1.0 : hyp.hand_top, bottom = EXPLICIT ? 0.0 : hyp.hand_bottom;
  const double baseline_const = 0.1;
  const double lc = left - 0.5 * (baseline_const - (right - left));
  const double sx = 1.0 / baseline_const, sy = 1.0 / (top - bottom);
  const double cellsize = 1.0 / (double)IS;
  // pass A: min over scaled y (learning.cpp:149)
  double mny = INFINITY;
  for (uint32_t t = tid; t < m; t += kRenderThreads) {
    if (EXPLICIT) {
      mny = fmin(mny, A.xpts[3 * (size_t)(br.offset + t) + 1]);
    } else {
      float4 p = A.g.pts[A.pool[br.offset + t]];
      double px = (double)__fsub_rn(p.x, q[0]), py = (double)__fsub_rn(p.y, q[1]), pz = (double)__fsub_rn(p.z, q[2]);
      double y = (FR[0][1] * px + FR[1][1] * py) + FR[2][1] * pz;
      mny = fmin(mny, sy * (y - bottom));
    }
  }
  mny = warp_min_d(mny);
  if ((tid & 31) == 0) atomicMin(&miny_s, dord(mny));
  __syncthreads();
  const double miny = (EXPLICIT && !A.aligns) ? 0.0 : ordd(miny_s);
  // pass B: cell of every point, histogram over the 3600 cells
  for (uint32_t t = tid; t < m; t += kRenderThreads) {
    uint32_t cell = 0xFFFFu;
    double xs, ys;
    if (EXPLICIT) {
      xs = A.xpts[3 * (size_t)(br.offset + t)];
      ys = A.aligns ? A.xpts[3 * (size_t)(br.offset + t) + 1] - miny : A.xpts[3 * (size_t)(br.offset + t) + 1];
    } else {
      float4 p = A.g.pts[A.pool[br.offset + t]];
      double px = (double)__fsub_rn(p.x, q[0]), py = (double)__fsub_rn(p.y, q[1]), pz = (double)__fsub_rn(p.z, q[2]);
      double x = (FR[0][0] * px + FR[1][0] * py) + FR[2][0] * pz;
      double y = (FR[0][1] * px + FR[1][1] * py) + FR[2][1] * pz;
      xs = sx * (x - lc);
      ys = sy * (y - bottom) - miny;
    }
    // floorVector(pts.row(0) / cellsize): true division, floor, int conversion
    double hc_ = floor(xs / cellsize), vc_ = floor(ys / cellsize);
    // cell_indices = h + v*60 evaluated in int; guard the int conversion range
    if (hc_ > -1e6 && hc_ < 1e6 && vc_ > -1e6 && vc_ < 1e6) {
      int ci = (int)hc_ + (int)vc_ * IS;
      if (ci >= 0 && ci < IS * IS) cell = (uint32_t)ci;  // else never matched
    }
    cellv[t] = (uint16_t)cell;
    if (cell != 0xFFFFu) atomicAdd(&cellc[cell], 1u);
  }
  __syncthreads();
  // exclusive prefix over the cells: kPer consecutive cells per thread, block scan of the partial sums
  {
    constexpr int kPer = (IS * IS + kRenderThreads - 1) / kRenderThreads;
    uint32_t loc[kPer], sum = 0;
#pragma unroll
    for (int k = 0; k < kPer; k++) {
      const int ci = tid * kPer + k;
      loc[k] = ci < IS * IS ? cellc[ci] : 0u;
      sum += loc[k];
    }
    uint32_t incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t v = __shfl_up_sync(AG2_FULL, incl, o);
      if ((tid & 31) >= o) incl += v;
    }
    if ((tid & 31) == 31) warp_tot[tid >> 5] = incl;
    __syncthreads();
    uint32_t base = 0;
    for (int wv = 0; wv < (tid >> 5); wv++) base += warp_tot[wv];
    uint32_t run = base + incl - sum;
#pragma unroll
    for (int k = 0; k < kPer; k++) {
      const int ci = tid * kPer + k;
      if (ci < IS * IS) cellc[ci] = run | (loc[k] << 16);
      run += loc[k];
    }
    if (tid == kRenderThreads - 1) cellc[IS * IS] = run;  // number of binned points
  }
  __syncthreads();
  // scatter into the cell runs (any order inside a run) with the radiusSearch key (d2, id)
  for (uint32_t t = tid; t < m; t += kRenderThreads) {
    const uint32_t cell = cellv[t];
    if (cell == 0xFFFFu) continue;
    const uint32_t old = atomicSub(&cellc[cell], 1u << 16);
    const uint32_t s = (old & 0xFFFFu) + (old >> 16) - 1u;
    if (EXPLICIT) {  // per-cell sums run in the order of the caller's list
      run_key[s] = (uint64_t)t;
      run_pos[s] = br.offset + t;
    } else {
      const uint32_t pos = (uint32_t)A.pool[br.offset + t];
      const float4 p = A.g.pts[pos];
      const float d2 = dist2_f(q[0] - p.x, q[1] - p.y, q[2] - p.z);
      run_key[s] = nn_key(d2, __float_as_uint(p.w));
      run_pos[s] = pos;
    }
    run_cell[s] = (uint16_t)cell;
  }
  __syncthreads();  // cellv is dead from here on: its memory becomes `ord`
  // The reference's per-cell double sums run in radiusSearch order (ascending (d2, id)); only the
  // order INSIDE a cell matters, so each run is ordered by counting smaller keys (runs are short).
  const uint32_t binned = cellc[IS * IS];
  for (uint32_t s = tid; s < binned; s += kRenderThreads) {
    const uint32_t cell = run_cell[s];
    const uint32_t a = cellc[cell] & 0xFFFFu, b = cellc[cell + 1] & 0xFFFFu;
    const uint64_t key = run_key[s];
    uint32_t r = 0;
    for (uint32_t u = a; u < b; u++) r += run_key[u] < key;
    ord[a + r] = (uint16_t)s;
  }
  __syncthreads();  // run_key is dead from here on: its memory (or region 1) becomes the image
  for (int i = tid; i < (int)(kImgBytes / 16); i += kRenderThreads) ((uint4*)img)[i] = make_uint4(0, 0, 0, 0);
  __syncthreads();
  // one thread per occupied cell sums its run in that order (learning.cpp:166-192); the cell value
  // goes through the tail's x255 rounding here (monotone, so it commutes with the 3x3 max)
  for (uint32_t cell = tid; cell < (uint32_t)(IS * IS); cell += kRenderThreads) {
    const uint32_t a = cellc[cell] & 0xFFFFu, b = cellc[cell + 1] & 0xFFFFu;
    if (a == b) continue;
    double ax = 0.0, ay = 0.0, az = 0.0;
    if (EXPLICIT) {
      for (uint32_t u = a; u < b; u++) {
        const double* nv = A.xnrm + 3 * (size_t)run_pos[ord[u]];
        ax += nv[0], ay += nv[1], az += nv[2];
      }
    } else {
      // the sums are sequential (their order is the reference's), the gathers need not be: four normals in flight
      for (uint32_t u0 = a; u0 < b; u0 += 4) {
        float4 nv[4];
#pragma unroll
        for (int k = 0; k < 4; k++) nv[k] = A.g.nrm[run_pos[ord[min(u0 + k, b - 1)]]];
#pragma unroll
        for (int k = 0; k < 4; k++) {
          if (u0 + k < b) {
            const double n0 = (double)nv[k].x, n1 = (double)nv[k].y, n2 = (double)nv[k].z;
            // normals_rot = frame_rot^T * normal (hand_search.cpp:360)
            ax += (FR[0][0] * n0 + FR[1][0] * n1) + FR[2][0] * n2;
            ay += (FR[0][1] * n0 + FR[1][1] * n1) + FR[2][1] * n2;
            az += (FR[0][2] * n0 + FR[1][2] * n1) + FR[2][2] * n2;
          }
        }
      }
    }
    double s = 1.0 / sqrt(ax * ax + (ay * ay + az * az));
    int row = (int)cell / IS, col = (int)cell % IS;
    // BGR2RGB: output channel k holds source channel 2 - k
    img[(IS - 1 - row + 1) * IP + col + 1] = make_uint2(to_u8(s * az) | (to_u8(s * ay) << 16), to_u8(s * ax));
  }
  __syncthreads();  // run_pos / ord / run_cell are dead from here on: region 2 becomes the output bytes
  // dilate 3x3 (in-bounds max; values are >= 0, so the zero border is neutral) -> u8 HWC
  for (int pix = tid; pix < IS * IS; pix += kRenderThreads) {
    const int r = pix / IS, c = pix % IS;
    const uint2* p0 = img + r * IP + c;
    uint2 m0, m1, m2;
    {
      const uint2 a = p0[0], b = p0[1], d = p0[2];
      m0 = make_uint2(__vimax3_u16x2(a.x, b.x, d.x), __vimax3_u16x2(a.y, b.y, d.y));
    }
    {
      const uint2 a = p0[IP], b = p0[IP + 1], d = p0[IP + 2];
      m1 = make_uint2(__vimax3_u16x2(a.x, b.x, d.x), __vimax3_u16x2(a.y, b.y, d.y));
    }
    {
      const uint2 a = p0[2 * IP], b = p0[2 * IP + 1], d = p0[2 * IP + 2];
      m2 = make_uint2(__vimax3_u16x2(a.x, b.x, d.x), __vimax3_u16x2(a.y, b.y, d.y));
    }
    const uint32_t w0 = __vimax3_u16x2(m0.x, m1.x, m2.x), w1 = __vimax3_u16x2(m0.y, m1.y, m2.y);
    outb[3 * pix] = (uint8_t)(w0 & 0xFFu);
    outb[3 * pix + 1] = (uint8_t)(w0 >> 16);
    outb[3 * pix + 2] = (uint8_t)(w1 & 0xFFu);
  }
  __syncthreads();
  uint4* out = (uint4*)(A.out + (size_t)image * AG2_IMAGE_BYTES);  // 10800 = 675 x 16, base 256-byte aligned
  for (int i = tid; i < (int)(kOutBytes / 16); i += kRenderThreads) out[i] = ((const uint4*)outb)[i];
  __syncthreads();  // every region is reused by the next image of this CTA
  }
}

}  // namespace

// Launches k_render for n images whose largest point set has max_points points.
template <bool EXPLICIT>
static int launch_render(ag2_ctx* c, RenderArgs A, size_t n, size_t max_points) {
  if (max_points >= 65536) {
    c->err = "render: more than 65535 points in a closing region";
    return AG2_E_CAPACITY;
  }
  static_assert(kOutBytes % 16 == 0 && kImgBytes % 16 == 0, "vector copies");
  const uint32_t cap_max = (std::max<uint32_t>((uint32_t)max_points, 8u) + 7u) & ~7u;
  // test hooks: AG2_RENDER_SPILL=1 forces the global-memory path, AG2_RENDER_SMALL_CAP=<n> the boundary
  // between the two size classes (small scenes then exercise both launches)
  const bool force_spill = getenv("AG2_RENDER_SPILL") && atoi(getenv("AG2_RENDER_SPILL")) != 0;
  uint32_t small_cap = kSmallCap;
  if (const char* e = getenv("AG2_RENDER_SMALL_CAP")) small_cap = (std::max(atoi(e), 8) + 7) & ~7;
  A.g = grid_dev(c);
  A.n = (uint32_t)n;
  auto launch = [&](cudaStream_t st, uint32_t cap, uint32_t m_lo, uint32_t m_hi, bool one_cta_per_image) -> int {
    auto kernel = k_render<EXPLICIT>;
    // Closing regions of more than ~12 000 points (fused multi-view clouds, c4) do not fit shared
    // memory: the sort arrays then live in global memory (L2 resident).
    const bool spill = force_spill || render_smem(cap, false).total > 180 * 1024;
    const size_t smem = render_smem(cap, spill).total;
    unsigned grid = (unsigned)n;
    if (spill || !one_cta_per_image) grid = (unsigned)std::min<size_t>(n, (size_t)c->sm_count * (spill ? 4 : 2));
    if (spill) AG2_CUDA(c, c->render_scratch.reserve((size_t)grid * cap * 16));
    AG2_CUDA(c, cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    A.cap = cap, A.m_lo = m_lo, A.m_hi = m_hi;
    A.scratch = spill ? c->render_scratch.p : nullptr;
    A.claim = nullptr;
    if (!one_cta_per_image) {
      A.claim = c->scalars.p + 62;
      AG2_CUDA(c, cudaMemsetAsync(A.claim, 0, 4, st));
    }
    kernel<<<grid, kRenderThreads, smem, st>>>(A, c->hc);
    AG2_LAUNCH_CHECK(c);
    return AG2_OK;
  };
  // Shared memory per CTA is sized by the LARGEST image of a launch, so one dense closing region
  // would cost every image its occupancy: images up to kSmallCap points run one CTA each (two CTAs =
  // 32 warps per SM), the larger ones in a second launch whose CTAs claim images from a counter and
  // skip the rest.  Boundary measured on c3: 2048 -> 1.41 ms, 3072 -> 1.28, 4096 -> 1.06, 5400 -> 1.07.
  // The second class is few, long images: it starts first, on an auxiliary stream, and the first
  // class fills the shared memory and issue slots it leaves free.
  if (cap_max <= small_cap) return launch(c->stream, cap_max, 0u, 0xFFFFFFFFu, true);
  AG2_CUDA(c, cudaEventRecord(c->ev_fork, c->stream));
  AG2_CUDA(c, cudaStreamWaitEvent(c->aux[0], c->ev_fork, 0));
  AG2_TRY(launch(c->aux[0], cap_max, small_cap + 1u, 0xFFFFFFFFu, false));
  AG2_CUDA(c, cudaEventRecord(c->ev_join[0], c->aux[0]));
  AG2_TRY(launch(c->stream, small_cap, 0u, small_cap, true));
  AG2_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_join[0], 0));
  return AG2_OK;
}

// Renders the hypotheses in slots d_list[0..n) into d_out (n * 10800 bytes).
int render_list(ag2_ctx* c, const uint32_t* d_list, size_t n, uint8_t* d_out) {
  if (n == 0) return AG2_OK;
  RenderArgs A = {};
  A.list = d_list;
  A.slots = c->slots.p;
  A.box_ref = c->box_ref.p;
  A.pool = c->box_pool.p;
  A.frames = c->frames.p;
  A.out = d_out;
  return launch_render<false>(c, A, n, c->max_in_box);
}

int render_compute(ag2_ctx* c) {
  phase_begin(c, 5);
  const size_t n = c->list_n[AG2_LIST_PRUNED];
  AG2_CUDA(c, c->images.reserve(std::max<size_t>(n, 1) * AG2_IMAGE_BYTES));
  AG2_TRY(render_list(c, c->list[AG2_LIST_PRUNED].p, n, c->images.p));
  c->have_images = true;
  phase_end(c, 5);
  return AG2_OK;
}

}  // namespace ag2

// Learning::createGraspImages (learning.cpp:4-33) for caller-chosen hypotheses of the last
// ag2_generate_hypotheses call, addressed by (sample_id, orientation_id).
extern "C" int ag2_render_hypotheses(ag2_ctx* c, const int32_t* sample_ids, const int32_t* orientation_ids, size_t n,
                                     uint8_t* out) {
  using namespace ag2;
  if (!c || (n && (!sample_ids || !orientation_ids || !out))) return AG2_E_INVALID;
  if (cudaSetDevice(c->device) != cudaSuccess) return AG2_E_NO_DEVICE;
  if (!c->have_hyps) return AG2_E_STATE;
  if (n == 0) return AG2_OK;
  const int R = c->p.num_orientations;
  std::vector<uint32_t> slots(n);
  for (size_t i = 0; i < n; i++) {
    if (sample_ids[i] < 0 || (size_t)sample_ids[i] >= c->num_samples || orientation_ids[i] < 0 || orientation_ids[i] >= R)
      return AG2_E_INVALID;
    slots[i] = (uint32_t)sample_ids[i] * (uint32_t)R + (uint32_t)orientation_ids[i];
  }
  // only emitted hypotheses own a closing region
  std::vector<ag2_sweep_record> sw(c->num_samples * (size_t)R);
  AG2_CUDA(c, cudaMemcpyAsync(sw.data(), c->sweeps.p, sw.size() * sizeof(ag2_sweep_record), cudaMemcpyDeviceToHost,
                              c->stream));
  AG2_CUDA(c, cudaStreamSynchronize(c->stream));
  for (size_t i = 0; i < n; i++)
    if (!sw[slots[i]].emitted) return AG2_E_INVALID;
  DevBuf<uint32_t> d_list;
  DevBuf<uint8_t> d_img;
  AG2_CUDA(c, d_list.reserve(n));
  AG2_CUDA(c, d_img.reserve(n * AG2_IMAGE_BYTES));
  AG2_CUDA(c, cudaMemcpyAsync(d_list.p, slots.data(), n * 4, cudaMemcpyHostToDevice, c->stream));
  AG2_TRY(render_list(c, d_list.p, n, d_img.p));
  AG2_CUDA(c, cudaMemcpyAsync(out, d_img.p, n * AG2_IMAGE_BYTES, cudaMemcpyDeviceToHost, c->stream));
  AG2_CUDA(c, cudaStreamSynchronize(c->stream));
  return AG2_OK;
}

// Learning::createGraspImage(points, normals, grasp) (learning.cpp:69-90; convertToImageRGB :143-209 with
// aligns = false, then convertTo(CV_8UC3, 255)): images of caller-supplied unit-box point sets.  This is
// the image generator of the ClassificationNode service (classification_node.cpp:75-141).
extern "C" int ag2_render_point_sets(ag2_ctx* c, const double* pts, const double* normals, const uint32_t* offsets,
                                     size_t n_images, int aligns, uint8_t* out) {
  using namespace ag2;
  if (!c || !offsets || (n_images && !out)) return AG2_E_INVALID;
  if (n_images == 0) return AG2_OK;
  const size_t total = offsets[n_images];
  size_t max_points = 0;
  for (size_t i = 0; i < n_images; i++) {
    if (offsets[i + 1] < offsets[i]) return AG2_E_INVALID;
    max_points = std::max<size_t>(max_points, offsets[i + 1] - offsets[i]);
  }
  if (total && (!pts || !normals)) return AG2_E_INVALID;
  if (cudaSetDevice(c->device) != cudaSuccess) return AG2_E_NO_DEVICE;
  DevBuf<double> d_pts, d_nrm;
  DevBuf<uint32_t> d_off;
  DevBuf<uint8_t> d_img;
  AG2_CUDA(c, d_pts.reserve(std::max<size_t>(3 * total, 1)));
  AG2_CUDA(c, d_nrm.reserve(std::max<size_t>(3 * total, 1)));
  AG2_CUDA(c, d_off.reserve(n_images + 1));
  AG2_CUDA(c, d_img.reserve(n_images * AG2_IMAGE_BYTES));
  if (total) {
    AG2_CUDA(c, cudaMemcpyAsync(d_pts.p, pts, 24 * total, cudaMemcpyHostToDevice, c->stream));
    AG2_CUDA(c, cudaMemcpyAsync(d_nrm.p, normals, 24 * total, cudaMemcpyHostToDevice, c->stream));
  }
  AG2_CUDA(c, cudaMemcpyAsync(d_off.p, offsets, 4 * (n_images + 1), cudaMemcpyHostToDevice, c->stream));
  RenderArgs A = {};
  A.slots = c->slots.p;  // unused in this mode
  A.xpts = d_pts.p, A.xnrm = d_nrm.p, A.xoff = d_off.p;
  A.aligns = aligns;
  A.out = d_img.p;
  // the kernel only reads GridDev fields on the hypothesis path; a context without a cloud is fine
  int rc = launch_render<true>(c, A, n_images, max_points);
  if (rc != AG2_OK) return rc;
  AG2_CUDA(c, cudaMemcpyAsync(out, d_img.p, n_images * AG2_IMAGE_BYTES, cudaMemcpyDeviceToHost, c->stream));
  AG2_CUDA(c, cudaStreamSynchronize(c->stream));
  return AG2_OK;
}
