// List building around the hand kernel: concatenation in (sample, orientation)
// order (hand_search.cpp:221-228), S13 prune (grasp_detector.cpp:363-395),
// mode filters (:170-221) and top-k selection (:238-252).  Integer / compare
// work over S*R slots: flag -> exclusive scan -> ordered compaction.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "dev_common.cuh"

namespace ag2 {
namespace {

constexpr int kT = 256;

__global__ void k_flag_emitted(const ag2_sweep_record* __restrict__ sw, size_t n, uint32_t* __restrict__ flag) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i < n) flag[i] = sw[i].emitted ? 1u : 0u;
}

struct PruneArgs {
  double half_width;  // 0.5 * outer_diameter_
  double min_aperture, max_aperture;
  float min_x, max_x, min_y, max_y, min_z;  // float arguments of pruneGraspsOnHandParameters
  int filter_half, enabled;
};

__device__ __forceinline__ bool prune_keep(const ag2_hypothesis& g, const PruneArgs& a) {
  if (!a.enabled) return true;
  if (a.filter_half && !g.half_antipodal) return false;
  double mn[3], mx[3];
  for (int d = 0; d < 3; d++) {
    double v0 = g.bottom[d] + a.half_width * g.binormal[d];
    double v1 = g.bottom[d] - a.half_width * g.binormal[d];
    double v2 = g.top[d] + a.half_width * g.binormal[d];
    double v3 = g.top[d] - a.half_width * g.binormal[d];
    double v4 = g.bottom[d] - 0.10 * g.approach[d];
    mn[d] = fmin(fmin(fmin(v0, v1), fmin(v2, v3)), v4);
    mx[d] = fmax(fmax(fmax(v0, v1), fmax(v2, v3)), v4);
  }
  double ap = g.width;
  return ap >= a.min_aperture && ap <= a.max_aperture && mn[2] >= (double)a.min_z && mn[1] >= (double)a.min_y &&
         mx[1] <= (double)a.max_y && mn[0] >= (double)a.min_x && mx[0] <= (double)a.max_x;
}

__global__ void k_flag_pruned(const ag2_sweep_record* __restrict__ sw, const ag2_hypothesis* __restrict__ slots,
                              size_t n, PruneArgs a, uint32_t* __restrict__ flag) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  flag[i] = (sw[i].emitted && prune_keep(slots[i], a)) ? 1u : 0u;
}

// mode: 0 keep all of `src`, 2 keep full_antipodal, 1 keep score >= thr (and mark)
__device__ __forceinline__ bool mode_keep(ag2_hypothesis& g, const float* __restrict__ logits, size_t i, int mode,
                                          double thr) {
  if (mode == AG2_MODE_GEOMETRIC) return g.full_antipodal != 0;
  if (mode == AG2_MODE_PREDICTION) {
    // grasp_detector.cpp:200: float difference stored in a double
    double score = (double)__fsub_rn(logits[2 * i + 1], logits[2 * i]);
    if (score >= thr) {
      g.score = score;
      g.full_antipodal = 1;
      return true;
    }
    return false;
  }
  return true;
}

__global__ void k_flag_mode(const uint32_t* __restrict__ src, size_t n, ag2_hypothesis* __restrict__ slots,
                            const float* __restrict__ logits, int mode, double thr, uint32_t* __restrict__ flag) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  flag[i] = mode_keep(slots[src[i]], logits, i, mode, thr) ? 1u : 0u;
}

// out[scan[i]] = src ? src[i] : i   for flagged i
__global__ void k_compact(const uint32_t* __restrict__ flag_scan, const uint32_t* __restrict__ src, size_t n,
                          uint32_t* __restrict__ out) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t a = flag_scan[i], b = flag_scan[i + 1];
  if (b != a) out[a] = src ? src[i] : (uint32_t)i;
}

__global__ void k_gather(const uint32_t* __restrict__ list, size_t n, const ag2_hypothesis* __restrict__ slots,
                         ag2_hypothesis* __restrict__ out) {
  // 192-byte records moved as 12 x 16 B per record
  size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  size_t rec = t / 12, part = t % 12;
  if (rec >= n) return;
  const uint4* s = reinterpret_cast<const uint4*>(&slots[list[rec]]);
  reinterpret_cast<uint4*>(&out[rec])[part] = s[part];
}

// complemented order-preserving code of a double: ascending key == descending score
__device__ __forceinline__ uint64_t score_key(double score) {
  long long o = __double_as_longlong(score);
  unsigned long long u = o >= 0 ? ((unsigned long long)o | 0x8000000000000000ull) : ~(unsigned long long)o;
  return ~u;
}

__global__ void k_score_keys(const uint32_t* __restrict__ list, size_t n, const ag2_hypothesis* __restrict__ slots,
                             uint64_t* __restrict__ keys) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  // descending score, ties keep list order (documented deviation D4: partial_sort is unstable)
  keys[i] = score_key(slots[list ? list[i] : (uint32_t)i].score);
}

// ---- ranking by (key asc, position asc): key = complemented order-preserving code of the
// score, so ascending key == descending score; equal scores keep list order (D4).
__device__ __forceinline__ bool kv_greater(uint64_t ka, uint32_t va, uint64_t kb, uint32_t vb) {
  return ka > kb || (ka == kb && va > vb);
}

constexpr int kSortTile = 1024;  // (8 + 4) B * 1024 = 12 KB of shared memory; a short network per tile, tiles side by side

// Sorts one tile of (key, pos) pairs in shared memory and writes its first `keep` pairs.
__global__ void __launch_bounds__(1024)
k_sort_tile(const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ pos_in, uint32_t n, uint32_t keep,
            uint64_t* __restrict__ keys_out, uint32_t* __restrict__ pos_out) {
  __shared__ uint64_t k_s[kSortTile];
  __shared__ uint32_t v_s[kSortTile];
  const uint32_t base = blockIdx.x * kSortTile;
  const uint32_t cnt = min((uint32_t)kSortTile, n - base);
  uint32_t n_pad = 2;
  while (n_pad < cnt) n_pad <<= 1;
  for (uint32_t t = threadIdx.x; t < n_pad; t += blockDim.x) {
    k_s[t] = t < cnt ? keys_in[base + t] : ~0ull;
    v_s[t] = t < cnt ? (pos_in ? pos_in[base + t] : base + t) : 0xFFFFFFFFu;
  }
  __syncthreads();
  for (uint32_t k = 2; k <= n_pad; k <<= 1)
    for (uint32_t j = k >> 1; j > 0; j >>= 1) {
      for (uint32_t t = threadIdx.x; t < (n_pad >> 1); t += blockDim.x) {
        uint32_t i = ((t & ~(j - 1)) << 1) | (t & (j - 1)), l = i | j;
        bool up = ((i & k) == 0);
        uint64_t ka = k_s[i], kb = k_s[l];
        uint32_t va = v_s[i], vb = v_s[l];
        if (kv_greater(ka, va, kb, vb) == up) {
          k_s[i] = kb, k_s[l] = ka;
          v_s[i] = vb, v_s[l] = va;
        }
      }
      __syncthreads();
    }
  const uint32_t m = min(keep, cnt);
  const uint32_t obase = blockIdx.x * keep;  // full tiles precede the ragged one
  for (uint32_t t = threadIdx.x; t < m; t += blockDim.x) {
    keys_out[obase + t] = k_s[t];
    pos_out[obase + t] = v_s[t];
  }
}

// One compare-exchange step of a global-memory bitonic sort (lists larger than a tile that
// must be sorted completely).
__global__ void k_bitonic_step(uint64_t* __restrict__ keys, uint32_t* __restrict__ pos, uint32_t n_pad, uint32_t k,
                               uint32_t j) {
  uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (n_pad >> 1)) return;
  uint32_t i = ((t & ~(j - 1)) << 1) | (t & (j - 1)), l = i | j;
  bool up = ((i & k) == 0);
  uint64_t ka = keys[i], kb = keys[l];
  uint32_t va = pos[i], vb = pos[l];
  if (kv_greater(ka, va, kb, vb) == up) {
    keys[i] = kb, keys[l] = ka;
    pos[i] = vb, pos[l] = va;
  }
}

__global__ void k_pad_keys(uint64_t* __restrict__ keys, uint32_t* __restrict__ pos, uint32_t n, uint32_t n_pad) {
  uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_pad) return;
  if (t >= n) keys[t] = ~0ull;
  pos[t] = t < n ? t : 0xFFFFFFFFu;
}

__global__ void k_take_ranked_rec(const uint32_t* __restrict__ pos, uint32_t k, const ag2_hypothesis* __restrict__ src,
                                  ag2_hypothesis* __restrict__ out) {
  uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t rec = t / 12, part = t % 12;
  if (rec >= k) return;
  reinterpret_cast<uint4*>(&out[rec])[part] = reinterpret_cast<const uint4*>(&src[pos[rec]])[part];
}

__global__ void k_take_ranked(const uint32_t* __restrict__ pos, uint32_t k, const uint32_t* __restrict__ list,
                              uint32_t* __restrict__ out) {
  uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < k) out[t] = list[pos[t]];
}

// ---- small lists: one CTA does flag -> scan -> ordered compaction in a single launch ----------
// With S*R in the thousands the multi-kernel path above is nothing but launch latency (five
// launches and a host round trip per list).
constexpr int kFT = 1024;
constexpr size_t kFusedMaxSlots = 1u << 16;  // S*R handled by the single-CTA kernel after the hand sweep
constexpr uint32_t kFusedMaxSelect = 4096;   // pruned hands ranked inside one CTA's shared memory

// exclusive position of this thread's flag among the CTA's flags; `total` = number of flags set
__device__ __forceinline__ uint32_t block_flag_scan(bool f, uint32_t* warp_sums /* [33] */, uint32_t& total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t ballot = __ballot_sync(AG2_FULL, f);
  if (lane == 0) warp_sums[warp] = __popc(ballot);
  __syncthreads();
  if (warp == 0) {
    const uint32_t v = warp_sums[lane];
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t t = __shfl_up_sync(AG2_FULL, incl, o);
      if (lane >= o) incl += t;
    }
    warp_sums[lane] = incl - v;
    if (lane == 31) warp_sums[32] = incl;
  }
  __syncthreads();
  const uint32_t r = warp_sums[warp] + __popc(ballot & ((1u << lane) - 1));
  total = warp_sums[32];
  __syncthreads();
  return r;
}

// hypotheses list (emitted slots) and pruned list (emitted slots that pass S13) in slot order.
// Thread t flags slots t, t + 1024, ... (coalesced), keeps its flags as bit masks, and ONE scan over
// the (tile, warp) ballot counts places every slot of every tile.
constexpr int kListTiles = (int)(kFusedMaxSlots / kFT);  // <= 64: one bit per tile in a 64-bit mask
__global__ void __launch_bounds__(kFT)
k_lists_after_hands(const ag2_sweep_record* __restrict__ sw, const ag2_hypothesis* __restrict__ slots, uint32_t n,
                    PruneArgs a, uint32_t* __restrict__ hyp_out, uint32_t* __restrict__ pruned_out,
                    uint32_t* __restrict__ counts /* [2] */) {
  __shared__ uint32_t wc_e[kListTiles * 32], wc_k[kListTiles * 32];  // per (tile, warp): emitted / kept
  __shared__ uint32_t wt_e[33], wt_k[33];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t tiles = (n + kFT - 1) / kFT;
  unsigned long long em_bits = 0, keep_bits = 0;
  for (uint32_t k = 0; k < tiles; k++) {
    const uint32_t i = k * kFT + threadIdx.x;
    const bool em = i < n && sw[i].emitted;
    const bool keep = em && prune_keep(slots[i], a);
    em_bits |= (unsigned long long)em << k;
    keep_bits |= (unsigned long long)keep << k;
    const uint32_t be = __ballot_sync(AG2_FULL, em), bk = __ballot_sync(AG2_FULL, keep);
    if (lane == 0) wc_e[k * 32 + warp] = __popc(be), wc_k[k * 32 + warp] = __popc(bk);
  }
  __syncthreads();
  // exclusive scan of the tiles * 32 counts: two consecutive entries per thread
  {
    const uint32_t m = tiles * 32, t0 = 2 * threadIdx.x;
    const uint32_t e0 = t0 < m ? wc_e[t0] : 0u, e1 = t0 + 1 < m ? wc_e[t0 + 1] : 0u;
    const uint32_t k0 = t0 < m ? wc_k[t0] : 0u, k1 = t0 + 1 < m ? wc_k[t0 + 1] : 0u;
    uint32_t ie = e0 + e1, ik = k0 + k1;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t te = __shfl_up_sync(AG2_FULL, ie, o), tk = __shfl_up_sync(AG2_FULL, ik, o);
      if (lane >= o) ie += te, ik += tk;
    }
    if (lane == 31) wt_e[warp] = ie, wt_k[warp] = ik;
    __syncthreads();
    if (warp == 0) {
      const uint32_t ve = wt_e[lane], vk = wt_k[lane];
      uint32_t se = ve, sk = vk;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t te = __shfl_up_sync(AG2_FULL, se, o), tk = __shfl_up_sync(AG2_FULL, sk, o);
        if (lane >= o) se += te, sk += tk;
      }
      wt_e[lane] = se - ve, wt_k[lane] = sk - vk;
      if (lane == 31) wt_e[32] = se, wt_k[32] = sk;
    }
    __syncthreads();
    const uint32_t xe = wt_e[warp] + ie - (e0 + e1), xk = wt_k[warp] + ik - (k0 + k1);
    if (t0 < m) wc_e[t0] = xe, wc_k[t0] = xk;
    if (t0 + 1 < m) wc_e[t0 + 1] = xe + e0, wc_k[t0 + 1] = xk + k0;
  }
  __syncthreads();
  for (uint32_t k = 0; k < tiles; k++) {
    const bool em = (em_bits >> k) & 1ull, keep = (keep_bits >> k) & 1ull;
    const uint32_t be = __ballot_sync(AG2_FULL, em), bk = __ballot_sync(AG2_FULL, keep);
    const uint32_t lt = (1u << lane) - 1u, i = k * kFT + threadIdx.x;
    if (em) hyp_out[wc_e[k * 32 + warp] + __popc(be & lt)] = i;
    if (keep) pruned_out[wc_k[k * 32 + warp] + __popc(bk & lt)] = i;
  }
  if (threadIdx.x == 0) counts[0] = wt_e[32], counts[1] = wt_k[32];
}

// mode filter over the pruned list -> antipodal list, then its top-k by (score desc, position asc)
// -> selected list.  n <= kFusedMaxSelect.
__global__ void __launch_bounds__(kFT)
k_select_small(const uint32_t* __restrict__ src, uint32_t n, ag2_hypothesis* __restrict__ slots,
               const float* __restrict__ logits, int mode, double thr, int num_selected,
               uint32_t* __restrict__ anti_out, uint32_t* __restrict__ sel_out, uint32_t* __restrict__ counts /* [2] */) {
  extern __shared__ __align__(16) unsigned char sel_smem[];
  uint64_t* k_s = reinterpret_cast<uint64_t*>(sel_smem);
  uint32_t* v_s = reinterpret_cast<uint32_t*>(sel_smem + (size_t)kFusedMaxSelect * 8);
  uint32_t* ws = v_s + kFusedMaxSelect;
  uint32_t n_a = 0;
  for (uint32_t t0 = 0; t0 < n; t0 += kFT) {
    const uint32_t i = t0 + threadIdx.x;
    bool keep = false;
    uint32_t slot = 0;
    if (i < n) {
      slot = src[i];
      keep = mode_keep(slots[slot], logits, i, mode, thr);
    }
    uint32_t tot;
    const uint32_t r = block_flag_scan(keep, ws, tot);
    if (keep) {
      anti_out[n_a + r] = slot;
      k_s[n_a + r] = score_key(slots[slot].score);
      v_s[n_a + r] = n_a + r;
    }
    n_a += tot;
  }
  uint32_t n_pad = 2;
  while (n_pad < n_a) n_pad <<= 1;
  for (uint32_t t = n_a + threadIdx.x; t < n_pad; t += kFT) k_s[t] = ~0ull, v_s[t] = 0xFFFFFFFFu;
  __syncthreads();
  for (uint32_t k = 2; k <= n_pad; k <<= 1)
    for (uint32_t j = k >> 1; j > 0; j >>= 1) {
      for (uint32_t t = threadIdx.x; t < (n_pad >> 1); t += kFT) {
        uint32_t i = ((t & ~(j - 1)) << 1) | (t & (j - 1)), l = i | j;
        bool up = ((i & k) == 0);
        uint64_t ka = k_s[i], kb = k_s[l];
        uint32_t va = v_s[i], vb = v_s[l];
        if (kv_greater(ka, va, kb, vb) == up) {
          k_s[i] = kb, k_s[l] = ka;
          v_s[i] = vb, v_s[l] = va;
        }
      }
      __syncthreads();
    }
  const uint32_t k = (num_selected > 0 && (uint32_t)num_selected < n_a) ? (uint32_t)num_selected : n_a;
  for (uint32_t t = threadIdx.x; t < k; t += kFT) sel_out[t] = anti_out[v_s[t]];
  if (threadIdx.x == 0) counts[0] = n_a, counts[1] = k;
}

}  // namespace

// flags (n entries, 0/1) -> compacted `out`; the count goes to pinned memory (*h_count is valid after the next
// synchronisation of the stream)
static int compact_list_async(ag2_ctx* c, size_t n, const uint32_t* src, DevBuf<uint32_t>& out, uint32_t* h_count) {
  cudaStream_t st = c->stream;
  AG2_CUDA(c, out.reserve(std::max<size_t>(n, 1)));
  *h_count = 0;
  if (n == 0) return AG2_OK;
  AG2_TRY(exclusive_scan_u32(c, c->flags.p, n));
  k_compact<<<(unsigned)((n + kT - 1) / kT), kT, 0, st>>>(c->flags.p, src, n, out.p);
  AG2_LAUNCH_CHECK(c);
  AG2_CUDA(c, cudaMemcpyAsync(h_count, c->flags.p + n, 4, cudaMemcpyDeviceToHost, st));
  return AG2_OK;
}
static int compact_list(ag2_ctx* c, size_t n, const uint32_t* src, DevBuf<uint32_t>& out, size_t* count) {
  AG2_TRY(compact_list_async(c, n, src, out, c->h_pin + 30));
  AG2_TRY(sync_and_finish(c));
  *count = c->h_pin[30];
  return AG2_OK;
}

static PruneArgs prune_args(const ag2_ctx* c) {
  PruneArgs a;
  a.half_width = 0.5 * c->p.hand_outer_diameter;
  a.min_aperture = c->p.min_aperture;
  a.max_aperture = c->p.max_aperture;
  a.min_x = (float)c->p.workspace[0];
  a.max_x = (float)c->p.workspace[1];
  a.min_y = (float)c->p.workspace[2];
  a.max_y = (float)c->p.workspace[3];
  a.min_z = c->min_z;
  a.filter_half = c->p.filter_half_grasps;
  a.enabled = c->p.prune;
  return a;
}

// test hook: AG2_LISTS_MULTIKERNEL=1 forces the multi-kernel list path
static bool fused_lists_allowed() {
  const char* e = getenv("AG2_LISTS_MULTIKERNEL");
  return !(e && atoi(e) != 0);
}

// h_pin[24] hypotheses, [25] pruned (single-launch path), [26] pruned (multi-kernel path)
int lists_launch(ag2_ctx* c) {
  const size_t SR = c->num_samples * (size_t)c->p.num_orientations;
  c->pruned_precomputed = c->selected_precomputed = false;
  for (int i = 0; i < 4; i++) c->list_rec_valid[i] = false;
  c->h_pin[24] = c->h_pin[25] = 0;
  if (SR && SR <= kFusedMaxSlots && fused_lists_allowed()) {
    // both lists in one launch; the prune stage then only publishes its count
    cudaStream_t st = c->stream;
    AG2_CUDA(c, c->list[AG2_LIST_HYPOTHESES].reserve(SR));
    AG2_CUDA(c, c->list[AG2_LIST_PRUNED].reserve(SR));
    k_lists_after_hands<<<1, kFT, 0, st>>>(c->sweeps.p, c->slots.p, (uint32_t)SR, prune_args(c),
                                           c->list[AG2_LIST_HYPOTHESES].p, c->list[AG2_LIST_PRUNED].p,
                                           c->scalars.p + 56);
    AG2_LAUNCH_CHECK(c);
    AG2_CUDA(c, cudaMemcpyAsync(c->h_pin + 24, c->scalars.p + 56, 8, cudaMemcpyDeviceToHost, st));
    c->pruned_precomputed = true;
    return AG2_OK;
  }
  AG2_CUDA(c, c->flags.reserve(SR + 1));
  if (SR) {
    k_flag_emitted<<<(unsigned)((SR + kT - 1) / kT), kT, 0, c->stream>>>(c->sweeps.p, SR, c->flags.p);
    AG2_LAUNCH_CHECK(c);
  }
  return compact_list_async(c, SR, nullptr, c->list[AG2_LIST_HYPOTHESES], c->h_pin + 24);
}

int lists_finish(ag2_ctx* c) {
  c->list_n[AG2_LIST_HYPOTHESES] = c->h_pin[24];
  c->st.num_hypotheses = (int64_t)c->h_pin[24];
  if (c->pruned_precomputed) c->pruned_count = c->h_pin[25];
  return AG2_OK;
}

int lists_build_after_hands(ag2_ctx* c) {
  AG2_TRY(lists_launch(c));
  AG2_TRY(sync_and_finish(c));
  return lists_finish(c);
}

int prune_launch(ag2_ctx* c) {
  phase_begin(c, 4);
  const size_t SR = c->num_samples * (size_t)c->p.num_orientations;
  c->selected_precomputed = false;
  if (!c->pruned_precomputed) {
    AG2_CUDA(c, c->flags.reserve(SR + 1));
    if (SR) {
      k_flag_pruned<<<(unsigned)((SR + kT - 1) / kT), kT, 0, c->stream>>>(c->sweeps.p, c->slots.p, SR, prune_args(c),
                                                                         c->flags.p);
      AG2_LAUNCH_CHECK(c);
    }
    AG2_TRY(compact_list_async(c, SR, nullptr, c->list[AG2_LIST_PRUNED], c->h_pin + 26));
  }
  phase_end(c, 4);
  return AG2_OK;
}

int prune_finish(ag2_ctx* c) {
  c->list_n[AG2_LIST_PRUNED] = c->pruned_precomputed ? c->pruned_count : (size_t)c->h_pin[26];
  c->list_rec_valid[AG2_LIST_PRUNED] = false;
  c->st.num_pruned = (int64_t)c->list_n[AG2_LIST_PRUNED];
  return AG2_OK;
}

int prune_compute(ag2_ctx* c) {
  AG2_TRY(prune_launch(c));
  AG2_TRY(sync_and_finish(c));
  return prune_finish(c);
}

// antipodal list from the pruned list (mode 1 needs logits of the pruned list)
int select_mode(ag2_ctx* c) {
  const size_t n = c->list_n[AG2_LIST_PRUNED];
  c->selected_precomputed = false;
  const bool clustered = c->p.min_inliers > 0 && c->p.antipodal_mode != AG2_MODE_NONE;
  if (n && n <= kFusedMaxSelect && !clustered && c->p.antipodal_mode != AG2_MODE_NONE && fused_lists_allowed()) {
    // filter + rank + take in one launch; select_top then only publishes the count
    cudaStream_t st = c->stream;
    AG2_CUDA(c, c->list[AG2_LIST_ANTIPODAL].reserve(n));
    AG2_CUDA(c, c->list[AG2_LIST_SELECTED].reserve(n));
    const size_t smem = (size_t)kFusedMaxSelect * 12 + 33 * 4;
    AG2_CUDA(c, cudaFuncSetAttribute(k_select_small, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_select_small<<<1, kFT, smem, st>>>(c->list[AG2_LIST_PRUNED].p, (uint32_t)n, c->slots.p, c->logits.p,
                                         c->p.antipodal_mode, c->p.min_score_diff, c->p.num_selected,
                                         c->list[AG2_LIST_ANTIPODAL].p, c->list[AG2_LIST_SELECTED].p,
                                         c->scalars.p + 58);
    AG2_LAUNCH_CHECK(c);
    // the two counts follow the kernel to pinned memory; finish_pending publishes them after the next
    // synchronisation (the end of ag2_detect / ag2_classify / ag2_select)
    AG2_CUDA(c, cudaMemcpyAsync(c->h_pin + 27, c->scalars.p + 58, 8, cudaMemcpyDeviceToHost, st));
    c->pend_select = true;
    c->selected_precomputed = true;
    for (int i = 0; i < 4; i++) c->list_rec_valid[i] = false;  // scores / flags changed in the slots
    return AG2_OK;
  }
  AG2_CUDA(c, c->flags.reserve(std::max<size_t>(n, c->num_samples * (size_t)c->p.num_orientations) + 1));
  if (n) {
    k_flag_mode<<<(unsigned)((n + kT - 1) / kT), kT, 0, c->stream>>>(
        c->list[AG2_LIST_PRUNED].p, n, c->slots.p, c->logits.p, c->p.antipodal_mode, c->p.min_score_diff, c->flags.p);
    AG2_LAUNCH_CHECK(c);
  }
  AG2_TRY(compact_list(c, n, c->list[AG2_LIST_PRUNED].p, c->list[AG2_LIST_ANTIPODAL], &c->list_n[AG2_LIST_ANTIPODAL]));
  for (int i = 0; i < 4; i++) c->list_rec_valid[i] = false;  // scores / flags changed in the slots
  c->st.num_antipodal = (int64_t)c->list_n[AG2_LIST_ANTIPODAL];
  return AG2_OK;
}

int select_top(ag2_ctx* c) {
  phase_begin(c, 7);
  cudaStream_t st = c->stream;
  size_t n = c->list_n[AG2_LIST_ANTIPODAL];
  if (c->selected_precomputed) {  // the single-launch path of select_mode has built the list; counts: finish_pending
    c->list_rec_valid[AG2_LIST_SELECTED] = false;
    phase_end(c, 7);
    return AG2_OK;
  }
  // 4. clusters (grasp_detector.cpp:229-236): the top-k then ranks the moved / re-scored hands
  const bool clustered = c->p.min_inliers > 0 && c->p.antipodal_mode != AG2_MODE_NONE;
  const ag2_hypothesis* cl_src = nullptr;
  if (clustered) {
    AG2_TRY(gather_list_records(c, AG2_LIST_ANTIPODAL));
    AG2_TRY(clusters_compute(c, c->list_rec[AG2_LIST_ANTIPODAL].p, n));
    n = c->list_n[AG2_LIST_CLUSTERS];
    cl_src = c->list_rec[AG2_LIST_CLUSTERS].p;
    AG2_CUDA(c, c->list_rec[AG2_LIST_SELECTED].reserve(std::max<size_t>(n, 1)));
  }
  AG2_CUDA(c, c->list[AG2_LIST_SELECTED].reserve(std::max<size_t>(n, 1)));
  if (c->p.antipodal_mode == AG2_MODE_NONE || n == 0) {
    if (n) AG2_CUDA(c, cudaMemcpyAsync(c->list[AG2_LIST_SELECTED].p, c->list[AG2_LIST_ANTIPODAL].p, n * 4,
                                       cudaMemcpyDeviceToDevice, st));
    c->list_n[AG2_LIST_SELECTED] = n;
  } else {
    // rank by (score desc, list position asc) on the device
    const size_t k = (c->p.num_selected > 0 && (size_t)c->p.num_selected < n) ? (size_t)c->p.num_selected : n;
    uint32_t n_pad = 2;
    while (n_pad < n) n_pad <<= 1;
    AG2_CUDA(c, c->sort_keys[0].reserve(n_pad));
    AG2_CUDA(c, c->sort_keys[1].reserve(n_pad));
    AG2_CUDA(c, c->sort_pos[0].reserve(n_pad));
    AG2_CUDA(c, c->sort_pos[1].reserve(n_pad));
    k_score_keys<<<(unsigned)((n + kT - 1) / kT), kT, 0, st>>>(clustered ? nullptr : c->list[AG2_LIST_ANTIPODAL].p, n,
                                                             clustered ? cl_src : c->slots.p, c->sort_keys[0].p);
    AG2_LAUNCH_CHECK(c);
    const uint32_t* ranked = nullptr;
    if (n <= (size_t)kSortTile || k <= (size_t)kSortTile / 2) {
      // tournament of tile sorts: the global top-k is inside the union of the tiles' top-k
      int cur = 0;
      size_t m = n;
      const uint32_t* pos_in = nullptr;
      for (;;) {
        const unsigned tiles = (unsigned)((m + kSortTile - 1) / kSortTile);
        const uint32_t keep = tiles == 1 ? (uint32_t)std::min(m, k) : (uint32_t)k;
        k_sort_tile<<<tiles, 1024, 0, st>>>(c->sort_keys[cur].p, pos_in, (uint32_t)m, keep, c->sort_keys[cur ^ 1].p,
                                           c->sort_pos[cur ^ 1].p);
        AG2_LAUNCH_CHECK(c);
        cur ^= 1;
        pos_in = c->sort_pos[cur].p;
        if (tiles == 1) break;
        const size_t rem = m - (size_t)(tiles - 1) * kSortTile;
        m = (size_t)(tiles - 1) * keep + std::min<size_t>(keep, rem);
      }
      ranked = c->sort_pos[cur].p;
    } else {
      k_pad_keys<<<(n_pad + kT - 1) / kT, kT, 0, st>>>(c->sort_keys[0].p, c->sort_pos[0].p, (uint32_t)n, n_pad);
      AG2_LAUNCH_CHECK(c);
      for (uint32_t kk = 2; kk <= n_pad; kk <<= 1)
        for (uint32_t j = kk >> 1; j > 0; j >>= 1) {
          k_bitonic_step<<<(n_pad / 2 + kT - 1) / kT, kT, 0, st>>>(c->sort_keys[0].p, c->sort_pos[0].p, n_pad, kk, j);
          AG2_LAUNCH_CHECK(c);
        }
      ranked = c->sort_pos[0].p;
    }
    if (clustered) {
      k_take_ranked_rec<<<(unsigned)((k * 12 + kT - 1) / kT), kT, 0, st>>>(ranked, (uint32_t)k, cl_src,
                                                                         c->list_rec[AG2_LIST_SELECTED].p);
    } else {
      k_take_ranked<<<(unsigned)((k + kT - 1) / kT), kT, 0, st>>>(ranked, (uint32_t)k, c->list[AG2_LIST_ANTIPODAL].p,
                                                                c->list[AG2_LIST_SELECTED].p);
    }
    AG2_LAUNCH_CHECK(c);
    c->list_n[AG2_LIST_SELECTED] = k;
  }
  // with clustering the selected list exists as records only (the hands were moved)
  c->list_rec_valid[AG2_LIST_SELECTED] = clustered;
  c->st.num_selected = (int64_t)c->list_n[AG2_LIST_SELECTED];
  phase_end(c, 7);
  return AG2_OK;
}

int gather_list_records(ag2_ctx* c, int which) {
  if (which < 0 || which > 5) return AG2_E_INVALID;
  if (which >= 4) return AG2_OK;  // importance / cluster lists are stored as records
  if (c->list_rec_valid[which]) return AG2_OK;
  const size_t n = c->list_n[which];
  AG2_CUDA(c, c->list_rec[which].reserve(std::max<size_t>(n, 1)));
  if (n) {
    size_t threads = n * 12;
    k_gather<<<(unsigned)((threads + kT - 1) / kT), kT, 0, c->stream>>>(c->list[which].p, n, c->slots.p,
                                                                      c->list_rec[which].p);
    AG2_LAUNCH_CHECK(c);
  }
  c->list_rec_valid[which] = true;
  return AG2_OK;
}

}  // namespace ag2
