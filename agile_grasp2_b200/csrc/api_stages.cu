// C ABI (include/ag2.h), part 2: stages, outputs, introspection.
#include <math.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "dev_common.cuh"

using namespace ag2;

namespace {

// host restatement of the d2 used for ordering (float, unfused; built with -fmad=false / no contraction)
inline float h_dist2(const float* q, const float* p) {
  volatile float dx = q[0] - p[0], dy = q[1] - p[1], dz = q[2] - p[2];
  volatile float a = dx * dx, b = dy * dy, cc = dz * dz;
  volatile float s = a + b;
  return s + cc;
}

__global__ void k_unsort_normals(const float4* __restrict__ nrm, const uint32_t* __restrict__ sorted_of_orig,
                                 size_t n, float* __restrict__ out) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t s = sorted_of_orig[i];
  float nanv = __int_as_float(0x7fc00000);
  float4 v = s == 0xFFFFFFFFu ? make_float4(nanv, nanv, nanv, 0.f) : nrm[s];
  out[3 * i] = v.x;
  out[3 * i + 1] = v.y;
  out[3 * i + 2] = v.z;
}

int ready(ag2_ctx* c) {
  if (!c) return AG2_E_INVALID;
  if (cudaSetDevice(c->device) != cudaSuccess) return AG2_E_NO_DEVICE;
  if (c->n == 0) return AG2_E_EMPTY_CLOUD;
  return AG2_OK;
}

}  // namespace

namespace ag2 {
void merge_rank_lists(const ag2_hypothesis* const* recs, const size_t* counts, int world, int num_selected, bool keep_all,
                      std::vector<ag2_hypothesis>& all) {
  all.clear();
  for (int r = 0; r < world; r++)
    for (size_t i = 0; i < counts[r]; i++) {
      all.push_back(recs[r][i]);
      all.back().sample_id = r + world * recs[r][i].sample_id;
    }
  std::stable_sort(all.begin(), all.end(), [](const ag2_hypothesis& a, const ag2_hypothesis& b) {
    return a.sample_id != b.sample_id ? a.sample_id < b.sample_id : a.orientation_id < b.orientation_id;
  });
  if (keep_all) return;
  // the device ranking's key (select.cu score_key): order-preserving code of the double, descending
  auto code = [](double v) {
    long long o;
    memcpy(&o, &v, 8);
    return o >= 0 ? ((unsigned long long)o | 0x8000000000000000ull) : ~(unsigned long long)o;
  };
  std::stable_sort(all.begin(), all.end(), [&](const ag2_hypothesis& a, const ag2_hypothesis& b) {
    return code(a.score) > code(b.score);
  });
  if (num_selected > 0 && (size_t)num_selected < all.size()) all.resize((size_t)num_selected);
}
}  // namespace ag2

extern "C" {

int ag2_compute_normals(ag2_ctx* c) {
  AG2_TRY(ready(c));
  AG2_TRY(grid_build(c));
  AG2_TRY(normals_compute(c));
  return sync_and_finish(c);  // capacity errors of the phase surface here
}

int ag2_generate_hypotheses(ag2_ctx* c) {
  AG2_TRY(ready(c));
  if (c->num_samples == 0) return AG2_E_NO_SAMPLES;
  AG2_TRY(grid_build(c));
  AG2_TRY(normals_compute(c));
  AG2_TRY(frames_compute(c));
  AG2_TRY(hands_compute(c));
  AG2_TRY(lists_build_after_hands(c));
  c->have_hyps = true;
  c->have_pruned = c->have_images = c->have_antipodal = c->have_selected = false;
  return AG2_OK;
}

int ag2_prune(ag2_ctx* c) {
  AG2_TRY(ready(c));
  if (!c->have_hyps) return AG2_E_STATE;
  AG2_TRY(prune_compute(c));
  c->have_pruned = true;
  c->have_images = c->have_antipodal = c->have_selected = false;
  return AG2_OK;
}

int ag2_render_images(ag2_ctx* c, uint8_t* out) {
  AG2_TRY(ready(c));
  if (!c->have_pruned) return AG2_E_STATE;
  AG2_TRY(render_compute(c));
  if (out && c->list_n[AG2_LIST_PRUNED]) {
    AG2_CUDA(c, cudaMemcpyAsync(out, c->images.p, c->list_n[AG2_LIST_PRUNED] * (size_t)AG2_IMAGE_BYTES,
                                cudaMemcpyDeviceToHost, c->stream));
    AG2_CUDA(c, cudaStreamSynchronize(c->stream));
  }
  return AG2_OK;
}

// the two stages below without their closing synchronisation (ag2_detect has one for the whole call)
static int classify_stage(ag2_ctx* c) {
  AG2_TRY(ready(c));
  if (!c->have_images) return AG2_E_STATE;
  if (!c->net.loaded) return AG2_E_NO_WEIGHTS;
  AG2_TRY(classify_compute(c));
  AG2_TRY(select_mode(c));
  c->have_antipodal = true;
  return AG2_OK;
}

static int select_stage(ag2_ctx* c) {
  AG2_TRY(ready(c));
  if (!c->have_pruned) return AG2_E_STATE;
  if (c->p.antipodal_mode != AG2_MODE_PREDICTION) {
    AG2_TRY(select_mode(c));
    c->have_antipodal = true;
  }
  if (!c->have_antipodal) return AG2_E_STATE;
  AG2_TRY(select_top(c));
  c->have_selected = true;
  return AG2_OK;
}

int ag2_classify(ag2_ctx* c) {
  AG2_TRY(classify_stage(c));
  return sync_and_finish(c);
}

int ag2_select(ag2_ctx* c) {
  AG2_TRY(select_stage(c));
  return sync_and_finish(c);
}

int ag2_detect(ag2_ctx* c) {
  AG2_TRY(ready(c));
  if (c->num_samples == 0) return AG2_E_NO_SAMPLES;
  // Same stages as the calls above, but the host waits only where it must: once for the size of the cell index
  // (grid_build), once after the hand sweep + lists + prune step (list sizes decide the next launches), once at the
  // end.  Everything in between is queued without a round trip; statistics and capacity checks ride on those waits.
  AG2_TRY(grid_build(c));
  AG2_TRY(normals_compute(c));
  AG2_TRY(frames_compute(c));
  int rc = kHandsRetry;
  for (int attempt = 0; attempt < 8 && rc == kHandsRetry; attempt++) {
    AG2_TRY(hands_launch(c));
    AG2_TRY(lists_launch(c));
    AG2_TRY(prune_launch(c));
    AG2_TRY(sync_and_finish(c));
    rc = hands_finish(c);
  }
  if (rc == kHandsRetry) rc = AG2_E_CAPACITY;
  AG2_TRY(rc);
  AG2_TRY(lists_finish(c));
  AG2_TRY(prune_finish(c));
  c->have_hyps = c->have_pruned = true;
  c->have_images = c->have_antipodal = c->have_selected = false;
  if (c->p.antipodal_mode == AG2_MODE_PREDICTION) {
    AG2_TRY(ag2_render_images(c, nullptr));
    AG2_TRY(classify_stage(c));
  }
  AG2_TRY(select_stage(c));
  return sync_and_finish(c);
}

// ---- outputs

int ag2_num_hypotheses(ag2_ctx* c, int which, size_t* n) {
  if (!c || !n || which < 0 || which > 5) return AG2_E_INVALID;
  *n = c->list_n[which];
  return AG2_OK;
}

int ag2_get_hypotheses_device(ag2_ctx* c, int which, const ag2_hypothesis** d_ptr, size_t* n) {
  if (!c || which < 0 || which > 5) return AG2_E_INVALID;
  cudaSetDevice(c->device);
  AG2_TRY(gather_list_records(c, which));
  if (d_ptr) *d_ptr = c->list_rec[which].p;
  if (n) *n = c->list_n[which];
  return AG2_OK;
}

int ag2_normals_shard_points(ag2_ctx* c, int world, size_t* shard_points) {
  if (!c || world < 1 || !shard_points) return AG2_E_INVALID;
  cudaSetDevice(c->device);
  AG2_TRY(grid_build(c));
  *shard_points = normals_shard_points((uint32_t)c->n_finite, 0, world);
  return AG2_OK;
}

int ag2_compute_normals_shard(ag2_ctx* c, int rank, int world, float* d_out, size_t cap_points, size_t* n_owned) {
  if (!c || world < 1 || rank < 0 || rank >= world || (cap_points && !d_out)) return AG2_E_INVALID;
  cudaSetDevice(c->device);
  return normals_shard_compute(c, rank, world, d_out, cap_points, n_owned);
}

int ag2_set_normals_shards(ag2_ctx* c, const float* d_all, size_t cap_points, int world) {
  if (!c || world < 1 || (cap_points && !d_all)) return AG2_E_INVALID;
  cudaSetDevice(c->device);
  return normals_shard_assemble(c, d_all, cap_points, world);
}

int ag2_copy_hypotheses_device(ag2_ctx* c, int which, ag2_hypothesis* d_dst, size_t cap, size_t* n) {
  if (!c || which < 0 || which > 5 || (cap && !d_dst)) return AG2_E_INVALID;
  const size_t cnt = c->list_n[which];
  if (n) *n = cnt;
  if (cnt > cap) return AG2_E_CAPACITY;
  cudaSetDevice(c->device);
  if (cnt) {
    AG2_TRY(gather_list_records(c, which));
    AG2_CUDA(c, cudaMemcpyAsync(d_dst, c->list_rec[which].p, cnt * sizeof(ag2_hypothesis), cudaMemcpyDeviceToDevice,
                                c->stream));
  }
  if (cap > cnt) AG2_CUDA(c, cudaMemsetAsync(d_dst + cnt, 0, (cap - cnt) * sizeof(ag2_hypothesis), c->stream));
  AG2_CUDA(c, cudaStreamSynchronize(c->stream));
  return AG2_OK;
}

int ag2_get_hypotheses(ag2_ctx* c, int which, ag2_hypothesis* out, size_t cap, size_t* n) {
  if (!c || which < 0 || which > 5) return AG2_E_INVALID;
  const size_t cnt = c->list_n[which];
  if (n) *n = cnt;
  if (cnt > cap) return AG2_E_CAPACITY;
  if (cnt == 0) return AG2_OK;
  if (!out) return AG2_E_INVALID;
  cudaSetDevice(c->device);
  AG2_TRY(gather_list_records(c, which));
  AG2_CUDA(c, cudaMemcpyAsync(out, c->list_rec[which].p, cnt * sizeof(ag2_hypothesis), cudaMemcpyDeviceToHost,
                              c->stream));
  return sync_and_finish(c);
}

int ag2_get_stats(ag2_ctx* c, ag2_stats* out) {
  if (!c || !out) return AG2_E_INVALID;
  cudaSetDevice(c->device);
  AG2_TRY(collect_phase_times(c));
  c->st.gpu_launches = c->launches;
  c->st.num_points = (int64_t)c->n;
  c->st.num_samples = (int64_t)c->num_samples;
  *out = c->st;
  return AG2_OK;
}

// ---- introspection

int ag2_get_normals(ag2_ctx* c, float* out) {
  AG2_TRY(ready(c));
  if (!out) return AG2_E_INVALID;
  if (!c->have_normals) return AG2_E_STATE;
  DevBuf<float> tmp;
  AG2_CUDA(c, tmp.reserve(3 * c->n));
  k_unsort_normals<<<(unsigned)((c->n + 255) / 256), 256, 0, c->stream>>>(c->nrm.p, c->sorted_of_orig.p, c->n, tmp.p);
  AG2_LAUNCH_CHECK(c);
  AG2_CUDA(c, cudaMemcpyAsync(out, tmp.p, 12 * c->n, cudaMemcpyDeviceToHost, c->stream));
  AG2_CUDA(c, cudaStreamSynchronize(c->stream));
  return AG2_OK;
}

int ag2_get_frames(ag2_ctx* c, ag2_frame* out, size_t cap) {
  AG2_TRY(ready(c));
  if (!out) return AG2_E_INVALID;
  if (!c->have_hyps) return AG2_E_STATE;
  if (c->num_samples > cap) return AG2_E_CAPACITY;
  AG2_CUDA(c, cudaMemcpyAsync(out, c->frames.p, c->num_samples * sizeof(ag2_frame), cudaMemcpyDeviceToHost,
                              c->stream));
  AG2_CUDA(c, cudaStreamSynchronize(c->stream));
  return AG2_OK;
}

int ag2_get_sweep_records(ag2_ctx* c, ag2_sweep_record* out, size_t cap) {
  AG2_TRY(ready(c));
  if (!out) return AG2_E_INVALID;
  if (!c->have_hyps) return AG2_E_STATE;
  const size_t SR = c->num_samples * (size_t)c->p.num_orientations;
  if (SR > cap) return AG2_E_CAPACITY;
  AG2_CUDA(c, cudaMemcpyAsync(out, c->sweeps.p, SR * sizeof(ag2_sweep_record), cudaMemcpyDeviceToHost, c->stream));
  AG2_CUDA(c, cudaStreamSynchronize(c->stream));
  return AG2_OK;
}

int ag2_radius_search(ag2_ctx* c, const float* q, double radius, int32_t* idx, float* d2, size_t cap, size_t* n) {
  AG2_TRY(ready(c));
  if (!q || !n || !(radius > 0)) return AG2_E_INVALID;
  std::vector<int32_t> vi;
  std::vector<float> vd;
  AG2_TRY(radius_search_device(c, q, radius, vi, vd));
  *n = vi.size();
  if (vi.size() > cap) return AG2_E_CAPACITY;
  if (idx) std::copy(vi.begin(), vi.end(), idx);
  if (d2) std::copy(vd.begin(), vd.end(), d2);
  return AG2_OK;
}

int ag2_get_in_box(ag2_ctx* c, int which, size_t h, int32_t* idx, size_t cap, size_t* n) {
  AG2_TRY(ready(c));
  if (!n || which < 0 || which > 3 || h >= c->list_n[which]) return AG2_E_INVALID;
  cudaStream_t st = c->stream;
  uint32_t slot = 0;
  AG2_CUDA(c, cudaMemcpyAsync(&slot, c->list[which].p + h, 4, cudaMemcpyDeviceToHost, st));
  AG2_CUDA(c, cudaStreamSynchronize(st));
  BoxRef br;
  ag2_hypothesis hyp;
  AG2_CUDA(c, cudaMemcpyAsync(&br, c->box_ref.p + slot, sizeof(br), cudaMemcpyDeviceToHost, st));
  AG2_CUDA(c, cudaMemcpyAsync(&hyp, c->slots.p + slot, sizeof(hyp), cudaMemcpyDeviceToHost, st));
  AG2_CUDA(c, cudaStreamSynchronize(st));
  *n = br.count;
  if (br.count > cap) return AG2_E_CAPACITY;
  if (br.count == 0) return AG2_OK;
  std::vector<int32_t> pos(br.count);
  AG2_CUDA(c, cudaMemcpy(pos.data(), c->box_pool.p + br.offset, br.count * 4, cudaMemcpyDeviceToHost));
  ag2_frame fr;
  AG2_CUDA(c, cudaMemcpy(&fr, c->frames.p + hyp.sample_id, sizeof(fr), cudaMemcpyDeviceToHost));
  float q[3] = {(float)fr.sample[0], (float)fr.sample[1], (float)fr.sample[2]};
  std::vector<std::pair<uint64_t, int32_t>> keyed(br.count);
  for (uint32_t i = 0; i < br.count; i++) {
    float4 p;
    AG2_CUDA(c, cudaMemcpy(&p, c->pts.p + pos[i], sizeof(p), cudaMemcpyDeviceToHost));
    float pp[3] = {p.x, p.y, p.z};
    float dd = h_dist2(q, pp);
    uint32_t bits, orig;
    memcpy(&bits, &dd, 4);
    memcpy(&orig, &p.w, 4);
    keyed[i] = {((uint64_t)bits << 32) | orig, (int32_t)orig};
  }
  std::sort(keyed.begin(), keyed.end());
  for (uint32_t i = 0; i < br.count; i++) idx[i] = keyed[i].second;
  return AG2_OK;
}

int ag2_get_logits(ag2_ctx* c, float* out, size_t cap_images) {
  AG2_TRY(ready(c));
  if (!out) return AG2_E_INVALID;
  if (!c->have_antipodal || c->p.antipodal_mode != AG2_MODE_PREDICTION) return AG2_E_STATE;
  const size_t n = c->list_n[AG2_LIST_PRUNED];
  if (n > cap_images) return AG2_E_CAPACITY;
  if (n) {
    AG2_CUDA(c, cudaMemcpyAsync(out, c->logits.p, n * 8, cudaMemcpyDeviceToHost, c->stream));
    AG2_CUDA(c, cudaStreamSynchronize(c->stream));
  }
  return AG2_OK;
}

int ag2_classify_images(ag2_ctx* c, const uint8_t* images, size_t n, float* logits_out) {
  if (!c || (n && (!images || !logits_out))) return AG2_E_INVALID;
  if (cudaSetDevice(c->device) != cudaSuccess) return AG2_E_NO_DEVICE;
  if (!c->net.loaded) return AG2_E_NO_WEIGHTS;
  if (n == 0) return AG2_OK;
  DevBuf<uint8_t> d_img;
  DevBuf<float> d_log;
  AG2_CUDA(c, d_img.reserve(n * AG2_IMAGE_BYTES));
  AG2_CUDA(c, d_log.reserve(2 * n));
  AG2_CUDA(c, cudaMemcpyAsync(d_img.p, images, n * AG2_IMAGE_BYTES, cudaMemcpyHostToDevice, c->stream));
  AG2_TRY(lenet_forward(c, d_img.p, n, d_log.p));
  AG2_CUDA(c, cudaMemcpyAsync(logits_out, d_log.p, n * 8, cudaMemcpyDeviceToHost, c->stream));
  AG2_CUDA(c, cudaStreamSynchronize(c->stream));
  return AG2_OK;
}

int ag2_merge_selected(const ag2_hypothesis* blocks, int world, size_t cap, int num_selected, ag2_hypothesis* out,
                       size_t* n_out) {
  if (!blocks || !out || !n_out || world < 1) return AG2_E_INVALID;
  std::vector<std::vector<ag2_hypothesis>> parts((size_t)world);
  std::vector<const ag2_hypothesis*> ptrs((size_t)world);
  std::vector<size_t> counts((size_t)world);
  for (int r = 0; r < world; r++) {
    for (size_t i = 0; i < cap; i++) {
      const ag2_hypothesis& h = blocks[(size_t)r * cap + i];
      if (h.approach[0] == 0.0 && h.approach[1] == 0.0 && h.approach[2] == 0.0) continue;  // padding
      parts[r].push_back(h);
    }
    ptrs[r] = parts[r].data();
    counts[r] = parts[r].size();
  }
  std::vector<ag2_hypothesis> all;
  ag2::merge_rank_lists(ptrs.data(), counts.data(), world, num_selected, false, all);
  std::copy(all.begin(), all.end(), out);
  *n_out = all.size();
  return AG2_OK;
}

}  // extern "C"
