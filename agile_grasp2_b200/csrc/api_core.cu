// C ABI (include/ag2.h), part 1: lifecycle, parameters, weights, inputs.
#include <math.h>
#include <string.h>

#include <new>

#include "dev_common.cuh"

namespace ag2 {

int fail(ag2_ctx* c, int code, const char* what, cudaError_t e, const char* file, int line) {
  if (c) {
    char buf[512];
    snprintf(buf, sizeof(buf), "%s: %s (%s:%d)", what, cudaGetErrorString(e), file, line);
    c->err = buf;
  }
  return code;
}

void phase_begin(ag2_ctx* c, int phase) {
  cudaEventRecord(c->ev[2 * phase], c->stream);
  c->ev_used[2 * phase] = true;
}
void phase_end(ag2_ctx* c, int phase) {
  cudaEventRecord(c->ev[2 * phase + 1], c->stream);
  c->ev_used[2 * phase + 1] = true;
}
// After a synchronisation of the ctx stream: folds the scalars that earlier stages sent to pinned memory
// into the statistics and reports the capacity errors they carry.
int finish_pending(ag2_ctx* c) {
  if (c->pend_normals) {
    c->pend_normals = false;
    unsigned long long sum;
    memcpy(&sum, c->h_pin + 2, 8);
    c->st.sum_neighbors_normals = (int64_t)sum;
    c->max_kn_hint = c->h_pin[0];
    if (c->h_pin[4]) {
      c->have_normals = false;
      c->err = "normals: a point has more than 13600 neighbours within normals_radius";
      return AG2_E_CAPACITY;
    }
  }
  if (c->pend_select) {
    c->pend_select = false;
    c->list_n[AG2_LIST_ANTIPODAL] = c->h_pin[27];
    c->list_n[AG2_LIST_SELECTED] = c->selected_count = c->h_pin[28];
    c->st.num_antipodal = (int64_t)c->h_pin[27];
    c->st.num_selected = (int64_t)c->h_pin[28];
  }
  if (c->pend_frames) {
    c->pend_frames = false;
    if (c->h_pin[5] > 16384u) {
      c->have_hyps = false;
      c->err = "frames: a sample has more than 16384 neighbours within nn_radius_taubin";
      return AG2_E_CAPACITY;
    }
  }
  return AG2_OK;
}

int sync_and_finish(ag2_ctx* c) {
  AG2_CUDA(c, cudaStreamSynchronize(c->stream));
  return finish_pending(c);
}

int collect_phase_times(ag2_ctx* c) {
  AG2_TRY(sync_and_finish(c));
  for (int ph = 0; ph < 8; ph++) {
    if (c->ev_used[2 * ph] && c->ev_used[2 * ph + 1]) {
      float ms = 0.f;
      if (cudaEventElapsedTime(&ms, c->ev[2 * ph], c->ev[2 * ph + 1]) == cudaSuccess) c->st.ms_phase[ph] = ms;
    }
    c->st.ms_gap[ph] = 0.0;
    if (ph > 0 && c->ev_used[2 * ph - 1] && c->ev_used[2 * ph]) {
      float ms = 0.f;
      if (cudaEventElapsedTime(&ms, c->ev[2 * ph - 1], c->ev[2 * ph]) == cudaSuccess && ms > 0.f) c->st.ms_gap[ph] = ms;
    }
  }
  return AG2_OK;
}

static bool params_ok(const ag2_params* p) {
  return p && p->num_orientations >= 1 && p->num_orientations <= AG2_MAX_ORIENTATIONS &&
         p->nn_radius_hands > 0 && p->nn_radius_taubin > 0 && p->normals_radius > 0 && p->finger_width > 0 &&
         p->hand_outer_diameter > p->finger_width && p->hand_depth > 0 && p->hand_height > 0 &&
         isfinite(p->init_bite);
}

void invalidate_results(ag2_ctx* c) {
  c->have_hyps = c->have_pruned = c->have_images = c->have_antipodal = c->have_selected = false;
  for (int i = 0; i < 4; i++) {
    c->list_n[i] = 0;
    c->list_rec_valid[i] = false;
  }
}

static void mat4_mul(const double* a, const double* b, double* o) {
  for (int i = 0; i < 4; i++)
    for (int j = 0; j < 4; j++) {
      double s = 0;
      for (int k = 0; k < 4; k++) s += a[4 * i + k] * b[4 * k + j];
      o[4 * i + j] = s;
    }
}
// general 4x4 inverse (Gauss-Jordan, partial pivoting)
static void mat4_inv(const double* m, double* o) {
  double a[4][8];
  for (int i = 0; i < 4; i++)
    for (int j = 0; j < 4; j++) {
      a[i][j] = m[4 * i + j];
      a[i][4 + j] = i == j ? 1.0 : 0.0;
    }
  for (int col = 0; col < 4; col++) {
    int piv = col;
    for (int r = col + 1; r < 4; r++)
      if (fabs(a[r][col]) > fabs(a[piv][col])) piv = r;
    for (int j = 0; j < 8; j++) {
      double t = a[col][j];
      a[col][j] = a[piv][j];
      a[piv][j] = t;
    }
    double d = a[col][col];
    for (int j = 0; j < 8; j++) a[col][j] /= d;
    for (int r = 0; r < 4; r++) {
      if (r == col) continue;
      double f = a[r][col];
      for (int j = 0; j < 8; j++) a[r][j] -= f * a[col][j];
    }
  }
  for (int i = 0; i < 4; i++)
    for (int j = 0; j < 4; j++) o[4 * i + j] = a[i][4 + j];
}

namespace {
__global__ void k_count_bad_indices(const int32_t* __restrict__ idx, size_t n, size_t limit, uint32_t* __restrict__ bad) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && (idx[i] < 0 || (size_t)idx[i] >= limit)) atomicAdd(bad, 1u);
}
}  // namespace
}  // namespace ag2

using namespace ag2;

extern "C" {

int ag2_abi_version(void) { return AG2_ABI_VERSION; }

const char* ag2_strerror(int code) {
  switch (code) {
    case AG2_OK: return "ok";
    case AG2_E_INVALID: return "invalid argument";
    case AG2_E_NO_DEVICE: return "no usable CUDA device (libag2 has no CPU fallback)";
    case AG2_E_CUDA: return "CUDA error";
    case AG2_E_EMPTY_CLOUD: return "empty point cloud";
    case AG2_E_NO_SAMPLES: return "no samples set";
    case AG2_E_NO_WEIGHTS: return "classifier weights not loaded";
    case AG2_E_STATE: return "stage called out of order";
    case AG2_E_IO: return "i/o or parse error";
    case AG2_E_CAPACITY: return "capacity exceeded";
    case AG2_E_NOMEM: return "out of memory";
  }
  return "unknown error";
}

const char* ag2_last_error(const ag2_ctx* c) { return c ? c->err.c_str() : ""; }

int ag2_default_params(ag2_params* p) {
  if (!p) return AG2_E_INVALID;
  memset(p, 0, sizeof(*p));
  // grasp_detector.cpp:17-80
  p->nn_radius_taubin = 0.01;
  p->nn_radius_hands = 0.1;
  p->num_orientations = 8;
  p->num_threads = 1;
  p->num_samples = 1000;
  p->finger_width = 0.01;
  p->hand_outer_diameter = 0.09;
  p->hand_depth = 0.06;
  p->hand_height = 0.02;
  p->init_bite = 0.015;
  // grasp_detector.cpp:111-127: cam_tf_left = base_tf * sqrt_tf^-1, cam_tf_right = base_tf * sqrt_tf
  const double base_tf[16] = {0, 0.445417, 0.895323, 0.215, 1, 0, 0, -0.015, 0, 0.895323, -0.445417, 0.23, 0, 0, 0, 1};
  const double sqrt_tf[16] = {0.9366, -0.0162, 0.3500, -0.2863, 0.0151, 0.9999, 0.0058, 0.0058,
                              -0.3501, -0.0002, 0.9367, 0.0554, 0, 0, 0, 1};
  double inv[16];
  mat4_inv(sqrt_tf, inv);
  mat4_mul(base_tf, inv, p->cam_tf_left);
  mat4_mul(base_tf, sqrt_tf, p->cam_tf_right);
  // launch/file_detect_grasps.launch:16
  const double ws[6] = {0.48, 0.92, -0.6, 0.4, -0.25, 1.0};
  memcpy(p->workspace, ws, sizeof(ws));
  p->antipodal_mode = AG2_MODE_PREDICTION;
  p->filter_half_grasps = 1;
  p->prune = 1;
  p->num_selected = 50;
  p->min_score_diff = 500.0;
  p->min_aperture = 0.03;
  p->max_aperture = 0.07;
  p->normals_radius = 0.01;
  p->rng_seed = 0;
  return AG2_OK;
}

int ag2_create(const ag2_params* p, int device, ag2_ctx** out) {
  if (!out || !params_ok(p)) return AG2_E_INVALID;
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0 || device < 0 || device >= ndev) {
    cudaGetLastError();
    return AG2_E_NO_DEVICE;
  }
  if (cudaSetDevice(device) != cudaSuccess) return AG2_E_NO_DEVICE;
  ag2_ctx* c = new (std::nothrow) ag2_ctx();
  if (!c) return AG2_E_NOMEM;
  c->p = *p;
  c->device = device;
  memset(&c->st, 0, sizeof(c->st));
  for (int i = 0; i < 18; i++) c->ev_used[i] = false;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) c->sm_count = prop.multiProcessorCount;
  if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) {
    delete c;
    return AG2_E_CUDA;
  }
  for (int i = 0; i < 18; i++) cudaEventCreate(&c->ev[i]);
  cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming);
  for (int i = 0; i < ag2_ctx::kAux; i++) {
    if (cudaStreamCreateWithFlags(&c->aux[i], cudaStreamNonBlocking) != cudaSuccess) {
      ag2_destroy(c);
      return AG2_E_CUDA;
    }
    cudaEventCreateWithFlags(&c->ev_join[i], cudaEventDisableTiming);
  }
  if (cudaHostAlloc((void**)&c->h_pin, kPinWords * sizeof(uint32_t), cudaHostAllocDefault) != cudaSuccess) {
    c->h_pin = nullptr;
    ag2_destroy(c);
    return AG2_E_NOMEM;
  }
  memset(c->h_pin, 0, kPinWords * sizeof(uint32_t));
  if (c->scalars.reserve(64) != cudaSuccess || build_hand_const(c) != AG2_OK) {
    ag2_destroy(c);
    return AG2_E_INVALID;
  }
  *out = c;
  return AG2_OK;
}

int ag2_destroy(ag2_ctx* c) {
  if (!c) return AG2_OK;
  cudaSetDevice(c->device);
  if (c->stream) {
    cudaStreamSynchronize(c->stream);
    cudaStreamDestroy(c->stream);
  }
  for (int i = 0; i < 18; i++) cudaEventDestroy(c->ev[i]);
  for (int i = 0; i < ag2_ctx::kAux; i++) {
    if (c->aux[i]) {
      cudaStreamSynchronize(c->aux[i]);
      cudaStreamDestroy(c->aux[i]);
    }
    if (c->ev_join[i]) cudaEventDestroy(c->ev_join[i]);
  }
  if (c->ev_fork) cudaEventDestroy(c->ev_fork);
  if (c->h_pin) cudaFreeHost(c->h_pin);
  delete c;
  return AG2_OK;
}

int ag2_set_params(ag2_ctx* c, const ag2_params* p) {
  if (!c || !params_ok(p)) return AG2_E_INVALID;
  bool regrid = p->normals_radius != c->p.normals_radius || p->nn_radius_hands != c->p.nn_radius_hands;
  c->p = *p;
  AG2_TRY(build_hand_const(c));
  if (regrid) {
    c->grid_built = false;
    c->have_normals = false;
  }
  invalidate_results(c);
  return AG2_OK;
}

int ag2_synchronize(ag2_ctx* c) {
  if (!c) return AG2_E_INVALID;
  return sync_and_finish(c);
}

int ag2_get_stream(ag2_ctx* c, void** stream) {
  if (!c || !stream) return AG2_E_INVALID;
  *stream = (void*)c->stream;
  return AG2_OK;
}

// ---- weights

int ag2_set_lenet_weights(ag2_ctx* c, const float* c1w, const float* c1b, const float* c2w, const float* c2b,
                          const float* i1w, const float* i1b, const float* i2w, const float* i2b) {
  if (!c || !c1w || !c1b || !c2w || !c2b || !i1w || !i1b || !i2w || !i2b) return AG2_E_INVALID;
  cudaSetDevice(c->device);
  const float* src[8] = {c1w, c1b, c2w, c2b, i1w, i1b, i2w, i2b};
  static const size_t sizes[8] = {20 * 75, 20, 50 * 500, 50, 500 * 7200, 500, 2 * 500, 2};
  for (int i = 0; i < 8; i++) c->net.host[i].assign(src[i], src[i] + sizes[i]);
  return lenet_upload(c);
}

int ag2_init_lenet_xavier(ag2_ctx* c, uint64_t seed) {
  if (!c) return AG2_E_INVALID;
  cudaSetDevice(c->device);
  xavier_fill(seed, c->net.host);
  return lenet_upload(c);
}

int ag2_load_lenet_caffemodel(ag2_ctx* c, const char* path) {
  if (!c || !path) return AG2_E_INVALID;
  cudaSetDevice(c->device);
  std::vector<float> w[8];
  int rc = load_caffemodel(path, w, c->err);
  if (rc) return rc;
  for (int i = 0; i < 8; i++) c->net.host[i].swap(w[i]);
  return lenet_upload(c);
}

int ag2_get_lenet_weights(ag2_ctx* c, float* c1w, float* c1b, float* c2w, float* c2b, float* i1w, float* i1b,
                          float* i2w, float* i2b) {
  if (!c) return AG2_E_INVALID;
  if (!c->net.loaded) return AG2_E_NO_WEIGHTS;
  float* dst[8] = {c1w, c1b, c2w, c2b, i1w, i1b, i2w, i2b};
  for (int i = 0; i < 8; i++)
    if (dst[i]) memcpy(dst[i], c->net.host[i].data(), c->net.host[i].size() * sizeof(float));
  return AG2_OK;
}

// ---- inputs

}  // extern "C"
namespace ag2 {
int cloud_reset(ag2_ctx* c, size_t n, int num_cams) {
  // sample INDICES were validated against the previous cloud: a new cloud drops them (sample points stay)
  if (!c->use_points) c->num_samples = 0;
  c->n = n;
  c->num_cams = num_cams;
  c->grid_built = false;
  c->have_normals = false;
  invalidate_results(c);
  int64_t launches_keep = 0;
  memset(&c->st, 0, sizeof(c->st));
  c->launches = launches_keep;
  c->st.num_points = (int64_t)n;
  c->st.num_samples = (int64_t)c->num_samples;
  for (int i = 0; i < 18; i++) c->ev_used[i] = false;
  return AG2_OK;
}
}  // namespace ag2
extern "C" {

int ag2_set_cloud(ag2_ctx* c, const float* xyz, size_t stride_bytes, size_t n, const uint8_t* cam_mask, int num_cams,
                  const float* normals) {
  if (!c || (n && !xyz) || stride_bytes < 12 || stride_bytes % 4 || num_cams < 1 || num_cams > 8)
    return AG2_E_INVALID;
  if (n >= 0x7fffffffu) return AG2_E_CAPACITY;
  cudaSetDevice(c->device);
  cloud_reset(c, n, num_cams);
  if (n == 0) return AG2_OK;
  AG2_CUDA(c, c->xyz_in.reserve(3 * n));
  AG2_CUDA(c, c->cam_in.reserve(n));
  if (stride_bytes == 12) {
    AG2_CUDA(c, cudaMemcpyAsync(c->xyz_in.p, xyz, 12 * n, cudaMemcpyHostToDevice, c->stream));
  } else {
    AG2_CUDA(c, cudaMemcpy2DAsync(c->xyz_in.p, 12, xyz, stride_bytes, 12, n, cudaMemcpyHostToDevice, c->stream));
  }
  if (cam_mask) {
    AG2_CUDA(c, cudaMemcpyAsync(c->cam_in.p, cam_mask, n, cudaMemcpyHostToDevice, c->stream));
  } else {
    AG2_CUDA(c, cudaMemsetAsync(c->cam_in.p, 1, n, c->stream));
  }
  c->normals_given = normals != nullptr;
  if (normals) {
    AG2_CUDA(c, c->normals_in.reserve(3 * n));
    AG2_CUDA(c, cudaMemcpyAsync(c->normals_in.p, normals, 12 * n, cudaMemcpyHostToDevice, c->stream));
  }
  return AG2_OK;
}

int ag2_set_cloud_device(ag2_ctx* c, const float* d_xyz, size_t n, const uint8_t* d_cam, int num_cams,
                         const float* d_normals) {
  if (!c || (n && !d_xyz) || num_cams < 1 || num_cams > 8) return AG2_E_INVALID;
  if (n >= 0x7fffffffu) return AG2_E_CAPACITY;
  cudaSetDevice(c->device);
  cloud_reset(c, n, num_cams);
  if (n == 0) return AG2_OK;
  AG2_CUDA(c, c->xyz_in.reserve(3 * n));
  AG2_CUDA(c, c->cam_in.reserve(n));
  AG2_CUDA(c, cudaMemcpyAsync(c->xyz_in.p, d_xyz, 12 * n, cudaMemcpyDeviceToDevice, c->stream));
  if (d_cam) {
    AG2_CUDA(c, cudaMemcpyAsync(c->cam_in.p, d_cam, n, cudaMemcpyDeviceToDevice, c->stream));
  } else {
    AG2_CUDA(c, cudaMemsetAsync(c->cam_in.p, 1, n, c->stream));
  }
  c->normals_given = d_normals != nullptr;
  if (d_normals) {
    AG2_CUDA(c, c->normals_in.reserve(3 * n));
    AG2_CUDA(c, cudaMemcpyAsync(c->normals_in.p, d_normals, 12 * n, cudaMemcpyDeviceToDevice, c->stream));
  }
  return AG2_OK;
}

int ag2_set_sample_indices(ag2_ctx* c, const int32_t* idx, size_t n) {
  if (!c || (n && !idx)) return AG2_E_INVALID;
  for (size_t i = 0; i < n; i++)
    if (idx[i] < 0 || (size_t)idx[i] >= c->n) return AG2_E_INVALID;
  cudaSetDevice(c->device);
  c->h_sample_idx.assign(idx, idx + n);
  c->h_sample_idx_stale = false;
  c->use_points = false;
  c->num_samples = n;
  c->st.num_samples = (int64_t)n;
  invalidate_results(c);
  if (n) {
    AG2_CUDA(c, c->sample_idx.reserve(n));
    AG2_CUDA(c, cudaMemcpyAsync(c->sample_idx.p, c->h_sample_idx.data(), 4 * n, cudaMemcpyHostToDevice, c->stream));
  }
  return AG2_OK;
}

int ag2_set_sample_indices_device(ag2_ctx* c, const int32_t* d_idx, size_t n) {
  if (!c || (n && !d_idx)) return AG2_E_INVALID;
  cudaSetDevice(c->device);
  if (n) {
    AG2_CUDA(c, c->sample_idx.reserve(n));
    AG2_CUDA(c, cudaMemsetAsync(c->scalars.p + 14, 0, 4, c->stream));
    k_count_bad_indices<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(d_idx, n, c->n, c->scalars.p + 14);
    AG2_LAUNCH_CHECK(c);
    AG2_CUDA(c, cudaMemcpyAsync(c->sample_idx.p, d_idx, 4 * n, cudaMemcpyDeviceToDevice, c->stream));
    uint32_t bad = 0;
    AG2_CUDA(c, cudaMemcpyAsync(&bad, c->scalars.p + 14, 4, cudaMemcpyDeviceToHost, c->stream));
    AG2_CUDA(c, cudaStreamSynchronize(c->stream));
    if (bad) return AG2_E_INVALID;
  }
  c->h_sample_idx.clear();
  c->h_sample_idx_stale = n > 0;  // fetched on demand by ag2_get_sample_indices
  c->use_points = false;
  c->num_samples = n;
  c->st.num_samples = (int64_t)n;
  invalidate_results(c);
  return AG2_OK;
}

int ag2_set_sample_points(ag2_ctx* c, const double* xyz, size_t n) {
  if (!c || (n && !xyz)) return AG2_E_INVALID;
  cudaSetDevice(c->device);
  c->h_sample_xyz.assign(xyz, xyz + 3 * n);
  c->use_points = true;
  c->num_samples = n;
  c->st.num_samples = (int64_t)n;
  invalidate_results(c);
  if (n) {
    AG2_CUDA(c, c->sample_xyz.reserve(3 * n));
    AG2_CUDA(c, cudaMemcpyAsync(c->sample_xyz.p, c->h_sample_xyz.data(), 24 * n, cudaMemcpyHostToDevice, c->stream));
  }
  return AG2_OK;
}

int ag2_set_sample_id_offset(ag2_ctx* c, int64_t first) {
  if (!c) return AG2_E_INVALID;
  c->sample_id_offset = first;
  c->sample_id_stride = 1;
  return AG2_OK;
}

int ag2_set_sample_id_map(ag2_ctx* c, int64_t first, int64_t stride) {
  if (!c || stride < 1) return AG2_E_INVALID;
  c->sample_id_offset = first;
  c->sample_id_stride = stride;
  return AG2_OK;
}

}  // extern "C"
