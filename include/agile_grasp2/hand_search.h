// Shim of the reference's HandSearch (include/agile_grasp2/hand_search.h:72-133,
// src/agile_grasp2/hand_search.cpp:4-80): same Parameters struct and generateHypotheses signature.
#ifndef AGILE_GRASP2_SHIM_HAND_SEARCH_H_
#define AGILE_GRASP2_SHIM_HAND_SEARCH_H_

#include <ag2.h>

#include <stdexcept>
#include <string>
#include <vector>

#include "cloud_camera.h"
#include "grasp_hypothesis.h"

namespace ag2shim {

inline void check(int rc, const char* what, const ag2_ctx* c = nullptr) {
  if (rc != AG2_OK)
    throw std::runtime_error(std::string(what) + ": " + ag2_strerror(rc) +
                             (c ? std::string(" ") + ag2_last_error(c) : std::string()));
}

// Uploads the processed cloud, camera sources, normals and samples of a CloudCamera.
// use_samples: xyz samples (getSamples) instead of indices, as in generateHypotheses(.., use_samples).
inline void upload(ag2_ctx* ctx, const CloudCamera& cc, bool use_samples) {
  const CloudCamera::PointCloudRGB& cloud = cc.getCloudProcessed();
  std::vector<uint8_t> mask = cc.cameraMask();
  std::vector<float> nrm;
  if (cc.getNormals().cols() == (int)cloud.size() && cloud.size() > 0) {
    nrm = ag2shim::to_floats(cc.getNormals());
  }
  check(ag2_set_cloud(ctx, cloud.xyz.data(), 12, cloud.size(), mask.data(), cc.getCameraSource().rows(),
                      nrm.empty() ? nullptr : nrm.data()),
        "ag2_set_cloud", ctx);
  if (use_samples) {
    check(ag2_set_sample_points(ctx, cc.getSamples().data(), (size_t)cc.getSamples().cols()),
          "ag2_set_sample_points", ctx);
  } else {
    std::vector<int32_t> idx(cc.getSampleIndices().begin(), cc.getSampleIndices().end());
    check(ag2_set_sample_indices(ctx, idx.data(), idx.size()), "ag2_set_sample_indices", ctx);
  }
}

inline std::vector<GraspHypothesis> fetch(ag2_ctx* ctx, int which) {
  size_t n = 0;
  check(ag2_num_hypotheses(ctx, which, &n), "ag2_num_hypotheses", ctx);
  std::vector<ag2_hypothesis> rec(n ? n : 1);
  check(ag2_get_hypotheses(ctx, which, rec.data(), rec.size(), &n), "ag2_get_hypotheses", ctx);
  std::vector<GraspHypothesis> out;
  out.reserve(n);
  for (size_t i = 0; i < n; i++) out.push_back(GraspHypothesis(rec[i], ctx));
  return out;
}

}  // namespace ag2shim

class HandSearch {
 public:
  struct Parameters {  // hand_search.h:72-91
    double nn_radius_taubin_, nn_radius_hands_;
    int num_orientations_, num_threads_, num_samples_;
    ag2shim::Matrix4d cam_tf_left_ = ag2shim::identity4(), cam_tf_right_ = ag2shim::identity4();
    double finger_width_, hand_outer_diameter_, hand_depth_, hand_height_, init_bite_;
  };

  HandSearch() : ctx_(nullptr) {
    ag2_default_params(&p_);
    p_.antipodal_mode = AG2_MODE_NONE;
    p_.prune = 0;
  }
  explicit HandSearch(const Parameters& params) : HandSearch() { setParameters(params); }
  // hand_search.h:114-118: hand geometry only; nn_radius_taubin_ = 0.03 and nn_radius_hands_ = 0.08 as there (the
  // reference leaves num_orientations_ uninitialised in this constructor; the default of 8 is kept)
  HandSearch(double finger_width, double hand_outer_diameter, double hand_depth, double hand_height, double init_bite,
             int num_threads, int num_samples)
      : HandSearch() {
    p_.finger_width = finger_width;
    p_.hand_outer_diameter = hand_outer_diameter;
    p_.hand_depth = hand_depth;
    p_.hand_height = hand_height;
    p_.init_bite = init_bite;
    p_.num_threads = num_threads;
    p_.num_samples = num_samples;
    p_.nn_radius_taubin = 0.03;
    p_.nn_radius_hands = 0.08;
  }
  ~HandSearch() { ag2_destroy(ctx_); }
  HandSearch(const HandSearch&) = delete;
  HandSearch& operator=(const HandSearch&) = delete;

  void setParameters(const Parameters& params) {  // hand_search.cpp:64-80
    p_.nn_radius_taubin = params.nn_radius_taubin_;
    p_.nn_radius_hands = params.nn_radius_hands_;
    p_.num_orientations = params.num_orientations_;
    p_.num_threads = params.num_threads_;
    p_.num_samples = params.num_samples_;
    p_.finger_width = params.finger_width_;
    p_.hand_outer_diameter = params.hand_outer_diameter_;
    p_.hand_depth = params.hand_depth_;
    p_.hand_height = params.hand_height_;
    p_.init_bite = params.init_bite_;
    ag2shim::row_major16(params.cam_tf_left_, p_.cam_tf_left);
    ag2shim::row_major16(params.cam_tf_right_, p_.cam_tf_right);
    apply();
  }
  // seed of the counter-based replacement of rand() (hand_search.cpp:130); the reference calls
  // std::srand(std::time(0)) in detect_grasps.cpp:9 instead
  void setSeed(uint64_t seed) {
    p_.rng_seed = seed;
    apply();
  }
  void setCamTfLeft(const ag2shim::Matrix4d& t) {
    ag2shim::row_major16(t, p_.cam_tf_left);
    apply();
  }
  void setCamTfRight(const ag2shim::Matrix4d& t) {
    ag2shim::row_major16(t, p_.cam_tf_right);
    apply();
  }

  // hand_search.cpp:4-61.  antipodal_mode / forces_PSD / plot flags are ignored by the reference too.
  std::vector<GraspHypothesis> generateHypotheses(const CloudCamera& cloud_cam, int /*antipodal_mode*/,
                                                  bool use_samples, bool /*forces_PSD*/ = false,
                                                  bool /*plots_normals*/ = false, bool /*plots_samples*/ = false) {
    if (cloud_cam.getCloudProcessed().size() == 0) return std::vector<GraspHypothesis>();
    ensure();
    ag2shim::upload(ctx_, cloud_cam, use_samples);
    int rc = ag2_generate_hypotheses(ctx_);
    if (rc == AG2_E_NO_SAMPLES) return std::vector<GraspHypothesis>();
    ag2shim::check(rc, "ag2_generate_hypotheses", ctx_);
    return ag2shim::fetch(ctx_, AG2_LIST_HYPOTHESES);
  }

  ag2_ctx* context() {
    ensure();
    return ctx_;
  }

 private:
  void ensure() {
    if (!ctx_) ag2shim::check(ag2_create(&p_, 0, &ctx_), "ag2_create");
  }
  void apply() {
    if (ctx_) ag2shim::check(ag2_set_params(ctx_, &p_), "ag2_set_params", ctx_);
  }
  ag2_params p_;
  ag2_ctx* ctx_;
};

#endif
