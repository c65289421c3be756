// Shims of the reference's GraspDetector (include/agile_grasp2/grasp_detector.h:72-126,
// src/agile_grasp2/grasp_detector.cpp:15-395), ImportanceSampling
// (include/agile_grasp2/importance_sampling.h:25-27, importance_sampling.cpp:30-112) and
// HandleSearch::findClusters (handle_search.cpp:4-80).  The ROS NodeHandle of the reference's
// constructors is replaced by a plain Params struct whose fields carry the ROS parameter names
// and the compiled-in defaults of grasp_detector.cpp:17-80.
#ifndef AGILE_GRASP2_SHIM_GRASP_DETECTOR_H_
#define AGILE_GRASP2_SHIM_GRASP_DETECTOR_H_

#include <ag2.h>

#include <algorithm>
#include <stdexcept>
#include <string>
#include <vector>

#include "hand_search.h"

class GraspDetector {
 public:
  struct Params {  // names = the ROS params read at grasp_detector.cpp:17-80
    std::vector<double> workspace;
    std::vector<double> camera_pose;  // 16 values row-major, empty: Baxter two-camera default (:108-137)
    int num_samples = 1000, num_threads = 1, num_orientations = 8;
    double nn_radius_taubin = 0.01, nn_radius_hands = 0.1;
    bool voxelize = true, filter_half_grasps = true;
    double finger_width = 0.01, hand_outer_diameter = 0.09, hand_depth = 0.06, hand_height = 0.02, init_bite = 0.015;
    int antipodal_mode = AG2_MODE_PREDICTION;
    std::string model_file, trained_file, label_file;  // trained_file empty: Xavier fillers (weights synthetic)
    double min_score_diff = 500.0;
    int min_inliers = 0, num_selected = 50;
    std::vector<double> gripper_width_range;  // [0.03, 0.07]
    uint64_t rng_seed = 0;
    bool prune = true;  // false when the sample_indices param is set (:150-160)
    // new: GPUs of this box the detect calls are sharded over (ag2_group_create; one host thread per GPU inside
    // the library, exchanges on NVLink peer memory).  1 = the single-GPU ctx.  Results do not depend on it.
    int num_gpus = 1;
  };

  explicit GraspDetector(const Params& prm) : prm_(prm), ctx_(nullptr), group_(nullptr), use_incoming_samples_(false) {
    ag2_default_params(&p_);
    if (prm.workspace.size() == 6)
      for (int i = 0; i < 6; i++) p_.workspace[i] = prm.workspace[i];
    if (prm.camera_pose.size() == 16)
      for (int i = 0; i < 16; i++) p_.cam_tf_left[i] = p_.cam_tf_right[i] = prm.camera_pose[i];
    p_.num_samples = prm.num_samples;
    p_.num_threads = prm.num_threads;
    p_.num_orientations = prm.num_orientations;
    p_.nn_radius_taubin = prm.nn_radius_taubin;
    p_.nn_radius_hands = prm.nn_radius_hands;
    p_.finger_width = prm.finger_width;
    p_.hand_outer_diameter = prm.hand_outer_diameter;
    p_.hand_depth = prm.hand_depth;
    p_.hand_height = prm.hand_height;
    p_.init_bite = prm.init_bite;
    p_.antipodal_mode = prm.antipodal_mode;
    p_.filter_half_grasps = prm.filter_half_grasps ? 1 : 0;
    p_.min_score_diff = prm.min_score_diff;
    p_.min_inliers = prm.min_inliers;
    p_.num_selected = prm.num_selected;
    p_.prune = prm.prune ? 1 : 0;
    p_.rng_seed = prm.rng_seed;
    if (prm.gripper_width_range.size() == 2) {
      p_.min_aperture = prm.gripper_width_range[0];
      p_.max_aperture = prm.gripper_width_range[1];
    }
    ag2shim::check(ag2_create(&p_, 0, &ctx_), "ag2_create");
    loadClassifier(ctx_);
    if (prm.num_gpus > 1) {
      int rc = ag2_group_create(&p_, prm.num_gpus, &group_);
      if (rc != AG2_OK) {
        ag2_destroy(ctx_);
        ag2shim::check(rc, "ag2_group_create");
      }
      for (int i = 0; i < prm.num_gpus; i++) {
        ag2_ctx* rc_ctx = nullptr;
        ag2_group_local_ctx(group_, i, &rc_ctx);
        loadClassifier(rc_ctx);
      }
    }
  }
  virtual ~GraspDetector() {
    ag2_group_destroy(group_);
    ag2_destroy(ctx_);
  }
  GraspDetector(const GraspDetector&) = delete;
  GraspDetector& operator=(const GraspDetector&) = delete;

  // Stand-ins for the two ROS messages the reference keeps sample requests in (msg/SamplesMsg.msg: geometry_msgs/Point[]
  // samples; msg/CloudIndexed.msg: std_msgs/Int64[] indices).  The setters below are templates, so the real message
  // types work as well.
  struct Point {
    double x, y, z;
  };
  struct SamplesMsg {
    std::vector<Point> samples;
  };

  // grasp_detector.h:111 / grasp_detector.cpp:353-360: cloud indices to sample at (CloudIndexed); they switch the
  // workspace filter, the voxelisation and the pruning off (:287, :150-160)
  template <class Msg>
  void setIndicesFromMsg(const Msg& msg) {
    indices_.resize(msg.indices.size());
    for (size_t i = 0; i < indices_.size(); i++) indices_[i] = (int)msg.indices[i].data;
  }
  void setIndices(const std::vector<int>& indices) { indices_ = indices; }
  // grasp_detector.h:113-116: sample points from an external source, used when use_incoming_samples is set
  template <class Msg>
  void setSamplesMsg(const Msg& msg) {
    samples_msg_.samples.resize(msg.samples.size());
    for (size_t i = 0; i < msg.samples.size(); i++)
      samples_msg_.samples[i] = Point{(double)msg.samples[i].x, (double)msg.samples[i].y, (double)msg.samples[i].z};
    have_samples_msg_ = true;
  }

  // grasp_detector.cpp:285-350: workspace filter, voxelisation (voxel_size_ = 0.003), sample selection
  void preprocessPointCloud(CloudCamera& cloud_cam) {
    if (!indices_.empty()) {  // :334-347: the cloud is used as it came
      cloud_cam.setSampleIndices(
          ag2shim::drawFromIndices(indices_, prm_.num_samples, cloud_cam.getCloudOriginal().size(), prm_.rng_seed));
      return;
    }
    cloud_cam.filterWorkspaceAndVoxelize(ctx_, prm_.voxelize, 0.003);
    const int n = (int)cloud_cam.getCloudProcessed().size();
    if (use_incoming_samples_) {
      if (!have_samples_msg_) return;  // xyz samples were set on the CloudCamera by the caller (setSamples)
      std::vector<ag2shim::SamplePoint> all;
      for (const Point& q : samples_msg_.samples) all.push_back(ag2shim::SamplePoint{{q.x, q.y, q.z}});
      SamplesMsg inside;  // :305-313: samples outside the workspace are removed
      for (const ag2shim::SamplePoint& q : ag2shim::samplesInWorkspace(all, p_.workspace))
        inside.samples.push_back(Point{q[0], q[1], q[2]});
      cloud_cam.subsampleSamples(inside, prm_.num_samples, prm_.rng_seed);  // :317
      return;
    }
    if (prm_.num_samples > n) {
      std::vector<int> all((size_t)n);
      for (int i = 0; i < n; i++) all[(size_t)i] = i;
      cloud_cam.setSampleIndices(all);
    } else {
      cloud_cam.subsampleUniformly(ctx_, prm_.num_samples, prm_.rng_seed);
    }
  }

  // grasp_detector.cpp:84-282
  virtual std::vector<GraspHypothesis> detectGraspPoses(const CloudCamera& cloud_cam, bool clusters_grasps = true) {
    if (cloud_cam.getCloudOriginal().size() == 0 || cloud_cam.getCloudProcessed().size() == 0)
      return std::vector<GraspHypothesis>();  // :86-91
    ag2_params p = p_;
    if (!clusters_grasps) p.min_inliers = 0;
    if (!indices_.empty()) p.prune = 0;  // :150-160: no pruning on caller-supplied indices
    if (group_ && !use_incoming_samples_ && indices_.empty() && p.min_inliers == 0 && cloud_cam.getNormals().cols() == 0)
      return detectOnGroup(cloud_cam, nullptr);
    ag2shim::check(ag2_set_params(ctx_, &p), "ag2_set_params", ctx_);
    ag2shim::upload(ctx_, cloud_cam, use_incoming_samples_);
    int rc = ag2_detect(ctx_);
    if (rc == AG2_E_NO_SAMPLES || rc == AG2_E_EMPTY_CLOUD) return std::vector<GraspHypothesis>();
    ag2shim::check(rc, "ag2_detect", ctx_);
    return ag2shim::fetch(ctx_, AG2_LIST_SELECTED);
  }

  void setUseIncomingSamples(bool b) { use_incoming_samples_ = b; }
  // grasp_detector.h:86-125
  static bool isScoreGreater(const GraspHypothesis& a, const GraspHypothesis& b) { return a.getScore() > b.getScore(); }
  bool getUseIncomingSamples() const { return use_incoming_samples_; }
  std::vector<double> getWorkspace() const { return std::vector<double>(p_.workspace, p_.workspace + 6); }
  int getNumSamples() const { return prm_.num_samples; }
  void setNumSamples(int num_samples) { prm_.num_samples = p_.num_samples = num_samples; }
  ag2_ctx* context() { return ctx_; }
  ag2_stats stats() {
    ag2_stats s;
    ag2shim::check(ag2_get_stats(ctx_, &s), "ag2_get_stats", ctx_);
    return s;
  }

 protected:
  void loadClassifier(ag2_ctx* c) {
    if (p_.antipodal_mode != AG2_MODE_PREDICTION) return;  // new Classifier(...), grasp_detector.cpp:55
    int rc = prm_.trained_file.empty() ? ag2_init_lenet_xavier(c, 1) : ag2_load_lenet_caffemodel(c, prm_.trained_file.c_str());
    if (rc != AG2_OK) {
      std::string msg = "GraspDetector: cannot load " + prm_.trained_file + ": " + ag2_last_error(c);
      ag2_group_destroy(group_);
      ag2_destroy(ctx_);
      throw std::runtime_error(msg);
    }
  }
  // the same call over all GPUs of the group (index samples, normals computed by the path); ip != 0: the
  // importance-sampling rounds
  std::vector<GraspHypothesis> detectOnGroup(const CloudCamera& cc, const ag2_is_params* ip) {
    const CloudCamera::PointCloudRGB& cloud = cc.getCloudProcessed();
    const std::vector<uint8_t> mask = cc.cameraMask();
    const std::vector<int32_t> idx(cc.getSampleIndices().begin(), cc.getSampleIndices().end());
    if (idx.empty()) return std::vector<GraspHypothesis>();
    const size_t per_call = p_.num_selected > 0 && p_.antipodal_mode != AG2_MODE_NONE ? (size_t)p_.num_selected
                                                                                       : idx.size() * (size_t)p_.num_orientations;
    size_t cap = per_call, n = 0;
    if (ip) cap = per_call + (size_t)ip->num_iterations * std::max<size_t>(per_call, (size_t)ip->num_samples * p_.num_orientations);
    std::vector<ag2_hypothesis> rec(cap ? cap : 1);
    int rc = ip ? ag2_group_detect_importance(group_, cloud.xyz.data(), cloud.size(), mask.data(), cc.getCameraSource().rows(),
                                              idx.data(), idx.size(), 0, ip, rec.data(), rec.size(), &n)
                : ag2_group_detect(group_, cloud.xyz.data(), cloud.size(), mask.data(), cc.getCameraSource().rows(),
                                   idx.data(), idx.size(), 0, rec.data(), rec.size(), &n);
    if (rc != AG2_OK) throw std::runtime_error(std::string("ag2_group_detect: ") + ag2_strerror(rc) + ": " + ag2_group_last_error(group_));
    std::vector<GraspHypothesis> out;
    out.reserve(n);
    for (size_t i = 0; i < n; i++) out.push_back(GraspHypothesis(rec[i], nullptr));
    return out;
  }

  Params prm_;
  ag2_params p_;
  ag2_ctx* ctx_;
  ag2_group* group_;
  bool use_incoming_samples_;
  bool have_samples_msg_ = false;
  SamplesMsg samples_msg_;
  std::vector<int> indices_;
};

class ImportanceSampling : public GraspDetector {
 public:
  explicit ImportanceSampling(const Params& prm) : GraspDetector(prm) { ag2_default_is_params(&ip_); }
  ag2_is_params& samplingParameters() { return ip_; }

  // importance_sampling.cpp:30-112 (the unconditional plot at :103 is left out: no graphics stack)
  std::vector<GraspHypothesis> detectGraspPoses(const CloudCamera& cloud_cam_in, bool /*clusters*/ = true) override {
    if (cloud_cam_in.getCloudProcessed().size() == 0) return std::vector<GraspHypothesis>();
    ip_.seed = prm_.rng_seed;
    if (group_ && !use_incoming_samples_ && cloud_cam_in.getNormals().cols() == 0) return detectOnGroup(cloud_cam_in, &ip_);
    ag2shim::check(ag2_set_params(ctx_, &p_), "ag2_set_params", ctx_);
    ag2shim::upload(ctx_, cloud_cam_in, use_incoming_samples_);
    ip_.seed = prm_.rng_seed;
    int rc = ag2_detect_importance(ctx_, &ip_);
    if (rc == AG2_E_NO_SAMPLES || rc == AG2_E_EMPTY_CLOUD) return std::vector<GraspHypothesis>();
    ag2shim::check(rc, "ag2_detect_importance", ctx_);
    return ag2shim::fetch(ctx_, p_.min_inliers > 0 ? AG2_LIST_CLUSTERS : AG2_LIST_IMPORTANCE);
  }

 private:
  ag2_is_params ip_;
};

#endif
