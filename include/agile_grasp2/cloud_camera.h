// Header-only shim with the reference's CloudCamera surface (include/agile_grasp2/cloud_camera.h
// :65-184, src/agile_grasp2/cloud_camera.cpp) on top of the C ABI in ag2.h.  Host container of the
// cloud, camera sources, optional normals and samples; filterWorkspace / voxelizeCloud /
// subsampleUniformly run on the GPU through ag2_preprocess_cloud / ag2_subsample_uniformly.
#ifndef AGILE_GRASP2_SHIM_CLOUD_CAMERA_H_
#define AGILE_GRASP2_SHIM_CLOUD_CAMERA_H_

#include <ag2.h>

#include <cstdio>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "sample_selection.h"
#include "types.h"

class CloudCamera {
 public:
  typedef ag2shim::PointCloud PointCloudRGB;

  CloudCamera() {}

  // CloudCamera(const PointCloudRGB::Ptr& cloud, int size_left_cloud), cloud_camera.cpp:34-51
  CloudCamera(const PointCloudRGB& cloud, int size_left_cloud) : original_(cloud), processed_(cloud) {
    initCameraSource((int)cloud.size(), size_left_cloud);
  }

  // with precomputed normals (PointCloudNormal overload, cloud_camera.cpp:4-31)
  CloudCamera(const PointCloudRGB& cloud, const ag2shim::Matrix3Xd& normals, int size_left_cloud)
      : original_(cloud), processed_(cloud), normals_(normals) {
    initCameraSource((int)cloud.size(), size_left_cloud);
  }

  // CloudCamera(const std::string& filename), cloud_camera.cpp:54-63 (pcl::io::loadPCDFile)
  explicit CloudCamera(const std::string& filename) {
    processed_ = loadPointCloudFromFile(filename, &normals_);
    original_ = processed_;
    camera_source_ = ag2shim::zeros_i(1, (int)processed_.size());
    for (int i = 0; i < camera_source_.cols(); i++) camera_source_(0, i) = 1;
  }

  // two files: left cloud followed by right cloud, cloud_camera.cpp:66-86
  CloudCamera(const std::string& filename_left, const std::string& filename_right) {
    PointCloudRGB l = loadPointCloudFromFile(filename_left, nullptr), r = loadPointCloudFromFile(filename_right, nullptr);
    processed_ = l;
    processed_.xyz.insert(processed_.xyz.end(), r.xyz.begin(), r.xyz.end());
    processed_.rgba.clear();
    original_ = processed_;
    initCameraSource((int)processed_.size(), (int)l.size());
    if (l.size() == processed_.size() && r.size() == 0) {
      for (int i = 0; i < camera_source_.cols(); i++) camera_source_(0, i) = 1;
    }
  }

  // camera_source_ as the bit mask the C ABI takes (bit c = seen by camera c)
  std::vector<uint8_t> cameraMask() const {
    std::vector<uint8_t> m((size_t)camera_source_.cols(), 0);
    for (int j = 0; j < camera_source_.cols(); j++)
      for (int i = 0; i < camera_source_.rows(); i++)
        if (camera_source_(i, j) == 1) m[(size_t)j] |= (uint8_t)(1u << i);
    return m;
  }

  // filterWorkspace + voxelizeCloud in one device pass; `ctx` is any ag2_ctx carrying the workspace.
  // The reference calls them separately (grasp_detector.cpp:294-302); both entry points are kept.
  void filterWorkspace(ag2_ctx* ctx) { preprocess(ctx, 1, 0, 0.003); }
  void voxelizeCloud(ag2_ctx* ctx, double cell_size) { preprocess(ctx, 0, 1, cell_size); }
  void filterWorkspaceAndVoxelize(ag2_ctx* ctx, bool voxelize, double cell_size) {
    preprocess(ctx, 1, voxelize ? 1 : 0, cell_size);
  }

  // subsampleUniformly, cloud_camera.cpp:171-178 (requires that ctx holds this processed cloud)
  void subsampleUniformly(ag2_ctx* ctx, int num_samples, uint64_t seed = 0) {
    check(ag2_subsample_uniformly(ctx, num_samples, seed), "ag2_subsample_uniformly");
    size_t n = 0;
    sample_indices_.resize((size_t)num_samples);
    check(ag2_get_sample_indices(ctx, sample_indices_.data(), sample_indices_.size(), &n), "ag2_get_sample_indices");
    sample_indices_.resize(n);
  }

  // subsampleSamples, cloud_camera.cpp:181-206: at most num_samples of the incoming sample points become the xyz
  // samples.  `Msg` is agile_grasp2::SamplesMsg or anything with samples[i].x / .y / .z; std::random_shuffle is replaced
  // by the counter-based shuffle of sample_selection.h.
  template <class Msg>
  void subsampleSamples(const Msg& msg, int num_samples, uint64_t seed = 0) {
    const std::vector<int> pos = ag2shim::subsampledPositions(msg.samples.size(), num_samples, seed);
    samples_.resize(3, (int)pos.size());
    for (int i = 0; i < (int)pos.size(); i++) {
      samples_(0, i) = msg.samples[(size_t)pos[(size_t)i]].x;
      samples_(1, i) = msg.samples[(size_t)pos[(size_t)i]].y;
      samples_(2, i) = msg.samples[(size_t)pos[(size_t)i]].z;
    }
  }

  const ag2shim::Matrix3Xd& getNormals() const { return normals_; }
  const ag2shim::MatrixXi& getCameraSource() const { return camera_source_; }
  const PointCloudRGB& getCloudProcessed() const { return processed_; }
  const PointCloudRGB& getCloudOriginal() const { return original_; }
  const std::vector<int>& getSampleIndices() const { return sample_indices_; }
  const ag2shim::Matrix3Xd& getSamples() const { return samples_; }
  void setSampleIndices(const std::vector<int>& sample_indices) { sample_indices_ = sample_indices; }
  void setSamples(const ag2shim::Matrix3Xd& samples) { samples_ = samples; }
  // setSamples(const agile_grasp2::SamplesMsg&), cloud_camera.cpp:243-251: any message type with samples[i].x / .y / .z
  template <class Msg, class = decltype(std::declval<const Msg&>().samples)>
  void setSamples(const Msg& msg) {
    samples_.resize(3, (int)msg.samples.size());
    for (int i = 0; i < (int)msg.samples.size(); i++) {
      samples_(0, i) = msg.samples[(size_t)i].x;
      samples_(1, i) = msg.samples[(size_t)i].y;
      samples_(2, i) = msg.samples[(size_t)i].z;
    }
  }

  // writeNormalsToFile, cloud_camera.cpp:209-220: one "nx,ny,nz" line per column, 17 significant digits as
  // boost::lexical_cast<std::string>(double) prints them
  static void writeNormalsToFile(const std::string& filename, const ag2shim::Matrix3Xd& normals) {
    std::FILE* f = std::fopen(filename.c_str(), "w");
    if (!f) throw std::runtime_error("writeNormalsToFile: cannot open " + filename);
    for (int i = 0; i < normals.cols(); i++) std::fprintf(f, "%.17g,%.17g,%.17g\n", normals(0, i), normals(1, i), normals(2, i));
    std::fclose(f);
  }

 private:
  static void check(int rc, const char* what) {
    if (rc != AG2_OK) throw std::runtime_error(std::string(what) + ": " + ag2_strerror(rc));
  }

  void initCameraSource(int n, int size_left_cloud) {
    if (size_left_cloud == n) {  // one camera: the reference leaves the row at zero here (:44-46)
      camera_source_ = ag2shim::zeros_i(1, n);
      for (int i = 0; i < n; i++) camera_source_(0, i) = 1;  // intent: the single camera sees every point
    } else {
      camera_source_ = ag2shim::zeros_i(2, n);
      for (int i = 0; i < size_left_cloud; i++) camera_source_(0, i) = 1;
      for (int i = size_left_cloud; i < n; i++) camera_source_(1, i) = 1;
    }
  }

  void preprocess(ag2_ctx* ctx, int filter_ws, int voxelize, double cell_size) {
    std::vector<uint8_t> mask = cameraMask();
    std::vector<float> nrm;
    if (normals_.cols() == (int)processed_.size() && normals_.cols() > 0) {
      nrm = ag2shim::to_floats(normals_);
    }
    check(ag2_preprocess_cloud(ctx, processed_.xyz.data(), 12, processed_.size(), mask.data(), camera_source_.rows(),
                               nrm.empty() ? nullptr : nrm.data(), filter_ws, voxelize, cell_size),
          "ag2_preprocess_cloud");
    size_t n = 0;
    check(ag2_get_cloud(ctx, nullptr, nullptr, (size_t)-1, &n), "ag2_get_cloud");
    processed_.xyz.resize(3 * n);
    processed_.rgba.clear();
    mask.resize(n);
    if (n) check(ag2_get_cloud(ctx, processed_.xyz.data(), mask.data(), n, &n), "ag2_get_cloud");
    ag2shim::MatrixXi cs = ag2shim::zeros_i((int)camera_source_.rows(), (int)n);
    for (size_t j = 0; j < n; j++)
      for (int i = 0; i < cs.rows(); i++) cs(i, (int)j) = (mask[j] >> i) & 1;
    camera_source_ = cs;
    if (voxelize || nrm.empty()) normals_ = ag2shim::Matrix3Xd();
    // (when only the workspace filter ran, precomputed normals stay inside the ctx, filtered alike)
  }

  static PointCloudRGB loadPointCloudFromFile(const std::string& filename, ag2shim::Matrix3Xd* normals) {
    PointCloudRGB cloud;
    size_t n = 0;
    int has_n = 0, has_c = 0;
    if (ag2_pcd_info(filename.c_str(), &n, &has_n, &has_c) != AG2_OK) return cloud;  // reference: empty cloud
    cloud.xyz.resize(3 * n);
    if (has_c) cloud.rgba.resize(n);
    std::vector<float> nrm(has_n && normals ? 3 * n : 0);
    if (ag2_pcd_read(filename.c_str(), cloud.xyz.data(), nrm.empty() ? nullptr : nrm.data(),
                     has_c ? cloud.rgba.data() : nullptr, n, &n) != AG2_OK) {
      cloud.xyz.clear();
      cloud.rgba.clear();
      return cloud;
    }
    if (!nrm.empty()) {
      normals->resize(3, (int)n);
      for (size_t i = 0; i < nrm.size(); i++) normals->d[i] = (double)nrm[i];
    }
    return cloud;
  }

  PointCloudRGB original_, processed_;
  ag2shim::MatrixXi camera_source_;
  ag2shim::Matrix3Xd normals_;
  std::vector<int> sample_indices_;
  ag2shim::Matrix3Xd samples_;
};

#endif
