// Minimal Eigen-free value types used by the header-only shims in this directory.  Where Eigen is
// available they map 1:1 (Eigen::Map<const Eigen::Vector3d>(v.data()) etc.); here they keep the
// shims compilable with nothing but a C++17 compiler and libag2.so.
#ifndef AGILE_GRASP2_SHIM_TYPES_H_
#define AGILE_GRASP2_SHIM_TYPES_H_

#include <array>
#include <cstdint>
#include <vector>

namespace ag2shim {

struct Vector3d {
  std::array<double, 3> v{{0, 0, 0}};
  Vector3d() = default;
  Vector3d(double x, double y, double z) : v{{x, y, z}} {}
  explicit Vector3d(const double* p) : v{{p[0], p[1], p[2]}} {}
  double& operator()(int i) { return v[i]; }
  double operator()(int i) const { return v[i]; }
  const double* data() const { return v.data(); }
  double* data() { return v.data(); }
};

// column-major 3 x N, like Eigen::Matrix3Xd
struct Matrix3Xd {
  std::vector<double> d;
  Matrix3Xd() = default;
  explicit Matrix3Xd(int cols) : d((size_t)3 * cols, 0.0) {}
  int rows() const { return 3; }
  int cols() const { return (int)(d.size() / 3); }
  void resize(int /*rows*/, int cols) { d.assign((size_t)3 * cols, 0.0); }
  double& operator()(int r, int c) { return d[(size_t)3 * c + r]; }
  double operator()(int r, int c) const { return d[(size_t)3 * c + r]; }
  const double* data() const { return d.data(); }
};

// row-major 4 x 4 homogeneous transform
struct Matrix4d {
  std::array<double, 16> m{{1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1}};
  double& operator()(int r, int c) { return m[4 * r + c]; }
  double operator()(int r, int c) const { return m[4 * r + c]; }
};

// rows x cols int matrix (camera_source_), column-major
struct MatrixXi {
  int r = 0, c = 0;
  std::vector<int> d;
  MatrixXi() = default;
  MatrixXi(int rows, int cols) : r(rows), c(cols), d((size_t)rows * cols, 0) {}
  int rows() const { return r; }
  int cols() const { return c; }
  int& operator()(int i, int j) { return d[(size_t)j * r + i]; }
  int operator()(int i, int j) const { return d[(size_t)j * r + i]; }
};

// 60 x 60 x 3 u8 image with the cv::Mat (HWC) layout
struct Image {
  int rows = 60, cols = 60;
  std::vector<uint8_t> data;
  Image() : data(60 * 60 * 3, 0) {}
  uint8_t at(int r, int c, int ch) const { return data[((size_t)r * cols + c) * 3 + ch]; }
};

// the xyz part of pcl::PointCloud<pcl::PointXYZRGBA>
struct PointCloud {
  std::vector<float> xyz;      // 3 per point
  std::vector<uint32_t> rgba;  // optional
  size_t size() const { return xyz.size() / 3; }
};

}  // namespace ag2shim
#endif
