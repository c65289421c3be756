// Value types of the header-only shims in this directory.  Where <Eigen/Dense> can be included (the reference's own
// build) Vector3d / Matrix3Xd / Matrix4d / MatrixXi ARE Eigen's types, so reference code passes its matrices
// unchanged; otherwise minimal Eigen-free stand-ins with the same members keep the shims compilable with nothing
// but a C++17 compiler and libag2.so (define AG2SHIM_NO_EIGEN to force them).  The shims touch the matrices only
// through operator()(r, c), rows(), cols(), resize(r, c), data() and the helpers at the end of this file.
#ifndef AGILE_GRASP2_SHIM_TYPES_H_
#define AGILE_GRASP2_SHIM_TYPES_H_

#include <array>
#include <cstdint>
#include <vector>

#if defined(__has_include) && !defined(AG2SHIM_NO_EIGEN)
#if __has_include(<Eigen/Dense>)
#include <Eigen/Dense>
#define AG2SHIM_HAVE_EIGEN 1
#endif
#endif

namespace ag2shim {

#ifdef AG2SHIM_HAVE_EIGEN
using Vector3d = Eigen::Vector3d;
using Matrix3Xd = Eigen::Matrix3Xd;  // column-major 3 x N
using Matrix4d = Eigen::Matrix4d;
using MatrixXi = Eigen::MatrixXi;
inline Matrix4d identity4() { return Matrix4d::Identity(); }
inline MatrixXi zeros_i(int rows, int cols) { return MatrixXi::Zero(rows, cols); }
#else

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

inline Matrix4d identity4() { return Matrix4d(); }
inline MatrixXi zeros_i(int rows, int cols) { return MatrixXi(rows, cols); }
#endif  // AG2SHIM_HAVE_EIGEN

// row-major copy of a 4 x 4 transform (the layout of ag2_params::cam_tf_*), whatever the storage order of the type
inline void row_major16(const Matrix4d& t, double* out) {
  for (int r = 0; r < 4; r++)
    for (int c = 0; c < 4; c++) out[4 * r + c] = t(r, c);
}
// 3 x N column-major doubles -> floats (xyz xyz ...)
inline std::vector<float> to_floats(const Matrix3Xd& m) {
  std::vector<float> out((size_t)3 * (size_t)m.cols());
  const double* d = m.data();
  for (size_t i = 0; i < out.size(); i++) out[i] = (float)d[i];
  return out;
}

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
