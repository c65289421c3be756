/*
 * ag2_oracle.cpp — CPU ORACLE for the grasp-hypothesis hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product (agile_grasp2_b200/, libag2.so)
 * may include, link, load or execute this file; it exists so tests/, smoke() and
 * bench.py's cpu_baseline / --impl reference legs can check and time-compare the
 * CUDA path.
 *
 * PARITY UNPINNED: the reference (gwding/agile_grasp2) ships no golden vectors,
 * known-answer tests or fixtures for this path, and none of its
 * translation units can be compiled here (each needs PCL / Eigen / OpenCV-C++ /
 * Caffe / ROS, none installed, no network).  This file is therefore a plain
 * C++17 RESTATEMENT of the reference's algorithm, stage by stage, following
 *   src/agile_grasp2/hand_search.cpp:4-61,83-235,319-426
 *   src/agile_grasp2/local_frame.cpp:4-13,26-59
 *   src/agile_grasp2/finger_hand.cpp:4-14,17-93,96-134,137-214,313-325
 *   src/agile_grasp2/antipodal.cpp:3-84
 *   src/agile_grasp2/grasp_detector.cpp:140-221,239-252,363-395
 *   src/agile_grasp2/learning.cpp:4-16,143-249
 *   src/agile_grasp2/caffe_classifier.cpp:94-198 + caffe/sigmoid_lenet.prototxt
 * Third-party arithmetic that is absent from /root/reference is restated from
 * the libraries' published behaviour (versions implied by ROS Indigo):
 *   PCL 1.7  KdTreeFLANN::radiusSearch / FLANN 1.8 L2_Simple      -> radius_search()
 *   PCL 1.7  computeMeanAndCovarianceMatrix, eigen33, computeRoots,
 *            solvePlaneParameters, flipNormalTowardsViewpoint      -> pcl_normal()
 *   Eigen 3.2 products / reductions (association orders noted inline)
 *   OpenCV 2.4 dilate / cvtColor / convertTo                        -> image_postprocess()
 *   Caffe Convolution / Pooling(MAX) / InnerProduct / ReLU          -> lenet_forward()
 * What CAN be pinned in this container is pinned by tests/ against real
 * libraries: neighbour sets against scipy.spatial.cKDTree, eigen-solves against
 * numpy.linalg.eigh, the image tail against cv2 (OpenCV 4.13), LeNet against
 * torch.nn.functional; fixtures generated that way live in tests/golden/.
 *
 * Documented deviations from the literal reference (all forced):
 *  D1 rand() (racy, time-seeded; hand_search.cpp:130) -> counter-based
 *     draw(seed, sample_id, j) below.
 *  D2 float atan2/cos/sin inside PCL computeRoots are evaluated in double and
 *     rounded to float (= correctly rounded float results; glibc's float
 *     versions are within 1 ulp of that).
 *  D3 Eigen::EigenSolver (general Hessenberg/QR) on the symmetric 3x3 scatter
 *     matrix -> cyclic Jacobi; identical up to rounding away from degenerate spectra.
 *  D4 std::partial_sort ties (unstable) -> ties keep list order.
 *
 * Build: g++ -O3 -fopenmp -ffp-contract=off -fno-fast-math (oracle/Makefile).
 */
#include "ag2_oracle.h"
#include "../include/ag2_detmath.h"  // shared with the device sampler by design (deviation D5)

#include <omp.h>

#include <algorithm>
#include <array>
#include <cmath>
#include <map>
#include <cstdint>
#include <cstring>
#include <limits>
#include <string>
#include <utility>
#include <vector>

namespace {

// ------------------------------------------------------------------ helpers

struct V3 {
  double x, y, z;
};
inline V3 v3(double x, double y, double z) { return V3{x, y, z}; }
// Eigen fixed-size 3-vector reductions unroll as a0 + (a1 + a2) (Redux.h
// redux_novec_unroller, Length 3 -> halves of 1 and 2).
inline double dot3(const V3& a, const V3& b) { return a.x * b.x + (a.y * b.y + a.z * b.z); }
inline double norm3(const V3& a) { return std::sqrt(dot3(a, a)); }
inline V3 cross3(const V3& a, const V3& b) {
  return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
inline V3 neg3(const V3& a) { return v3(-a.x, -a.y, -a.z); }

// D1: replacement for rand() % n in hand_search.cpp:130.
inline uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
inline uint32_t draw(uint64_t seed, int64_t sample_id, int j, uint32_t n) {
  uint64_t h = splitmix64(seed ^ splitmix64(((uint64_t)sample_id << 8) ^ (uint64_t)j));
  return (uint32_t)(h % (uint64_t)n);
}

struct Neighbor {
  float d2;
  int32_t idx;
  bool operator<(const Neighbor& o) const {  // FLANN DistanceIndex::operator<
    return d2 < o.d2 || (d2 == o.d2 && idx < o.idx);
  }
};

// ------------------------------------------------------------------ kd-tree
// Bucketed kd-tree (leaf size 15 like PCL's KdTreeFLANN default) used for the
// timed baseline.  Membership rule is the exact FLANN/PCL one; pruning is
// conservative (double, with slack) so the tree returns exactly the brute-force set.
struct KdTree {
  struct Node {
    int32_t left, right;  // children or -1
    int32_t lo, hi;       // leaf range in perm
    int32_t dim;
    float split;
  };
  std::vector<Node> nodes;
  std::vector<int32_t> perm;
  std::vector<float> pts;  // reordered xyz (reorder_ = true in FLANN's single index)
  const float* xyz = nullptr;

  void build(const float* p, size_t n) {
    xyz = p;
    perm.clear();
    nodes.clear();
    perm.reserve(n);
    for (size_t i = 0; i < n; i++) {
      // PCL convertCloudToArray: non-finite points never enter the tree.
      if (std::isfinite(p[3 * i]) && std::isfinite(p[3 * i + 1]) && std::isfinite(p[3 * i + 2]))
        perm.push_back((int32_t)i);
    }
    if (!perm.empty()) {
      nodes.reserve(perm.size() / 4 + 16);
      build_rec(0, (int32_t)perm.size());
    }
    pts.resize(perm.size() * 3);
    for (size_t i = 0; i < perm.size(); i++)
      for (int d = 0; d < 3; d++) pts[3 * i + d] = p[3 * (size_t)perm[i] + d];
  }
  int32_t build_rec(int32_t lo, int32_t hi) {
    int32_t id = (int32_t)nodes.size();
    nodes.push_back(Node{-1, -1, lo, hi, 0, 0.f});
    if (hi - lo <= 15) return id;
    float mn[3], mx[3];
    for (int d = 0; d < 3; d++) mn[d] = mx[d] = xyz[3 * (size_t)perm[lo] + d];
    for (int32_t i = lo + 1; i < hi; i++)
      for (int d = 0; d < 3; d++) {
        float v = xyz[3 * (size_t)perm[i] + d];
        mn[d] = std::min(mn[d], v);
        mx[d] = std::max(mx[d], v);
      }
    int dim = 0;
    for (int d = 1; d < 3; d++)
      if (mx[d] - mn[d] > mx[dim] - mn[dim]) dim = d;
    if (!(mx[dim] > mn[dim])) return id;  // all points coincide
    int32_t mid = lo + (hi - lo) / 2;
    const float* base = xyz;
    std::nth_element(perm.begin() + lo, perm.begin() + mid, perm.begin() + hi,
                     [base, dim](int32_t a, int32_t b) {
                       float va = base[3 * (size_t)a + dim], vb = base[3 * (size_t)b + dim];
                       return va < vb || (va == vb && a < b);
                     });
    float split = xyz[3 * (size_t)perm[mid] + dim];
    int32_t l = build_rec(lo, mid);
    int32_t r = build_rec(mid, hi);
    nodes[id].left = l;
    nodes[id].right = r;
    nodes[id].dim = dim;
    nodes[id].split = split;
    return id;
  }
  void query(const float q[3], float r2, std::vector<Neighbor>& out) const {
    if (nodes.empty()) return;
    query_rec(0, q, r2, (double)r2 * (1.0 + 1e-5) + 1e-30, out);
  }
  void query_rec(int32_t id, const float q[3], float r2, double r2_slack,
                 std::vector<Neighbor>& out) const {
    const Node& nd = nodes[id];
    if (nd.left < 0) {
      for (int32_t i = nd.lo; i < nd.hi; i++) {
        const float* p = &pts[3 * (size_t)i];
        // FLANN L2_Simple: result += diff*diff, sequential over x,y,z, float.
        float dx = q[0] - p[0], dy = q[1] - p[1], dz = q[2] - p[2];
        float d2 = (dx * dx + dy * dy) + dz * dz;
        if (d2 < r2) out.push_back(Neighbor{d2, perm[i]});  // strict <, RadiusResultSet::addPoint
      }
      return;
    }
    double diff = (double)q[nd.dim] - (double)nd.split;
    int32_t near = diff < 0 ? nd.left : nd.right;
    int32_t far = diff < 0 ? nd.right : nd.left;
    query_rec(near, q, r2, r2_slack, out);
    if (diff * diff <= r2_slack) query_rec(far, q, r2, r2_slack, out);
  }
};

// ------------------------------------------------------------------ PCL normal (S1)

// pcl::computeRoots2 (common/eigen.hpp)
inline void compute_roots2(float b, float c, float roots[3]) {
  roots[0] = 0.f;
  float d = (float)(b * b - 4.0 * c);  // Scalar(b*b - 4.0*c): double arithmetic then cast
  if (d < 0.0f) d = 0.0f;
  float sd = std::sqrt(d);
  roots[2] = 0.5f * (b + sd);
  roots[1] = 0.5f * (b - sd);
}

// pcl::computeRoots for a symmetric 3x3 float matrix (row-major m[9]).
inline void compute_roots(const float m[9], float roots[3]) {
  float m00 = m[0], m01 = m[1], m02 = m[2], m11 = m[4], m12 = m[5], m22 = m[8];
  float c0 = m00 * m11 * m22 + 2.f * m01 * m02 * m12 - m00 * m12 * m12 - m11 * m02 * m02 -
             m22 * m01 * m01;
  float c1 = m00 * m11 - m01 * m01 + m00 * m22 - m02 * m02 + m11 * m22 - m12 * m12;
  float c2 = m00 + m11 + m22;
  if (std::fabs(c0) < std::numeric_limits<float>::epsilon()) {
    compute_roots2(c2, c1, roots);
    return;
  }
  const float s_inv3 = (float)(1.0 / 3.0);
  const float s_sqrt3 = std::sqrt(3.0f);
  float c2_over_3 = c2 * s_inv3;
  float a_over_3 = (c1 - c2 * c2_over_3) * s_inv3;
  if (a_over_3 > 0.f) a_over_3 = 0.f;
  float half_b = 0.5f * (c0 + c2_over_3 * (2.f * c2_over_3 * c2_over_3 - c1));
  float q = half_b * half_b + a_over_3 * a_over_3 * a_over_3;
  if (q > 0.f) q = 0.f;
  float rho = std::sqrt(-a_over_3);
  // D2: trig in double, rounded to float.
  float theta = (float)std::atan2((double)std::sqrt(-q), (double)half_b) * s_inv3;
  float cos_theta = (float)std::cos((double)theta);
  float sin_theta = (float)std::sin((double)theta);
  roots[0] = c2_over_3 + 2.f * rho * cos_theta;
  roots[1] = c2_over_3 - rho * (cos_theta + s_sqrt3 * sin_theta);
  roots[2] = c2_over_3 - rho * (cos_theta - s_sqrt3 * sin_theta);
  if (roots[0] >= roots[1]) std::swap(roots[0], roots[1]);
  if (roots[1] >= roots[2]) {
    std::swap(roots[1], roots[2]);
    if (roots[0] >= roots[1]) std::swap(roots[0], roots[1]);
  }
  if (roots[0] <= 0.f) compute_roots2(c2, c1, roots);
}

// pcl::eigen33(mat, eigenvalue, eigenvector): smallest eigenpair.
inline void eigen33(const float cov[9], float& eigenvalue, float ev[3]) {
  float scale = 0.f;
  for (int i = 0; i < 9; i++) scale = std::max(scale, std::fabs(cov[i]));
  if (scale <= std::numeric_limits<float>::min()) scale = 1.0f;
  float m[9];
  for (int i = 0; i < 9; i++) m[i] = cov[i] / scale;
  float roots[3];
  compute_roots(m, roots);
  eigenvalue = roots[0] * scale;
  m[0] -= roots[0];
  m[4] -= roots[0];
  m[8] -= roots[0];
  const float* r0 = &m[0];
  const float* r1 = &m[3];
  const float* r2 = &m[6];
  auto cross = [](const float* a, const float* b, float* o) {
    o[0] = a[1] * b[2] - a[2] * b[1];
    o[1] = a[2] * b[0] - a[0] * b[2];
    o[2] = a[0] * b[1] - a[1] * b[0];
  };
  float v1[3], v2[3], v3_[3];
  cross(r0, r1, v1);
  cross(r0, r2, v2);
  cross(r1, r2, v3_);
  auto sq = [](const float* a) { return a[0] * a[0] + (a[1] * a[1] + a[2] * a[2]); };
  float l1 = sq(v1), l2 = sq(v2), l3 = sq(v3_);
  const float* best;
  float len;
  if (l1 >= l2 && l1 >= l3) {
    best = v1;
    len = l1;
  } else if (l2 >= l1 && l2 >= l3) {
    best = v2;
    len = l2;
  } else {
    best = v3_;
    len = l3;
  }
  float s = std::sqrt(len);
  for (int i = 0; i < 3; i++) ev[i] = best[i] / s;
}

// One point normal: computeMeanAndCovarianceMatrix (single-pass float moments in
// neighbour-list order) + solvePlaneParameters + flipNormalTowardsViewpoint(0,0,0).
inline void pcl_normal(const float* xyz, const std::vector<Neighbor>& nn, const float p[3],
                       float out[3]) {
  const float nan = std::numeric_limits<float>::quiet_NaN();
  if (nn.size() < 3) {
    out[0] = out[1] = out[2] = nan;
    return;
  }
  float a[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  for (const Neighbor& nb : nn) {
    const float* q = &xyz[3 * (size_t)nb.idx];
    a[0] += q[0] * q[0];
    a[1] += q[0] * q[1];
    a[2] += q[0] * q[2];
    a[3] += q[1] * q[1];
    a[4] += q[1] * q[2];
    a[5] += q[2] * q[2];
    a[6] += q[0];
    a[7] += q[1];
    a[8] += q[2];
  }
  // Eigen 3.2 DenseBase::operator/=(scalar) multiplies by the reciprocal.
  float inv = 1.0f / (float)nn.size();
  for (int i = 0; i < 9; i++) a[i] *= inv;
  float cov[9];
  cov[0] = a[0] - a[6] * a[6];
  cov[1] = a[1] - a[6] * a[7];
  cov[2] = a[2] - a[6] * a[8];
  cov[4] = a[3] - a[7] * a[7];
  cov[5] = a[4] - a[7] * a[8];
  cov[8] = a[5] - a[8] * a[8];
  cov[3] = cov[1];
  cov[6] = cov[2];
  cov[7] = cov[5];
  float eval, n[3];
  eigen33(cov, eval, n);
  // flipNormalTowardsViewpoint, viewpoint (0,0,0) (hand_search.cpp:88)
  float vx = 0.f - p[0], vy = 0.f - p[1], vz = 0.f - p[2];
  float cos_theta = (vx * n[0] + vy * n[1]) + vz * n[2];
  if (cos_theta < 0) {
    n[0] *= -1;
    n[1] *= -1;
    n[2] *= -1;
  }
  out[0] = n[0];
  out[1] = n[1];
  out[2] = n[2];
}

// ------------------------------------------------------------------ local frame (S2)

// D3: cyclic Jacobi eigen-decomposition of a symmetric 3x3 (double).
// a: in (symmetric), destroyed; d: eigenvalues; v: eigenvectors in columns (v[r][c]).
inline void jacobi3(double a[3][3], double d[3], double v[3][3]) {
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) v[i][j] = (i == j) ? 1.0 : 0.0;
  for (int sweep = 0; sweep < 32; sweep++) {
    double off = std::fabs(a[0][1]) + std::fabs(a[0][2]) + std::fabs(a[1][2]);
    if (off == 0.0) break;
    for (int p = 0; p < 2; p++)
      for (int q = p + 1; q < 3; q++) {
        double apq = a[p][q];
        if (apq == 0.0) continue;
        double g = 100.0 * std::fabs(apq);
        if (sweep > 3 && std::fabs(a[p][p]) + g == std::fabs(a[p][p]) &&
            std::fabs(a[q][q]) + g == std::fabs(a[q][q])) {
          a[p][q] = a[q][p] = 0.0;
          continue;
        }
        double h = a[q][q] - a[p][p];
        double t;
        if (std::fabs(h) + g == std::fabs(h)) {
          t = apq / h;
        } else {
          double theta = 0.5 * h / apq;
          t = 1.0 / (std::fabs(theta) + std::sqrt(1.0 + theta * theta));
          if (theta < 0.0) t = -t;
        }
        double c = 1.0 / std::sqrt(1.0 + t * t);
        double s = t * c;
        // A <- J^T A J with J the rotation in the (p,q) plane.
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

// LocalFrame::findAverageNormalAxis (local_frame.cpp:26-59).  N: n unit normals.
inline void average_normal_axis(const std::vector<V3>& N, const V3& sample, const V3& cam_origin,
                                V3& normal, V3& binormal, V3& axis) {
  const int n = (int)N.size();
  // 1. M = N N^T (sum over drawn normals in draw order), min eigenvector.
  double M[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
  for (int j = 0; j < n; j++) {
    const double c[3] = {N[j].x, N[j].y, N[j].z};
    for (int a = 0; a < 3; a++)
      for (int b = 0; b < 3; b++) M[a][b] += c[a] * c[b];
  }
  double d[3], v[3][3];
  jacobi3(M, d, v);
  int min_index = 0;
  for (int i = 1; i < 3; i++)
    if (d[i] < d[min_index]) min_index = i;  // minCoeff: first minimum
  V3 c = v3(v[0][min_index], v[1][min_index], v[2][min_index]);
  // 2. column of (N^T N)^.6 with the largest column sum (local_frame.cpp:42).
  int max_index = 0;
  double best = -std::numeric_limits<double>::infinity();
  for (int j = 0; j < n; j++) {
    double s = 0.0;
    for (int i = 0; i < n; i++) {
      double gij = (N[i].x * N[j].x + N[i].y * N[j].y) + N[i].z * N[j].z;
      double g2 = gij * gij;
      s += (g2 * g2) * g2;  // pow(g, 6)
    }
    if (s > best) {  // maxCoeff: first maximum
      best = s;
      max_index = j;
    }
  }
  const V3& nm = N[max_index];
  // normpartial = (I - c c^T) n_max, row-by-row coefficient product.
  double cc[3] = {c.x, c.y, c.z};
  double nv[3] = {nm.x, nm.y, nm.z};
  double np[3];
  for (int a = 0; a < 3; a++) {
    double P0 = (a == 0 ? 1.0 : 0.0) - cc[a] * cc[0];
    double P1 = (a == 1 ? 1.0 : 0.0) - cc[a] * cc[1];
    double P2 = (a == 2 ? 1.0 : 0.0) - cc[a] * cc[2];
    np[a] = (P0 * nv[0] + P1 * nv[1]) + P2 * nv[2];
  }
  V3 npv = v3(np[0], np[1], np[2]);
  double len = norm3(npv);
  normal = v3(npv.x / len, npv.y / len, npv.z / len);
  // 3. binormal, 4. orientation toward the majority camera.
  binormal = cross3(c, normal);
  V3 s2s = v3(sample.x - cam_origin.x, sample.y - cam_origin.y, sample.z - cam_origin.z);
  if (dot3(normal, s2s) > 0) normal = neg3(normal);
  if (dot3(binormal, s2s) > 0) binormal = neg3(binormal);
  axis = cross3(normal, binormal);
}

// ------------------------------------------------------------------ tables (S5, S6, S8)

// Eigen 3.2 LinSpaced (random access): low + i*step, step = (high-low)/(size-1).
inline void finger_spacing(double fw, double od, double fs[20]) {
  double step = ((od - fw) - 0.0) / 9.0;
  for (int i = 0; i < 10; i++) {
    double h = 0.0 + (double)i * step;
    fs[i] = (h - od) + fw;  // (fs_half.array() - od + fw), finger_hand.cpp:12
    fs[10 + i] = h;
  }
}
inline int deepen_depths(double init_bite, double hand_depth, double* out, int cap) {
  int n = 0;
  // finger_hand.cpp:122: repeated double addition.
  for (double depth = init_bite + 0.005; depth <= hand_depth; depth += 0.005) {
    if (n < cap) out[n] = depth;
    n++;
  }
  return n;
}
inline void orientation_angles(int R, double* out) {
  // hand_search.cpp:179-180
  double low = -1.0 * M_PI / 2.0, high = M_PI / 2.0;
  double step = (high - low) / (double)R;  // size-1 == R
  for (int i = 0; i < R; i++) out[i] = low + (double)i * step;
}

// ------------------------------------------------------------------ image tail (S14)

inline uint8_t saturate_u8(float v) {
  // cv::saturate_cast<uchar>(float): cvRound (round-half-even) then clamp.
  long r = std::lrintf(v);
  return (uint8_t)(r < 0 ? 0 : (r > 255 ? 255 : r));
}
inline void image_postprocess(const float* img, uint8_t* out) {
  const int S = AG2_IMAGE_SIZE;
  for (int r = 0; r < S; r++)
    for (int c = 0; c < S; c++)
      for (int ch = 0; ch < 3; ch++) {
        // cv::dilate, 3x3 MORPH_RECT, default border (ignored): max over in-bounds window.
        float m = -std::numeric_limits<float>::infinity();
        for (int dr = -1; dr <= 1; dr++)
          for (int dc = -1; dc <= 1; dc++) {
            int rr = r + dr, cc = c + dc;
            if (rr < 0 || rr >= S || cc < 0 || cc >= S) continue;
            m = std::max(m, img[(rr * S + cc) * 3 + ch]);
          }
        // cvtColor BGR2RGB: channel ch -> 2-ch; convertTo(CV_8UC3, 255.0).
        out[(r * S + c) * 3 + (2 - ch)] = saturate_u8(m * 255.0f);
      }
}

// ------------------------------------------------------------------ LeNet (S15)

struct LeNet {
  bool loaded = false;
  std::vector<float> c1w, c1b, c2w, c2b, i1w, i1b, i2w, i2b;
};

// conv (cross-correlation, stride 1, no pad) via im2col + GEMM like Caffe.
static void conv_im2col(const float* in, int C, int H, int W, const float* w, const float* b, int M,
                        int KS, float* col, float* out) {
  const int OH = H - KS + 1, OW = W - KS + 1, P = OH * OW, K = C * KS * KS;
  for (int c = 0; c < C; c++)
    for (int ky = 0; ky < KS; ky++)
      for (int kx = 0; kx < KS; kx++) {
        float* dst = col + (size_t)((c * KS + ky) * KS + kx) * P;
        for (int y = 0; y < OH; y++)
          std::memcpy(dst + y * OW, in + ((size_t)c * H + y + ky) * W + kx, sizeof(float) * OW);
      }
  for (int m = 0; m < M; m++) {
    float* o = out + (size_t)m * P;
    for (int p = 0; p < P; p++) o[p] = 0.f;
    for (int k = 0; k < K; k++) {
      float wk = w[(size_t)m * K + k];
      const float* cp = col + (size_t)k * P;
      for (int p = 0; p < P; p++) o[p] += wk * cp[p];
    }
    float bb = b[m];  // Caffe adds the bias after the GEMM
    for (int p = 0; p < P; p++) o[p] += bb;
  }
}
static void maxpool2(const float* in, int C, int H, int W, float* out) {
  const int OH = H / 2, OW = W / 2;
  for (int c = 0; c < C; c++)
    for (int y = 0; y < OH; y++)
      for (int x = 0; x < OW; x++) {
        const float* p = in + ((size_t)c * H + 2 * y) * W + 2 * x;
        out[((size_t)c * OH + y) * OW + x] = std::max(std::max(p[0], p[1]), std::max(p[W], p[W + 1]));
      }
}
// images: n HWC u8; logits: n x 2.
static void lenet_forward(const LeNet& net, const uint8_t* images, size_t n, float* logits,
                          int num_threads) {
#pragma omp parallel num_threads(num_threads)
  {
    std::vector<float> in(3 * 60 * 60), col((size_t)500 * 576 > (size_t)75 * 3136 ? 500 * 576 : 75 * 3136),
        c1(20 * 56 * 56), p1(20 * 28 * 28), c2(50 * 24 * 24), p2(50 * 12 * 12), h(500);
#pragma omp for schedule(dynamic, 4)
    for (long long im = 0; im < (long long)n; im++) {
      const uint8_t* src = images + (size_t)im * AG2_IMAGE_BYTES;
      // PreprocessBatch: convertTo(CV_32FC3) + cv::split -> planar, no mean, no scale.
      for (int px = 0; px < 3600; px++)
        for (int ch = 0; ch < 3; ch++) in[ch * 3600 + px] = (float)src[px * 3 + ch];
      conv_im2col(in.data(), 3, 60, 60, net.c1w.data(), net.c1b.data(), 20, 5, col.data(), c1.data());
      maxpool2(c1.data(), 20, 56, 56, p1.data());
      conv_im2col(p1.data(), 20, 28, 28, net.c2w.data(), net.c2b.data(), 50, 5, col.data(), c2.data());
      maxpool2(c2.data(), 50, 24, 24, p2.data());
      for (int o = 0; o < 500; o++) {
        const float* wr = &net.i1w[(size_t)o * 7200];
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};  // blocked dot product (BLAS-like)
        for (int k = 0; k < 7200; k += 8)
          for (int u = 0; u < 8; u++) acc[u] += wr[k + u] * p2[k + u];
        float s = ((acc[0] + acc[1]) + (acc[2] + acc[3])) + ((acc[4] + acc[5]) + (acc[6] + acc[7]));
        s += net.i1b[o];
        h[o] = s > 0.f ? s : 0.f;  // ReLU in place
      }
      for (int o = 0; o < 2; o++) {
        const float* wr = &net.i2w[(size_t)o * 500];
        float s = 0.f;
        for (int k = 0; k < 500; k++) s += wr[k] * h[k];
        logits[2 * im + o] = s + net.i2b[o];
      }
    }
  }
}

// ------------------------------------------------------------------ context

struct Hyp {
  ag2_hypothesis rec;
  std::vector<int32_t> in_box;   // cloud ids, reference list order
  std::vector<double> pts;       // scaled in-box points, 3 per point (points_for_learning_)
  std::vector<double> nrm;       // rotated in-box normals
};

}  // namespace

struct ag2o_ctx {
  ag2_params p;
  std::vector<float> xyz;  // packed
  std::vector<uint8_t> cam;
  int num_cams = 1;
  bool have_normals = false, normals_given = false;
  std::vector<float> normals;
  bool tree_built = false;
  KdTree tree;
  bool use_points = false;
  std::vector<int32_t> sample_idx;
  std::vector<double> sample_xyz;
  int64_t sample_id_offset = 0, sample_id_stride = 1;
  std::vector<ag2_frame> frames;
  std::vector<ag2_sweep_record> sweeps;
  std::vector<Hyp> hyps;
  std::vector<size_t> pruned, antipodal, selected;  // indices into hyps
  std::vector<double> scores;                      // per hyps entry (after classify)
  std::vector<uint8_t> images;                     // pruned x 10800
  std::vector<float> logits;                       // pruned x 2
  bool have_hyps = false, have_pruned = false, have_images = false, have_antipodal = false;
  LeNet net;
  ag2_stats st;
  std::vector<ag2_hypothesis> is_hands;            // importance sampling result (all rounds)
  std::vector<std::vector<double>> is_round_samples;
  std::vector<ag2_hypothesis> clusters;            // findClusters result
  std::vector<ag2_hypothesis> selected_rec;        // selected list when clustering moved the hands
  bool selected_is_rec = false;
  size_t num_samples() const { return use_points ? sample_xyz.size() / 3 : sample_idx.size(); }
  int threads() const { return p.num_threads > 0 ? p.num_threads : 1; }
};

namespace {

void radius_search(const ag2o_ctx* c, const float q[3], double radius, std::vector<Neighbor>& out) {
  out.clear();
  float r2 = (float)(radius * radius);  // PCL: static_cast<float>(radius * radius)
  c->tree.query(q, r2, out);
  std::sort(out.begin(), out.end());
}

void radius_search_brute(const ag2o_ctx* c, const float q[3], double radius,
                         std::vector<Neighbor>& out) {
  out.clear();
  float r2 = (float)(radius * radius);
  size_t n = c->xyz.size() / 3;
  for (size_t i = 0; i < n; i++) {
    const float* p = &c->xyz[3 * i];
    if (!(std::isfinite(p[0]) && std::isfinite(p[1]) && std::isfinite(p[2]))) continue;
    float dx = q[0] - p[0], dy = q[1] - p[1], dz = q[2] - p[2];
    float d2 = (dx * dx + dy * dy) + dz * dz;
    if (d2 < r2) out.push_back(Neighbor{d2, (int32_t)i});
  }
  std::sort(out.begin(), out.end());
}

void ensure_tree(ag2o_ctx* c) {
  if (c->tree_built) return;
  double t0 = omp_get_wtime();
  c->tree.build(c->xyz.data(), c->xyz.size() / 3);
  c->tree_built = true;
  c->st.ms_phase[0] = (omp_get_wtime() - t0) * 1e3;
}

// hand_search.cpp:83-94
void compute_normals(ag2o_ctx* c) {
  ensure_tree(c);
  double t0 = omp_get_wtime();
  size_t n = c->xyz.size() / 3;
  c->normals.assign(3 * n, 0.f);
  long long sum_kn = 0;
#pragma omp parallel num_threads(c->threads()) reduction(+ : sum_kn)
  {
    std::vector<Neighbor> nn;
#pragma omp for schedule(dynamic, 256)
    for (long long i = 0; i < (long long)n; i++) {
      const float* p = &c->xyz[3 * (size_t)i];
      float* o = &c->normals[3 * (size_t)i];
      if (!(std::isfinite(p[0]) && std::isfinite(p[1]) && std::isfinite(p[2]))) {
        o[0] = o[1] = o[2] = std::numeric_limits<float>::quiet_NaN();
        continue;
      }
      radius_search(c, p, c->p.normals_radius, nn);
      sum_kn += (long long)nn.size();
      pcl_normal(c->xyz.data(), nn, p, o);
    }
  }
  c->have_normals = true;
  c->st.sum_neighbors_normals = sum_kn;
  c->st.ms_phase[1] = (omp_get_wtime() - t0) * 1e3;
}

void sample_float(const ag2o_ctx* c, size_t s, float q[3]) {
  if (c->use_points) {
    for (int d = 0; d < 3; d++) q[d] = (float)c->sample_xyz[3 * s + d];  // hand_search.cpp:262-264
  } else {
    const float* p = &c->xyz[3 * (size_t)c->sample_idx[s]];
    q[0] = p[0];
    q[1] = p[1];
    q[2] = p[2];
  }
}

// hand_search.cpp:97-170 / :238-317
void compute_frames(ag2o_ctx* c) {
  double t0 = omp_get_wtime();
  size_t S = c->num_samples();
  c->frames.assign(S, ag2_frame{});
  V3 cam_origin[2] = {v3(c->p.cam_tf_left[3], c->p.cam_tf_left[7], c->p.cam_tf_left[11]),
                      v3(c->p.cam_tf_right[3], c->p.cam_tf_right[7], c->p.cam_tf_right[11])};
#pragma omp parallel num_threads(c->threads())
  {
    std::vector<Neighbor> nn;
    std::vector<V3> N;
#pragma omp for schedule(dynamic, 8)
    for (long long s = 0; s < (long long)S; s++) {
      ag2_frame& f = c->frames[(size_t)s];
      float q[3];
      sample_float(c, (size_t)s, q);
      f.sample[0] = q[0];
      f.sample[1] = q[1];
      f.sample[2] = q[2];
      radius_search(c, q, c->p.nn_radius_taubin, nn);
      f.num_neighbors = (int32_t)nn.size();
      if (nn.empty()) continue;  // sample dropped (hand_search.cpp:122,157-163)
      int n = std::min(50, (int)nn.size());
      N.resize(n);
      int num_source[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      for (int j = 0; j < n; j++) {
        uint32_t r = draw(c->p.rng_seed, c->sample_id_offset + c->sample_id_stride * (int64_t)s, j, (uint32_t)nn.size());
        int32_t id = nn[r].idx;
        const float* nf = &c->normals[3 * (size_t)id];
        N[j] = v3((double)nf[0], (double)nf[1], (double)nf[2]);
        for (int cam = 0; cam < c->num_cams; cam++)
          if ((c->cam[(size_t)id] >> cam) & 1) num_source[cam]++;
      }
      int majority = 0;
      for (int cam = 1; cam < c->num_cams; cam++)
        if (num_source[cam] > num_source[majority]) majority = cam;
      for (int j = 0; j < n; j++) {
        double mag = std::sqrt(N[j].x * N[j].x + (N[j].y * N[j].y + N[j].z * N[j].z));
        N[j] = v3(N[j].x / mag, N[j].y / mag, N[j].z / mag);
      }
      V3 nrm, bin, axis;
      V3 sample = v3((double)q[0], (double)q[1], (double)q[2]);
      average_normal_axis(N, sample, cam_origin[majority > 1 ? 1 : majority], nrm, bin, axis);
      f.valid = 1;
      f.majority_cam = majority;
      f.normal[0] = nrm.x, f.normal[1] = nrm.y, f.normal[2] = nrm.z;
      f.binormal[0] = bin.x, f.binormal[1] = bin.y, f.binormal[2] = bin.z;
      f.curvature_axis[0] = axis.x, f.curvature_axis[1] = axis.y, f.curvature_axis[2] = axis.z;
    }
  }
  long long nf = 0;
  for (const ag2_frame& f : c->frames) nf += f.valid;
  c->st.num_frames = nf;
  c->st.ms_phase[2] = (omp_get_wtime() - t0) * 1e3;
}

// FingerHand state (finger_hand.h:48-223), restated.
struct Hand {
  double fw, od, depth;
  double fs[20];
  bool fingers[20];
  bool hand[10];
  double top = 0, bottom = 0, left = 0, right = 0, center = 0, surface = 0;

  Hand(double fw_, double od_, double depth_) : fw(fw_), od(od_), depth(depth_) {
    finger_spacing(fw, od, fs);
    for (bool& f : fingers) f = false;
    for (bool& h : hand) h = false;
  }
  // finger_hand.cpp:17-93; px/py: rotated x and y of the cropped points.
  void evaluate_fingers(const std::vector<double>& px, const std::vector<double>& py, double bite,
                        int idx) {
    top = bite;
    bottom = bite - depth;
    for (bool& f : fingers) f = false;
    std::vector<double> cx;
    for (size_t i = 0; i < py.size(); i++) {
      if (py[i] < bite) {
        cx.push_back(px[i]);
        if (py[i] < bottom) return;  // back of the hand would collide
      }
    }
    if (cx.empty()) return;
    for (int i = 0; i < 20; i++) {
      if (idx != -1 && !(i == idx || i == 10 + idx)) continue;
      int num_in_gap = 0;
      for (double x : cx)
        if (x > fs[i] && x < fs[i] + fw) num_in_gap++;
      if (num_in_gap == 0) fingers[i] = true;
    }
  }
  int num_fingers() const {
    int s = 0;
    for (bool f : fingers) s += f;
    return s;
  }
  void evaluate_hand() {  // finger_hand.cpp:313-325
    for (int i = 0; i < 10; i++) hand[i] = fingers[i] && fingers[10 + i];
  }
  // finger_hand.cpp:96-134; returns the number of successful deepen steps.
  int deepen(const std::vector<double>& px, const std::vector<double>& py, double min_depth,
             double max_depth, int& eroded) {
    std::vector<int> hand_idx;
    for (int i = 0; i < 10; i++)
      if (hand[i]) hand_idx.push_back(i);
    eroded = -1;
    if (hand_idx.empty()) return 0;
    eroded = hand_idx[(int)std::ceil(hand_idx.size() / 2.0) - 1];
    Hand nh = *this, last = *this;
    int steps = 0;
    for (double d = min_depth + 0.005; d <= max_depth; d += 0.005) {
      nh.evaluate_fingers(px, py, d, eroded);
      if (nh.num_fingers() < 2) break;
      nh.evaluate_hand();
      last = nh;
      steps++;
    }
    *this = last;
    for (int i = 0; i < 10; i++) hand[i] = (i == eroded);
    return steps;
  }
};

// Antipodal::evaluateGrasp (antipodal.cpp:8-84). pts/nrm: 3 doubles per point.
int antipodal_eval(const std::vector<double>& pts, const std::vector<double>& nrm, double thresh) {
  size_t m = pts.size() / 3;
  double minx = std::numeric_limits<double>::infinity(), maxx = -minx;
  for (size_t i = 0; i < m; i++) {
    minx = std::min(minx, pts[3 * i]);
    maxx = std::max(maxx, pts[3 * i]);
  }
  double cos_fc = std::cos(30.0 * M_PI / 180.0);
  double lthr = minx + thresh, rthr = maxx - thresh;
  bool anyl = false, anyr = false;
  double inf = std::numeric_limits<double>::infinity();
  double lmaxy = -inf, lminy = inf, lmaxz = -inf, lminz = inf;
  double rmaxy = -inf, rminy = inf, rmaxz = -inf, rminz = inf;
  for (size_t i = 0; i < m; i++) {
    double x = pts[3 * i], y = pts[3 * i + 1], z = pts[3 * i + 2];
    // l^T n = (-1*nx + 0*ny) + 0*nz ; r^T n = (1*nx + 0*ny) + 0*nz
    double ln = (-1.0 * nrm[3 * i] + 0.0 * nrm[3 * i + 1]) + 0.0 * nrm[3 * i + 2];
    double rn = (1.0 * nrm[3 * i] + 0.0 * nrm[3 * i + 1]) + 0.0 * nrm[3 * i + 2];
    if (ln > cos_fc && x < lthr) {
      anyl = true;
      lmaxy = std::max(lmaxy, y), lminy = std::min(lminy, y);
      lmaxz = std::max(lmaxz, z), lminz = std::min(lminz, z);
    }
    if (rn > cos_fc && x > rthr) {
      anyr = true;
      rmaxy = std::max(rmaxy, y), rminy = std::min(rminy, y);
      rmaxz = std::max(rmaxz, z), rminz = std::min(rminz, z);
    }
  }
  int result = 0;
  if (anyl || anyr) result = 1;
  if (anyl && anyr) {
    double top_y = std::min(lmaxy, rmaxy), bot_y = std::max(lminy, rminy);
    double top_z = std::min(lmaxz, rmaxz), bot_z = std::max(lminz, rminz);
    if (top_y > bot_y && top_z > bot_z) result = 2;
  }
  return result;
}

// hand_search.cpp:173-235 + :319-426 for one sample.
void hands_for_sample(const ag2o_ctx* c, size_t s, const double* angles, std::vector<Hyp>& out,
                      ag2_sweep_record* sweep, long long& sumK, long long& sumKc,
                      std::vector<Neighbor>& nn) {
  const ag2_params& P = c->p;
  const int R = P.num_orientations;
  for (int r = 0; r < R; r++) {
    sweep[r] = ag2_sweep_record{};
    sweep[r].slot = -1;
    sweep[r].depth_step = -1;
    sweep[r].num_in_box = -1;
  }
  const ag2_frame& f = c->frames[s];
  if (!f.valid) return;
  float q[3] = {(float)f.sample[0], (float)f.sample[1], (float)f.sample[2]};
  radius_search(c, q, P.nn_radius_hands, nn);
  if (nn.empty()) return;
  sumK += (long long)nn.size();
  const size_t K = nn.size();
  // centered neighbourhood: float subtraction, then cast (hand_search.cpp:209-210)
  // frame = [normal | binormal | curvature_axis] (hand_search.cpp:325-326)
  double F[3][3];
  for (int a = 0; a < 3; a++) {
    F[a][0] = f.normal[a];
    F[a][1] = f.binormal[a];
    F[a][2] = f.curvature_axis[a];
  }
  std::vector<double> pc, nc;  // cropped points / normals, 3 per point
  std::vector<int32_t> idc;
  pc.reserve(K * 3 / 2);
  nc.reserve(K * 3 / 2);
  for (size_t j = 0; j < K; j++) {
    const float* pf = &c->xyz[3 * (size_t)nn[j].idx];
    double px = (double)(pf[0] - q[0]), py = (double)(pf[1] - q[1]), pz = (double)(pf[2] - q[2]);
    // row 2 of frame^T * p
    double zf = (F[0][2] * px + F[1][2] * py) + F[2][2] * pz;
    if (zf > -1.0 * P.hand_height && zf < P.hand_height) {
      const float* nf = &c->normals[3 * (size_t)nn[j].idx];
      pc.push_back(px), pc.push_back(py), pc.push_back(pz);
      nc.push_back((double)nf[0]), nc.push_back((double)nf[1]), nc.push_back((double)nf[2]);
      idc.push_back(nn[j].idx);
    }
  }
  const size_t k = idc.size();
  sumKc += (long long)k;
  Hand fh(P.finger_width, P.hand_outer_diameter, P.hand_depth);
  std::vector<double> rx(k), ry(k), rz(k), nx(k), ny(k), nz(k);
  for (int r = 0; r < R; r++) {
    double ca = std::cos(angles[r]), sa = std::sin(angles[r]);
    double rot[3][3] = {{ca, -1.0 * sa, 0.0}, {sa, ca, 0.0}, {0.0, 0.0, 1.0}};
    double FR[3][3];  // frame_rot = frame * rot
    for (int a = 0; a < 3; a++)
      for (int b = 0; b < 3; b++)
        FR[a][b] = (F[a][0] * rot[0][b] + F[a][1] * rot[1][b]) + F[a][2] * rot[2][b];
    for (size_t i = 0; i < k; i++) {  // frame_rot^T * points, frame_rot^T * normals
      const double* p = &pc[3 * i];
      const double* n = &nc[3 * i];
      rx[i] = (FR[0][0] * p[0] + FR[1][0] * p[1]) + FR[2][0] * p[2];
      ry[i] = (FR[0][1] * p[0] + FR[1][1] * p[1]) + FR[2][1] * p[2];
      rz[i] = (FR[0][2] * p[0] + FR[1][2] * p[1]) + FR[2][2] * p[2];
      nx[i] = (FR[0][0] * n[0] + FR[1][0] * n[1]) + FR[2][0] * n[2];
      ny[i] = (FR[0][1] * n[0] + FR[1][1] * n[1]) + FR[2][1] * n[2];
      nz[i] = (FR[0][2] * n[0] + FR[1][2] * n[1]) + FR[2][2] * n[2];
    }
    fh.evaluate_fingers(rx, ry, P.init_bite, -1);
    ag2_sweep_record& sw = sweep[r];
    for (int i = 0; i < 20; i++) sw.fingers |= (uint32_t)fh.fingers[i] << i;
    if (!(fh.num_fingers() > 2)) continue;  // hand_search.cpp:366
    fh.evaluate_hand();
    int nh = 0;
    for (int i = 0; i < 10; i++) {
      sw.hand |= (uint16_t)((uint16_t)fh.hand[i] << i);
      nh += fh.hand[i];
    }
    if (!(nh > 0)) continue;
    int e;
    int steps = fh.deepen(rx, ry, P.init_bite, P.hand_depth, e);
    sw.slot = (int8_t)e;
    sw.depth_step = (int8_t)steps;
    // computePointsInClosingRegion (finger_hand.cpp:137-180)
    fh.left = fh.fs[e] + fh.fw;
    fh.right = fh.fs[10 + e];
    fh.center = 0.5 * (fh.left + fh.right);
    double surf = std::numeric_limits<double>::infinity();
    for (size_t i = 0; i < k; i++) surf = std::min(surf, ry[i]);
    fh.surface = surf;
    std::vector<size_t> box;
    for (size_t i = 0; i < k; i++)
      if (ry[i] < fh.top && rx[i] > fh.left && rx[i] < fh.right) box.push_back(i);
    sw.num_in_box = (int32_t)box.size();
    if (box.empty()) continue;  // hand_search.cpp:377-381
    Hyp h;
    ag2_hypothesis& g = h.rec;
    std::memset(&g, 0, sizeof(g));
    // calculateGraspParameters (finger_hand.cpp:183-214): frame*pos + sample
    double smp[3] = {f.sample[0], f.sample[1], f.sample[2]};
    for (int a = 0; a < 3; a++) {
      g.surface[a] = ((FR[a][0] * fh.center + FR[a][1] * fh.surface) + FR[a][2] * 0.0) + smp[a];
      g.bottom[a] = ((FR[a][0] * fh.center + FR[a][1] * fh.bottom) + FR[a][2] * 0.0) + smp[a];
      g.top[a] = ((FR[a][0] * fh.center + FR[a][1] * fh.top) + FR[a][2] * 0.0) + smp[a];
      g.binormal[a] = FR[a][0];
      g.approach[a] = FR[a][1];
      g.axis[a] = FR[a][2];
    }
    double mnx = std::numeric_limits<double>::infinity(), mxx = -mnx;
    for (size_t i : box) {
      mnx = std::min(mnx, rx[i]);
      mxx = std::max(mxx, rx[i]);
    }
    g.width = mxx - mnx;
    // unit-box scaling (hand_search.cpp:399-409)
    double baseline_const = 0.1;
    double left_const = fh.left - 0.5 * (baseline_const - (fh.right - fh.left));
    double lower[3] = {left_const, fh.bottom, -1.0 * P.hand_height};
    double scales[3] = {1.0 / baseline_const, 1.0 / (fh.top - fh.bottom), 1.0 / (2.0 * P.hand_height)};
    h.pts.resize(3 * box.size());
    h.nrm.resize(3 * box.size());
    h.in_box.resize(box.size());
    for (size_t j = 0; j < box.size(); j++) {
      size_t i = box[j];
      h.pts[3 * j] = scales[0] * (rx[i] - lower[0]);
      h.pts[3 * j + 1] = scales[1] * (ry[i] - lower[1]);
      h.pts[3 * j + 2] = scales[2] * (rz[i] - lower[2]);
      h.nrm[3 * j] = nx[i], h.nrm[3 * j + 1] = ny[i], h.nrm[3 * j + 2] = nz[i];
      h.in_box[j] = idc[i];
    }
    int ap = antipodal_eval(h.pts, h.nrm, 0.003);  // hand_search.cpp:415-418
    g.half_antipodal = (ap == 1 || ap == 2);
    g.full_antipodal = (ap == 2);
    g.score = 0.0;
    g.sample_id = (int32_t)s;
    g.orientation_id = r;
    g.finger_slot = (uint8_t)e;
    g.depth_step = (uint8_t)steps;
    g.num_in_box = (int32_t)box.size();
    g.hand_top = fh.top;
    g.hand_bottom = fh.bottom;
    sw.antipodal = (uint8_t)ap;
    sw.emitted = 1;
    out.push_back(std::move(h));
  }
}

void evaluate_hands(ag2o_ctx* c) {
  double t0 = omp_get_wtime();
  const int R = c->p.num_orientations;
  size_t S = c->num_samples();
  std::vector<double> angles(R);
  orientation_angles(R, angles.data());
  std::vector<std::vector<Hyp>> lists(S);
  c->sweeps.assign(S * (size_t)R, ag2_sweep_record{});
  long long sumK = 0, sumKc = 0;
#pragma omp parallel num_threads(c->threads()) reduction(+ : sumK, sumKc)
  {
    std::vector<Neighbor> nn;
#pragma omp for schedule(dynamic, 1)
    for (long long s = 0; s < (long long)S; s++)
      hands_for_sample(c, (size_t)s, angles.data(), lists[(size_t)s], &c->sweeps[(size_t)s * R], sumK,
                       sumKc, nn);
  }
  c->hyps.clear();
  long long sum_m = 0;
  for (size_t s = 0; s < S; s++)
    for (Hyp& h : lists[s]) {
      sum_m += h.rec.num_in_box;
      c->hyps.push_back(std::move(h));
    }
  c->have_hyps = true;
  c->have_pruned = c->have_images = c->have_antipodal = false;
  c->st.num_hypotheses = (int64_t)c->hyps.size();
  c->st.sum_neighbors_hands = sumK;
  c->st.sum_cropped = sumKc;
  c->st.sum_in_box = sum_m;
  c->st.ms_phase[3] = (omp_get_wtime() - t0) * 1e3;
}

// grasp_detector.cpp:148-160,363-395
void prune(ag2o_ctx* c) {
  double t0 = omp_get_wtime();
  c->pruned.clear();
  if (!c->p.prune) {
    for (size_t i = 0; i < c->hyps.size(); i++) c->pruned.push_back(i);
  } else {
    // pcl::getMinMax3D (float); arguments of pruneGraspsOnHandParameters are floats.
    float min_zf = std::numeric_limits<float>::max();
    size_t n = c->xyz.size() / 3;
    for (size_t i = 0; i < n; i++) {
      const float* p = &c->xyz[3 * i];
      if (!(std::isfinite(p[0]) && std::isfinite(p[1]) && std::isfinite(p[2]))) continue;
      min_zf = std::min(min_zf, p[2]);
    }
    float min_x = (float)c->p.workspace[0], max_x = (float)c->p.workspace[1];
    float min_y = (float)c->p.workspace[2], max_y = (float)c->p.workspace[3];
    float min_z = min_zf;
    for (size_t i = 0; i < c->hyps.size(); i++) {
      const ag2_hypothesis& g = c->hyps[i].rec;
      if (c->p.filter_half_grasps && !g.half_antipodal) continue;
      double half_width = 0.5 * c->p.hand_outer_diameter;
      double x[5], y[5], z[5];
      double* arr[3] = {x, y, z};
      for (int a = 0; a < 3; a++) {
        arr[a][0] = g.bottom[a] + half_width * g.binormal[a];
        arr[a][1] = g.bottom[a] - half_width * g.binormal[a];
        arr[a][2] = g.top[a] + half_width * g.binormal[a];
        arr[a][3] = g.top[a] - half_width * g.binormal[a];
        arr[a][4] = g.bottom[a] - 0.10 * g.approach[a];
      }
      auto mn = [](const double* v) { return *std::min_element(v, v + 5); };
      auto mx = [](const double* v) { return *std::max_element(v, v + 5); };
      double aperture = g.width;
      if (aperture >= c->p.min_aperture && aperture <= c->p.max_aperture && mn(z) >= min_z &&
          mn(y) >= min_y && mx(y) <= max_y && mn(x) >= min_x && mx(x) <= max_x)
        c->pruned.push_back(i);
    }
  }
  c->have_pruned = true;
  c->have_images = c->have_antipodal = false;
  c->st.num_pruned = (int64_t)c->pruned.size();
  c->st.ms_phase[4] = (omp_get_wtime() - t0) * 1e3;
}

// Learning::convertToImageRGB(ins, aligns) + convertTo (learning.cpp:143-209,16) for m unit-box points
// pts / nrm (3 doubles each)
void render_points(const double* pts, const double* nrm, size_t m, bool aligns, uint8_t* out);

void render_one(const Hyp& h, uint8_t* out) { render_points(h.pts.data(), h.nrm.data(), h.in_box.size(), true, out); }

void render_points(const double* h_pts, const double* h_nrm, size_t m, bool aligns, uint8_t* out) {
  const int S = AG2_IMAGE_SIZE;
  double miny = 0.0;
  if (aligns) {
    miny = std::numeric_limits<double>::infinity();
    for (size_t j = 0; j < m; j++) miny = std::min(miny, h_pts[3 * j + 1]);
  }
  double cellsize = 1.0 / (double)S;
  std::vector<int> cell(m);
  for (size_t j = 0; j < m; j++) {
    double y = aligns ? h_pts[3 * j + 1] - miny : h_pts[3 * j + 1];
    double hcd = std::floor(h_pts[3 * j] / cellsize), vcd = std::floor(y / cellsize);
    // floorVector -> int (learning.cpp:212-218); far-away points never match a cell index
    if (!(hcd > -1e6 && hcd < 1e6 && vcd > -1e6 && vcd < 1e6)) {
      cell[j] = -1;
      continue;
    }
    int hc = (int)hcd;
    int vc = (int)vcd;
    cell[j] = hc + vc * S;
  }
  // The reference scans all points per cell (O(3600 m)); the accumulation order per
  // cell is list order either way.
  std::vector<double> acc(3 * S * S, 0.0);
  std::vector<int> cnt(S * S, 0);
  for (size_t j = 0; j < m; j++) {
    int ci = cell[j];
    if (ci < 0 || ci >= S * S) continue;  // never matched by the i-loop (learning.cpp:166-174)
    acc[3 * ci] += h_nrm[3 * j];
    acc[3 * ci + 1] += h_nrm[3 * j + 1];
    acc[3 * ci + 2] += h_nrm[3 * j + 2];
    cnt[ci]++;
  }
  std::vector<float> img(3 * S * S, 0.f);
  for (int i = 0; i < S * S; i++) {
    if (cnt[i] == 0) continue;
    double ax = acc[3 * i], ay = acc[3 * i + 1], az = acc[3 * i + 2];
    double s = 1.0 / std::sqrt(ax * ax + (ay * ay + az * az));
    int row = i / S, col = i % S;
    float* px = &img[((S - 1 - row) * S + col) * 3];
    px[0] = (float)std::fabs(s * ax);
    px[1] = (float)std::fabs(s * ay);
    px[2] = (float)std::fabs(s * az);
  }
  image_postprocess(img.data(), out);
}

void render(ag2o_ctx* c) {
  double t0 = omp_get_wtime();
  size_t H = c->pruned.size();
  c->images.assign(H * (size_t)AG2_IMAGE_BYTES, 0);
#pragma omp parallel for num_threads(c->threads()) schedule(dynamic, 4)
  for (long long i = 0; i < (long long)H; i++)
    render_one(c->hyps[c->pruned[(size_t)i]], &c->images[(size_t)i * AG2_IMAGE_BYTES]);
  c->have_images = true;
  c->st.ms_phase[5] = (omp_get_wtime() - t0) * 1e3;
}

void classify(ag2o_ctx* c) {
  double t0 = omp_get_wtime();
  size_t H = c->pruned.size();
  c->logits.assign(2 * H, 0.f);
  lenet_forward(c->net, c->images.data(), H, c->logits.data(), c->threads());
  c->antipodal.clear();
  c->scores.assign(c->hyps.size(), 0.0);
  for (size_t i = 0; i < H; i++) {
    // grasp_detector.cpp:200: float difference stored in a double
    double score = (double)(c->logits[2 * i + 1] - c->logits[2 * i]);
    if (score >= c->p.min_score_diff) {
      size_t hi = c->pruned[i];
      c->scores[hi] = score;
      c->hyps[hi].rec.score = score;
      c->hyps[hi].rec.full_antipodal = 1;  // setFullAntipodal(true)
      c->antipodal.push_back(hi);
    }
  }
  c->have_antipodal = true;
  c->st.num_antipodal = (int64_t)c->antipodal.size();
  c->st.ms_phase[6] = (omp_get_wtime() - t0) * 1e3;
}

void select_geometric(ag2o_ctx* c) {
  c->antipodal.clear();
  for (size_t hi : c->pruned)
    if (c->p.antipodal_mode == AG2_MODE_NONE || c->hyps[hi].rec.full_antipodal) c->antipodal.push_back(hi);
  c->have_antipodal = true;
  c->st.num_antipodal = (int64_t)c->antipodal.size();
}

// grasp_detector.cpp:238-252 (D4: stable on ties)
// HandleSearch::findClusters(hand_list, remove_inliers = false), handle_search.cpp:4-80
void find_clusters(const std::vector<ag2_hypothesis>& hand_list, int min_inliers, std::vector<ag2_hypothesis>& hands_out) {
  const double AXIS_ALIGN_ANGLE_THRESH = 15.0 * M_PI / 180.0;
  const double AXIS_ALIGN_DIST_THRESH = 0.005;
  const double MAX_DIST_THRESH = 0.05;
  hands_out.clear();
  for (size_t i = 0; i < hand_list.size(); i++) {
    const ag2_hypothesis& hi = hand_list[i];
    int num_inliers = 0;
    double gpd[3] = {0.0, 0.0, 0.0};
    double outer[3][3];
    for (int r = 0; r < 3; r++)
      for (int q = 0; q < 3; q++) outer[r][q] = hi.axis[r] * hi.axis[q];
    double avg_score = 0.0;
    for (size_t j = 0; j < hand_list.size(); j++) {
      if (i == j) continue;
      const ag2_hypothesis& hj = hand_list[j];
      double axis_aligned = dot3(v3(hi.axis[0], hi.axis[1], hi.axis[2]), v3(hj.axis[0], hj.axis[1], hj.axis[2]));
      bool axis_aligned_binary = std::fabs(axis_aligned) > std::cos(AXIS_ALIGN_ANGLE_THRESH);
      V3 dp = v3(hi.bottom[0] - hj.bottom[0], hi.bottom[1] - hj.bottom[1], hi.bottom[2] - hj.bottom[2]);
      bool delta_pos_mag_binary = norm3(dp) <= MAX_DIST_THRESH;
      double P[3][3];
      for (int r = 0; r < 3; r++)
        for (int q = 0; q < 3; q++) P[r][q] = (r == q ? 1.0 : 0.0) - outer[r][q];
      V3 pr = v3((P[0][0] * dp.x + P[0][1] * dp.y) + P[0][2] * dp.z, (P[1][0] * dp.x + P[1][1] * dp.y) + P[1][2] * dp.z,
                 (P[2][0] * dp.x + P[2][1] * dp.y) + P[2][2] * dp.z);
      bool delta_pos_proj_mag_binary = norm3(pr) <= AXIS_ALIGN_DIST_THRESH;
      if (axis_aligned_binary && delta_pos_mag_binary && delta_pos_proj_mag_binary) {
        num_inliers++;
        for (int a = 0; a < 3; a++) gpd[a] += hj.bottom[a];
        avg_score += hj.score;
      }
    }
    if (num_inliers >= min_inliers) {
      ag2_hypothesis hand = hi;
      for (int a = 0; a < 3; a++) {
        double d = gpd[a] / (double)num_inliers - hi.bottom[a];
        hand.surface[a] = hi.surface[a] + d;
        hand.bottom[a] = hi.bottom[a] + d;
        hand.top[a] = hi.top[a] + d;
      }
      hand.score = avg_score / (double)num_inliers;
      hands_out.push_back(hand);
    }
  }
}

void select_top(ag2o_ctx* c) {
  double t0 = omp_get_wtime();
  c->selected = c->antipodal;
  c->selected_is_rec = false;
  if (c->p.min_inliers > 0 && c->p.antipodal_mode != AG2_MODE_NONE) {  // grasp_detector.cpp:229-252
    std::vector<ag2_hypothesis> src;
    for (size_t i : c->antipodal) src.push_back(c->hyps[i].rec);
    find_clusters(src, c->p.min_inliers, c->clusters);
    c->selected_rec = c->clusters;
    std::stable_sort(c->selected_rec.begin(), c->selected_rec.end(),
                     [](const ag2_hypothesis& a, const ag2_hypothesis& b) { return a.score > b.score; });
    if (c->p.num_selected > 0 && c->selected_rec.size() > (size_t)c->p.num_selected)
      c->selected_rec.resize((size_t)c->p.num_selected);
    c->selected_is_rec = true;
    c->selected.clear();
    c->st.num_selected = (int64_t)c->selected_rec.size();
    c->st.ms_phase[7] = (omp_get_wtime() - t0) * 1e3;
    return;
  }
  if (c->p.antipodal_mode != AG2_MODE_NONE) {
    std::stable_sort(c->selected.begin(), c->selected.end(), [c](size_t a, size_t b) {
      return c->hyps[a].rec.score > c->hyps[b].rec.score;
    });
    if (c->p.num_selected > 0 && c->selected.size() > (size_t)c->p.num_selected)
      c->selected.resize((size_t)c->p.num_selected);
  }
  c->st.num_selected = (int64_t)c->selected.size();
  c->st.ms_phase[7] = (omp_get_wtime() - t0) * 1e3;
}

const std::vector<size_t>* list_of(ag2o_ctx* c, int which, std::vector<size_t>& tmp) {
  switch (which) {
    case AG2_LIST_HYPOTHESES:
      tmp.resize(c->hyps.size());
      for (size_t i = 0; i < tmp.size(); i++) tmp[i] = i;
      return &tmp;
    case AG2_LIST_PRUNED:
      return &c->pruned;
    case AG2_LIST_ANTIPODAL:
      return &c->antipodal;
    case AG2_LIST_SELECTED:
      return &c->selected;
  }
  return nullptr;
}

bool params_ok(const ag2_params* p) {
  return p && p->num_orientations >= 1 && p->num_orientations <= AG2_MAX_ORIENTATIONS &&
         p->nn_radius_hands > 0 && p->nn_radius_taubin > 0 && p->normals_radius > 0 &&
         p->finger_width > 0 && p->hand_outer_diameter > p->finger_width && p->hand_depth > 0 &&
         p->hand_height > 0;
}

}  // namespace

// ------------------------------------------------------------------ C ABI

extern "C" {

int ag2o_create(const ag2_params* p, ag2o_ctx** out) {
  if (!params_ok(p) || !out) return AG2_E_INVALID;
  ag2o_ctx* c = new ag2o_ctx();
  c->p = *p;
  std::memset(&c->st, 0, sizeof(c->st));
  *out = c;
  return AG2_OK;
}
int ag2o_destroy(ag2o_ctx* c) {
  delete c;
  return AG2_OK;
}
int ag2o_set_params(ag2o_ctx* c, const ag2_params* p) {
  if (!c || !params_ok(p)) return AG2_E_INVALID;
  c->p = *p;
  return AG2_OK;
}
int ag2o_set_lenet_weights(ag2o_ctx* c, const float* c1w, const float* c1b, const float* c2w,
                           const float* c2b, const float* i1w, const float* i1b, const float* i2w,
                           const float* i2b) {
  if (!c || !c1w || !c1b || !c2w || !c2b || !i1w || !i1b || !i2w || !i2b) return AG2_E_INVALID;
  c->net.c1w.assign(c1w, c1w + 20 * 75);
  c->net.c1b.assign(c1b, c1b + 20);
  c->net.c2w.assign(c2w, c2w + 50 * 500);
  c->net.c2b.assign(c2b, c2b + 50);
  c->net.i1w.assign(i1w, i1w + 500 * 7200);
  c->net.i1b.assign(i1b, i1b + 500);
  c->net.i2w.assign(i2w, i2w + 2 * 500);
  c->net.i2b.assign(i2b, i2b + 2);
  c->net.loaded = true;
  return AG2_OK;
}
int ag2o_set_cloud(ag2o_ctx* c, const float* xyz, size_t stride, size_t n, const uint8_t* cam,
                   int num_cams, const float* normals) {
  if (!c || (n && !xyz) || stride < 12 || num_cams < 1 || num_cams > 8) return AG2_E_INVALID;
  c->xyz.resize(3 * n);
  for (size_t i = 0; i < n; i++) {
    const float* p = (const float*)((const char*)xyz + i * stride);
    c->xyz[3 * i] = p[0], c->xyz[3 * i + 1] = p[1], c->xyz[3 * i + 2] = p[2];
  }
  c->cam.assign(n, 1);
  if (cam) std::memcpy(c->cam.data(), cam, n);
  c->num_cams = num_cams;
  c->normals_given = normals != nullptr;
  c->have_normals = c->normals_given;
  if (normals) c->normals.assign(normals, normals + 3 * n);
  else c->normals.clear();
  c->tree_built = false;
  c->have_hyps = c->have_pruned = c->have_images = c->have_antipodal = false;
  std::memset(&c->st, 0, sizeof(c->st));
  c->st.num_points = (int64_t)n;
  return AG2_OK;
}
int ag2o_set_sample_indices(ag2o_ctx* c, const int32_t* idx, size_t n) {
  if (!c || (n && !idx)) return AG2_E_INVALID;
  size_t N = c->xyz.size() / 3;
  for (size_t i = 0; i < n; i++)
    if (idx[i] < 0 || (size_t)idx[i] >= N) return AG2_E_INVALID;
  c->sample_idx.assign(idx, idx + n);
  c->use_points = false;
  c->have_hyps = false;
  c->st.num_samples = (int64_t)n;
  return AG2_OK;
}
int ag2o_set_sample_points(ag2o_ctx* c, const double* xyz, size_t n) {
  if (!c || (n && !xyz)) return AG2_E_INVALID;
  c->sample_xyz.assign(xyz, xyz + 3 * n);
  c->use_points = true;
  c->have_hyps = false;
  c->st.num_samples = (int64_t)n;
  return AG2_OK;
}
int ag2o_set_sample_id_offset(ag2o_ctx* c, int64_t first) {
  if (!c) return AG2_E_INVALID;
  c->sample_id_offset = first;
  c->sample_id_stride = 1;
  return AG2_OK;
}
int ag2o_set_sample_id_map(ag2o_ctx* c, int64_t first, int64_t stride) {
  if (!c || stride < 1) return AG2_E_INVALID;
  c->sample_id_offset = first;
  c->sample_id_stride = stride;
  return AG2_OK;
}

int ag2o_compute_normals(ag2o_ctx* c) {
  if (!c) return AG2_E_INVALID;
  if (c->xyz.empty()) return AG2_E_EMPTY_CLOUD;
  ensure_tree(c);
  if (!c->normals_given) compute_normals(c);
  return AG2_OK;
}
// test helper (no counterpart in ag2.h): normals of selected points only, the same per-point code
// as compute_normals, so that multi-million-point clouds can be spot-checked in seconds
int ag2o_point_normals(ag2o_ctx* c, const int32_t* idx, size_t n, float* out) {
  if (!c || !idx || !out) return AG2_E_INVALID;
  if (c->xyz.empty()) return AG2_E_EMPTY_CLOUD;
  ensure_tree(c);
  const size_t N = c->xyz.size() / 3;
#pragma omp parallel num_threads(c->threads())
  {
    std::vector<Neighbor> nn;
#pragma omp for schedule(dynamic, 8)
    for (long long k = 0; k < (long long)n; k++) {
      float* o = out + 3 * (size_t)k;
      if (idx[k] < 0 || (size_t)idx[k] >= N) {
        o[0] = o[1] = o[2] = std::numeric_limits<float>::quiet_NaN();
        continue;
      }
      const float* p = &c->xyz[3 * (size_t)idx[k]];
      radius_search(c, p, c->p.normals_radius, nn);
      pcl_normal(c->xyz.data(), nn, p, o);
    }
  }
  return AG2_OK;
}
int ag2o_generate_hypotheses(ag2o_ctx* c) {
  if (!c) return AG2_E_INVALID;
  if (c->xyz.empty()) return AG2_E_EMPTY_CLOUD;
  if (c->num_samples() == 0) return AG2_E_NO_SAMPLES;
  ensure_tree(c);
  if (!c->have_normals) compute_normals(c);
  compute_frames(c);
  evaluate_hands(c);
  return AG2_OK;
}
int ag2o_prune(ag2o_ctx* c) {
  if (!c) return AG2_E_INVALID;
  if (!c->have_hyps) return AG2_E_STATE;
  prune(c);
  return AG2_OK;
}
int ag2o_render_images(ag2o_ctx* c, uint8_t* out) {
  if (!c) return AG2_E_INVALID;
  if (!c->have_pruned) return AG2_E_STATE;
  render(c);
  if (out && !c->images.empty()) std::memcpy(out, c->images.data(), c->images.size());
  return AG2_OK;
}
int ag2o_classify(ag2o_ctx* c) {
  if (!c) return AG2_E_INVALID;
  if (!c->have_images) return AG2_E_STATE;
  if (!c->net.loaded) return AG2_E_NO_WEIGHTS;
  classify(c);
  return AG2_OK;
}
int ag2o_select(ag2o_ctx* c) {
  if (!c) return AG2_E_INVALID;
  if (!c->have_pruned) return AG2_E_STATE;
  if (c->p.antipodal_mode != AG2_MODE_PREDICTION) select_geometric(c);
  if (!c->have_antipodal) return AG2_E_STATE;
  select_top(c);
  return AG2_OK;
}
int ag2o_detect(ag2o_ctx* c) {
  int rc = ag2o_generate_hypotheses(c);
  if (rc) return rc;
  if ((rc = ag2o_prune(c))) return rc;
  if (c->p.antipodal_mode == AG2_MODE_PREDICTION) {
    if ((rc = ag2o_render_images(c, nullptr))) return rc;
    if ((rc = ag2o_classify(c))) return rc;
  }
  return ag2o_select(c);
}

// ---- pre-processing: CloudCamera::filterWorkspace / voxelizeCloud / subsampleUniformly
// (cloud_camera.cpp:89-121, :124-168, :171-178), literal form with std::set
int ag2o_preprocess_cloud(ag2o_ctx* c, const float* xyz, size_t stride, size_t n, const uint8_t* cam, int num_cams,
                          const float* normals, int filter_workspace, int voxelize, double voxel_size) {
  if (!c || (n && !xyz) || stride < 12) return AG2_E_INVALID;
  std::vector<float> pts, nrm;
  std::vector<uint8_t> cm;
  const double* ws = c->p.workspace;
  for (size_t i = 0; i < n; i++) {  // filterWorkspace, :92-99
    const float* p = (const float*)((const char*)xyz + i * stride);
    if (!filter_workspace || (p[0] > ws[0] && p[0] < ws[1] && p[1] > ws[2] && p[1] < ws[3] && p[2] > ws[4] && p[2] < ws[5])) {
      pts.insert(pts.end(), p, p + 3);
      cm.push_back(cam ? cam[i] : (uint8_t)1);
      if (normals && !voxelize) nrm.insert(nrm.end(), normals + 3 * i, normals + 3 * i + 3);
    }
  }
  if (voxelize && !pts.empty()) {  // voxelizeCloud, :124-168
    const size_t m = pts.size() / 3;
    float mn[3] = {INFINITY, INFINITY, INFINITY};
    for (size_t i = 0; i < m; i++)
      for (int d = 0; d < 3; d++)
        if (pts[3 * i + d] < mn[d]) mn[d] = pts[3 * i + d];
    const float cell = (float)voxel_size;  // Vector3f / double: the scalar is converted to float
    struct Less {
      bool operator()(const std::array<int, 3>& a, const std::array<int, 3>& b) const {
        for (int i = 0; i < 3; i++)
          if (a[i] != b[i]) return a[i] < b[i];
        return false;
      }
    };
    std::map<std::array<int, 3>, size_t, Less> bins;  // the std::set (the mapped value is unused)
    std::vector<size_t> idx_cam_source;               // :139-141: points that open a voxel, in scan order
    for (size_t i = 0; i < m; i++) {
      std::array<int, 3> v;
      bool finite = true;
      for (int d = 0; d < 3; d++) {
        float f = std::floor((pts[3 * i + d] - mn[d]) / cell);
        if (!(f == f)) finite = false;
        v[d] = finite ? (int)f : 0;
      }
      if (finite && bins.insert(std::make_pair(v, i)).second) idx_cam_source.push_back(i);
    }
    std::vector<float> vp;
    std::vector<uint8_t> vc;
    size_t vi = 0;
    for (const auto& kv : bins) {
      for (int d = 0; d < 3; d++) {
        float t = (float)kv.first[d] * cell;
        vp.push_back(t + mn[d]);
      }
      // :147-153 as written: the i-th voxel of the SORTED set takes the camera source of the i-th opener in
      // scan order (idx_cam_source[i]), not of its own first point
      vc.push_back(cm[idx_cam_source[vi++]]);
    }
    pts.swap(vp);
    cm.swap(vc);
  }
  return ag2o_set_cloud(c, pts.data(), 12, pts.size() / 3, cm.data(), num_cams, nrm.empty() ? nullptr : nrm.data());
}

int ag2o_get_cloud(ag2o_ctx* c, float* xyz, uint8_t* cam, size_t cap, size_t* n) {
  if (!c || !n) return AG2_E_INVALID;
  *n = c->xyz.size() / 3;
  if (*n > cap) return AG2_E_CAPACITY;
  if (xyz) std::copy(c->xyz.begin(), c->xyz.end(), xyz);
  if (cam) std::copy(c->cam.begin(), c->cam.end(), cam);
  return AG2_OK;
}

int ag2o_subsample_uniformly(ag2o_ctx* c, int num_samples, uint64_t seed) {
  if (!c || num_samples < 0) return AG2_E_INVALID;
  const size_t N0 = c->xyz.size() / 3;
  if ((size_t)num_samples > N0) return AG2_E_INVALID;
  std::vector<int32_t> out((size_t)num_samples);
  auto unif = [seed](uint64_t k) {
    uint64_t x = splitmix64(seed ^ (0xD1B54A32D192ED03ull * (k + 1)));
    return (float)((double)(x >> 11) * (1.0 / 9007199254740992.0));
  };
  // pcl::RandomSample::applyFilter (Vitter's algorithm A), restated
  size_t N = N0, top = N0 - (size_t)num_samples, index = 0, i = 0;
  uint64_t draw = 0;
  for (size_t m = (size_t)num_samples; m >= 2; m--) {
    float V = unif(draw++);
    size_t S = 0;
    float quot = (float)top / (float)N;
    while (quot > V) {
      S++;
      top--;
      N--;
      quot = quot * (float)top / (float)N;
    }
    index += S;
    out[i++] = (int32_t)index++;
    N--;
  }
  if (num_samples >= 1) {
    index += (size_t)((float)N * unif(draw++));
    if (index >= N0) index = N0 - 1;
    out[i++] = (int32_t)index;
  }
  return ag2o_set_sample_indices(c, out.data(), out.size());
}

int ag2o_get_sample_indices(ag2o_ctx* c, int32_t* idx, size_t cap, size_t* n) {
  if (!c || !n) return AG2_E_INVALID;
  *n = c->sample_idx.size();
  if (*n > cap) return AG2_E_CAPACITY;
  std::copy(c->sample_idx.begin(), c->sample_idx.end(), idx);
  return AG2_OK;
}

// ---- ImportanceSampling::detectGraspPoses, importance_sampling.cpp:30-112 (literal form of the
// samplers :115-157; RNG replaced by the documented counter-based stream D5, see importance.cu)
namespace {
inline uint64_t is_hash(uint64_t seed, uint64_t round, uint64_t cand, uint64_t j) {
  uint64_t k = splitmix64(seed ^ 0x1234ABCD5678EF01ull);
  k = splitmix64(k + round);
  k = splitmix64(k + cand);
  return splitmix64(k + j);
}
inline double is_u01(uint64_t h) { return ((double)(h >> 11) + 0.5) * (1.0 / 9007199254740992.0); }
inline double is_normal(uint64_t seed, uint64_t round, uint64_t cand, int comp) {
  double u1 = is_u01(is_hash(seed, round, cand, 2 * (uint64_t)comp));
  double u2 = is_u01(is_hash(seed, round, cand, 2 * (uint64_t)comp + 1));
  // Box-Muller on the exactly reproducible log / cos of include/ag2_detmath.h (deviation D5): the same source is
  // compiled into the device generator, which is what makes the candidates identical bit for bit
  return ag2_det_normal(u1, u2);
}
// the sample batch of one round, importance_sampling.cpp:47-86 with the samplers :115-157
void is_draw_round(ag2o_ctx* c, const ag2_is_params* ip, int it, const ag2_hypothesis* hands, size_t H,
                   double* samples) {
  const int ns = ip->num_samples;
  const int num_rand_samples = (int)(ip->prob_rand_samples * ns);  // :47
  const int num_gauss_samples = ns - num_rand_samples;
  const double sigma = ip->radius;
  const double term = 1.0 / std::sqrt(std::pow(2.0 * M_PI, 3.0) * std::pow(sigma, 3.0));  // :53
  const size_t npts = c->xyz.size() / 3;
  uint64_t cand = 0;
  int j = 0;
  while (j < num_gauss_samples) {  // :132-156 (method 2) / :118-124 (method 1)
    const size_t idx = (size_t)(is_hash(ip->seed, (uint64_t)it, cand, 6) % (uint64_t)H);
    double x[3];
    for (int a = 0; a < 3; a++) x[a] = hands[idx].surface[a] + is_normal(ip->seed, (uint64_t)it, cand, a) * sigma;
    cand++;
    bool take = true;
    if (ip->method == 2) {
      double maxp = 0;
      for (size_t k = 0; k < H; k++) {
        V3 d = v3(x[0] - hands[k].surface[0], x[1] - hands[k].surface[1], x[2] - hands[k].surface[2]);
        double p = dot3(d, d);
        p = term * std::exp((-1.0 / (2.0 * sigma)) * p);
        if (p > maxp) maxp = p;
      }
      V3 d = v3(x[0] - hands[idx].surface[0], x[1] - hands[idx].surface[1], x[2] - hands[idx].surface[2]);
      double p = dot3(d, d);
      p = term * std::exp((-1.0 / (2.0 * sigma)) * p);
      take = p >= maxp;
    }
    if (take) {
      for (int a = 0; a < 3; a++) samples[3 * (size_t)j + a] = x[a];
      j++;
    }
  }
  for (int jj = num_gauss_samples; jj < ns; jj++) {  // :79-86
    size_t r = (size_t)(is_hash(ip->seed, (uint64_t)it, 0xFFFFFFFFull + (uint64_t)(jj - num_gauss_samples), 7) %
                        (uint64_t)npts);
    for (int a = 0; a < 3; a++) samples[3 * (size_t)jj + a] = (double)c->xyz[3 * r + a];
  }
}
void is_append_selected(ag2o_ctx* c, int32_t sample_offset) {
  for (size_t i : c->selected) {
    ag2_hypothesis h = c->hyps[i].rec;
    h.sample_id += sample_offset;
    c->is_hands.push_back(h);
  }
}
}  // namespace

int ag2o_default_is_params(ag2_is_params* p) {
  if (!p) return AG2_E_INVALID;
  std::memset(p, 0, sizeof(*p));
  p->num_iterations = 5;
  p->num_samples = 50;
  p->num_init_samples = 100;
  p->method = 2;
  p->prob_rand_samples = 0.3;
  p->radius = 0.02;
  return AG2_OK;
}

int ag2o_detect_importance(ag2o_ctx* c, const ag2_is_params* ip) {
  if (!c || !ip) return AG2_E_INVALID;
  c->is_hands.clear();
  c->is_round_samples.clear();
  struct NoCluster {  // the rounds call detectGraspPoses(cloud_cam, false)
    ag2o_ctx* c;
    int saved;
    explicit NoCluster(ag2o_ctx* cc) : c(cc), saved(cc->p.min_inliers) { c->p.min_inliers = 0; }
    ~NoCluster() { c->p.min_inliers = saved; }
  } no_cluster(c);
  int rc = ag2o_detect(c);  // :38
  if (rc) return rc;
  int32_t sample_offset = 0;
  is_append_selected(c, sample_offset);
  sample_offset += (int32_t)c->num_samples();
  if (c->is_hands.empty()) return AG2_OK;
  const int ns = ip->num_samples;
  for (int it = 0; it < ip->num_iterations; it++) {
    std::vector<double> samples((size_t)3 * ns);
    is_draw_round(c, ip, it, c->is_hands.data(), c->is_hands.size(), samples.data());
    c->is_round_samples.push_back(samples);
    if ((rc = ag2o_set_sample_points(c, samples.data(), (size_t)ns))) return rc;  // :89
    if ((rc = ag2o_detect(c))) return rc;                                          // :90
    is_append_selected(c, sample_offset);                                          // :91
    sample_offset += ns;
  }
  if (no_cluster.saved > 0) find_clusters(c->is_hands, no_cluster.saved, c->clusters);  // :105-109
  return AG2_OK;
}

int ag2o_draw_importance_samples(ag2o_ctx* c, const ag2_is_params* ip, int round, const ag2_hypothesis* hands,
                                 size_t n_hands, double* xyz_out) {
  if (!c || !ip || !hands || !xyz_out || n_hands == 0 || round < 0) return AG2_E_INVALID;
  if (c->xyz.empty()) return AG2_E_EMPTY_CLOUD;
  is_draw_round(c, ip, round, hands, n_hands, xyz_out);
  return AG2_OK;
}

int ag2o_find_clusters(ag2o_ctx* c, int which_src) {
  if (!c) return AG2_E_INVALID;
  size_t n = 0;
  int rc = ag2o_num_hypotheses(c, which_src, &n);
  if (rc) return rc;
  std::vector<ag2_hypothesis> src(std::max<size_t>(n, 1));
  if ((rc = ag2o_get_hypotheses(c, which_src, src.data(), src.size(), &n))) return rc;
  src.resize(n);
  std::vector<ag2_hypothesis> out;
  find_clusters(src, c->p.min_inliers, out);
  c->clusters.swap(out);
  return AG2_OK;
}

int ag2o_get_importance_samples(ag2o_ctx* c, int round, double* xyz, size_t cap, size_t* n) {
  if (!c || !n || round < 0 || (size_t)round >= c->is_round_samples.size()) return AG2_E_INVALID;
  const std::vector<double>& v = c->is_round_samples[round];
  *n = v.size() / 3;
  if (v.size() / 3 > cap) return AG2_E_CAPACITY;
  std::copy(v.begin(), v.end(), xyz);
  return AG2_OK;
}

int ag2o_num_hypotheses(ag2o_ctx* c, int which, size_t* n) {
  if (!c || !n) return AG2_E_INVALID;
  if (which == AG2_LIST_IMPORTANCE || which == AG2_LIST_CLUSTERS) {
    *n = which == AG2_LIST_IMPORTANCE ? c->is_hands.size() : c->clusters.size();
    return AG2_OK;
  }
  if (which == AG2_LIST_SELECTED && c->selected_is_rec) {
    *n = c->selected_rec.size();
    return AG2_OK;
  }
  std::vector<size_t> tmp;
  const std::vector<size_t>* l = list_of(c, which, tmp);
  if (!l) return AG2_E_INVALID;
  *n = l->size();
  return AG2_OK;
}
int ag2o_get_hypotheses(ag2o_ctx* c, int which, ag2_hypothesis* out, size_t cap, size_t* n) {
  if (!c) return AG2_E_INVALID;
  if (which == AG2_LIST_IMPORTANCE || which == AG2_LIST_CLUSTERS || (which == AG2_LIST_SELECTED && c->selected_is_rec)) {
    const std::vector<ag2_hypothesis>& v =
        which == AG2_LIST_IMPORTANCE ? c->is_hands : (which == AG2_LIST_CLUSTERS ? c->clusters : c->selected_rec);
    if (n) *n = v.size();
    if (v.size() > cap) return AG2_E_CAPACITY;
    for (size_t i = 0; i < v.size(); i++) out[i] = v[i];
    return AG2_OK;
  }
  std::vector<size_t> tmp;
  const std::vector<size_t>* l = list_of(c, which, tmp);
  if (!l) return AG2_E_INVALID;
  if (n) *n = l->size();
  if (l->size() > cap) return AG2_E_CAPACITY;
  for (size_t i = 0; i < l->size(); i++) out[i] = c->hyps[(*l)[i]].rec;
  return AG2_OK;
}
int ag2o_get_stats(ag2o_ctx* c, ag2_stats* out) {
  if (!c || !out) return AG2_E_INVALID;
  *out = c->st;
  return AG2_OK;
}
int ag2o_get_normals(ag2o_ctx* c, float* out) {
  if (!c || !out) return AG2_E_INVALID;
  if (!c->have_normals) return AG2_E_STATE;
  std::memcpy(out, c->normals.data(), c->normals.size() * sizeof(float));
  return AG2_OK;
}
int ag2o_get_frames(ag2o_ctx* c, ag2_frame* out, size_t cap) {
  if (!c || !out) return AG2_E_INVALID;
  if (c->frames.size() > cap) return AG2_E_CAPACITY;
  std::copy(c->frames.begin(), c->frames.end(), out);
  return AG2_OK;
}
int ag2o_get_sweep_records(ag2o_ctx* c, ag2_sweep_record* out, size_t cap) {
  if (!c || !out) return AG2_E_INVALID;
  if (c->sweeps.size() > cap) return AG2_E_CAPACITY;
  std::copy(c->sweeps.begin(), c->sweeps.end(), out);
  return AG2_OK;
}
int ag2o_radius_search(ag2o_ctx* c, const float* q, double radius, int brute, int32_t* idx, float* d2,
                       size_t cap, size_t* n) {
  if (!c || !q || !n) return AG2_E_INVALID;
  std::vector<Neighbor> nn;
  if (brute) {
    radius_search_brute(c, q, radius, nn);
  } else {
    ensure_tree(c);
    radius_search(c, q, radius, nn);
  }
  *n = nn.size();
  if (nn.size() > cap) return AG2_E_CAPACITY;
  for (size_t i = 0; i < nn.size(); i++) {
    if (idx) idx[i] = nn[i].idx;
    if (d2) d2[i] = nn[i].d2;
  }
  return AG2_OK;
}
int ag2o_get_in_box(ag2o_ctx* c, int which, size_t h, int32_t* idx, size_t cap, size_t* n) {
  if (!c || !n) return AG2_E_INVALID;
  std::vector<size_t> tmp;
  const std::vector<size_t>* l = list_of(c, which, tmp);
  if (!l || h >= l->size()) return AG2_E_INVALID;
  const Hyp& hy = c->hyps[(*l)[h]];
  *n = hy.in_box.size();
  if (hy.in_box.size() > cap) return AG2_E_CAPACITY;
  std::copy(hy.in_box.begin(), hy.in_box.end(), idx);
  return AG2_OK;
}
// Learning::createGraspImage(points, normals, grasp) (learning.cpp:69-90): explicit unit-box point sets
int ag2o_render_point_sets(ag2o_ctx* c, const double* pts, const double* normals, const uint32_t* offsets,
                           size_t n_images, int aligns, uint8_t* out) {
  if (!c || !offsets || (n_images && !out)) return AG2_E_INVALID;
  if (offsets[n_images] && (!pts || !normals)) return AG2_E_INVALID;
#pragma omp parallel for num_threads(c->threads()) schedule(dynamic, 4)
  for (long long i = 0; i < (long long)n_images; i++)
    render_points(pts + 3 * (size_t)offsets[i], normals + 3 * (size_t)offsets[i], offsets[i + 1] - offsets[i], aligns != 0,
                  out + (size_t)i * AG2_IMAGE_BYTES);
  return AG2_OK;
}

int ag2o_get_logits(ag2o_ctx* c, float* out, size_t cap_images) {
  if (!c || !out) return AG2_E_INVALID;
  if (c->logits.size() / 2 > cap_images) return AG2_E_CAPACITY;
  std::copy(c->logits.begin(), c->logits.end(), out);
  return AG2_OK;
}
int ag2o_classify_images(ag2o_ctx* c, const uint8_t* images, size_t n, float* logits_out) {
  if (!c || (n && (!images || !logits_out))) return AG2_E_INVALID;
  if (!c->net.loaded) return AG2_E_NO_WEIGHTS;
  lenet_forward(c->net, images, n, logits_out, c->threads());
  return AG2_OK;
}

// test hooks: the exactly reproducible log / cos / normal draw of include/ag2_detmath.h (deviation D5), n values each
int ag2o_detmath(const double* u1, const double* u2, size_t n, double* log_u1, double* cos2pi_u2, double* normal) {
  for (size_t i = 0; i < n; i++) {
    log_u1[i] = ag2_det_log(u1[i]);
    cos2pi_u2[i] = ag2_det_cos2pi(u2[i]);
    normal[i] = ag2_det_normal(u1[i], u2[i]);
  }
  return AG2_OK;
}
int ag2o_eigen33(const float* cov9, float* eigenvalue, float* ev3) {
  eigen33(cov9, *eigenvalue, ev3);
  return AG2_OK;
}
int ag2o_average_normal_axis(const double* normals, int n, const double* sample3,
                             const double* cam_origin3, double* normal3, double* binormal3,
                             double* axis3) {
  std::vector<V3> N(n);
  for (int i = 0; i < n; i++) N[i] = v3(normals[3 * i], normals[3 * i + 1], normals[3 * i + 2]);
  V3 a, b, cax;
  average_normal_axis(N, v3(sample3[0], sample3[1], sample3[2]),
                      v3(cam_origin3[0], cam_origin3[1], cam_origin3[2]), a, b, cax);
  normal3[0] = a.x, normal3[1] = a.y, normal3[2] = a.z;
  binormal3[0] = b.x, binormal3[1] = b.y, binormal3[2] = b.z;
  axis3[0] = cax.x, axis3[1] = cax.y, axis3[2] = cax.z;
  return AG2_OK;
}
int ag2o_image_postprocess(const float* img, uint8_t* out) {
  image_postprocess(img, out);
  return AG2_OK;
}
int ag2o_finger_spacing(double fw, double od, double* out20) {
  finger_spacing(fw, od, out20);
  return AG2_OK;
}
int ag2o_deepen_depths(double init_bite, double hand_depth, double* out, int cap, int* n) {
  *n = deepen_depths(init_bite, hand_depth, out, cap);
  return AG2_OK;
}
int ag2o_orientation_angles(int R, double* out) {
  orientation_angles(R, out);
  return AG2_OK;
}

}  // extern "C"
