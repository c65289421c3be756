// Host-side sample selection of the reference's pre-processing, the part that is plain
// index arithmetic and needs no device: the workspace filter on incoming sample points
// (GraspDetector::preprocessPointCloud, grasp_detector.cpp:305-313), CloudCamera::subsampleSamples
// (cloud_camera.cpp:181-206) and the draw from caller-supplied indices (grasp_detector.cpp:334-347).
// The reference draws with std::random_shuffle / rand(); here the draws come from a counter-based stream
// (splitmix64 of seed and position, deviation D5 of DESIGN.md) so that a run can be repeated.  Depends on nothing
// but the standard library: tests/test_cpp_shim.py checks it on the CPU against a numpy restatement.
#ifndef AGILE_GRASP2_SHIM_SAMPLE_SELECTION_H_
#define AGILE_GRASP2_SHIM_SAMPLE_SELECTION_H_

#include <array>
#include <cstddef>
#include <cstdint>
#include <utility>
#include <vector>

namespace ag2shim {

typedef std::array<double, 3> SamplePoint;

inline uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
// draw number `k` of purpose `what` from `seed`
inline uint64_t selection_draw(uint64_t seed, uint64_t what, uint64_t k) {
  return splitmix64(splitmix64(seed ^ (0xA5A5F00DBAADC0DEull + what)) + k);
}

// grasp_detector.cpp:305-313: the samples STRICTLY inside the workspace [xmin xmax ymin ymax zmin zmax], order kept
inline std::vector<SamplePoint> samplesInWorkspace(const std::vector<SamplePoint>& samples, const double ws[6]) {
  std::vector<SamplePoint> out;
  for (const SamplePoint& p : samples)
    if (p[0] > ws[0] && p[0] < ws[1] && p[1] > ws[2] && p[1] < ws[3] && p[2] > ws[4] && p[2] < ws[5]) out.push_back(p);
  return out;
}

// cloud_camera.cpp:181-206: positions (into the incoming list) of the samples that are kept.  Fewer incoming samples
// than num_samples: all of them in order (:198-205).  Otherwise the first num_samples entries of a shuffled
// 0..n-1 (:185-196); the shuffle is Fisher-Yates from the back, position i swapped with draw i % (i + 1).
inline std::vector<int> subsampledPositions(size_t n_incoming, int num_samples, uint64_t seed) {
  std::vector<int> seq(n_incoming);
  for (size_t i = 0; i < n_incoming; i++) seq[i] = (int)i;
  if (num_samples < 0 || (size_t)num_samples >= n_incoming) return seq;
  for (size_t i = n_incoming - 1; i > 0; i--) {
    size_t j = (size_t)(selection_draw(seed, 1, (uint64_t)i) % (uint64_t)(i + 1));
    std::swap(seq[i], seq[j]);
  }
  seq.resize((size_t)num_samples);
  return seq;
}

// grasp_detector.cpp:334-347: with caller-supplied cloud indices, num_samples draws WITH replacement when
// num_samples differs from their number and is below the cloud size, else the indices as they are
inline std::vector<int> drawFromIndices(const std::vector<int>& indices, int num_samples, size_t cloud_size, uint64_t seed) {
  if (indices.empty() || num_samples < 0) return indices;
  if ((size_t)num_samples != indices.size() && (size_t)num_samples < cloud_size) {
    std::vector<int> out((size_t)num_samples);
    for (int i = 0; i < num_samples; i++)
      out[(size_t)i] = indices[(size_t)(selection_draw(seed, 2, (uint64_t)i) % (uint64_t)indices.size())];
    return out;
  }
  return indices;
}

}  // namespace ag2shim

#endif  // AGILE_GRASP2_SHIM_SAMPLE_SELECTION_H_
