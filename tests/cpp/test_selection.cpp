// Host-only check of include/agile_grasp2/sample_selection.h and CloudCamera::subsampleSamples: prints the choices
// for the cases tests/test_cpp_shim.py restates in numpy.  Needs no GPU (no ag2_ctx is created).
#include <agile_grasp2/cloud_camera.h>
#include <agile_grasp2/grasp_hypothesis.h>
#include <agile_grasp2/sample_selection.h>

#include <cstdio>
#include <cstdlib>

struct Pt {
  double x, y, z;
};
struct Msg {
  std::vector<Pt> samples;
};

int main(int argc, char** argv) {
  if (argc < 5) return 2;
  const size_t n = (size_t)std::atoll(argv[1]);
  const int k = std::atoi(argv[2]);
  const uint64_t seed = std::strtoull(argv[3], nullptr, 10);
  const size_t cloud = (size_t)std::atoll(argv[4]);
  std::printf("positions");
  for (int v : ag2shim::subsampledPositions(n, k, seed)) std::printf(" %d", v);
  std::printf("\n");
  std::vector<int> idx(n);
  for (size_t i = 0; i < n; i++) idx[i] = (int)(7 * i + 3);
  std::printf("indices");
  for (int v : ag2shim::drawFromIndices(idx, k, cloud, seed)) std::printf(" %d", v);
  std::printf("\n");
  // sample points on a lattice around the workspace [0,1] x [0,2] x [-1,0]: the bounds are strict
  const double ws[6] = {0.0, 1.0, 0.0, 2.0, -1.0, 0.0};
  std::vector<ag2shim::SamplePoint> pts;
  for (size_t i = 0; i < n; i++)
    pts.push_back(ag2shim::SamplePoint{{0.25 * (double)(i % 7) - 0.25, 0.5 * (double)(i % 6) - 0.5, -0.25 * (double)(i % 5)}});
  std::vector<ag2shim::SamplePoint> in = ag2shim::samplesInWorkspace(pts, ws);
  std::printf("inside %zu\n", in.size());
  Msg msg;
  for (const ag2shim::SamplePoint& p : in) msg.samples.push_back(Pt{p[0], p[1], p[2]});
  CloudCamera cc;
  cc.subsampleSamples(msg, k, seed);
  std::printf("samples");
  for (int c = 0; c < cc.getSamples().cols(); c++)
    std::printf(" %.17g %.17g %.17g", cc.getSamples()(0, c), cc.getSamples()(1, c), cc.getSamples()(2, c));
  std::printf("\n");
  // setSamples(msg) keeps every sample in order; writeNormalsToFile prints one "x,y,z" line per column
  cc.setSamples(msg);
  std::printf("all %d\n", cc.getSamples().cols());
  if (argc > 5) CloudCamera::writeNormalsToFile(argv[5], cc.getSamples());
  // GraspHypothesis: convertToGraspMsg narrows width and score to float32; writeHandsToFile's line format
  if (argc > 6) {
    std::vector<GraspHypothesis> hands;
    for (int h = 0; h < 3; h++) {
      ag2_hypothesis r = ag2_hypothesis();
      for (int a = 0; a < 3; a++) {
        r.bottom[a] = 0.1 * (h + 1) + a / 3.0;
        r.surface[a] = -0.2 * (h + 1) + a / 7.0;
        r.top[a] = 0.3 + a;
        r.axis[a] = a == 0;
        r.approach[a] = a == 1;
        r.binormal[a] = a == 2;
      }
      r.width = 0.05 + h / 300.0;
      r.score = 1.0 / 3.0 + h;
      hands.push_back(GraspHypothesis(r, nullptr));
    }
    GraspHypothesis::writeHandsToFile(argv[6], hands);
    GraspHypothesis::GraspMsg m = hands[1].convertToGraspMsg();
    std::printf("msg %.17g %.17g %.9g %.9g\n", m.bottom.x, m.surface.z, (double)m.width.data, (double)m.score.data);
  }
  return 0;
}
