// Two-GPU C++ test of the multi-GPU entry points behind the C ABI (include/ag2.h ag2_group_*) and of the shim's
// GraspDetector::Params::num_gpus: the caller of the reference (grasp_detection_node.cpp:109-115) hands one
// CloudCamera to GraspDetector::detectGraspPoses; with num_gpus = 2 the same call is sharded over two GPUs by
// host threads inside libag2 and must return the single-GPU list byte for byte.
//   usage: test_group <cloud.pcd> <num_samples> [n_gpus]      (exit 0 = equal; prints the counts)
#include <agile_grasp2/cloud_camera.h>
#include <agile_grasp2/grasp_detector.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>

static bool same(const std::vector<GraspHypothesis>& a, const std::vector<GraspHypothesis>& b) {
  if (a.size() != b.size()) return false;
  for (size_t i = 0; i < a.size(); i++)
    if (memcmp(&a[i].record(), &b[i].record(), sizeof(ag2_hypothesis)) != 0) return false;
  return true;
}

int main(int argc, char** argv) {
  if (argc < 3) {
    fprintf(stderr, "usage: %s <cloud.pcd> <num_samples> [n_gpus]\n", argv[0]);
    return 2;
  }
  const int n_gpus = argc > 3 ? atoi(argv[3]) : 2;
  GraspDetector::Params prm;
  prm.workspace = {-10, 10, -10, 10, -10, 10};
  prm.num_samples = atoi(argv[2]);
  prm.voxelize = false;
  prm.init_bite = 0.01;
  prm.min_score_diff = -1e9;
  prm.num_selected = 20;
  prm.rng_seed = 5;
  try {
    GraspDetector one(prm);
    prm.num_gpus = n_gpus;
    GraspDetector many(prm);
    CloudCamera cc(argv[1], "");
    one.preprocessPointCloud(cc);
    const std::vector<GraspHypothesis> want = one.detectGraspPoses(cc, false);
    bool ok = !want.empty();
    for (int rep = 0; rep < 3; rep++) {
      const std::vector<GraspHypothesis> got = many.detectGraspPoses(cc, false);
      const bool eq = same(want, got);
      printf("detectGraspPoses over %d GPUs, call %d: %zu grasps, %s\n", n_gpus, rep, got.size(),
             eq ? "equal to one GPU" : "DIFFERENT from one GPU");
      ok = ok && eq;
    }
    // importance sampling over the group
    ImportanceSampling is_one(([&] { GraspDetector::Params q = prm; q.num_gpus = 1; return q; })());
    ImportanceSampling is_many(prm);
    is_one.samplingParameters().num_iterations = is_many.samplingParameters().num_iterations = 2;
    is_one.samplingParameters().num_samples = is_many.samplingParameters().num_samples = 40;
    const std::vector<GraspHypothesis> w2 = is_one.detectGraspPoses(cc);
    const std::vector<GraspHypothesis> g2 = is_many.detectGraspPoses(cc);
    const bool eq2 = same(w2, g2);
    printf("ImportanceSampling over %d GPUs: %zu grasps, %s\n", n_gpus, g2.size(), eq2 ? "equal to one GPU" : "DIFFERENT from one GPU");
    ok = ok && eq2 && !w2.empty();
    printf("TEST_GROUP %s\n", ok ? "OK" : "FAILED");
    return ok ? 0 : 1;
  } catch (const std::exception& e) {
    fprintf(stderr, "test_group: %s\n", e.what());
    return 3;
  }
}
