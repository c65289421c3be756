// The reference's test_cnn flow (src/tests/test_cnn.cpp:89-151: load cloud, filter + voxelise, hand search,
// grasp images, classify) and the detect_grasps_file flow (src/nodes/grasp_detection_node.cpp:98-120) written
// against the header-only shims in include/agile_grasp2/, with assertions.  Usage:
//   test_shim <cloud.pcd> <out.bin>
// out.bin: uint64 n_processed, uint64 n_samples, int32 samples[], uint64 n_hands, ag2_hypothesis hands[],
//          uint64 n_pairs, double pair_scores[] (ClassificationNode flow on the first two hands)
// (compared with the oracle by tests/test_cpp_shim.py).
#include <agile_grasp2/caffe_classifier.h>
#include <agile_grasp2/classification_node.h>
#include <agile_grasp2/cloud_camera.h>
#include <agile_grasp2/grasp_detector.h>
#include <agile_grasp2/hand_search.h>
#include <agile_grasp2/handle_search.h>
#include <agile_grasp2/importance_sampling.h>
#include <agile_grasp2/learning.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#define CHECK(cond)                                                        \
  do {                                                                     \
    if (!(cond)) {                                                         \
      std::fprintf(stderr, "CHECK failed %s:%d: %s\n", __FILE__, __LINE__, #cond); \
      return 1;                                                            \
    }                                                                      \
  } while (0)

int main(int argc, char** argv) {
  if (argc < 3) {
    std::fprintf(stderr, "usage: %s cloud.pcd out.bin\n", argv[0]);
    return 2;
  }
  try {
    // ---- detect_grasps_file flow
    CloudCamera cloud_cam(argv[1]);
    CHECK(cloud_cam.getCloudOriginal().size() > 1000);
    GraspDetector::Params prm;
    prm.workspace = {0.48, 0.92, -0.6, 0.4, -0.25, 1.0};  // launch/file_detect_grasps.launch:16
    prm.num_samples = 64;
    prm.num_threads = 4;
    prm.init_bite = 0.01;
    prm.antipodal_mode = AG2_MODE_PREDICTION;
    prm.min_score_diff = -1e9;
    prm.num_selected = 20;
    prm.rng_seed = 5;
    GraspDetector det(prm);
    det.preprocessPointCloud(cloud_cam);
    const size_t n_proc = cloud_cam.getCloudProcessed().size();
    CHECK(n_proc > 1000 && n_proc < cloud_cam.getCloudOriginal().size());
    CHECK(cloud_cam.getSampleIndices().size() == 64);
    std::vector<GraspHypothesis> hands = det.detectGraspPoses(cloud_cam);
    CHECK(!hands.empty() && hands.size() <= 20);
    for (size_t i = 1; i < hands.size(); i++) CHECK(hands[i - 1].getScore() >= hands[i].getScore());
    for (const GraspHypothesis& h : hands) {
      CHECK(h.isFullAntipodal());  // accepted by the classifier, grasp_detector.cpp:203
      CHECK(h.getGraspWidth() >= 0.03 && h.getGraspWidth() <= 0.07);
    }
    ag2_stats st = det.stats();
    CHECK(st.gpu_launches > 0 && st.num_points == (int64_t)n_proc);

    // ---- test_cnn flow: HandSearch -> Learning -> Classifier, one piece at a time
    HandSearch::Parameters hp;
    hp.nn_radius_taubin_ = 0.01, hp.nn_radius_hands_ = 0.1, hp.num_orientations_ = 8, hp.num_threads_ = 2;
    hp.num_samples_ = 64, hp.finger_width_ = 0.01, hp.hand_outer_diameter_ = 0.09, hp.hand_depth_ = 0.06;
    hp.hand_height_ = 0.02, hp.init_bite_ = 0.01;
    ag2_params dp;
    ag2_default_params(&dp);
    for (int i = 0; i < 16; i++) hp.cam_tf_left_.m[i] = dp.cam_tf_left[i], hp.cam_tf_right_.m[i] = dp.cam_tf_right[i];
    HandSearch hand_search(hp);
    hand_search.setSeed(prm.rng_seed);  // same neighbour draws as the detector
    std::vector<GraspHypothesis> hyps = hand_search.generateHypotheses(cloud_cam, 0, false);
    CHECK(hyps.size() >= hands.size());
    std::vector<GraspHypothesis> some(hyps.begin(), hyps.begin() + std::min<size_t>(hyps.size(), 6));
    Learning learn(60, 2);
    std::vector<ag2shim::Image> images = learn.createGraspImages(some, ag2shim::Matrix3Xd());
    CHECK(images.size() == some.size());
    size_t nonzero = 0;
    for (uint8_t b : images[0].data) nonzero += b != 0;
    CHECK(nonzero > 0);
    Classifier classifier("", "", "");  // Xavier fillers, seed 1 == the detector's weights
    std::vector<std::vector<Prediction> > pred = classifier.ClassifyBatch(images, 2);
    CHECK(pred.size() == images.size() && pred[0].size() == 2);
    // a hypothesis that the detector selected must carry score = logit[1] - logit[0] of its own image
    for (const GraspHypothesis& h : hands)
      for (size_t k = 0; k < hyps.size(); k++)
        if (hyps[k].record().sample_id == h.record().sample_id &&
            hyps[k].record().orientation_id == h.record().orientation_id) {
          std::vector<ag2shim::Image> one = learn.createGraspImages(std::vector<GraspHypothesis>(1, hyps[k]), ag2shim::Matrix3Xd());
          std::vector<Prediction> pr = classifier.Classify(one[0]);
          double score = (double)(pr[1].second - pr[0].second);
          // fp32 summation order depends on the batch split: tolerance on the logits' magnitude
          CHECK(std::fabs(score - h.getScore()) <= 1e-3 + 1e-5 * (std::fabs(pr[0].second) + std::fabs(pr[1].second)));
        }

    // ---- importance sampling + clustering through the same surface
    GraspDetector::Params ip = prm;
    ip.antipodal_mode = AG2_MODE_NONE;
    ImportanceSampling is(ip);
    is.samplingParameters().num_iterations = 2;
    is.samplingParameters().num_samples = 30;
    std::vector<GraspHypothesis> is_hands = is.detectGraspPoses(cloud_cam);
    CHECK(is_hands.size() > hands.size());
    std::vector<uint8_t> wire = GraspHypothesis::toGraspListMsg(hands, "base", 1, 2, 3);
    CHECK(wire.size() == 16 + 4 + 4 + hands.size() * 152);

    // ---- ClassificationNode service flow (classification_node.cpp:26-171): two pairs of per-face densities
    ClassificationNode cls_node("", "", "");
    std::vector<GraspHypothesis> two(hands.begin(), hands.begin() + std::min<size_t>(hands.size(), 2));
    std::vector<float> densities;
    for (size_t k = 0; k < 2 * two.size() * 6; k++) densities.push_back(k % 5 == 3 ? 0.0f : 0.35f + 0.11f * (float)(k % 7));
    std::vector<double> pair_scores = cls_node.classify(two, densities);
    CHECK(pair_scores.size() == 2);
    for (double s : pair_scores) CHECK(std::isfinite(s));

    // ---- dump for the oracle comparison
    FILE* f = std::fopen(argv[2], "wb");
    CHECK(f != nullptr);
    uint64_t v = n_proc;
    std::fwrite(&v, 8, 1, f);
    v = cloud_cam.getSampleIndices().size();
    std::fwrite(&v, 8, 1, f);
    for (int s : cloud_cam.getSampleIndices()) {
      int32_t s32 = s;
      std::fwrite(&s32, 4, 1, f);
    }
    v = hands.size();
    std::fwrite(&v, 8, 1, f);
    for (const GraspHypothesis& h : hands) std::fwrite(&h.record(), sizeof(ag2_hypothesis), 1, f);
    v = pair_scores.size();
    std::fwrite(&v, 8, 1, f);
    std::fwrite(pair_scores.data(), 8, pair_scores.size(), f);
    std::fclose(f);
    std::printf("test_shim ok: %zu processed points, %zu hypotheses, %zu selected, %zu importance hands\n", n_proc,
                hyps.size(), hands.size(), is_hands.size());
  } catch (const std::exception& e) {
    std::fprintf(stderr, "exception: %s\n", e.what());
    return 1;
  }
  return 0;
}
