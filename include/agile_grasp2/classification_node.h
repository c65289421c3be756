// Shim of the reference's ClassificationNode service logic (src/nodes/classification_node.cpp:26-171):
// from per-face point densities of a candidate grasp pair it fills the unit box with points on the
// three visible faces, renders the 60 x 60 image (Learning::createGraspImage, aligns = false) and
// returns the LeNet score.  The ROS service plumbing (agile_grasp2::Classify request / response) is
// replaced by plain vectors; the arithmetic is host-side bookkeeping, image and score come from libag2.
#ifndef AGILE_GRASP2_SHIM_CLASSIFICATION_NODE_H_
#define AGILE_GRASP2_SHIM_CLASSIFICATION_NODE_H_

#include <cmath>
#include <string>
#include <vector>

#include "learning.h"

class ClassificationNode {
 public:
  // classification_node.cpp:6-22 (model / weights / labels come from ROS params there)
  ClassificationNode(const std::string& model_file, const std::string& trained_file, const std::string& label_file,
                     int num_threads = 1)
      : classifier_(model_file, trained_file, label_file), learning_(60, num_threads) {}

  // classificationCallback, classification_node.cpp:26-72: densities holds, per pair and per grasp,
  // 2 * NUM_FACES values; returns the mean score over the grasps of every pair.
  std::vector<double> classify(const std::vector<GraspHypothesis>& grasps, const std::vector<float>& densities) {
    const int NUM_FACES = 3;
    const int num_grasps = (int)grasps.size();
    const int per_pair = num_grasps * NUM_FACES * 2;
    std::vector<double> scores;
    if (per_pair == 0) return scores;
    const int num_pairs = (int)densities.size() / per_pair;
    scores.resize((size_t)num_pairs);
    for (int i = 0; i < num_pairs; i++) {
      float score = 0.0f;
      for (int j = 0; j < num_grasps; j++) {
        const int idx_start = i * per_pair + j * NUM_FACES * 2;
        score += calculatePredictionScore(grasps, densities, idx_start, idx_start + NUM_FACES * 2, j);
      }
      scores[(size_t)i] = score / (double)num_grasps;
    }
    return scores;
  }

  // classification_node.cpp:75-171
  float calculatePredictionScore(const std::vector<GraspHypothesis>& grasps, const std::vector<float>& densities,
                                 int idx_start, int /*idx_end*/, int grasp_idx) {
    ag2shim::Matrix3Xd points, normals;
    fillUnitBox(grasps[(size_t)grasp_idx], densities, idx_start, points, normals);
    ag2shim::Image img = learning_.createGraspImage(points, normals, grasps[(size_t)grasp_idx]);
    std::vector<Prediction> p = classifier_.Classify(img, false);
    return p[1].second - p[0].second;
  }

  // the point generator of calculatePredictionScore (classification_node.cpp:77-133), exposed for tests
  static void fillUnitBox(const GraspHypothesis& grasp, const std::vector<float>& densities, int idx_start,
                          ag2shim::Matrix3Xd& points_in_box, ag2shim::Matrix3Xd& normals_in_box) {
    const int NUM_SPACINGS = 6;
    const float MIN_SPACINGS[2] = {0.04f, 0.02f};
    const float ROW_LENGTH = 0.5f, COLUMN_LENGTH = 0.9f;
    const float LENGTH[6] = {COLUMN_LENGTH, COLUMN_LENGTH, COLUMN_LENGTH, COLUMN_LENGTH, ROW_LENGTH, ROW_LENGTH};
    const float X_START[6] = {0.25f, 0.25f, 0.75f, 0.75f, 0.25f, 0.25f};
    const float Y_START = 0.1f;
    const int VAR[6] = {1, 1, 1, 1, 0, 0};
    const ag2shim::Vector3d approach = grasp.getApproach(), axis = grasp.getAxis(), binormal = grasp.getBinormal();
    // normals = frame^T * [binormal, -binormal, -approach], frame = [binormal approach axis]
    const ag2shim::Vector3d fcol[3] = {binormal, approach, axis};
    double ncol[3][3];  // ncol[k] = the k-th face normal in the hand frame
    for (int k = 0; k < 3; k++) {
      const double sgn = k == 0 ? 1.0 : -1.0;
      const ag2shim::Vector3d& src = k == 2 ? approach : binormal;
      for (int r = 0; r < 3; r++) {
        double s = 0.0;
        for (int a = 0; a < 3; a++) s += fcol[r](a) * (sgn * src(a));
        ncol[k][r] = s;
      }
    }
    int total = 0;
    int num_pts[6];
    double spacings[6] = {0, 0, 0, 0, 0, 0};
    for (int i = 0; i < NUM_SPACINGS; i++) {
      const float dens = densities[(size_t)(idx_start + i)];
      if (dens > 0) {
        spacings[i] = MIN_SPACINGS[(int)std::floor(i / 4.0)] / dens;
        num_pts[i] = (int)std::floor(LENGTH[i] / spacings[i]) + 1;
        total += num_pts[i];
      } else {
        num_pts[i] = 0;
      }
    }
    points_in_box.resize(3, total);
    normals_in_box.resize(3, total);
    int k = 0;
    for (int i = 0; i < NUM_SPACINGS; i++) {
      double p[3] = {X_START[i], Y_START, 0.0};
      for (int j = 0; j < num_pts[i]; j++) {
        for (int r = 0; r < 3; r++) {
          normals_in_box(r, k) = ncol[(int)std::floor(i / 2.0)][r];
          points_in_box(r, k) = p[r];
        }
        p[VAR[i]] += spacings[i];
        k++;
      }
    }
  }

 private:
  Classifier classifier_;
  Learning learning_;
};

#endif
