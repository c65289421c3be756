// Shim of HandleSearch::findClusters (include/agile_grasp2/handle_search.h:66-67,
// handle_search.cpp:4-80, remove_inliers = false).  findHandles is outside the path.
#ifndef AGILE_GRASP2_SHIM_HANDLE_SEARCH_H_
#define AGILE_GRASP2_SHIM_HANDLE_SEARCH_H_

#include "grasp_detector.h"

class HandleSearch {
 public:
  HandleSearch() : min_inliers_(0) {}
  void setMinInliers(int m) { min_inliers_ = m; }
  int getMinInliers() const { return min_inliers_; }

  // Clusters list `which_src` (AG2_LIST_*) held by the detector's context.
  std::vector<GraspHypothesis> findClusters(GraspDetector& det, int which_src) {
    ag2_ctx* ctx = det.context();  // min_inliers comes from the detector's parameters
    ag2shim::check(ag2_find_clusters(ctx, which_src), "ag2_find_clusters", ctx);
    return ag2shim::fetch(ctx, AG2_LIST_CLUSTERS);
  }

 private:
  int min_inliers_;
};

#endif
