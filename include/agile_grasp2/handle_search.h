// Shim of HandleSearch::findClusters (include/agile_grasp2/handle_search.h:66-67,
// handle_search.cpp:4-80, remove_inliers = false).  findHandles is outside the path.
#ifndef AGILE_GRASP2_SHIM_HANDLE_SEARCH_H_
#define AGILE_GRASP2_SHIM_HANDLE_SEARCH_H_

#include "grasp_detector.h"

class HandleSearch {
 public:
  HandleSearch() : min_inliers_(-1) {}
  void setMinInliers(int m) { min_inliers_ = m; }  // handle_search.h:83-86
  int getMinInliers() const { return min_inliers_; }

  // Clusters list `which_src` (AG2_LIST_*) held by the detector's context, with this object's min_inliers when
  // setMinInliers was called (else the detector's own parameter).
  std::vector<GraspHypothesis> findClusters(GraspDetector& det, int which_src) {
    ag2_ctx* ctx = det.context();
    if (min_inliers_ >= 0) ag2shim::check(ag2_set_min_inliers(ctx, min_inliers_), "ag2_set_min_inliers", ctx);
    ag2shim::check(ag2_find_clusters(ctx, which_src), "ag2_find_clusters", ctx);
    return ag2shim::fetch(ctx, AG2_LIST_CLUSTERS);
  }

 private:
  int min_inliers_;  // -1: not set here
};

#endif
